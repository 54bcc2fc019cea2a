// structure_learning.hpp -- the two drivers that consume the hot path.
//   pc                 src/structure_learning/pc.cpp:16-86
//   exhaustive_search  src/structure_learning/exhaustive_search.cpp:14-165
// The drivers are written against Evaluator / Comm (engine.hpp) so that the
// same code runs on one GPU, on N GPUs (tuples sharded over ranks) and, in the
// CPU test-suite only, against a checker-backed evaluator.
#pragma once
#include <cstdint>
#include <functional>
#include <vector>

#include "engine.hpp"
#include "graph.hpp"
#include "types.hpp"

namespace obn {

struct PcStats {
  std::vector<uint64_t> tests_per_depth;   // CI tests evaluated (after de-duplication)
  std::vector<uint64_t> batches_per_depth;
  double enumerate_ms = 0, evaluate_ms = 0, resolve_ms = 0;   // host wall time of the three driver phases
  // pipelined rounds: host time with NO round in flight (depth starts and ends: nothing hides it), and time the
  // host spent blocked in collect (the device was the one being waited for)
  double exposed_ms = 0, collect_wait_ms = 0;
  bool record = false;                     // keep every evaluated test in `recorded`
  TupleBatch recorded;
  uint64_t total_tests() const {
    uint64_t s = 0;
    for (auto v : tests_per_depth) s += v;
    return s;
  }
};

// `decide` answers a batch of CI questions: p-values for statistical tests
// (independent iff p >= level).  cards are only needed to size nothing here;
// the driver is purely combinatorial.
struct PcOptions {
  size_t max_cond_vars = 0;
  bool stable = true;
  int kind = OBN_CI_G2;
  double level = 0.05;
  size_t target_batch = 8192;     // tests per round the driver aims for
  // stable mode: keep the evaluator one round ahead of the host -- round r+1 is enumerated and submitted
  // (chained: the evaluator drops the tests of edges round r decides) before round r is collected
  // (Evaluator::pc_round_submit / pc_round_collect); the learned graph does not depend on it
  bool overlap = false;
  // multi-GPU: a round is sharded over ranks (+ one all-reduce) only when it holds at least this
  // many variable slots (sum of tuple widths); smaller rounds are evaluated redundantly by every
  // rank, which costs less than the collective's latency and needs no exchange at all
  size_t shard_min_slots = 0;
};

PDAG pc_skeleton_and_colliders(size_t n_vars, Evaluator& ev, const Comm& comm, const PcOptions& opt,
                               PcStats* stats);

// OracleGraph variant (d-separation on a known graph; host only, no data)
PDAG pc_oracle(size_t n_vars, const PDAG& truth, size_t max_cond_vars, bool stable);

struct EsOptions {
  double ess = 1.0;
  size_t max_parents = 0;
  int exact_order = 2;            // 0 fast, 1 exact (error if tables do not fit), 2 auto
  std::vector<uint32_t> cards;    // cardinalities (device driver: cost model of the lattice mode)
  double shard_min_bytes = 0;     // multi-GPU: shard a level only if it has this many algorithmic bytes
  size_t dp_shard_min_subsets = 4096;   // multi-GPU lattice mode: shard the DP by child from the first level this large
  int force_lattice = -1;         // -1 auto (by estimated counting time), 0 never, 1 always (exact_order != 1 only)
};
struct EsStats {
  uint64_t tables = 0, scores = 0;
  double best_score = 0.0;
  double enumerate_ms = 0, score_ms = 0, gather_ms = 0, dp_ms = 0;
  bool lattice = false;           // scores came from the marginalisation lattice (fast mode)
};

PDAG exhaustive_search_driver(size_t n_vars, Evaluator& ev, const Comm& comm, const EsOptions& opt,
                              EsStats* stats);

// The product path: scores AND the DP tables stay in HBM (obn_es_* in the C ABI); the host only
// enumerates the subsets of each level and backtracks at the end.  Multi-GPU: the level's subsets
// are split into `world` equal ranges, each rank scores its range, one in-place all-gather of the
// level's score buffer over NVLink, then every rank runs the (cheap) DP for the whole level.
PDAG exhaustive_search_device(obn_dataset* ds, const Comm& comm, const EsOptions& opt, EsStats* stats);

// Multi-chain simulated annealing (src/structure_learning/simulated_annealing.cpp:15-159): all
// chains in lockstep, one batch of new (child, parent set) scores per step (simulated_annealing.cpp).
struct SaOptions {
  double ess = 1.0;
  size_t max_parents = 0;
  size_t max_iters = 10000;
  double init_temp = 1.0;
  double cooling_rate = 0.9995;
  bool deterministic = false;     // std::set vs std::unordered_set parent sets (iteration order feeds the moves)
  uint64_t seed = 0;
  size_t num_chains = 0;          // 0 = one per host thread, like the reference's omp_get_max_threads()
};
struct SaStats {
  size_t chains = 0, proposals = 0, evaluated = 0, batches = 0, best_chain = 0;
  double best_score = 0.0, evaluate_ms = 0.0;
};
PDAG simulated_annealing_driver(size_t n_vars, Evaluator& ev, const Comm& comm, const SaOptions& opt, SaStats* stats);

}  // namespace obn
