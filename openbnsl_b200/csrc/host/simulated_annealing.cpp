// simulated_annealing.cpp -- multi-chain simulated annealing over DAGs with the
// reference's semantics (src/structure_learning/simulated_annealing.cpp:15-159),
// as a consumer of the batched BDeu call.
//
// The reference runs one chain per OpenMP thread; every step of a chain draws a
// child, proposes one parent to add or remove and evaluates ONE local score by a
// pass over all N rows (`:77-81`).  Chains are independent, so here all chains
// advance in LOCKSTEP: step `it` of every chain is proposed on the host, the
// distinct (child, parent set) pairs not seen before go to the GPU as one batch
// (`Evaluator::child_scores`, exact-order mode => bit-identical scores), and
// every chain then takes its accept/reject decision.  A chain consumes its own
// `std::mt19937_64` in exactly the reference's order (`:37-40,44,53,61,75,85`),
// with the same standard-library distributions and containers, so for the same
// seed each chain visits the same states and the returned DAG is identical.
// Scores are memoised per (child, parent set): late in a run most proposals are
// revisits and never reach the GPU.
//
// Multi-GPU: chains are independent objects -- rank r owns chains r, r+world, ...;
// one all-gather of (best score, adjacency) per search picks the winner.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <random>
#include <set>
#include <thread>
#include <unordered_map>
#include <unordered_set>

#include "structure_learning.hpp"

namespace obn {
namespace {

template <bool Det>
using ParentSet = std::conditional_t<Det, std::set<std::size_t>, std::unordered_set<std::size_t>>;   // pdag_with_adjlist.h:5-8

// pdag_with_adjlist.h:42-58: is dst reachable from src along parent links (i.e. is dst an ancestor of src)
template <bool Det>
bool has_path(const std::vector<ParentSet<Det>>& parents, std::size_t src, std::size_t dst) {
  std::vector<bool> visited(parents.size(), false);
  std::vector<std::size_t> stack{src};
  while (!stack.empty()) {
    const std::size_t v = stack.back();
    stack.pop_back();
    if (v == dst) return true;
    if (visited[v]) continue;
    visited[v] = true;
    for (const auto& u : parents[v])
      if (!visited[u]) stack.push_back(u);
  }
  return false;
}

struct KeyHash {
  size_t operator()(const std::vector<uint32_t>& k) const {
    uint64_t h = 0x9e3779b97f4a7c15ull;
    for (uint32_t v : k) h = (h ^ v) * 0x100000001b3ull + (h >> 29);
    return (size_t)h;
  }
};
// key = {child, parents ascending}
using ScoreCache = std::unordered_map<std::vector<uint32_t>, double, KeyHash>;

template <bool Det>
struct Chain {
  std::vector<ParentSet<Det>> parents, best_parents;
  std::vector<double> ls;
  double curr = 0, best = 0;
  std::mt19937_64 rng;
  std::uniform_real_distribution<double> uni{0.0, 1.0};
  std::uniform_int_distribution<std::size_t> pick_node;
  // the step's proposal
  bool has = false;
  std::size_t child = 0;
  std::vector<std::size_t> new_pa;
  const double* new_ls = nullptr;      // into the cache (node addresses are stable)
  Chain(std::size_t n, uint64_t seed) : parents(n), best_parents(n), ls(n, 0.0), rng(seed), pick_node(0, n - 1) {}
};

// simulated_annealing.cpp:43-75: draw the child and the move; false = no move this step (`:72`)
template <bool Det>
bool propose(Chain<Det>& c, std::size_t n, std::size_t max_parents) {
  c.child = c.pick_node(c.rng);
  const auto& pa = c.parents[c.child];
  bool add;
  if (pa.empty())
    add = true;
  else if (pa.size() == max_parents)
    add = false;
  else
    add = c.uni(c.rng) < 0.5;
  c.new_pa.assign(pa.begin(), pa.end());
  if (add) {
    std::vector<std::size_t> cand(n);
    std::iota(cand.begin(), cand.end(), 0);
    cand.erase(std::remove(cand.begin(), cand.end(), c.child), cand.end());
    std::shuffle(cand.begin(), cand.end(), c.rng);
    for (auto p : cand) {
      if (pa.count(p)) continue;
      if (has_path<Det>(c.parents, p, c.child)) continue;
      c.new_pa.push_back(p);
      std::sort(c.new_pa.begin(), c.new_pa.end());
      return true;
    }
    return false;
  }
  std::uniform_int_distribution<std::size_t> pick_par(0, c.new_pa.size() - 1);
  c.new_pa.erase(c.new_pa.begin() + (std::ptrdiff_t)pick_par(c.rng));
  return true;
}

template <bool Det>
PDAG run(std::size_t n, Evaluator& ev, const Comm& comm, const SaOptions& opt, SaStats* stats) {
  using clock = std::chrono::steady_clock;
  const std::size_t n_chains = opt.num_chains;
  ScoreCache cache;
  TupleBatch batch;
  std::vector<uint32_t> batch_child;
  std::vector<double*> batch_slot;
  std::vector<double> scores;
  double eval_ms = 0;
  size_t evaluated = 0, proposals = 0, batches = 0;
  auto flush = [&]() {
    if (!batch.size()) return;
    const auto t0 = clock::now();
    scores.resize(batch.size());
    ev.child_scores(batch, batch_child.data(), opt.ess, /*exact_order=*/1, scores.data());
    for (size_t i = 0; i < batch.size(); ++i) *batch_slot[i] = scores[i];
    evaluated += batch.size();
    ++batches;
    batch.clear();
    batch_child.clear();
    batch_slot.clear();
    eval_ms += std::chrono::duration<double, std::milli>(clock::now() - t0).count();
  };
  // score of (child | pa): cached, or queued for the next flush; returns where it will be
  std::vector<uint32_t> key;
  auto request = [&](std::size_t child, const std::vector<std::size_t>& pa) -> const double* {
    key.clear();
    key.push_back((uint32_t)child);
    for (auto p : pa) key.push_back((uint32_t)p);
    std::sort(key.begin() + 1, key.end());          // after a REMOVE new_pa is in the container's order (`:74-75`)
    auto ins = cache.emplace(key, 0.0);
    if (ins.second) {
      std::vector<uint32_t> vars(key.begin() + 1, key.end());       // `:77-79`: parents + child, ascending
      vars.insert(std::upper_bound(vars.begin(), vars.end(), (uint32_t)child), (uint32_t)child);
      batch.var_ids.insert(batch.var_ids.end(), vars.begin(), vars.end());
      batch.offsets.push_back(batch.var_ids.size());
      batch_child.push_back((uint32_t)child);
      batch_slot.push_back(&ins.first->second);
    }
    return &ins.first->second;
  };

  // this rank's chains (chain c uses seed + c * 1234567, `:131,149`)
  std::vector<std::size_t> mine;
  for (std::size_t c = (std::size_t)comm.rank; c < n_chains; c += (std::size_t)comm.world) mine.push_back(c);
  std::vector<Chain<Det>> chains;
  chains.reserve(mine.size());
  for (std::size_t c : mine) chains.emplace_back(n, opt.seed + (uint64_t)c * 1234567ULL);

  // `:27-35`: empty graph, ls[v] = score(v | {})
  std::vector<const double*> empty_ls(n);
  const std::vector<std::size_t> none;
  for (std::size_t v = 0; v < n; ++v) empty_ls[v] = request(v, none);
  flush();
  for (auto& c : chains) {
    for (std::size_t v = 0; v < n; ++v) c.ls[v] = *empty_ls[v];
    c.curr = std::accumulate(c.ls.begin(), c.ls.end(), 0.0);
    c.best = c.curr;
  }

  double T = opt.init_temp;
  for (std::size_t it = 0; it < opt.max_iters; ++it, T *= opt.cooling_rate) {
    for (auto& c : chains) {
      c.has = propose<Det>(c, n, opt.max_parents);
      if (c.has) {
        c.new_ls = request(c.child, c.new_pa);
        ++proposals;
      }
    }
    flush();
    for (auto& c : chains) {
      if (!c.has) continue;
      const double new_ls = *c.new_ls;
      const double delta = new_ls - c.ls[c.child];
      if (delta >= 0.0 || std::exp(delta / T) > c.uni(c.rng)) {                     // `:85`
        const ParentSet<Det> np(c.new_pa.begin(), c.new_pa.end());                   // `:86-87`: range-constructed, then
        c.parents[c.child] = np;                                                     // copy-assigned (set_parents)
        c.ls[c.child] = new_ls;
        c.curr += delta;
        if (c.curr > c.best) {
          c.best = c.curr;
          c.best_parents = c.parents;
        }
      }
    }
  }

  // the best chain: first maximum in chain order (`:135-137`), over all ranks
  std::vector<double> best_of(n_chains, -1e100);
  for (size_t i = 0; i < mine.size(); ++i) best_of[mine[i]] = chains[i].best;
  std::vector<uint8_t> adj_of;                                   // adjacency of every chain's best graph, [chain][v][p]
  const size_t per_rank = (n_chains + (size_t)comm.world - 1) / (size_t)comm.world;
  if (comm.world > 1) {
    // fixed-size record per rank: per_rank x (score, n*n adjacency bytes)
    const size_t rec = sizeof(double) + n * n;
    std::vector<uint8_t> send(per_rank * rec, 0), recv(per_rank * rec * (size_t)comm.world, 0);
    for (size_t i = 0; i < per_rank; ++i) {
      double s = -1e100;
      if (i < chains.size()) {
        s = chains[i].best;
        for (std::size_t v = 0; v < n; ++v)
          for (auto p : chains[i].best_parents[v]) send[i * rec + sizeof(double) + v * n + p] = 1;
      }
      std::memcpy(&send[i * rec], &s, sizeof(double));
    }
    comm.allgather(send.data(), recv.data(), send.size());
    adj_of.assign(n_chains * n * n, 0);
    for (int r = 0; r < comm.world; ++r)
      for (size_t i = 0; i < per_rank; ++i) {
        const size_t c = (size_t)r + i * (size_t)comm.world;
        if (c >= n_chains) continue;
        const uint8_t* p = &recv[((size_t)r * per_rank + i) * rec];
        std::memcpy(&best_of[c], p, sizeof(double));
        std::memcpy(&adj_of[c * n * n], p + sizeof(double), n * n);
      }
  }
  const size_t best_idx = (size_t)std::distance(best_of.begin(), std::max_element(best_of.begin(), best_of.end()));
  PDAG result(n);
  if (comm.world > 1) {
    for (std::size_t v = 0; v < n; ++v)
      for (std::size_t p = 0; p < n; ++p)
        if (adj_of[best_idx * n * n + v * n + p]) result.add_edge(p, v);
  } else if (!chains.empty()) {
    for (std::size_t v = 0; v < n; ++v)                                            // to_pdag, pdag_with_adjlist.h:64-72
      for (auto p : chains[best_idx].best_parents[v]) result.add_edge(p, v);
  }
  if (stats) {
    stats->chains = n_chains;
    stats->proposals = proposals;
    stats->evaluated = evaluated;
    stats->batches = batches;
    stats->evaluate_ms = eval_ms;
    stats->best_score = n_chains ? best_of[best_idx] : 0.0;
    stats->best_chain = best_idx;
  }
  return result;
}

}  // namespace

PDAG simulated_annealing_driver(std::size_t n_vars, Evaluator& ev, const Comm& comm, const SaOptions& opt_in, SaStats* stats) {
  SaOptions opt = opt_in;
  if (opt.max_parents >= n_vars) throw std::invalid_argument("max_parents out of range");   // `:109-111`
  if (opt.num_chains == 0) {                                                               // `:113-115`: omp_get_max_threads()
    const char* e = std::getenv("OMP_NUM_THREADS");
    const long v = e && *e ? std::atol(e) : 0;
    opt.num_chains = v > 0 ? (std::size_t)v : std::max(1u, std::thread::hardware_concurrency());
  }
  return opt.deterministic ? run<true>(n_vars, ev, comm, opt, stats) : run<false>(n_vars, ev, comm, opt, stats);
}

}  // namespace obn
