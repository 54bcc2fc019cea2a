// exhaustive_search.cpp -- Silander-Myllymaki DP over the subset lattice with
// the reference's tie-breaking (src/structure_learning/exhaustive_search.cpp:14-165),
// restructured level-synchronously: all subsets of one size form ONE batch of
// tuples whose |W| local scores each come from one table (the reference's own
// sharing, :55-72).  Multi-GPU: a level's subsets are sharded over ranks in
// Gosper order and the scores are all-gathered once per level.
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>

#include "structure_learning.hpp"

namespace obn {

namespace {

// include/utils/next_combmask.h:41-46
inline bool next_combmask(uint64_t& comb, int n) {
  if (comb == 0) return false;
  const uint64_t x = comb & (~comb + 1);
  const uint64_t y = comb + x;
  comb = (((comb & ~y) / x) >> 1) | y;
  return comb < ((uint64_t)1 << n);
}

}  // namespace

PDAG exhaustive_search_driver(size_t n_vars, Evaluator& ev, const Comm& comm, const EsOptions& opt,
                              EsStats* stats) {
  const size_t n = n_vars;
  if (n == 0) return PDAG(0);
  if (opt.max_parents > n - 1) throw std::invalid_argument("max_parents must be in [0, num_of_vars-1]");
  if (n > 64) throw std::invalid_argument("num_of_vars must be in [0, 64]");
  if (n > 30) throw std::runtime_error("exhaustive_search: 2^n DP tables for n > 30 do not fit memory");
  const size_t max_varset = opt.max_parents + 1;
  const size_t n_combs = (size_t)1 << n;

  std::vector<double> best_gs(n_combs, 0.0);
  std::vector<uint8_t> best_ch(n_combs, 0xff);
  std::vector<double> bl_score(n_combs * n);        // [W_without_child * n + child]
  std::vector<uint32_t> bl_ps(n_combs * n);
  const double ninf = -std::numeric_limits<double>::infinity();

  TupleBatch batch;
  std::vector<double> scores, gathered;
  for (size_t size = 0; size <= n; ++size) {                       // :47
    const bool calc_ls = size <= max_varset;                       // :48
    // ---- the level's tuples, in Gosper order (:51,:129) --------------------
    if (calc_ls && size > 0) {
      batch.clear();
      uint64_t W = ((uint64_t)1 << size) - 1;
      do {
        for (uint64_t m = W; m; m &= m - 1) batch.var_ids.push_back((uint32_t)__builtin_ctzll(m));
        batch.offsets.push_back(batch.var_ids.size());
      } while (next_combmask(W, (int)n));
      const size_t T = batch.size();
      scores.assign(T * size, 0.0);
      if (comm.world > 1) {
        size_t b, e;
        shard_range(T, comm.rank, comm.world, &b, &e);
        const size_t chunk = (T + comm.world - 1) / comm.world;   // equal-size slots for all-gather
        std::vector<double> mine(chunk * size, 0.0);
        ev.subset_scores(batch, b, e, opt.ess, opt.exact_order, mine.data());
        gathered.assign(chunk * size * comm.world, 0.0);
        comm.allgather(mine.data(), gathered.data(), chunk * size * sizeof(double));
        for (int r = 0; r < comm.world; ++r) {
          size_t rb, re;
          shard_range(T, r, comm.world, &rb, &re);
          std::memcpy(scores.data() + rb * size, gathered.data() + (size_t)r * chunk * size,
                      (re - rb) * size * sizeof(double));
        }
      } else {
        ev.subset_scores(batch, 0, T, opt.ess, opt.exact_order, scores.data());
      }
      if (stats) {
        stats->tables += T;
        stats->scores += T * size;
      }
    } else if (calc_ls && stats) {
      stats->tables += 1;   // the empty subset's table (N rows in one cell); no child, no score
    }
    // ---- DP for this level (:59-116) -----------------------------------------
    uint64_t W = ((uint64_t)1 << size) - 1;
    size_t t = 0;
    do {
      double bgs = 0;
      int bch = -1;
      size_t ci = 0;
      for (uint64_t m = W; m; m &= m - 1, ++ci) {
        const uint32_t child = (uint32_t)__builtin_ctzll(m);
        const uint64_t pa = W - ((uint64_t)1 << child);
        const double ls = calc_ls ? scores[t * size + ci] : ninf;        // :65-73
        double bls = ls;
        uint32_t bps = (uint32_t)pa;
        for (uint64_t q = pa; q; q &= q - 1) {                           // :75-87
          const uint64_t sub = pa - (q & (~q + 1));
          const double s = bl_score[sub * n + child];
          if (s >= bls) {
            bls = s;
            bps = bl_ps[sub * n + child];
          }
        }
        bl_score[pa * n + child] = bls;
        bl_ps[pa * n + child] = bps;
        const double gs = best_gs[pa] + bls;                             // :89-94
        if (bgs >= 0 || gs > bgs) {
          bgs = gs;
          bch = (int)child;
        }
      }
      best_gs[W] = bgs;
      best_ch[W] = (uint8_t)bch;
      ++t;
    } while (next_combmask(W, (int)n));
  }

  PDAG g(n);                                                            // :133-148
  uint64_t W = n_combs - 1;
  for (size_t i = 0; i < n; ++i) {
    const uint8_t ch = best_ch[W];
    const uint64_t pa = W - ((uint64_t)1 << ch);
    const uint32_t ps = bl_ps[pa * n + ch];
    for (size_t j = 0; j < n; ++j)
      if ((ps >> j) & 1u) g.add_edge(j, ch);
    W = pa;
  }
  if (stats) stats->best_score = best_gs[n_combs - 1];
  return g;
}

PDAG exhaustive_search_device(obn_dataset* ds, const Comm& comm, const EsOptions& opt, EsStats* stats) {
  const size_t n = obn_dataset_n_vars(ds);
  if (n == 0) return PDAG(0);
  obn_ctx* ctx = Engine::instance().ctx();
  obn_es* es = nullptr;
  check(obn_es_begin(ctx, ds, opt.ess, opt.max_parents, opt.exact_order, &es));
  struct Guard {
    obn_es* e;
    ~Guard() { obn_es_end(e); }
  } guard{es};
  using clk = std::chrono::steady_clock;
  auto ms = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
  const size_t n_rows = obn_dataset_n_rows(ds);
  const size_t max_varset = opt.max_parents + 1;
  // Fast mode for deep searches: every F(S) from the marginalisation lattice over the joint table
  // (one pass over the data) instead of one counted table per subset.  Chosen when exact-order
  // scores are not demanded and the lattice (~12 bytes per cell of every table) is cheaper than
  // counting (~1 row visit per row per table at ~1e12 /s).
  bool lattice = false;
  if (opt.exact_order != 1 && opt.cards.size() == n) {
    double count_s = 0, binom = 1, lat_cells = 1;
    for (size_t k = 0; k <= std::min(n, max_varset); ++k) {
      if (k) binom = binom * (double)(n - k + 1) / (double)k;
      count_s += binom * (double)n_rows / 1e12;
    }
    for (size_t v = 0; v < n; ++v) lat_cells *= 1.0 + (double)opt.cards[v];
    const double lattice_s = lat_cells * 12.0 / 3e12 + 0.05;
    lattice = lattice_s < count_s && (opt.exact_order == 0 || count_s > 2.0);
    if (opt.force_lattice >= 0) lattice = opt.force_lattice != 0;   // tests / A-B runs
  }
  if (lattice) {
    const auto t0 = clk::now();
    const int rc = obn_es_lattice(es, comm.rank, comm.world);
    if (rc == OBN_ERR_UNSUPPORTED) lattice = false;            // joint table too large: count per subset
    else check(rc);
    if (lattice && comm.world > 1) {
      const auto t1 = clk::now();
      void* p = nullptr;
      size_t per = 0;
      check(obn_es_f_buffer(es, &p, &per));
      comm.allgather_dev((uintptr_t)p, per * sizeof(double));
      if (stats) stats->gather_ms += ms(t1, clk::now());
    }
    if (stats) {
      stats->score_ms += ms(t0, clk::now());
      stats->lattice = lattice;
    }
  }
  std::vector<uint64_t> masks;
  bool dp_by_child = false;             // lattice mode on several GPUs: the DP columns are sharded by child
  for (size_t size = 0; size <= n; ++size) {                        // exhaustive_search.cpp:47
    const auto t0 = clk::now();
    masks.clear();
    size_t T = 1;
    const bool scored = size <= max_varset && size > 0;              // :48
    if (lattice || !scored) {             // the device DP walks its own level lists: nothing to enumerate
      for (size_t i = 1; i <= size; ++i) T = T * (n - size + i) / i;             // C(n, size), exact stepwise
    } else {
      uint64_t W = ((uint64_t)1 << size) - 1;                          // :51
      do masks.push_back(W); while (next_combmask(W, (int)n));         // :129 (Gosper order)
      T = masks.size();
    }
    const bool calc_ls = size <= max_varset && size > 0;             // :48
    const bool sharded = comm.world > 1 && calc_ls && !lattice &&
                         (double)T * (double)size * (double)n_rows >= opt.shard_min_bytes && T >= (size_t)comm.world;
    const size_t chunk = sharded ? (T + comm.world - 1) / comm.world : T;
    check(obn_es_level_begin(es, (lattice || !scored) ? nullptr : masks.data(), T, (int)size, sharded ? chunk * comm.world : T));
    const auto t1 = clk::now();
    if (calc_ls) {
      if (sharded) check(obn_es_level_scores(es, std::min(T, chunk * comm.rank), std::min(T, chunk * (comm.rank + 1))));
      else check(obn_es_level_scores(es, 0, T));
    }
    const auto t2 = clk::now();
    if (sharded) {
      void* p = nullptr;
      size_t nd = 0;
      check(obn_es_level_buffer(es, &p, &nd));
      comm.allgather_dev((uintptr_t)p, chunk * size * sizeof(double));
    }
    const auto t3 = clk::now();
    // :59-116.  Lattice mode on several GPUs: from the first level that is worth a collective on,
    // every rank runs the DP for its own children only and the sink choice merges through one
    // all-gather of (sum, child) per level (obn_es_level_dp_part).
    if (lattice && comm.world > 1 && (dp_by_child || T >= opt.dp_shard_min_subsets)) {
      dp_by_child = true;
      if (comm.native) {                // the library's communicator: the whole level stays on its stream
        check(obn_es_level_dp_sharded(es));
      } else {
        void* parts = nullptr;
        size_t per = 0;
        check(obn_es_level_dp_part(es, comm.rank, comm.world, &parts, &per));
        comm.allgather_dev((uintptr_t)parts, per);
        check(obn_es_level_dp_commit(es, comm.world));
      }
    } else {
      check(obn_es_level_dp(es));
    }
    const auto t4 = clk::now();
    if (stats) {
      if (size <= max_varset) {
        stats->tables += T;
        stats->scores += T * size;
      }
      stats->enumerate_ms += ms(t0, t1);
      stats->score_ms += ms(t1, t2);
      stats->gather_ms += ms(t2, t3);
      stats->dp_ms += ms(t3, t4);
    }
  }
  std::vector<uint64_t> pm(n, 0);
  double best = 0;
  check(obn_es_finish(es, pm.data(), &best));                         // :133-148
  if (dp_by_child) {                    // every rank reported the parent sets of its own children
    std::vector<uint64_t> all(n * (size_t)comm.world, 0);
    comm.allgather(pm.data(), all.data(), n * sizeof(uint64_t));
    for (size_t v = 0; v < n; ++v) {
      pm[v] = 0;
      for (int r = 0; r < comm.world; ++r) pm[v] |= all[(size_t)r * n + v];
    }
  }
  PDAG g(n);
  for (size_t v = 0; v < n; ++v)
    for (size_t j = 0; j < n; ++j)
      if ((pm[v] >> j) & 1u) g.add_edge(j, v);
  if (stats) stats->best_score = best;
  return g;
}

}  // namespace obn
