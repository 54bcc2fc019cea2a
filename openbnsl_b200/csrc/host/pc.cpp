// pc.cpp -- PC / PC-stable skeleton + collider orientation with the reference's
// semantics (src/structure_learning/pc.cpp:16-86), restructured for batching:
//
//  * stable mode: all CI questions of one depth depend only on the snapshot
//    graph, so they are independent.  The driver evaluates them in ROUNDS: each
//    round takes the next `quota` candidate sets of every still-undecided
//    edge (in the reference's enumeration order: (x,y) combinations first,
//    then (y,x); pc.cpp:31-45), ships them as ONE batch to the evaluator, and
//    resolves every edge at its FIRST passing candidate -- exactly the set the
//    serial reference would have stopped at.  The quota grows geometrically,
//    which keeps batches large without evaluating the whole frontier of edges
//    that separate early.
//  * non-stable mode mutates the graph between pairs (pc.cpp:25,40), so pairs
//    are sequential; each pair's candidates form one batch.
//  * (v,u|Z) is the same test as (u,v|Z) (citest.h:48 swaps), so candidates of
//    the second direction that were already asked in the first are skipped.
//  * multi-GPU: each round's batch is sharded over ranks; one all-reduce(min)
//    of "first passing candidate index per edge" per round.
#include <algorithm>
#include <chrono>
#include <climits>

#include "structure_learning.hpp"

namespace obn {

namespace {

// lexicographic k-combinations of m items (include/utils/gen_comb.h:5-38)
struct CombIter {
  int m = 0, k = 0;
  std::vector<int> idx;
  bool valid = false;
  void start(int m_, int k_) {
    m = m_;
    k = k_;
    idx.resize(k);
    for (int i = 0; i < k; ++i) idx[i] = i;
    valid = k <= m;
  }
  void next() {
    int i = k - 1;
    while (i >= 0 && idx[i] == m - k + i) --i;
    if (i < 0) {
      valid = false;
      return;
    }
    ++idx[i];
    for (int j = i + 1; j < k; ++j) idx[j] = idx[j - 1] + 1;
  }
};

struct PairCursor {
  uint32_t x = 0, y = 0;
  std::vector<uint32_t> neigh[2];   // adj(x)\y, adj(y)\x  (ascending)
  std::vector<char> in_first;       // membership of adj(x)\y, for de-duplication
  CombIter it;
  int dir = 0;
  bool done = false;                // no candidates left
  bool removed = false;

  void begin(size_t k, size_t n_vars) {
    in_first.assign(n_vars, 0);
    for (auto v : neigh[0]) in_first[v] = 1;
    dir = 0;
    it.start((int)neigh[0].size(), (int)k);
    settle(k);
  }
  // move to a valid, non-duplicate candidate (or mark done)
  void settle(size_t k) {
    for (;;) {
      if (!it.valid) {
        if (dir == 1) {
          done = true;
          return;
        }
        dir = 1;
        it.start((int)neigh[1].size(), (int)k);
        continue;
      }
      if (dir == 1) {
        bool dup = true;               // Z subset of adj(x)\y  => asked already in direction 0
        for (int i = 0; i < it.k; ++i)
          if (!in_first[neigh[1][it.idx[i]]]) {
            dup = false;
            break;
          }
        if (it.k == 0) dup = true;     // Z = {} is always shared
        if (dup && (int)neigh[0].size() >= it.k) {
          it.next();
          continue;
        }
      }
      return;
    }
  }
  void current(std::vector<uint32_t>& z) const {
    z.clear();
    for (int i = 0; i < it.k; ++i) z.push_back(neigh[dir][it.idx[i]]);
  }
  void advance(size_t k) {
    it.next();
    settle(k);
  }
};

using Sepsets = std::vector<std::vector<uint32_t>>;   // index x*n+y, valid iff has[x*n+y]

void orient_colliders(AdjMatGraph& g, const Sepsets& sep) {   // pc.cpp:54-74
  const size_t n = g.num_vars();
  std::vector<uint32_t> neigh;
  for (size_t y = 0; y < n; ++y) {
    bool changed = true;
    while (changed) {
      changed = false;
      g.undirected_neighbors_without(y, (size_t)-1, neigh);
      for (size_t i = 0; i + 1 < neigh.size(); ++i)
        for (size_t j = i + 1; j < neigh.size(); ++j) {
          const size_t x = neigh[i], z = neigh[j];
          if (g.is_adjacent(x, z)) continue;
          const auto& s = sep[x * n + z];
          if (std::find(s.begin(), s.end(), (uint32_t)y) == s.end()) {
            g.orient_edge(x, y);
            g.orient_edge(z, y);
            changed = true;
          }
        }
    }
  }
}

// decide(batch, zs_flat/off are inside batch) -> first[pair_slot] = index of the
// first independent candidate of that pair within the batch, INT_MAX if none.
using DecideFn = std::function<void(const TupleBatch&, const std::vector<uint32_t>& pair_of_test,
                                    size_t n_slots, std::vector<int32_t>& first)>;

void append_test(TupleBatch& b, uint32_t u, uint32_t v, const std::vector<uint32_t>& z,
                 std::vector<uint32_t>& tmp) {
  tmp.assign(z.begin(), z.end());
  tmp.push_back(u);
  tmp.push_back(v);
  std::sort(tmp.begin(), tmp.end());                 // pc.cpp:34-37
  b.var_ids.insert(b.var_ids.end(), tmp.begin(), tmp.end());
  b.offsets.push_back(b.var_ids.size());
  b.x.push_back(u);
  b.y.push_back(v);
}

// The same decision in two halves (Evaluator::pc_round_submit / pc_round_collect): submit enqueues a round and
// returns, collect waits for the OLDEST round in flight.  Two rounds may be in flight; chain = the evaluator
// skips the tests of edges the round submitted before this one decided.
struct AsyncDecide {
  std::function<void(const TupleBatch&, const std::vector<uint32_t>& pair_of_test, size_t n_slots, bool chain)> submit;
  std::function<void(size_t n_slots, std::vector<int32_t>& first)> collect;
};

PDAG run_pc(size_t n, size_t max_cond_vars, bool stable, size_t target_batch, const DecideFn& decide,
            PcStats* stats, const AsyncDecide* async = nullptr) {
  AdjMatGraph g(n);
  Sepsets sep(n * n);
  if (stats) {
    stats->recorded.clear();
    stats->tests_per_depth.assign(max_cond_vars + 1, 0);
    stats->batches_per_depth.assign(max_cond_vars + 1, 0);
  }
  using clk = std::chrono::steady_clock;
  auto ms_since = [](clk::time_point t0) { return std::chrono::duration<double, std::milli>(clk::now() - t0).count(); };
  if (stats) stats->enumerate_ms = stats->evaluate_ms = stats->resolve_ms = stats->exposed_ms = stats->collect_wait_ms = 0;
  clk::time_point t_phase = clk::now();
  clk::time_point t_empty = clk::now();       // since when no round is in flight
  auto lap = [&](double PcStats::*slot) {
    if (stats) stats->*slot += ms_since(t_phase);
    t_phase = clk::now();
  };
  struct FastCursor {
    uint32_t x = 0, y = 0, m[2] = {0, 0}, skip[2] = {0, 0};
    uint8_t dir = 0;
    bool valid = false, done = false, removed = false;
  };
  // working storage kept per host thread: a run allocates nothing after the first one of its size
  // (a C2 run is 0.8 ms in all; growing these vectors from empty was a measurable part of it)
  struct Scratch {
    std::vector<FastCursor> cursors;
    std::vector<uint32_t> idx_arena, pair_of_test;
    std::vector<std::vector<uint32_t>> adj;
    TupleBatch batch, ring[3];
    std::vector<uint32_t> ring_pairs[3];
  };
  static thread_local Scratch W;
  std::vector<FastCursor>& cursors = W.cursors;
  std::vector<uint32_t>& idx_arena = W.idx_arena;
  std::vector<std::vector<uint32_t>>& adj = W.adj;
  adj.resize(n);
  TupleBatch& batch = W.batch;
  std::vector<uint32_t>& pair_of_test = W.pair_of_test;
  batch.clear();
  std::vector<uint32_t> z, tmp;
  std::vector<std::vector<uint32_t>> z_of_test;
  std::vector<int32_t> first;
  // accounts a round; skip(t): test t was left out by the evaluator (chained rounds)
  auto note_batch = [&](size_t k, const TupleBatch& b, const std::function<bool(size_t)>* skip) {
    if (!stats) return;
    size_t asked = b.size();
    if (skip) {
      asked = 0;
      for (size_t t = 0; t < b.size(); ++t) asked += !(*skip)(t);
    }
    stats->tests_per_depth[k] += asked;
    stats->batches_per_depth[k] += 1;
    if (!stats->record) return;
    TupleBatch& r = stats->recorded;
    for (size_t t = 0; t < b.size(); ++t) {
      if (skip && (*skip)(t)) continue;
      r.var_ids.insert(r.var_ids.end(), b.var_ids.begin() + b.offsets[t], b.var_ids.begin() + b.offsets[t + 1]);
      r.offsets.push_back(r.var_ids.size());
      r.x.push_back(b.x[t]);
      r.y.push_back(b.y[t]);
    }
  };

  for (size_t k = 0; k <= max_cond_vars; ++k) {              // pc.cpp:22 (no early exit)
    if (stable) {
      // All pairs see the graph as it was when the depth started (pc.cpp:24-25): snapshot the
      // adjacency lists once.  Cursors are allocation-free: candidate j of direction d of pair
      // (x,y) is the j-th lexicographic k-combination of adj(u)\{v} (gen_comb.h:5-38), kept as
      // k indices into the snapshot list of u (skipping v's position).
      const AdjMatGraph snap = g;
      for (size_t v = 0; v < n; ++v) snap.undirected_neighbors_without(v, (size_t)-1, adj[v]);
      cursors.clear();
      for (size_t x = 0; x < n; ++x)
        for (size_t y = x + 1; y < n; ++y) {
          if (!snap.has_undirected_edge_unchecked(x, y)) continue;
          FastCursor c;
          c.x = (uint32_t)x;
          c.y = (uint32_t)y;
          c.m[0] = (uint32_t)adj[x].size() - 1;
          c.m[1] = (uint32_t)adj[y].size() - 1;
          if (c.m[0] < k && c.m[1] < k) continue;
          if (k) {                          // position of the partner in each list (nothing is indexed at depth 0)
            c.skip[0] = (uint32_t)(std::lower_bound(adj[x].begin(), adj[x].end(), (uint32_t)y) - adj[x].begin());
            c.skip[1] = (uint32_t)(std::lower_bound(adj[y].begin(), adj[y].end(), (uint32_t)x) - adj[y].begin());
          }
          cursors.push_back(c);
        }
      idx_arena.assign(cursors.size() * std::max<size_t>(k, 1), 0);
      auto item = [&](const FastCursor& c, int dir, uint32_t i) -> uint32_t {   // i-th element of adj(u)\{v}
        const std::vector<uint32_t>& a = adj[dir ? c.y : c.x];
        return a[i + (i >= c.skip[dir] ? 1u : 0u)];
      };
      auto start_dir = [&](FastCursor& c, uint32_t* idx, int dir) {
        c.dir = (uint8_t)dir;
        c.valid = k <= c.m[dir];
        for (size_t i = 0; i < k; ++i) idx[i] = (uint32_t)i;
      };
      auto next_comb = [&](FastCursor& c, uint32_t* idx) {
        const uint32_t m = c.m[c.dir];
        int i = (int)k - 1;
        while (i >= 0 && idx[i] == m - (uint32_t)k + (uint32_t)i) --i;
        if (i < 0) { c.valid = false; return; }
        ++idx[i];
        for (size_t j = (size_t)i + 1; j < k; ++j) idx[j] = idx[j - 1] + 1;
      };
      // move to a valid candidate that was not already asked in direction 0 (citest.h:48 makes
      // (y,x|Z) the same test as (x,y|Z) whenever Z is also a subset of adj(x)\{y})
      auto settle = [&](FastCursor& c, uint32_t* idx) {
        for (;;) {
          if (!c.valid) {
            if (c.dir == 1) { c.done = true; return; }
            start_dir(c, idx, 1);
            continue;
          }
          if (c.dir == 1 && c.m[0] >= k) {
            bool dup = true;
            for (size_t i = 0; i < k && dup; ++i) dup = snap.has_undirected_edge_unchecked(c.x, item(c, 1, idx[i]));
            if (dup) { next_comb(c, idx); continue; }
          }
          return;
        }
      };
      for (size_t pi = 0; pi < cursors.size(); ++pi) {
        uint32_t* idx = idx_arena.data() + pi * std::max<size_t>(k, 1);
        start_dir(cursors[pi], idx, 0);
        settle(cursors[pi], idx);
      }
      auto count_live = [&]() {
        size_t live = 0;
        for (const auto& c : cursors) live += (!c.done && !c.removed);
        return live;
      };
      // the next `quota` candidates of every undecided edge, in the reference's enumeration order.  A round may be
      // cut into windows of edges: starts at cursor pi_begin, stops in front of the first edge that would begin
      // beyond max_tests tests, and returns where it stopped (cursors.size(): the sweep is complete).
      auto build_round = [&](TupleBatch& out, std::vector<uint32_t>& pair_out, size_t quota, size_t pi_begin = 0,
                             size_t max_tests = (size_t)-1) -> size_t {
        out.clear();
        pair_out.clear();
        // raw writes into storage sized once per window (a push_back per field and test was a third of the loop)
        const size_t w = k + 2;
        size_t cap = 0, n_out = 0;
        uint32_t *vp = nullptr, *xp = nullptr, *yp = nullptr, *pp = nullptr;
        uint64_t* op = nullptr;
        auto grow = [&](size_t need) {
          cap = std::max<size_t>(need, cap ? 2 * cap : 4096);
          out.var_ids.resize(cap * w);
          out.offsets.resize(cap + 1);
          out.x.resize(cap);
          out.y.resize(cap);
          pair_out.resize(cap);
          vp = out.var_ids.data(); op = out.offsets.data(); xp = out.x.data(); yp = out.y.data(); pp = pair_out.data();
        };
        auto finish = [&]() {
          out.var_ids.resize(n_out * w);
          out.offsets.resize(n_out + 1);
          out.x.resize(n_out);
          out.y.resize(n_out);
          pair_out.resize(n_out);
        };
        grow(4096);
        op[0] = 0;
        for (size_t pi = pi_begin; pi < cursors.size(); ++pi) {
          FastCursor& c = cursors[pi];
          if (c.done || c.removed) continue;
          if (n_out >= max_tests) {
            finish();
            return pi;
          }
          uint32_t* idx = idx_arena.data() + pi * std::max<size_t>(k, 1);
          if (n_out + quota > cap) grow(n_out + quota);
          for (size_t q = 0; q < quota && !c.done; ++q) {
            // tuple = sort(Z + {x,y}) (pc.cpp:34-37): Z is ascending already, merge x < y into it
            const uint32_t* a = adj[c.dir ? c.y : c.x].data();
            const uint32_t skip = c.skip[c.dir];
            uint32_t* wr = vp + n_out * w;
            uint32_t pend[2] = {c.x, c.y};
            int np = 0;
            for (size_t i = 0; i < k; ++i) {
              const uint32_t z = a[idx[i] + (idx[i] >= skip ? 1u : 0u)];
              while (np < 2 && pend[np] < z) *wr++ = pend[np++];
              *wr++ = z;
            }
            while (np < 2) *wr++ = pend[np++];
            xp[n_out] = c.dir ? c.y : c.x;
            yp[n_out] = c.dir ? c.x : c.y;
            pp[n_out] = (uint32_t)pi;
            ++n_out;
            op[n_out] = (uint64_t)n_out * w;
            next_comb(c, idx);
            settle(c, idx);
          }
        }
        finish();
        return cursors.size();
      };
      // every edge is resolved at its FIRST passing candidate (pc.cpp:40-42)
      auto resolve_round = [&](const TupleBatch& b) {
        for (size_t pi = 0; pi < cursors.size(); ++pi) {
          if (first[pi] == INT_MAX) continue;
          FastCursor& c = cursors[pi];
          if (c.removed) continue;
          c.removed = true;
          g.remove_undirected_edge(c.x, c.y);
          std::vector<uint32_t>& sz = sep[(size_t)c.x * n + c.y];
          sz.clear();
          const size_t t = (size_t)first[pi];
          for (uint64_t i = b.offsets[t]; i < b.offsets[t + 1]; ++i)
            if (b.var_ids[i] != c.x && b.var_ids[i] != c.y) sz.push_back(b.var_ids[i]);
          sep[(size_t)c.y * n + c.x] = sz;
        }
      };
      size_t quota = 1;
      size_t live = count_live();
      if (!async) {
        if (live) {
          quota = std::max<size_t>(quota, target_batch / live);
          build_round(batch, pair_of_test, quota);
        }
        while (live && batch.size()) {
          lap(&PcStats::enumerate_ms);
          decide(batch, pair_of_test, cursors.size(), first);
          quota *= 2;
          lap(&PcStats::evaluate_ms);
          note_batch(k, batch, nullptr);
          resolve_round(batch);
          live = count_live();
          if (live) {
            quota = std::max<size_t>(quota, target_batch / live);
            build_round(batch, pair_of_test, quota);
          }
          lap(&PcStats::resolve_ms);
        }
      } else if (live) {
        // Pipelined rounds (asynchronous evaluators).  The candidates of a depth depend only on the snapshot,
        // so the next round can be enumerated -- and handed to the evaluator -- before this one has answered:
        // it is built as if this round removed no edge, and the evaluator itself drops the tests of the edges
        // this round does remove (chain), which is what the serial reference does by stopping at an edge's first
        // independent candidate (pc.cpp:40-42).  The device always has the next round queued behind the one it
        // is working on; the host enumerates, packs and resolves in its shadow.  Batch boundaries do not change
        // what an edge resolves at.  Ring of three rounds: being resolved / in flight behind it / being built.
        // Sweeps and windows: a sweep hands every undecided edge its next `quota` candidates (the quota doubles
        // from sweep to sweep, as in the serial loop).  The FIRST batch of a depth is only a window of the first
        // sweep (a quarter of target_batch): the device starts after a short build, and the rest of the sweep is
        // enumerated and packed in its shadow.  Edges are independent within a depth, so cutting a sweep anywhere
        // changes nothing an edge resolves at.
        size_t head = 0, n_flight = 0, sweep_pos = cursors.size();
        // an exception (a CUDA error, a Python evaluator that raises) must not leave rounds in flight in the
        // evaluator: the next call on the same ctx would be refused
        struct Drain {
          const AsyncDecide* a;
          size_t* n;
          size_t slots;
          ~Drain() {
            std::vector<int32_t> sink;
            while (*n) {
              --*n;
              try { a->collect(slots, sink); } catch (...) {}
            }
          }
        } drain{async, &n_flight, cursors.size()};
        bool first_sweep = true, first_window = true;
        auto next_window = [&](TupleBatch& out, std::vector<uint32_t>& pairs) -> bool {
          for (int attempt = 0; attempt < 2; ++attempt) {
            if (sweep_pos >= cursors.size()) {            // start the next sweep
              const size_t l2 = count_live();
              if (!l2) return false;
              if (!first_sweep) quota *= 2;
              first_sweep = false;
              quota = std::max<size_t>(quota, target_batch / l2);
              sweep_pos = 0;
            }
            const size_t budget = first_window ? std::max<size_t>(1, target_batch / 4) : (size_t)-1;
            sweep_pos = build_round(out, pairs, quota, sweep_pos, budget);
            if (out.size()) {
              first_window = false;
              return true;
            }
          }
          return false;
        };
        bool more = next_window(W.ring[0], W.ring_pairs[0]);
        lap(&PcStats::enumerate_ms);
        if (more) {
          async->submit(W.ring[0], W.ring_pairs[0], cursors.size(), false);
          n_flight = 1;
          if (stats) stats->exposed_ms += ms_since(t_empty);
        }
        lap(&PcStats::evaluate_ms);
        while (n_flight) {
          if (more && n_flight < 2) {
            const size_t nxt = (head + n_flight) % 3;
            more = next_window(W.ring[nxt], W.ring_pairs[nxt]);
            lap(&PcStats::enumerate_ms);
            if (more) {
              async->submit(W.ring[nxt], W.ring_pairs[nxt], cursors.size(), true);
              ++n_flight;
            }
            lap(&PcStats::evaluate_ms);
          }
          const TupleBatch& b = W.ring[head];
          const std::vector<uint32_t>& pairs = W.ring_pairs[head];
          const clk::time_point t_wait = clk::now();
          async->collect(cursors.size(), first);
          if (stats) stats->collect_wait_ms += ms_since(t_wait);
          if (n_flight == 1) t_empty = clk::now();
          lap(&PcStats::evaluate_ms);
          // an edge that is marked removed by now was decided by the round before this one: its tests were skipped
          const std::function<bool(size_t)> skipped = [&](size_t t) { return cursors[pairs[t]].removed; };
          note_batch(k, b, &skipped);
          resolve_round(b);
          head = (head + 1) % 3;
          --n_flight;
          lap(&PcStats::resolve_ms);
        }
      }
    } else {
      for (size_t x = 0; x < n; ++x)
        for (size_t y = x + 1; y < n; ++y) {
          if (!g.has_undirected_edge(x, y)) continue;
          PairCursor p;
          p.x = (uint32_t)x;
          p.y = (uint32_t)y;
          g.undirected_neighbors_without(x, y, p.neigh[0]);   // live graph
          g.undirected_neighbors_without(y, x, p.neigh[1]);
          p.begin(k, n);
          if (p.done) continue;
          batch.clear();
          pair_of_test.clear();
          z_of_test.clear();
          while (!p.done) {
            p.current(z);
            const uint32_t u = p.dir ? p.y : p.x, v = p.dir ? p.x : p.y;
            append_test(batch, u, v, z, tmp);
            pair_of_test.push_back(0);
            z_of_test.push_back(z);
            p.advance(k);
          }
          decide(batch, pair_of_test, 1, first);
          note_batch(k, batch, nullptr);
          if (first[0] != INT_MAX) {
            g.remove_undirected_edge(x, y);
            sep[x * n + y] = z_of_test[(size_t)first[0]];
            sep[y * n + x] = sep[x * n + y];
          }
        }
    }
  }
  lap(&PcStats::enumerate_ms);
  orient_colliders(g, sep);
  lap(&PcStats::resolve_ms);
  // g.apply_meeks_rules(false) is disabled in the reference (pc.cpp:84)
  return g.to_pdag();
}

}  // namespace

PDAG pc_skeleton_and_colliders(size_t n_vars, Evaluator& ev, const Comm& comm, const PcOptions& opt,
                               PcStats* stats) {
  std::vector<double> p;
  TupleBatch mine;                   // this rank's share of a sharded round (host-side comm only)
  auto sharded = [&](const TupleBatch& b) { return comm.world > 1 && b.var_ids.size() >= opt.shard_min_slots; };
  DecideFn decide = [&](const TupleBatch& b, const std::vector<uint32_t>& pair_of_test, size_t n_slots,
                        std::vector<int32_t>& first) {
    const bool shard = sharded(b);
    // The product path (obn_pc_round): the decision is taken on the device, a sharded round is merged
    // by the library's own all-reduce(min) on the device buffer, only first[] comes back.  A sharded
    // round over a host-side communicator (test seams) falls through to the generic path below.
    if (!(shard && !comm.native) && ev.pc_round(b, pair_of_test, n_slots, opt.kind, opt.level, shard, first)) return;
    first.assign(n_slots, INT_MAX);
    if (!shard) {
      p.resize(b.size());
      ev.ci_pvalues(b, 0, b.size(), opt.kind, p.data());
      for (size_t t = 0; t < b.size(); ++t)
        if (p[t] >= opt.level) {                              // citest.h:147
          int32_t& f = first[pair_of_test[t]];
          if ((int32_t)t < f) f = (int32_t)t;
        }
      return;
    }
    // Interleaved shares (test t belongs to rank t mod world): neighbouring tests have similar
    // table sizes, so every rank gets the same mix; contiguous ranges would not.
    mine.clear();
    for (size_t t = (size_t)comm.rank; t < b.size(); t += (size_t)comm.world) {
      mine.var_ids.insert(mine.var_ids.end(), b.var_ids.begin() + b.offsets[t], b.var_ids.begin() + b.offsets[t + 1]);
      mine.offsets.push_back(mine.var_ids.size());
      mine.x.push_back(b.x[t]);
      mine.y.push_back(b.y[t]);
    }
    p.resize(mine.size());
    ev.ci_pvalues(mine, 0, mine.size(), opt.kind, p.data());
    for (size_t i = 0; i < mine.size(); ++i)
      if (p[i] >= opt.level) {
        const size_t t = (size_t)comm.rank + i * (size_t)comm.world;
        int32_t& f = first[pair_of_test[t]];
        if ((int32_t)t < f) f = (int32_t)t;
      }
    comm.allreduce_min_i32(first.data(), first.size());
  };
  // The same decision in two halves, so that the device is kept one round ahead of the host.  A sharded
  // round works too when the communicator is the library's (the collective is enqueued behind the kernels at
  // submit); a host-side communicator would need its own split: not pipelined.
  AsyncDecide async;
  async.submit = [&](const TupleBatch& b, const std::vector<uint32_t>& pair_of_test, size_t n_slots, bool chain) {
    if (!ev.pc_round_submit(b, pair_of_test, n_slots, opt.kind, opt.level, sharded(b), chain))
      throw std::runtime_error("pc: the evaluator does not take asynchronous rounds (overlap=False runs them one by one)");
  };
  async.collect = [&](size_t n_slots, std::vector<int32_t>& first) {
    ev.pc_round_collect(first);
    first.resize(n_slots, INT_MAX);
  };
  const bool overlap = opt.overlap && opt.stable && (comm.world == 1 || comm.native);
  return run_pc(n_vars, opt.max_cond_vars, opt.stable, opt.target_batch, decide, stats, overlap ? &async : nullptr);
}

PDAG pc_oracle(size_t n_vars, const PDAG& truth, size_t max_cond_vars, bool stable) {
  DecideFn decide = [&](const TupleBatch& b, const std::vector<uint32_t>& pair_of_test, size_t n_slots,
                        std::vector<int32_t>& first) {
    first.assign(n_slots, INT_MAX);
    std::vector<size_t> z;
    for (size_t t = 0; t < b.size(); ++t) {
      if (first[pair_of_test[t]] != INT_MAX) continue;
      z.clear();
      for (uint64_t i = b.offsets[t]; i < b.offsets[t + 1]; ++i)
        if (b.var_ids[i] != b.x[t] && b.var_ids[i] != b.y[t]) z.push_back(b.var_ids[i]);
      size_t xx = b.x[t], yy = b.y[t];
      if (xx >= yy) std::swap(xx, yy);                       // citest.h:48
      if (is_d_separated(truth, xx, yy, z)) first[pair_of_test[t]] = (int32_t)t;
    }
  };
  return run_pc(n_vars, max_cond_vars, stable, 1 << 14, decide, nullptr);
}

}  // namespace obn
