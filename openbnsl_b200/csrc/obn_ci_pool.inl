// obn_ci_pool.inl -- host-side planner of MODE_CI_POOL: CI tests of the same
// (x, y) pair whose conditioning sets are drawn from one neighbourhood share ONE joint table over
// (union of the conditioning sets, x, y) -- one pass over the rows and one ATOMS per row for all
// of them -- and every test's own (Z, x, y) table is summed out of it in the kernel's epilogue.
// This is the PC frontier's shape: pc.cpp:27-47 tests (x, y | Z) for every k-subset Z of
// adj(x) - {y}.  Pure host code (no CUDA calls): obn_ci_pool_plan exposes it to the CPU tests.
// Included by obn_capi.cu before obn_batch.inl.

namespace obn {

// Checks shared by every batch entry point; messages are the reference's where it has one
// (contingency_table.h:37-38, citest.h:47,59-61, local_score.h:27-30).
static int check_tuple(const uint32_t* cards, size_t n_vars, const uint32_t* v, size_t k, size_t t, bool ci,
                       bool one_child, uint32_t ra, uint32_t rb, uint64_t* cells_out) {
  if (k > (size_t)kMaxTupleVars)
    return fail(OBN_ERR_UNSUPPORTED, "tuple %zu has %zu variables (max %d)", t, k, kMaxTupleVars);
  uint64_t cl = 1;
  bool has_a = false, has_b = false;
  for (size_t i = 0; i < k; ++i) {
    if (v[i] >= n_vars) return fail(OBN_ERR_RANGE, "variable id %u out of range", v[i]);
    if (i && v[i - 1] > v[i]) return fail(OBN_ERR_INVALID, "var_ids must be sorted");
    cl *= cards[v[i]];
    if (cl >= (1ull << 31)) return fail(OBN_ERR_UNSUPPORTED, "tuple %zu: dense table has >= 2^31 cells", t);
    if (v[i] == ra) has_a = true;
    if (v[i] == rb) has_b = true;
  }
  if (ci) {
    if (ra == rb) return fail(OBN_ERR_INVALID, "x and y must differ");
    if (!has_a || !has_b) return fail(OBN_ERR_INVALID, "x or y not contained in contingency table");
  } else if (one_child && !has_a) {
    return fail(OBN_ERR_INVALID, "child_var not present in contingency table");
  }
  *cells_out = cl;
  return OBN_OK;
}

struct CiPoolPlan {
  std::vector<uint32_t> var_ids;    // CSR of the pools: ascending ids, x and y included
  std::vector<uint64_t> offsets;    // [pools + 1]
  std::vector<uint32_t> x, y;       // per pool, x < y (citest.h:48)
  std::vector<uint32_t> mem_off;    // [pools + 1]
  std::vector<uint32_t> members;    // per member: test id, keep mask (bit i = i-th conditioning variable of the pool)
  size_t pools() const { return x.size(); }
};

// Greedy, order-preserving: a test joins the open pool of its pair while the union of the
// conditioning sets keeps the joint table within max_cells, the pool within max_z conditioning
// variables (<= 30: the mask is a 32-bit word) and kPoolMaxMembers tests; otherwise the pool is
// closed and the test opens the next one.  A test whose own table exceeds max_cells, that is wider
// than max_z or that repeats a variable id is passed through with its own variable list (it may
// use any table tier; nothing is summed out).
static const uint32_t kPoolMaxMembers = kPoolMembers;   // count_kernel keeps (stat, dof) of a pool's members in shared memory

// Working storage of the planner, kept per host thread so that a PC run (one plan per round)
// does not allocate (about 100 ns per test with per-pool vectors and a std::map, about 40 like this).
struct CiPoolScratch {
  struct Pool {
    uint32_t x, y, nz, head, tail, count;
    uint64_t zcells;               // product of the cards of z
    bool single;                   // one test emitted with its own variable list
  };
  std::vector<Pool> pools;
  std::vector<uint32_t> z;         // `stride` slots per pool: its conditioning variables, ascending
  std::vector<uint32_t> next;      // per test: next member of its pool
  // pair -> its open pool: open addressing, only consulted when the pair changes
  std::vector<uint64_t> hkey;
  std::vector<uint32_t> hval;
  size_t hmask = 0;
  void hash_reset(size_t n) {
    size_t cap = 64;
    while (cap < 2 * n) cap <<= 1;
    hkey.assign(cap, ~0ull);
    hval.resize(cap);
    hmask = cap - 1;
  }
  size_t slot(uint64_t key) const {
    size_t h = (size_t)((key * 0x9E3779B97F4A7C15ull) >> 20) & hmask;
    while (hkey[h] != ~0ull && hkey[h] != key) h = (h + 1) & hmask;
    return h;
  }
};

static int plan_ci_pools(const uint32_t* cards, size_t n_vars, const uint32_t* var_ids, const uint64_t* offsets,
                         size_t n, const uint32_t* xs, const uint32_t* ys, uint32_t max_cells, uint32_t max_z,
                         CiPoolPlan& P) {
  if (max_z > 30) max_z = 30;                 // keep masks are 32-bit words
  if (max_z < 1) max_z = 1;
  static thread_local CiPoolScratch W;
  const size_t stride = max_z;
  W.pools.clear();
  W.pools.reserve(n);
  W.hash_reset(n);
  W.z.resize(n * stride);                     // capacity persists across calls; slots are written before they are read
  W.next.resize(n);
  auto new_pool = [&](uint32_t x, uint32_t y, uint32_t t, bool single) -> uint32_t {
    W.pools.push_back({x, y, 0u, t, t, 1u, 1ull, single});
    W.next[t] = 0xffffffffu;
    return (uint32_t)(W.pools.size() - 1);
  };
  uint64_t last_key = ~0ull;                  // the tests of a pair arrive back to back (pc.cpp:27-47)
  uint32_t last_pool = 0;
  uint32_t zt[kMaxTupleVars], un[2 * kMaxTupleVars];
  for (size_t t = 0; t < n; ++t) {
    const uint64_t o0 = offsets[t], o1 = offsets[t + 1];
    if (o1 < o0) return fail(OBN_ERR_INVALID, "offsets must be non-decreasing");
    const uint32_t* ids = var_ids + o0;
    const size_t k = (size_t)(o1 - o0);
    const uint32_t x = std::min(xs[t], ys[t]), y = std::max(xs[t], ys[t]);
    // one scan: the checks of check_tuple, the test's conditioning set, its table size
    bool ok = k <= (size_t)kMaxTupleVars && x != y, plain = true, hx = false, hy = false;
    uint64_t cl = 1;
    uint32_t nzt = 0;
    for (size_t i = 0; ok && i < k; ++i) {
      const uint32_t v = ids[i];
      if (v >= n_vars || (i && ids[i - 1] > v)) { ok = false; break; }
      if (i && ids[i - 1] == v) plain = false;
      cl *= cards[v];
      if (cl >= (1ull << 31)) { ok = false; break; }
      if (v == x) hx = true;
      else if (v == y) hy = true;
      else zt[nzt++] = v;
    }
    if (!ok || !hx || !hy) {                  // let check_tuple name the problem
      uint64_t dummy;
      if (int rc = check_tuple(cards, n_vars, ids, k, t, true, false, xs[t], ys[t], &dummy)) return rc;
      return fail(OBN_ERR_INVALID, "tuple %zu: inconsistent CI test", t);
    }
    if (!plain || cl > max_cells || nzt > max_z) {   // repeated id, too large or too wide to share: passed through
      new_pool(x, y, (uint32_t)t, true);
      continue;
    }
    const uint64_t key = (uint64_t)x * n_vars + y;
    uint32_t pi = 0xffffffffu;
    size_t hs = (size_t)-1;
    if (key == last_key) {
      pi = last_pool;
    } else {
      hs = W.slot(key);
      if (W.hkey[hs] == key) pi = W.hval[hs];
    }
    if (pi != 0xffffffffu && W.pools[pi].count < kPoolMaxMembers) {
      // sorted union of the pool's and the test's conditioning sets; the table grows by the new cards
      CiPoolScratch::Pool& p = W.pools[pi];
      uint32_t* pz = W.z.data() + (size_t)pi * stride;
      uint32_t a = 0, b = 0, nu = 0;
      uint64_t ucells = p.zcells;
      while (a < p.nz || b < nzt) {
        if (b == nzt || (a < p.nz && pz[a] < zt[b])) un[nu++] = pz[a++];
        else if (a == p.nz || zt[b] < pz[a]) { ucells *= cards[zt[b]]; un[nu++] = zt[b++]; }
        else { un[nu++] = pz[a++]; ++b; }
      }
      if (nu <= max_z && ucells * cards[x] * cards[y] <= max_cells) {
        if (nu != p.nz) {
          std::copy(un, un + nu, pz);
          p.nz = nu;
          p.zcells = ucells;
        }
        W.next[p.tail] = (uint32_t)t;
        W.next[t] = 0xffffffffu;
        p.tail = (uint32_t)t;
        ++p.count;
        last_key = key;
        last_pool = pi;
        continue;
      }
    }
    pi = new_pool(x, y, (uint32_t)t, false);
    std::copy(zt, zt + nzt, W.z.data() + (size_t)pi * stride);
    W.pools[pi].nz = nzt;
    W.pools[pi].zcells = cl / ((uint64_t)cards[x] * cards[y]);
    if (hs == (size_t)-1) hs = W.slot(key);
    W.hkey[hs] = key;
    W.hval[hs] = pi;
    last_key = key;
    last_pool = pi;
  }
  // ---- emit ------------------------------------------------------------------------------
  const size_t n_pools = W.pools.size();
  P.var_ids.clear(); P.offsets.clear(); P.mem_off.clear();
  P.x.resize(n_pools);
  P.y.resize(n_pools);
  P.members.resize(2 * n);
  P.var_ids.reserve(offsets[n]);
  P.offsets.reserve(n_pools + 1);
  P.mem_off.reserve(n_pools + 1);
  P.offsets.push_back(0);
  P.mem_off.push_back(0);
  size_t nm = 0;
  for (size_t q = 0; q < n_pools; ++q) {
    const CiPoolScratch::Pool& p = W.pools[q];
    P.x[q] = p.x;
    P.y[q] = p.y;
    if (p.single) {                 // the test's own variable list, nothing summed out
      const uint32_t t = p.head;
      P.var_ids.insert(P.var_ids.end(), var_ids + offsets[t], var_ids + offsets[t + 1]);
      P.members[2 * nm] = t;
      P.members[2 * nm + 1] = 0xffffffffu;
      ++nm;
    } else {
      const uint32_t* pz = W.z.data() + q * stride;
      bool dx = false, dy = false;  // merge x and y into the ascending list
      for (uint32_t i = 0; i < p.nz; ++i) {
        if (!dx && p.x < pz[i]) { P.var_ids.push_back(p.x); dx = true; }
        if (!dy && p.y < pz[i]) { P.var_ids.push_back(p.y); dy = true; }
        P.var_ids.push_back(pz[i]);
      }
      if (!dx) P.var_ids.push_back(p.x);
      if (!dy) P.var_ids.push_back(p.y);
      for (uint32_t t = p.head; t != 0xffffffffu; t = W.next[t]) {
        uint32_t mask = 0, j = 0;
        for (uint64_t i = offsets[t]; i < offsets[t + 1]; ++i) {
          const uint32_t v = var_ids[i];
          if (v == p.x || v == p.y) continue;
          while (pz[j] != v) ++j;   // both ascending, the test's set is contained in the pool's
          mask |= 1u << j;
        }
        P.members[2 * nm] = t;
        P.members[2 * nm + 1] = mask;
        ++nm;
      }
    }
    P.offsets.push_back(P.var_ids.size());
    P.mem_off.push_back((uint32_t)nm);
  }
  return OBN_OK;
}

}  // namespace obn

extern "C" {

// Diagnostics, host only (include/openbnsl_b200.h): the plan obn_ci_tests would use.
int obn_ci_pool_plan(const uint32_t* cards, size_t n_vars, const uint32_t* var_ids, const uint64_t* offsets,
                     size_t n_tests, const uint32_t* x, const uint32_t* y, uint32_t max_cells, uint32_t max_z,
                     uint32_t* pool_var_ids, uint64_t* pool_offsets, uint32_t* pool_of_test,
                     uint32_t* keep_of_test, size_t* n_pools) {
  if (!cards || !offsets || !n_pools || (n_tests && (!x || !y))) return fail(OBN_ERR_INVALID, "NULL argument");
  static thread_local CiPoolPlan P;
  if (int rc = plan_ci_pools(cards, n_vars, var_ids, offsets, n_tests, x, y, max_cells, max_z, P)) return rc;
  *n_pools = P.pools();
  if (pool_offsets) memcpy(pool_offsets, P.offsets.data(), P.offsets.size() * sizeof(uint64_t));
  if (pool_var_ids && !P.var_ids.empty()) memcpy(pool_var_ids, P.var_ids.data(), P.var_ids.size() * sizeof(uint32_t));
  for (size_t q = 0; q < P.pools(); ++q)
    for (uint32_t m = P.mem_off[q]; m < P.mem_off[q + 1]; ++m) {
      if (pool_of_test) pool_of_test[P.members[2 * m]] = (uint32_t)q;
      if (keep_of_test) keep_of_test[P.members[2 * m]] = P.members[2 * m + 1];
    }
  return OBN_OK;
}

// Diagnostics, host only: the kernel's marginalisation passes (obn_kernels.cuh: MargPasses +
// marg_cell, the same functions the device runs) applied to a joint table on the host.
int obn_ci_pool_marginalise(const int32_t* joint, uint32_t cells, const uint32_t* cards_kernel_order, int n_z,
                            uint32_t keep, int32_t* out, uint32_t* out_cells) {
  if (!joint || !cards_kernel_order || !out || !out_cells || n_z < 0 || n_z > 30)
    return fail(OBN_ERR_INVALID, "invalid argument");
  std::vector<int32_t> a(joint, joint + cells), b;
  MargPasses it(cells);
  uint32_t w, c, n_out;
  while (it.next(cards_kernel_order, n_z, keep, w, c, n_out)) {
    b.resize(n_out);
    for (uint32_t j = 0; j < n_out; ++j) b[j] = marg_cell(a.data(), j, w, c);
    a.swap(b);
  }
  *out_cells = it.cur;
  memcpy(out, a.data(), (size_t)it.cur * sizeof(int32_t));
  return OBN_OK;
}

}  // extern "C"
