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
// conditioning sets keeps the joint table within max_cells, the pool within max_z conditioning variables (<= 30: the
// mask is a 32-bit word) and kPoolMaxMembers tests;
// otherwise the pool is closed and the test opens the next one.  A test whose own table exceeds
// max_cells is a pool of its own (it may use any table tier; nothing is summed out).  Tests with
// a repeated variable id are left alone as well (the kernel counts such a table as it is given).
static const uint32_t kPoolMaxMembers = kPoolMembers;   // count_kernel keeps (stat, dof) of a pool's members in shared memory

static int plan_ci_pools(const uint32_t* cards, size_t n_vars, const uint32_t* var_ids, const uint64_t* offsets,
                         size_t n, const uint32_t* xs, const uint32_t* ys, uint32_t max_cells, uint32_t max_z,
                         CiPoolPlan& P) {
  if (max_z > 30) max_z = 30;                 // keep masks are 32-bit words
  struct Pool {
    uint32_t x, y;
    std::vector<uint32_t> z;        // union of the members' conditioning sets, ascending
    std::vector<uint32_t> tests;
    bool verbatim;                  // a single test passed through with its own variable list
  };
  std::vector<Pool> pools;
  pools.reserve(n);
  std::map<uint64_t, size_t> open;  // pair -> index of its open pool
  uint64_t last_key = ~0ull;        // the tests of a pair arrive back to back (pc.cpp:27-47): skip the map
  size_t last_pool = 0;
  std::vector<uint32_t> z, u;
  for (size_t t = 0; t < n; ++t) {
    const uint64_t o0 = offsets[t], o1 = offsets[t + 1];
    if (o1 < o0) return fail(OBN_ERR_INVALID, "offsets must be non-decreasing");
    uint64_t cl = 0;
    if (int rc = check_tuple(cards, n_vars, var_ids + o0, (size_t)(o1 - o0), t, true, false, xs[t], ys[t], &cl))
      return rc;
    const uint32_t x = std::min(xs[t], ys[t]), y = std::max(xs[t], ys[t]);
    z.clear();
    bool plain = true;              // strictly ascending, x and y once each
    for (uint64_t i = o0; i < o1; ++i) {
      if (i > o0 && var_ids[i] == var_ids[i - 1]) plain = false;
      if (var_ids[i] != x && var_ids[i] != y) z.push_back(var_ids[i]);
    }
    if (!plain) {
      pools.push_back({x, y, std::vector<uint32_t>(var_ids + o0, var_ids + o1), {(uint32_t)t}, true});
      continue;
    }
    if (cl > max_cells) {
      pools.push_back({x, y, z, {(uint32_t)t}, false});
      continue;
    }
    const uint64_t key = (uint64_t)x * n_vars + y;
    size_t pi = (size_t)-1;
    if (key == last_key) {
      pi = last_pool;
    } else {
      auto it = open.find(key);
      if (it != open.end()) pi = it->second;
    }
    if (pi != (size_t)-1 && pools[pi].tests.size() < kPoolMaxMembers) {
      Pool& p = pools[pi];
      u.clear();
      std::set_union(p.z.begin(), p.z.end(), z.begin(), z.end(), std::back_inserter(u));
      uint64_t ucells = (uint64_t)cards[x] * cards[y];
      for (size_t i = 0; i < u.size() && ucells <= max_cells; ++i) ucells *= cards[u[i]];
      if (ucells <= max_cells && u.size() <= max_z) {
        if (u.size() != p.z.size()) p.z = u;
        p.tests.push_back((uint32_t)t);
        last_key = key;
        last_pool = pi;
        continue;
      }
    }
    open[key] = pools.size();
    last_key = key;
    last_pool = pools.size();
    pools.push_back({x, y, z, {(uint32_t)t}, false});
  }
  // ---- emit ------------------------------------------------------------------------------
  P = CiPoolPlan{};
  P.offsets.push_back(0);
  P.mem_off.push_back(0);
  for (const Pool& p : pools) {
    if (p.verbatim) {               // the test's own variable list, nothing summed out
      P.var_ids.insert(P.var_ids.end(), p.z.begin(), p.z.end());
      P.x.push_back(p.x);
      P.y.push_back(p.y);
      P.members.push_back(p.tests[0]);
      P.members.push_back(0xffffffffu);
    } else {
      bool px = false, py = false;  // merge x and y into the ascending list
      for (uint32_t v : p.z) {
        if (!px && p.x < v) { P.var_ids.push_back(p.x); px = true; }
        if (!py && p.y < v) { P.var_ids.push_back(p.y); py = true; }
        P.var_ids.push_back(v);
      }
      if (!px) P.var_ids.push_back(p.x);
      if (!py) P.var_ids.push_back(p.y);
      P.x.push_back(p.x);
      P.y.push_back(p.y);
      for (uint32_t t : p.tests) {
        uint32_t mask = 0;
        size_t j = 0;
        for (uint64_t i = offsets[t]; i < offsets[t + 1]; ++i) {
          const uint32_t v = var_ids[i];
          if (v == p.x || v == p.y) continue;
          while (p.z[j] != v) ++j;  // both ascending, the test's set is contained in the pool's
          mask |= 1u << j;
        }
        P.members.push_back(t);
        P.members.push_back(mask);
      }
    }
    P.offsets.push_back(P.var_ids.size());
    P.mem_off.push_back((uint32_t)(P.members.size() / 2));
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
  CiPoolPlan P;
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
