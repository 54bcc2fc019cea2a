// obn_ci_groups.inl -- host-side planner of the CI path: the tests of a batch that are asked over
// the SAME variable set share one contingency table.
//
// The reference builds one table per test (pc.cpp:38: ContingencyTable(sort(Z + {x, y}))) and
// citest() then locates x and y inside it (citest.h:66-70), i.e. the table does not depend on which
// two of its variables are tested.  A PC frontier asks the same set many times: for a star
// x - {y, z1 .. zk} every choice of the tested neighbour gives a question over {x, y, z1 .. zk}
// (pc.cpp:27-47: (x, y | Z) for every k-subset Z of adj(x) - {y}, for every neighbour y), so a
// depth-k round holds k+1 or more tests per distinct set.  One pass over the rows, one increment
// per row, serves all of them; the epilogue reads the table once per member with that member's
// (x, y) digits (obn_kernels.cuh: CiGeom).  Counts are the very same integers the reference's
// table holds, so dof is identical and the statistic is the same sum over the same slices.
//
// Pure host code (no CUDA calls): obn_ci_group_plan exposes it to the CPU tests.
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

struct CiGroupPlan {
  std::vector<uint32_t> var_ids;    // CSR of the groups: the shared variable set, ascending
  std::vector<uint64_t> offsets;    // [groups + 1]
  std::vector<uint32_t> mem_off;    // [groups + 1]
  std::vector<uint32_t> members;    // per member: test id, px | py << 8 (digits of x < y in the set; citest.h:48,63-67)
  size_t groups() const { return mem_off.empty() ? 0 : mem_off.size() - 1; }
};

// Working storage of the planner, kept per host thread so that a PC run (one plan per round)
// does not allocate.
struct CiGroupScratch {
  std::vector<uint32_t> head, tail, count, first_test;   // per group
  std::vector<uint32_t> next, role;                      // per test: next member of its group, px | py << 8
  std::vector<uint32_t> hval;                            // open addressing: slot -> newest group of a variable set
  std::vector<uint64_t> hkey;
  size_t hmask = 0;
  void hash_reset(size_t n) {
    size_t cap = 64;
    while (cap < 2 * n) cap <<= 1;
    if (hkey.size() != cap) {
      hkey.assign(cap, ~0ull);
      hval.resize(cap);
    } else {
      std::fill(hkey.begin(), hkey.end(), ~0ull);
    }
    hmask = cap - 1;
  }
};

// Order-preserving: groups appear in the order of their first test, members in batch order; a
// group takes at most kGroupMembers tests (the kernel keeps their statistics in shared memory),
// further tests over the same set open another group.
//
// rank / world: a sharded round.  Every rank plans the same batch but keeps only the variable sets whose
// hash falls to it, so all tests over one set land on one rank (the sharing survives the split), the
// ranks agree without talking, and each pays for planning and packing its own share only.  Test ids
// stay those of the whole batch.
static int plan_ci_groups(const uint32_t* cards, size_t n_vars, const uint32_t* var_ids, const uint64_t* offsets,
                          size_t n, const uint32_t* xs, const uint32_t* ys, CiGroupPlan& P, int rank = 0, int world = 1) {
  static thread_local CiGroupScratch W;
  W.hash_reset(n);
  W.head.clear(); W.tail.clear(); W.count.clear(); W.first_test.clear();
  W.next.resize(n);
  W.role.resize(n);
  for (size_t t = 0; t < n; ++t) {
    const uint64_t o0 = offsets[t], o1 = offsets[t + 1];
    if (o1 < o0) return fail(OBN_ERR_INVALID, "offsets must be non-decreasing");
    const uint32_t* ids = var_ids + o0;
    const size_t k = (size_t)(o1 - o0);
    const uint32_t x = std::min(xs[t], ys[t]), y = std::max(xs[t], ys[t]);     // citest.h:48
    // one scan: the checks of check_tuple, the digits of x and y (first occurrence, citest.h:63-64), the hash
    bool ok = k <= (size_t)kMaxTupleVars && x != y;
    uint64_t cl = 1, h = 0x9E3779B97F4A7C15ull ^ k;
    uint32_t px = 0xffu, py = 0xffu;
    for (size_t i = 0; ok && i < k; ++i) {
      const uint32_t v = ids[i];
      if (v >= n_vars || (i && ids[i - 1] > v)) { ok = false; break; }
      cl *= cards[v];
      if (cl >= (1ull << 31)) { ok = false; break; }
      if (v == x && px == 0xffu) px = (uint32_t)i;
      else if (v == y && py == 0xffu) py = (uint32_t)i;
      h = (h ^ v) * 0xFF51AFD7ED558CCDull;
      h ^= h >> 29;
    }
    if (!ok || px == 0xffu || py == 0xffu) {     // let check_tuple name the problem
      uint64_t dummy;
      if (int rc = check_tuple(cards, n_vars, ids, k, t, true, false, xs[t], ys[t], &dummy)) return rc;
      return fail(OBN_ERR_INVALID, "tuple %zu: inconsistent CI test", t);
    }
    if (world > 1 && (int)((h >> 40) % (uint64_t)world) != rank) continue;      // another rank's table
    W.role[t] = px | (py << 8);
    W.next[t] = 0xffffffffu;
    if (h == ~0ull) h = 0;
    size_t s = (size_t)(h >> 17) & W.hmask;
    uint32_t gi = 0xffffffffu;
    for (;; s = (s + 1) & W.hmask) {
      if (W.hkey[s] == ~0ull) break;
      if (W.hkey[s] != h) continue;
      const uint32_t ft = W.first_test[W.hval[s]];
      const uint64_t f0 = offsets[ft];
      if (offsets[ft + 1] - f0 == k && std::equal(ids, ids + k, var_ids + f0)) { gi = W.hval[s]; break; }
    }
    if (gi != 0xffffffffu && W.count[gi] < (uint32_t)kGroupMembers) {
      W.next[W.tail[gi]] = (uint32_t)t;
      W.tail[gi] = (uint32_t)t;
      ++W.count[gi];
      continue;
    }
    const uint32_t ng = (uint32_t)W.head.size();   // a new set, or the set's newest group is full
    W.head.push_back((uint32_t)t);
    W.tail.push_back((uint32_t)t);
    W.count.push_back(1u);
    W.first_test.push_back((uint32_t)t);
    W.hkey[s] = h;
    W.hval[s] = ng;
  }
  // ---- emit ------------------------------------------------------------------------------
  const size_t n_groups = W.head.size();
  P.var_ids.clear(); P.offsets.clear(); P.mem_off.clear();
  P.members.resize(2 * n + 2);
  P.var_ids.reserve(offsets[n]);
  P.offsets.reserve(n_groups + 1);
  P.mem_off.reserve(n_groups + 1);
  P.offsets.push_back(0);
  P.mem_off.push_back(0);
  size_t nm = 0;
  for (size_t q = 0; q < n_groups; ++q) {
    const uint32_t ft = W.first_test[q];
    P.var_ids.insert(P.var_ids.end(), var_ids + offsets[ft], var_ids + offsets[ft + 1]);
    for (uint32_t t = W.head[q]; t != 0xffffffffu; t = W.next[t]) {
      P.members[2 * nm] = t;
      P.members[2 * nm + 1] = W.role[t];
      ++nm;
    }
    P.offsets.push_back(P.var_ids.size());
    P.mem_off.push_back((uint32_t)nm);
  }
  return OBN_OK;
}

}  // namespace obn

extern "C" {

// Diagnostics, host only (include/openbnsl_b200.h): the plan obn_ci_tests / obn_pc_round use.
int obn_ci_group_plan(const uint32_t* cards, size_t n_vars, const uint32_t* var_ids, const uint64_t* offsets,
                      size_t n_tests, const uint32_t* x, const uint32_t* y, int rank, int world, uint32_t* group_var_ids,
                      uint64_t* group_offsets, uint32_t* group_of_test, uint32_t* digits_of_test, size_t* n_groups) {
  if (!cards || !offsets || !n_groups || (n_tests && (!x || !y))) return fail(OBN_ERR_INVALID, "NULL argument");
  if (n_tests && offsets[n_tests] && !var_ids) return fail(OBN_ERR_INVALID, "var_ids is NULL");
  if (world < 1 || rank < 0 || rank >= world) return fail(OBN_ERR_INVALID, "bad rank/world %d/%d", rank, world);
  static thread_local CiGroupPlan P;
  if (int rc = plan_ci_groups(cards, n_vars, var_ids, offsets, n_tests, x, y, P, rank, world)) return rc;
  *n_groups = P.groups();
  if (group_of_test) std::fill(group_of_test, group_of_test + n_tests, 0xffffffffu);      // tests of other ranks' tables
  if (group_offsets) memcpy(group_offsets, P.offsets.data(), P.offsets.size() * sizeof(uint64_t));
  if (group_var_ids && !P.var_ids.empty()) memcpy(group_var_ids, P.var_ids.data(), P.var_ids.size() * sizeof(uint32_t));
  for (size_t q = 0; q < P.groups(); ++q)
    for (uint32_t m = P.mem_off[q]; m < P.mem_off[q + 1]; ++m) {
      if (group_of_test) group_of_test[P.members[2 * m]] = (uint32_t)q;
      if (digits_of_test) digits_of_test[P.members[2 * m]] = P.members[2 * m + 1];
    }
  return OBN_OK;
}

}  // extern "C"
