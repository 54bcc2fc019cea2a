// obn_es.inl -- exhaustive search with the DP tables resident in HBM
// (include/openbnsl_b200.h, "exhaustive search on the device").  Included at the
// end of obn_capi.cu, after obn_batch.inl.

namespace obn {
struct EsLevels { uint64_t off[34]; };     // off[k] = number of subsets with fewer than k variables, k = 0 .. n+1
}

struct obn_es {
  obn_ctx* c = nullptr;
  const obn_dataset* ds = nullptr;
  int n = 0;
  double ess = 1.0;
  size_t max_parents = 0;
  int exact = 0;
  // DP tables (device).  bl_* are indexed [(parent candidates + child) mask * n + child] (es_dp_warp_kernel).
  double* bl_score = nullptr;
  uint32_t* bl_ps = nullptr;
  double* best_gs = nullptr;
  uint8_t* best_ch = nullptr;
  // marginalisation lattice (fast mode): F(S) per subset, block-major (obn_lattice.inl)
  double* F = nullptr;
  bool use_F = false;
  obn::EsLevels levels{};
  int h = 0, L = 0;
  size_t f_per_rank = 0;
  const uint32_t* perm = nullptr;   // lattice digit order != variable order: three 1024-entry tables, subset mask ->
                                    // mask in digit order = perm[m & 1023] | perm[1024 + (m >> 10 & 1023)] | perm[2048 + (m >> 20)]
  // all device buffers of a session come from c->es_pool (reused by the next session)
  std::vector<uint64_t> masks;
  size_t T = 0, buffer_doubles = 0;
  int k = 0;
  bool calc = false;
  bool from_lists = false;        // this level's subsets come from the device level lists (no host masks)
  int own_first = 0, own_end = 64;  // children whose DP columns this rank holds (obn_es_level_dp_part)
};

namespace obn {

// exhaustive_search.cpp:59-116 for one subset W per WARP, lane c = candidate child c.
//
// bl_score / bl_ps rows are indexed by the subset that PRODUCED them: entry [W * n + c] is the best
// (score, parent set) of child c over the parent sets inside W - {c} (the reference keeps the same
// value under the key (W - {c}, c), :33-36).  The candidates of (W, c) are its own parent set W - {c}
// and the entries [ (W - {j}) * n + c ] for j in W - {c} (:75-87), so for a fixed j all lanes read ONE
// row: k coalesced row reads and one row write per subset instead of k (k - 1) scattered sectors.
// Ties are broken exactly like the serial loops: `>=` in ascending j, then the sink scan in
// ascending child order with the reference's `bgs >= 0 || gs > bgs`.
//   LIST   uint32_t (lattice level lists) or unsigned long long (the host's Gosper lists)
//   USE_F  local scores from the lattice, ls(c | W - c) = F(W) - F(W - c), F block-major:
//          index = slot[S & (2^h - 1)] << L | S >> h;  else scores[t * k + rank of c in W]
__device__ __forceinline__ size_t es_f_index(uint32_t S, const uint32_t* perm, const uint32_t* slot, int h, int L) {
  if (perm) S = __ldg(perm + (S & 1023u)) | __ldg(perm + 1024 + ((S >> 10) & 1023u)) | __ldg(perm + 2048 + (S >> 20));
  return ((size_t)slot[S & ((1u << h) - 1u)] << L) | (S >> h);
}

// Multi-GPU: the DP shards by CHILD.  Entry [W * n + c] depends only on entries of the same child, so
// rank r keeps the columns of its children (a contiguous range) and never needs the others'; only
// the sink choice needs every child, and it merges through one all-gather of (best own sum, child)
// per level (EsPart) followed by es_dp_commit_kernel.  A group of 2^gshift lanes handles one subset,
// lane q = child own_first + q; unsharded: one warp per subset, all n children.
struct EsPart {
  double gs;      // best best_gs[W - c] + bls over this rank's children of W (-inf if it has none)
  int32_t ch;     // that child (INT_MAX if none)
  int32_t pad;
};

template <typename LIST, bool USE_F>
__global__ void __launch_bounds__(128) es_dp_warp_kernel(const LIST* list, size_t T, int k, int n, bool calc, const double* scores,
                                                       const double* F, const uint32_t* slot, const uint32_t* perm, int h, int L,
                                                       double* bl_score, uint32_t* bl_ps, double* best_gs, uint8_t* best_ch,
                                                       int own_first, int own_count, int gshift, EsPart* part) {
  const int lane = threadIdx.x & 31, width = 1 << gshift, q = lane & (width - 1);
  const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const size_t t = warp * (size_t)(32 >> gshift) + (size_t)(lane >> gshift);
  const bool in_range = t < T;                            // (no early return: the sink scan shuffles over full warps)
  const uint32_t W = in_range ? (uint32_t)list[t] : 0u;
  const int c = own_first + q;
  const bool active = in_range && q < own_count && ((W >> c) & 1u);
  const uint32_t pa = W & ~(1u << c);
  double bls = -INFINITY;
  uint32_t bps = pa;
  if (active && calc) {                                   // :65-73
    if (USE_F) {
      bls = F[es_f_index(W, perm, slot, h, L)] - F[es_f_index(pa, perm, slot, h, L)];
    } else {
      bls = scores[t * (size_t)k + __popc(W & ((1u << c) - 1u))];
    }
  }
  for (uint32_t m = W; m; m &= m - 1) {                   // :75-87, the same j for every lane of the subset
    const int j = __ffs((int)m) - 1;
    if (active && j != c) {
      const size_t at = (size_t)(W ^ (1u << j)) * n + c;
      const double s = bl_score[at];
      if (s >= bls) {
        bls = s;
        bps = bl_ps[at];
      }
    }
  }
  double gs = 0.0;
  if (active) {
    bl_score[(size_t)W * n + c] = bls;
    bl_ps[(size_t)W * n + c] = bps;
    gs = best_gs[pa] + bls;                               // :89-94
  }
  double bgs = 0.0;
  int bch = -1;
  const int base = lane & ~(width - 1);
  const uint32_t act = __ballot_sync(0xffffffffu, active);
  for (int qq = 0; qq < own_count; ++qq) {                // ascending child order
    const double g = __shfl_sync(0xffffffffu, gs, base + qq);
    if (((act >> (base + qq)) & 1u) && (bgs >= 0 || g > bgs)) {
      bgs = g;
      bch = own_first + qq;
    }
  }
  if (q == 0 && in_range) {                               // :115-116
    if (part) {
      part[t] = EsPart{bch >= 0 ? bgs : -INFINITY, bch >= 0 ? bch : INT_MAX, 0};
    } else {
      best_gs[W] = bgs;
      best_ch[W] = (uint8_t)bch;
    }
  }
}

// parts[r * T + t] = rank r's (sum, child) for subset list[t]; ranks own ascending child ranges, so the
// serial scan over children (:89-94) is a scan over ranks that only replaces on a strictly larger sum.
template <typename LIST>
__global__ void es_dp_commit_kernel(const LIST* list, size_t T, int world, const EsPart* parts, double* best_gs, uint8_t* best_ch) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  double bgs = 0.0;
  int bch = -1;
  for (int r = 0; r < world; ++r) {
    const EsPart p = parts[(size_t)r * T + t];
    if (p.ch != INT_MAX && (bgs >= 0 || p.gs > bgs)) {
      bgs = p.gs;
      bch = p.ch;
    }
  }
  const uint32_t W = (uint32_t)list[t];
  best_gs[W] = bgs;
  best_ch[W] = (uint8_t)bch;
}

// The same merge without gathering every rank's pair: ranks own ascending child ranges and the serial scan
// keeps the FIRST child that reaches the largest sum, so the choice is (max over ranks of gs, then the
// smallest child among the ranks whose gs equals that max) -- two all-reduces of 8 and 4 bytes per subset
// (reduced inside the NVSwitch) instead of an all-gather of 16 bytes per subset and rank.
template <int STEP>
__global__ void es_dp_reduce_kernel(size_t T, const EsPart* part, double* gs_max, int32_t* ch_min) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const EsPart p = part[t];
  if (STEP == 0) gs_max[t] = p.gs;                                              // -> all-reduce(max)
  else ch_min[t] = (p.ch != INT_MAX && p.gs == gs_max[t]) ? p.ch : INT_MAX;     // -> all-reduce(min)
}
template <typename LIST>
__global__ void es_dp_commit2_kernel(const LIST* list, size_t T, const double* gs_max, const int32_t* ch_min, double* best_gs,
                                     uint8_t* best_ch) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const uint32_t W = (uint32_t)list[t];
  const int32_t ch = ch_min[t];
  best_gs[W] = (ch == INT_MAX) ? 0.0 : gs_max[t];
  best_ch[W] = (uint8_t)((ch == INT_MAX) ? -1 : ch);
}

// Level lists for the lattice mode (no host mask lists there):
// lists[off[k] + rank] = W for every subset W: rank = position of W among the k-subsets in
// increasing numeric order = sum_i C(p_i, i+1) over its set bits p_0 < p_1 < ... (combinatorial
// number system), which is also the reference's Gosper order (next_combmask.h:41-46).
__global__ void es_level_lists_kernel(size_t combs, int n, EsLevels lv, uint32_t* lists) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= combs) return;
  const uint32_t W = (uint32_t)t;
  uint64_t rank = 0;
  int i = 0;
  for (uint32_t m = W; m; m &= m - 1) {
    const int p = __ffs((int)m) - 1;
    ++i;                                           // C(p, i)
    uint64_t c = 1;
    if (i > p) c = 0;
    else
      for (int j = 1; j <= i; ++j) c = c * (uint64_t)(p - i + j) / (uint64_t)j;
    rank += c;
  }
  lists[lv.off[__popc(W)] + rank] = W;
}


// Level lists on the device (they depend on n only; kept in the ctx pool for the next search).
static int es_ensure_lists(obn_es* e) {
  obn_ctx* c = e->c;
  EsPool& pool = c->es_pool;
  const int n = e->n;
  const size_t combs = (size_t)1 << n;
  const void* before = pool.d_lists.p;
  if (int rc = pool.d_lists.reserve(combs * sizeof(uint32_t))) return rc;
  if (pool.d_lists.p != before) pool.lists_n = 0;
  EsLevels lv{};
  uint64_t off = 0, binom = 1;
  for (int k = 0; k <= n; ++k) {
    if (k) binom = binom * (uint64_t)(n - k + 1) / (uint64_t)k;
    lv.off[k] = off;
    off += binom;
  }
  lv.off[n + 1] = off;
  e->levels = lv;
  if (pool.lists_n != n) {
    (void)cudaGetLastError();
    es_level_lists_kernel<<<(unsigned)((combs + 255) / 256), 256, 0, c->stream>>>(combs, n, lv, pool.d_lists.as<uint32_t>());
    OBN_CUDA(cudaGetLastError());
    OBN_CUDA(cudaStreamSynchronize(c->stream));
    ++c->prof.total_launches;
    pool.lists_n = n;
  }
  return OBN_OK;
}


// exhaustive_search.cpp:133-148 on the device: walk the sink choices from the full set down, one dependent read
// per variable (2n tiny loads instead of 2n + 1 blocking copies over PCIe).  out[0 .. n) = parent set of every
// child this rank owns (0 elsewhere), out[n] = bits of the best score, out[n + 1] = 0 or 1 + the step that found
// no sink choice (the search did not process every level).
__global__ void es_backtrack_kernel(int n, const uint8_t* __restrict__ best_ch, const uint32_t* __restrict__ bl_ps,
                                    const double* __restrict__ best_gs, int own_first, int own_end,
                                    unsigned long long* __restrict__ out) {
  if (threadIdx.x || blockIdx.x) return;
  uint32_t W = (uint32_t)(((uint64_t)1 << n) - 1);
  for (int v = 0; v < n; ++v) out[v] = 0ull;
  out[n] = (unsigned long long)__double_as_longlong(best_gs[W]);
  out[n + 1] = 0ull;
  for (int i = 0; i < n; ++i) {
    const uint32_t ch = best_ch[W];
    if (ch >= (uint32_t)n) {
      out[n + 1] = 1ull + (unsigned long long)i;
      return;
    }
    if ((int)ch >= own_first && (int)ch < own_end) out[ch] = bl_ps[(size_t)W * n + ch];     // row of W = pa + ch
    W &= ~(1u << ch);
  }
}

}  // namespace obn

extern "C" {

int obn_es_begin(obn_ctx* c, const obn_dataset* ds, double ess, size_t max_parents, int exact_order, obn_es** out) {
  if (!c || !ds || !out) return fail(OBN_ERR_INVALID, "obn_es_begin: NULL argument");
  *out = nullptr;
  if (ds->ctx != c) return fail(OBN_ERR_INVALID, "dataset belongs to another ctx");
  const size_t n = ds->n_vars;
  if (n == 0) return fail(OBN_ERR_INVALID, "exhaustive search needs at least one variable");
  if (max_parents > n - 1) return fail(OBN_ERR_INVALID, "max_parents must be in [0, num_of_vars-1]");   // :18-19
  if (n > 64) return fail(OBN_ERR_INVALID, "num_of_vars must be in [0, 64]");                          // :22-23
  if (n > 28) return fail(OBN_ERR_UNSUPPORTED, "exhaustive search: the n * 2^n DP tables for n > 28 exceed HBM");
  OBN_CUDA(cudaSetDevice(c->device));
  obn_es* e = new obn_es();
  e->c = c;
  e->ds = ds;
  e->n = (int)n;
  e->ess = ess;
  e->max_parents = max_parents;
  e->exact = exact_order;
  const size_t combs = (size_t)1 << n;
  EsPool& P = c->es_pool;
  if (P.in_use) {
    delete e;
    return fail(OBN_ERR_INVALID, "another exhaustive-search session is open on this ctx");
  }
  int rc = P.d_bl_score.reserve(combs * n * sizeof(double));
  if (!rc) rc = P.d_bl_ps.reserve(combs * n * sizeof(uint32_t));
  if (!rc) rc = P.d_best_gs.reserve(combs * sizeof(double));
  if (!rc) rc = P.d_best_ch.reserve(combs);
  if (rc) {
    delete e;
    return rc == OBN_ERR_OOM ? fail(OBN_ERR_OOM, "exhaustive search: DP tables for %zu variables do not fit in HBM", n) : rc;
  }
  P.in_use = true;
  e->bl_score = P.d_bl_score.as<double>();
  e->bl_ps = P.d_bl_ps.as<uint32_t>();
  e->best_gs = P.d_best_gs.as<double>();
  e->best_ch = P.d_best_ch.as<uint8_t>();
  OBN_CUDA(cudaMemsetAsync(e->best_gs, 0, sizeof(double), c->stream));      // best_gs_tbl[empty] = 0 (:42-44)
  OBN_CUDA(cudaMemsetAsync(e->best_ch, 0xff, combs, c->stream));            // "not processed yet" marker for finish
  *out = e;
  return OBN_OK;
}

void obn_es_end(obn_es* e) {
  if (!e) return;
  cudaSetDevice(e->c->device);
  cudaStreamSynchronize(e->c->stream);
  e->c->es_pool.in_use = false;          // the buffers stay with the ctx for the next session
  delete e;
}

int obn_es_level_begin(obn_es* e, const uint64_t* masks, size_t n_subsets, int k, size_t buffer_subsets) {
  if (!e) return fail(OBN_ERR_INVALID, "obn_es_level_begin: NULL argument");
  if (k < 0 || k > e->n) return fail(OBN_ERR_INVALID, "subset size %d out of range", k);
  if (buffer_subsets < n_subsets) buffer_subsets = n_subsets;
  obn_ctx* c = e->c;
  OBN_CUDA(cudaSetDevice(c->device));
  e->T = n_subsets;
  e->k = k;
  e->calc = (size_t)k <= e->max_parents + 1;                                      // :48
  e->from_lists = false;
  if (e->use_F) return OBN_OK;                 // lattice mode: F and the level lists are already on the device
  if (!masks) {
    // a level that needs no scores (k = 0 or k > max_parents + 1, exhaustive_search.cpp:48,65): the DP
    // walks the device level list, nothing to enumerate or upload
    if (e->calc && k > 0) return fail(OBN_ERR_INVALID, "obn_es_level_begin: masks are required for a scored level");
    if (int rc = es_ensure_lists(e)) return rc;
    if (n_subsets != (size_t)(e->levels.off[k + 1] - e->levels.off[k]))
      return fail(OBN_ERR_INVALID, "obn_es_level_begin: level %d has %llu subsets", k,
                  (unsigned long long)(e->levels.off[k + 1] - e->levels.off[k]));
    e->from_lists = true;
    e->masks.clear();
    return OBN_OK;
  }
  e->masks.assign(masks, masks + n_subsets);
  e->k = k;
  e->calc = (size_t)k <= e->max_parents + 1;                                      // :48
  e->buffer_doubles = buffer_subsets * (size_t)std::max(k, 1);
  if (int rc = c->es_pool.d_masks.reserve(std::max<size_t>(n_subsets * 8, 16))) return rc;
  if (int rc = c->es_pool.d_level.reserve(std::max<size_t>(e->buffer_doubles * 8, 16))) return rc;
  if (n_subsets)
    OBN_CUDA(cudaMemcpyAsync(c->es_pool.d_masks.p, masks, n_subsets * 8, cudaMemcpyHostToDevice, c->stream));
  OBN_CUDA(cudaStreamSynchronize(c->stream));
  c->prof.h2d_bytes += (double)(n_subsets * 8);
  return OBN_OK;
}

int obn_es_level_scores(obn_es* e, size_t begin, size_t end) {
  if (!e) return fail(OBN_ERR_INVALID, "obn_es_level_scores: NULL argument");
  if (end > e->T) end = e->T;
  if (begin >= end || !e->calc || e->k == 0 || e->use_F) return OBN_OK;
  const size_t k = (size_t)e->k;
  const size_t chunk = std::max<size_t>(1, ((size_t)1 << 21) / k);      // ~2M variable slots per batch
  std::vector<uint32_t> var_ids;
  std::vector<uint64_t> offsets;
  for (size_t b = begin; b < end; b += chunk) {
    const size_t m = std::min(chunk, end - b);
    var_ids.resize(m * k);
    offsets.resize(m + 1);
    size_t o = 0;
    for (size_t t = 0; t < m; ++t) {
      offsets[t] = o;
      for (uint64_t w = e->masks[b + t]; w; w &= w - 1) var_ids[o++] = (uint32_t)__builtin_ctzll(w);
      if (o != (t + 1) * k) return fail(OBN_ERR_INVALID, "subset %zu of the level does not have %zu variables", b + t, k);
    }
    offsets[m] = o;
    BatchSpec S{e->c, e->ds, var_ids.data(), offsets.data(), m, MODE_BDEU_ALL};
    S.ess = e->ess;
    S.exact = e->exact;
    S.scores_dev = e->c->es_pool.d_level.as<double>() + b * k;
    if (int rc = run_batch(S)) return rc;
  }
  return OBN_OK;
}

int obn_es_level_buffer(obn_es* e, void** scores_dev, size_t* n_doubles) {
  if (!e || !scores_dev || !n_doubles) return fail(OBN_ERR_INVALID, "obn_es_level_buffer: NULL argument");
  *scores_dev = e->c->es_pool.d_level.p;
  *n_doubles = e->buffer_doubles;
  return OBN_OK;
}

// one DP launch for the level: children [own_first, own_first + own_count), results to the tables
// (part == nullptr) or to `part` (child-sharded)
static int es_launch_dp(obn_es* e, int own_first, int own_count, obn::EsPart* part) {
  obn_ctx* c = e->c;
  int gshift = 0;
  while ((1 << gshift) < own_count) ++gshift;
  const size_t per_cta = (size_t)4 * (32 >> gshift);                 // subsets per 128-thread CTA
  const unsigned grid = (unsigned)((e->T + per_cta - 1) / per_cta);
  (void)cudaGetLastError();
  if (e->use_F) {
    es_dp_warp_kernel<uint32_t, true><<<grid, 128, 0, c->stream>>>(
        c->es_pool.d_lists.as<uint32_t>() + e->levels.off[e->k], e->T, e->k, e->n, e->calc, nullptr, e->F,
        c->es_pool.d_slot.as<uint32_t>(), e->perm, e->h, e->L, e->bl_score, e->bl_ps, e->best_gs, e->best_ch, own_first, own_count,
        gshift, part);
  } else if (e->from_lists) {
    es_dp_warp_kernel<uint32_t, false><<<grid, 128, 0, c->stream>>>(
        c->es_pool.d_lists.as<uint32_t>() + e->levels.off[e->k], e->T, e->k, e->n, false, nullptr, nullptr, nullptr, nullptr,
        0, 0, e->bl_score, e->bl_ps, e->best_gs, e->best_ch, own_first, own_count, gshift, part);
  } else {
    es_dp_warp_kernel<unsigned long long, false><<<grid, 128, 0, c->stream>>>(
        c->es_pool.d_masks.as<unsigned long long>(), e->T, e->k, e->n, e->calc && e->k, c->es_pool.d_level.as<double>(), nullptr,
        nullptr, nullptr, 0, 0, e->bl_score, e->bl_ps, e->best_gs, e->best_ch, own_first, own_count, gshift, part);
  }
  OBN_CUDA(cudaGetLastError());
  ++c->prof.total_launches;
  return OBN_OK;
}

int obn_es_level_dp(obn_es* e) {
  if (!e) return fail(OBN_ERR_INVALID, "obn_es_level_dp: NULL argument");
  if (!e->T) return OBN_OK;
  obn_ctx* c = e->c;
  OBN_CUDA(cudaSetDevice(c->device));
  if (e->own_first != 0 || e->own_end < e->n)
    return fail(OBN_ERR_INVALID, "obn_es_level_dp after obn_es_level_dp_part: the DP columns are sharded by child");
  if (int rc = es_launch_dp(e, 0, e->n, nullptr)) return rc;
  OBN_CUDA(cudaStreamSynchronize(c->stream));
  return OBN_OK;
}

int obn_es_level_dp_part(obn_es* e, int rank, int world, void** parts_dev, size_t* bytes_per_rank) {
  if (!e || !parts_dev || !bytes_per_rank) return fail(OBN_ERR_INVALID, "obn_es_level_dp_part: NULL argument");
  if (world < 1 || rank < 0 || rank >= world) return fail(OBN_ERR_INVALID, "bad rank/world");
  if (!e->use_F) return fail(OBN_ERR_UNSUPPORTED, "the child-sharded DP needs the lattice mode (obn_es_lattice)");
  obn_ctx* c = e->c;
  OBN_CUDA(cudaSetDevice(c->device));
  const int first = (int)((size_t)e->n * (size_t)rank / (size_t)world), end = (int)((size_t)e->n * (size_t)(rank + 1) / (size_t)world);
  if (e->own_first != 0 || e->own_end < e->n) {
    if (e->own_first != first || e->own_end != end) return fail(OBN_ERR_INVALID, "rank/world changed inside a search");
  }
  e->own_first = first;
  e->own_end = end;
  const size_t bytes = std::max<size_t>(e->T, 1) * sizeof(obn::EsPart);
  if (int rc = c->es_pool.d_level.reserve(bytes * (size_t)world)) return rc;
  *parts_dev = c->es_pool.d_level.p;
  *bytes_per_rank = e->T * sizeof(obn::EsPart);
  if (!e->T) return OBN_OK;
  obn::EsPart* mine = reinterpret_cast<obn::EsPart*>(c->es_pool.d_level.as<char>() + (size_t)rank * e->T * sizeof(obn::EsPart));
  if (int rc = es_launch_dp(e, first, end - first, mine)) return rc;
  OBN_CUDA(cudaStreamSynchronize(c->stream));
  return OBN_OK;
}

// One level of the child-sharded DP entirely on the ctx's stream: this rank's children, the two
// all-reduces over the ctx's communicator, the commit -- no host synchronisation in between (the next
// level's kernels are ordered behind it on the same stream).
int obn_es_level_dp_sharded(obn_es* e) {
  if (!e) return fail(OBN_ERR_INVALID, "obn_es_level_dp_sharded: NULL argument");
  if (!e->use_F) return fail(OBN_ERR_UNSUPPORTED, "the child-sharded DP needs the lattice mode (obn_es_lattice)");
  obn_ctx* c = e->c;
  if (!c->comm || c->comm_world <= 1) return fail(OBN_ERR_COMM, "obn_es_level_dp_sharded needs a communicator (obn_comm_init)");
  OBN_CUDA(cudaSetDevice(c->device));
  const int rank = c->comm_rank, world = c->comm_world;
  const int first = (int)((size_t)e->n * (size_t)rank / (size_t)world), end = (int)((size_t)e->n * (size_t)(rank + 1) / (size_t)world);
  if (e->own_first != 0 || e->own_end < e->n) {
    if (e->own_first != first || e->own_end != end) return fail(OBN_ERR_INVALID, "rank/world changed inside a search");
  }
  e->own_first = first;
  e->own_end = end;
  if (!e->T) return OBN_OK;
  const size_t T = e->T;
  if (int rc = c->es_pool.d_level.reserve(T * 32 + 64)) return rc;
  obn::EsPart* part = c->es_pool.d_level.as<obn::EsPart>();
  double* gs_max = reinterpret_cast<double*>(c->es_pool.d_level.as<char>() + T * 16);
  int32_t* ch_min = reinterpret_cast<int32_t*>(c->es_pool.d_level.as<char>() + T * 24);
  if (int rc = es_launch_dp(e, first, end - first, part)) return rc;
  const unsigned grid = (unsigned)((T + 255) / 256);
  NcclApi* api = nccl_api();
  if (!api) return fail(OBN_ERR_COMM, "NCCL is not available");
  es_dp_reduce_kernel<0><<<grid, 256, 0, c->stream>>>(T, part, gs_max, ch_min);
  OBN_NCCL(api, api->AllReduce(gs_max, gs_max, T, ncclDouble, ncclMax, (ncclComm_t)c->comm, c->stream));
  es_dp_reduce_kernel<1><<<grid, 256, 0, c->stream>>>(T, part, gs_max, ch_min);
  OBN_NCCL(api, api->AllReduce(ch_min, ch_min, T, ncclInt32, ncclMin, (ncclComm_t)c->comm, c->stream));
  es_dp_commit2_kernel<uint32_t><<<grid, 256, 0, c->stream>>>(c->es_pool.d_lists.as<uint32_t>() + e->levels.off[e->k], T, gs_max,
                                                              ch_min, e->best_gs, e->best_ch);
  OBN_CUDA(cudaGetLastError());
  c->prof.total_launches += 3;
  c->prof.collectives += 2;
  return OBN_OK;
}

int obn_es_level_dp_commit(obn_es* e, int world) {
  if (!e) return fail(OBN_ERR_INVALID, "obn_es_level_dp_commit: NULL argument");
  if (!e->use_F || world < 1) return fail(OBN_ERR_INVALID, "obn_es_level_dp_commit without obn_es_level_dp_part");
  if (!e->T) return OBN_OK;
  obn_ctx* c = e->c;
  OBN_CUDA(cudaSetDevice(c->device));
  (void)cudaGetLastError();
  es_dp_commit_kernel<uint32_t><<<(unsigned)((e->T + 255) / 256), 256, 0, c->stream>>>(
      c->es_pool.d_lists.as<uint32_t>() + e->levels.off[e->k], e->T, world, c->es_pool.d_level.as<obn::EsPart>(), e->best_gs,
      e->best_ch);
  OBN_CUDA(cudaGetLastError());
  OBN_CUDA(cudaStreamSynchronize(c->stream));
  ++c->prof.total_launches;
  return OBN_OK;
}

int obn_es_finish(obn_es* e, uint64_t* parent_masks, double* best_score) {
  if (!e || !parent_masks) return fail(OBN_ERR_INVALID, "obn_es_finish: NULL argument");
  obn_ctx* c = e->c;
  OBN_CUDA(cudaSetDevice(c->device));
  const int n = e->n;
  for (int v = 0; v < n; ++v) parent_masks[v] = 0;
  // one tiny kernel behind whatever is still enqueued on the ctx's stream (the levels of a child-sharded DP are
  // only enqueued, obn_es_level_dp_sharded), one copy back
  if (int rc = c->d_scalar.reserve((size_t)(n + 2) * 8)) return rc;
  unsigned long long* d_out = c->d_scalar.as<unsigned long long>();
  es_backtrack_kernel<<<1, 32, 0, c->stream>>>(n, e->best_ch, e->bl_ps, e->best_gs, e->own_first, e->own_end, d_out);
  OBN_CUDA(cudaGetLastError());
  unsigned long long h_out[66];
  OBN_CUDA(cudaMemcpyAsync(h_out, d_out, (size_t)(n + 2) * 8, cudaMemcpyDeviceToHost, c->stream));
  OBN_CUDA(cudaStreamSynchronize(c->stream));
  ++c->prof.total_launches;
  if (h_out[n + 1]) return fail(OBN_ERR_INVALID, "obn_es_finish called before every level was processed");
  for (int v = 0; v < n; ++v) parent_masks[v] = h_out[v];
  if (best_score) memcpy(best_score, &h_out[n], 8);
  return OBN_OK;
}

}  // extern "C"
