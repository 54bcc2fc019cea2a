// obn_lattice.inl -- host orchestration of the marginalisation lattice (obn_lattice.cuh).
// Included by obn_capi.cu after obn_es.inl.
//
// Variables 0..h-1 ("high", the most significant digits of the key, contingency_table.h:76-85)
// select a BLOCK; variables h..n-1 ("low") span a sub-lattice inside each block.
//   1. joint table J of all n variables: one counting launch (the only pass over the data);
//   2. root lattice: R_H = table of (H + all low variables) for every H subset of the high
//      variables, level by level from J, all kept (<= prod(1+c_high) * cells(low) cells);
//   3. per block H: tables of H + T for every T subset of the low variables, level by level
//      from R_H, two levels resident at a time (ping-pong), F(H+T) accumulated on the way.
// F is stored block-major: F[slot(H) * 2^L + maskT], so that the blocks of one rank are one
// contiguous slice for the all-gather.

namespace obn {

struct LatticePlan {
  int n = 0, h = 0, L = 0;
  std::vector<uint32_t> card;
  uint64_t cells_low = 1;                 // prod cards of the low variables
  std::vector<uint64_t> cells_high;       // [2^h] prod cards of the block's high variables
  std::vector<uint64_t> root_off;         // [2^h] cell offset of R_H in the root buffer
  uint64_t root_cells = 0;
  std::vector<uint32_t> slot;             // [2^h] block -> slot (block-major F index, rank-contiguous)
  // low lattice, shared by all blocks (units are multiplied by cells_high[H])
  std::vector<std::vector<MargTask>> level_tasks;   // [L] tasks producing level k (|T| = k) from level k+1
  std::vector<uint64_t> level_units;      // [L+1] sum of cells_low(T) over |T| = k
  std::vector<uint32_t> level_max_units;  // [L+1] max cells_low(T) over |T| = k
  // root lattice
  std::vector<std::vector<MargTask>> root_tasks;    // [h] tasks producing |H| = k roots from |H| = k+1
  std::vector<uint32_t> root_level_max;   // [h+1] max cells of a root at level k, in units of cells_low
};

static bool plan_lattice(const std::vector<uint32_t>& card, size_t budget_bytes, int world, LatticePlan& P) {
  const int n = (int)card.size();
  P.n = n;
  P.card = card;
  double joint = 1;
  for (int v = 0; v < n; ++v) joint *= card[v];
  if (joint >= 2147483648.0) return false;                          // dense joint table must fit int32 indexing
  // the largest L whose two largest adjacent levels (times all high cells) fit the budget
  int best_L = -1;
  for (int L = std::min(n, 20); L >= 1; --L) {
    const int h = n - L;
    if (h > 16) break;
    if (world > 1 && ((1 << h) % world != 0 || (1 << h) < world)) continue;
    // multi-GPU balance: blocks are dealt whole, so the heaviest block (all high variables
    // present) must be a small share of the total: prod c_high / prod (1 + c_high) <= 1 / (2 world)
    if (world > 1 && L > 1) {
      double heavy = 1, total = 1;
      for (int v = 0; v < h; ++v) { heavy *= card[v]; total *= 1.0 + card[v]; }
      if (heavy / total > 1.0 / (2.0 * world)) continue;
    }
    // units per level by DP over the low variables: poly = prod (1 + c x)
    std::vector<double> lvl(L + 1, 0.0);
    lvl[0] = 1;
    for (int i = 0; i < L; ++i)
      for (int k = i + 1; k >= 1; --k) lvl[k] += lvl[k - 1] * card[h + i];
    double cells_all_high = 1, roots = 1, low = 1;
    for (int v = 0; v < h; ++v) { cells_all_high *= card[v]; roots *= 1.0 + card[v]; }
    for (int v = h; v < n; ++v) low *= card[v];
    double pair = 0, maxu = 0;
    for (int k = 0; k < L; ++k) pair = std::max(pair, lvl[k] + lvl[k + 1]);
    for (int k = 0; k <= L; ++k) maxu = std::max(maxu, lvl[k]);
    if (maxu >= 4294967296.0) continue;
    const double bytes = 4.0 * (pair * cells_all_high + roots * low);
    if (bytes <= (double)budget_bytes) { best_L = L; break; }
  }
  if (best_L < 0) return false;
  const int L = best_L, h = n - L;
  P.L = L;
  P.h = h;
  P.cells_low = 1;
  for (int v = h; v < n; ++v) P.cells_low *= card[v];
  const uint32_t B = 1u << h;
  P.cells_high.assign(B, 1);
  for (uint32_t H = 0; H < B; ++H)
    for (int v = 0; v < h; ++v)
      if (H >> v & 1) P.cells_high[H] *= card[v];
  // root storage, level-major descending (allHigh = J first)
  P.root_off.assign(B, 0);
  {
    uint64_t off = 0;
    for (int k = h; k >= 0; --k)
      for (uint32_t H = 0; H < B; ++H)
        if (__builtin_popcount(H) == k) { P.root_off[H] = off; off += P.cells_high[H] * P.cells_low; }
    P.root_cells = off;
    if (off >= ((uint64_t)1 << 32)) return false;                   // MargTask offsets are 32-bit
  }
  // slots: heaviest blocks dealt round-robin over the ranks, each rank's blocks contiguous
  {
    std::vector<uint32_t> order(B);
    for (uint32_t H = 0; H < B; ++H) order[H] = H;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return P.cells_high[a] > P.cells_high[b]; });
    P.slot.assign(B, 0);
    const uint32_t per = B / (uint32_t)world;
    std::vector<uint32_t> fill(world, 0);
    for (uint32_t i = 0; i < B; ++i) {                      // snake order: 0..w-1, w-1..0, ... (no rank always first)
      const uint32_t round = i / (uint32_t)world, pos = i % (uint32_t)world;
      const uint32_t r = (round & 1u) ? (uint32_t)world - 1u - pos : pos;
      P.slot[order[i]] = r * per + fill[r]++;
    }
  }
  // low lattice: O(1) work per subset (products by peeling the lowest bit, tasks placed by rank)
  const uint32_t NT = 1u << L;
  std::vector<uint32_t> cl(NT, 1), off(NT, 0);
  for (uint32_t T = 1; T < NT; ++T) cl[T] = cl[T & (T - 1)] * card[h + __builtin_ctz(T)];
  P.level_units.assign(L + 1, 0);
  P.level_max_units.assign(L + 1, 0);
  std::vector<uint32_t> level_count(L + 1, 0);
  for (uint32_t T = 0; T < NT; ++T) {
    const int k = __builtin_popcount(T);
    off[T] = (uint32_t)P.level_units[k];
    P.level_units[k] += cl[T];
    P.level_max_units[k] = std::max(P.level_max_units[k], cl[T]);
    ++level_count[k];
  }
  std::vector<uint32_t> pre(L + 1, 1);                              // pre[i] = prod of the first i low cards
  for (int i = 0; i < L; ++i) pre[i + 1] = pre[i] * card[h + i];
  P.level_tasks.assign(L, {});
  for (int k = 0; k < L; ++k) P.level_tasks[k].resize(level_count[k]);
  std::vector<uint32_t> fill(L + 1, 0);
  for (uint32_t T = 0; T + 1 < NT; ++T) {
    const int k = __builtin_popcount(T);
    const int i = __builtin_ctz(~T);                                // lowest absent low variable
    MargTask& t = P.level_tasks[k][fill[k]++];
    t.src_off = off[T | (1u << i)];
    t.dst_off = off[T];
    t.hi = pre[i];                                                  // bits below i are all present in T
    t.r = card[h + i];
    const uint32_t lo = cl[T & ~((2u << i) - 1u)];                  // T's variables above i
    t.lo = lo;
    t.f_index = T;
    t.lo_shift = (t.r == 2 && lo >= 4 && (lo & (lo - 1)) == 0) ? __builtin_ctz(lo) : -1;
    t.pad1 = 0;
  }
  // root lattice (offsets in cells, mul = 1; hi/lo include the low variables' cells in lo)
  P.root_tasks.assign(std::max(h, 1), {});
  P.root_level_max.assign(h + 1, 0);
  for (uint32_t H = 0; H < B; ++H) {
    const int k = __builtin_popcount(H);
    P.root_level_max[k] = (uint32_t)std::max<uint64_t>(P.root_level_max[k], P.cells_high[H]);
    if (k == h) continue;
    const int i = __builtin_ctz(~H);
    const uint32_t S = H | (1u << i);
    MargTask t{};
    t.src_off = (uint32_t)P.root_off[S];
    t.dst_off = (uint32_t)P.root_off[H];
    uint64_t hi = 1, lo = P.cells_low;
    for (int j = 0; j < i; ++j) hi *= card[j];
    for (int j = i + 1; j < h; ++j)
      if (H >> j & 1) lo *= card[j];
    if (lo >= ((uint64_t)1 << 32)) return false;
    t.hi = (uint32_t)hi;
    t.r = card[i];
    t.lo = (uint32_t)lo;
    t.f_index = P.slot[H] * NT + (NT - 1);
    t.lo_shift = (t.r == 2 && lo >= 4 && (lo & (lo - 1)) == 0) ? __builtin_ctzll(lo) : -1;
    P.root_tasks[k].push_back(t);
  }
  return true;
}

}  // namespace obn

namespace obn {

static int launch_marg_level(obn_ctx* c, const MargTask* d_tasks, size_t n_tasks, uint64_t max_cells, uint64_t mul,
                             const int32_t* src, int32_t* dst, double ess, const TermTable& tt, DevBuf& partials,
                             double* F, uint64_t f_base) {
  const uint32_t chunks = (uint32_t)((max_cells + kMargChunk - 1) / kMargChunk);
  for (size_t b = 0; b < n_tasks; b += 65535) {
    const uint32_t m = (uint32_t)std::min<size_t>(65535, n_tasks - b);
    if (int rc = partials.reserve((size_t)m * chunks * sizeof(double))) return rc;
    (void)cudaGetLastError();
    marg_kernel<<<dim3(chunks, m), kMargThreads, 0, c->stream>>>(d_tasks + b, mul, src, dst, ess, tt, partials.as<double>());
    OBN_CUDA(cudaGetLastError());
    marg_reduce_kernel<<<(m + 127) / 128, 128, 0, c->stream>>>(d_tasks + b, m, chunks, partials.as<double>(), F, f_base);
    OBN_CUDA(cudaGetLastError());
    c->prof.total_launches += 2;
    c->prof.count_launches += 1;
  }
  return OBN_OK;
}

}  // namespace obn

extern "C" {

int obn_es_lattice(obn_es* e, int rank, int world) {
  if (!e) return fail(OBN_ERR_INVALID, "obn_es_lattice: NULL argument");
  if (world < 1 || rank < 0 || rank >= world) return fail(OBN_ERR_INVALID, "bad rank/world");
  obn_ctx* c = e->c;
  const obn_dataset* ds = e->ds;
  OBN_CUDA(cudaSetDevice(c->device));
  const int n = e->n;
  size_t free_b = 0, total_b = 0;
  OBN_CUDA(cudaMemGetInfo(&free_b, &total_b));
  const size_t pooled = c->es_pool.d_root.cap + c->es_pool.d_pp[0].cap + c->es_pool.d_pp[1].cap;   // reusable
  const size_t budget = std::min<size_t>((size_t)((free_b + pooled) * 0.7), (size_t)env_int("OBN_LATTICE_BUDGET_GB", 24) << 30);
  const bool dbg = env_int("OBN_DEBUG", 0) != 0;
  auto t_dbg = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (!dbg) return;
    cudaStreamSynchronize(c->stream);
    const auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "[obn_es_lattice] %-14s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_dbg).count());
    t_dbg = now;
  };
  LatticePlan P;
  if (!plan_lattice(ds->cards, budget, world, P))
    return fail(OBN_ERR_UNSUPPORTED, "marginalisation lattice does not fit (joint table >= 2^31 cells or too little memory)");
  lap("plan");
  const int h = P.h, L = P.L;
  const uint32_t B = 1u << h, NT = 1u << L;
  if (dbg) fprintf(stderr, "[obn_es_lattice] n=%d h=%d L=%d budget=%.1f GB\n", n, h, L, budget / 1073741824.0);
  const size_t combs = (size_t)1 << n;

  // ---- device buffers ---------------------------------------------------------------------
  uint64_t cells_all_high = P.cells_high[B - 1], buf_units = 0;
  for (int k = 0; k < L; ++k) buf_units = std::max<uint64_t>(buf_units, P.level_units[k]);
  if (int rc = c->es_pool.d_F.reserve(combs * sizeof(double))) return rc;
  e->F = c->es_pool.d_F.as<double>();
  if (int rc = c->es_pool.d_root.reserve(std::max<uint64_t>(P.root_cells, 4) * 4)) return rc;
  for (int i = 0; i < 2; ++i)
    if (int rc = c->es_pool.d_pp[i].reserve(std::max<uint64_t>(buf_units * cells_all_high, 4) * 4)) return rc;
  lap("alloc");
  // tasks: low levels then root levels, one upload
  size_t n_all = 1;
  std::vector<size_t> low_at(L + 1, 0), root_at(h + 1, 0);
  for (int k = 0; k < L; ++k) { low_at[k] = n_all; n_all += P.level_tasks[k].size(); }
  for (int k = 0; k < h; ++k) { root_at[k] = n_all; n_all += P.root_tasks[k].size(); }
  const size_t joint_task_at = 0;
  if (int rc = c->es_pool.d_tasks.reserve(n_all * sizeof(MargTask))) return rc;
  if (int rc = c->h_in.reserve(n_all * sizeof(MargTask))) return rc;               // page-locked staging
  {
    MargTask* all = c->h_in.as<MargTask>();
    MargTask t{};
    t.f_index = P.slot[B - 1] * NT + (NT - 1);
    t.lo_shift = -1;
    all[0] = t;
    for (int k = 0; k < L; ++k)
      if (!P.level_tasks[k].empty()) memcpy(all + low_at[k], P.level_tasks[k].data(), P.level_tasks[k].size() * sizeof(MargTask));
    for (int k = 0; k < h; ++k)
      if (!P.root_tasks[k].empty()) memcpy(all + root_at[k], P.root_tasks[k].data(), P.root_tasks[k].size() * sizeof(MargTask));
  }
  OBN_CUDA(cudaMemcpyAsync(c->es_pool.d_tasks.p, c->h_in.p, n_all * sizeof(MargTask), cudaMemcpyHostToDevice, c->stream));
  OBN_CUDA(cudaStreamSynchronize(c->stream));                                        // h_in is reused by run_batch
  if (int rc = c->es_pool.d_slot.reserve(B * 4)) return rc;
  OBN_CUDA(cudaMemcpyAsync(c->es_pool.d_slot.p, P.slot.data(), B * 4, cudaMemcpyHostToDevice, c->stream));
  e->h = h;
  e->L = L;

  lap("tasks upload");
  // ---- term tables: one per distinct cell count (device libm; this is the fast mode) ---------
  TermTable tt{nullptr, nullptr, 0, 0};
  {
    // distinct cell counts of all subsets = distinct products over sub-multisets of the cards
    std::vector<uint32_t> distinct{1};
    for (int v = 0; v < n && distinct.size() <= 4096; ++v) {
      const size_t m = distinct.size();
      for (size_t i = 0; i < m; ++i) {
        const uint64_t p = (uint64_t)distinct[i] * ds->cards[v];
        if (p < ((uint64_t)1 << 31)) distinct.push_back((uint32_t)p);
      }
      std::sort(distinct.begin(), distinct.end());
      distinct.erase(std::unique(distinct.begin(), distinct.end()), distinct.end());
    }
    const size_t stride = ds->n_rows + 1;
    if (distinct.size() <= 256 && distinct.size() * stride * 8 <= ((size_t)2 << 30)) {
      std::vector<double> a(distinct.size());
      for (size_t i = 0; i < distinct.size(); ++i) a[i] = e->ess / (double)(size_t)distinct[i];
      if (int rc = c->es_pool.d_lut_a.reserve(a.size() * 8)) return rc;
      if (int rc = c->es_pool.d_lutF.reserve(a.size() * stride * 8)) return rc;
      OBN_CUDA(cudaMemcpyAsync(c->es_pool.d_lut_a.p, a.data(), a.size() * 8, cudaMemcpyHostToDevice, c->stream));
      OBN_CUDA(cudaStreamSynchronize(c->stream));          // `a` leaves scope
      (void)cudaGetLastError();
      lut_fill_kernel<<<dim3((unsigned)((stride + 255) / 256), (unsigned)a.size()), 256, 0, c->stream>>>(
          c->es_pool.d_lutF.as<double>(), c->es_pool.d_lut_a.as<double>(), stride);
      OBN_CUDA(cudaGetLastError());
      tt = TermTable{c->es_pool.d_lutF.as<double>(), c->es_pool.d_lut_a.as<double>(), (int)a.size(), stride};
    }
    // accounting: what the reference would have read for the subsets whose scores are used:
    // sum over k <= max_parents+1 of C(n,k) * k * N bytes
    double alg = 0, binom = 1;
    for (int k = 1; k <= n && (size_t)k <= e->max_parents + 1; ++k) {
      binom = binom * (double)(n - k + 1) / (double)k;
      alg += binom * (double)k * (double)ds->n_rows;
    }
    c->prof.algorithmic_bytes += alg / world;
  }

  lap("term tables");
  // ---- 1. joint table (the only pass over the data) -------------------------------------------
  int32_t* root = c->es_pool.d_root.as<int32_t>();
  {
    const uint64_t cells = P.cells_high[B - 1] * P.cells_low;
    int32_t* J = root + P.root_off[B - 1];
    OBN_CUDA(cudaMemsetAsync(J, 0, cells * 4, c->stream));
    JointCards jc{};
    for (int v = 0; v < n; ++v) jc.c[v] = ds->cards[v];
    const uint32_t n_vec = (uint32_t)((ds->n_rows + 15) >> 4);
    (void)cudaGetLastError();
    joint_count_kernel<<<std::max(1u, (n_vec + 127) / 128), 128, 0, c->stream>>>(ds->data, ds->stride, (uint32_t)ds->n_rows, n, jc, J);
    OBN_CUDA(cudaGetLastError());
    ++c->prof.total_launches;
    ++c->prof.count_launches;
  }
  cudaEvent_t e0 = c->ev0, e1 = c->ev1;
  const MargTask* d_tasks = c->es_pool.d_tasks.as<MargTask>();
  // The lattice is thousands of small dependent launches (blocks x levels x 2): capture them into
  // ONE CUDA graph so that the GPU never waits for the host thread (which shares its cores with
  // the other ranks' processes).  All buffers are reserved before the capture starts.
  {
    size_t need = (size_t)((P.cells_high[B - 1] * P.cells_low + kMargChunk - 1) / kMargChunk);
    for (int k = 0; k < h; ++k)
      need = std::max(need, std::min<size_t>(65535, P.root_tasks[k].size()) *
                                (size_t)(((uint64_t)P.root_level_max[k] * P.cells_low + kMargChunk - 1) / kMargChunk));
    for (int k = 0; k < L; ++k)
      need = std::max(need, std::min<size_t>(65535, P.level_tasks[k].size()) *
                                (size_t)(((uint64_t)P.level_max_units[k] * cells_all_high + kMargChunk - 1) / kMargChunk));
    if (int rc = c->es_pool.d_partials.reserve(need * sizeof(double))) return rc;
  }
  const bool use_graph = env_int("OBN_LATTICE_GRAPH", 1) != 0;
  OBN_CUDA(cudaStreamSynchronize(c->stream));
  if (use_graph) OBN_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
  else OBN_CUDA(cudaEventRecord(e0, c->stream));
  const bool dbg_saved = dbg;
  auto lap_if = [&](const char* what) { if (!use_graph) lap(what); };
  (void)dbg_saved;
  {
    const uint64_t cells = P.cells_high[B - 1] * P.cells_low;
    const uint32_t chunks = (uint32_t)((cells + kMargChunk - 1) / kMargChunk);
    if (int rc = c->es_pool.d_partials.reserve((size_t)chunks * 8)) return rc;
    (void)cudaGetLastError();
    table_f_kernel<<<chunks, kMargThreads, 0, c->stream>>>(root + P.root_off[B - 1], cells, e->ess, tt, c->es_pool.d_partials.as<double>());
    OBN_CUDA(cudaGetLastError());
    marg_reduce_kernel<<<1, 128, 0, c->stream>>>(d_tasks + joint_task_at, 1, chunks, c->es_pool.d_partials.as<double>(), e->F, 0);
    OBN_CUDA(cudaGetLastError());
  }
  lap_if("joint");
  // ---- 2. root lattice --------------------------------------------------------------------------
  for (int k = h - 1; k >= 0; --k)
    if (int rc = launch_marg_level(c, d_tasks + root_at[k], P.root_tasks[k].size(), (uint64_t)P.root_level_max[k] * P.cells_low,
                                   1, root, root, e->ess, tt, c->es_pool.d_partials, e->F, 0)) {
      if (use_graph) { cudaGraph_t g = nullptr; cudaStreamEndCapture(c->stream, &g); if (g) cudaGraphDestroy(g); }
      return rc;
    }
  lap_if("roots");
  // ---- 3. blocks of this rank ---------------------------------------------------------------------
  const uint32_t per = B / (uint32_t)world;
  for (uint32_t H = 0; H < B; ++H) {
    const uint32_t s = P.slot[H];
    if (s / per != (uint32_t)rank) continue;
    const uint64_t mul = P.cells_high[H];
    for (int k = L - 1; k >= 0; --k) {
      const int32_t* src = (k == L - 1) ? root + P.root_off[H] : c->es_pool.d_pp[(k + 1) & 1].as<int32_t>();
      if (int rc = launch_marg_level(c, d_tasks + low_at[k], P.level_tasks[k].size(), (uint64_t)P.level_max_units[k] * mul, mul,
                                     src, c->es_pool.d_pp[k & 1].as<int32_t>(), e->ess, tt, c->es_pool.d_partials, e->F, (uint64_t)s * NT)) {
        if (use_graph) { cudaGraph_t g = nullptr; cudaStreamEndCapture(c->stream, &g); if (g) cudaGraphDestroy(g); }
        return rc;
      }
    }
  }
  if (use_graph) {
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    OBN_CUDA(cudaStreamEndCapture(c->stream, &graph));
    cudaError_t ge = cudaGraphInstantiate(&exec, graph, 0);
    if (ge != cudaSuccess) {
      cudaGraphDestroy(graph);
      return fail(OBN_ERR_CUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(ge));
    }
    lap("graph build");
    OBN_CUDA(cudaEventRecord(e0, c->stream));
    ge = cudaGraphLaunch(exec, c->stream);
    if (ge == cudaSuccess) ge = cudaEventRecord(e1, c->stream);
    if (ge == cudaSuccess) ge = cudaStreamSynchronize(c->stream);
    cudaGraphExecDestroy(exec);
    cudaGraphDestroy(graph);
    if (ge != cudaSuccess) return fail(OBN_ERR_CUDA, "lattice graph launch failed: %s", cudaGetErrorString(ge));
  } else {
    OBN_CUDA(cudaEventRecord(e1, c->stream));
    OBN_CUDA(cudaStreamSynchronize(c->stream));
  }
  lap("blocks");
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  c->prof.count_kernel_ms += ms;
  // level lists for the DP: subsets grouped by size, ascending inside a level (= Gosper order)
  {
    if (int rc = c->es_pool.d_masks.reserve(combs * sizeof(uint32_t))) return rc;
    EsLevels lv{};
    uint64_t off = 0, binom = 1;
    for (int k = 0; k <= n; ++k) {
      if (k) binom = binom * (uint64_t)(n - k + 1) / (uint64_t)k;
      lv.off[k] = off;
      off += binom;
    }
    lv.off[n + 1] = off;
    e->levels = lv;
    (void)cudaGetLastError();
    es_level_lists_kernel<<<(unsigned)((combs + 255) / 256), 256, 0, c->stream>>>(combs, n, lv, c->es_pool.d_masks.as<uint32_t>());
    OBN_CUDA(cudaGetLastError());
    OBN_CUDA(cudaStreamSynchronize(c->stream));
    ++c->prof.total_launches;
  }
  e->use_F = true;
  e->f_per_rank = (size_t)per * NT;
  return OBN_OK;
}

int obn_es_f_buffer(obn_es* e, void** f_dev, size_t* doubles_per_rank) {
  if (!e || !f_dev || !doubles_per_rank) return fail(OBN_ERR_INVALID, "obn_es_f_buffer: NULL argument");
  *f_dev = e->F;
  *doubles_per_rank = e->f_per_rank;
  return OBN_OK;
}

}  // extern "C"
