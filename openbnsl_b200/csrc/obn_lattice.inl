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
  int bot = 0, Lm = 0;                    // the last `bot` low variables live in shared memory (0 or kBotL); Lm = L - bot
  uint64_t cells_bot = 1;                 // prod cards of the bottom variables (a factor of every mid-tier unit)
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

static bool plan_lattice(const std::vector<uint32_t>& card, size_t budget_bytes, int world, bool allow_bottom, int max_L_opt, LatticePlan& P) {
  const int n = (int)card.size();
  P.n = n;
  P.card = card;
  // bottom tier: the last kBotL variables, when they are all binary
  int bot = (allow_bottom && n >= kBotL) ? kBotL : 0;
  for (int v = n - bot; v < n; ++v)
    if (card[v] != 2) bot = 0;
  const int max_L = std::min(n, std::max(1, max_L_opt));
  if (max_L < bot) bot = 0;
  const double cells_bot = bot ? (double)kBotCells : 1.0;
  double joint = 1;
  for (int v = 0; v < n; ++v) joint *= card[v];
  if (joint >= 2147483648.0) return false;                          // dense joint table must fit int32 indexing
  // the largest L whose two largest adjacent levels (times all high cells) fit the budget
  int best_L = -1;
  for (int L = max_L; L >= std::max(1, bot); --L) {
    const int h = n - L;
    const int Lm = L - bot;
    if (h > 16) break;
    if (world > 1 && ((1 << h) % world != 0 || (1 << h) < world)) continue;
    // multi-GPU balance: blocks are dealt whole, so the heaviest block (all high variables
    // present) must be a small share of the total: prod c_high / prod (1 + c_high) <= 1 / (2 world)
    if (world > 1 && L > 1) {
      double heavy = 1, total = 1;
      for (int v = 0; v < h; ++v) { heavy *= card[v]; total *= 1.0 + card[v]; }
      if (heavy / total > 1.0 / (2.0 * world)) continue;
    }
    // units per level by DP over the low (mid-tier) variables: poly = cells_bot * prod (1 + c x)
    std::vector<double> lvl(L + 1, 0.0);
    lvl[0] = cells_bot;
    for (int i = 0; i < Lm; ++i)
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
  const int L = best_L, h = n - L, Lm = L - bot;
  P.L = L;
  P.h = h;
  P.bot = bot;
  P.Lm = Lm;
  P.cells_bot = bot ? kBotCells : 1;
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
  // low lattice over the mid-tier variables h..h+Lm-1; every table also holds all bottom variables
  // (a factor cells_bot on the fast side).  O(1) work per subset.
  const uint32_t NT = 1u << L, NM = 1u << Lm;
  const uint32_t cb = (uint32_t)P.cells_bot;
  std::vector<uint32_t> cl(NM, cb), off(NM, 0);
  for (uint32_t T = 1; T < NM; ++T) cl[T] = cl[T & (T - 1)] * card[h + __builtin_ctz(T)];
  P.level_units.assign(Lm + 1, 0);
  P.level_max_units.assign(Lm + 1, 0);
  std::vector<uint32_t> level_count(Lm + 1, 0);
  for (uint32_t T = 0; T < NM; ++T) {
    const int k = __builtin_popcount(T);
    off[T] = (uint32_t)P.level_units[k];
    P.level_units[k] += cl[T];
    P.level_max_units[k] = std::max(P.level_max_units[k], cl[T]);
    ++level_count[k];
  }
  std::vector<uint32_t> pre(Lm + 1, 1);                             // pre[i] = prod of the first i low cards
  for (int i = 0; i < Lm; ++i) pre[i + 1] = pre[i] * card[h + i];
  P.level_tasks.assign(std::max(Lm, 1), {});
  for (int k = 0; k < Lm; ++k) P.level_tasks[k].resize(level_count[k]);
  std::vector<uint32_t> fill(Lm + 1, 0);
  for (uint32_t T = 0; T + 1 < NM; ++T) {
    const int k = __builtin_popcount(T);
    const int i = __builtin_ctz(~T);                                // lowest absent low variable
    MargTask& t = P.level_tasks[k][fill[k]++];
    t.src_off = off[T | (1u << i)];
    t.dst_off = off[T];
    t.hi = pre[i];                                                  // bits below i are all present in T
    t.r = card[h + i];
    const uint32_t lo = cl[T & ~((2u << i) - 1u)];                  // T's variables above i (and the bottom tier)
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

static int launch_marg_level(obn_ctx* c, cudaStream_t st, const MargTask* d_tasks, size_t n_tasks, uint64_t max_cells, uint64_t mul,
                             const int32_t* src, int32_t* dst, double ess, const TermTable& tt, DevBuf& partials,
                             double* F, uint64_t f_base) {
  const uint32_t chunks = (uint32_t)((max_cells + kMargChunk - 1) / kMargChunk);
  for (size_t b = 0; b < n_tasks; b += 65535) {
    const uint32_t m = (uint32_t)std::min<size_t>(65535, n_tasks - b);
    (void)cudaGetLastError();
    if (F == nullptr) {                                   // tables only: the bottom tier computes F
      marg_kernel<<<dim3(chunks, m), kMargThreads, 0, st>>>(d_tasks + b, mul, src, dst, ess, tt, nullptr);
      OBN_CUDA(cudaGetLastError());
      c->prof.total_launches += 1;
      c->prof.count_launches += 1;
      continue;
    }
    if (int rc = partials.reserve((size_t)m * chunks * sizeof(double))) return rc;
    marg_kernel<<<dim3(chunks, m), kMargThreads, 0, st>>>(d_tasks + b, mul, src, dst, ess, tt, partials.as<double>());
    OBN_CUDA(cudaGetLastError());
    marg_reduce_kernel<<<(m + 127) / 128, 128, 0, st>>>(d_tasks + b, m, chunks, partials.as<double>(), F, f_base);
    OBN_CUDA(cudaGetLastError());
    c->prof.total_launches += 2;
    c->prof.count_launches += 1;
  }
  return OBN_OK;
}

// The per-(level, thread) program of bottom_lattice_kernel (obn_lattice.cuh): which two source quads
// a thread adds and where the result goes.  A quad = the four cells that differ in the two fastest
// digits; quads are indexed by (upper pattern U = which of the other six bottom variables are kept,
// cell of U's table).  Tables are stored by level (popcount of U) descending, patterns ascending;
// the parent of U is U with its highest absent bit i set, and digit i sits above
// sh = popcount(U & (2^i - 1)) kept digits.
struct BottomProgram {
  std::vector<uint32_t> words;     // [kBotUpper][kBotThreads] levels 5..0
  std::vector<uint16_t> patterns;  // [kBotSlots * kBotThreads] U of quad q (phase 2 takes quads tid, tid+256, tid+512)
};
static bool bottom_program(BottomProgram& out) {
  const uint32_t NU = 1u << kBotUpper;
  std::vector<std::vector<uint32_t>> pats(kBotUpper + 1);
  for (uint32_t U = 0; U < NU; ++U) pats[__builtin_popcount(U)].push_back(U);
  std::vector<uint32_t> qoff(NU, 0);
  uint32_t quads = 0;
  for (int j = kBotUpper; j >= 0; --j) {
    for (size_t r = 0; r < pats[j].size(); ++r) qoff[pats[j][r]] = quads + (uint32_t)(r << j);
    quads += (uint32_t)(pats[j].size() << j);
  }
  if (quads != (uint32_t)kBotQuads) return false;
  out.words.assign((size_t)kBotUpper * kBotThreads, kBotIdleWord);
  out.patterns.assign((size_t)kBotSlots * kBotThreads, kBotNoPattern);
  for (int j = kBotUpper; j >= 0; --j) {
    const size_t items = pats[j].size() << j;
    if (items > (size_t)kBotThreads) return false;
    for (size_t e = 0; e < items; ++e) {
      const uint32_t U = pats[j][e >> j], cidx = (uint32_t)(e & (((size_t)1 << j) - 1));
      out.patterns[qoff[U] + cidx] = (uint16_t)U;
      if (j == kBotUpper) continue;                                   // level 6 is the chunk itself
      const int i = 31 - __builtin_clz((~U) & (NU - 1u));
      const uint32_t sh = (uint32_t)__builtin_popcount(U & ((1u << i) - 1u));
      const uint32_t src0 = qoff[U | (1u << i)] + ((cidx >> sh) << (sh + 1)) + (cidx & ((1u << sh) - 1u));
      out.words[(size_t)(kBotUpper - 1 - j) * kBotThreads + e] = bot_word(src0, sh, qoff[U] + cidx);
    }
  }
  return true;
}

// F of all 2^8 bottom patterns of every table of one level (the tables are the tasks' `dst`)
static int launch_bottom_level(obn_ctx* c, cudaStream_t st, const MargTask* d_tasks, size_t n_tasks, uint64_t max_cells, uint64_t mul,
                               const int32_t* base, const uint32_t* d_prog, const FixTable& ft, double ess, double* F,
                               uint64_t f_base, int f_shift) {
  const uint16_t* d_pat = reinterpret_cast<const uint16_t*>(d_prog + (size_t)kBotUpper * kBotThreads);
  const uint64_t chunks = max_cells >> kBotL;
  // chunks per CTA: a CTA costs about 1.3 chunks of fixed work (program copy, term-table lookup,
  // the flush of its twelve sums), so one-chunk CTAs more than double the cost of a small table.
  // Pick the power of two that minimises waves x (chunks per CTA + fixed cost), 3 CTAs per SM.
  const double resident = (double)c->sm_count * 3.0;
  uint64_t cpc = 1;
  double best_cost = 1e300;
  for (uint64_t cand = 1; cand <= 64 && cand <= std::max<uint64_t>(chunks, 1); cand <<= 1) {
    const double ctas = (double)n_tasks * (double)((chunks + cand - 1) / cand);
    const double cost = std::ceil(ctas / resident) * ((double)cand + 1.3);
    if (cost < best_cost * 0.999) {
      best_cost = cost;
      cpc = cand;
    }
  }
  const uint32_t groups = (uint32_t)((chunks + cpc - 1) / cpc);
  for (size_t b = 0; b < n_tasks; b += 65535) {
    const uint32_t m = (uint32_t)std::min<size_t>(65535, n_tasks - b);
    (void)cudaGetLastError();
    auto* Fx = reinterpret_cast<unsigned long long*>(F);
    if (ft.lut)
      bottom_lattice_kernel<true><<<dim3(groups, m), kBotThreads, 0, st>>>(d_tasks + b, mul, base, d_prog, d_pat, ft, ess, Fx,
                                                                                 f_base, f_shift, (uint32_t)cpc);
    else
      bottom_lattice_kernel<false><<<dim3(groups, m), kBotThreads, 0, st>>>(d_tasks + b, mul, base, d_prog, d_pat, ft, ess, Fx,
                                                                                  f_base, f_shift, (uint32_t)cpc);
    OBN_CUDA(cudaGetLastError());
    c->prof.total_launches += 1;
    c->prof.count_launches += 1;
  }
  return OBN_OK;
}

// The joint table over all variables: the only pass over the data in lattice mode.
static int joint_count(obn_ctx* c, const obn_dataset* ds, const std::vector<uint32_t>& order, int32_t* J, uint64_t cells) {
  OBN_CUDA(cudaMemsetAsync(J, 0, cells * 4, c->stream));
  JointCards jc{};
  for (size_t i = 0; i < ds->n_vars; ++i) {
    jc.v[i] = order[i];
    jc.c[i] = ds->cards[order[i]];
  }
  const uint32_t n_vec = (uint32_t)((ds->n_rows + 15) >> 4);
  (void)cudaGetLastError();
  joint_count_kernel<<<std::max(1u, (n_vec + 127) / 128), 128, 0, c->stream>>>(ds->data, ds->stride, (uint32_t)ds->n_rows, (int)ds->n_vars, jc, J);
  OBN_CUDA(cudaGetLastError());
  ++c->prof.total_launches;
  ++c->prof.count_launches;
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
  const size_t pooled = c->es_pool.d_root.cap + c->es_pool.d_pp[0].cap + c->es_pool.d_pp[1].cap + c->es_pool.d_pp[2].cap +
                        c->es_pool.d_pp[3].cap;   // reusable
  // The plan (tier split, F layout, all-gather slice) follows from the budget, so every rank must plan
  // from the SAME number: the minimum over the ranks (one small all-reduce); without a communicator a
  // multi-rank caller gets the fixed cap alone, independent of what happens to be free on each GPU.
  uint64_t budget64 = (uint64_t)c->tune.lattice_budget_gb << 30;
  if (world == 1 || c->comm) budget64 = std::min<uint64_t>(budget64, (uint64_t)((free_b + pooled) * 0.7));
  if (world > 1 && c->comm)
    if (int rc = comm_min_u64_host(c, &budget64)) return rc;
  const size_t budget = (size_t)budget64;
  const bool dbg = c->tune.debug != 0;
  auto t_dbg = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (!dbg) return;
    cudaStreamSynchronize(c->stream);
    const auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "[obn_es_lattice] %-14s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_dbg).count());
    t_dbg = now;
  };
  const bool use_graph = c->tune.lattice_graph != 0;
  const size_t combs = (size_t)1 << n;
  EsPool& pool = c->es_pool;
  LatticeGraphCache& G = pool.lattice;
  cudaEvent_t e0 = c->ev0, e1 = c->ev1;
  // what the captured graph depends on: the search shape and every pooled buffer it points into
  auto make_key = [&]() {
    std::vector<uint64_t> k;
    uint64_t ess_bits = 0;
    memcpy(&ess_bits, &e->ess, 8);
    k.insert(k.end(), {(uint64_t)n, (uint64_t)ds->n_rows, ess_bits, (uint64_t)rank, (uint64_t)world,
                       (uint64_t)budget, (uint64_t)c->tune.lattice_bottom,
                       (uint64_t)c->tune.lattice_max_l, (uint64_t)c->tune.lattice_no_lut,
                       (uint64_t)c->tune.lattice_reorder, (uint64_t)c->tune.lattice_lanes});
    for (int v = 0; v < n; ++v) k.push_back(ds->cards[v]);
    for (const DevBuf* d : {&pool.d_F, &pool.d_root, &pool.d_pp[0], &pool.d_pp[1], &pool.d_pp[2], &pool.d_pp[3], &pool.d_tasks, &pool.d_partials, &pool.d_lut_a, &pool.d_lutF, &pool.d_prog})
      k.push_back((uint64_t)(uintptr_t)d->p);
    return k;
  };
  // Digit order of the joint table.  The bottom tier needs the 8 fastest digits binary: when the last
  // 8 variables are not all binary but the data set has 8 binary variables, the 8 highest-numbered
  // binary variables become the fastest digits (everything else keeps its relative order).  Subset
  // masks are translated to digit order by three small tables (es_f_index in obn_es.inl).
  std::vector<uint32_t> order(n);
  for (int v = 0; v < n; ++v) order[v] = (uint32_t)v;
  if (c->tune.lattice_bottom != 0 && c->tune.lattice_reorder != 0 && n >= kBotL) {
    bool tail_binary = true;
    for (int v = n - kBotL; v < n; ++v) tail_binary &= ds->cards[v] == 2;
    std::vector<uint32_t> bin;
    for (int v = n - 1; v >= 0 && bin.size() < (size_t)kBotL; --v)
      if (ds->cards[v] == 2) bin.push_back((uint32_t)v);
    if (!tail_binary && bin.size() == (size_t)kBotL) {
      std::sort(bin.begin(), bin.end());
      order.clear();
      for (int v = 0; v < n; ++v)
        if (!std::binary_search(bin.begin(), bin.end(), (uint32_t)v)) order.push_back((uint32_t)v);
      order.insert(order.end(), bin.begin(), bin.end());
    }
  }
  std::vector<uint32_t> cards_p(n);
  for (int i = 0; i < n; ++i) cards_p[i] = ds->cards[order[i]];
  const bool hit = use_graph && G.exec && G.key == make_key();
  if (dbg) fprintf(stderr, "[obn_es_lattice] graph cache %s\n", hit ? "hit" : "miss");
  if (!hit) {
    G.clear();
    LatticePlan P;
    if (!plan_lattice(cards_p, budget, world, c->tune.lattice_bottom != 0, c->tune.lattice_max_l, P))
      return fail(OBN_ERR_UNSUPPORTED, "marginalisation lattice does not fit (joint table >= 2^31 cells or too little memory)");
    lap("plan");
    const int h = P.h, L = P.L, Lm = P.Lm;
    const bool bot = P.bot != 0;
    const uint32_t B = 1u << h, NT = 1u << L;
    if (dbg) fprintf(stderr, "[obn_es_lattice] n=%d h=%d L=%d bottom=%d budget=%.1f GB\n", n, h, L, P.bot, budget / 1073741824.0);

    // ---- device buffers (all reserved before the capture starts) ----------------------------
    uint64_t cells_all_high = P.cells_high[B - 1], buf_units = 0;
    for (int k = 0; k < Lm; ++k) buf_units = std::max<uint64_t>(buf_units, P.level_units[k]);
    if (int rc = pool.d_F.reserve(combs * sizeof(double))) return rc;
    if (int rc = pool.d_root.reserve(std::max<uint64_t>(P.root_cells, 4) * 4)) return rc;
    // Blocks are independent once their roots exist: with the bottom tier they alternate between two
    // stream lanes (own ping-pong pair each), so the small launches at the ends of one block's levels
    // overlap with the other lane's (matters when there are many small blocks, i.e. on several GPUs).
    const uint64_t pp_bytes = std::max<uint64_t>(buf_units * cells_all_high, 4) * 4;
    const int lanes = (bot && c->tune.lattice_lanes >= 2 && B / (uint32_t)world >= 2 &&
                       4 * pp_bytes + P.root_cells * 4 <= budget) ? 2 : 1;
    for (int i = 0; i < 2 * lanes; ++i)
      if (int rc = pool.d_pp[i].reserve(pp_bytes)) return rc;
    if (!bot) {
      size_t need = (size_t)((P.cells_high[B - 1] * P.cells_low + kMargChunk - 1) / kMargChunk);
      for (int k = 0; k < h; ++k)
        need = std::max(need, std::min<size_t>(65535, P.root_tasks[k].size()) *
                                  (size_t)(((uint64_t)P.root_level_max[k] * P.cells_low + kMargChunk - 1) / kMargChunk));
      for (int k = 0; k < Lm; ++k)
        need = std::max(need, std::min<size_t>(65535, P.level_tasks[k].size()) *
                                  (size_t)(((uint64_t)P.level_max_units[k] * cells_all_high + kMargChunk - 1) / kMargChunk));
      if (int rc = pool.d_partials.reserve(need * sizeof(double))) return rc;
    }
    lap("alloc");
    // tasks: low levels then root levels, one upload
    size_t n_all = 1;
    std::vector<size_t> low_at(Lm + 1, 0), root_at(h + 1, 0);
    for (int k = 0; k < Lm; ++k) { low_at[k] = n_all; n_all += P.level_tasks[k].size(); }
    for (int k = 0; k < h; ++k) { root_at[k] = n_all; n_all += P.root_tasks[k].size(); }
    const size_t joint_task_at = 0;
    if (int rc = pool.d_tasks.reserve(n_all * sizeof(MargTask))) return rc;
    if (int rc = c->st[0].h_in.reserve(n_all * sizeof(MargTask))) return rc;               // page-locked staging
    {
      MargTask* all = c->st[0].h_in.as<MargTask>();
      MargTask t{};
      t.f_index = P.slot[B - 1] * NT + (NT - 1);
      t.lo_shift = -1;
      if (bot) {
        // bottom tier: task 0 describes a block's root table R_H (base = R_H, mul = cells of H)
        t.hi = 1;
        t.lo = (uint32_t)P.cells_low;
        t.f_index = (1u << Lm) - 1u;
      }
      all[0] = t;
      for (int k = 0; k < Lm; ++k)
        if (!P.level_tasks[k].empty()) memcpy(all + low_at[k], P.level_tasks[k].data(), P.level_tasks[k].size() * sizeof(MargTask));
      for (int k = 0; k < h; ++k)
        if (!P.root_tasks[k].empty()) memcpy(all + root_at[k], P.root_tasks[k].data(), P.root_tasks[k].size() * sizeof(MargTask));
    }
    OBN_CUDA(cudaMemcpyAsync(pool.d_tasks.p, c->st[0].h_in.p, n_all * sizeof(MargTask), cudaMemcpyHostToDevice, c->stream));
    OBN_CUDA(cudaStreamSynchronize(c->stream));                                        // h_in is reused by run_batch
    if (int rc = pool.d_slot.reserve(B * 4)) return rc;
    OBN_CUDA(cudaMemcpyAsync(pool.d_slot.p, P.slot.data(), B * 4, cudaMemcpyHostToDevice, c->stream));
    OBN_CUDA(cudaStreamSynchronize(c->stream));                                        // P.slot is a local
    if (bot) {
      BottomProgram bp;
      if (!bottom_program(bp)) return fail(OBN_ERR_UNSUPPORTED, "bottom program does not match the kernel's level table");
      if (int rc = pool.d_prog.reserve(bp.words.size() * 4 + bp.patterns.size() * 2)) return rc;
      OBN_CUDA(cudaMemcpy(pool.d_prog.p, bp.words.data(), bp.words.size() * 4, cudaMemcpyHostToDevice));
      OBN_CUDA(cudaMemcpy(pool.d_prog.as<char>() + bp.words.size() * 4, bp.patterns.data(), bp.patterns.size() * 2, cudaMemcpyHostToDevice));
    }
    bool permuted = false;
    for (int i = 0; i < n; ++i) permuted |= order[i] != (uint32_t)i;
    if (permuted) {
      // mask translation tables: bit v of a subset mask -> bit (digit of variable v)
      std::vector<uint32_t> digit_of(n), tab(3 * 1024, 0);
      for (int i = 0; i < n; ++i) digit_of[order[i]] = (uint32_t)i;
      for (int t = 0; t < 3; ++t)
        for (uint32_t m = 0; m < 1024; ++m) {
          uint32_t out = 0;
          for (int b = 0; b < 10; ++b)
            if ((m >> b & 1u) && t * 10 + b < n) out |= 1u << digit_of[t * 10 + b];
          tab[t * 1024 + m] = out;
        }
      if (int rc = pool.d_perm.reserve(tab.size() * 4)) return rc;
      OBN_CUDA(cudaMemcpy(pool.d_perm.p, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice));
    }
    lap("tasks upload");

    // ---- term tables: one per distinct cell count (device libm; this is the fast mode) -------
    TermTable tt{nullptr, nullptr, 0, 0};
    FixTable ft{nullptr, nullptr, 0, 0, 1.0};
    if (bot) {
      // fixed-point scale: |F| <= N * (ln(N + ess) + max |lgamma(a)| + 1) must stay below 2^62
      double joint_cells = 1;
      for (int v = 0; v < n; ++v) joint_cells *= ds->cards[v];
      const double lg = std::max(std::fabs(std::lgamma(e->ess / joint_cells)), std::fabs(std::lgamma(e->ess)));
      const double bound = (double)ds->n_rows * (std::log((double)ds->n_rows + e->ess + 1.0) + lg + 1.0) + 1.0;
      const int shift = std::max(0, std::min(40, 62 - (int)std::ceil(std::log2(bound))));
      ft.scale = std::ldexp(1.0, shift);
    }
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
      if (distinct.size() <= 256 && distinct.size() * stride * 8 <= ((size_t)2 << 30) && !c->tune.lattice_no_lut) {
        std::vector<double> a(distinct.size());
        for (size_t i = 0; i < distinct.size(); ++i) a[i] = e->ess / (double)(size_t)distinct[i];
        if (int rc = pool.d_lut_a.reserve(a.size() * 8)) return rc;
        if (int rc = pool.d_lutF.reserve(a.size() * stride * 8)) return rc;
        OBN_CUDA(cudaMemcpyAsync(pool.d_lut_a.p, a.data(), a.size() * 8, cudaMemcpyHostToDevice, c->stream));
        OBN_CUDA(cudaStreamSynchronize(c->stream));          // `a` leaves scope
        (void)cudaGetLastError();
        const dim3 grid((unsigned)((stride + 255) / 256), (unsigned)a.size());
        if (bot) {
          lut_fill_fix_kernel<<<grid, 256, 0, c->stream>>>(pool.d_lutF.as<long long>(), pool.d_lut_a.as<double>(), stride, ft.scale);
          ft.lut = pool.d_lutF.as<long long>();
          ft.lut_a = pool.d_lut_a.as<double>();
          ft.n_lut = (int)a.size();
          ft.stride = stride;
        } else {
          lut_fill_kernel<<<grid, 256, 0, c->stream>>>(pool.d_lutF.as<double>(), pool.d_lut_a.as<double>(), stride);
          tt = TermTable{pool.d_lutF.as<double>(), pool.d_lut_a.as<double>(), (int)a.size(), stride};
        }
        OBN_CUDA(cudaGetLastError());
      }
    }
    lap("term tables");

    // ---- the lattice itself: thousands of small dependent launches (blocks x levels x 2), captured
    // into ONE CUDA graph so that the GPU never waits for the host thread (which shares its cores
    // with the other ranks' processes) ------------------------------------------------------------
    int32_t* root = pool.d_root.as<int32_t>();
    double* F = pool.d_F.as<double>();
    const MargTask* d_tasks = pool.d_tasks.as<MargTask>();
    const uint64_t launches_before = c->prof.total_launches;
    OBN_CUDA(cudaStreamSynchronize(c->stream));
    if (use_graph) OBN_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    else {
      // eager path (debugging): the joint table has to exist before the first launch
      if (int rc = joint_count(c, ds, order, root + P.root_off[B - 1], P.cells_high[B - 1] * P.cells_low)) return rc;
      OBN_CUDA(cudaEventRecord(e0, c->stream));
    }
    auto abort_capture = [&]() {
      if (!use_graph) return;
      cudaGraph_t g = nullptr;
      cudaStreamEndCapture(c->stream, &g);
      if (g) cudaGraphDestroy(g);
    };
    auto lap_if = [&](const char* what) { if (!use_graph) lap(what); };
    const uint32_t per = B / (uint32_t)world;
    const size_t slice = (size_t)per * NT;                       // this rank's F words
    const uint32_t* d_prog = pool.d_prog.as<uint32_t>();
    if (bot) {
      if (cudaMemsetAsync(F + (size_t)rank * slice, 0, slice * 8, c->stream) != cudaSuccess) {
        abort_capture();
        return fail(OBN_ERR_CUDA, "F memset failed");
      }
    } else {
      const uint64_t cells = P.cells_high[B - 1] * P.cells_low;
      const uint32_t chunks = (uint32_t)((cells + kMargChunk - 1) / kMargChunk);
      (void)cudaGetLastError();
      table_f_kernel<<<chunks, kMargThreads, 0, c->stream>>>(root + P.root_off[B - 1], cells, e->ess, tt, pool.d_partials.as<double>());
      marg_reduce_kernel<<<1, 128, 0, c->stream>>>(d_tasks + joint_task_at, 1, chunks, pool.d_partials.as<double>(), F, 0);
      if (cudaGetLastError() != cudaSuccess) { abort_capture(); return fail(OBN_ERR_CUDA, "joint table launch failed"); }
      c->prof.total_launches += 2;
    }
    lap_if("joint");
    // root lattice
    for (int k = h - 1; k >= 0; --k)
      if (int rc = launch_marg_level(c, c->stream, d_tasks + root_at[k], P.root_tasks[k].size(), (uint64_t)P.root_level_max[k] * P.cells_low,
                                     1, root, root, e->ess, tt, pool.d_partials, bot ? nullptr : F, 0)) {
        abort_capture();
        return rc;
      }
    lap_if("roots");
    // blocks of this rank
    if (lanes > 1) {                                            // fork the second lane
      if (cudaEventRecord(c->ev_fork, c->stream) != cudaSuccess || cudaStreamWaitEvent(c->aux[0], c->ev_fork, 0) != cudaSuccess) {
        abort_capture();
        return fail(OBN_ERR_CUDA, "lattice lane fork failed");
      }
    }
    uint32_t n_mine = 0;
    for (uint32_t H = 0; H < B; ++H) {
      const uint32_t s = P.slot[H];
      if (s / per != (uint32_t)rank) continue;
      const int lane = (int)(n_mine++ % (uint32_t)lanes);
      cudaStream_t st = lane ? c->aux[0] : c->stream;
      DevBuf* pp = &pool.d_pp[2 * lane];
      const uint64_t mul = P.cells_high[H];
      if (bot)                                                  // F of R_H and of its 255 bottom sub-patterns
        if (int rc = launch_bottom_level(c, st, d_tasks, 1, mul * P.cells_low, mul, root + P.root_off[H], d_prog, ft, e->ess, F,
                                         (uint64_t)s * NT, Lm)) {
          abort_capture();
          return rc;
        }
      for (int k = Lm - 1; k >= 0; --k) {
        const int32_t* src = (k == Lm - 1) ? root + P.root_off[H] : pp[(k + 1) & 1].as<int32_t>();
        int rc = launch_marg_level(c, st, d_tasks + low_at[k], P.level_tasks[k].size(), (uint64_t)P.level_max_units[k] * mul, mul, src,
                                   pp[k & 1].as<int32_t>(), e->ess, tt, pool.d_partials, bot ? nullptr : F, (uint64_t)s * NT);
        if (rc == OBN_OK && bot)
          rc = launch_bottom_level(c, st, d_tasks + low_at[k], P.level_tasks[k].size(), (uint64_t)P.level_max_units[k] * mul, mul,
                                   pp[k & 1].as<int32_t>(), d_prog, ft, e->ess, F, (uint64_t)s * NT, Lm);
        if (rc != OBN_OK) {
          abort_capture();
          return rc;
        }
      }
    }
    if (lanes > 1) {                                            // join
      if (cudaEventRecord(c->ev_join[0], c->aux[0]) != cudaSuccess || cudaStreamWaitEvent(c->stream, c->ev_join[0], 0) != cudaSuccess) {
        abort_capture();
        return fail(OBN_ERR_CUDA, "lattice lane join failed");
      }
    }
    if (bot) {
      (void)cudaGetLastError();
      fix_to_double_kernel<<<(unsigned)((slice + 255) / 256), 256, 0, c->stream>>>(F + (size_t)rank * slice, slice, 1.0 / ft.scale);
      if (cudaGetLastError() != cudaSuccess) { abort_capture(); return fail(OBN_ERR_CUDA, "fix_to_double launch failed"); }
      ++c->prof.total_launches;
    }
    G.h = h;
    G.L = L;
    G.per = per;
    G.order = order;
    G.permuted = permuted;
    G.joint_off = P.root_off[B - 1];
    G.joint_cells = P.cells_high[B - 1] * P.cells_low;
    G.launches = c->prof.total_launches - launches_before;
    if (use_graph) {
      cudaGraph_t graph = nullptr;
      OBN_CUDA(cudaStreamEndCapture(c->stream, &graph));
      const cudaError_t ge = cudaGraphInstantiate(&G.exec, graph, 0);
      cudaGraphDestroy(graph);
      if (ge != cudaSuccess) {
        G.exec = nullptr;
        return fail(OBN_ERR_CUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(ge));
      }
      G.key = make_key();
      lap("graph build");
    } else {
      OBN_CUDA(cudaEventRecord(e1, c->stream));
      OBN_CUDA(cudaStreamSynchronize(c->stream));
    }
  } else {
    c->prof.total_launches += G.launches;
  }
  e->F = pool.d_F.as<double>();
  e->h = G.h;
  e->L = G.L;
  e->perm = G.permuted ? pool.d_perm.as<uint32_t>() : nullptr;
  {
    // accounting: what the reference would have read for the subsets whose scores are used:
    // sum over k <= max_parents+1 of C(n,k) * k * N bytes
    double alg = 0, binom = 1;
    for (int k = 1; k <= n && (size_t)k <= e->max_parents + 1; ++k) {
      binom = binom * (double)(n - k + 1) / (double)k;
      alg += binom * (double)k * (double)ds->n_rows;
    }
    c->prof.algorithmic_bytes += alg / world;
  }
  if (use_graph) {
    // the joint table is the only pass over the data; then the whole lattice is one graph launch
    if (int rc = joint_count(c, ds, G.order, pool.d_root.as<int32_t>() + G.joint_off, G.joint_cells)) return rc;
    OBN_CUDA(cudaEventRecord(e0, c->stream));
    cudaError_t ge = cudaGraphLaunch(G.exec, c->stream);
    if (ge == cudaSuccess) ge = cudaEventRecord(e1, c->stream);
    if (ge == cudaSuccess) ge = cudaStreamSynchronize(c->stream);
    if (ge != cudaSuccess) {
      G.clear();
      return fail(OBN_ERR_CUDA, "lattice graph launch failed: %s", cudaGetErrorString(ge));
    }
  }
  lap("blocks");
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  c->prof.count_kernel_ms += ms;
  // level lists for the DP: subsets grouped by size, ascending inside a level (= Gosper order)
  if (int rc = es_ensure_lists(e)) return rc;
  e->use_F = true;
  e->f_per_rank = (size_t)G.per << G.L;
  return OBN_OK;
}

int obn_lattice_bottom_program(uint32_t* words, size_t n_words, uint16_t* patterns, size_t n_patterns, int* geometry) {
  if (!geometry) return fail(OBN_ERR_INVALID, "obn_lattice_bottom_program: NULL argument");
  geometry[0] = kBotL;
  geometry[1] = kBotThreads;
  geometry[2] = kBotQuads;
  geometry[3] = kBotTabQuads;
  geometry[4] = kBotUpper;
  geometry[5] = kBotSlots;
  BottomProgram bp;
  if (!bottom_program(bp)) return fail(OBN_ERR_UNSUPPORTED, "bottom program does not match the kernel's level table");
  if ((words && n_words < bp.words.size()) || (patterns && n_patterns < bp.patterns.size()))
    return fail(OBN_ERR_INVALID, "obn_lattice_bottom_program: buffers too small (%zu words, %zu patterns)", bp.words.size(),
                bp.patterns.size());
  if (words) std::memcpy(words, bp.words.data(), bp.words.size() * 4);
  if (patterns) std::memcpy(patterns, bp.patterns.data(), bp.patterns.size() * 2);
  return OBN_OK;
}

int obn_es_f_buffer(obn_es* e, void** f_dev, size_t* doubles_per_rank) {
  if (!e || !f_dev || !doubles_per_rank) return fail(OBN_ERR_INVALID, "obn_es_f_buffer: NULL argument");
  *f_dev = e->F;
  *doubles_per_rank = e->f_per_rank;
  return OBN_OK;
}

}  // extern "C"
