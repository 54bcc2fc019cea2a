// obn_batch.inl -- host-side batch dispatcher: validates a CSR tuple batch,
// sorts the tuples into kernel tiers by table size, uploads ONE packed input
// buffer, launches one counting kernel per tier, downloads ONE packed output.
// Included at the end of obn_capi.cu.

namespace obn {

struct BatchSpec {
  obn_ctx* c;
  const obn_dataset* ds;
  const uint32_t* var_ids;
  const uint64_t* offsets;
  size_t n;
  int mode;
  const uint32_t* role_a = nullptr;
  const uint32_t* role_b = nullptr;
  int kind = 0;
  double ess = 1.0;
  int exact = 0;
  // host outputs
  double* stat = nullptr;
  uint64_t* dof = nullptr;
  double* p = nullptr;
  double* scores = nullptr;
  double* scores_dev = nullptr;   // BDEU_ALL: leave the scores on the device at this address (no D2H)
  int32_t* counts = nullptr;
  int32_t* counts_dev = nullptr;  // COUNTS: leave the dense tables on the device at this address (no D2H)
  const uint64_t* counts_off = nullptr;
  // MODE_CI: the tuples are GROUPS (obn_ci_groups.inl: the tests asked over one variable set share
  // its table), the outputs are per TEST
  size_t n_out = 0;                 // number of tests
  const uint32_t* mem_off = nullptr;    // [n + 1]
  const uint32_t* members = nullptr;    // 2 words per member
  // PC round (obn_pc_round): decision on the device, first[slot] = first passing test of the slot
  const uint32_t* slot_of_test = nullptr;   // [n_out]
  size_t n_slots = 0;
  double level = 0.05;
  int32_t* first = nullptr;         // host, [n_slots]
  bool shard = false;               // split the groups over the ranks of the ctx's communicator + all-reduce(min)
  // CI modes: return once everything is enqueued (obn_ci_tests_submit / obn_pc_round_submit);
  // the matching collect waits and hands the results out
  bool defer = false;
  // PC rounds: skip the tests of edges the round submitted before this one decided (BatchArgs::prev_first)
  bool chain = false;
};

// Table-size classes.  One launch covers the tuples of one class, so that the
// dynamic shared memory of the launch (sized by the class's largest table) is
// within 4x of what each tuple needs and occupancy stays high.
//   class  cells        kernel
//   0      <= 64        <8,  SMEM, scale4>   four 8-bit keys per register, key = byte offset
//   1      <= 256       <8,  SMEM, plain>
//   2..4   <= 1Ki/4Ki/16Ki  <16, SMEM, scale4>
//   5      <= smem      <32, SMEM, plain>
//   6      <= 1.5 Mi    <32, CHUNK, plain>   cell range cut into shared-memory chunks, merged into global scratch
//   7      larger       <32, GLOBAL, plain>  table in L2-resident global scratch
static const int kNumClasses = 8;
static const int kClassSmemMax = 5, kClassChunk = 6, kClassGlobal = 7;
static const uint32_t kClassCells[5] = {64, 256, 1024, 4096, 16384};
// class 6 (TIER_CHUNK): the table is cut into chunks of this many cells (176 KB of shared memory), at most kMaxChunks of them
static const uint32_t kChunkCells = 44 * 1024;
static const uint32_t kMaxChunks = 32;
// log2 of the shared-memory table copies per class (0: none): 32 / 32 / 16 / 4 copies keep the copies of one
// table within 8 / 32 / 64 / 64 KB (A/B: profiles/r2_ab_replication.jsonl)
static constexpr int kRepShift[5] = {5, 5, 4, 2, 0};

static size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

template <int KW, int TIER, bool SCALE4, int RSH = -1>
static cudaError_t launch_count(const BatchArgs& a, unsigned grid, int nthr, size_t smem, cudaStream_t s) {
  // cudaFuncAttributeMaxDynamicSharedMemorySize is per device and a process may hold one ctx per
  // device: set it whenever a launch needs more than the default 48 KB (a few microseconds)
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(count_kernel<KW, TIER, SCALE4, RSH>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return e;
  }
  (void)cudaGetLastError();      // a launch reports only its own error, not a stale one
  count_kernel<KW, TIER, SCALE4, RSH><<<grid, nthr, smem, s>>>(a);
  return cudaGetLastError();
}

static const size_t kScratchMaxBytes = (size_t)8 << 30;

// Results of a finished batch: staging buffer -> the caller's arrays, profile counters.
static void finish_batch(obn_ctx* c, const obn_ctx::Staging& G, double* stat, uint64_t* dof, double* p, int32_t* first) {
  const obn_ctx::Pending& P = G.pending;
  const unsigned char* hout = G.h_out.as<unsigned char>();
  unsigned long long skipped[4] = {0, 0, 0, 0};
  if (P.n_slots) {
    if (first) {
      const int32_t* f = reinterpret_cast<const int32_t*>(hout);
      for (size_t i = 0; i < P.n_slots; ++i) first[i] = f[i] >= kFirstNone ? INT32_MAX : f[i];
    }
    // a chained round: what the device left out (tests of edges the previous round decided) was not evaluated
    if (P.chained) memcpy(skipped, hout + ((P.n_slots * 4 + 7) & ~(size_t)7), sizeof(skipped));
  } else {
    const size_t n = P.n_results;
    if (stat) memcpy(stat, hout, n * 8);
    if (dof) memcpy(dof, hout + n * 8, n * 8);
    if (p) memcpy(p, hout + n * 16, n * 8);
  }
  if (P.timed) {
    float ms = 0;
    cudaEventElapsedTime(&ms, G.e0, G.e1);
    c->prof.count_kernel_ms += ms;
  }
  c->prof.h2d_bytes += (double)P.in_bytes;
  c->prof.d2h_bytes += (double)P.out_bytes;
  c->prof.algorithmic_bytes += P.alg_bytes - (double)skipped[1] * P.n_rows;
  c->prof.scanned_bytes += P.scanned_bytes - (double)skipped[3] * P.n_rows;
  c->prof.increments += P.increments - (double)skipped[2] * P.n_rows;
  c->prof.tuples += P.n_out - (size_t)skipped[0];
  c->prof.tables += P.tables - (size_t)skipped[2];
}

static int run_batch(BatchSpec& S) {
  obn_ctx* c = S.c;
  const obn_dataset* ds = S.ds;
  if (!c || !ds) return fail(OBN_ERR_INVALID, "ctx or dataset is NULL");
  if (ds->ctx != c) return fail(OBN_ERR_INVALID, "dataset belongs to another ctx");
  if (S.defer ? c->q_count == obn_ctx::kStagings : c->q_count != 0)
    return fail(OBN_ERR_INVALID, S.defer ? "two submitted batches are waiting to be collected"
                                         : "a submitted batch is waiting to be collected");
  const int g_idx = (c->q_head + c->q_count) % obn_ctx::kStagings;
  obn_ctx::Staging& G = c->st[g_idx];
  const size_t n = S.n;
  const bool ci = (S.mode == MODE_CI);
  const bool pc_round = ci && S.n_slots != 0;
  const size_t n_out = ci ? S.n_out : n;
  // S.shard (PC rounds): the planner already kept this rank's share of the groups (plan_ci_groups); what is
  // left to do here is the all-reduce of the per-edge minima after the kernels
  const bool merge_ranks = S.shard && c->comm && c->comm_world > 1;
  if (n == 0 && !merge_ranks) {
    if (pc_round && S.first) std::fill(S.first, S.first + S.n_slots, INT32_MAX);
    return OBN_OK;
  }
  if (!S.offsets) return fail(OBN_ERR_INVALID, "offsets is NULL");
  if (n >= 0xffffffffull) return fail(OBN_ERR_UNSUPPORTED, "more than 2^32-2 tuples in one batch");
  const size_t total_vars = S.offsets[n];
  if (total_vars && !S.var_ids) return fail(OBN_ERR_INVALID, "var_ids is NULL");
  if ((S.mode == MODE_BDEU_ONE) && !S.role_a) return fail(OBN_ERR_INVALID, "child is NULL");
  OBN_CUDA(cudaSetDevice(c->device));
  const auto t_begin = std::chrono::steady_clock::now();

  // ---- validate + classify --------------------------------------------------
  // (CI groups were validated test by test in plan_ci_groups)
  const size_t budget = c->smem_optin > 4096 ? c->smem_optin - 2048 : 46 * 1024;
  static thread_local std::vector<uint32_t> cells, rxy, perm;
  static thread_local std::vector<uint8_t> group;
  cells.resize(n);
  group.resize(n);
  rxy.assign(n, 0);
  size_t group_count[kNumClasses] = {};
  double alg_bytes = 0, scanned_bytes = 0, increments = 0;
  size_t n_mine = 0, tests_mine = 0;
  if (ds->n_vars > (1u << 24)) return fail(OBN_ERR_UNSUPPORTED, "more than 2^24 variables");
  for (size_t t = 0; t < n; ++t) {
    const uint64_t o0 = S.offsets[t], o1 = S.offsets[t + 1];
    if (o1 < o0) return fail(OBN_ERR_INVALID, "offsets must be non-decreasing");
    const size_t k = o1 - o0;
    uint64_t cl = 1;
    uint32_t c1 = 0, c2 = 0;             // CI: the two largest cardinalities bound rx + ry of every member
    if (ci) {
      for (size_t i = 0; i < k; ++i) {
        const uint32_t cd = ds->cards[S.var_ids[o0 + i]];
        cl *= cd;
        if (cd > c1) { c2 = c1; c1 = cd; } else if (cd > c2) c2 = cd;
      }
    } else if (int rc = check_tuple(ds->cards.data(), ds->n_vars, S.var_ids + o0, k, t, false, S.mode == MODE_BDEU_ONE,
                                    S.role_a ? S.role_a[t] : 0xffffffffu, 0xffffffffu, &cl)) {
      return rc;
    }
    cells[t] = (uint32_t)cl;
    int g = 0;
    while (g < kClassSmemMax && cl > kClassCells[g]) ++g;
    if (g < c->tune.min_class) g = c->tune.min_class;
    if (g == kClassSmemMax && cl * 4 + 16 + 8 * 512 * 4 > budget)
      g = (c->tune.chunk && cl <= (uint64_t)kChunkCells * kMaxChunks) ? kClassChunk : kClassGlobal;
    // epilogue scratch per warp: the margins, and for tables in global memory one staged slice (ci_slices_any)
    if (ci) rxy[t] = c1 + c2 + (g >= kClassChunk ? c1 * c2 : 0u);
    group[t] = (uint8_t)g;
    ++group_count[g];
    ++n_mine;
    const double members = ci ? (double)(S.mem_off[t + 1] - S.mem_off[t]) : 1.0;
    tests_mine += (size_t)members;
    alg_bytes += members * (double)k * (double)ds->n_rows;      // the unit is the TEST
    scanned_bytes += (double)k * (double)ds->n_rows;
    increments += (double)ds->n_rows;
  }

  // ---- exact-order BDeu: host-libm lgamma tables, cached in the ctx across calls ------------
  const bool bdeu = (S.mode == MODE_BDEU_ONE || S.mode == MODE_BDEU_ALL);
  bool exact = bdeu && S.exact;   // 1 = required, 2 = if the tables fit
  std::vector<uint32_t> lut_ijk, lut_ij;
  const size_t lut_stride = ds->n_rows + 1;
  if (exact) {
    LutCache& lc = c->lut;
    if (lc.stride != lut_stride) lc.clear();                 // another sample count: other tables
    lc.stride = lut_stride;
    for (int attempt = 0;; ++attempt) {
      std::vector<double> pending;
      std::map<uint64_t, uint32_t> pending_ids;
      auto id_of = [&](double a) {
        uint64_t bits;
        memcpy(&bits, &a, 8);
        auto it = lc.ids.find(bits);
        if (it != lc.ids.end()) return it->second;
        auto jt = pending_ids.find(bits);
        if (jt != pending_ids.end()) return jt->second;
        const uint32_t id = (uint32_t)(lc.rows + pending.size());
        pending_ids[bits] = id;
        pending.push_back(a);
        return id;
      };
      lut_ijk.resize(n);
      lut_ij.assign(total_vars, 0);
      uint32_t last_cells = 0, last_ijk = 0;
      for (size_t t = 0; t < n; ++t) {
        const double a_ijk = S.ess / (double)(size_t)cells[t];      // local_score.h:41
        if (cells[t] != last_cells) { last_ijk = id_of(a_ijk); last_cells = cells[t]; }
        lut_ijk[t] = last_ijk;
        uint32_t last_card = 0, last_id = 0;
        for (uint64_t i = S.offsets[t]; i < S.offsets[t + 1]; ++i) {
          const uint32_t cd = ds->cards[S.var_ids[i]];
          if (cd != last_card) { last_id = id_of(a_ijk * (double)(size_t)cd); last_card = cd; }   // :45
          lut_ij[i] = last_id;
        }
      }
      const size_t want_rows = lc.rows + pending.size();
      if (want_rows * lut_stride * sizeof(double) > (size_t)c->tune.lut_max_mb << 20) {
        // the cap is hit by what EARLIER batches left in the cache: evict everything and number this
        // batch's tables afresh, so that the outcome depends on this batch alone, not on the ctx's history
        if (attempt == 0 && lc.rows) {
          lc.clear();
          continue;
        }
        if (S.exact != 2)
          return fail(OBN_ERR_UNSUPPORTED, "exact_order needs %zu lgamma tables of %zu entries", want_rows, lut_stride);
        exact = false;
        ++c->prof.exact_fallbacks;
      } else if (!pending.empty()) {
        if (int rc = lc.append(c, pending)) return rc;
        for (auto& kv : pending_ids) lc.ids[kv.first] = kv.second;
      }
      break;
    }
  }

  // ---- packed input buffer ---------------------------------------------------
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align16(off + bytes); return o; };
  if (total_vars >= 0xffffffffull) return fail(OBN_ERR_UNSUPPORTED, "more than 2^32-2 variable slots in one batch");
  const size_t o_var = take(total_vars * 4), o_hdr = take(n_mine * 16);
  const size_t o_roles = take(S.role_a ? n_mine * 8 : 0);
  const size_t o_soff = take(n_mine * 8);
  const size_t o_coff = take(S.mode == MODE_COUNTS ? n * 8 : 0);
  const size_t o_lijk = take(exact ? n * 4 : 0), o_lij = take(exact ? total_vars * 4 : 0);
  const size_t n_members = ci ? S.mem_off[n] : 0;
  const size_t o_moff = take(ci ? (n + 1) * 4 : 0), o_mem = take(n_members * 8);
  const size_t o_slot = take(pc_round ? n_out * 4 : 0);
  const size_t in_bytes = off;
  if (int rc = G.h_in.reserve(std::max<size_t>(in_bytes, 16))) return rc;
  if (int rc = G.d_in.reserve(std::max<size_t>(in_bytes, 16))) return rc;
  unsigned char* hin = G.h_in.as<unsigned char>();
  {
    uint32_t* desc = reinterpret_cast<uint32_t*>(hin + o_var);
    for (size_t i = 0; i < total_vars; ++i) desc[i] = S.var_ids[i] | ((ds->cards[S.var_ids[i]] - 1u) << 24);
  }
  if (ci) {
    memcpy(hin + o_moff, S.mem_off, (n + 1) * 4);
    memcpy(hin + o_mem, S.members, n_members * 8);
    if (pc_round) memcpy(hin + o_slot, S.slot_of_test, n_out * 4);
  }
  if (exact) {
    memcpy(hin + o_lijk, lut_ijk.data(), n * 4);
    if (total_vars) memcpy(hin + o_lij, lut_ij.data(), total_vars * 4);
  }
  perm.resize(n_mine);
  uint64_t* soff = reinterpret_cast<uint64_t*>(hin + o_soff);
  size_t gstart[kNumClasses + 1];
  gstart[0] = 0;
  for (int g = 0; g < kNumClasses; ++g) gstart[g + 1] = gstart[g] + group_count[g];
  {
    size_t fill[kNumClasses];
    for (int g = 0; g < kNumClasses; ++g) fill[g] = gstart[g];
    for (size_t t = 0; t < n; ++t) perm[fill[group[t]]++] = (uint32_t)t;
  }
  {
    uint32_t* hdr = reinterpret_cast<uint32_t*>(hin + o_hdr);
    uint32_t* roles = S.role_a ? reinterpret_cast<uint32_t*>(hin + o_roles) : nullptr;
    for (size_t q = 0; q < n_mine; ++q) {
      const uint32_t t = perm[q];
      hdr[4 * q + 0] = t;
      hdr[4 * q + 1] = (uint32_t)S.offsets[t];
      hdr[4 * q + 2] = (uint32_t)(S.offsets[t + 1] - S.offsets[t]);
      hdr[4 * q + 3] = cells[t];
      if (roles) {
        roles[2 * q + 0] = S.role_a[t];
        roles[2 * q + 1] = 0xffffffffu;
      }
    }
  }

  // ---- output buffer ----------------------------------------------------------
  size_t out_bytes = 0, counts_total = 0;
  const size_t first_bytes = (S.n_slots * 4 + 7) & ~(size_t)7;
  if (pc_round) out_bytes = first_bytes + 32;          // first[n_slots], then the four counters of a chained round
  else if (ci) out_bytes = n_out * 24;
  else if (S.mode == MODE_BDEU_ONE) out_bytes = n * 8;
  else if (S.mode == MODE_BDEU_ALL) out_bytes = S.scores_dev ? 0 : total_vars * 8;
  else {
    uint64_t* coff = reinterpret_cast<uint64_t*>(hin + o_coff);
    for (size_t t = 0; t < n; ++t) { coff[t] = counts_total; counts_total += cells[t]; }
    out_bytes = S.counts_dev ? 0 : counts_total * 4;
  }
  if (int rc = G.d_out.reserve(std::max<size_t>(out_bytes, 16))) return rc;

  // ---- plan launches: one per class (chunked only when the scratch would not fit)
  struct Launch { int g; size_t begin, end; int nthr, splits; uint32_t vps; size_t smem; size_t scratch_ints; size_t scratch_base; bool barrier_before; uint32_t rep_shift; uint32_t wstride; uint32_t row_splits; uint32_t epi_ctas; size_t partial_base; };
  std::vector<Launch> launches;
  const uint32_t n_vec = (uint32_t)((ds->n_rows + 15) >> 4);
  size_t max_scratch_ints = 0, partial_doubles = 0;
  const size_t target_threads = (size_t)c->sm_count * (size_t)c->tune.threads_per_sm;
  for (int g = 0; g < kNumClasses; ++g) {
    if (!group_count[g]) continue;
    const bool global_tier = (g == kClassGlobal), chunk_tier = (g == kClassChunk);
    size_t i = gstart[g];
    while (i < gstart[g + 1]) {
      const size_t remaining = gstart[g + 1] - i;
      uint32_t max_cells = 0, max_rxy = 0;
      for (size_t q = i; q < gstart[g + 1]; ++q) {
        max_cells = std::max(max_cells, cells[perm[q]]);
        max_rxy = std::max(max_rxy, rxy[perm[q]]);
      }
      // CTA width / row splits (benchmarks/kernel_sweep.py, profiles/r1_kernel_sweep.txt): the
      // best geometry keeps about 768 threads per SM busy in total, with the widest CTA
      // that allows; row splits only pay when a launch has too few tuples to do that
      // (each split repeats the prologue and adds a merge through L2).
      // table copies (obn_count_kernel.cuh): fixed per shared-memory class, compile-time constants of the kernels
      const uint32_t rep_shift = (c->tune.rep && g <= 4 && kRepShift[g] > 0) ? (uint32_t)kRepShift[g] : 0u;
      const size_t table_bytes = chunk_tier ? (size_t)kChunkCells * 4 : global_tier ? 0
                                 : align16((size_t)max_cells * 4) + (rep_shift ? ((size_t)max_cells << rep_shift) * 4 : 0);
      const size_t tpt = target_threads / remaining;
      // the 32-bit-key kernels (tables of > 16 Ki cells: one CTA per SM by shared memory) run up to 512 threads
      const size_t maxthr = g >= kClassSmemMax ? 512 : 256;
      size_t p2 = 32;
      while (p2 * 2 <= tpt && p2 < maxthr) p2 <<= 1;
      // shared memory may allow only a few CTAs per SM: each must then be wide enough to keep the SM busy
      const size_t ctas_by_smem = std::max<size_t>(1, std::min<size_t>(32, (c->smem_optin + 1024) / (table_bytes + 2048)));
      while (p2 * ctas_by_smem < (size_t)c->tune.threads_per_sm && p2 < maxthr) p2 <<= 1;
      // copies pay with wide CTAs: eight warps on one set of copies (benchmarks/micro/smem_incr.cu variant 9:
      // 23 increments/clk/SM at 256 threads against 19.7 with one warp per CTA)
      if (rep_shift) p2 = 256;
      int nthr = (int)p2;
      int splits = 1;
      if (nthr == (int)maxthr) {
        // global tier: every CTA adds straight into the one global table, a split costs no merge
        const uint32_t min_vps = global_tier ? 256u : std::max<uint32_t>(1024u, max_cells / 4u);
        splits = (int)std::min<size_t>(tpt / maxthr, std::max<uint32_t>(1u, n_vec / min_vps));
        splits = std::max(1, splits);
        // a launch of only a few waves ends with SMs idling behind its last CTAs: cut the tuples into row
        // ranges until the launch has min_waves waves of CTAs (the merge is one pass over the table through L2)
        if (!global_tier && c->tune.min_waves > 0) {
          const size_t ctas_per_sm = std::max<size_t>(1, std::min<size_t>(ctas_by_smem, 2048 / maxthr));
          const size_t slots = (size_t)c->sm_count * ctas_per_sm;
          const size_t want = ((size_t)c->tune.min_waves * slots + remaining - 1) / remaining;
          const size_t cap = std::max<uint32_t>(1u, n_vec / std::max<uint32_t>(min_vps, 4096u));   // >= 64 Ki rows per range
          splits = (int)std::max<size_t>((size_t)splits, std::min(want, cap));
        }
      }
      uint32_t row_splits = 1;
      if (chunk_tier) {
        // CTA = (chunk, row range): enough row ranges for about three CTAs per SM over the launch
        nthr = 512;
        const uint32_t chunks = (max_cells + kChunkCells - 1) / kChunkCells;
        const size_t want = (3 * (size_t)c->sm_count + remaining * chunks - 1) / (remaining * chunks);
        row_splits = (uint32_t)std::max<size_t>(1, std::min<size_t>(want, std::max<uint32_t>(1u, n_vec / 2048u)));
        splits = (int)(chunks * row_splits);
      }
      if (c->tune.force_threads) nthr = c->tune.force_threads;
      if (c->tune.force_splits && !chunk_tier)
        splits = std::max(1, std::min<int>(c->tune.force_splits, (int)std::max<uint32_t>(1u, n_vec)));
      const int nw = nthr / 32;
      size_t smem = table_bytes + (size_t)nw * max_rxy * 4;
      smem = std::max<size_t>(align16(smem), 16);
      const bool need_scratch = splits > 1 || global_tier || chunk_tier;
      size_t j = i, scratch_ints = 0;
      if (need_scratch) {
        for (j = i; j < gstart[g + 1]; ++j) {
          const size_t add = (cells[perm[j]] + 3u) & ~3u;
          if (scratch_ints && (scratch_ints + add) * 4 > kScratchMaxBytes) break;
          soff[j] = scratch_ints;
          scratch_ints += add;
        }
      } else {
        for (j = i; j < gstart[g + 1]; ++j) soff[j] = 0;
      }
      uint32_t vps = chunk_tier ? (n_vec + row_splits - 1) / row_splits : (n_vec + splits - 1) / splits;
      if (vps == 0) vps = 1;
      // CI tables in global memory: the epilogue is a second launch of epi_ctas CTAs per tuple (ci_epilogue_kernel)
      uint32_t epi_ctas = 0;
      if (ci && (global_tier || chunk_tier))
        epi_ctas = (uint32_t)std::max<size_t>(1, std::min<size_t>(64, (4 * (size_t)c->sm_count + (j - i) - 1) / std::max<size_t>(1, j - i)));
      launches.push_back({g, i, j, nthr, splits, vps, smem, scratch_ints, 0, false, rep_shift, max_rxy, row_splits, epi_ctas, partial_doubles});
      partial_doubles += (size_t)(j - i) * kGroupMembers * epi_ctas * 2;
      max_scratch_ints = std::max(max_scratch_ints, scratch_ints);
      i = j;
    }
  }
  if (c->tune.largest_first) std::reverse(launches.begin(), launches.end());
  // launches of one batch run concurrently on the ctx's streams, so each gets its own slice of
  // the scratch; when the slices would not fit, a barrier serialises and the slices restart at 0
  {
    size_t base = 0, need = 0;
    for (Launch& L : launches) {
      if (base && (base + L.scratch_ints) * 4 > kScratchMaxBytes) {
        L.barrier_before = true;
        base = 0;
      }
      L.scratch_base = base;
      base += L.scratch_ints;
      need = std::max(need, base);
    }
    max_scratch_ints = need;
  }
  if (max_scratch_ints)
    if (int rc = c->d_scratch.reserve(max_scratch_ints * 4, true)) return rc;
  if (int rc = c->d_tickets.reserve(std::max<size_t>(n_mine, 4) * 4, true)) return rc;
  if (partial_doubles)
    if (int rc = c->d_partials.reserve(partial_doubles * 8)) return rc;

  const auto t_packed = std::chrono::steady_clock::now();
  if (in_bytes) {
    if (c->q_count) {
      // another batch is running: copy beside it (this staging's buffers are free -- its last batch was collected)
      OBN_CUDA(cudaMemcpyAsync(G.d_in.p, hin, in_bytes, cudaMemcpyHostToDevice, c->up_stream));
      OBN_CUDA(cudaEventRecord(G.uploaded, c->up_stream));
      OBN_CUDA(cudaStreamWaitEvent(c->stream, G.uploaded, 0));
    } else {
      OBN_CUDA(cudaMemcpyAsync(G.d_in.p, hin, in_bytes, cudaMemcpyHostToDevice, c->stream));
    }
  }

  // ---- launch ------------------------------------------------------------------
  unsigned char* din = G.d_in.as<unsigned char>();
  unsigned char* dout = G.d_out.as<unsigned char>();
  BatchArgs a{};
  a.data = ds->data; a.stride = ds->stride; a.n_rows = (uint32_t)ds->n_rows;
  a.data4 = c->tune.pack4 ? ds->data4 : nullptr; a.stride4 = ds->stride4;
  a.desc = reinterpret_cast<const uint32_t*>(din + o_var);
  a.mode = S.mode; a.kind = S.kind; a.ess = S.ess;
  if (exact) {
    a.lut = c->lut.dev; a.lut_stride = lut_stride;
    a.lut_ijk = reinterpret_cast<const uint32_t*>(din + o_lijk);
    a.lut_ij = reinterpret_cast<const uint32_t*>(din + o_lij);
  }
  if (ci) {
    a.mem_off = reinterpret_cast<const uint32_t*>(din + o_moff);
    a.members = reinterpret_cast<const uint2*>(din + o_mem);
    if (pc_round) {
      a.first = reinterpret_cast<int32_t*>(dout);
      a.slot_of_test = reinterpret_cast<const uint32_t*>(din + o_slot);
      a.level = S.level;
      OBN_CUDA(cudaMemsetAsync(dout, 0x7f, first_bytes, c->stream));      // kFirstNone in every slot
      if (S.chain) {
        // the round submitted before this one, still in flight or collected last: its first[] is in the other staging
        if (c->last_round < 0 || c->last_round == g_idx || c->last_round_slots != S.n_slots)
          return fail(OBN_ERR_INVALID, "chain: no previous PC round over the same %zu edge slots to continue", S.n_slots);
        a.prev_first = c->st[c->last_round].d_out.as<int32_t>();
        a.skipped = reinterpret_cast<unsigned long long*>(dout + first_bytes);
        OBN_CUDA(cudaMemsetAsync(dout + first_bytes, 0, 32, c->stream));
      }
    } else {
      a.stat = reinterpret_cast<double*>(dout);
      a.dof = reinterpret_cast<unsigned long long*>(dout + n_out * 8);
      a.pval = S.p ? reinterpret_cast<double*>(dout + n_out * 16) : nullptr;
    }
  }
  else if (bdeu) a.scores = S.scores_dev ? S.scores_dev : reinterpret_cast<double*>(dout);
  else { a.counts_out = S.counts_dev ? S.counts_dev : reinterpret_cast<int32_t*>(dout); a.counts_off = reinterpret_cast<const uint64_t*>(din + o_coff); }
  a.scratch = c->d_scratch.as<int32_t>();
  a.tickets = c->d_tickets.as<unsigned int>();

  // fork/join over the ctx's auxiliary streams: launch i of a segment goes to stream i % (1+kAuxStreams)
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (c->profile_on) {
    e0 = G.e0;
    e1 = G.e1;
    OBN_CUDA(cudaEventRecord(e0, c->stream));
  }
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> dbg_events;
  struct DbgDesc { int g, tuples, nthr, splits, smem; };
  std::vector<DbgDesc> dbg_desc;
  const int n_streams = std::max(1, std::min(1 + obn_ctx::kAuxStreams, c->tune.streams));
  bool aux_used[obn_ctx::kAuxStreams] = {};
  bool forked = false;
  int slot = 0;
  auto join = [&]() -> int {
    for (int j = 0; j < obn_ctx::kAuxStreams; ++j)
      if (aux_used[j]) {
        OBN_CUDA(cudaEventRecord(c->ev_join[j], c->aux[j]));
        OBN_CUDA(cudaStreamWaitEvent(c->stream, c->ev_join[j], 0));
        aux_used[j] = false;
      }
    forked = false;
    slot = 0;
    return OBN_OK;
  };
  for (const Launch& L : launches) {
    if (L.barrier_before)
      if (int rc = join()) return rc;
    cudaStream_t st = c->stream;
    const int sidx = slot % n_streams;
    ++slot;
    if (sidx > 0) {
      if (!forked) {
        OBN_CUDA(cudaEventRecord(c->ev_fork, c->stream));
        forked = true;
      }
      if (!aux_used[sidx - 1]) {
        OBN_CUDA(cudaStreamWaitEvent(c->aux[sidx - 1], c->ev_fork, 0));
        aux_used[sidx - 1] = true;
      }
      st = c->aux[sidx - 1];
    }
    a.hdr = reinterpret_cast<const uint4*>(din + o_hdr) + L.begin;
    a.roles = S.role_a ? reinterpret_cast<const uint2*>(din + o_roles) + L.begin : nullptr;
    a.scratch_off = reinterpret_cast<const uint64_t*>(din + o_soff) + L.begin;
    a.scratch = c->d_scratch.as<int32_t>() + L.scratch_base;
    a.tickets = c->d_tickets.as<unsigned int>() + L.begin;
    a.splits = L.splits;
    a.vec_per_split = L.vps;
    a.wstride = L.wstride;
    a.chunk_cells = kChunkCells;
    a.row_splits = L.row_splits;
    a.epi_ctas = L.epi_ctas;
    a.partials = L.epi_ctas ? c->d_partials.as<double>() + L.partial_base : nullptr;
    const unsigned grid = (unsigned)((L.end - L.begin) * (size_t)L.splits);
    cudaError_t e;
    if (L.rep_shift) {
      if (L.g == 0) e = launch_count<8, TIER_SMEM, true, kRepShift[0]>(a, grid, L.nthr, L.smem, st);
      else if (L.g == 1) e = launch_count<8, TIER_SMEM, false, kRepShift[1]>(a, grid, L.nthr, L.smem, st);
      else if (L.g == 2) e = launch_count<16, TIER_SMEM, true, kRepShift[2]>(a, grid, L.nthr, L.smem, st);
      else e = launch_count<16, TIER_SMEM, true, kRepShift[3]>(a, grid, L.nthr, L.smem, st);
    } else
    if (L.g == 0) e = launch_count<8, TIER_SMEM, true>(a, grid, L.nthr, L.smem, st);
    else if (L.g == 1) e = launch_count<8, TIER_SMEM, false>(a, grid, L.nthr, L.smem, st);
    else if (L.g <= 4) e = launch_count<16, TIER_SMEM, true>(a, grid, L.nthr, L.smem, st);
    else if (L.g == kClassSmemMax) e = launch_count<32, TIER_SMEM, false>(a, grid, L.nthr, L.smem, st);
    else if (L.g == kClassChunk) e = launch_count<32, TIER_CHUNK, false>(a, grid, L.nthr, L.smem, st);
    else e = launch_count<32, TIER_GLOBAL, false>(a, grid, L.nthr, L.smem, st);
    if (e != cudaSuccess) return fail(OBN_ERR_CUDA, "count_kernel launch failed: %s", cudaGetErrorString(e));
    if (L.epi_ctas) {
      ci_epilogue_kernel<<<(unsigned)((L.end - L.begin) * L.epi_ctas), 256, (size_t)8 * L.wstride * 4 + 16, st>>>(a);
      e = cudaGetLastError();
      if (e != cudaSuccess) return fail(OBN_ERR_CUDA, "ci_epilogue_kernel launch failed: %s", cudaGetErrorString(e));
      ++c->prof.total_launches;
    }
    if (c->tune.debug >= 2) {               // launch log: [begin, end) of every launch relative to the batch's first event
      cudaEvent_t b0, b1;
      cudaEventCreate(&b0);
      cudaEventCreate(&b1);
      // (recorded AFTER the launch on the same stream: b1 - e0 is the launch's completion time; b0 is unused)
      cudaEventRecord(b1, st);
      dbg_events.push_back({b0, b1});
      dbg_desc.push_back({L.g, (int)(L.end - L.begin), L.nthr, L.splits, (int)L.smem});
    }
    ++c->prof.count_launches;
    ++c->prof.total_launches;
  }
  if (int rc = join()) return rc;
  if (c->profile_on) OBN_CUDA(cudaEventRecord(e1, c->stream));

  // ---- sharded PC round: merge the ranks' first-passing indices, still on the device -----------
  if (pc_round && merge_ranks)
    if (int rc = comm_allreduce_min_i32(c, reinterpret_cast<int32_t*>(dout), S.n_slots)) return rc;

  // ---- download ------------------------------------------------------------------
  if (int rc = G.h_out.reserve(std::max<size_t>(out_bytes, 16))) return rc;
  if (out_bytes) OBN_CUDA(cudaMemcpyAsync(G.h_out.p, dout, out_bytes, cudaMemcpyDeviceToHost, c->stream));
  // whatever this staging held of an earlier PC round is gone now
  if (pc_round) {
    c->last_round = g_idx;
    c->last_round_slots = S.n_slots;
  } else if (c->last_round == g_idx) {
    c->last_round = -1;
  }
  obn_ctx::Pending& P = G.pending;
  P = obn_ctx::Pending();
  P.n_out = ci ? tests_mine : n;
  P.n_results = n_out;
  P.n_slots = pc_round ? S.n_slots : 0;
  P.out_bytes = out_bytes;
  P.in_bytes = in_bytes;
  P.alg_bytes = alg_bytes;
  P.scanned_bytes = scanned_bytes;
  P.increments = increments;
  P.n_rows = (double)ds->n_rows;
  P.tables = n_mine;
  P.want_p = S.p != nullptr;
  P.timed = c->profile_on;
  P.chained = pc_round && S.chain;
  if (S.defer) {                    // CI modes only (obn_ci_tests_submit / obn_pc_round_submit)
    OBN_CUDA(cudaEventRecord(G.done, c->stream));
    ++c->q_count;
    const auto now = std::chrono::steady_clock::now();
    c->prof.host_prep_ms += std::chrono::duration<double, std::milli>(t_packed - t_begin).count();
    c->prof.host_wait_ms += std::chrono::duration<double, std::milli>(now - t_packed).count();
    return OBN_OK;
  }
  OBN_CUDA(cudaStreamSynchronize(c->stream));
  if (!dbg_events.empty() && e0) {
    cudaDeviceSynchronize();
    fprintf(stderr, "[obn batch] %zu tuples, %zu launches:", n, dbg_events.size());
    for (size_t i = 0; i < dbg_events.size(); ++i) {
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, dbg_events[i].second);
      fprintf(stderr, "  c%d x%d t%d s%d smem%dK done@%.3f", dbg_desc[i].g, dbg_desc[i].tuples, dbg_desc[i].nthr, dbg_desc[i].splits,
              dbg_desc[i].smem >> 10, ms);
      cudaEventDestroy(dbg_events[i].first);
      cudaEventDestroy(dbg_events[i].second);
    }
    float tot = 0;
    cudaEventElapsedTime(&tot, e0, e1);
    fprintf(stderr, "  all@%.3f ms\n", tot);
  }
  const auto t_synced = std::chrono::steady_clock::now();
  c->prof.host_prep_ms += std::chrono::duration<double, std::milli>(t_packed - t_begin).count();
  c->prof.host_wait_ms += std::chrono::duration<double, std::milli>(t_synced - t_packed).count();
  const unsigned char* hout = G.h_out.as<unsigned char>();
  if (ci) {
    finish_batch(c, G, S.stat, S.dof, S.p, S.first);
    return OBN_OK;
  }
  if (S.mode == MODE_BDEU_ONE) {
    memcpy(S.scores, hout, n * 8);
  } else if (S.mode == MODE_BDEU_ALL) {
    if (total_vars && !S.scores_dev) memcpy(S.scores, hout, total_vars * 8);
  } else {
    const int32_t* src = reinterpret_cast<const int32_t*>(hout);
    size_t o = 0;
    for (size_t t = 0; t < n && !S.counts_dev; ++t) {
      memcpy(S.counts + S.counts_off[t], src + o, (size_t)cells[t] * 4);
      o += cells[t];
    }
  }
  finish_batch(c, G, nullptr, nullptr, nullptr, nullptr);
  return OBN_OK;
}

}  // namespace obn

extern "C" {

int obn_count_tables(obn_ctx* c, const obn_dataset* ds, const uint32_t* var_ids, const uint64_t* offsets,
                     size_t n_tuples, int32_t* counts_out, const uint64_t* out_offsets) {
  if (n_tuples && (!counts_out || !out_offsets)) return fail(OBN_ERR_INVALID, "counts_out/out_offsets is NULL");
  BatchSpec S{c, ds, var_ids, offsets, n_tuples, MODE_COUNTS};
  S.counts = counts_out;
  S.counts_off = out_offsets;
  return run_batch(S);
}

int obn_bdeu_scores(obn_ctx* c, const obn_dataset* ds, const uint32_t* var_ids, const uint64_t* offsets,
                    size_t n_tuples, const uint32_t* child, double ess, int exact_order, double* scores_out) {
  if (n_tuples && !scores_out) return fail(OBN_ERR_INVALID, "scores_out is NULL");
  BatchSpec S{c, ds, var_ids, offsets, n_tuples, MODE_BDEU_ONE};
  S.role_a = child;
  S.ess = ess;
  S.exact = exact_order;
  S.scores = scores_out;
  return run_batch(S);
}

int obn_bdeu_subset_scores(obn_ctx* c, const obn_dataset* ds, const uint32_t* var_ids,
                           const uint64_t* offsets, size_t n_tuples, double ess, int exact_order,
                           double* scores_out) {
  if (n_tuples && !scores_out) return fail(OBN_ERR_INVALID, "scores_out is NULL");
  BatchSpec S{c, ds, var_ids, offsets, n_tuples, MODE_BDEU_ALL};
  S.ess = ess;
  S.exact = exact_order;
  S.scores = scores_out;
  return run_batch(S);
}

// Every CI entry point: plan the groups (tests over one variable set share its table), then one batch.
static int ci_tests_impl(obn_ctx* c, const obn_dataset* ds, const uint32_t* var_ids, const uint64_t* offsets,
                         size_t n_tuples, const uint32_t* x, const uint32_t* y, int kind, double* stat_out,
                         uint64_t* dof_out, double* p_out, bool defer, const uint32_t* slot_of_test, size_t n_slots,
                         double level, int shard, int32_t* first_out, bool chain = false) {
  if (kind != OBN_CI_CHI2 && kind != OBN_CI_G2) return fail(OBN_ERR_INVALID, "Unsupported CI test type");
  if (!c || !ds) return fail(OBN_ERR_INVALID, "ctx or dataset is NULL");
  if (ds->ctx != c) return fail(OBN_ERR_INVALID, "dataset belongs to another ctx");
  if (n_tuples && (!offsets || !x || !y)) return fail(OBN_ERR_INVALID, "offsets, x or y is NULL");
  if (n_tuples && offsets[n_tuples] && !var_ids) return fail(OBN_ERR_INVALID, "var_ids is NULL");
  if (n_tuples >= 0x7f000000ull) return fail(OBN_ERR_UNSUPPORTED, "more than 2^31-2^24 tests in one batch");
  if (n_slots) {
    if (!slot_of_test && n_tuples) return fail(OBN_ERR_INVALID, "slot_of_test is NULL");
    for (size_t t = 0; t < n_tuples; ++t)
      if (slot_of_test[t] >= n_slots) return fail(OBN_ERR_RANGE, "slot_of_test[%zu] = %u out of range", t, slot_of_test[t]);
  }
  static thread_local CiGroupPlan P;            // capacity reused across the rounds of a PC run
  static const uint64_t zero_off[1] = {0};
  const bool sharded = shard != 0 && n_slots && c->comm && c->comm_world > 1;
  if (int rc = plan_ci_groups(ds->cards.data(), ds->n_vars, var_ids, n_tuples ? offsets : zero_off, n_tuples, x, y, P,
                              sharded ? c->comm_rank : 0, sharded ? c->comm_world : 1))
    return rc;
  BatchSpec S{c, ds, P.var_ids.data(), P.offsets.data(), P.groups(), MODE_CI};
  S.kind = kind;
  S.stat = stat_out;
  S.dof = dof_out;
  S.p = p_out;
  S.n_out = n_tuples;
  S.mem_off = P.mem_off.data();
  S.members = P.members.data();
  S.slot_of_test = slot_of_test;
  S.n_slots = n_slots;
  S.level = level;
  S.first = first_out;
  S.shard = shard != 0;
  S.defer = defer;
  S.chain = chain && n_slots;
  return run_batch(S);
}

int obn_ci_tests(obn_ctx* c, const obn_dataset* ds, const uint32_t* var_ids, const uint64_t* offsets,
                 size_t n_tuples, const uint32_t* x, const uint32_t* y, int kind, double* stat_out,
                 uint64_t* dof_out, double* p_out) {
  return ci_tests_impl(c, ds, var_ids, offsets, n_tuples, x, y, kind, stat_out, dof_out, p_out, false, nullptr, 0, 0.0,
                       0, nullptr);
}

int obn_pc_round(obn_ctx* c, const obn_dataset* ds, const uint32_t* var_ids, const uint64_t* offsets, size_t n_tests,
                 const uint32_t* x, const uint32_t* y, const uint32_t* slot_of_test, size_t n_slots, int kind,
                 double level, int shard, int32_t* first_out) {
  if (!n_slots) return OBN_OK;
  if (!first_out) return fail(OBN_ERR_INVALID, "first_out is NULL");
  return ci_tests_impl(c, ds, var_ids, offsets, n_tests, x, y, kind, nullptr, nullptr, nullptr, false, slot_of_test,
                       n_slots, level, shard, first_out);
}

// The same two calls split in halves, so that the caller's host thread can prepare its next batch
// while the device runs this one: submit plans, packs, uploads and enqueues everything (the input
// arrays may be reused as soon as it returns); collect waits for the OLDEST submitted batch and writes its
// results.  Two batches may be in flight per ctx; every other batch call fails until both are collected.
int obn_ci_tests_submit(obn_ctx* c, const obn_dataset* ds, const uint32_t* var_ids, const uint64_t* offsets,
                        size_t n_tuples, const uint32_t* x, const uint32_t* y, int kind, int want_p) {
  static double dummy;              // run_batch only asks whether p is wanted
  return ci_tests_impl(c, ds, var_ids, offsets, n_tuples, x, y, kind, nullptr, nullptr, want_p ? &dummy : nullptr, true,
                       nullptr, 0, 0.0, 0, nullptr);
}

int obn_pc_round_submit(obn_ctx* c, const obn_dataset* ds, const uint32_t* var_ids, const uint64_t* offsets,
                        size_t n_tests, const uint32_t* x, const uint32_t* y, const uint32_t* slot_of_test,
                        size_t n_slots, int kind, double level, int shard, int chain) {
  if (!n_slots) return OBN_OK;
  return ci_tests_impl(c, ds, var_ids, offsets, n_tests, x, y, kind, nullptr, nullptr, nullptr, true, slot_of_test,
                       n_slots, level, shard, nullptr, chain != 0);
}

static int collect_impl(obn_ctx* c, double* stat_out, uint64_t* dof_out, double* p_out, int32_t* first_out, bool round) {
  if (!c) return fail(OBN_ERR_INVALID, "ctx is NULL");
  if (!c->q_count) return OBN_OK;     // nothing was submitted (an empty batch): nothing to hand out
  obn_ctx::Staging& G = c->st[c->q_head];           // the oldest batch in flight
  const obn_ctx::Pending& P = G.pending;
  if (round != (P.n_slots != 0)) return fail(OBN_ERR_INVALID, "the batch in flight was submitted by the other submit call");
  if (p_out && !P.want_p) return fail(OBN_ERR_INVALID, "p-values were not requested at submit");
  OBN_CUDA(cudaSetDevice(c->device));
  const auto t0 = std::chrono::steady_clock::now();
  c->q_head = (c->q_head + 1) % obn_ctx::kStagings;
  --c->q_count;
  OBN_CUDA(cudaEventSynchronize(G.done));
  c->prof.host_wait_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  finish_batch(c, G, stat_out, dof_out, p_out, first_out);
  return OBN_OK;
}

int obn_ci_tests_collect(obn_ctx* c, double* stat_out, uint64_t* dof_out, double* p_out) {
  return collect_impl(c, stat_out, dof_out, p_out, nullptr, false);
}

int obn_pc_round_collect(obn_ctx* c, int32_t* first_out) {
  return collect_impl(c, nullptr, nullptr, nullptr, first_out, true);
}

}  // extern "C"
