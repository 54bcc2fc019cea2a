// obn_lattice.cuh -- kernels of the marginalisation lattice (exhaustive search, fast mode).
//
// The reference builds one contingency table per subset W by a pass over all N rows
// (exhaustive_search.cpp:55-57 -> contingency_table.h:52-55).  The table of W is also the
// table of W + {v} summed over v, so once the joint table of ALL variables has been counted
// (one pass over the data) every other table follows by integer adds: sum_k C(n,k) * 2^(k+1)
// cell reads instead of 2^n * N row visits (n = 24 binary, N = 1M: 5.6e11 vs 1.7e13).
// Counts stay bit-exact (integer sums of the same rows).
//
// BDeu decomposes over tables (local_score.h:38-61 with a_ij = a_ijk * r_child = ess / cells(pa)):
//     score(child | pa) = F(pa + child) - F(pa),
//     F(S) = sum over cells c of S:  lgamma(N_c + ess/cells(S)) - lgamma(ess/cells(S))
// so ONE scalar per subset is kept; it is accumulated in the same kernel that produces the
// table.  (The association of the fp64 sum differs from the reference's sequential order:
// this is the 1e-9-relative "fast" mode; exact-order mode keeps the counting path.)
#pragma once
#include "obn_count_kernel.cuh"

namespace obn {

// One marginalisation task: dst = src summed over one digit.  Offsets and `hi` are in units that
// the launch scales by `mul` (the cell count of the block's fixed high variables), so one task
// list serves every block.
struct MargTask {
  uint32_t src_off;    // * mul -> first cell of the source table (relative to the launch's base)
  uint32_t dst_off;    // * mul -> first cell of the destination table
  uint32_t hi;         // * mul -> cells above the removed digit
  uint32_t r;          // cardinality of the removed variable
  uint32_t lo;         // cells below the removed digit
  uint32_t f_index;    // where F(dst subset) goes, relative to the launch's F base
  int32_t lo_shift;    // log2(lo) when lo is a power of two >= 4 and r == 2 (vector fast path), else -1
  uint32_t pad1;
};

constexpr int kMargThreads = 256;
constexpr int kMargCellsPerThread = 8;
constexpr int kMargChunk = kMargThreads * kMargCellsPerThread;

// lgamma(n + a) - lgamma(a) from a device-resident table when one exists for this a
struct TermTable {
  const double* lut;      // [n_lut][stride] or null
  const double* lut_a;    // [n_lut] the a of each table, ascending cells <=> descending a
  int n_lut;
  size_t stride;
};

__device__ __forceinline__ int find_lut(const TermTable& tt, double a) {
  for (int i = 0; i < tt.n_lut; ++i)
    if (tt.lut_a[i] == a) return i;
  return -1;
}

// grid = (chunks of the largest destination table, tasks).  partials[task * gridDim.x + chunk].
__global__ void __launch_bounds__(kMargThreads, 8) marg_kernel(const MargTask* tasks, uint64_t mul, const int32_t* src_base,
                                                            int32_t* dst_base, double ess, TermTable tt,
                                                            double* partials) {
  __shared__ double red[32];
  const MargTask T = tasks[blockIdx.y];
  const uint64_t hi = (uint64_t)T.hi * mul, lo = T.lo, r = T.r;
  const uint64_t cells = hi * lo;
  const uint64_t c0 = (uint64_t)blockIdx.x * kMargChunk;
  double acc = 0.0;
  if (c0 < cells) {
    const int32_t* src = src_base + (uint64_t)T.src_off * mul;
    int32_t* dst = dst_base + (uint64_t)T.dst_off * mul;
    const bool want_f = partials != nullptr;                    // the bottom tier computes F itself
    const double a = ess / (double)cells;                       // local_score.h:41 for the table `dst`
    const int s_lut = want_f ? find_lut(tt, a) : -1;            // uniform over the CTA
    const double* lut = s_lut >= 0 ? tt.lut + (size_t)s_lut * tt.stride : nullptr;
    const double lg_a = (lut || !want_f) ? 0.0 : lgamma(a);
    if (T.lo_shift >= 2 && (cells & 3) == 0 && (((uintptr_t)src | (uintptr_t)dst) & 15) == 0) {
      // removed digit is binary and everything below it is a power of two >= 4 cells wide:
      // four consecutive destination cells = two 16-byte loads, one 16-byte store, no divisions
      const int sh = T.lo_shift;
      const uint64_t lo_mask = lo - 1;
#pragma unroll
      for (int u = 0; u < kMargCellsPerThread / 4; ++u) {
        const uint64_t c = c0 + ((uint64_t)u * kMargThreads + threadIdx.x) * 4;
        if (c < cells) {
          const uint64_t i = c >> sh, j = c & lo_mask;
          const int32_t* p = src + ((i << 1) << sh) + j;
          const int4 x = *reinterpret_cast<const int4*>(p);
          const int4 y = *reinterpret_cast<const int4*>(p + lo);
          const int4 n = make_int4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
          *reinterpret_cast<int4*>(dst + c) = n;
          if (!want_f) {
          } else if (lut) {
            if (n.x) acc += lut[n.x];
            if (n.y) acc += lut[n.y];
            if (n.z) acc += lut[n.z];
            if (n.w) acc += lut[n.w];
          } else {
            if (n.x) acc += lgamma((double)n.x + a) - lg_a;
            if (n.y) acc += lgamma((double)n.y + a) - lg_a;
            if (n.z) acc += lgamma((double)n.z + a) - lg_a;
            if (n.w) acc += lgamma((double)n.w + a) - lg_a;
          }
        }
      }
    } else {
#pragma unroll
      for (int u = 0; u < kMargCellsPerThread; ++u) {
        const uint64_t c = c0 + (uint64_t)u * kMargThreads + threadIdx.x;
        if (c < cells) {
          uint64_t i, j;
          if (cells * r < 0xffffffffull) {                  // 32-bit division when it is enough
            i = (uint32_t)c / (uint32_t)lo;
            j = (uint32_t)c - (uint32_t)i * (uint32_t)lo;
          } else {
            i = c / lo;
            j = c - i * lo;
          }
          const int32_t* p = src + (i * r) * lo + j;
          int n = 0;
          for (uint32_t s = 0; s < r; ++s) n += p[(uint64_t)s * lo];
          dst[c] = n;
          if (n && want_f) acc += lut ? lut[n] : (lgamma((double)n + a) - lg_a);
        }
      }
    }
  }
  if (partials == nullptr) return;
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partials[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = acc;
}

// F of a table that already exists (the joint table): same partial layout, one "task".
__global__ void __launch_bounds__(kMargThreads) table_f_kernel(const int32_t* tab, uint64_t cells, double ess, TermTable tt,
                                                               double* partials) {
  __shared__ double red[32];
  __shared__ int s_lut;
  const double a = ess / (double)cells;
  if (threadIdx.x == 0) s_lut = find_lut(tt, a);
  __syncthreads();
  const double* lut = s_lut >= 0 ? tt.lut + (size_t)s_lut * tt.stride : nullptr;
  const double lg_a = lut ? 0.0 : lgamma(a);
  double acc = 0.0;
  const uint64_t c0 = (uint64_t)blockIdx.x * kMargChunk;
#pragma unroll
  for (int u = 0; u < kMargCellsPerThread; ++u) {
    const uint64_t c = c0 + (uint64_t)u * kMargThreads + threadIdx.x;
    if (c < cells) {
      const int n = tab[c];
      if (n) acc += lut ? lut[n] : (lgamma((double)n + a) - lg_a);
    }
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc;
}

// F[f_base + task.f_index] = sum of the task's partials in chunk order (deterministic).
__global__ void marg_reduce_kernel(const MargTask* tasks, uint32_t n_tasks, uint32_t chunks, const double* partials,
                                   double* F, uint64_t f_base) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_tasks) return;
  const double* p = partials + (size_t)t * chunks;
  double s = 0.0;
  for (uint32_t i = 0; i < chunks; ++i) s += p[i];
  F[f_base + tasks[t].f_index] = s;
}

// The joint table of all n variables, counted straight into its final place (zeroed by the
// caller): one thread per 16-row vector, RED.E.ADD into L2.  Padding rows are all-zero codes.
struct JointCards { uint32_t c[kMaxTupleVars]; };
__global__ void joint_count_kernel(const uint8_t* data, size_t stride, uint32_t n_rows, int n, JointCards cards,
                                   int32_t* tab) {
  const uint32_t n_vec = (n_rows + 15u) >> 4;
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < n_vec) {
    const size_t off = (size_t)v << 4;
    KeyVec<32> key;
    key.init(ld_codes16(data + off));
    for (int i = 1; i < n; ++i) key.push(ld_codes16(data + (size_t)i * stride + off), cards.c[i]);
    count16<32, TIER_GLOBAL, false>(key, reinterpret_cast<unsigned char*>(tab));
  }
  if (v == 0) {
    const int pad = (int)((n_vec << 4) - n_rows);
    if (pad) atomicAdd(&tab[0], -pad);
  }
}

// lut[id][m] = lgamma(m + a_id) - lgamma(a_id), m = 0 .. stride-1 (device libm: fast mode only)
__global__ void lut_fill_kernel(double* lut, const double* lut_a, size_t stride) {
  const size_t m = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= stride) return;
  const double a = lut_a[blockIdx.y];
  lut[(size_t)blockIdx.y * stride + m] = m ? lgamma((double)m + a) - lgamma(a) : 0.0;
}

// ---- bottom tier: the sub-lattice of the last 8 (binary) variables, in shared memory ---------------
//
// Most of the lattice's 3^n cells sit in mid-size tables, and producing every table in HBM costs
// two reads and one write per cell (3.4 TB for n = 24).  The last kBotL variables are the fastest
// digits of every key (contingency_table.h:76-85), so a table that contains all of them is a
// sequence of 2^kBotL-cell CHUNKS, and the tables of its 2^kBotL sub-patterns (which bottom
// variables stay) are sums inside each chunk.  One CTA therefore streams chunks through shared
// memory, builds the 3^kBotL = 6561 marginal cells of each chunk with one add per cell (the same
// "sum out the highest absent digit of the parent" rule as the HBM tiers), and accumulates
// F(pattern) for all 256 patterns.  Only tables holding all bottom variables ever exist in HBM:
// 3^(n-8) * 256 cells instead of 3^n.
//
// Work of one chunk is a fixed PROGRAM (built on the host, bottom_program() in obn_lattice.inl):
// level k = 7..0 (bottom variables kept) is a few SLOTS, one item per thread and slot:
//   quad slot    four consecutive cells of one pattern whose summed-out digit sits above >= 2 kept
//                digits: two 16-byte LDS, four adds, one 16-byte STS;
//   scalar slot  one cell (the patterns where that digit is among the two lowest, and levels < 3).
// Threads without an item read two always-zero cells and write to a dump cell, so a slot is
// branch-free and the slots of a level pipeline; whole idle warps skip the slot.
//
// F is accumulated in 64-bit FIXED POINT (term * 2^shift, rounded once per term): integer adds are
// associative, so lane-private accumulators, warp reductions and RED.ADD.64 into the F array give
// the same bits whatever the chunk -> CTA split or the rank count, and the sum carries no
// accumulation error at all (better than any fp64 summation order).
constexpr int kBotL = 8;
constexpr int kBotCells = 1 << kBotL;      // cells of a chunk = threads of the CTA
constexpr int kBotThreads = kBotCells;
constexpr int kBotExpanded = 6561;         // 3^8 marginal cells of one chunk
constexpr int kBotZero = 6564;             // tab[kBotZero .. +7] stay 0 (sources of idle threads)
constexpr int kBotDump = 6572;             // tab[kBotDump .. +3]: where idle threads store
constexpr int kBotTab = 6576;
constexpr int kBotSlots = 16;
// slot table (must match bottom_program(); the host checks it):
//   level  quad slots (items)     scalar slot (items)
//   8      -                      0 (256, the chunk itself)
//   7      1 (192)                2 (256)
//   6      3, 4 (400)             5 (192)
//   5      6, 7 (416)             8 (128)
//   4      9 (256)                10 (96)
//   3      11 (100)               12 (48)
//   2      -                      13 (112)
//   1      -                      14 (16)
//   0      -                      15 (1)
// quad word:   source quad (11 bits) | log2(distance to the second source) << 11 | destination quad << 14
// scalar word: source cell (13 bits) | log2(distance) << 13 | destination cell << 16
__host__ __device__ constexpr uint32_t bot_quad_word(uint32_t src0, uint32_t sh, uint32_t dst) {
  return (src0 >> 2) | (sh << 11) | ((dst >> 2) << 14);
}
__host__ __device__ constexpr uint32_t bot_cell_word(uint32_t src0, uint32_t sh, uint32_t dst) {
  return src0 | (sh << 13) | (dst << 16);
}
constexpr uint32_t kBotIdleQuad = bot_quad_word(kBotZero, 2, kBotDump);
constexpr uint32_t kBotIdleCell = bot_cell_word(kBotZero, 0, kBotDump);
constexpr uint16_t kBotNoPattern = 0xffff;

struct FixTable {
  const long long* lut;   // [n_lut][stride] round(term * 2^shift), or null
  const double* lut_a;    // [n_lut]
  int n_lut;
  size_t stride;
  double scale;           // 2^shift
};

__device__ __noinline__ long long fix_term(int n, double a, double scale) {
  return n ? __double2ll_rn((lgamma((double)n + a) - lgamma(a)) * scale) : 0ll;
}

// true for the warps that hold at least one of the slot's ITEMS items
#define OBN_BOT_ACTIVE(ITEMS) ((tid & ~31) < (ITEMS))

template <int SLOT, int ITEMS>
__device__ __forceinline__ void bot_quad_load(const uint32_t* prog, const int32_t* tab, int tid, int4& x, uint32_t& dst) {
  x = make_int4(0, 0, 0, 0);
  dst = kBotDump;
  if (OBN_BOT_ACTIVE(ITEMS)) {
    const uint32_t w = prog[SLOT * kBotThreads + tid];
    const uint32_t s0 = (w & 0x7ffu) << 2;
    const int4 a = *reinterpret_cast<const int4*>(tab + s0);
    const int4 b = *reinterpret_cast<const int4*>(tab + s0 + (1u << ((w >> 11) & 7u)));
    x = make_int4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
    dst = (w >> 14) << 2;
  }
}
template <int ITEMS>
__device__ __forceinline__ void bot_quad_store(int32_t* tab, int tid, const int4& x, uint32_t dst) {
  if (OBN_BOT_ACTIVE(ITEMS)) *reinterpret_cast<int4*>(tab + dst) = x;
}
template <int SLOT, int ITEMS>
__device__ __forceinline__ void bot_cell_load(const uint32_t* prog, const int32_t* tab, int tid, int& x, uint32_t& dst) {
  x = 0;
  dst = kBotDump;
  if (OBN_BOT_ACTIVE(ITEMS)) {
    const uint32_t w = prog[SLOT * kBotThreads + tid];
    const uint32_t s0 = w & 0x1fffu;
    x = tab[s0] + tab[s0 + (1u << ((w >> 13) & 7u))];
    dst = w >> 16;
  }
}
template <int ITEMS>
__device__ __forceinline__ void bot_cell_store(int32_t* tab, int tid, int x, uint32_t dst) {
  if (OBN_BOT_ACTIVE(ITEMS)) tab[dst] = x;
}

// terms of a slot's cells: table gathers (issued here, added one level later), or libm when this
// cell count has no table
template <bool LUT, int ITEMS>
__device__ __forceinline__ long long bot_quad_terms(const long long* lut, double a, double scale, int tid, const int4& x) {
  long long g = 0;
  if (OBN_BOT_ACTIVE(ITEMS)) {
    if (LUT) {
      const long long g0 = __ldg(lut + x.x), g1 = __ldg(lut + x.y), g2 = __ldg(lut + x.z), g3 = __ldg(lut + x.w);
      g = (g0 + g1) + (g2 + g3);
    } else {
      g = fix_term(x.x, a, scale) + fix_term(x.y, a, scale) + fix_term(x.z, a, scale) + fix_term(x.w, a, scale);
    }
  }
  return g;
}
template <bool LUT, int ITEMS>
__device__ __forceinline__ long long bot_cell_term(const long long* lut, double a, double scale, int tid, int x) {
  long long g = 0;
  if (OBN_BOT_ACTIVE(ITEMS)) g = LUT ? __ldg(lut + x) : fix_term(x, a, scale);
  return g;
}

// reduce a slot over the W lanes that share a pattern, then one RED.ADD.64 per (slot, pattern)
template <int SLOT, int ITEMS, int W>
__device__ __forceinline__ void bot_flush(const uint16_t* g_pat, long long a, unsigned long long* f, int f_shift, int tid) {
  if (!OBN_BOT_ACTIVE(ITEMS)) return;
#pragma unroll
  for (int o = W / 2; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
  const uint32_t pattern = __ldg(g_pat + SLOT * kBotThreads + tid);
  // pattern bit j = variable n-1-j = bit (L-1-j) of the low mask
  if (pattern != kBotNoPattern && (tid & (W - 1)) == 0)
    atomicAdd(f + ((size_t)(__brev(pattern) >> 24) << f_shift), (unsigned long long)a);
}

// grid = (chunk groups, tasks).  The table of task t is its `dst` (cells = hi * mul * lo, a multiple
// of 256); F words of its 256 patterns are Fx[f_base + f_index + (pattern mask << f_shift)].
// LUT = every cell count has a term table (the normal case); otherwise libm per cell.
template <bool LUT>
__global__ void __launch_bounds__(kBotThreads, 3) bottom_lattice_kernel(const MargTask* tasks, uint64_t mul, const int32_t* base,
                                                                       const uint32_t* g_prog, const uint16_t* g_pat, FixTable ft,
                                                                       double ess, unsigned long long* Fx, uint64_t f_base,
                                                                       int f_shift, uint32_t chunks_per_cta) {
  __shared__ uint32_t prog[kBotSlots * kBotThreads];
  __shared__ __align__(16) int32_t tab[kBotTab];
  __shared__ const long long* s_lut[kBotL + 1];
  __shared__ double s_a[kBotL + 1];
  const int tid = threadIdx.x;
  const MargTask T = tasks[blockIdx.y];
  const uint64_t cells = (uint64_t)T.hi * mul * T.lo;
  const uint64_t chunks = cells >> kBotL;
  const uint64_t c_begin = (uint64_t)blockIdx.x * chunks_per_cta;
  if (c_begin >= chunks) return;
  const uint64_t c_end = min(chunks, c_begin + chunks_per_cta);
#pragma unroll
  for (int s = 1; s < kBotSlots; ++s) prog[s * kBotThreads + tid] = __ldg(g_prog + s * kBotThreads + tid);
  if (tid < kBotTab - kBotZero) tab[kBotZero + tid] = 0;
  if (tid <= kBotL) {
    // pattern of popcount k: cells(table) / 2^(8-k) cells; a = ess / cells (local_score.h:41)
    const double a = ess / (double)(cells >> (kBotL - tid));
    int id = -1;
    for (int i = 0; i < ft.n_lut; ++i)
      if (ft.lut_a[i] == a) { id = i; break; }
    s_a[tid] = a;
    s_lut[tid] = id >= 0 ? ft.lut + (size_t)id * ft.stride : nullptr;
  }
  __syncthreads();
  const int32_t* src = base + (uint64_t)T.dst_off * mul + (c_begin << kBotL);
  long long acc[kBotSlots];
#pragma unroll
  for (int s = 0; s < kBotSlots; ++s) acc[s] = 0;
  int next = __ldg(src + tid);
  const double scale = ft.scale;
  for (uint64_t ch = c_begin; ch < c_end; ++ch) {
    const int c8 = next;
    src += kBotCells;
    if (ch + 1 < c_end) next = __ldg(src + tid);
    tab[tid] = c8;
    const long long t0 = bot_cell_term<LUT, 256>(s_lut[8], s_a[8], scale, tid, c8);
    __syncthreads();
    int4 qa, qb;
    int ca;
    uint32_t da, db, dc;
    // level 7
    bot_quad_load<1, 192>(prog, tab, tid, qa, da);
    bot_cell_load<2, 256>(prog, tab, tid, ca, dc);
    const long long t1 = bot_quad_terms<LUT, 192>(s_lut[7], s_a[7], scale, tid, qa);
    const long long t2 = bot_cell_term<LUT, 256>(s_lut[7], s_a[7], scale, tid, ca);
    bot_quad_store<192>(tab, tid, qa, da);
    bot_cell_store<256>(tab, tid, ca, dc);
    acc[0] += t0;
    __syncthreads();
    // level 6
    bot_quad_load<3, 256>(prog, tab, tid, qa, da);
    bot_quad_load<4, 144>(prog, tab, tid, qb, db);
    bot_cell_load<5, 192>(prog, tab, tid, ca, dc);
    const long long t3 = bot_quad_terms<LUT, 256>(s_lut[6], s_a[6], scale, tid, qa);
    const long long t4 = bot_quad_terms<LUT, 144>(s_lut[6], s_a[6], scale, tid, qb);
    const long long t5 = bot_cell_term<LUT, 192>(s_lut[6], s_a[6], scale, tid, ca);
    bot_quad_store<256>(tab, tid, qa, da);
    bot_quad_store<144>(tab, tid, qb, db);
    bot_cell_store<192>(tab, tid, ca, dc);
    acc[1] += t1;
    acc[2] += t2;
    __syncthreads();
    // level 5
    bot_quad_load<6, 256>(prog, tab, tid, qa, da);
    bot_quad_load<7, 160>(prog, tab, tid, qb, db);
    bot_cell_load<8, 128>(prog, tab, tid, ca, dc);
    const long long t6 = bot_quad_terms<LUT, 256>(s_lut[5], s_a[5], scale, tid, qa);
    const long long t7 = bot_quad_terms<LUT, 160>(s_lut[5], s_a[5], scale, tid, qb);
    const long long t8 = bot_cell_term<LUT, 128>(s_lut[5], s_a[5], scale, tid, ca);
    bot_quad_store<256>(tab, tid, qa, da);
    bot_quad_store<160>(tab, tid, qb, db);
    bot_cell_store<128>(tab, tid, ca, dc);
    acc[3] += t3;
    acc[4] += t4;
    acc[5] += t5;
    __syncthreads();
    // level 4
    bot_quad_load<9, 256>(prog, tab, tid, qa, da);
    bot_cell_load<10, 96>(prog, tab, tid, ca, dc);
    const long long t9 = bot_quad_terms<LUT, 256>(s_lut[4], s_a[4], scale, tid, qa);
    const long long t10 = bot_cell_term<LUT, 96>(s_lut[4], s_a[4], scale, tid, ca);
    bot_quad_store<256>(tab, tid, qa, da);
    bot_cell_store<96>(tab, tid, ca, dc);
    acc[6] += t6;
    acc[7] += t7;
    acc[8] += t8;
    __syncthreads();
    // level 3
    bot_quad_load<11, 100>(prog, tab, tid, qa, da);
    bot_cell_load<12, 48>(prog, tab, tid, ca, dc);
    const long long t11 = bot_quad_terms<LUT, 100>(s_lut[3], s_a[3], scale, tid, qa);
    const long long t12 = bot_cell_term<LUT, 48>(s_lut[3], s_a[3], scale, tid, ca);
    bot_quad_store<100>(tab, tid, qa, da);
    bot_cell_store<48>(tab, tid, ca, dc);
    acc[9] += t9;
    acc[10] += t10;
    __syncthreads();
    // level 2
    bot_cell_load<13, 112>(prog, tab, tid, ca, dc);
    const long long t13 = bot_cell_term<LUT, 112>(s_lut[2], s_a[2], scale, tid, ca);
    bot_cell_store<112>(tab, tid, ca, dc);
    acc[11] += t11;
    acc[12] += t12;
    __syncthreads();
    // level 1
    bot_cell_load<14, 16>(prog, tab, tid, ca, dc);
    const long long t14 = bot_cell_term<LUT, 16>(s_lut[1], s_a[1], scale, tid, ca);
    bot_cell_store<16>(tab, tid, ca, dc);
    acc[13] += t13;
    __syncthreads();
    // level 0
    bot_cell_load<15, 1>(prog, tab, tid, ca, dc);
    const long long t15 = bot_cell_term<LUT, 1>(s_lut[0], s_a[0], scale, tid, ca);
    acc[14] += t14;
    acc[15] += t15;
    // (the next chunk's first store is tab[0..255]; level 0 read level 1's cells only, and a
    // barrier separates that store from the next level-7 loads)
  }
  unsigned long long* f = Fx + f_base + T.f_index;
  bot_flush<0, 256, 32>(g_pat, acc[0], f, f_shift, tid);
  bot_flush<1, 192, 32>(g_pat, acc[1], f, f_shift, tid);      // level 7 quads: 32 quads per pattern
  bot_flush<2, 256, 32>(g_pat, acc[2], f, f_shift, tid);
  bot_flush<3, 256, 16>(g_pat, acc[3], f, f_shift, tid);      // level 6 quads: 16 per pattern
  bot_flush<4, 144, 16>(g_pat, acc[4], f, f_shift, tid);
  bot_flush<5, 192, 32>(g_pat, acc[5], f, f_shift, tid);
  bot_flush<6, 256, 8>(g_pat, acc[6], f, f_shift, tid);       // level 5 quads: 8 per pattern
  bot_flush<7, 160, 8>(g_pat, acc[7], f, f_shift, tid);
  bot_flush<8, 128, 32>(g_pat, acc[8], f, f_shift, tid);
  bot_flush<9, 256, 4>(g_pat, acc[9], f, f_shift, tid);       // level 4 quads: 4 per pattern
  bot_flush<10, 96, 16>(g_pat, acc[10], f, f_shift, tid);
  bot_flush<11, 100, 2>(g_pat, acc[11], f, f_shift, tid);     // level 3 quads: 2 per pattern
  bot_flush<12, 48, 8>(g_pat, acc[12], f, f_shift, tid);
  bot_flush<13, 112, 4>(g_pat, acc[13], f, f_shift, tid);
  bot_flush<14, 16, 2>(g_pat, acc[14], f, f_shift, tid);
  bot_flush<15, 1, 1>(g_pat, acc[15], f, f_shift, tid);
}
#undef OBN_BOT_ACTIVE

// F words of one rank's slice: fixed point -> double, in place
__global__ void fix_to_double_kernel(double* F, size_t n, double inv_scale) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) F[i] = (double)reinterpret_cast<const long long*>(F)[i] * inv_scale;
}

// lutx[id][m] = round((lgamma(m + a_id) - lgamma(a_id)) * scale)
__global__ void lut_fill_fix_kernel(long long* lut, const double* lut_a, size_t stride, double scale) {
  const size_t m = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= stride) return;
  const double a = lut_a[blockIdx.y];
  lut[(size_t)blockIdx.y * stride + m] = m ? __double2ll_rn((lgamma((double)m + a) - lgamma(a)) * scale) : 0ll;
}

}  // namespace obn
