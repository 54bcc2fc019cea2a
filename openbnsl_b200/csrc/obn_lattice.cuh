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
// c[i] / v[i]: cardinality / dataset column of the variable at digit i (digit 0 most significant)
struct JointCards { uint32_t c[kMaxTupleVars]; uint32_t v[kMaxTupleVars]; };
__global__ void joint_count_kernel(const uint8_t* data, size_t stride, uint32_t n_rows, int n, JointCards cards,
                                   int32_t* tab) {
  const uint32_t n_vec = (n_rows + 15u) >> 4;
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < n_vec) {
    const size_t off = (size_t)v << 4;
    KeyVec<32> key;
    key.init(ld_codes16(data + (size_t)cards.v[0] * stride + off));
    for (int i = 1; i < n; ++i) key.push(ld_codes16(data + (size_t)cards.v[i] * stride + off), cards.c[i]);
    count16<32, TIER_GLOBAL, false, -1>(key, reinterpret_cast<unsigned char*>(tab), 0u, 0u);
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

// ---- bottom tier: the sub-lattice of the last 8 (binary) variables, on chip --------------------------
//
// Most of the lattice's 3^n cells sit in mid-size tables, and producing every table in HBM costs
// two reads and one write per cell (3.4 TB for n = 24).  The last kBotL variables are the fastest
// digits of every key (contingency_table.h:76-85), so a table that contains all of them is a
// sequence of 2^kBotL-cell CHUNKS, and the tables of its 2^kBotL sub-patterns (which bottom
// variables stay) are sums inside each chunk.  One CTA streams chunks through the SM and
// accumulates F(pattern) for all 256 patterns of the table.  Only tables holding all bottom
// variables ever exist in HBM: 3^(n-8) * 256 cells instead of 3^n.
//
// A chunk's 3^8 = 6561 marginal cells are split by the two fastest digits.  A QUAD is the four
// cells that differ only in those two digits; there are 3^6 = 729 quads, one per (upper pattern U =
// which of the other six bottom variables are kept, cell of U's table).  Per chunk:
//   phase 1  the quads are built level by level in shared memory with the same "sum out the highest
//            absent digit of the parent" rule as the HBM tiers: two 16-byte LDS, four adds, one
//            16-byte STS per quad, conflict-free.  Which quad a thread builds at each level is a
//            fixed PROGRAM (bottom_program() in obn_lattice.inl); idle threads of a partly filled
//            warp read always-zero quads and store to a dump quad (branch-free), idle warps skip.
//   phase 2  a flat, barrier-free pass: thread t takes quads t, t+256, t+512, derives the five cells
//            that sum out one or both of the two fastest digits in registers (nothing depends on
//            them, so they never touch shared memory) and gathers the nine terms: 27 independent
//            table lookups per thread in flight, twelve accumulators per thread.
// The quad buffer is double-buffered, so fast warps start the next chunk while slow ones finish
// phase 2.
//
constexpr int kBotL = 8;
constexpr int kBotCells = 1 << kBotL;      // cells of a chunk
constexpr int kBotThreads = 256;
constexpr int kBotUpper = kBotL - 2;       // digits handled through shared memory
constexpr int kBotQuads = 729;             // 3^6 quads of one chunk
constexpr int kBotZeroQuad = kBotQuads;    // quads 729, 730 stay 0 (sources of idle threads)
constexpr int kBotDumpQuad = kBotQuads + 2;
constexpr int kBotTabQuads = kBotQuads + 3;
constexpr int kBotLevels = kBotUpper + 1;  // j = 6 (the chunk itself) .. 0
// items of level j = C(6,j) 2^j:  j  6   5    4    3    2   1   0
//                                    64  192  240  160  60  12  1
// program word: source quad (10 bits) | log2(distance to the second source quad) << 10 | destination quad << 13
__host__ __device__ constexpr uint32_t bot_word(uint32_t src0, uint32_t sh, uint32_t dst) {
  return src0 | (sh << 10) | (dst << 13);
}
constexpr uint32_t kBotIdleWord = bot_word(kBotZeroQuad, 0, kBotDumpQuad);
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

constexpr int kBotSlots = 3;               // quads per thread in phase 2: ceil(729 / 256)

// phase 1, one level: build this thread's quad from its two source quads and store it
template <int J>
__device__ __forceinline__ void bot_level(const uint32_t* prog, int4* tab, int tid) {
  const uint32_t w = prog[(kBotUpper - 1 - J) * kBotThreads + tid];
  const uint32_t s0 = w & 0x3ffu;
  const int4 a = tab[s0], b = tab[s0 + (1u << ((w >> 10) & 7u))];
  tab[w >> 13] = make_int4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}
#define OBN_BOT_LEVEL(J, ITEMS)                               \
  if ((tid & ~31) < (ITEMS)) bot_level<J>(prog, tab, tid);    \
  __syncthreads();

// phase 2, one quad (cells x = {00, 01, 10, 11} in the two fastest digits; its upper pattern keeps
// j variables): acc[0] both digits kept (4 cells, level j+2), acc[1] the fastest summed out,
// acc[2] the second fastest summed out (2 cells each, level j+1), acc[3] both (1 cell, level j).
// o2/o1/o0 = first element of the term table of level j+2 / j+1 / j.
__device__ __forceinline__ void bot_quad_lut(const long long* lut, uint32_t o2, uint32_t o1, uint32_t o0, const int4& x,
                                             long long (&acc)[4]) {
  const int m0 = x.x + x.y, m1 = x.z + x.w, n0 = x.x + x.z, n1 = x.y + x.w;
  const long long a0 = __ldg(lut + (o2 + (uint32_t)x.x)), a1 = __ldg(lut + (o2 + (uint32_t)x.y));
  const long long a2 = __ldg(lut + (o2 + (uint32_t)x.z)), a3 = __ldg(lut + (o2 + (uint32_t)x.w));
  const long long b0 = __ldg(lut + (o1 + (uint32_t)m0)), b1 = __ldg(lut + (o1 + (uint32_t)m1));
  const long long c0 = __ldg(lut + (o1 + (uint32_t)n0)), c1 = __ldg(lut + (o1 + (uint32_t)n1));
  const long long d0 = __ldg(lut + (o0 + (uint32_t)(m0 + m1)));
  acc[0] += (a0 + a1) + (a2 + a3);
  acc[1] += b0 + b1;
  acc[2] += c0 + c1;
  acc[3] += d0;
}
__device__ __forceinline__ void bot_quad_libm(double a2, double a1, double a0, double scale, const int4& x, long long (&acc)[4]) {
  const int m0 = x.x + x.y, m1 = x.z + x.w, n0 = x.x + x.z, n1 = x.y + x.w;
  acc[0] += fix_term(x.x, a2, scale) + fix_term(x.y, a2, scale) + fix_term(x.z, a2, scale) + fix_term(x.w, a2, scale);
  acc[1] += fix_term(m0, a1, scale) + fix_term(m1, a1, scale);
  acc[2] += fix_term(n0, a1, scale) + fix_term(n1, a1, scale);
  acc[3] += fix_term(m0 + m1, a0, scale);
}

// grid = (chunk groups, tasks).  The table of task t is its `dst` (cells = hi * mul * lo, a multiple
// of 256); F words of its 256 patterns are Fx[f_base + f_index + (pattern mask << f_shift)].
// LUT = every cell count has a term table (the normal case); otherwise libm per cell.
template <bool LUT>
__global__ void __launch_bounds__(kBotThreads, 3) bottom_lattice_kernel(const MargTask* tasks, uint64_t mul, const int32_t* base,
                                                                       const uint32_t* g_prog, const uint16_t* g_pat, FixTable ft,
                                                                       double ess, unsigned long long* Fx, uint64_t f_base,
                                                                       int f_shift, uint32_t chunks_per_cta) {
  __shared__ uint32_t prog[kBotUpper * kBotThreads];
  __shared__ int4 tabs[2][kBotTabQuads];
  __shared__ uint32_t s_off[kBotL + 1];
  __shared__ double s_a[kBotL + 1];
  const int tid = threadIdx.x;
  const MargTask T = tasks[blockIdx.y];
  const uint64_t cells = (uint64_t)T.hi * mul * T.lo;
  const uint64_t chunks = cells >> kBotL;
  const uint64_t c_begin = (uint64_t)blockIdx.x * chunks_per_cta;
  if (c_begin >= chunks) return;
  const uint64_t c_end = min(chunks, c_begin + chunks_per_cta);
#pragma unroll
  for (int s = 0; s < kBotUpper; ++s) prog[s * kBotThreads + tid] = __ldg(g_prog + s * kBotThreads + tid);
  if (tid < 2 * (kBotTabQuads - kBotQuads))
    tabs[tid / (kBotTabQuads - kBotQuads)][kBotQuads + tid % (kBotTabQuads - kBotQuads)] = make_int4(0, 0, 0, 0);
  if (tid <= kBotL) {
    // pattern of popcount k: cells(table) / 2^(8-k) cells; a = ess / cells (local_score.h:41)
    const double a = ess / (double)(cells >> (kBotL - tid));
    int id = 0;
    for (int i = 0; i < ft.n_lut; ++i)
      if (ft.lut_a[i] == a) { id = i; break; }
    s_a[tid] = a;
    s_off[tid] = (uint32_t)((size_t)id * ft.stride);
  }
  __syncthreads();
  // phase-2 items of this thread: quads tid, tid + 256, tid + 512 (the last slot is ragged)
  uint32_t quad[kBotSlots], upper[kBotSlots], o2[kBotSlots], o1[kBotSlots], o0[kBotSlots];
#pragma unroll
  for (int s = 0; s < kBotSlots; ++s) {
    const uint32_t q = (uint32_t)(s * kBotThreads + tid);
    upper[s] = q < (uint32_t)kBotQuads ? __ldg(g_pat + q) : (uint32_t)kBotNoPattern;
    quad[s] = q < (uint32_t)kBotQuads ? q : (uint32_t)kBotZeroQuad;
    const int j = upper[s] == kBotNoPattern ? 0 : __popc(upper[s]);
    o2[s] = s_off[j + 2];
    o1[s] = s_off[j + 1];
    o0[s] = s_off[j];
  }
  const int4* src = reinterpret_cast<const int4*>(base + (uint64_t)T.dst_off * mul + (c_begin << kBotL));
  long long acc[kBotSlots][4];
#pragma unroll
  for (int s = 0; s < kBotSlots; ++s)
#pragma unroll
    for (int l = 0; l < 4; ++l) acc[s][l] = 0;
  const double scale = ft.scale;
  int4 next = make_int4(0, 0, 0, 0);
  if (tid < 64) next = __ldg(src + tid);
  for (uint64_t ch = c_begin; ch < c_end; ++ch) {
    int4* tab = tabs[(ch - c_begin) & 1];
    if (tid < 64) {                       // level 6: the chunk itself
      tab[tid] = next;
      src += kBotCells / 4;
      if (ch + 1 < c_end) next = __ldg(src + tid);
    }
    __syncthreads();
    OBN_BOT_LEVEL(5, 192)
    OBN_BOT_LEVEL(4, 240)
    OBN_BOT_LEVEL(3, 160)
    OBN_BOT_LEVEL(2, 60)
    OBN_BOT_LEVEL(1, 12)
    OBN_BOT_LEVEL(0, 1)
    // phase 2 (no barrier after it: the next chunk builds in the other buffer, and the one after
    // that is at least seven barriers away)
#pragma unroll
    for (int s = 0; s < kBotSlots; ++s) {
      const int4 x = tab[quad[s]];
      if (LUT) {
        bot_quad_lut(ft.lut, o2[s], o1[s], o0[s], x, acc[s]);
      } else {
        const int j = upper[s] == kBotNoPattern ? 0 : __popc(upper[s]);
        bot_quad_libm(s_a[j + 2], s_a[j + 1], s_a[j], scale, x, acc[s]);
      }
    }
  }
  // flush: quads of one upper pattern are an aligned power-of-two run of lanes, so a butterfly that
  // only adds partners of the same pattern leaves the run's total in every lane of the run
  unsigned long long* f = Fx + f_base + T.f_index;
#pragma unroll
  for (int s = 0; s < kBotSlots; ++s) {
    const uint32_t U = upper[s];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const bool same = __shfl_xor_sync(0xffffffffu, U, o) == U;
#pragma unroll
      for (int l = 0; l < 4; ++l) {
        const long long other = __shfl_xor_sync(0xffffffffu, acc[s][l], o);
        if (same) acc[s][l] += other;
      }
    }
    const uint32_t left = __shfl_up_sync(0xffffffffu, U, 1);
    if (U != kBotNoPattern && ((tid & 31) == 0 || left != U)) {
      // pattern bit b = variable n-1-b = bit (L-1-b) of the low mask; bits 1, 0 are the two fastest digits
      const uint32_t p = U << 2;
      atomicAdd(f + ((size_t)(__brev(p | 3u) >> 24) << f_shift), (unsigned long long)acc[s][0]);
      atomicAdd(f + ((size_t)(__brev(p | 2u) >> 24) << f_shift), (unsigned long long)acc[s][1]);   // fastest digit summed out
      atomicAdd(f + ((size_t)(__brev(p | 1u) >> 24) << f_shift), (unsigned long long)acc[s][2]);
      atomicAdd(f + ((size_t)(__brev(p) >> 24) << f_shift), (unsigned long long)acc[s][3]);
    }
  }
}
#undef OBN_BOT_LEVEL

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
