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
    const double a = ess / (double)cells;                       // local_score.h:41 for the table `dst`
    const int s_lut = find_lut(tt, a);                          // uniform over the CTA
    const double* lut = s_lut >= 0 ? tt.lut + (size_t)s_lut * tt.stride : nullptr;
    const double lg_a = lut ? 0.0 : lgamma(a);
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
          if (lut) {
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
          if (n) acc += lut ? lut[n] : (lgamma((double)n + a) - lg_a);
        }
      }
    }
  }
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

}  // namespace obn
