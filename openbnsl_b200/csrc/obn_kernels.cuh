// obn_kernels.cuh -- the batched contingency-counting kernels (sm_100a) and
// their fused epilogues.  One CTA owns one (tuple, row-split); see DESIGN.md.
//
// Replaces, per tuple, the reference's
//   ContingencyTable<true> ctor      include/base/contingency_table.h:34-56
//   calculate_local_score (BDeu)     include/score/local_score.h:22-62
//   citest (chi-square / G-square)   include/citest/citest.h:73-137
#pragma once
#include "obn_device.cuh"

namespace obn {

enum Mode : int { MODE_COUNTS = 0, MODE_CI = 1, MODE_BDEU_ONE = 2, MODE_BDEU_ALL = 3,
                  MODE_CI_POOL = 4 };   // CI tests of one (x, y) pair sharing a joint table
enum Tier : int { TIER_SMEM = 0, TIER_GLOBAL = 1 };

struct BatchArgs {
  // dataset
  const uint8_t* data;      // column-major codes, column v at data + v*stride
  size_t stride;            // bytes between columns (multiple of 128)
  uint32_t n_rows;
  // batch (device copies of the CSR arrays)
  const uint32_t* desc;     // [offsets[n]]  var_id | (card-1) << 24, ascending var ids per tuple
  const uint4* hdr;         // per launch tuple: {tuple id, desc offset, #vars, #cells}
  const uint2* roles;       // per launch tuple: CI {x, y} with x < y; BDEU_ONE {child, -}
  int mode;
  int kind;                 // CI: 0 chi2, 1 G2
  double ess;
  // exact-order BDeu (host-libm lgamma tables); null => fast mode
  const double* lut;        // [n_luts][lut_stride]
  size_t lut_stride;
  const uint32_t* lut_ijk;  // [n_tuples]
  const uint32_t* lut_ij;   // CSR-shaped like var_ids
  // outputs (device)
  double* stat;
  unsigned long long* dof;
  double* pval;             // CI: p = pchisq(stat, dof) (citest.h:139); may be null
  double* scores;           // BDEU_ONE: [tuple]   BDEU_ALL: [offsets[t] + i]
  int32_t* counts_out;      // MODE_COUNTS
  const uint64_t* counts_off;
  // row splitting / global tier
  int splits;               // CTAs per tuple
  uint32_t vec_per_split;   // 16-row vectors per split
  int32_t* scratch;         // zero on entry, zero on exit
  const uint64_t* scratch_off;  // [launch tuples]
  unsigned int* tickets;        // [launch tuples], zero on entry and exit
  // MODE_CI_POOL: the launch tuple is a POOL -- the joint table over (union of the members'
  // conditioning sets, x, y); member m of pool q is the CI test members[m].x whose conditioning
  // set is the subset members[m].y (bit i = i-th conditioning digit of the pool, kernel order)
  const uint32_t* mem_off;  // [pools + 1], indexed by pool id (hdr.x)
  const uint2* members;
  uint32_t pool_cells;      // joint tables above this size are single tests: no pass buffers
  int wide_prefetch;        // POOL kernels, tuples of more than six columns: 1 scan_rows_wide, 2 scan_rows_tma
  uint32_t stage_off;       // scan_rows_tma: byte offset of the tile ring in dynamic shared memory (128-aligned)
};

// ---------------------------------------------------------------------------
// epilogues.  `tab` is the dense int32 table in KERNEL digit order (generic
// pointer: shared or global).  All threads of the CTA participate.
// ---------------------------------------------------------------------------

// chi-square / G-square with per-slice observed-margin dof and Yates
// (citest.h:86-137).  Kernel digit order puts x then y as the two fastest
// digits, so slice s is the contiguous block tab[s*rx*ry ...) laid out
// x_state * ry + y_state exactly like the reference's slice (citest.h:83).

// one cell's contribution (citest.h:116-131)
__device__ __forceinline__ double ci_cell(double obs, int ri, int cj, double n_total, bool yates, int kind) {
  const double ex = (double)ri * (double)cj / n_total;              // :116
  if (yates) {                                                      // :120-125
    const double diff = ex - obs;
    const double dir = (diff > 0) ? 1.0 : (diff < 0) ? -1.0 : 0.0;
    obs += fmin(0.5, fabs(diff)) * dir;
  }
  if (kind == 0) {
    const double df = obs - ex;
    return df * df / ex;                                            // :127-129
  }
  return (obs > 0) ? 2.0 * obs * log(obs / ex) : 0.0;               // :130-131
}

// Small slices (rx, ry <= 4, the common case for the named networks): ONE LANE
// per slice, the slice held in registers.
__device__ __forceinline__ void ci_slices_small(const int32_t* tab, uint32_t n_slices, uint32_t rx, uint32_t ry,
                                                int kind, double& stat, unsigned long long& dof) {
  const uint32_t sl = rx * ry;
  for (uint32_t s = threadIdx.x; s < n_slices; s += blockDim.x) {
    const int32_t* t = tab + (size_t)s * sl;
    int o[4][4], rs[4] = {0, 0, 0, 0}, cs[4] = {0, 0, 0, 0};
    int tot = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int v = ((uint32_t)i < rx && (uint32_t)j < ry) ? t[i * ry + j] : 0;
        o[i][j] = v;
        rs[i] += v;
        cs[j] += v;
        tot += v;
      }
    if (tot == 0) continue;                                             // citest.h:100
    int robs = 0, cobs = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) { robs += (rs[i] > 0); cobs += (cs[i] > 0); }
    const unsigned d = (unsigned)(robs - 1) * (unsigned)(cobs - 1);     // :109
    const bool yates = (d == 1);                                        // :110
    dof += d;
    const double n_total = (double)tot;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (rs[i] > 0 && cs[j] > 0) stat += ci_cell((double)o[i][j], rs[i], cs[j], n_total, yates, kind);
  }
}

// Any slice shape: one warp per slice, margins in per-warp scratch.
__device__ __forceinline__ void ci_slices_any(const int32_t* tab, uint32_t n_slices, uint32_t rx, uint32_t ry,
                                              int kind, int32_t* wscratch, double& stat,
                                              unsigned long long& dof) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const uint32_t sl = rx * ry;
  int32_t* rs = wscratch;        // [rx]
  int32_t* cs = wscratch + rx;   // [ry]
  for (uint32_t s = warp; s < n_slices; s += nw) {
    const int32_t* t = tab + (size_t)s * sl;
    int tot = 0, robs = 0, cobs = 0;
    for (uint32_t i = lane; i < rx; i += 32) {
      int a = 0;
      for (uint32_t j = 0; j < ry; ++j) a += t[i * ry + j];
      rs[i] = a;
      tot += a;
      robs += (a > 0);
    }
    for (uint32_t j = lane; j < ry; j += 32) {
      int a = 0;
      for (uint32_t i = 0; i < rx; ++i) a += t[i * ry + j];
      cs[j] = a;
      cobs += (a > 0);
    }
    tot = warp_sum(tot);
    robs = warp_sum(robs);
    cobs = warp_sum(cobs);
    __syncwarp();
    if (tot != 0) {                                        // citest.h:100
      const unsigned d = (unsigned)(robs - 1) * (unsigned)(cobs - 1);   // :109
      const bool yates = (d == 1);                         // :110
      if (lane == 0) dof += d;
      const double n_total = (double)tot;
      for (uint32_t c = lane; c < sl; c += 32) {
        const uint32_t i = c / ry, j = c - i * ry;
        const int ri = rs[i], cj = cs[j];
        if (ri > 0 && cj > 0) stat += ci_cell((double)t[c], ri, cj, n_total, yates, kind);
      }
    }
    __syncwarp();
  }
}

__device__ __forceinline__ void epilogue_ci(const int32_t* tab, uint32_t cells, uint32_t rx,
                                            uint32_t ry, int kind, int32_t* wscratch,
                                            double* red_d, unsigned long long* red_u,
                                            double* stat_out, unsigned long long* dof_out, double* p_out) {
  const uint32_t n_slices = cells / (rx * ry);
  double stat = 0.0;
  unsigned long long dof = 0;
  if (rx <= 4 && ry <= 4) ci_slices_small(tab, n_slices, rx, ry, kind, stat, dof);
  else ci_slices_any(tab, n_slices, rx, ry, kind, wscratch, stat, dof);
  stat = block_sum(stat, red_d);
  dof = block_sum(dof, red_u);
  if (threadIdx.x == 0) {
    *stat_out = stat;
    *dof_out = dof;
    if (p_out) *p_out = pchisq(stat, dof);               // citest.h:139
  }
}

// ---------------------------------------------------------------------------
// MODE_CI_POOL: the table of (Z, x, y) from the joint table of (P, x, y), Z a subset of P, by
// summing out the digits of P - Z one at a time, most significant first (so the weight of the
// digit being removed is still its weight in the joint table).  Integer sums of the same rows:
// the result IS the table the counting loop would have produced for (Z, x, y).
// ---------------------------------------------------------------------------

// one output cell of one pass: src viewed as [hi][c][w], out[j] = sum_d src[hi][d][lo], j = hi*w + lo
__host__ __device__ __forceinline__ int32_t marg_cell(const int32_t* src, uint32_t j, uint32_t w, uint32_t c) {
  const uint32_t hi = j / w, lo = j - hi * w;
  const int32_t* p = src + (size_t)hi * c * w + lo;
  int32_t a = 0;
  for (uint32_t d = 0; d < c; ++d) a += p[(size_t)d * w];
  return a;
}

// The pass schedule, shared by the kernel and its host replay (obn_ci_pool_marginalise).
// card[0 .. n_z) are the pool's conditioning digits in kernel order; `cells` the joint table's.
struct MargPasses {
  uint32_t w, cur;          // weight of the digit last visited in the joint table; cells of the current table
  int i;
  __host__ __device__ MargPasses(uint32_t cells) : w(cells), cur(cells), i(0) {}
  // next removed digit: its weight and cardinality, and the cells of the table after the pass
  __host__ __device__ bool next(const uint32_t* card, int n_z, uint32_t keep, uint32_t& pw, uint32_t& pc,
                                uint32_t& n_out) {
    while (i < n_z) {
      const uint32_t c = card[i];
      w /= c;                                 // weight of digit i in the joint table
      const bool kept = (keep >> i) & 1u;
      ++i;
      if (kept || c == 1u) continue;          // summing out a one-state digit changes nothing
      // the digits removed so far were all more significant, the block below digit i is untouched:
      // its weight in the current table is still w
      cur /= c;
      pw = w; pc = c; n_out = cur;
      return true;
    }
    return false;
  }
};

// All threads of the CTA.  buf0 holds >= cells/2 ints, buf1 >= cells/4 (every pass at least halves
// the table: one-state digits are skipped); returns the table (tab itself when nothing is removed) and its size, visible to
// every thread.
__device__ __forceinline__ const int32_t* marginalise(const int32_t* tab, uint32_t cells, const uint32_t* card,
                                                      int n_z, uint32_t keep, int32_t* buf0, int32_t* buf1,
                                                      uint32_t& out_cells) {
  const int32_t* src = tab;
  int32_t* dst = buf0;
  MargPasses it(cells);
  uint32_t w, c, n_out;
  while (it.next(card, n_z, keep, w, c, n_out)) {
    for (uint32_t j = threadIdx.x; j < n_out; j += blockDim.x) dst[j] = marg_cell(src, j, w, c);
    __syncthreads();
    src = dst;
    dst = (dst == buf0) ? buf1 : buf0;
  }
  out_cells = it.cur;
  return src;
}

__device__ __forceinline__ double bdeu_term(int n, double a, double lg_a) {
  return n ? (lgamma((double)n + a) - lg_a) : 0.0;   // local_score.h:52-53
}

// BDeu, fast mode: parallel sums (warp-shuffle reductions).  Scores of the
// children ci in [c_begin, c_end) of the table are written to out[ci - c_begin].
__device__ __forceinline__ void epilogue_bdeu_fast(const int32_t* tab, uint32_t cells, int k,
                                                   const uint32_t* card, double ess, int c_begin,
                                                   int c_end, double* red_d, double* out) {
  const double a_ijk = ess / (double)cells;             // local_score.h:41
  const double lg_ijk = lgamma(a_ijk);
  double t1 = 0.0;
  for (uint32_t c = threadIdx.x; c < cells; c += blockDim.x) t1 += bdeu_term(tab[c], a_ijk, lg_ijk);
  t1 = block_sum(t1, red_d);
  for (int ci = c_begin; ci < c_end; ++ci) {
    const uint32_t rc = card[ci];
    uint32_t w = 1;
    for (int i = k - 1; i > ci; --i) w *= card[i];
    const double a_ij = a_ijk * (double)rc;             // :45
    const double lg_ij = lgamma(a_ij);
    const uint32_t n_cfg = cells / rc;
    double t2 = 0.0;
    for (uint32_t j = threadIdx.x; j < n_cfg; j += blockDim.x) {
      const uint32_t hi = j / w, lo = j - hi * w;
      const int32_t* p = tab + (size_t)hi * rc * w + lo;
      int n_ij = 0;
      for (uint32_t c = 0; c < rc; ++c) n_ij += p[(size_t)c * w];
      t2 += bdeu_term(n_ij, a_ij, lg_ij);
    }
    t2 = block_sum(t2, red_d);
    if (threadIdx.x == 0) out[ci - c_begin] = t1 - t2;
  }
}

// BDeu, exact-order mode: ONE thread per child performs the reference's
// sequence of fp64 additions (all N_ijk terms in ascending key order, then
// all N_ij terms in ascending parent-key order, local_score.h:51-60) on terms
// taken from host-libm tables lut[m] = lgamma(m + a) - lgamma(a), so the sum
// is bit-identical to the reference's.
__device__ __forceinline__ void epilogue_bdeu_exact(const int32_t* tab, uint32_t cells, int k,
                                                    const uint32_t* card, const double* lut_ijk,
                                                    const double* lut_base, size_t lut_stride,
                                                    const uint32_t* lut_ij_ids, int c_begin,
                                                    int c_end, double* out) {
  const int ci = c_begin + (int)threadIdx.x;
  if (ci >= c_end) return;
  double ls = 0.0;
  for (uint32_t c = 0; c < cells; ++c) {
    const int n = tab[c];
    if (n) ls += lut_ijk[n];
  }
  const uint32_t rc = card[ci];
  uint32_t w = 1;
  for (int i = k - 1; i > ci; --i) w *= card[i];
  const double* lut_ij = lut_base + (size_t)lut_ij_ids[ci] * lut_stride;
  const uint32_t n_hi = cells / (rc * w);
  for (uint32_t hi = 0; hi < n_hi; ++hi)
    for (uint32_t lo = 0; lo < w; ++lo) {
      const int32_t* p = tab + (size_t)hi * rc * w + lo;
      int n_ij = 0;
      for (uint32_t c = 0; c < rc; ++c) n_ij += p[(size_t)c * w];
      if (n_ij) ls -= lut_ij[n_ij];
    }
  out[ci - c_begin] = ls;
}

}  // namespace obn
