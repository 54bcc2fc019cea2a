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

enum Mode : int { MODE_COUNTS = 0, MODE_CI = 1, MODE_BDEU_ONE = 2, MODE_BDEU_ALL = 3 };
enum Tier : int { TIER_SMEM = 0, TIER_GLOBAL = 1, TIER_CHUNK = 2 };

// "no test of this edge passed" in the device-side first[] array of a PC round: the array is
// initialised with cudaMemsetAsync(0x7f), so every value >= kFirstNone means none
constexpr int32_t kFirstNone = 0x7f000000;

struct BatchArgs {
  // dataset
  const uint8_t* data;      // column-major codes, column v at data + v*stride
  size_t stride;            // bytes between columns (multiple of 128)
  const uint8_t* data4;     // the same matrix with two rows per byte for the columns of <= 16 states (null: none);
  size_t stride4;           // byte j of 16-byte chunk c of a column = row 32c + j | row 32c + 16 + j << 4
  uint32_t n_rows;
  // batch (device copies of the CSR arrays)
  const uint32_t* desc;     // [offsets[n]]  var_id | (card-1) << 24, ascending var ids per tuple
  const uint4* hdr;         // per launch tuple: {tuple id, desc offset, #vars, #cells}
  const uint2* roles;       // per launch tuple: BDEU_ONE {child, -}; unused otherwise
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
  uint32_t chunk_cells;         // TIER_CHUNK: cells of one shared-memory chunk of the table
  uint32_t row_splits;          // TIER_CHUNK: splits = chunks * row_splits, CTA (chunk, row range) = (split / row_splits, split % row_splits)
  // tables in global memory (TIER_CHUNK / TIER_GLOBAL), CI mode: the epilogue is its own launch of
  // epi_ctas CTAs per tuple (ci_epilogue_kernel); each takes a slice range and leaves partial (stat, dof)
  // per member in `partials`, the last one to arrive sums them in CTA order
  uint32_t epi_ctas;            // 0: epilogue inside count_kernel (last CTA of the tuple)
  double* partials;             // [launch tuple][kGroupMembers][epi_ctas] {stat, dof as double bits}
  uint32_t wstride;             // ints of epilogue scratch per warp (rx + ry, plus rx * ry for tables in global memory)
  // MODE_CI: the launch tuple is a GROUP -- the CI tests of the batch that are asked over the same
  // variable set share ONE table (counted once, in reference key order); member m of group q is the
  // test members[m].x whose x is digit (members[m].y & 0xff) and y digit (members[m].y >> 8) of the
  // table, every other digit conditions (obn_ci_groups.inl)
  const uint32_t* mem_off;  // [groups + 1], indexed by group id (hdr.x)
  const uint2* members;
  // PC rounds (obn_pc_round): the decision p >= level (citest.h:147) is taken on the device and only
  // the first passing test of every edge leaves it: first[slot_of_test[t]] = min(t) over passing tests
  const uint32_t* slot_of_test;   // null unless a PC round
  int32_t* first;
  double level;
  // Chained PC rounds (obn_pc_round_submit, chain != 0): first[] of the round submitted before this one, over the
  // same edge slots.  A test whose edge that round decided is skipped -- the serial reference had stopped at that
  // candidate (pc.cpp:40-42) -- and a table none of whose tests is left is not counted at all.  The host enqueues
  // this round before it has seen those results, so the device never waits for the host between two rounds.
  const int32_t* prev_first;      // null unless chained
  unsigned long long* skipped;    // [4]: tests skipped, sum of their tuple sizes, tables skipped, sum of their tuple sizes
};

// ---------------------------------------------------------------------------
// epilogues.  `tab` is the dense int32 table in REFERENCE KEY ORDER (first variable most
// significant, contingency_table.h:76-85; generic pointer: shared or global).  All threads of the
// CTA participate.
// ---------------------------------------------------------------------------

// chi-square / G-square with per-slice observed-margin dof and Yates (citest.h:86-137).
// The reference groups the cells by slice_key = key - x_state*w_x - y_state*w_y (citest.h:77-78);
// here the same slices are walked directly in the dense table.  With x at digit px and y at digit
// py (px < py: citest.h:48 orders them and the table's digits ascend), a key splits as
//     key = hi * (rx*wx) + xs * wx + mid * (ry*wy) + ys * wy + lo
// (hi: digits above x, mid: between x and y, lo: below y), so slice s = (hi, mid, lo) starts at
// hi*(rx*wx) + mid*(ry*wy) + lo and its cell (xs, ys) sits xs*wx + ys*wy further on -- laid out
// x_state * ry + y_state exactly like the reference's slice vector (citest.h:83).  One table can
// therefore serve every (x, y | rest) question over its variable set.
struct CiGeom {
  uint32_t rx, ry, wx, wy;
  uint32_t n_lo, n_mid, mid_stride, hi_stride, n_slices;
  __device__ __forceinline__ CiGeom(uint32_t cells, uint32_t rx_, uint32_t ry_, uint32_t wx_, uint32_t wy_)
      : rx(rx_), ry(ry_), wx(wx_), wy(wy_) {
    n_lo = wy;
    mid_stride = ry * wy;
    n_mid = wx / mid_stride;
    hi_stride = rx * wx;
    n_slices = cells / (rx * ry);
  }
  __device__ __forceinline__ uint32_t base(uint32_t s) const {
    if (n_lo == 1u && n_mid == 1u) return s * hi_stride;      // x, y are the two fastest digits
    const uint32_t t = s / n_lo, lo = s - t * n_lo;
    const uint32_t hi = t / n_mid, mid = t - hi * n_mid;
    return hi * hi_stride + mid * mid_stride + lo;
  }
};

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
__device__ __forceinline__ void ci_slices_small(const int32_t* tab, const CiGeom& g, int kind, double& stat,
                                                unsigned long long& dof, uint32_t s_begin, uint32_t s_end) {
  const uint32_t rx = g.rx, ry = g.ry;
  for (uint32_t s = s_begin + threadIdx.x; s < s_end; s += blockDim.x) {
    const int32_t* t = tab + g.base(s);
    int o[4][4], rs[4] = {0, 0, 0, 0}, cs[4] = {0, 0, 0, 0};
    int tot = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int v = ((uint32_t)i < rx && (uint32_t)j < ry) ? t[i * g.wx + j * g.wy] : 0;
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

// Any slice shape: one warp per slice, margins in per-warp scratch (rx + ry ints).
// STAGE (tables in global memory): the slice is first gathered into the warp's scratch (rx * ry more
// ints) with all its loads in flight at once -- walking a slice cell by cell through L2 is a chain of
// dependent round trips, which made the last CTA of a 400 k-cell table the longest thing in a PC run.
template <bool STAGE>
__device__ __forceinline__ void ci_slices_any(const int32_t* tab, const CiGeom& g, int kind, int32_t* wscratch,
                                              double& stat, unsigned long long& dof, uint32_t s_begin, uint32_t s_end) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const uint32_t rx = g.rx, ry = g.ry, sl = rx * ry;
  int32_t* rs = wscratch;        // [rx]
  int32_t* cs = wscratch + rx;   // [ry]
  int32_t* buf = wscratch + rx + ry;   // [rx * ry], STAGE only
  const uint32_t wx = STAGE ? ry : g.wx, wy = STAGE ? 1u : g.wy;
  for (uint32_t s = s_begin + warp; s < s_end; s += nw) {
    const int32_t* t = tab + g.base(s);
    if (STAGE) {
      for (uint32_t c = lane; c < sl; c += 32) {
        const uint32_t i = c / ry, j = c - i * ry;
        buf[c] = __ldcg(t + (size_t)i * g.wx + (size_t)j * g.wy);
      }
      __syncwarp();
      t = buf;
    }
    int tot = 0, robs = 0, cobs = 0;
    for (uint32_t i = lane; i < rx; i += 32) {
      int a = 0;
      for (uint32_t j = 0; j < ry; ++j) a += t[i * wx + j * wy];
      rs[i] = a;
      tot += a;
      robs += (a > 0);
    }
    for (uint32_t j = lane; j < ry; j += 32) {
      int a = 0;
      for (uint32_t i = 0; i < rx; ++i) a += t[i * wx + j * wy];
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
        if (ri > 0 && cj > 0) stat += ci_cell((double)t[i * wx + j * wy], ri, cj, n_total, yates, kind);
      }
    }
    __syncwarp();
  }
}

// Statistic and dof of (x at digit px, y at digit py | every other digit) from the table; the
// result is valid in thread 0.
template <bool GLOBAL_TABLE>
__device__ __forceinline__ void epilogue_ci(const int32_t* tab, const CiGeom& g, int kind, int32_t* wscratch,
                                            double* red_d, unsigned long long* red_u, double& stat_out,
                                            unsigned long long& dof_out, uint32_t s_begin = 0, uint32_t s_end = 0xffffffffu) {
  double stat = 0.0;
  unsigned long long dof = 0;
  if (s_end > g.n_slices) s_end = g.n_slices;             // (default: every slice)
  if (g.rx <= 4 && g.ry <= 4) ci_slices_small(tab, g, kind, stat, dof, s_begin, s_end);
  else ci_slices_any<GLOBAL_TABLE>(tab, g, kind, wscratch, stat, dof, s_begin, s_end);
  stat_out = block_sum(stat, red_d);
  dof_out = block_sum(dof, red_u);
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
