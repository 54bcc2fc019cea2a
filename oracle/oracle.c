/* oracle/oracle.c -- TEST INFRASTRUCTURE, not product code.
 *
 * CPU restatement, in plain C, of the reference's algorithm for the hot path:
 * contingency counting -> BDeu local score / chi-square & G-square CI test ->
 * exhaustive search and PC.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library, and
 * only as the checker.  Every function cites the reference lines it follows
 * (paths relative to /root/reference).
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks this file against
 * fixtures generated from the unmodified reference compiled here
 * (oracle/build_ref.py -> oracle/_ref, fixtures by tests/golden/make_golden.py),
 * and tests/test_oracle_vs_ref.py against the live reference module when
 * oracle/_ref is present.  Two caveats, both stated in DESIGN.md:
 *   - p-values go through oracle/gammainc.h because Boost.Math is absent;
 *   - the reference sums CI statistics over Z-slices in libstdc++
 *     unordered_map order (citest.h:74,88); here slices are visited in
 *     ascending key order, so `stat` agrees to rounding, not bit for bit.
 *
 * Data layout everywhere: column-major uint8 codes, data[v * stride + row].
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "gammainc.h"

#define OBO_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------
 * include/utils/next_combmask.h:41-46 (Gosper's hack) */
OBO_API int obo_next_combmask(uint64_t* comb, int n) {
  uint64_t c = *comb;
  if (c == 0) return 0;
  uint64_t x = c & (~c + 1);
  uint64_t y = c + x;
  c = (((c & ~y) / x) >> 1) | y;
  *comb = c;
  if (n >= 64) return y != 0; /* reference is limited to n <= 63 in practice */
  return c < ((uint64_t)1 << n);
}

/* include/utils/combmask2vec.h:15-24: ascending bit indices */
OBO_API int obo_combmask2vec(uint64_t mask, uint32_t* out) {
  int k = 0;
  while (mask) {
    out[k++] = (uint32_t)__builtin_ctzll(mask);
    mask &= mask - 1;
  }
  return k;
}

/* ------------------------------------------------------------------------
 * include/base/contingency_table.h:34-56,76-85.
 * Dense counts in reference key order: key = sum row[var_ids[i]] * w_i with
 * the LAST variable least significant.  The reference keeps only observed
 * cells in a std::map; a dense array whose zero cells are skipped by the
 * consumers is the same table.  Returns -1 if var_ids is not ascending
 * (contingency_table.h:37-38; duplicates are "sorted" and therefore accepted
 * by the reference, so they are accepted here too). */
OBO_API int obo_count_table(const uint8_t* data, size_t n_rows, size_t stride,
                            const uint32_t* cards_all, const uint32_t* var_ids,
                            int k, int64_t* counts) {
  size_t w[64];
  size_t mult = 1;
  for (int i = 1; i < k; ++i)
    if (var_ids[i] < var_ids[i - 1]) return -1;
  for (int i = k - 1; i >= 0; --i) {
    w[i] = mult;
    mult *= cards_all[var_ids[i]];
  }
  memset(counts, 0, mult * sizeof(int64_t));
  for (size_t r = 0; r < n_rows; ++r) {
    size_t key = 0;
    for (int i = 0; i < k; ++i) key += (size_t)data[var_ids[i] * stride + r] * w[i];
    ++counts[key];
  }
  return 0;
}

/* ------------------------------------------------------------------------
 * include/score/local_score.h:22-62 (BDeu).  `cards` are the k cardinalities
 * of the table's variables, `child_idx` the child's position in the table.
 * Term order is the reference's: all N_ijk terms in ascending key order, then
 * all N_ij terms in ascending parent-key order (std::map iteration). */
OBO_API double obo_bdeu_from_counts(const int64_t* counts, const uint32_t* cards,
                                    int k, int child_idx, double ess) {
  size_t total = 1;
  for (int i = 0; i < k; ++i) total *= cards[i];
  const size_t child_card = cards[child_idx];
  size_t w_child = 1;
  for (int i = k - 1; i > child_idx; --i) w_child *= cards[i];
  const double a_ijk = ess / (double)total;            /* local_score.h:41 */
  const double a_ij = a_ijk * (double)child_card;      /* :45 */
  const double lg_ijk = lgamma(a_ijk), lg_ij = lgamma(a_ij);
  double ls = 0.0;
  for (size_t key = 0; key < total; ++key) {           /* :51-56 */
    if (counts[key] == 0) continue;
    ls += lgamma((double)counts[key] + a_ijk) - lg_ijk;
  }
  /* parent keys ascending == (hi, lo) ascending with the child digit zeroed */
  const size_t n_hi = total / (w_child * child_card);
  for (size_t hi = 0; hi < n_hi; ++hi) {
    for (size_t lo = 0; lo < w_child; ++lo) {
      int64_t n_ij = 0;
      for (size_t c = 0; c < child_card; ++c)
        n_ij += counts[(hi * child_card + c) * w_child + lo];
      if (n_ij == 0) continue;
      ls -= lgamma((double)n_ij + a_ij) - lg_ij;       /* :57-60 */
    }
  }
  return ls;
}

/* ------------------------------------------------------------------------
 * include/citest/citest.h:73-139.  kind: 0 = chi-square, 1 = G-square.
 * idx_x / idx_y are positions in the table (the reference swaps so that
 * x < y, citest.h:48; the slice layout is x_state * n_y + y_state, :83).
 * Every other variable of the table is a conditioning variable. */
OBO_API int obo_citest_from_counts(const int64_t* counts, const uint32_t* cards,
                                   int k, int idx_x, int idx_y, int kind,
                                   double* stat_out, uint64_t* dof_out,
                                   double* p_out) {
  if (idx_x == idx_y) return -1;
  if (idx_x > idx_y) { int t = idx_x; idx_x = idx_y; idx_y = t; }
  size_t w[64], total = 1;
  for (int i = k - 1; i >= 0; --i) { w[i] = total; total *= cards[i]; }
  const size_t n_x = cards[idx_x], n_y = cards[idx_y];
  const size_t wx = w[idx_x], wy = w[idx_y];
  double* row_sum = (double*)malloc(sizeof(double) * (n_x + n_y));
  double* col_sum = row_sum + n_x;
  double stat = 0.0;
  uint64_t dof = 0;
  for (size_t base = 0; base < total; ++base) {
    /* slice keys = keys whose x and y digits are zero */
    if ((base / wx) % n_x != 0 || (base / wy) % n_y != 0) continue;
    double n_total = 0;
    for (size_t i = 0; i < n_x; ++i) row_sum[i] = 0;
    for (size_t j = 0; j < n_y; ++j) col_sum[j] = 0;
    for (size_t i = 0; i < n_x; ++i)
      for (size_t j = 0; j < n_y; ++j) {
        const double obs = (double)counts[base + i * wx + j * wy];
        row_sum[i] += obs; col_sum[j] += obs; n_total += obs;   /* :89-98 */
      }
    if (n_total == 0) continue;                                  /* :100 */
    size_t r_obs = 0, c_obs = 0;
    for (size_t i = 0; i < n_x; ++i) r_obs += row_sum[i] > 0;
    for (size_t j = 0; j < n_y; ++j) c_obs += col_sum[j] > 0;
    const size_t d = (r_obs - 1) * (c_obs - 1);                 /* :109 */
    const int yates = (d == 1);
    dof += d;
    for (size_t i = 0; i < n_x; ++i) {
      if (!(row_sum[i] > 0)) continue;
      for (size_t j = 0; j < n_y; ++j) {
        if (!(col_sum[j] > 0)) continue;
        double obs = (double)counts[base + i * wx + j * wy];
        const double ex = row_sum[i] * col_sum[j] / n_total;    /* :116 */
        if (yates) {                                            /* :120-125 */
          const double diff = ex - obs;
          const double dir = (diff > 0) ? 1.0 : (diff < 0) ? -1.0 : 0.0;
          obs += fmin(0.5, fabs(diff)) * dir;
        }
        if (kind == 0) {
          const double df = obs - ex;
          stat += df * df / ex;                                 /* :127-129 */
        } else {
          if (obs > 0) stat += 2.0 * obs * log(obs / ex);       /* :130-131 */
        }
      }
    }
  }
  free(row_sum);
  *stat_out = stat;
  *dof_out = dof;
  *p_out = obo_pchisq(stat, dof);                                /* :139 */
  return 0;
}

OBO_API double obo_pchisq_export(double x, uint64_t dof) { return obo_pchisq(x, dof); }
