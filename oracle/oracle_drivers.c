/* oracle/oracle_drivers.c -- TEST INFRASTRUCTURE, not product code.
 *
 * CPU restatement of the two drivers that consume the hot path:
 *   exhaustive_search  src/structure_learning/exhaustive_search.cpp:14-165
 *   pc                 src/structure_learning/pc.cpp:16-86
 * Same header rules as oracle.c (only tests / smoke / bench's CPU legs may
 * load it; parity pinned against the compiled reference).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define OBO_API __attribute__((visibility("default")))

int obo_next_combmask(uint64_t* comb, int n);
int obo_combmask2vec(uint64_t mask, uint32_t* out);
int obo_count_table(const uint8_t* data, size_t n_rows, size_t stride,
                    const uint32_t* cards_all, const uint32_t* var_ids, int k,
                    int64_t* counts);
double obo_bdeu_from_counts(const int64_t* counts, const uint32_t* cards, int k,
                            int child_idx, double ess);
int obo_citest_from_counts(const int64_t* counts, const uint32_t* cards, int k,
                           int idx_x, int idx_y, int kind, double* stat_out,
                           uint64_t* dof_out, double* p_out);

/* ------------------------------------------------------------------------
 * exhaustive_search.cpp:14-148.  parent_masks[v] = bitmask of v's parents in
 * the learned DAG.  Optionally exports the DP tables for parity tests:
 *   best_gs_out[2^n], best_ch_out[2^n]  (exhaustive_search.cpp:115-116)
 *   ls_out[n * 2^n]: the raw local score ls of (child | W \ child) stored at
 *   [child * 2^n + (W \ child)], NaN where not computed. */
OBO_API int obo_exhaustive_search(const uint8_t* data, size_t n_rows, size_t stride,
                                  const uint32_t* cards_all, int n, double ess,
                                  int max_parents, uint64_t* parent_masks,
                                  double* best_gs_out, uint8_t* best_ch_out,
                                  double* ls_out) {
  if (max_parents < 0 || max_parents > n - 1) return -1;   /* :18-19 */
  if (n > 30) return -2; /* oracle memory guard (reference allows n <= 64) */
  const size_t n_combs = (size_t)1 << n;
  double* best_gs = (double*)calloc(n_combs, sizeof(double));
  uint8_t* best_ch = (uint8_t*)calloc(n_combs, 1);
  double* bl_score = (double*)calloc(n_combs * (size_t)n, sizeof(double));
  uint64_t* bl_ps = (uint64_t*)calloc(n_combs * (size_t)n, sizeof(uint64_t));
  int64_t* counts = NULL;
  size_t counts_cap = 0;
  if (ls_out)
    for (size_t i = 0; i < n_combs * (size_t)n; ++i) ls_out[i] = NAN;
  const int max_varset = max_parents + 1;
  for (int size = 0; size <= n; ++size) {                   /* :47 */
    const int calc_ls = size <= max_varset;                 /* :48 */
    uint64_t W = ((uint64_t)1 << size) - 1;                 /* :51 */
    do {
      uint32_t vars[64], cards[64];
      const int k = obo_combmask2vec(W, vars);
      if (calc_ls) {                                        /* :55-57 */
        size_t cells = 1;
        for (int i = 0; i < k; ++i) { cards[i] = cards_all[vars[i]]; cells *= cards[i]; }
        if (cells > counts_cap) {
          free(counts);
          counts = (int64_t*)malloc(cells * sizeof(int64_t));
          counts_cap = cells;
        }
        obo_count_table(data, n_rows, stride, cards_all, vars, k, counts);
      }
      double bgs = 0;
      int bch = -1;
      for (int ci = 0; ci < k; ++ci) {                      /* :61 */
        const uint32_t child = vars[ci];
        const uint64_t pa = W - ((uint64_t)1 << child);
        double ls = -INFINITY;
        if (calc_ls) {
          ls = obo_bdeu_from_counts(counts, cards, k, ci, ess);
          if (ls_out) ls_out[(size_t)child * n_combs + pa] = ls;
        }
        double bls = ls;
        uint64_t bps = pa;
        for (int pi = 0; pi < k; ++pi) {                    /* :75-87 */
          if (pi == ci) continue;
          const uint64_t sub = pa - ((uint64_t)1 << vars[pi]);
          const double s = bl_score[(size_t)child * n_combs + sub];
          if (s >= bls) { bls = s; bps = bl_ps[(size_t)child * n_combs + sub]; }
        }
        bl_score[(size_t)child * n_combs + pa] = bls;
        bl_ps[(size_t)child * n_combs + pa] = bps;
        const double gs = best_gs[pa] + bls;                /* :89-94 */
        if (bgs >= 0 || gs > bgs) { bgs = gs; bch = (int)child; }
      }
      best_gs[W] = bgs;
      best_ch[W] = (uint8_t)bch;
    } while (obo_next_combmask(&W, n));                     /* :129 */
  }
  for (int v = 0; v < n; ++v) parent_masks[v] = 0;
  uint64_t W = (n == 64) ? ~(uint64_t)0 : (((uint64_t)1 << n) - 1);
  for (int i = 0; i < n; ++i) {                             /* :133-148 */
    const uint8_t ch = best_ch[W];
    const uint64_t pa = W - ((uint64_t)1 << ch);
    parent_masks[ch] = bl_ps[(size_t)ch * n_combs + pa];
    W = pa;
  }
  if (best_gs_out) memcpy(best_gs_out, best_gs, n_combs * sizeof(double));
  if (best_ch_out) memcpy(best_ch_out, best_ch, n_combs);
  free(best_gs); free(best_ch); free(bl_score); free(bl_ps); free(counts);
  return 0;
}

/* ------------------------------------------------------------------------
 * gen_comb.h:5-38: k-combinations of items[0..m) in lexicographic order.
 * Iterator form: idx[] holds positions; returns 0 when exhausted. */
static int comb_first(int* idx, int m, int k) {
  if (k > m) return 0;
  for (int i = 0; i < k; ++i) idx[i] = i;
  return 1;
}
static int comb_next(int* idx, int m, int k) {
  int i = k - 1;
  while (i >= 0 && idx[i] == m - k + i) --i;
  if (i < 0) return 0;
  ++idx[i];
  for (int j = i + 1; j < k; ++j) idx[j] = idx[j - 1] + 1;
  return 1;
}

/* adjacency as a dense byte matrix a[u*n+v] = 1 iff u -> v is present
 * (pdag_with_adjmat.cpp:27-36: undirected = both directions). */
static int und_neighbors_without(const uint8_t* a, int n, int v, int excl, uint32_t* out) {
  int m = 0;                                   /* pdag_with_adjmat.cpp:110-117 */
  for (int u = 0; u < n; ++u)
    if (u != excl && a[v * n + u] && a[u * n + v]) out[m++] = (uint32_t)u;
  return m;
}

/* pc.cpp:16-86.  kind 0 chi2 / 1 G2.  adj_out[u*n+v] = 1 iff the returned PDAG
 * has_edge(u, v).  sepset_out (optional) n*n*n bytes: [x][y][var].
 * n_tests_out[depth] (optional, max_cond+1 entries) = CI tests executed. */
OBO_API int obo_pc(const uint8_t* data, size_t n_rows, size_t stride,
                   const uint32_t* cards_all, int n, int kind, double level,
                   int max_cond, int stable, uint8_t* adj_out,
                   uint8_t* sepset_out, uint64_t* n_tests_out) {
  const size_t nn = (size_t)n * n;
  uint8_t* g = (uint8_t*)malloc(nn);
  uint8_t* snap = (uint8_t*)malloc(nn);
  uint8_t* sep = (uint8_t*)calloc(nn * (size_t)n, 1);
  uint32_t* neigh = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
  int64_t* counts = NULL;
  size_t counts_cap = 0;
  for (int u = 0; u < n; ++u)
    for (int v = 0; v < n; ++v) g[u * n + v] = (u != v);    /* complete graph */
  for (int k = 0; k <= max_cond; ++k) {                     /* pc.cpp:22 */
    memcpy(snap, g, nn);
    const uint8_t* gref = stable ? snap : g;                /* :24-25 */
    uint64_t tests = 0;
    for (int x = 0; x < n; ++x)
      for (int y = x + 1; y < n; ++y) {
        if (!(g[x * n + y] && g[y * n + x])) continue;      /* :30 */
        int removed = 0;
        for (int dir = 0; dir < 2 && !removed; ++dir) {     /* :31 */
          const int u = dir ? y : x, v = dir ? x : y;
          const int m = und_neighbors_without(gref, n, u, v, neigh);
          int idx[64];
          if (k > 62 || !comb_first(idx, m, k)) continue;
          do {
            uint32_t vars[64], cards[64];
            for (int i = 0; i < k; ++i) vars[i] = neigh[idx[i]];
            vars[k] = (uint32_t)u; vars[k + 1] = (uint32_t)v;
            const int kk = k + 2;
            for (int i = 1; i < kk; ++i) {                  /* sort, :37 */
              uint32_t t = vars[i]; int j = i - 1;
              while (j >= 0 && vars[j] > t) { vars[j + 1] = vars[j]; --j; }
              vars[j + 1] = t;
            }
            size_t cells = 1; int ix = -1, iy = -1;
            for (int i = 0; i < kk; ++i) {
              cards[i] = cards_all[vars[i]]; cells *= cards[i];
              if ((int)vars[i] == u) ix = i;
              if ((int)vars[i] == v) iy = i;
            }
            if (cells > counts_cap) {
              free(counts);
              counts = (int64_t*)malloc(cells * sizeof(int64_t));
              counts_cap = cells;
            }
            obo_count_table(data, n_rows, stride, cards_all, vars, kk, counts);
            double st, p; uint64_t dof;
            obo_citest_from_counts(counts, cards, kk, ix, iy, kind, &st, &dof, &p);
            ++tests;
            if (p >= level) {                               /* citest.h:147 */
              g[x * n + y] = g[y * n + x] = 0;              /* pc.cpp:40 */
              for (int i = 0; i < k; ++i) {
                sep[((size_t)x * n + y) * n + neigh[idx[i]]] = 1;
                sep[((size_t)y * n + x) * n + neigh[idx[i]]] = 1;
              }
              removed = 1;
              break;
            }
          } while (comb_next(idx, m, k));
        }
      }
    if (n_tests_out) n_tests_out[k] = tests;
  }
  /* orient_colliders, pc.cpp:54-74 */
  for (int y = 0; y < n; ++y) {
    int changed = 1;
    while (changed) {
      changed = 0;
      const int m = und_neighbors_without(g, n, y, -1, neigh);
      for (int i = 0; i + 1 < m; ++i)
        for (int j = i + 1; j < m; ++j) {
          const int x = (int)neigh[i], z = (int)neigh[j];
          if (g[x * n + z] || g[z * n + x]) continue;       /* is_adjacent */
          if (!sep[((size_t)x * n + z) * n + y]) {
            /* orient_edge(x,y): only if still undirected (adjmat.cpp:66-69) */
            if (g[x * n + y] && g[y * n + x]) g[y * n + x] = 0;
            if (g[z * n + y] && g[y * n + z]) g[y * n + z] = 0;
            changed = 1;
          }
        }
    }
  }
  memcpy(adj_out, g, nn);
  if (sepset_out) memcpy(sepset_out, sep, nn * (size_t)n);
  free(g); free(snap); free(sep); free(neigh); free(counts);
  return 0;
}
