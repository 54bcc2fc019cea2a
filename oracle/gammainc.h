/* oracle/gammainc.h -- TEST INFRASTRUCTURE, not product code.
 *
 * Regularised lower incomplete gamma P(a, x) in plain C, used by
 *   (1) oracle/oracle.c            (the CPU restatement of the hot path) and
 *   (2) oracle/ref_shim/boost/...  (the stand-in for Boost.Math that lets the
 *       UNMODIFIED reference compile here; Boost is absent from this image).
 *
 * The reference calls boost::math::cdf(chi_squared(dof), x) at
 * include/citest/citest.h:17-21, i.e. P(dof/2, x/2).  Boost (unpinned distro
 * version, Dockerfile:39) is a third-party dependency that is not under
 * /root/reference; this file restates the published textbook algorithm
 * (power series for x < a+1, modified-Lentz continued fraction for the upper
 * tail otherwise; Abramowitz & Stegun 6.5.29 / 6.5.31).
 */
#ifndef OBN_ORACLE_GAMMAINC_H
#define OBN_ORACLE_GAMMAINC_H
#include <math.h>

static inline double obo_gamma_p(double a, double x) {
  if (!(x > 0.0)) return 0.0;
  if (isinf(x)) return 1.0;
  const double lg = lgamma(a);
  const double lead = exp(-x + a * log(x) - lg);
  if (x < a + 1.0) {
    double ap = a, sum = 1.0 / a, del = sum;
    for (int n = 0; n < 200000; ++n) {
      ap += 1.0;
      del *= x / ap;
      sum += del;
      if (fabs(del) < fabs(sum) * 1e-17) break;
    }
    double p = sum * lead;
    return p > 1.0 ? 1.0 : p;
  }
  const double tiny = 1e-300;
  double b = x + 1.0 - a, c = 1.0 / tiny, d = 1.0 / b, h = d;
  for (int i = 1; i < 200000; ++i) {
    const double an = -(double)i * ((double)i - a);
    b += 2.0;
    d = an * d + b;
    if (fabs(d) < tiny) d = tiny;
    c = b + an / c;
    if (fabs(c) < tiny) c = tiny;
    d = 1.0 / d;
    const double del = d * c;
    h *= del;
    if (fabs(del - 1.0) < 1e-16) break;
  }
  double q = lead * h;
  if (q > 1.0) q = 1.0;
  return 1.0 - q;
}

/* citest.h:17-21 */
static inline double obo_pchisq(double x, unsigned long long dof) {
  if (dof == 0) return (x == 0.0 ? 1.0 : 0.0);
  return 1.0 - obo_gamma_p(0.5 * (double)dof, 0.5 * x);
}
#endif
