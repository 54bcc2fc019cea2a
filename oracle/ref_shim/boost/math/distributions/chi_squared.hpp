// oracle/ref_shim -- TEST INFRASTRUCTURE.  Stand-in for the one Boost.Math
// entry point the reference uses (include/citest/citest.h:17-21), so that the
// UNMODIFIED reference sources compile in an image without Boost.
//
// Besides computing the cdf it records the (dof, x) of the most recent call in
// thread-local variables.  The reference keeps `stat` and `dof` as locals of
// citest<>() and returns only a bool; the recorder is how oracle/ref_harness.cpp
// reads them back without touching reference code.
#pragma once
#include "../../../../gammainc.h"

namespace boost {
namespace math {
struct chi_squared {
  double k;
  explicit chi_squared(double dof) : k(dof) {}
};
namespace obn_shim {
inline thread_local double last_x = -1.0;
inline thread_local double last_k = -1.0;
inline thread_local unsigned long long n_calls = 0;
}  // namespace obn_shim
inline double cdf(const chi_squared& d, double x) {
  obn_shim::last_x = x;
  obn_shim::last_k = d.k;
  ++obn_shim::n_calls;
  return obo_gamma_p(0.5 * d.k, 0.5 * x);
}
}  // namespace math
}  // namespace boost
