// obn_device.cuh -- device-side helpers shared by the counting kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace obn {

constexpr int kMaxTupleVars = 32;   // a dense table has <= 2^31 cells => <= 31 non-trivial vars

// ---------------------------------------------------------------------------
// chi-square survival function p = 1 - P(dof/2, x/2), the arithmetic of
// pchisq (reference include/citest/citest.h:17-21, which calls Boost.Math).
// Power series for x < a+1, modified-Lentz continued fraction otherwise.
// ---------------------------------------------------------------------------
// Returns the UPPER tail Q = 1 - P.  In the continued-fraction branch Q is returned as computed
// (the reference's 1 - cdf round trip, citest.h:20, flushes p below 2^-53 to 0; keeping Q differs
// from it by at most 1.2e-16 absolute and cannot move a decision p >= level).
__device__ __forceinline__ double gamma_q(double a, double x) {
  if (!(x > 0.0)) return 1.0;
  if (isinf(x)) return 0.0;
  const double lead = exp(-x + a * log(x) - lgamma(a));
  if (x < a + 1.0) {
    double ap = a, sum = 1.0 / a, del = sum;
    for (int n = 0; n < 200000; ++n) {
      ap += 1.0;
      del *= x / ap;
      sum += del;
      if (fabs(del) < fabs(sum) * 1e-17) break;
    }
    const double p = sum * lead;
    return p > 1.0 ? 0.0 : 1.0 - p;
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
  const double q = lead * h;
  return q > 1.0 ? 1.0 : q;
}

__device__ __forceinline__ double pchisq(double x, unsigned long long dof) {
  if (dof == 0) return (x == 0.0 ? 1.0 : 0.0);
  return gamma_q(0.5 * (double)dof, 0.5 * x);
}

// ---------------------------------------------------------------------------
// block-wide sums by warp shuffles (+ one shared-memory hop between warps)
// ---------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// `red` must hold 32 elements of T; result valid in every thread.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();               // protect `red` from a previous use
  if (lane == 0) red[w] = v;
  __syncthreads();
  T r = (lane < nw) ? red[lane] : T(0);
  r = warp_sum(r);
  return r;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// 16-byte read-only load of 16 category codes
__device__ __forceinline__ uint4 ld_codes16(const uint8_t* p) {
  return __ldg(reinterpret_cast<const uint4*>(p));
}

}  // namespace obn
