// smem_incr.cu -- microbenchmark: how fast can one SM apply histogram
// increments?  Decides the counter layout of the counting kernels.
// Variants (all take the same pseudo-random cell stream, `cells` bins):
//   0 u16 lane-private  [cell][T]            (round-1 first layout)
//   1 u16 lane-private  [cell/2][T][2]       (one 32-bit word per lane)
//   2 u8  lane-private  [cell][T]
//   3 u8  lane-private  [cell/4][T][4]       (one 32-bit word per lane)
//   4 u32 lane-private  [cell][T]
//   5 ATOMS on a warp-private int32 table
//   6 ATOMS on a CTA-shared int32 table
//   7 __match_any_sync aggregated ATOMS on a CTA-shared table
//   8 variant 3 with two independent increments in flight (pairing)
//   9 ATOMS on a CTA-shared table replicated R times, copy = lane & (R-1): word (cell * R + copy), so the
//     lanes of one instruction spread over the banks whatever their cells (R = 32: conflict-free)
// usage: smem_incr <cells> <threads> <ctas_per_sm> <skew_pct> [R]
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

// 16 keys per iteration are bit-fields of 8 LCG words (2 ALU ops per key);
// cells must be a power of two <= 65536.  skew != 0 ANDs two fields (P(bit)=1/4).
struct Keys {
  uint32_t w[8];
  __device__ __forceinline__ void refill(uint32_t& s) {
#pragma unroll
    for (int i = 0; i < 8; ++i) { s = s * 1664525u + 1013904223u; w[i] = s ^ (s >> 15); }
  }
  __device__ __forceinline__ uint32_t get(int u, uint32_t mask, uint32_t skew) const {
    uint32_t f = (w[u >> 1] >> ((u & 1) * 16)) & mask;
    if (skew) f &= (w[(u >> 1) ^ 5] >> (((u & 1) ^ 1) * 16));
    return f;
  }
};

template <int V>
__global__ void k(uint32_t cells, uint32_t iters, uint32_t skew, unsigned long long* out, uint32_t R = 1) {
  extern __shared__ __align__(16) unsigned char sm[];
  const uint32_t tid = threadIdx.x, T = blockDim.x, lane = tid & 31, warp = tid >> 5;
  size_t bytes;
  if (V == 0 || V == 1) bytes = (size_t)((cells + 1) & ~1u) * T * 2;
  else if (V == 2 || V == 3 || V == 8) bytes = (size_t)((cells + 3) & ~3u) * T;
  else if (V == 4) bytes = (size_t)cells * T * 4;
  else if (V == 5) bytes = (size_t)cells * 4 * (T / 32);
  else if (V == 9) bytes = (size_t)cells * 4 * R;
  else bytes = (size_t)cells * 4;
  const uint32_t copy4 = (lane & (R - 1)) * 4, r4 = R * 4;
  for (uint32_t i = tid; i < bytes / 4; i += T) reinterpret_cast<uint32_t*>(sm)[i] = 0;
  __syncthreads();
  uint32_t s = blockIdx.x * 7919u + tid * 104729u + 1u;
  uint16_t* h16 = reinterpret_cast<uint16_t*>(sm);
  uint8_t* h8 = sm;
  uint32_t* h32 = reinterpret_cast<uint32_t*>(sm);
  unsigned flushes = 0;
  const uint32_t mask = cells - 1;
  Keys K;
  for (uint32_t it = 0; it < iters; ++it) {
    K.refill(s);
#pragma unroll
    for (int u = 0; u < 16; u += (V == 8 ? 2 : 1)) {
      const uint32_t c = K.get(u, mask, skew);
      if (V == 0) h16[c * T + tid] += 1;
      else if (V == 1) h16[((c >> 1) * T + tid) * 2 + (c & 1)] += 1;
      else if (V == 2) { uint8_t v = h8[c * T + tid] + 1; h8[c * T + tid] = v; flushes += (v == 0); }
      else if (V == 3) { uint8_t* p = &h8[((c >> 2) * T + tid) * 4 + (c & 3)]; uint8_t v = *p + 1; *p = v; flushes += (v == 0); }
      else if (V == 4) h32[c * T + tid] += 1;
      else if (V == 5) atomicAdd(&h32[warp * cells + c], 1u);
      else if (V == 6) atomicAdd(&h32[c], 1u);
      else if (V == 9) atomicAdd(reinterpret_cast<uint32_t*>(sm + c * r4 + copy4), 1u);
      else if (V == 7) {
        const unsigned m = __match_any_sync(0xffffffffu, c);
        if ((m & ((1u << lane) - 1)) == 0) atomicAdd(&h32[c], (unsigned)__popc(m));
      } else if (V == 8) {
        const uint32_t c1 = K.get(u + 1, mask, skew);
        uint8_t* p0 = &h8[((c >> 2) * T + tid) * 4 + (c & 3)];
        uint8_t* p1 = &h8[((c1 >> 2) * T + tid) * 4 + (c1 & 3)];
        const uint32_t v0 = *p0, v1 = *p1;
        const uint32_t n0 = v0 + 1, n1 = (p0 == p1) ? v0 + 2 : v1 + 1;
        *p0 = (uint8_t)n0;
        *p1 = (uint8_t)n1;
        flushes += ((n0 & 255u) == 0) + ((n1 & 255u) == 0);
      }
    }
  }
  __syncthreads();
  unsigned long long acc = flushes;
  for (uint32_t i = tid; i < bytes / 4; i += T) acc += reinterpret_cast<uint32_t*>(sm)[i];
  if (acc == 0xdeadbeefull) out[0] = acc;
}

template <int V>
static void run(uint32_t cells, int T, int occ, uint32_t skew, uint32_t R = 1) {
  size_t bytes;
  if (V == 0 || V == 1) bytes = (size_t)((cells + 1) & ~1u) * T * 2;
  else if (V == 2 || V == 3 || V == 8) bytes = (size_t)((cells + 3) & ~3u) * T;
  else if (V == 4) bytes = (size_t)cells * T * 4;
  else if (V == 5) bytes = (size_t)cells * 4 * (T / 32);
  else if (V == 9) bytes = (size_t)cells * 4 * R;
  else bytes = (size_t)cells * 4;
  bytes = (bytes + 15) & ~(size_t)15;
  if (bytes > 227 * 1024) { printf("  v%d: smem too large\n", V); return; }
  cudaFuncSetAttribute(k<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  int real_occ = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&real_occ, k<V>, T, bytes);
  if (real_occ < 1) { printf("  v%d: does not fit\n", V); return; }
  int use = occ < real_occ ? occ : real_occ;
  unsigned long long* out;
  cudaMalloc(&out, 8);
  const uint32_t iters = 2000;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<V><<<148 * use, T, bytes>>>(cells, 10, skew, out, R);
  cudaEventRecord(e0);
  k<V><<<148 * use, T, bytes>>>(cells, iters, skew, out, R);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  const double incs = 148.0 * use * T * 16.0 * iters;
  int clk = 0;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  if (V == 9) printf("  R=%u", R);
  printf("  v%d cells=%u T=%d ctas/sm=%d smem=%zuB: %.1f Ginc/s  (%.2f inc/clk/SM at %d MHz nominal) err=%s\n", V, cells, T, use,
         bytes, incs / ms * 1e-6, incs / ms * 1e-6 * 1e9 / (148.0 * clk * 1e3), clk / 1000, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out);
}

int main(int argc, char** argv) {
  const uint32_t cells = argc > 1 ? atoi(argv[1]) : 64;
  const int T = argc > 2 ? atoi(argv[2]) : 256;
  const int occ = argc > 3 ? atoi(argv[3]) : 8;
  const uint32_t skew = argc > 4 ? atoi(argv[4]) : 0;
  printf("cells=%u T=%d occ<=%d skew=%u/128\n", cells, T, occ, skew);
  if (argc > 5) {                       // replication sweep only
    run<6>(cells, T, occ, skew);
    for (uint32_t R = 2; R <= 32; R *= 2) run<9>(cells, T, occ, skew, R);
    return 0;
  }
  run<0>(cells, T, occ, skew); run<1>(cells, T, occ, skew); run<2>(cells, T, occ, skew);
  run<3>(cells, T, occ, skew); run<4>(cells, T, occ, skew); run<5>(cells, T, occ, skew);
  run<6>(cells, T, occ, skew); run<7>(cells, T, occ, skew); run<8>(cells, T, occ, skew);
  return 0;
}
