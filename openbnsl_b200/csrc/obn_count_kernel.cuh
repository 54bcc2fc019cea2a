// obn_count_kernel.cuh -- the counting kernel proper.
//
//   TIER_PRIVATE  every thread owns a private column of 16-bit counters in
//                 shared memory (hist16[cell][thread]); increments are plain
//                 LDS/IADD/STS with no atomics and no bank conflicts (thread t
//                 always hits bank t/2).  Reduced over threads into an int32
//                 table before the epilogue.
//   TIER_SHARED   one int32 histogram per CTA, shared-memory atomics.
//   TIER_GLOBAL   int32 table in global scratch (L2 atomics) for tables that
//                 do not fit shared memory.
//
// Keys are built from 16-byte vector loads (16 rows per column per load) with
// packed arithmetic: KW=8 keeps four 8-bit keys per register (tables of <=256
// cells), KW=16 two 16-bit keys, KW=32 one.
#pragma once
#include "obn_kernels.cuh"

namespace obn {

struct CtaTuple {            // per-CTA tuple description in shared memory
  const uint8_t* col[kMaxTupleVars];   // column pointers in KERNEL digit order
  uint32_t card[kMaxTupleVars];        // cards in kernel digit order
  uint32_t ref_card[kMaxTupleVars];    // cards in ascending-var (reference) order
  int k;
  uint32_t cells;
  uint32_t rx, ry;
  int child_pos;             // BDEU_ONE: position of the child in reference order
  unsigned int ticket;
};

template <int KW>
struct KeyVec;               // keys of 16 consecutive rows
template <>
struct KeyVec<8> {
  uint32_t w[4];
  __device__ __forceinline__ void zero() { w[0] = w[1] = w[2] = w[3] = 0; }
  __device__ __forceinline__ void push(const uint4& c, uint32_t card) {
    w[0] = w[0] * card + c.x; w[1] = w[1] * card + c.y;
    w[2] = w[2] * card + c.z; w[3] = w[3] * card + c.w;
  }
  __device__ __forceinline__ uint32_t get(int r) const { return (w[r >> 2] >> ((r & 3) * 8)) & 0xffu; }
};
template <>
struct KeyVec<16> {
  uint32_t w[8];
  __device__ __forceinline__ void zero() {
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = 0;
  }
  __device__ __forceinline__ void push(const uint4& c, uint32_t card) {
    const uint32_t s[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      w[2 * i] = w[2 * i] * card + __byte_perm(s[i], 0, 0x4140);       // bytes 0,1 -> halves
      w[2 * i + 1] = w[2 * i + 1] * card + __byte_perm(s[i], 0, 0x4342);  // bytes 2,3 -> halves
    }
  }
  __device__ __forceinline__ uint32_t get(int r) const { return (w[r >> 1] >> ((r & 1) * 16)) & 0xffffu; }
};
template <>
struct KeyVec<32> {
  uint32_t w[16];
  __device__ __forceinline__ void zero() {
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = 0;
  }
  __device__ __forceinline__ void push(const uint4& c, uint32_t card) {
    const uint32_t s[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = w[i] * card + ((s[i >> 2] >> ((i & 3) * 8)) & 0xffu);
  }
  __device__ __forceinline__ uint32_t get(int r) const { return w[r]; }
};

template <int KW, int TIER>
__global__ void __launch_bounds__(256) count_kernel(const BatchArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ CtaTuple T;
  __shared__ double red_d[32];
  __shared__ unsigned long long red_u[32];

  const int tid = threadIdx.x, nthr = blockDim.x;
  const uint32_t lt = blockIdx.x / A.splits;       // launch-local tuple
  const uint32_t split = blockIdx.x - lt * A.splits;
  const uint32_t t = A.perm[lt];

  // ---- tuple prologue: digit order, column pointers, cells ----------------
  if (tid == 0) {
    const uint64_t o0 = A.offsets[t], o1 = A.offsets[t + 1];
    const int k = (int)(o1 - o0);
    T.k = k;
    uint32_t xa = 0xffffffffu, ya = 0xffffffffu;
    if (A.mode == MODE_CI) {
      xa = A.role_a[t]; ya = A.role_b[t];
      if (xa > ya) { const uint32_t s = xa; xa = ya; ya = s; }   // citest.h:48
    }
    int n = 0;
    uint32_t cells = 1;
    T.child_pos = -1;
    for (int i = 0; i < k; ++i) {
      const uint32_t v = A.var_ids[o0 + i];
      const uint32_t c = A.cards[v];
      T.ref_card[i] = c;
      cells *= c;
      if (A.mode == MODE_BDEU_ONE && v == A.role_a[t]) T.child_pos = i;
      if (v == xa || v == ya) continue;
      T.col[n] = A.data + (size_t)v * A.stride;
      T.card[n] = c;
      ++n;
    }
    if (A.mode == MODE_CI) {
      T.col[n] = A.data + (size_t)xa * A.stride; T.card[n] = A.cards[xa]; ++n;
      T.col[n] = A.data + (size_t)ya * A.stride; T.card[n] = A.cards[ya]; ++n;
      T.rx = A.cards[xa]; T.ry = A.cards[ya];
    }
    T.cells = cells;
  }
  __syncthreads();
  const int k = T.k;
  const uint32_t cells = T.cells;

  // ---- shared-memory carve-up ---------------------------------------------
  int32_t* tab;                     // dense int32 table the epilogue reads
  uint16_t* hist16 = nullptr;       // TIER_PRIVATE only
  int32_t* wscratch;                // per-warp epilogue scratch
  if (TIER == TIER_PRIVATE) {
    tab = reinterpret_cast<int32_t*>(smem_raw);
    const uint32_t tab_bytes = (cells * 4u + 15u) & ~15u;
    hist16 = reinterpret_cast<uint16_t*>(smem_raw + tab_bytes);
    wscratch = reinterpret_cast<int32_t*>(hist16);       // aliased: hist16 is dead by then
    const uint32_t n16 = (cells * (uint32_t)nthr * 2u) / 16u;
    uint4* z = reinterpret_cast<uint4*>(hist16);
    for (uint32_t i = tid; i < n16; i += nthr) z[i] = make_uint4(0, 0, 0, 0);
  } else if (TIER == TIER_SHARED) {
    tab = reinterpret_cast<int32_t*>(smem_raw);
    const uint32_t tab_bytes = (cells * 4u + 15u) & ~15u;
    wscratch = reinterpret_cast<int32_t*>(smem_raw + tab_bytes);
    for (uint32_t i = tid; i < cells; i += nthr) tab[i] = 0;
  } else {
    tab = A.scratch + A.scratch_off[lt];
    wscratch = reinterpret_cast<int32_t*>(smem_raw);
  }
  __syncthreads();

  // ---- main loop over this split's 16-row vectors -------------------------
  const uint32_t n_full = A.n_rows >> 4;
  const uint32_t v_begin = split * A.vec_per_split;
  uint32_t v_end = v_begin + A.vec_per_split;
  if (v_end > n_full) v_end = n_full;
  uint16_t* my16 = hist16 + tid;
  for (uint32_t v = v_begin + tid; v < v_end; v += nthr) {
    KeyVec<KW> key;
    key.zero();
    const size_t off = (size_t)v << 4;
    for (int i = 0; i < k; ++i) key.push(ld_codes16(T.col[i] + off), T.card[i]);
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const uint32_t cell = key.get(r);
      if (TIER == TIER_PRIVATE) my16[cell * nthr] += 1;
      else if (TIER == TIER_SHARED) atomicAdd(&tab[cell], 1);
      else atomicAdd(&tab[cell], 1);
    }
  }
  // tail rows (n_rows % 16) belong to the last split; one thread, scalar.
  // TIER_PRIVATE: thread 0 adds them to ITS private counters.
  if (split == (uint32_t)A.splits - 1 && tid == 0) {
    for (uint32_t r = n_full << 4; r < A.n_rows; ++r) {
      uint32_t cell = 0;
      for (int i = 0; i < k; ++i) cell = cell * T.card[i] + T.col[i][r];
      if (TIER == TIER_PRIVATE) my16[cell * nthr] += 1;
      else atomicAdd(&tab[cell], 1);
    }
  }
  __syncthreads();

  // ---- TIER_PRIVATE: reduce the per-thread counters into `tab` -------------
  if (TIER == TIER_PRIVATE) {
    const int lane = tid & 31, warp = tid >> 5, nw = nthr >> 5;
    const uint32_t words = (uint32_t)nthr >> 1;          // u32 words per cell row
    for (uint32_t c = warp; c < cells; c += nw) {
      const uint32_t* row = reinterpret_cast<const uint32_t*>(hist16 + (size_t)c * nthr);
      int s = 0;
      for (uint32_t i = lane; i < words; i += 32) {
        const uint32_t x = row[i];
        s += (int)(x & 0xffffu) + (int)(x >> 16);
      }
      s = warp_sum(s);
      if (lane == 0) tab[c] = s;
    }
    __syncthreads();
  }

  // ---- row splits: merge through global scratch, last CTA continues -------
  if (A.splits > 1 || TIER == TIER_GLOBAL) {
    int32_t* g = A.scratch + A.scratch_off[lt];
    if (TIER != TIER_GLOBAL) {
      for (uint32_t c = tid; c < cells; c += nthr) {
        const int v = tab[c];
        if (v) atomicAdd(&g[c], v);
      }
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) T.ticket = atomicAdd(&A.tickets[lt], 1u);
    __syncthreads();
    if (T.ticket != (unsigned)A.splits - 1) return;
    __threadfence();
    if (TIER != TIER_GLOBAL) {
      for (uint32_t c = tid; c < cells; c += nthr) {
        tab[c] = __ldcg(&g[c]);
        g[c] = 0;
      }
      if (tid == 0) A.tickets[lt] = 0;
      __syncthreads();
    }
  }

  // ---- fused epilogue -------------------------------------------------------
  if (A.mode == MODE_CI) {
    epilogue_ci(tab, cells, T.rx, T.ry, A.kind, wscratch + (tid >> 5) * (T.rx + T.ry),
                red_d, red_u, &A.stat[t], &A.dof[t]);
  } else if (A.mode == MODE_BDEU_ONE || A.mode == MODE_BDEU_ALL) {
    const int c0 = (A.mode == MODE_BDEU_ONE) ? T.child_pos : 0;
    const int c1 = (A.mode == MODE_BDEU_ONE) ? T.child_pos + 1 : k;
    double* out = (A.mode == MODE_BDEU_ONE) ? &A.scores[t] : &A.scores[A.offsets[t]];
    if (A.lut) {
      epilogue_bdeu_exact(tab, cells, k, T.ref_card, A.lut + (size_t)A.lut_ijk[t] * A.lut_stride,
                          A.lut, A.lut_stride, A.lut_ij + A.offsets[t], c0, c1, out);
    } else {
      epilogue_bdeu_fast(tab, cells, k, T.ref_card, A.ess, c0, c1, red_d, out);
    }
  } else {
    int32_t* out = A.counts_out + A.counts_off[t];
    for (uint32_t c = tid; c < cells; c += nthr) out[c] = tab[c];
  }

  if (TIER == TIER_GLOBAL) {      // leave the scratch table zeroed for the next launch
    __syncthreads();
    for (uint32_t c = tid; c < cells; c += nthr) tab[c] = 0;
    if (tid == 0) A.tickets[lt] = 0;
  }
}

// p = pchisq(stat, dof) for a batch (citest.h:17-21,139)
__global__ void pvalue_kernel(const double* stat, const unsigned long long* dof, double* p, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = pchisq(stat[i], dof[i]);
}

}  // namespace obn
