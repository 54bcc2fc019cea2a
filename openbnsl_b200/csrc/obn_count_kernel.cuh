// obn_count_kernel.cuh -- the counting kernel proper (sm_100a).
//
// One CTA owns one (tuple, row-split).  Every row of the split is turned into a
// mixed-radix cell index and counted into ONE int32 table per CTA:
//
//   TIER_SMEM    table in shared memory, incremented with `atomicAdd(p, 1)`,
//                which ptxas lowers to ATOMS.POPC.INC.32 -- the hardware form of
//                warp-aggregated increments (lanes that hit the same cell are
//                popcounted and applied as one add).  Measured on B200
//                (benchmarks/micro/smem_incr.cu, profiles/r1_micro_smem_incr.txt):
//                16 increments/clk/SM, against 6-9 for lane-private LDS/IADD/STS
//                counters and 0.5-3 for __match_any_sync aggregation in software.
//   TIER_CHUNK   tables of up to 32 shared-memory-fulls: the cell range is cut into chunks, CTA
//                (chunk, row range) scans its rows, builds the full key and counts only the rows
//                whose cell falls into its chunk (in shared memory, same increments), then adds
//                the chunk into the global table.  Every row is keyed once per chunk but counted
//                once; hot cells stay in shared memory, where POPC.INC merges them.
//   TIER_GLOBAL  table in global scratch (L2 reductions), for tables beyond that.  Same-address
//                reductions serialise in L2: on skewed data (a PC frontier's 5-variable tables)
//                39 tables of 55 k - 420 k cells took 5.4 ms this way -- a quarter of a C5 run --
//                which is why TIER_CHUNK exists.
//
// Category codes are read as 16-byte vectors (16 rows of one column per load,
// coalesced: consecutive lanes read consecutive 16-byte chunks) and combined
// with packed arithmetic: KW=8 keeps four 8-bit keys per register (tables of
// <= 256 cells), KW=16 two 16-bit keys, KW=32 one.  With SCALE4 the packed key
// is pre-multiplied by 4 so that an extracted field IS the byte offset of the
// cell (one ALU op + one ATOMS per row).
//
// Rows beyond n_rows up to the next multiple of 16 are zero padding in every
// column (obn_dataset_upload), i.e. they land in cell 0; the CTA that owns the
// last vector subtracts them again.
#pragma once
#include "obn_kernels.cuh"

namespace obn {

constexpr int kGroupMembers = 32;   // CI tests that may share one table (plan_ci_groups splits larger groups)

struct CtaTuple {            // per-CTA tuple description in shared memory
  const uint8_t* col[kMaxTupleVars];   // column pointers, ascending var ids (reference key order)
  uint32_t card[kMaxTupleVars];        // their cardinalities
  // column groups (kernel order): consecutive columns whose cards multiply to <= 256 are folded
  // with packed 8-bit arithmetic (4 rows per IMAD) and only the group keys are widened
  uint32_t gcells[kMaxTupleVars];      // cells of the group that ends at column i (0: not a group end)
  uint32_t weight[kMaxTupleVars];      // radix weight of digit i (contingency_table.h:43-49: the last digit has weight 1)
  int packed;                // every column of the tuple has <= 16 states and the data set carries nibble-packed columns
  int child_pos;             // BDEU_ONE: position of the child in reference order
  unsigned int ticket;
  uint32_t live;                       // CI: bit m = member m of the group is still asked (all ones unless a chained PC round)
};

template <int KW>
struct KeyVec;               // keys of 16 consecutive rows
template <>
struct KeyVec<8> {
  uint32_t w[4];
  __device__ __forceinline__ void from8(const KeyVec<8>& p) { *this = p; }
  __device__ __forceinline__ void push8(const KeyVec<8>& p, uint32_t cells) {      // only one group when KW == 8
#pragma unroll
    for (int i = 0; i < 4; ++i) w[i] = w[i] * cells + p.w[i];
  }
  __device__ __forceinline__ void init(const uint4& c) { w[0] = c.x; w[1] = c.y; w[2] = c.z; w[3] = c.w; }
  __device__ __forceinline__ void push(const uint4& c, uint32_t card) {
    w[0] = w[0] * card + c.x; w[1] = w[1] * card + c.y;
    w[2] = w[2] * card + c.z; w[3] = w[3] * card + c.w;
  }
  __device__ __forceinline__ void scale4() {
#pragma unroll
    for (int i = 0; i < 4; ++i) w[i] <<= 2;
  }
  __device__ __forceinline__ uint32_t get(int r) const { return __byte_perm(w[r >> 2], 0, 0x4440 + (r & 3)); }
};
template <>
struct KeyVec<16> {
  uint32_t w[8];
  __device__ __forceinline__ void init(const uint4& c) {
    const uint32_t s[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      w[2 * i] = __byte_perm(s[i], 0, 0x4140);        // bytes 0,1 -> halves
      w[2 * i + 1] = __byte_perm(s[i], 0, 0x4342);    // bytes 2,3 -> halves
    }
  }
  __device__ __forceinline__ void push(const uint4& c, uint32_t card) {
    const uint32_t s[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      w[2 * i] = w[2 * i] * card + __byte_perm(s[i], 0, 0x4140);
      w[2 * i + 1] = w[2 * i + 1] * card + __byte_perm(s[i], 0, 0x4342);
    }
  }
  __device__ __forceinline__ void scale4() {
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] <<= 2;
  }
  __device__ __forceinline__ void from8(const KeyVec<8>& p) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      w[2 * i] = __byte_perm(p.w[i], 0, 0x4140);
      w[2 * i + 1] = __byte_perm(p.w[i], 0, 0x4342);
    }
  }
  __device__ __forceinline__ void push8(const KeyVec<8>& p, uint32_t cells) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      w[2 * i] = w[2 * i] * cells + __byte_perm(p.w[i], 0, 0x4140);
      w[2 * i + 1] = w[2 * i + 1] * cells + __byte_perm(p.w[i], 0, 0x4342);
    }
  }
  __device__ __forceinline__ uint32_t get(int r) const { return (r & 1) ? (w[r >> 1] >> 16) : (w[r >> 1] & 0xffffu); }
};
template <>
struct KeyVec<32> {
  uint32_t w[16];
  __device__ __forceinline__ void init(const uint4& c) {
    const uint32_t s[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = __byte_perm(s[i >> 2], 0, 0x4440 + (i & 3));
  }
  __device__ __forceinline__ void push(const uint4& c, uint32_t card) {
    const uint32_t s[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = w[i] * card + __byte_perm(s[i >> 2], 0, 0x4440 + (i & 3));
  }
  __device__ __forceinline__ void scale4() {
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] <<= 2;
  }
  __device__ __forceinline__ void from8(const KeyVec<8>& p) {
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = __byte_perm(p.w[i >> 2], 0, 0x4440 + (i & 3));
  }
  __device__ __forceinline__ void push8(const KeyVec<8>& p, uint32_t cells) {
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = w[i] * cells + __byte_perm(p.w[i >> 2], 0, 0x4440 + (i & 3));
  }
  __device__ __forceinline__ uint32_t get(int r) const { return w[r]; }
};

// Replicated tables (REP kernels).  ATOMS applies one address per bank per pass, so the lanes of one
// instruction that hit DIFFERENT cells of the same bank serialise: 1.5 passes per instruction at 64
// cells, about 2.8 at 256 and 3.4 at 1 Ki on the PC frontiers (ncu: atom wavefronts per instruction),
// which is what held the mid-size classes at 40 % of the increment ceiling.  With R copies of the
// table interleaved cell-major -- word (cell * R + (lane & (R-1))) -- lane l can only meet lanes
// l + R, l + 2R, ... in its bank: at R = 16 at most two addresses per bank, which fits the two
// clocks an ATOMS instruction occupies anyway (benchmarks/micro/smem_incr.cu variant 9).  The copies
// are summed into one compact table before the epilogue.
//
// RSH = log2(R) is a compile-time constant of the kernel (-1: a single table): the address of an
// increment is then `lane_base + (field << shift)`, one LEA, where lane_base = table + 4 * (lane & (R-1))
// is computed once per thread.  (With R as a run-time value ptxas spent four ALU instructions per
// increment on it and the <= 64-cell class became issue-bound at 82 % issue-active.)

// 16 increments for one key vector.  `lane_base` is this lane's copy of the table as a byte pointer.
template <int KW, int TIER, bool SCALE4, int RSH>
__device__ __forceinline__ void count16(KeyVec<KW>& key, unsigned char* lane_base, uint32_t lo, uint32_t span) {
  if (SCALE4) key.scale4();
  constexpr int kShift = (SCALE4 ? 0 : 2) + (RSH > 0 ? RSH : 0);
  if (TIER == TIER_CHUNK) {
    const uint32_t a0 = (uint32_t)__cvta_generic_to_shared(lane_base);
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const uint32_t f = key.get(r) - lo;             // cells outside [lo, lo + span) belong to another CTA
      if (f < span) asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(a0 + (f << 2)) : "memory");
    }
  } else if (TIER == TIER_GLOBAL) {
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      // non-returning reduction (REDG.E.ADD): a plain atomicAdd compiles to ATOMG here, which makes
      // every row wait for an L2 round trip it has no use for
      asm volatile("red.global.add.u32 [%0], 1;" ::"l"(lane_base + ((size_t)key.get(r) << kShift)) : "memory");
    }
  } else {
    // one 32-bit shared-window address per lane, so that an increment is extract + LEA + ATOMS.POPC.INC
    const uint32_t a0 = (uint32_t)__cvta_generic_to_shared(lane_base);
#pragma unroll
    for (int r = 0; r < 16; ++r)
      asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(a0 + (key.get(r) << kShift)) : "memory");
  }
}

// Fold one column into the running key: inside a group with packed 8-bit arithmetic, across
// groups in the key's own width.
template <int KW>
__device__ __forceinline__ void fold_column(KeyVec<8>& part, KeyVec<KW>& key, bool& have_part, bool& have_key,
                                            const uint4& c, uint32_t card, uint32_t gcells) {
  if (!have_part) {
    part.init(c);
    have_part = true;
  } else {
    part.push(c, card);
  }
  if (gcells) {                       // this column closes its group
    if (!have_key) {
      key.from8(part);
      have_key = true;
    } else {
      key.push8(part, gcells);
    }
    have_part = false;
  }
}

// Row scan with the column count known at compile time: column pointers and
// cards live in registers, and the loads of the NEXT 16-row vector are issued
// before the increments of the current one (register double buffering).
// Nibble-packed columns (BatchArgs::data4): a 16-byte load holds 32 rows of a column whose cardinality is <= 16
// -- byte j of chunk c carries row 32c + j in its low and row 32c + 16 + j in its high nibble -- so the two
// halves unpack into the same byte vectors the unpacked loop builds its keys from.
__device__ __forceinline__ uint4 nibbles(const uint4& p, int half) {
  const uint32_t m = 0x0f0f0f0fu;
  return half ? make_uint4((p.x >> 4) & m, (p.y >> 4) & m, (p.z >> 4) & m, (p.w >> 4) & m)
              : make_uint4(p.x & m, p.y & m, p.z & m, p.w & m);
}

template <int KW, int TIER, bool SCALE4, int RSH, int K, bool PACK>
__device__ __forceinline__ void scan_rows_k(const CtaTuple& T, unsigned char* base, uint32_t v_begin,
                                            uint32_t v_end, uint32_t tid, uint32_t nthr, uint32_t lo, uint32_t span) {
  const uint8_t* col[K];
  uint32_t card[K], gcells[K];
#pragma unroll
  for (int i = 0; i < K; ++i) { col[i] = T.col[i]; card[i] = T.card[i]; gcells[i] = T.gcells[i]; }
  uint4 cur[K], nxt[K];
  uint32_t v = v_begin + tid;
  if (v < v_end) {
#pragma unroll
    for (int i = 0; i < K; ++i) cur[i] = ld_codes16(col[i] + ((size_t)v << 4));
  }
  for (; v < v_end; v += nthr) {
    const uint32_t vn = v + nthr;
    if (vn < v_end) {
#pragma unroll
      for (int i = 0; i < K; ++i) nxt[i] = ld_codes16(col[i] + ((size_t)vn << 4));
    }
#pragma unroll
    for (int half = 0; half < (PACK ? 2 : 1); ++half) {
      uint4 c[K];
#pragma unroll
      for (int i = 0; i < K; ++i) c[i] = PACK ? nibbles(cur[i], half) : cur[i];
      KeyVec<KW> key;
      if (KW == 8) {                    // tables of <= 256 cells: one group, all packed
        key.init(c[0]);
#pragma unroll
        for (int i = 1; i < K; ++i) key.push(c[i], card[i]);
      } else {
        KeyVec<8> part;
        bool have_part = false, have_key = false;
#pragma unroll
        for (int i = 0; i < K; ++i) fold_column<KW>(part, key, have_part, have_key, c[i], card[i], gcells[i]);
      }
      count16<KW, TIER, SCALE4, RSH>(key, base, lo, span);
    }
#pragma unroll
    for (int i = 0; i < K; ++i) cur[i] = nxt[i];
  }
}

// Any column count (descriptors stay in shared memory): columns are loaded four at a time so
// that four 16-byte loads are in flight per thread.
template <int KW, int TIER, bool SCALE4, int RSH, bool PACK>
__device__ __forceinline__ void scan_rows_any(const CtaTuple& T, int k, unsigned char* base, uint32_t v_begin,
                                              uint32_t v_end, uint32_t tid, uint32_t nthr, uint32_t lo, uint32_t span) {
  for (uint32_t v = v_begin + tid; v < v_end; v += nthr) {
    const size_t off = (size_t)v << 4;
#pragma unroll
    for (int half = 0; half < (PACK ? 2 : 1); ++half) {     // (the packed columns are read once per half: L1 hits)
      KeyVec<KW> key;
      KeyVec<8> part;
      bool have_part = false, have_key = false;
      for (int i = 0; i < k; i += 4) {
        uint4 c[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (i + j < k) {
            c[j] = ld_codes16(T.col[i + j] + off);
            if (PACK) c[j] = nibbles(c[j], half);
          }
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (i + j < k) fold_column<KW>(part, key, have_part, have_key, c[j], T.card[i + j], T.gcells[i + j]);
      }
      count16<KW, TIER, SCALE4, RSH>(key, base, lo, span);
    }
  }
}

// Chained PC rounds: which members of group t are still asked.  Called by all 32 lanes of warp 0; lane m looks at
// member m (a group has at most kGroupMembers = 32 tests).  The previous round's first[] is final before this kernel
// starts (stream order), so every CTA of the tuple, and the epilogue kernel, take the same decision.
__device__ __forceinline__ uint32_t live_members(const BatchArgs& A, uint32_t t, uint32_t lane, uint32_t* n_mem_out) {
  const uint32_t m_begin = __ldg(&A.mem_off[t]), n_mem = __ldg(&A.mem_off[t + 1]) - m_begin;
  bool live = false;
  if (lane < n_mem) {
    const uint32_t id = __ldg(&A.members[m_begin + lane]).x;
    live = __ldcg(&A.prev_first[__ldg(&A.slot_of_test[id])]) >= kFirstNone;
  }
  *n_mem_out = n_mem;
  return __ballot_sync(0xffffffffu, live);
}

template <int KW, int TIER, bool SCALE4, int RSH = -1>
__global__ void __launch_bounds__((KW == 32 ? 512 : 256), (KW == 8 ? 4 : KW == 32 ? 1 : 3)) count_kernel(const BatchArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ CtaTuple T;
  __shared__ double red_d[32];
  __shared__ unsigned long long red_u[32];

  const uint32_t tid = threadIdx.x, nthr = blockDim.x;
  const uint32_t lt = blockIdx.x / A.splits;       // launch-local tuple
  const uint32_t split = blockIdx.x - lt * A.splits;
  // per-launch-tuple header, written by the host in launch order:
  //   x = tuple id in the batch, y = offset of its descriptors, z = #variables, w = #cells
  const uint4 hdr = __ldg(&A.hdr[lt]);
  const uint32_t t = hdr.x;
  const int k = (int)hdr.z;
  const uint32_t cells = hdr.w;

  // ---- tuple prologue (one warp, one lane per variable) ---------------------
  // desc[i] = var_id | (card-1) << 24, ascending var ids = the digits of the reference key, most
  // significant first (contingency_table.h:43-49), which is also the order the table is built in.
  if (tid < 32) {
    const uint32_t child = (A.mode == MODE_BDEU_ONE) ? __ldg(&A.roles[lt]).x : 0xffffffffu;
    const bool live = (int)tid < k;
    const uint32_t d = live ? __ldg(&A.desc[hdr.y + tid]) : 0u;
    const uint32_t v = d & 0xffffffu, c = (d >> 24) + 1u;
    const bool packed = A.data4 != nullptr && k > 0 && !__any_sync(0xffffffffu, live && c > 16u);
    if (live) {
      T.col[tid] = packed ? A.data4 + (size_t)v * A.stride4 : A.data + (size_t)v * A.stride;
      T.card[tid] = c;
    }
    if (tid == 0) T.packed = packed;
    if (A.prev_first) {                            // chained PC round: drop the tests of edges the previous round decided
      uint32_t n_mem;
      const uint32_t mask = live_members(A, t, tid, &n_mem);
      if (tid == 0) {
        T.live = mask;
        const uint32_t n_dead = n_mem - (uint32_t)__popc(mask);
        if (split == 0 && n_dead) {                // accounting: one CTA per tuple
          atomicAdd(&A.skipped[0], (unsigned long long)n_dead);
          atomicAdd(&A.skipped[1], (unsigned long long)n_dead * (unsigned)k);
          if (!mask) {
            atomicAdd(&A.skipped[2], 1ull);
            atomicAdd(&A.skipped[3], (unsigned long long)k);
          }
        }
      }
    } else if (tid == 0) {
      T.live = 0xffffffffu;
    }
    const unsigned m_child = __ballot_sync(0xffffffffu, live && v == child);
    __syncwarp();
    if (tid == 0) {
      T.child_pos = m_child ? (__ffs(m_child) - 1) : -1;
      uint32_t g = 1;                              // greedy groups of <= 256 cells
      for (int i = 0; i < k; ++i) {
        g *= T.card[i];
        const bool close = (i + 1 == k) || g * T.card[i + 1] > 256u;
        T.gcells[i] = close ? g : 0u;
        if (close) g = 1;
      }
      uint32_t w = 1;
      for (int i = k - 1; i >= 0; --i) {
        T.weight[i] = w;
        w *= T.card[i];
      }
    }
  }

  // ---- table + epilogue scratch (zeroed while the descriptor loads are in flight)
  int32_t* tab;                     // the table the epilogue reads (one copy)
  int32_t* tab_rep = nullptr;       // RSH > 0: the R interleaved copies the row loop counts into
  int32_t* tab_chunk = nullptr;     // TIER_CHUNK: this CTA's cell range [chunk_lo, chunk_lo + chunk_span) in shared memory
  uint32_t chunk_lo = 0, chunk_span = 0;
  int32_t* wscratch;                // per-warp epilogue scratch
  constexpr uint32_t R = RSH > 0 ? (1u << RSH) : 1u;
  if (TIER == TIER_SMEM) {
    size_t o = 0;
    if (R > 1u) {
      tab_rep = reinterpret_cast<int32_t*>(smem_raw);
      o = (size_t)cells * R * 4u;
      for (uint32_t i = tid; i < cells * R; i += nthr) tab_rep[i] = 0;
    }
    tab = reinterpret_cast<int32_t*>(smem_raw + o);
    wscratch = reinterpret_cast<int32_t*>(smem_raw + o + (((size_t)cells * 4u + 15u) & ~(size_t)15u));
    if (!tab_rep)
      for (uint32_t i = tid; i < cells; i += nthr) tab[i] = 0;
  } else if (TIER == TIER_CHUNK) {
    tab = A.scratch + A.scratch_off[lt];
    chunk_lo = (split / A.row_splits) * A.chunk_cells;
    chunk_span = chunk_lo < cells ? min(A.chunk_cells, cells - chunk_lo) : 0u;
    tab_chunk = reinterpret_cast<int32_t*>(smem_raw);
    wscratch = reinterpret_cast<int32_t*>(smem_raw + (size_t)A.chunk_cells * 4u);
    for (uint32_t i = tid; i < chunk_span; i += nthr) tab_chunk[i] = 0;
  } else {
    tab = A.scratch + A.scratch_off[lt];
    wscratch = reinterpret_cast<int32_t*>(smem_raw);
  }
  __syncthreads();
  if (A.mode == MODE_CI && T.live == 0u) return;     // nothing of this table is asked any more (nothing was touched yet)

  // ---- main loop over this split's 16-row vectors -----------------------------
  // a "vector" is what one 16-byte load of a column holds: 16 rows, or 32 of a nibble-packed column
  const bool packed = T.packed != 0;
  const int vshift = packed ? 5 : 4;
  const uint32_t n_vec = (A.n_rows + (packed ? 31u : 15u)) >> vshift;
  const uint32_t vps = packed ? (A.vec_per_split + 1u) >> 1 : A.vec_per_split;
  const uint32_t v_begin = (TIER == TIER_CHUNK ? split % A.row_splits : split) * vps;
  uint32_t v_end = v_begin + vps;
  if (v_end > n_vec) v_end = n_vec;
  unsigned char* base = reinterpret_cast<unsigned char*>(tab_chunk ? tab_chunk : tab_rep ? tab_rep : tab) + 4u * (tid & (R - 1u));   // this lane's copy
#define OBN_SCAN(PACK)                                                                                              \
  switch (k) {                                                                                                     \
    case 0: break;                                                                                                 \
    case 1: scan_rows_k<KW, TIER, SCALE4, RSH, 1, PACK>(T, base, v_begin, v_end, tid, nthr, chunk_lo, chunk_span); break; \
    case 2: scan_rows_k<KW, TIER, SCALE4, RSH, 2, PACK>(T, base, v_begin, v_end, tid, nthr, chunk_lo, chunk_span); break; \
    case 3: scan_rows_k<KW, TIER, SCALE4, RSH, 3, PACK>(T, base, v_begin, v_end, tid, nthr, chunk_lo, chunk_span); break; \
    case 4: scan_rows_k<KW, TIER, SCALE4, RSH, 4, PACK>(T, base, v_begin, v_end, tid, nthr, chunk_lo, chunk_span); break; \
    case 5: scan_rows_k<KW, TIER, SCALE4, RSH, 5, PACK>(T, base, v_begin, v_end, tid, nthr, chunk_lo, chunk_span); break; \
    case 6: scan_rows_k<KW, TIER, SCALE4, RSH, 6, PACK>(T, base, v_begin, v_end, tid, nthr, chunk_lo, chunk_span); break; \
    default: scan_rows_any<KW, TIER, SCALE4, RSH, PACK>(T, k, base, v_begin, v_end, tid, nthr, chunk_lo, chunk_span); break; \
  }
  if (packed) { OBN_SCAN(true) } else { OBN_SCAN(false) }
#undef OBN_SCAN
  if (tab_rep) {                     // sum the copies into the one table everything below reads
    __syncthreads();
    for (uint32_t c = tid; c < cells; c += nthr) {
      const int32_t* p = tab_rep + (size_t)c * R;
      int a = 0;
      for (uint32_t r = 0; r < R; ++r) a += p[(r + tid) & (R - 1u)];     // staggered: lanes start in different banks
      tab[c] = a;
    }
    __syncthreads();
  }
  if (tid == 0 && (TIER != TIER_CHUNK || (chunk_lo == 0 && chunk_span))) {     // (cell 0 lives in chunk 0)
    int32_t* cell0 = TIER == TIER_CHUNK ? tab_chunk : tab;
    if (k == 0) {
      // the empty tuple: one cell holding every row of the split (contingency_table.h:52-55)
      const uint32_t r0 = v_begin << vshift;
      uint32_t r1 = v_end << vshift;
      if (r1 > A.n_rows) r1 = A.n_rows;
      if (r1 > r0) atomicAdd(cell0, (int)(r1 - r0));
    } else if (v_end == n_vec && v_end > v_begin) {
      const int pad = (int)((n_vec << vshift) - A.n_rows);     // zero-padding rows counted into cell 0
      if (pad) atomicAdd(cell0, -pad);
    }
  }
  __syncthreads();

  // ---- row splits: merge through global scratch, last CTA continues ---------
  if (A.splits > 1 || TIER != TIER_SMEM) {
    int32_t* g = A.scratch + A.scratch_off[lt];
    if (TIER == TIER_CHUNK) {
      for (uint32_t c = tid; c < chunk_span; c += nthr) {
        const int v = tab_chunk[c];
        if (v) atomicAdd(&g[chunk_lo + c], v);
      }
    }
    if (TIER == TIER_SMEM) {
      for (uint32_t c = tid; c < cells; c += nthr) {
        const int v = tab[c];
        if (v) atomicAdd(&g[c], v);
      }
    }
    if (TIER != TIER_SMEM && A.epi_ctas) return;          // ci_epilogue_kernel reads the table after this launch
    __threadfence();
    __syncthreads();
    if (tid == 0) T.ticket = atomicAdd(&A.tickets[lt], 1u);
    __syncthreads();
    if (T.ticket != (unsigned)A.splits - 1) return;
    __threadfence();
    if (TIER == TIER_SMEM) {
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
    // one table, every test asked over its variable set: member (px, py) reads the table with x at
    // digit px and y at digit py (CiGeom).  The p-values wait until every member has its statistic
    // and are then evaluated by one thread per member instead of serially by thread 0.
    __shared__ double m_stat[kGroupMembers];
    __shared__ unsigned long long m_dof[kGroupMembers];
    __shared__ uint32_t m_id[kGroupMembers];
    const uint32_t m_begin = __ldg(&A.mem_off[t]), m_end = __ldg(&A.mem_off[t + 1]);
    const uint32_t live = T.live;
    for (uint32_t m = m_begin; m < m_end; ++m) {
      if (!((live >> (m - m_begin)) & 1u)) continue;
      const uint2 mem = __ldg(&A.members[m]);
      const uint32_t px = mem.y & 0xffu, py = (mem.y >> 8) & 0xffu;
      const CiGeom g(cells, T.card[px], T.card[py], T.weight[px], T.weight[py]);
      double stat;
      unsigned long long dof;
      epilogue_ci<TIER != TIER_SMEM>(tab, g, A.kind, wscratch + (tid >> 5) * A.wstride, red_d, red_u, stat, dof);
      if (tid == 0) {
        m_stat[m - m_begin] = stat;
        m_dof[m - m_begin] = dof;
        m_id[m - m_begin] = mem.x;
      }
    }
    __syncthreads();
    if (tid < m_end - m_begin && ((live >> tid) & 1u)) {
      const uint32_t id = m_id[tid];
      if (A.stat) A.stat[id] = m_stat[tid];
      if (A.dof) A.dof[id] = m_dof[tid];
      if (A.pval || A.first) {
        const double p = pchisq(m_stat[tid], m_dof[tid]);                  // citest.h:139
        if (A.pval) A.pval[id] = p;
        if (A.first && p >= A.level) atomicMin(&A.first[__ldg(&A.slot_of_test[id])], (int32_t)id);   // :147
      }
    }
  } else if (A.mode == MODE_BDEU_ONE || A.mode == MODE_BDEU_ALL) {
    const int c0 = (A.mode == MODE_BDEU_ONE) ? T.child_pos : 0;
    const int c1 = (A.mode == MODE_BDEU_ONE) ? T.child_pos + 1 : k;
    double* out = (A.mode == MODE_BDEU_ONE) ? &A.scores[t] : &A.scores[hdr.y];
    if (A.lut) {
      epilogue_bdeu_exact(tab, cells, k, T.card, A.lut + (size_t)A.lut_ijk[t] * A.lut_stride,
                          A.lut, A.lut_stride, A.lut_ij + hdr.y, c0, c1, out);
    } else {
      epilogue_bdeu_fast(tab, cells, k, T.card, A.ess, c0, c1, red_d, out);
    }
  } else {
    int32_t* out = A.counts_out + A.counts_off[t];
    for (uint32_t c = tid; c < cells; c += nthr) out[c] = tab[c];
  }

  if (TIER != TIER_SMEM) {        // leave the scratch table zeroed for the next launch
    __syncthreads();
    for (uint32_t c = tid; c < cells; c += nthr) tab[c] = 0;
    if (tid == 0) A.tickets[lt] = 0;
  }
}

// The CI epilogue of tables that live in global memory, as its own launch: A.epi_ctas CTAs per tuple, CTA e
// takes the slices [e, e+1) * ceil(n_slices / epi_ctas) of every member, so a 400 k-cell table with ten members
// is read by the whole GPU instead of by the one CTA that happened to finish counting last (6 ms of a 21 ms C5
// run before).  Partial sums are combined by the last CTA to arrive, in CTA order: deterministic.
__global__ void __launch_bounds__(256) ci_epilogue_kernel(const BatchArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ CtaTuple T;
  __shared__ double red_d[32];
  __shared__ unsigned long long red_u[32];
  const uint32_t tid = threadIdx.x, nthr = blockDim.x;
  const uint32_t lt = blockIdx.x / A.epi_ctas, e = blockIdx.x - lt * A.epi_ctas;
  const uint4 hdr = __ldg(&A.hdr[lt]);
  const uint32_t t = hdr.x, cells = hdr.w;
  const int k = (int)hdr.z;
  if (tid < 32) {
    if ((int)tid < k) T.card[tid] = (__ldg(&A.desc[hdr.y + tid]) >> 24) + 1u;
    __syncwarp();
    if (tid == 0) {
      uint32_t w = 1;
      for (int i = k - 1; i >= 0; --i) {
        T.weight[i] = w;
        w *= T.card[i];
      }
    }
    uint32_t mask = 0xffffffffu, n_mem_unused;
    if (A.prev_first) mask = live_members(A, t, tid, &n_mem_unused);
    if (tid == 0) T.live = mask;
  }
  __syncthreads();
  const uint32_t live = T.live;
  if (live == 0u) return;                                  // the counting kernel skipped this table: it is still all zero
  int32_t* tab = A.scratch + A.scratch_off[lt];
  int32_t* wscratch = reinterpret_cast<int32_t*>(smem_raw);
  const uint32_t m_begin = __ldg(&A.mem_off[t]), m_end = __ldg(&A.mem_off[t + 1]), n_mem = m_end - m_begin;
  double* part = A.partials + (size_t)lt * kGroupMembers * A.epi_ctas * 2;
  for (uint32_t m = 0; m < n_mem; ++m) {
    if (!((live >> m) & 1u)) continue;
    const uint2 mem = __ldg(&A.members[m_begin + m]);
    const uint32_t px = mem.y & 0xffu, py = (mem.y >> 8) & 0xffu;
    const CiGeom g(cells, T.card[px], T.card[py], T.weight[px], T.weight[py]);
    const uint32_t per = (g.n_slices + A.epi_ctas - 1) / A.epi_ctas;
    const uint32_t s0 = min(g.n_slices, e * per), s1 = min(g.n_slices, s0 + per);
    double stat;
    unsigned long long dof;
    epilogue_ci<true>(tab, g, A.kind, wscratch + (tid >> 5) * A.wstride, red_d, red_u, stat, dof, s0, s1);
    if (tid == 0) {
      part[((size_t)m * A.epi_ctas + e) * 2] = stat;
      part[((size_t)m * A.epi_ctas + e) * 2 + 1] = __longlong_as_double((long long)dof);
    }
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) T.ticket = atomicAdd(&A.tickets[lt], 1u);
  __syncthreads();
  if (T.ticket != A.epi_ctas - 1) return;
  __threadfence();
  if (tid < n_mem && ((live >> tid) & 1u)) {               // one thread per member: CTA order, then p and the decision
    double stat = 0.0;
    unsigned long long dof = 0;
    for (uint32_t q = 0; q < A.epi_ctas; ++q) {
      stat += __ldcg(&part[((size_t)tid * A.epi_ctas + q) * 2]);
      dof += (unsigned long long)__double_as_longlong(__ldcg(&part[((size_t)tid * A.epi_ctas + q) * 2 + 1]));
    }
    const uint32_t id = __ldg(&A.members[m_begin + tid]).x;
    if (A.stat) A.stat[id] = stat;
    if (A.dof) A.dof[id] = dof;
    if (A.pval || A.first) {
      const double p = pchisq(stat, dof);                                  // citest.h:139
      if (A.pval) A.pval[id] = p;
      if (A.first && p >= A.level) atomicMin(&A.first[__ldg(&A.slot_of_test[id])], (int32_t)id);   // :147
    }
  }
  for (uint32_t c = tid; c < cells; c += nthr) tab[c] = 0;         // the scratch table goes back zeroed
  if (tid == 0) A.tickets[lt] = 0;
}

}  // namespace obn
