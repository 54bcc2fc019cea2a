// obn_static.inl -- obn_count_table_static: ONE contingency table counted without any allocation.
//
// No counterpart in the reference; added for callers that
// cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid
// it).  It is the global-table variant of the counting kernel (obn_count_kernel.cuh, TIER_GLOBAL:
// CTAs merge their partial counts into a table in global memory, the last CTA hands it on) with
// that global table living in a file-scope __device__ array instead of a grow-only cudaMalloc'ed
// buffer of the ctx.  Inputs and outputs are the caller's DEVICE buffers; the call needs no ctx.
//
//   limit     OBN_STATIC_MAX_CELLS (65536) cells and kStaticMaxVars (16) variables per table; the
//             array (256 KiB + one 256-byte control block) is held by every GPU that loads the
//             library, until the process exits
//   sharing   every call on a device uses the SAME array.  It is cleared at the start of every call
//             (cudaMemsetAsync on the caller's stream), so a call never sees the previous call's
//             counts; calls that overlap on the GPU (different streams, no dependency) race.
//   no alloc  the call issues cudaSetDevice, cudaMemsetAsync and one kernel launch.  The first call
//             per device also resolves the array's device address (cudaGetSymbolAddress -- the
//             array's NAME in host code is a host-side stand-in, not a device pointer) and reads
//             the SM count; make that first call outside stream capture.
// Included by obn_capi.cu (same translation unit as the array and the kernel).
#include <mutex>

namespace obn {

constexpr int kStaticMaxVars = 16;
constexpr uint32_t kStaticMaxCells = OBN_STATIC_MAX_CELLS;
constexpr uint32_t kStaticCtrlWords = 64;                 // 256-byte control block behind the table
constexpr uint32_t kStaticSmemCells = 8192;               // tables up to here are privatised per CTA (32 KiB)
constexpr int kStaticMaxDevices = 64;

static __device__ __align__(256) uint32_t g_static_scratch[kStaticMaxCells + kStaticCtrlWords];

struct StaticArgs {
  const uint8_t* col[kStaticMaxVars];   // column pointers, ascending var ids (reference key order)
  uint32_t weight[kStaticMaxVars];      // radix weight of digit i (contingency_table.h:43-49: the last digit has weight 1)
  uint32_t k, cells;
  uint64_t n_rows;
  uint32_t* scratch;                    // device address of g_static_scratch (cudaGetSymbolAddress)
  int32_t* out;
};

// VEC: every column pointer is 4-byte aligned -> four rows per 32-bit load.
template <bool SMEM, bool VEC>
__global__ void __launch_bounds__(256) count_table_static_kernel(const StaticArgs A) {
  __shared__ uint32_t tab_s[SMEM ? kStaticSmemCells : 1];
  __shared__ bool s_last;
  uint32_t* const gtab = A.scratch;
  if (SMEM) {
    for (uint32_t i = threadIdx.x; i < A.cells; i += blockDim.x) tab_s[i] = 0u;
    __syncthreads();
  }
  uint32_t* const tab = SMEM ? tab_s : gtab;
  const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t nthr = (uint64_t)gridDim.x * blockDim.x;
  if (VEC) {
    const uint64_t n4 = A.n_rows >> 2;
    for (uint64_t q = tid; q < n4; q += nthr) {
      uint32_t key0 = 0, key1 = 0, key2 = 0, key3 = 0;
      for (uint32_t j = 0; j < A.k; ++j) {
        const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(A.col[j]) + q);
        const uint32_t wt = A.weight[j];
        key0 += (w & 0xffu) * wt;
        key1 += ((w >> 8) & 0xffu) * wt;
        key2 += ((w >> 16) & 0xffu) * wt;
        key3 += (w >> 24) * wt;
      }
      atomicAdd(&tab[key0], 1u);
      atomicAdd(&tab[key1], 1u);
      atomicAdd(&tab[key2], 1u);
      atomicAdd(&tab[key3], 1u);
    }
  }
  for (uint64_t r = (VEC ? (A.n_rows & ~(uint64_t)3) : 0) + tid; r < A.n_rows; r += nthr) {
    uint32_t key = 0;
    for (uint32_t j = 0; j < A.k; ++j) key += (uint32_t)__ldg(A.col[j] + r) * A.weight[j];
    atomicAdd(&tab[key], 1u);
  }
  if (SMEM) {
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < A.cells; i += blockDim.x)
      if (const uint32_t v = tab_s[i]) atomicAdd(&gtab[i], v);
  }
  // the last CTA to arrive hands the table to the caller's buffer
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(&gtab[kStaticMaxCells], 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (uint32_t i = threadIdx.x; i < A.cells; i += blockDim.x) A.out[i] = (int32_t)__ldcg(&gtab[i]);
}

struct StaticDevice {
  bool ready = false;
  uint32_t* scratch = nullptr;
  int sm_count = 0;
};
static StaticDevice g_static_dev[kStaticMaxDevices];
static std::mutex g_static_mutex;

// Once per device, with that device current.
static int static_setup(int device, StaticDevice** out) {
  std::lock_guard<std::mutex> lock(g_static_mutex);
  StaticDevice& D = g_static_dev[device];
  if (!D.ready) {
    void* p = nullptr;
    OBN_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    if (!p) return fail(OBN_ERR_CUDA, "cudaGetSymbolAddress returned a NULL address for the static scratch");
    OBN_CUDA(cudaDeviceGetAttribute(&D.sm_count, cudaDevAttrMultiProcessorCount, device));
    D.scratch = static_cast<uint32_t*>(p);
    D.ready = true;
  }
  *out = &D;
  return OBN_OK;
}

struct DeviceGuard {
  int prev = -1;
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

}  // namespace obn

extern "C" {

int obn_count_table_static(int device, void* stream, const uint8_t* data_dev, size_t n_rows, size_t row_stride,
                           const uint32_t* cols, const uint32_t* cards, size_t n_cols, int32_t* counts_out_dev,
                           size_t n_cells) {
  using namespace obn;
  if (device < 0 || device >= kStaticMaxDevices) return fail(OBN_ERR_INVALID, "device %d out of range", device);
  if (!counts_out_dev || (n_cols && (!cols || !cards)) || (n_cols && n_rows && !data_dev))
    return fail(OBN_ERR_INVALID, "obn_count_table_static: NULL argument");
  if (n_cols > (size_t)kStaticMaxVars)
    return fail(OBN_ERR_UNSUPPORTED, "obn_count_table_static: %zu variables (the static variant takes at most %d)",
                n_cols, kStaticMaxVars);
  if (n_cols && n_rows > row_stride) return fail(OBN_ERR_INVALID, "row_stride %zu < n_rows %zu", row_stride, n_rows);
  if (n_rows >= (1ull << 31)) return fail(OBN_ERR_UNSUPPORTED, "more than 2^31 - 1 rows (int32 counts)");
  uint64_t cells = 1;
  for (size_t i = 0; i < n_cols; ++i) {
    if (i && cols[i - 1] > cols[i]) return fail(OBN_ERR_INVALID, "var_ids must be sorted");     // contingency_table.h:37-38
    if (cards[i] == 0 || cards[i] > 256) return fail(OBN_ERR_INVALID, "cards[%zu] = %u is not in 1..256", i, cards[i]);
    cells *= cards[i];
    if (cells > kStaticMaxCells)
      return fail(OBN_ERR_UNSUPPORTED,
                  "obn_count_table_static: the table has more than %u cells, the size of the library's static scratch "
                  "(OBN_STATIC_MAX_CELLS); use obn_count_tables",
                  kStaticMaxCells);
  }
  if (cells != n_cells)
    return fail(OBN_ERR_INVALID, "n_cells is %zu but the cardinalities give %llu cells", n_cells, (unsigned long long)cells);

  DeviceGuard guard;
  int cur = -1;
  OBN_CUDA(cudaGetDevice(&cur));
  if (cur != device) {
    OBN_CUDA(cudaSetDevice(device));
    guard.prev = cur;
  }
  StaticDevice* D = nullptr;
  if (int rc = static_setup(device, &D)) return rc;

  StaticArgs A{};
  bool vec = true;
  uint32_t w = 1;
  for (size_t i = n_cols; i-- > 0;) {
    A.col[i] = data_dev + (size_t)cols[i] * row_stride;
    A.weight[i] = w;
    w *= cards[i];
    if (reinterpret_cast<uintptr_t>(A.col[i]) & 3u) vec = false;
  }
  A.k = (uint32_t)n_cols;
  A.cells = (uint32_t)cells;
  A.n_rows = n_rows;
  A.scratch = D->scratch;
  A.out = counts_out_dev;

  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // the array keeps the last call's table and ticket: both start at zero
  OBN_CUDA(cudaMemsetAsync(D->scratch, 0, sizeof(uint32_t) * (kStaticMaxCells + kStaticCtrlWords), s));
  const bool smem = cells <= kStaticSmemCells;
  const uint64_t per_thread = vec ? 16 : 4;                       // rows per thread before the grid stops growing
  uint64_t grid = (n_rows + 256 * per_thread - 1) / (256 * per_thread);
  grid = std::max<uint64_t>(1, std::min<uint64_t>(grid, (uint64_t)D->sm_count * 8));
  if (smem && vec) count_table_static_kernel<true, true><<<(unsigned)grid, 256, 0, s>>>(A);
  else if (smem) count_table_static_kernel<true, false><<<(unsigned)grid, 256, 0, s>>>(A);
  else if (vec) count_table_static_kernel<false, true><<<(unsigned)grid, 256, 0, s>>>(A);
  else count_table_static_kernel<false, false><<<(unsigned)grid, 256, 0, s>>>(A);
  OBN_CUDA(cudaGetLastError());
  return OBN_OK;
}

}  // extern "C"
