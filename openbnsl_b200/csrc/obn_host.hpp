// obn_host.hpp -- host-side plumbing of the C-ABI library (ctx, buffers, errors).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <map>
#include <string>
#include <vector>

#include "../../include/openbnsl_b200.h"

namespace obn {

extern thread_local std::string g_last_error;

inline int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
  return code;
}

#define OBN_CUDA(call)                                                              \
  do {                                                                              \
    cudaError_t e_ = (call);                                                        \
    if (e_ != cudaSuccess)                                                          \
      return obn::fail(e_ == cudaErrorMemoryAllocation ? OBN_ERR_OOM : OBN_ERR_CUDA, \
                       "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),      \
                       __FILE__, __LINE__);                                         \
  } while (0)

// grow-only device buffer
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes, bool zero_new = false) {
    if (bytes <= cap) return OBN_OK;
    size_t want = bytes + bytes / 4 + 256;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    OBN_CUDA(cudaMalloc(&p, want));
    cap = want;
    if (zero_new) {      // rare (grow only); the ctx streams are non-blocking, so fence explicitly
      OBN_CUDA(cudaMemset(p, 0, want));
      OBN_CUDA(cudaDeviceSynchronize());
    }
    return OBN_OK;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <typename T>
  T* as() const { return reinterpret_cast<T*>(p); }
};

// grow-only pinned host buffer (staging for async H2D / D2H)
struct PinBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes) {
    if (bytes <= cap) return OBN_OK;
    size_t want = bytes + bytes / 4 + 256;
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
    OBN_CUDA(cudaMallocHost(&p, want));
    cap = want;
    return OBN_OK;
  }
  void release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
  }
  template <typename T>
  T* as() const { return reinterpret_cast<T*>(p); }
};

}  // namespace obn

namespace obn {
// Host-libm tables lut[id][m] = lgamma(m + a) - lgamma(a), m = 0..n_rows, for the exact-order
// BDeu epilogue (local_score.h:53,59); kept on the device across calls, keyed by the bits of a.
struct LutCache {
  std::map<uint64_t, uint32_t> ids;
  size_t stride = 0, rows = 0, cap_rows = 0, lut_stride_of_alloc = 0;
  double* dev = nullptr;
  void clear() {
    ids.clear();
    rows = 0;
  }
  int append(obn_ctx* c, const std::vector<double>& avals);   // obn_capi.cu
  void release() {
    if (dev) cudaFree(dev);
    dev = nullptr;
    cap_rows = rows = 0;
    ids.clear();
  }
};
}  // namespace obn

namespace obn {
// Device buffers of exhaustive-search sessions (obn_es.inl, obn_lattice.inl).  They belong to the
// ctx, grow only, and are reused by the next session: a search does not pay cudaMalloc / cudaFree
// of tens of GB (which also serialise between the per-GPU processes of one box).
// The instantiated lattice graph of the last search shape (cards, rows, ess, rank/world): a repeated
// search on the same ctx (bootstrap resampling, ess sweeps re-key it) skips plan, uploads and capture.
struct LatticeGraphCache {
  std::vector<uint64_t> key;
  cudaGraphExec_t exec = nullptr;
  int h = 0, L = 0;
  uint32_t per = 0;
  uint64_t joint_off = 0, joint_cells = 0;
  uint64_t launches = 0;
  bool permuted = false;           // digit order differs from the variable order (d_perm holds the mask translation)
  std::vector<uint32_t> order;     // order[i] = variable at digit i
  void clear() {
    if (exec) cudaGraphExecDestroy(exec);
    exec = nullptr;
    key.clear();
  }
};

struct EsPool {
  bool in_use = false;
  LatticeGraphCache lattice;
  int lists_n = 0;                 // d_lists holds the level lists (subsets grouped by size) of this many variables
  DevBuf d_bl_score, d_bl_ps, d_best_gs, d_best_ch, d_F;
  DevBuf d_root, d_pp[4], d_tasks, d_slot, d_partials, d_lut_a, d_lutF, d_masks, d_level, d_prog, d_lists, d_perm;
  void release() {
    lattice.clear();
    for (DevBuf* b : {&d_bl_score, &d_bl_ps, &d_best_gs, &d_best_ch, &d_F, &d_root, &d_pp[0], &d_pp[1], &d_pp[2], &d_pp[3], &d_tasks, &d_slot,
                      &d_partials, &d_lut_a, &d_lutF, &d_masks, &d_level, &d_prog, &d_lists, &d_perm})
      b->release();
  }
  size_t bytes() const {
    size_t s = 0;
    for (const DevBuf* b : {&d_bl_score, &d_bl_ps, &d_best_gs, &d_best_ch, &d_F, &d_root, &d_pp[0], &d_pp[1], &d_pp[2], &d_pp[3], &d_tasks,
                            &d_slot, &d_partials, &d_lut_a, &d_lutF, &d_masks, &d_level, &d_prog, &d_lists, &d_perm})
      s += b->cap;
    return s;
  }
};
}  // namespace obn

struct obn_dataset {
  obn_ctx* ctx = nullptr;
  uint8_t* data = nullptr;      // device, column-major, padded
  uint8_t* data4 = nullptr;     // device, behind `data`: two rows per byte for the columns of <= 16 states (null: not built)
  size_t stride4 = 0;
  size_t capacity = 0;          // bytes allocated behind `data`
  std::vector<uint32_t> cards;  // host copy
  size_t n_vars = 0, n_rows = 0, stride = 0;
};

struct obn_ctx {
  int device = 0;
  int sm_count = 148;
  size_t smem_optin = 0;
  cudaStream_t stream = nullptr;
  // the launches of one batch (one per table-size class) overlap on these
  static constexpr int kAuxStreams = 7;
  cudaStream_t aux[kAuxStreams] = {};
  cudaEvent_t ev_fork = nullptr, ev_join[kAuxStreams] = {};
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // uploads of a batch that is submitted while another one is running go here, so that the copy overlaps
  // the running batch's kernels instead of queueing behind them
  cudaStream_t up_stream = nullptr;
  obn::LutCache lut;
  obn::EsPool es_pool;
  obn::DevBuf d_scratch;   // int32, kept all-zero between launches
  obn::DevBuf d_tickets;   // uint32, kept all-zero between launches
  obn::DevBuf d_partials;  // per-CTA partial statistics of the split CI epilogue (ci_epilogue_kernel)
  obn::DevBuf d_scalar;    // staging of small host values that go through a collective
  uint8_t* ds_cache = nullptr;  // device buffer of the last destroyed dataset, reused by the next upload
  size_t ds_cache_bytes = 0;
  // What run_batch remembers of a batch between enqueueing it and handing out its results.
  struct Pending {
    size_t n_out = 0;             // units this rank evaluated (tests / tuples)
    size_t n_results = 0;         // CI: tests of the batch (= results handed out)
    size_t n_slots = 0;           // PC round: edges of the round (0: a plain CI batch)
    size_t out_bytes = 0, in_bytes = 0, tables = 0;
    double alg_bytes = 0, scanned_bytes = 0, increments = 0, n_rows = 0;
    bool want_p = false, timed = false, chained = false;
  };
  // Batch staging: one packed upload and one packed download per batch.  Two sets, so that a second
  // CI batch can be packed, uploaded and enqueued while the first is still running (obn_ci_tests_submit /
  // obn_pc_round_submit; the PC driver keeps the device one round ahead of the host).  The batches of a
  // ctx run in stream order, so the kernels' scratch (d_scratch, d_tickets, d_partials) stays single.
  struct Staging {
    obn::DevBuf d_in, d_out;
    obn::PinBuf h_in, h_out;
    cudaEvent_t e0 = nullptr, e1 = nullptr, done = nullptr;   // kernel phase [e0, e1]; done: results are in h_out
    cudaEvent_t uploaded = nullptr;                            // d_in holds the batch (copy on obn_ctx::up_stream)
    Pending pending;
  };
  static constexpr int kStagings = 2;
  Staging st[kStagings];
  int q_head = 0, q_count = 0;    // FIFO of submitted batches: st[q_head] is the oldest
  int last_round = -1;            // staging that holds first[] of the PC round submitted last (chained rounds read it)
  size_t last_round_slots = 0;
  // launch policy of the counting kernels (obn_ctx_set_option; defaults from the sweep in
  // profiles/r1_kernel_sweep.txt)
  struct Tuning {
    int threads_per_sm = 768;     // resident threads per SM a launch aims for
    int force_threads = 0;        // != 0: CTA width of every launch (sweeps)
    int force_splits = 0;         // != 0: row splits of every launch (sweeps)
    int streams = 4;              // launches of one batch are forked over this many streams (<= 1 + kAuxStreams)
    int largest_first = 0;        // launch the classes of a batch from the largest tables down (the long poles start first)
    int rep = 1;                  // replicated shared-memory tables (obn_count_kernel.cuh) on / off (A/B)
    int pack4 = 1;                // build and use nibble-packed columns (two rows per byte) for variables of <= 16 states
    int chunk = 1;                // tables beyond shared memory: chunked shared-memory counting (TIER_CHUNK) instead of L2 reductions
    int min_waves = 0;            // row-split the tuples of a launch until it has this many waves of CTAs (0: off)
    int min_class = 0;            // tables of the shared-memory classes below this one are counted by this one (fewer launches per batch; A/B)
    int lut_max_mb = 2048;        // cap of the lgamma tables kept for the exact-order BDeu epilogue (tests lower it)
    // marginalisation lattice (obn_lattice.inl)
    int lattice_budget_gb = 24, lattice_max_l = 20, lattice_bottom = 1, lattice_reorder = 1, lattice_graph = 1,
        lattice_lanes = 2, lattice_no_lut = 0, debug = 0;
  } tune;
  // communicator between the per-GPU processes of one job (obn_comm_init): NCCL, owned by the ctx
  void* comm = nullptr;           // ncclComm_t
  int comm_rank = 0, comm_world = 1;
  // profiling
  bool profile_on = false;
  obn_profile prof = {};
};
