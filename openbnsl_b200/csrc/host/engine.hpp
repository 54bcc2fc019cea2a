// engine.hpp -- C++ face of the C ABI for the host drivers: RAII handles,
// status -> exception mapping, the tuple-batch evaluator interface and the
// communicator interface used for multi-GPU sharding.
#pragma once
#include <cstdint>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "openbnsl_b200.h"

namespace obn {

// status -> the exception classes the reference throws for the same condition
inline void check(int rc) {
  if (rc == OBN_OK) return;
  const std::string msg = obn_last_error();
  switch (rc) {
    case OBN_ERR_INVALID: throw std::invalid_argument(msg);
    case OBN_ERR_RANGE: throw std::out_of_range(msg);
    case OBN_ERR_OOM: throw std::bad_alloc();
    default: throw std::runtime_error(msg);
  }
}

// One ctx per process (one process per GPU).  Device = $OBN_DEVICE, else
// $LOCAL_RANK, else 0.  Creation fails loudly when no GPU is usable.
class Engine {
 public:
  static Engine& instance();
  obn_ctx* ctx() const { return ctx_; }
  int device() const { return device_; }
  ~Engine();

 private:
  Engine();
  obn_ctx* ctx_ = nullptr;
  int device_ = 0;
};

class DeviceDataset {
 public:
  DeviceDataset(const uint8_t* col_major, size_t n_vars, size_t n_rows, size_t stride,
                const uint32_t* cards);
  ~DeviceDataset();
  DeviceDataset(const DeviceDataset&) = delete;
  DeviceDataset& operator=(const DeviceDataset&) = delete;
  obn_dataset* handle() const { return ds_; }

 private:
  obn_dataset* ds_ = nullptr;
};

// A CSR batch of tuples (ascending var ids inside each tuple).
struct TupleBatch {
  std::vector<uint32_t> var_ids;
  std::vector<uint64_t> offsets{0};
  std::vector<uint32_t> x, y;        // CI roles
  size_t size() const { return offsets.size() - 1; }
  void clear() {
    var_ids.clear();
    offsets.assign(1, 0);
    x.clear();
    y.clear();
  }
};

// What the drivers need from the hot path.  The product implementation is
// GpuEvaluator (the C ABI); tests inject a checker-backed one to exercise the
// driver/sharding logic on machines without a GPU.
struct Evaluator {
  virtual ~Evaluator() = default;
  // p-values of tuples [begin, end) of the batch -> p[0 .. end-begin)
  virtual void ci_pvalues(const TupleBatch& b, size_t begin, size_t end, int kind, double* p) = 0;
  // One PC round answered as a whole (obn_pc_round): first[s] = the first test of edge slot s whose
  // p >= level, INT32_MAX if none.  shard: split the round over the ranks of the library's own
  // communicator (collective).  false = not supported, the driver decides from ci_pvalues.
  virtual bool pc_round(const TupleBatch&, const std::vector<uint32_t>& /*slot_of_test*/, size_t /*n_slots*/,
                        int /*kind*/, double /*level*/, bool /*shard*/, std::vector<int32_t>& /*first*/) { return false; }
  // The same round in two halves (obn_pc_round_submit / _collect): submit returns while the answer is being
  // computed, collect waits for the OLDEST submitted round; two rounds may be in flight.  chain: the tests of
  // edges the round submitted before this one decided are skipped by the evaluator (same slot numbering).
  // false = not supported.
  virtual bool pc_round_submit(const TupleBatch&, const std::vector<uint32_t>& /*slot_of_test*/, size_t /*n_slots*/,
                               int /*kind*/, double /*level*/, bool /*shard*/, bool /*chain*/) { return false; }
  virtual void pc_round_collect(std::vector<int32_t>& /*first*/) {}
  // scores of every child of tuples [begin, end) -> out[offsets[t]-offsets[begin] + i]
  virtual void subset_scores(const TupleBatch& b, size_t begin, size_t end, double ess, int exact_order,
                             double* out) = 0;
  // out[t] = BDeu score of child[t] given the other variables of tuple t, for every tuple of the batch
  virtual void child_scores(const TupleBatch& b, const uint32_t* child, double ess, int exact_order, double* out) = 0;
};

class GpuEvaluator : public Evaluator {
 public:
  explicit GpuEvaluator(obn_dataset* ds) : ds_(ds) {}
  void ci_pvalues(const TupleBatch& b, size_t begin, size_t end, int kind, double* p) override;
  bool pc_round(const TupleBatch& b, const std::vector<uint32_t>& slot_of_test, size_t n_slots, int kind, double level,
                bool shard, std::vector<int32_t>& first) override;
  bool pc_round_submit(const TupleBatch& b, const std::vector<uint32_t>& slot_of_test, size_t n_slots, int kind,
                       double level, bool shard, bool chain) override;
  void pc_round_collect(std::vector<int32_t>& first) override;
  void subset_scores(const TupleBatch& b, size_t begin, size_t end, double ess, int exact_order,
                     double* out) override;
  void child_scores(const TupleBatch& b, const uint32_t* child, double ess, int exact_order, double* out) override;

 private:
  obn_dataset* ds_;
  std::vector<uint64_t> tmp_off_;
  size_t pending_slots_[2] = {0, 0};   // FIFO of the rounds in flight
  int pending_head_ = 0, pending_count_ = 0;
};

// Collectives between the per-GPU processes.  Default = single rank.
struct Comm {
  int rank = 0;
  int world = 1;
  // the library's own NCCL communicator (obn_comm_init): sharded PC rounds reduce on the device inside
  // obn_pc_round, the all-gathers below go through obn_comm_allgather_inplace
  bool native = false;
  // in-place elementwise MIN over ranks
  std::function<void(int32_t*, size_t)> allreduce_min_i32;
  // gather `bytes` bytes from every rank into recv (world * bytes), rank order
  std::function<void(const void*, void*, size_t)> allgather;
  // in-place all-gather of a DEVICE buffer of world * bytes_per_rank bytes: rank r contributes
  // the r-th slice (NCCL over NVLink; the data never visits the host)
  std::function<void(uintptr_t, size_t)> allgather_dev;
};

// The communicator the ctx of this process owns (after obn_comm_init); world == 1 without one.
Comm native_comm();

// contiguous shard [begin, end) of n items for rank r of w
inline void shard_range(size_t n, int r, int w, size_t* begin, size_t* end) {
  const size_t base = n / (size_t)w, rem = n % (size_t)w;
  *begin = (size_t)r * base + std::min<size_t>((size_t)r, rem);
  *end = *begin + base + ((size_t)r < rem ? 1 : 0);
}

}  // namespace obn
