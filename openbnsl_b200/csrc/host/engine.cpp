#include "engine.hpp"

#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <cstring>

namespace obn {

Engine::Engine() {
  const char* d = std::getenv("OBN_DEVICE");
  if (!d) d = std::getenv("LOCAL_RANK");
  device_ = d ? std::atoi(d) : 0;
  check(obn_ctx_create(device_, &ctx_));
}
Engine::~Engine() {
  // Deliberately no obn_ctx_destroy at process exit: the CUDA runtime may
  // already be torn down when static destructors run.
}
Engine& Engine::instance() {
  static Engine e;
  return e;
}

DeviceDataset::DeviceDataset(const uint8_t* col_major, size_t n_vars, size_t n_rows, size_t stride,
                             const uint32_t* cards) {
  check(obn_dataset_upload(Engine::instance().ctx(), col_major, n_vars, n_rows, stride, cards, &ds_));
}
DeviceDataset::~DeviceDataset() { obn_dataset_destroy(ds_); }

// rebased offsets for a sub-range of a batch
static const uint64_t* rebase(const TupleBatch& b, size_t begin, size_t end, std::vector<uint64_t>& tmp) {
  if (begin == 0) return b.offsets.data();
  tmp.resize(end - begin + 1);
  const uint64_t o = b.offsets[begin];
  for (size_t i = begin; i <= end; ++i) tmp[i - begin] = b.offsets[i] - o;
  return tmp.data();
}

void GpuEvaluator::ci_pvalues(const TupleBatch& b, size_t begin, size_t end, int kind, double* p) {
  if (end <= begin) return;
  const uint64_t* off = rebase(b, begin, end, tmp_off_);
  check(obn_ci_tests(Engine::instance().ctx(), ds_, b.var_ids.data() + b.offsets[begin], off, end - begin,
                     b.x.data() + begin, b.y.data() + begin, kind, nullptr, nullptr, p));
}

bool GpuEvaluator::pc_round(const TupleBatch& b, const std::vector<uint32_t>& slot_of_test, size_t n_slots, int kind,
                            double level, bool shard, std::vector<int32_t>& first) {
  first.resize(n_slots);
  check(obn_pc_round(Engine::instance().ctx(), ds_, b.var_ids.data(), b.offsets.data(), b.size(), b.x.data(), b.y.data(),
                     slot_of_test.data(), n_slots, kind, level, shard ? 1 : 0, first.data()));
  return true;
}
bool GpuEvaluator::pc_round_submit(const TupleBatch& b, const std::vector<uint32_t>& slot_of_test, size_t n_slots, int kind,
                                   double level, bool shard, bool chain) {
  check(obn_pc_round_submit(Engine::instance().ctx(), ds_, b.var_ids.data(), b.offsets.data(), b.size(), b.x.data(),
                            b.y.data(), slot_of_test.data(), n_slots, kind, level, shard ? 1 : 0, chain ? 1 : 0));
  pending_slots_[(pending_head_ + pending_count_) % 2] = n_slots;
  ++pending_count_;
  return true;
}
void GpuEvaluator::pc_round_collect(std::vector<int32_t>& first) {
  first.assign(pending_count_ ? pending_slots_[pending_head_] : 0, INT32_MAX);
  if (pending_count_) {
    pending_head_ = (pending_head_ + 1) % 2;
    --pending_count_;
  }
  check(obn_pc_round_collect(Engine::instance().ctx(), first.data()));
}

Comm native_comm() {
  obn_ctx* c = Engine::instance().ctx();
  Comm m;
  m.rank = obn_comm_rank(c);
  m.world = obn_comm_world(c);
  m.native = true;
  m.allreduce_min_i32 = [c](int32_t* p, size_t n) { check(obn_comm_allreduce_min_i32(c, p, n)); };
  m.allgather_dev = [c](uintptr_t p, size_t bytes) { check(obn_comm_allgather_inplace(c, reinterpret_cast<void*>(p), bytes)); };
  m.allgather = [c](const void* s, void* recv, size_t bytes) { check(obn_comm_allgather_host(c, s, recv, bytes)); };
  return m;
}

void GpuEvaluator::subset_scores(const TupleBatch& b, size_t begin, size_t end, double ess, int exact_order,
                                 double* out) {
  if (end <= begin) return;
  const uint64_t* off = rebase(b, begin, end, tmp_off_);
  check(obn_bdeu_subset_scores(Engine::instance().ctx(), ds_, b.var_ids.data() + b.offsets[begin], off,
                               end - begin, ess, exact_order, out));
}

void GpuEvaluator::child_scores(const TupleBatch& b, const uint32_t* child, double ess, int exact_order, double* out) {
  if (!b.size()) return;
  check(obn_bdeu_scores(Engine::instance().ctx(), ds_, b.var_ids.data(), b.offsets.data(), b.size(), child, ess,
                        exact_order, out));
}

}  // namespace obn
