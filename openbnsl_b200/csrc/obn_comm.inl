// obn_comm.inl -- the communicator of a multi-GPU job, owned by the ctx.
//
// One process per GPU; the tuples of a PC round / the blocks of a lattice are sharded over the
// ranks and the only exchanges are
//   ncclAllReduce(min, int32[#edges])   per PC round   (first passing candidate of every edge)
//   ncclAllGather                       per ES level / once per lattice (scores, F, DP sink choices)
// both on DEVICE buffers of the library, on the ctx's stream: between the counting kernel and the
// collective nothing visits the host.  The reference has no collective to cite; the
// contract is SURVEY 8b's.
//
// NCCL is bound at run time (dlopen): a process that already holds libnccl.so.2 (torch's copy)
// shares that one, otherwise the system library is loaded.  Single-GPU use needs no NCCL at all.
// The 128-byte unique id travels between the processes by whatever the host has (bench.py: a
// torch.distributed store; MPI, a file or a socket work as well) -- obn_comm_unique_id on rank 0,
// obn_comm_init on every rank.
// Included by obn_capi.cu before obn_batch.inl.
#include <dlfcn.h>
#include <nccl.h>

namespace obn {

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
  std::string where;
};

static NcclApi* nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api.lib ? &api : nullptr;
  tried = true;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_LOCAL);     // the copy this process already uses
  api.where = "already loaded";
  if (!h) {
    h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    api.where = "system";
  }
  if (!h) return nullptr;
  auto sym = [&](const char* n) { return dlsym(h, n); };
  api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
  api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
  api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
  api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
  api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
  api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
  api.GetVersion = (decltype(api.GetVersion))sym("ncclGetVersion");
  if (!api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.AllReduce || !api.AllGather ||
      !api.GetErrorString)
    return nullptr;
  api.lib = h;
  return &api;
}

#define OBN_NCCL(api, call)                                                                       \
  do {                                                                                            \
    ncclResult_t r_ = (call);                                                                     \
    if (r_ != ncclSuccess)                                                                        \
      return obn::fail(OBN_ERR_COMM, "%s failed: %s (%s:%d)", #call, (api)->GetErrorString(r_),  \
                       __FILE__, __LINE__);                                                       \
  } while (0)

// in place, on the ctx's stream; the caller keeps enqueueing behind it
static int comm_allreduce_min_i32(obn_ctx* c, int32_t* dev, size_t count) {
  NcclApi* api = nccl_api();
  if (!api || !c->comm) return fail(OBN_ERR_COMM, "the ctx has no communicator (obn_comm_init)");
  OBN_NCCL(api, api->AllReduce(dev, dev, count, ncclInt32, ncclMin, (ncclComm_t)c->comm, c->stream));
  ++c->prof.collectives;
  return OBN_OK;
}

// A host scalar every rank must agree on (e.g. the memory budget a lattice is planned from).
static int comm_min_u64_host(obn_ctx* c, uint64_t* value) {
  if (!c->comm || c->comm_world <= 1) return OBN_OK;
  NcclApi* api = nccl_api();
  if (!api) return fail(OBN_ERR_COMM, "NCCL is not available");
  if (int rc = c->d_scalar.reserve(16)) return rc;
  OBN_CUDA(cudaMemcpyAsync(c->d_scalar.p, value, 8, cudaMemcpyHostToDevice, c->stream));
  OBN_NCCL(api, api->AllReduce(c->d_scalar.p, c->d_scalar.p, 1, ncclUint64, ncclMin, (ncclComm_t)c->comm, c->stream));
  OBN_CUDA(cudaMemcpyAsync(value, c->d_scalar.p, 8, cudaMemcpyDeviceToHost, c->stream));
  OBN_CUDA(cudaStreamSynchronize(c->stream));
  ++c->prof.collectives;
  return OBN_OK;
}

}  // namespace obn

extern "C" {

int obn_comm_unique_id(uint8_t* id_out) {
  if (!id_out) return fail(OBN_ERR_INVALID, "id_out is NULL");
  NcclApi* api = nccl_api();
  if (!api) return fail(OBN_ERR_COMM, "libnccl.so.2 could not be loaded: %s", dlerror());
  ncclUniqueId id;
  OBN_NCCL(api, api->GetUniqueId(&id));
  static_assert(sizeof(id) == OBN_COMM_ID_BYTES, "ncclUniqueId size");
  memcpy(id_out, &id, sizeof(id));
  return OBN_OK;
}

int obn_comm_init(obn_ctx* c, const uint8_t* id_bytes, int rank, int world) {
  if (!c || !id_bytes) return fail(OBN_ERR_INVALID, "obn_comm_init: NULL argument");
  if (world < 1 || rank < 0 || rank >= world) return fail(OBN_ERR_INVALID, "bad rank/world %d/%d", rank, world);
  if (c->comm) return fail(OBN_ERR_INVALID, "the ctx already has a communicator");
  NcclApi* api = nccl_api();
  if (!api) return fail(OBN_ERR_COMM, "libnccl.so.2 could not be loaded: %s", dlerror());
  OBN_CUDA(cudaSetDevice(c->device));
  ncclUniqueId id;
  memcpy(&id, id_bytes, sizeof(id));
  ncclComm_t comm = nullptr;
  OBN_NCCL(api, api->CommInitRank(&comm, world, id, rank));
  c->comm = comm;
  c->comm_rank = rank;
  c->comm_world = world;
  return OBN_OK;
}

void obn_comm_destroy(obn_ctx* c) {
  if (!c || !c->comm) return;
  NcclApi* api = nccl_api();
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  if (api) api->CommDestroy((ncclComm_t)c->comm);
  c->comm = nullptr;
  c->comm_rank = 0;
  c->comm_world = 1;
}

int obn_comm_rank(const obn_ctx* c) { return c ? c->comm_rank : 0; }
int obn_comm_world(const obn_ctx* c) { return c ? c->comm_world : 1; }

// In-place all-gather of a device buffer of world * bytes_per_rank bytes: rank r contributes slice r.
int obn_comm_allgather_inplace(obn_ctx* c, void* dev, size_t bytes_per_rank) {
  if (!c || (!dev && bytes_per_rank)) return fail(OBN_ERR_INVALID, "obn_comm_allgather_inplace: NULL argument");
  if (!c->comm || c->comm_world <= 1) return OBN_OK;
  NcclApi* api = nccl_api();
  if (!api) return fail(OBN_ERR_COMM, "NCCL is not available");
  OBN_CUDA(cudaSetDevice(c->device));
  const auto t0 = std::chrono::steady_clock::now();
  OBN_NCCL(api, api->AllGather((const char*)dev + (size_t)c->comm_rank * bytes_per_rank, dev, bytes_per_rank, ncclInt8,
                               (ncclComm_t)c->comm, c->stream));
  OBN_CUDA(cudaStreamSynchronize(c->stream));
  c->prof.collective_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  ++c->prof.collectives;
  return OBN_OK;
}

// All-gather of small HOST payloads (the backtracked parent sets, a search's winner): staged
// through the device.
int obn_comm_allgather_host(obn_ctx* c, const void* send, void* recv, size_t bytes_per_rank) {
  if (!c || ((!send || !recv) && bytes_per_rank)) return fail(OBN_ERR_INVALID, "obn_comm_allgather_host: NULL argument");
  if (!bytes_per_rank) return OBN_OK;
  if (!c->comm || c->comm_world <= 1) {
    memcpy(recv, send, bytes_per_rank);
    return OBN_OK;
  }
  OBN_CUDA(cudaSetDevice(c->device));
  const size_t slot = (bytes_per_rank + 15) & ~(size_t)15;
  if (int rc = c->d_scalar.reserve(slot * (size_t)c->comm_world)) return rc;
  char* d = c->d_scalar.as<char>();
  OBN_CUDA(cudaMemcpyAsync(d + slot * (size_t)c->comm_rank, send, bytes_per_rank, cudaMemcpyHostToDevice, c->stream));
  if (int rc = obn_comm_allgather_inplace(c, d, slot)) return rc;
  for (int r = 0; r < c->comm_world; ++r)
    OBN_CUDA(cudaMemcpyAsync((char*)recv + bytes_per_rank * (size_t)r, d + slot * (size_t)r, bytes_per_rank,
                             cudaMemcpyDeviceToHost, c->stream));
  OBN_CUDA(cudaStreamSynchronize(c->stream));
  return OBN_OK;
}

// In-place elementwise MIN over the ranks of a HOST int32 array (staged through the device; for
// host-side drivers that keep their own per-edge state).
int obn_comm_allreduce_min_i32(obn_ctx* c, int32_t* host, size_t count) {
  if (!c || (!host && count)) return fail(OBN_ERR_INVALID, "obn_comm_allreduce_min_i32: NULL argument");
  if (!c->comm || c->comm_world <= 1 || !count) return OBN_OK;
  OBN_CUDA(cudaSetDevice(c->device));
  if (int rc = c->d_scalar.reserve(count * 4)) return rc;
  OBN_CUDA(cudaMemcpyAsync(c->d_scalar.p, host, count * 4, cudaMemcpyHostToDevice, c->stream));
  if (int rc = comm_allreduce_min_i32(c, c->d_scalar.as<int32_t>(), count)) return rc;
  OBN_CUDA(cudaMemcpyAsync(host, c->d_scalar.p, count * 4, cudaMemcpyDeviceToHost, c->stream));
  OBN_CUDA(cudaStreamSynchronize(c->stream));
  return OBN_OK;
}

}  // extern "C"
