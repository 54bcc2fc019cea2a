// obn_capi.cu -- the extern "C" layer (include/openbnsl_b200.h).  This is the
// translation unit that takes the place of the reference's src/template_cuda.cu.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <iterator>
#include <map>
#include <thread>
#include <unordered_map>

#include "obn_lattice.cuh"
#include "obn_host.hpp"

namespace obn {
thread_local std::string g_last_error;
}
using namespace obn;

// Fill the new rows on the host (glibc lgamma, the reference's libm; threaded for long tables),
// grow the device table if needed, and upload.
int obn::LutCache::append(obn_ctx* c, const std::vector<double>& avals) {
  const size_t add = avals.size();
  std::vector<double> host(add * stride);
  const size_t total = add * stride;
  auto fill = [&](size_t lo, size_t hi) {
    for (size_t q = lo; q < hi; ++q) {
      const size_t j = q / stride, m = q - j * stride;
      int sg;
      const double a = avals[j];
      host[q] = lgamma_r((double)m + a, &sg) - lgamma_r(a, &sg);      // local_score.h:53,:59
    }
  };
  unsigned nt = total > (1u << 16) ? std::min(16u, std::max(1u, std::thread::hardware_concurrency())) : 1u;
  if (nt <= 1) {
    fill(0, total);
  } else {
    std::vector<std::thread> th;
    const size_t step = (total + nt - 1) / nt;
    for (unsigned i = 0; i < nt; ++i) th.emplace_back(fill, std::min(total, i * step), std::min(total, (i + 1) * step));
    for (auto& t : th) t.join();
  }
  if (rows + add > cap_rows || lut_stride_of_alloc != stride) {
    const size_t new_cap = std::max(rows + add, cap_rows * 2);
    double* nd = nullptr;
    OBN_CUDA(cudaMalloc((void**)&nd, new_cap * stride * sizeof(double)));
    if (dev && rows && lut_stride_of_alloc == stride)
      OBN_CUDA(cudaMemcpyAsync(nd, dev, rows * stride * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
    OBN_CUDA(cudaStreamSynchronize(c->stream));
    if (dev) cudaFree(dev);
    dev = nd;
    cap_rows = new_cap;
    lut_stride_of_alloc = stride;
  }
  OBN_CUDA(cudaMemcpyAsync(dev + rows * stride, host.data(), total * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  OBN_CUDA(cudaStreamSynchronize(c->stream));
  rows += add;
  return OBN_OK;
}

namespace obn {
// Two rows per byte (see BatchArgs::data4): one thread per 16-byte chunk = 32 rows of one column.
__global__ void pack4_kernel(const uint8_t* data, size_t stride, uint8_t* data4, size_t stride4, size_t n_vars) {
  const size_t chunks = stride4 >> 4;
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= chunks * n_vars) return;
  const size_t v = i / chunks, ch = i - v * chunks;
  const uint4 lo = *reinterpret_cast<const uint4*>(data + v * stride + ch * 32);
  const uint4 hi = *reinterpret_cast<const uint4*>(data + v * stride + ch * 32 + 16);
  const uint32_t m = 0x0f0f0f0fu;
  *reinterpret_cast<uint4*>(data4 + v * stride4 + ch * 16) =
      make_uint4((lo.x & m) | ((hi.x & m) << 4), (lo.y & m) | ((hi.y & m) << 4), (lo.z & m) | ((hi.z & m) << 4),
                 (lo.w & m) | ((hi.w & m) << 4));
}
}  // namespace obn

extern "C" {

const char* obn_last_error(void) { return g_last_error.c_str(); }
const char* obn_version(void) { return "openbnsl_b200 0.1 (sm_100a)"; }

int obn_ctx_create(int device, obn_ctx** out) {
  if (!out) return fail(OBN_ERR_INVALID, "obn_ctx_create: out is NULL");
  *out = nullptr;
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0)
    return fail(OBN_ERR_CUDA, "no usable CUDA device (%s); this library has no CPU fallback",
                e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  if (device < 0 || device >= n_dev)
    return fail(OBN_ERR_INVALID, "device %d out of range [0,%d)", device, n_dev);
  OBN_CUDA(cudaSetDevice(device));
  obn_ctx* c = new obn_ctx();
  c->device = device;
  int v = 0;
  OBN_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device));
  c->sm_count = v;
  OBN_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
  c->smem_optin = (size_t)v;
  OBN_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  for (int j = 0; j < obn_ctx::kAuxStreams; ++j) {
    OBN_CUDA(cudaStreamCreateWithFlags(&c->aux[j], cudaStreamNonBlocking));
    OBN_CUDA(cudaEventCreateWithFlags(&c->ev_join[j], cudaEventDisableTiming));
  }
  OBN_CUDA(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
  OBN_CUDA(cudaEventCreate(&c->ev0));
  OBN_CUDA(cudaEventCreate(&c->ev1));
  for (obn_ctx::Staging& G : c->st) {
    OBN_CUDA(cudaEventCreate(&G.e0));
    OBN_CUDA(cudaEventCreate(&G.e1));
    OBN_CUDA(cudaEventCreateWithFlags(&G.done, cudaEventDisableTiming));
    OBN_CUDA(cudaEventCreateWithFlags(&G.uploaded, cudaEventDisableTiming));
  }
  OBN_CUDA(cudaStreamCreateWithFlags(&c->up_stream, cudaStreamNonBlocking));
  *out = c;
  return OBN_OK;
}

void obn_ctx_destroy(obn_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  obn_comm_destroy(c);
  for (DevBuf* b : {&c->d_scratch, &c->d_tickets, &c->d_scalar, &c->d_partials}) b->release();
  for (obn_ctx::Staging& G : c->st) {
    G.d_in.release();
    G.d_out.release();
    G.h_in.release();
    G.h_out.release();
    if (G.e0) cudaEventDestroy(G.e0);
    if (G.e1) cudaEventDestroy(G.e1);
    if (G.done) cudaEventDestroy(G.done);
    if (G.uploaded) cudaEventDestroy(G.uploaded);
  }
  if (c->up_stream) cudaStreamDestroy(c->up_stream);
  c->lut.release();
  c->es_pool.release();
  if (c->ds_cache) cudaFree(c->ds_cache);
  for (int j = 0; j < obn_ctx::kAuxStreams; ++j) {
    cudaStreamDestroy(c->aux[j]);
    cudaEventDestroy(c->ev_join[j]);
  }
  cudaEventDestroy(c->ev_fork);
  cudaEventDestroy(c->ev0);
  cudaEventDestroy(c->ev1);
  cudaStreamDestroy(c->stream);
  delete c;
}

void* obn_ctx_stream(obn_ctx* c) { return c ? (void*)c->stream : nullptr; }

int obn_ctx_set_option(obn_ctx* c, const char* name, double value) {
  if (!c || !name) return fail(OBN_ERR_INVALID, "obn_ctx_set_option: NULL argument");
  const std::string k(name);
  const int v = (int)value;
  if (k == "threads_per_sm") c->tune.threads_per_sm = std::max(32, v);
  else if (k == "force_threads") c->tune.force_threads = v;
  else if (k == "force_splits") c->tune.force_splits = v;
  else if (k == "streams") c->tune.streams = std::max(1, v);
  else if (k == "largest_first") c->tune.largest_first = v;
  else if (k == "rep") c->tune.rep = v;
  else if (k == "chunk") c->tune.chunk = v;
  else if (k == "pack4") c->tune.pack4 = v;
  else if (k == "min_waves") c->tune.min_waves = std::max(0, v);
  else if (k == "min_class") c->tune.min_class = std::max(0, std::min(4, v));
  else if (k == "lut_max_mb") c->tune.lut_max_mb = std::max(1, v);
  else if (k == "lattice_budget_gb") c->tune.lattice_budget_gb = std::max(1, v);
  else if (k == "lattice_max_l") c->tune.lattice_max_l = std::max(1, v);
  else if (k == "lattice_bottom") c->tune.lattice_bottom = v;
  else if (k == "lattice_reorder") c->tune.lattice_reorder = v;
  else if (k == "lattice_graph") c->tune.lattice_graph = v;
  else if (k == "lattice_lanes") c->tune.lattice_lanes = v;
  else if (k == "lattice_no_lut") c->tune.lattice_no_lut = v;
  else if (k == "debug") c->tune.debug = v;
  else return fail(OBN_ERR_INVALID, "unknown option '%s'", name);
  return OBN_OK;
}

int obn_dataset_upload(obn_ctx* c, const uint8_t* col_major, size_t n_vars, size_t n_rows,
                       size_t row_stride, const uint32_t* cards, obn_dataset** out) {
  if (!c || !out || !cards || (!col_major && n_vars * n_rows))
    return fail(OBN_ERR_INVALID, "obn_dataset_upload: NULL argument");
  *out = nullptr;
  if (row_stride < n_rows) return fail(OBN_ERR_INVALID, "row_stride < n_rows");
  if (n_rows >= (1ull << 31)) return fail(OBN_ERR_UNSUPPORTED, "n_rows must be < 2^31 (int32 counts)");
  if (n_vars >= (1ull << 31)) return fail(OBN_ERR_UNSUPPORTED, "too many variables");
  for (size_t v = 0; v < n_vars; ++v)
    if (cards[v] == 0 || cards[v] > 256)
      return fail(OBN_ERR_INVALID, "cardinality of variable %zu is %u, must be in [1,256]", v, cards[v]);
  OBN_CUDA(cudaSetDevice(c->device));
  obn_dataset* d = new obn_dataset();
  d->ctx = c;
  d->n_vars = n_vars;
  d->n_rows = n_rows;
  d->stride = ((n_rows + 127) / 128) * 128;
  if (d->stride == 0) d->stride = 128;
  d->cards.assign(cards, cards + n_vars);
  // behind the byte matrix: the nibble-packed copy (half the size), when any variable can use it
  bool any_small = false;
  for (size_t v = 0; v < n_vars; ++v) any_small |= cards[v] <= 16;
  const size_t byte_bytes = (std::max<size_t>(d->stride * n_vars, 128) + 255) & ~(size_t)255;
  const bool pack = c->tune.pack4 && any_small && n_rows > 0;
  d->stride4 = pack ? d->stride / 2 : 0;
  const size_t bytes = byte_bytes + (pack ? d->stride4 * n_vars : 0);
  if (c->ds_cache && c->ds_cache_bytes >= bytes) {       // reuse the buffer of the last destroyed dataset
    d->data = c->ds_cache;
    d->capacity = c->ds_cache_bytes;
    c->ds_cache = nullptr;
    c->ds_cache_bytes = 0;
  } else {
    cudaError_t e = cudaMalloc((void**)&d->data, bytes);
    if (e != cudaSuccess) {
      delete d;
      return fail(OBN_ERR_OOM, "cudaMalloc(%zu) for the dataset failed: %s", bytes, cudaGetErrorString(e));
    }
    d->capacity = bytes;
  }
  // padding bytes are zero: rows n_rows .. stride of every column count into cell 0 and are
  // subtracted again by the kernel; vector reads never leave the column
  if (n_vars == 0 || n_rows == 0) {
    OBN_CUDA(cudaMemsetAsync(d->data, 0, bytes, c->stream));
  } else {
    if (d->stride > n_rows)
      OBN_CUDA(cudaMemset2DAsync(d->data + n_rows, d->stride, 0, d->stride - n_rows, n_vars, c->stream));
    OBN_CUDA(cudaMemcpy2DAsync(d->data, d->stride, col_major, row_stride, n_rows, n_vars,
                               cudaMemcpyHostToDevice, c->stream));
  }
  if (pack) {
    d->data4 = d->data + byte_bytes;
    const size_t work = (d->stride4 >> 4) * n_vars;
    obn::pack4_kernel<<<(unsigned)((work + 255) / 256), 256, 0, c->stream>>>(d->data, d->stride, d->data4, d->stride4, n_vars);
    OBN_CUDA(cudaGetLastError());
  }
  OBN_CUDA(cudaStreamSynchronize(c->stream));
  c->prof.h2d_bytes += (double)(n_vars * n_rows);
  *out = d;
  return OBN_OK;
}

void obn_dataset_destroy(obn_dataset* d) {
  if (!d) return;
  obn_ctx* c = d->ctx;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  if (c->ds_cache) cudaFree(c->ds_cache);              // keep exactly one buffer for the next upload
  c->ds_cache = d->data;
  c->ds_cache_bytes = d->capacity;
  delete d;
}

int obn_host_pin(obn_ctx* c, const void* p, size_t bytes) {
  if (!c || !p || !bytes) return fail(OBN_ERR_INVALID, "obn_host_pin: NULL argument");
  OBN_CUDA(cudaSetDevice(c->device));
  cudaError_t e = cudaHostRegister(const_cast<void*>(p), bytes, cudaHostRegisterDefault);
  if (e == cudaErrorHostMemoryAlreadyRegistered) {
    cudaGetLastError();
    return OBN_OK;
  }
  if (e != cudaSuccess) return fail(OBN_ERR_CUDA, "cudaHostRegister failed: %s", cudaGetErrorString(e));
  return OBN_OK;
}

int obn_host_unpin(obn_ctx* c, const void* p) {
  if (!c || !p) return fail(OBN_ERR_INVALID, "obn_host_unpin: NULL argument");
  cudaSetDevice(c->device);
  cudaError_t e = cudaHostUnregister(const_cast<void*>(p));
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(OBN_ERR_CUDA, "cudaHostUnregister failed: %s", cudaGetErrorString(e));
  }
  return OBN_OK;
}

size_t obn_dataset_n_vars(const obn_dataset* d) { return d ? d->n_vars : 0; }
size_t obn_dataset_n_rows(const obn_dataset* d) { return d ? d->n_rows : 0; }

int obn_profile_enable(obn_ctx* c, int on) {
  if (!c) return fail(OBN_ERR_INVALID, "ctx is NULL");
  c->profile_on = on != 0;
  return OBN_OK;
}
int obn_profile_read(obn_ctx* c, obn_profile* out, int reset) {
  if (!c || !out) return fail(OBN_ERR_INVALID, "NULL argument");
  *out = c->prof;
  if (reset) c->prof = obn_profile{};
  return OBN_OK;
}

}  // extern "C"

#include "obn_ci_groups.inl"
#include "obn_comm.inl"
#include "obn_batch.inl"
#include "obn_es.inl"
#include "obn_lattice.inl"
#include "obn_static.inl"
