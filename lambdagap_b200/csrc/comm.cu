// comm.cu -- NCCL over NVLink 5 / NVSwitch in place of the reference's socket / MPI linkers.
//
// Reference: Network::Allreduce / ReduceScatter (src/network/network.cpp:68-97,227-323) over Linkers
// (src/network/linkers_socket.cpp, linkers_mpi.cpp) with host byte buffers and C reducer callbacks
// (include/LightGBM/bin.h:48-82).  Here the payloads are exact int64 histograms that never leave HBM.
#include "comm.cuh"

#include <dlfcn.h>

#include <cstring>
#include <mutex>

namespace gbl {

// Minimal NCCL ABI (nccl.h of NCCL 2.x): only what is called here.
typedef struct {
  char internal[128];
} ncclUniqueId_t;
typedef void* ncclComm_h;
enum { kNcclSuccess = 0 };
enum { kNcclInt64 = 4, kNcclFloat32 = 7, kNcclFloat64 = 8 };  // ncclDataType_t
enum { kNcclSum = 0, kNcclMax = 2 };                          // ncclRedOp_t

struct NcclApi {
  int (*GetUniqueId)(ncclUniqueId_t*);
  int (*CommInitRank)(ncclComm_h*, int, ncclUniqueId_t, int);
  int (*CommDestroy)(ncclComm_h);
  int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_h, cudaStream_t);
  const char* (*GetErrorString)(int);
  bool loaded = false;
};

static NcclApi g_nccl;
static std::mutex g_nccl_mutex;

static NcclApi& nccl() {
  std::lock_guard<std::mutex> lock(g_nccl_mutex);
  if (g_nccl.loaded) return g_nccl;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (h == nullptr) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (h == nullptr) fatal("NCCL not found (dlopen libnccl.so.2): %s", dlerror());
  g_nccl.GetUniqueId = reinterpret_cast<decltype(g_nccl.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
  g_nccl.CommInitRank = reinterpret_cast<decltype(g_nccl.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
  g_nccl.CommDestroy = reinterpret_cast<decltype(g_nccl.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
  g_nccl.AllReduce = reinterpret_cast<decltype(g_nccl.AllReduce)>(dlsym(h, "ncclAllReduce"));
  g_nccl.GetErrorString = reinterpret_cast<decltype(g_nccl.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommDestroy || !g_nccl.AllReduce)
    fatal("NCCL symbols missing in libnccl.so.2");
  g_nccl.loaded = true;
  return g_nccl;
}

#define GBL_NCCL(call)                                                                                  \
  do {                                                                                                  \
    int r__ = (call);                                                                                   \
    if (r__ != kNcclSuccess)                                                                            \
      fatal("NCCL error %d (%s) at %s:%d", r__,                                                         \
            nccl().GetErrorString ? nccl().GetErrorString(r__) : "?", __FILE__, __LINE__);              \
  } while (0)

void comm_get_unique_id(uint8_t out[128]) {
  ncclUniqueId_t id;
  GBL_NCCL(nccl().GetUniqueId(&id));
  std::memcpy(out, id.internal, 128);
}

Comm* comm_create(const uint8_t id_bytes[128], int num_ranks, int rank, int device) {
  if (num_ranks < 1 || rank < 0 || rank >= num_ranks) fatal("comm: bad rank %d of %d", rank, num_ranks);
  DeviceGuard guard(device);
  ncclUniqueId_t id;
  std::memcpy(id.internal, id_bytes, 128);
  Comm* c = new Comm();
  c->num_ranks = num_ranks;
  c->rank = rank;
  c->device = device;
  ncclComm_h h = nullptr;
  GBL_NCCL(nccl().CommInitRank(&h, num_ranks, id, rank));
  c->nccl_comm = h;
  return c;
}

void comm_free(Comm* c) {
  if (c == nullptr) return;
  if (c->nccl_comm != nullptr) nccl().CommDestroy(c->nccl_comm);
  delete c;
}

static int to_op(RedOp op) { return op == RedOp::kSum ? kNcclSum : kNcclMax; }

void comm_allreduce_i64(Comm* c, long long* buf, size_t count, RedOp op, cudaStream_t s) {
  GBL_NCCL(nccl().AllReduce(buf, buf, count, kNcclInt64, to_op(op), c->nccl_comm, s));
}
void comm_allreduce_f32(Comm* c, float* buf, size_t count, RedOp op, cudaStream_t s) {
  GBL_NCCL(nccl().AllReduce(buf, buf, count, kNcclFloat32, to_op(op), c->nccl_comm, s));
}
void comm_allreduce_f64(Comm* c, double* buf, size_t count, RedOp op, cudaStream_t s) {
  GBL_NCCL(nccl().AllReduce(buf, buf, count, kNcclFloat64, to_op(op), c->nccl_comm, s));
}

}  // namespace gbl

using namespace gbl;

extern "C" {

int gbl_comm_get_unique_id(uint8_t out_id[128]) {
  GBL_API_BEGIN();
  if (out_id == nullptr) fatal("gbl_comm_get_unique_id: null argument");
  comm_get_unique_id(out_id);
  GBL_API_END();
}

int gbl_comm_create(const uint8_t id[128], int32_t num_ranks, int32_t rank, int32_t device_id, gbl_comm** out) {
  GBL_API_BEGIN();
  if (id == nullptr || out == nullptr) fatal("gbl_comm_create: null argument");
  *out = reinterpret_cast<gbl_comm*>(comm_create(id, num_ranks, rank, device_id));
  GBL_API_END();
}

int gbl_comm_allreduce_host_f64(gbl_comm* comm, double* inout_host, int32_t n, int32_t op) {
  GBL_API_BEGIN();
  Comm* c = reinterpret_cast<Comm*>(comm);
  if (c == nullptr || inout_host == nullptr || n <= 0) fatal("gbl_comm_allreduce_host_f64: bad argument");
  DeviceGuard guard(c->device);
  double* d = nullptr;
  GBL_CUDA(cudaMalloc(&d, sizeof(double) * n));
  GBL_CUDA(cudaMemcpy(d, inout_host, sizeof(double) * n, cudaMemcpyHostToDevice));
  comm_allreduce_f64(c, d, n, op == 1 ? RedOp::kMax : RedOp::kSum, 0);
  GBL_CUDA(cudaStreamSynchronize(0));
  GBL_CUDA(cudaMemcpy(inout_host, d, sizeof(double) * n, cudaMemcpyDeviceToHost));
  cudaFree(d);
  GBL_API_END();
}

int gbl_comm_info(const gbl_comm* comm, int32_t* out_num_ranks, int32_t* out_rank) {
  GBL_API_BEGIN();
  const Comm* c = reinterpret_cast<const Comm*>(comm);
  if (c == nullptr) fatal("gbl_comm_info: null communicator");
  if (out_num_ranks) *out_num_ranks = c->num_ranks;
  if (out_rank) *out_rank = c->rank;
  GBL_API_END();
}

int gbl_comm_free(gbl_comm* comm) {
  GBL_API_BEGIN();
  comm_free(reinterpret_cast<Comm*>(comm));
  GBL_API_END();
}

}  // extern "C"
