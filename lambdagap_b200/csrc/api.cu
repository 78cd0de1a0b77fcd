// api.cu -- error plumbing and small device-memory helpers of the C ABI.
//
// Error convention mirrors the reference C API: every entry point returns 0 / -1 and keeps the message in a
// thread-local buffer (API_BEGIN/API_END + LGBM_GetLastError, src/c_api.cpp:38-54); exceptions never cross
// the extern "C" boundary.
#include <cstring>

#include <cstdlib>

#include "common.cuh"

namespace gbl {

std::atomic<long long> g_kernel_launches{0};

bool debug_sync_enabled() {
  static const bool on = [] {
    const char* e = std::getenv("GBL_DEBUG_SYNC");
    return e != nullptr && e[0] == '1';
  }();
  return on;
}
static thread_local char g_last_error[1024] = "everything is fine";

void set_last_error(const char* msg) {
  std::strncpy(g_last_error, msg, sizeof(g_last_error) - 1);
  g_last_error[sizeof(g_last_error) - 1] = '\0';
}

void fatal(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  throw GblError(buf);
}

__global__ void fill_f64_kernel(double* dst, double v, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    dst[i] = v;
}

}  // namespace gbl

using namespace gbl;

extern "C" {

const char* gbl_last_error(void) { return g_last_error; }

int gbl_device_count(int* out_count) {
  GBL_API_BEGIN();
  if (out_count == nullptr) fatal("gbl_device_count: null argument");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    cudaGetLastError();
    n = 0;
  }
  int usable = 0;
  for (int d = 0; d < n; ++d) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, d) == cudaSuccess && p.major == 10) ++usable;
  }
  *out_count = usable;
  GBL_API_END();
}

int64_t gbl_kernel_launch_count(void) { return g_kernel_launches.load(); }

int gbl_device_malloc(int32_t device_id, size_t bytes, void** out_dev) {
  GBL_API_BEGIN();
  if (out_dev == nullptr) fatal("gbl_device_malloc: null argument");
  DeviceGuard guard(device_id);
  GBL_CUDA(cudaMalloc(out_dev, bytes));
  GBL_API_END();
}

int gbl_device_free(int32_t device_id, void* ptr_dev) {
  GBL_API_BEGIN();
  DeviceGuard guard(device_id);
  GBL_CUDA(cudaFree(ptr_dev));
  GBL_API_END();
}

int gbl_host_register(void* ptr_host, size_t bytes) {
  GBL_API_BEGIN();
  if (ptr_host == nullptr || bytes == 0) fatal("gbl_host_register: null argument");
  const cudaError_t e = cudaHostRegister(ptr_host, bytes, cudaHostRegisterPortable);
  if (e == cudaErrorHostMemoryAlreadyRegistered) {
    cudaGetLastError();
  } else if (e != cudaSuccess) {
    cudaGetLastError();
    fatal("gbl_host_register: %s", cudaGetErrorString(e));
  }
  GBL_API_END();
}

int gbl_host_unregister(void* ptr_host) {
  GBL_API_BEGIN();
  if (ptr_host != nullptr && cudaHostUnregister(ptr_host) != cudaSuccess) cudaGetLastError();
  GBL_API_END();
}

int gbl_memcpy_h2d(int32_t device_id, void* dst_dev, const void* src_host, size_t bytes) {
  GBL_API_BEGIN();
  DeviceGuard guard(device_id);
  GBL_CUDA(cudaMemcpy(dst_dev, src_host, bytes, cudaMemcpyHostToDevice));
  GBL_API_END();
}

int gbl_memcpy_d2h(int32_t device_id, void* dst_host, const void* src_dev, size_t bytes) {
  GBL_API_BEGIN();
  DeviceGuard guard(device_id);
  GBL_CUDA(cudaMemcpy(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost));
  GBL_API_END();
}

int gbl_device_fill_f64(int32_t device_id, double* dst_dev, double value, int64_t n) {
  GBL_API_BEGIN();
  DeviceGuard guard(device_id);
  if (n > 0) {
    const int grid = static_cast<int>(std::min<int64_t>((n + 255) / 256, kNumSM * 8));
    GBL_LAUNCH(fill_f64_kernel, grid, 256, 0, 0, dst_dev, value, static_cast<long long>(n));
    GBL_CUDA(cudaDeviceSynchronize());
  }
  GBL_API_END();
}

int gbl_device_synchronize(int32_t device_id) {
  GBL_API_BEGIN();
  DeviceGuard guard(device_id);
  GBL_CUDA(cudaDeviceSynchronize());
  GBL_API_END();
}

}  // extern "C"
