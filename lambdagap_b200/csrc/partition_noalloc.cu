// partition_noalloc.cu -- gbl_partition_rows_noalloc: the stable two-way row partition for callers that must not
// allocate device memory.
//
// Who this is for.  Two kinds of caller cannot let a library call cudaMalloc on their behalf:
//   * hosts that hand every byte of HBM to their own pool allocator at start-up (a training framework, a
//     database engine, another booster sharing the GPU): a cudaMalloc from inside a library then fails, or
//     fragments what the pool thought it owned;
//   * callers on a latency-critical path: cudaMalloc / cudaFree synchronise the whole device, which stalls
//     every other stream the process is running.
// The regular learner (partition.cu) keeps its per-CTA left counts in a cudaMalloc'ed buffer owned by the learner
// handle.  This entry point does the same partition with that scratch in a fixed array that is part of the loaded
// library image instead, so the call itself allocates nothing: no cudaMalloc, no cudaFree, no library routine
// that allocates.  The caller passes every input and output buffer.
//
// The price of a static array, stated plainly:
//   * a fixed ceiling: at most kNoAllocMaxRows rows per call (larger requests are refused with an error);
//   * the array (kNoAllocScratchInts * 4 bytes = 16.25 KB) exists once per GPU that loads this library and stays
//     until the process exits;
//   * every call on a GPU shares it.  It still holds the previous call's numbers, so each call first clears the
//     cells that must start at zero.  Two calls that overlap on the same GPU (different streams, or different host
//     threads) race on it -- issue them one after the other.
//   * first use on a device: the first call made while a given device is current looks up the array's device
//     address (cudaGetSymbolAddress) and remembers it.  Make that first call before starting any stream capture.
#include <mutex>

#include "common.cuh"

namespace gbl {

constexpr int kNoAllocThreads = 256;
constexpr int kNoAllocRowsPerCta = 4096;
constexpr int kNoAllocMaxRows = 1 << 24;                                   // 16,777,216 rows per call
constexpr int kNoAllocMaxCtas = kNoAllocMaxRows / kNoAllocRowsPerCta;      // 4096
constexpr int kNoAllocHeader = 64;                                         // ints; cell 0 = running left total
constexpr int kNoAllocScratchInts = kNoAllocHeader + kNoAllocMaxCtas;      // 4160 ints = 16,640 bytes
constexpr int kNoAllocMaxDevices = 64;

// The scratch itself.  In host code this name is only a stand-in for the symbol, NOT a device address: it must
// never be passed to a kernel directly (the kernel would fault, or on systems where the GPU can reach host memory
// silently use the wrong bytes).  Host code below resolves the real address with cudaGetSymbolAddress.
static __device__ __align__(256) int g_noalloc_scratch[kNoAllocScratchInts];

static std::mutex g_noalloc_mutex;
static int* g_noalloc_addr[kNoAllocMaxDevices] = {nullptr};

// Set-up, once per device, with that device current: resolve the device address of the static array.
static int* noalloc_scratch_for_current_device() {
  int dev = -1;
  GBL_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) fatal("gbl_partition_rows_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (g_noalloc_addr[dev] == nullptr) {
    int* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
      fatal("gbl_partition_rows_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
    }
    g_noalloc_addr[dev] = ptr;
  }
  return g_noalloc_addr[dev];
}

struct NoAllocRule {
  const uint8_t* col;
  int bits;
  int threshold;
  int missing_bin;  // -1 = none
  int default_left;
};

// same decision as DenseBin::SplitInner in mapper-bin terms (see partition.cu)
__device__ inline bool noalloc_goes_left(const NoAllocRule& r, int row) {
  const int bin = read_col_bin(r.col, r.bits, row);
  return bin == r.missing_bin ? (r.default_left != 0) : bin <= r.threshold;
}

// pass 1: per-CTA left counts -> scratch[header + cta]; running total -> scratch[0] (cleared by the host first)
__global__ void __launch_bounds__(kNoAllocThreads) noalloc_count_kernel(NoAllocRule rule, const int* __restrict__ indices,
                                                                       int n, int* __restrict__ scratch) {
  const int c_begin = blockIdx.x * kNoAllocRowsPerCta;
  const int c_end = min(n, c_begin + kNoAllocRowsPerCta);
  int cnt = 0;
  for (int i = c_begin + threadIdx.x; i < c_end; i += kNoAllocThreads) {
    const int row = indices != nullptr ? indices[i] : i;
    cnt += noalloc_goes_left(rule, row) ? 1 : 0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  __shared__ int s[kNoAllocThreads / 32];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int tot = 0;
    for (int w = 0; w < kNoAllocThreads / 32; ++w) tot += s[w];
    scratch[kNoAllocHeader + blockIdx.x] = tot;
    atomicAdd(&scratch[0], tot);
  }
}

// pass 2: stable scatter (left block first, both blocks keep their order) + left count for the caller
__global__ void __launch_bounds__(kNoAllocThreads) noalloc_scatter_kernel(NoAllocRule rule, const int* __restrict__ indices,
                                                                         int n, const int* __restrict__ scratch,
                                                                         int* __restrict__ out, int* __restrict__ out_left) {
  __shared__ int s_red[kNoAllocThreads / 32];
  __shared__ int s_wl[kNoAllocThreads / 32];
  __shared__ int s_prefix;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int before = 0;
  for (int b = tid; b < static_cast<int>(blockIdx.x); b += kNoAllocThreads) before += scratch[kNoAllocHeader + b];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
  if (lane == 0) s_red[warp] = before;
  __syncthreads();
  if (tid == 0) {
    int v = 0;
    for (int w = 0; w < kNoAllocThreads / 32; ++w) v += s_red[w];
    s_prefix = v;
  }
  __syncthreads();
  const int left_total = scratch[0];
  if (blockIdx.x == 0 && tid == 0) *out_left = left_total;
  const int c_begin = blockIdx.x * kNoAllocRowsPerCta;
  const int c_end = min(n, c_begin + kNoAllocRowsPerCta);
  int left_run = s_prefix;
  int right_run = c_begin - s_prefix;
  for (int base = c_begin; base < c_end; base += kNoAllocThreads) {
    const int i = base + tid;
    const bool valid = i < c_end;
    int row = 0;
    bool left = false;
    if (valid) {
      row = indices != nullptr ? indices[i] : i;
      left = noalloc_goes_left(rule, row);
    }
    const unsigned bl = __ballot_sync(0xffffffffu, valid && left);
    const unsigned bv = __ballot_sync(0xffffffffu, valid);
    if (lane == 0) s_wl[warp] = __popc(bl);
    __syncthreads();
    int wl_before = 0, tile_left = 0;
#pragma unroll
    for (int w = 0; w < kNoAllocThreads / 32; ++w) {
      const int cw = s_wl[w];
      if (w < warp) wl_before += cw;
      tile_left += cw;
    }
    const unsigned lt = (1u << lane) - 1u;
    if (valid) {
      if (left) {
        out[left_run + wl_before + __popc(bl & lt)] = row;
      } else {
        out[left_total + right_run + warp * 32 - wl_before + __popc((bv & ~bl) & lt)] = row;
      }
    }
    left_run += tile_left;
    right_run += min(kNoAllocThreads, c_end - base) - tile_left;
    __syncthreads();
  }
}

}  // namespace gbl

using namespace gbl;

extern "C" {

int gbl_partition_rows_noalloc(const gbl_dataset* ds_, int32_t feature, uint32_t threshold_bin, int32_t default_left,
                               const int32_t* indices_dev, int32_t num_indices, int32_t* out_indices_dev,
                               int32_t* out_left_count_dev, void* stream) {
  GBL_API_BEGIN();
  const Dataset* ds = reinterpret_cast<const Dataset*>(ds_);
  if (ds == nullptr || out_indices_dev == nullptr || out_left_count_dev == nullptr)
    fatal("gbl_partition_rows_noalloc: null argument");
  if (feature < 0 || feature >= ds->num_features) fatal("gbl_partition_rows_noalloc: feature %d out of range", feature);
  if (num_indices < 0) fatal("gbl_partition_rows_noalloc: negative row count");
  if (num_indices > kNoAllocMaxRows)
    fatal("gbl_partition_rows_noalloc: %d rows exceed the fixed scratch of this entry point (at most %d rows per call); "
          "use the learner, which sizes its own scratch", num_indices, kNoAllocMaxRows);
  if (indices_dev == nullptr && num_indices > ds->num_data) fatal("gbl_partition_rows_noalloc: more rows than the dataset holds");
  if (indices_dev != nullptr && num_indices > 0) {
    // the scatter pass reads `indices` while other CTAs already write `out`: an in-place (or overlapping) call would
    // silently corrupt the result
    const uintptr_t a = reinterpret_cast<uintptr_t>(indices_dev), b = reinterpret_cast<uintptr_t>(out_indices_dev);
    const uintptr_t bytes = sizeof(int32_t) * static_cast<uintptr_t>(num_indices);
    if (a < b + bytes && b < a + bytes)
      fatal("gbl_partition_rows_noalloc: indices_dev and out_indices_dev overlap; the partition is not in place, pass two distinct buffers");
  }
  DeviceGuard guard(ds->device);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int* scratch = noalloc_scratch_for_current_device();  // first call on this device resolves the address
  // the running total must start at zero; the per-CTA cells are overwritten before they are read
  GBL_CUDA(cudaMemsetAsync(scratch, 0, sizeof(int) * kNoAllocHeader, s));
  if (num_indices == 0) {
    GBL_CUDA(cudaMemsetAsync(out_left_count_dev, 0, sizeof(int), s));
  } else {
    const FeatMeta& m = ds->meta_host[feature];
    NoAllocRule rule;
    rule.col = ds->col_bins + m.col_off;
    rule.bits = m.bits;
    rule.threshold = static_cast<int>(threshold_bin);
    rule.missing_bin = -1;
    if (m.missing_type == GBL_MISSING_ZERO) rule.missing_bin = m.default_bin;
    if (m.missing_type == GBL_MISSING_NAN) rule.missing_bin = m.num_bin - 1;
    rule.default_left = default_left;
    const int ctas = (num_indices + kNoAllocRowsPerCta - 1) / kNoAllocRowsPerCta;
    GBL_LAUNCH(noalloc_count_kernel, ctas, kNoAllocThreads, 0, s, rule, indices_dev, num_indices, scratch);
    GBL_LAUNCH(noalloc_scatter_kernel, ctas, kNoAllocThreads, 0, s, rule, indices_dev, num_indices, scratch,
               out_indices_dev, out_left_count_dev);
  }
  GBL_API_END();
}

}  // extern "C"
