// histogram.cu -- per-leaf gradient/hessian histogram construction (the dominant kernel).
//
// Reference: Dataset::ConstructHistograms (include/LightGBM/dataset.h:735-766, src/io/dataset.cpp:1262-1454)
//            -> DenseBin::ConstructHistogramInner / ConstructHistogramIntInner (src/io/dense_bin.hpp:99-222):
//            hist[bin(row, f)] += (grad[row], hess[row]) for every row of the leaf and every feature f.
//
// B200 design.  On sm_100a the only native shared-memory atomic is the 32-bit integer ATOMS.ADD (f32 / f64 /
// u64 shared atomics compile to ATOMS.CAST.SPIN loops), so both modes accumulate 32-bit integers:
//   * lane = feature, warp = row: one warp handles a 32-feature tile of one row at a time; lane f adds into
//     smem word  hist[bin][f], i.e. bank f -> a warp-wide ATOMS is bank-conflict free whatever the bin
//     distribution (skewed / zero-inflated features cost the same as uniform ones);
//   * bins of the row come from the row-major mirror: the warp reads one aligned 32-byte sector per
//     (row, tile) -- gather-friendly for sparse leaves, streaming for dense ones;
//   * quantized mode: one ATOMS per cell on a packed (grad << 16) + hess word, flushed to HBM as the
//     reference's packed int64 (grad << 32 | hess) with 64-bit RED atomics before the 16-bit halves can overflow;
//   * float mode: gradients are exact power-of-two fixed point (|v| <= 2^30); each of grad/hess is one returning
//     ATOMS on a low word plus a rare (|v| / 2^32) carry ATOMS on a high word -> exact 64-bit CTA partial sums,
//     flushed with 64-bit RED atomics.  Sums are order-independent, so histograms are deterministic and
//     identical for any grid shape or GPU count.
#include "kernels.cuh"

namespace gbl {

constexpr int kHistThreadsQ = 256;
constexpr int kHistThreadsF = 512;
constexpr int kMinChunk = 2048;  // rows per CTA below which fewer CTAs are used
constexpr int kUnroll = 8;

struct HistJob {
  const int* indices;  // nullptr -> identity
  int begin;
  int count;
};

__device__ inline HistJob resolve_job(const LearnerDev& d, const int* explicit_indices, int explicit_count) {
  HistJob j;
  if (explicit_count >= 0) {
    j.indices = explicit_indices;
    j.begin = 0;
    j.count = explicit_count;
    return j;
  }
  const TreeState& t = *d.ts;
  if (t.done || t.skip_hist || t.smaller < 0) {
    j.indices = nullptr;
    j.begin = 0;
    j.count = 0;
    return j;
  }
  const LeafState& L = d.leaves[t.smaller];
  j.indices = d.idx[L.buf];
  j.begin = L.begin;
  j.count = L.count;
  return j;
}

__device__ inline int chunk_rows(int n, int grid_x, int cap) {
  int c = (n + grid_x - 1) / grid_x;
  c = max(c, kMinChunk);
  c = (c + 255) & ~255;
  return min(c, cap);
}

// ------------------------------------------------------------------------------------------ quantized mode
// smem: int32 [256 bins][32 lanes]
__global__ void __launch_bounds__(kHistThreadsQ) hist_quant_kernel(LearnerDev d, long long* __restrict__ out,
                                                                   const int* explicit_indices, int explicit_count,
                                                                   int flush_cap) {
  extern __shared__ int sh[];
  const HistJob job = resolve_job(d, explicit_indices, explicit_count);
  if (job.count == 0) return;
  const int tile = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  constexpr int kWarps = kHistThreadsQ / 32;
  const bool lane_active = tile * kTile + lane < d.p.num_features;
  const int chunk = chunk_rows(job.count, gridDim.x, flush_cap);
  const int num_chunks = (job.count + chunk - 1) / chunk;
  if (static_cast<int>(blockIdx.x) >= num_chunks) return;

  for (int e = threadIdx.x; e < kMaxBin * kTile; e += kHistThreadsQ) sh[e] = 0;
  __syncthreads();

  const uint8_t* __restrict__ rp = d.row_bins + tile * kTile + lane;
  const int64_t rs = d.row_stride;
  long long* __restrict__ dst = out + static_cast<long long>(tile) * kMaxBin * kTile;

  for (int c = blockIdx.x; c < num_chunks; c += gridDim.x) {
    const int c_begin = c * chunk;
    const int c_end = min(job.count, c_begin + chunk);
    for (int base = c_begin + warp * 32; base < c_end; base += kWarps * 32) {
      const int my = base + lane;
      int r = -1;
      int v = 0;
      if (my < c_end) {
        r = job.indices != nullptr ? job.indices[job.begin + my] : my;
        const int q = d.gh_q[r];                  // int16: high byte grad (signed), low byte hess
        v = ((q >> 8) << 16) + (q & 0xff);        // packed (grad << 16) + hess
      }
      const int rows = min(32, c_end - base);
      for (int j0 = 0; j0 < rows; j0 += kUnroll) {
        int bins[kUnroll];
        int vals[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          const int rj = __shfl_sync(0xffffffffu, r, (j0 + u) & 31);
          vals[u] = __shfl_sync(0xffffffffu, v, (j0 + u) & 31);
          const bool ok = (j0 + u < rows);
          bins[u] = ok ? static_cast<int>(rp[static_cast<int64_t>(rj) * rs]) : -1;
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          if (bins[u] >= 0 && lane_active && vals[u] != 0) atomicAdd(&sh[bins[u] * kTile + lane], vals[u]);
        }
      }
    }
    __syncthreads();
    // flush + re-zero: widen (g16 | h16) to the reference's (g32 | h32) packing
    for (int e = threadIdx.x; e < kMaxBin * kTile; e += kHistThreadsQ) {
      const int s = sh[e];
      if (s != 0) {
        const long long g = static_cast<long long>(s >> 16);
        const long long h = static_cast<long long>(s & 0xffff);
        atomicAdd(reinterpret_cast<unsigned long long*>(dst + e), static_cast<unsigned long long>((g << 32) + h));
        sh[e] = 0;
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------ float mode
// smem: 4 planes of int32 [256][32]: g_lo, g_hi, h_lo, h_hi
__global__ void __launch_bounds__(kHistThreadsF) hist_float_kernel(LearnerDev d, long long* __restrict__ out,
                                                                   const int* explicit_indices, int explicit_count) {
  extern __shared__ int sh[];
  unsigned* __restrict__ g_lo = reinterpret_cast<unsigned*>(sh);
  int* __restrict__ g_hi = sh + kMaxBin * kTile;
  unsigned* __restrict__ h_lo = reinterpret_cast<unsigned*>(sh + 2 * kMaxBin * kTile);
  int* __restrict__ h_hi = sh + 3 * kMaxBin * kTile;
  const HistJob job = resolve_job(d, explicit_indices, explicit_count);
  if (job.count == 0) return;
  const int tile = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  constexpr int kWarps = kHistThreadsF / 32;
  const bool lane_active = tile * kTile + lane < d.p.num_features;
  const int chunk = chunk_rows(job.count, gridDim.x, 1 << 30);
  const int num_chunks = (job.count + chunk - 1) / chunk;
  if (static_cast<int>(blockIdx.x) >= num_chunks) return;

  for (int e = threadIdx.x; e < 4 * kMaxBin * kTile; e += kHistThreadsF) sh[e] = 0;
  __syncthreads();

  const uint8_t* __restrict__ rp = d.row_bins + tile * kTile + lane;
  const int64_t rs = d.row_stride;

  for (int c = blockIdx.x; c < num_chunks; c += gridDim.x) {
    const int c_begin = c * chunk;
    const int c_end = min(job.count, c_begin + chunk);
    for (int base = c_begin + warp * 32; base < c_end; base += kWarps * 32) {
      const int my = base + lane;
      int r = -1;
      int2 v = make_int2(0, 0);
      if (my < c_end) {
        r = job.indices != nullptr ? job.indices[job.begin + my] : my;
        v = d.gh_fix[r];
      }
      const int rows = min(32, c_end - base);
      for (int j0 = 0; j0 < rows; j0 += kUnroll) {
        int bins[kUnroll];
        int gv[kUnroll];
        int hv[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          const int rj = __shfl_sync(0xffffffffu, r, (j0 + u) & 31);
          gv[u] = __shfl_sync(0xffffffffu, v.x, (j0 + u) & 31);
          hv[u] = __shfl_sync(0xffffffffu, v.y, (j0 + u) & 31);
          const bool ok = (j0 + u < rows);
          bins[u] = ok ? static_cast<int>(rp[static_cast<int64_t>(rj) * rs]) : -1;
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          if (bins[u] >= 0 && lane_active) {
            const int a = bins[u] * kTile + lane;
            const unsigned gl = static_cast<unsigned>(gv[u]);
            const unsigned og = atomicAdd(&g_lo[a], gl);
            const int gd = (gv[u] >> 31) + ((og + gl) < og ? 1 : 0);  // sign extension + carry
            if (gd != 0) atomicAdd(&g_hi[a], gd);
            const unsigned hl = static_cast<unsigned>(hv[u]);
            const unsigned oh = atomicAdd(&h_lo[a], hl);
            const int hd = (hv[u] >> 31) + ((oh + hl) < oh ? 1 : 0);
            if (hd != 0) atomicAdd(&h_hi[a], hd);
          }
        }
      }
    }
  }
  __syncthreads();
  longlong2* __restrict__ dst = reinterpret_cast<longlong2*>(out) + static_cast<long long>(tile) * kMaxBin * kTile;
  for (int e = threadIdx.x; e < kMaxBin * kTile; e += kHistThreadsF) {
    const long long g = (static_cast<long long>(g_hi[e]) << 32) + static_cast<long long>(g_lo[e]);
    const long long h = (static_cast<long long>(h_hi[e]) << 32) + static_cast<long long>(h_lo[e]);
    if (g != 0) atomicAdd(reinterpret_cast<unsigned long long*>(&dst[e].x), static_cast<unsigned long long>(g));
    if (h != 0) atomicAdd(reinterpret_cast<unsigned long long*>(&dst[e].y), static_cast<unsigned long long>(h));
  }
}

static const int kSmemQ = kMaxBin * kTile * 4;       // 32 KB
static const int kSmemF = 4 * kMaxBin * kTile * 4;   // 128 KB

void histogram_setup() {
  GBL_CUDA(cudaFuncSetAttribute(hist_float_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemF));
  GBL_CUDA(cudaFuncSetAttribute(hist_quant_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemQ));
}

void launch_histogram(const LearnerDev& d, int slot, const int* explicit_indices, int explicit_count,
                      long long* explicit_out, cudaStream_t s) {
  long long* out = explicit_out != nullptr ? explicit_out : d.hist_pool + static_cast<long long>(slot) * d.slot_stride;
  const int tiles = d.p.num_tiles;
  if (d.p.quantized) {
    // 16-bit halves hold sums of at most flush_cap rows: |g| <= bins/2, h <= bins (gradient_discretizer.cpp:105-127)
    int flush_cap = (65535 / std::max(1, d.p.num_grad_quant_bins)) & ~255;
    flush_cap = std::max(256, std::min(flush_cap, 1 << 20));
    const int gx = std::max(1, (kNumSM * 6 + tiles - 1) / tiles);
    GBL_LAUNCH(hist_quant_kernel, dim3(gx, tiles), kHistThreadsQ, kSmemQ, s, d, out, explicit_indices, explicit_count,
               flush_cap);
  } else {
    const int gx = std::max(1, (kNumSM + tiles - 1) / tiles);
    GBL_LAUNCH(hist_float_kernel, dim3(gx, tiles), kHistThreadsF, kSmemF, s, d, out, explicit_indices, explicit_count);
  }
}

}  // namespace gbl
