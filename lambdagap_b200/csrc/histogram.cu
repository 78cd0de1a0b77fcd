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
#include <cstdlib>

#include "kernels.cuh"

namespace gbl {

constexpr int kHistThreads = 1024;
constexpr int kHistWarps = kHistThreads / 32;
constexpr int kMinChunk = 1024;   // rows per CTA below which fewer CTAs are used
constexpr int kPlane = kMaxBin * kTile;  // words per smem plane

// Non-returning shared-memory atomic add on a 32-bit shared-window address (SASS: ATOMS.ADD RZ).
__device__ __forceinline__ void red_shared_add(unsigned addr, int v) {
  asm volatile("red.shared.add.s32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
// Hides where a value came from so that ptxas keeps it in one register (pair) instead of re-deriving it from
// uniform + per-lane parts on every use (keeps the per-row address math at one IMAD.WIDE / one LEA).
__device__ __forceinline__ const uint8_t* opaque_ptr(const uint8_t* p) {
  asm volatile("" : "+l"(p));
  return p;
}
__device__ __forceinline__ unsigned opaque_u32(unsigned v) {
  asm volatile("" : "+r"(v));
  return v;
}

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

// Rows per chunk: one chunk per CTA when it fits under the flush cap, otherwise k equal rounds of chunks so that
// every CTA gets the same number of chunks (no tail wave).
__device__ inline int chunk_rows(int n, int grid_x, int cap) {
  int c = (n + grid_x - 1) / grid_x;
  if (c > cap) {
    const int k = (c + cap - 1) / cap;
    c = (n + grid_x * k - 1) / (grid_x * k);
  }
  c = max(c, kMinChunk);
  c = (c + 255) & ~255;
  return min(c, cap);
}

// ------------------------------------------------------------------------------------------ quantized mode
// smem: int32 hist[256 bins][32 lanes] + per-warp staging of 32 rows {row, packed (grad << 16) + hess}.
// Hot loop per row and warp: LDS.64 (row info, broadcast) + LDG.U8 (32 bins = one sector) + ATOMS.ADD.
__global__ void __launch_bounds__(kHistThreads, 1) hist_quant_kernel(LearnerDev d, long long* __restrict__ out,
                                                                     const int* explicit_indices, int explicit_count,
                                                                     int flush_cap) {
  extern __shared__ int sh[];
  int2* stage_all = reinterpret_cast<int2*>(sh + kPlane);
  const HistJob job = resolve_job(d, explicit_indices, explicit_count);
  if (job.count == 0) return;
  const int tile = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int chunk = chunk_rows(job.count, gridDim.x, flush_cap);
  const int num_chunks = (job.count + chunk - 1) / chunk;
  if (static_cast<int>(blockIdx.x) >= num_chunks) return;

  for (int e = threadIdx.x; e < kPlane; e += kHistThreads) sh[e] = 0;
  __syncthreads();

  const uint8_t* rp = opaque_ptr(d.row_bins + tile * kTile + lane);
  const unsigned rs = static_cast<unsigned>(d.row_stride);
  int2* st = stage_all + warp * 32;
  const unsigned col = opaque_u32(static_cast<unsigned>(__cvta_generic_to_shared(sh + lane)));
  const int lanes_used = min(kTile, d.p.num_features - tile * kTile);
  long long* __restrict__ dst = out + static_cast<long long>(tile) * kPlane;

  for (int c = blockIdx.x; c < num_chunks; c += gridDim.x) {
    const int c_begin = c * chunk;
    const int c_end = min(job.count, c_begin + chunk);
    for (int base = c_begin + warp * 32; base < c_end; base += kHistWarps * 32) {
      const int my = base + lane;
      int2 t = make_int2(0, 0);  // padding rows add 0 to bin(row 0)
      if (my < c_end) {
        const int r = job.indices != nullptr ? job.indices[job.begin + my] : my;
        const int q = d.gh_q[r];                           // int16: high byte grad (signed), low byte hess
        t = make_int2(r, ((q >> 8) << 16) + (q & 0xff));   // packed (grad << 16) + hess
      }
      st[lane] = t;
      __syncwarp();
#pragma unroll
      for (int j0 = 0; j0 < 32; j0 += 16) {
        int bins[16];
        int vals[16];
#pragma unroll
        for (int u = 0; u < 16; ++u) {
          const int2 q = st[j0 + u];
          vals[u] = q.y;
          bins[u] = __ldg(rp + static_cast<size_t>(static_cast<unsigned>(q.x)) * rs);
        }
#pragma unroll
        for (int u = 0; u < 16; ++u) red_shared_add(col + bins[u] * (kTile * 4), vals[u]);
      }
      __syncwarp();
    }
    __syncthreads();
    // flush + re-zero: widen (g16 | h16) to the reference's (g32 | h32) packing
    for (int e = threadIdx.x; e < kPlane; e += kHistThreads) {
      const int s = sh[e];
      if (s != 0) {
        sh[e] = 0;
        if ((e & 31) < lanes_used) {
          const long long g = static_cast<long long>(s >> 16);
          const long long h = static_cast<long long>(s & 0xffff);
          atomicAdd(reinterpret_cast<unsigned long long*>(dst + e), static_cast<unsigned long long>((g << 32) + h));
        }
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------ float mode
// smem: 4 planes of int32 [256][32]: g_lo, g_hi, h_lo, h_hi (16-bit limbs of the 2^k fixed-point value, so no
// carries are needed: lo limbs sum < 2^16 * 65536 rows, hi limbs < 2^15 * 65536 rows) + per-warp staging.
// Hot loop per row and warp: LDS.128 + LDG.U8 + 4 ATOMS.ADD.
constexpr int kFloatFlushCap = 65536;

__global__ void __launch_bounds__(kHistThreads, 1) hist_float_kernel(LearnerDev d, long long* __restrict__ out,
                                                                     const int* explicit_indices, int explicit_count) {
  extern __shared__ int sh[];
  int4* stage_all = reinterpret_cast<int4*>(sh + 4 * kPlane);
  const HistJob job = resolve_job(d, explicit_indices, explicit_count);
  if (job.count == 0) return;
  const int tile = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int chunk = chunk_rows(job.count, gridDim.x, kFloatFlushCap);
  const int num_chunks = (job.count + chunk - 1) / chunk;
  if (static_cast<int>(blockIdx.x) >= num_chunks) return;

  for (int e = threadIdx.x; e < 4 * kPlane; e += kHistThreads) sh[e] = 0;
  __syncthreads();

  const uint8_t* rp = opaque_ptr(d.row_bins + tile * kTile + lane);
  const unsigned rs = static_cast<unsigned>(d.row_stride);
  int4* st = stage_all + warp * 32;
  const unsigned col = opaque_u32(static_cast<unsigned>(__cvta_generic_to_shared(sh + lane)));
  const int lanes_used = min(kTile, d.p.num_features - tile * kTile);
  longlong2* __restrict__ dst = reinterpret_cast<longlong2*>(out) + static_cast<long long>(tile) * kPlane;

  for (int c = blockIdx.x; c < num_chunks; c += gridDim.x) {
    const int c_begin = c * chunk;
    const int c_end = min(job.count, c_begin + chunk);
    for (int base = c_begin + warp * 32; base < c_end; base += kHistWarps * 32) {
      const int my = base + lane;
      int4 t = make_int4(0, 0, 0, 0);
      if (my < c_end) {
        const int r = job.indices != nullptr ? job.indices[job.begin + my] : my;
        const int2 v = d.gh_fix[r];
        t = make_int4(r, v.x, v.y, 0);
      }
      st[lane] = t;
      __syncwarp();
#pragma unroll
      for (int j0 = 0; j0 < 32; j0 += 8) {
        int bins[8];
        int gv[8];
        int hv[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int4 q = st[j0 + u];
          gv[u] = q.y;
          hv[u] = q.z;
          bins[u] = __ldg(rp + static_cast<size_t>(static_cast<unsigned>(q.x)) * rs);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const unsigned a = col + bins[u] * (kTile * 4);
          red_shared_add(a, gv[u] & 0xffff);
          red_shared_add(a + kPlane * 4, gv[u] >> 16);
          red_shared_add(a + 2 * kPlane * 4, hv[u] & 0xffff);
          red_shared_add(a + 3 * kPlane * 4, hv[u] >> 16);
        }
      }
      __syncwarp();
    }
    __syncthreads();
    for (int e = threadIdx.x; e < kPlane; e += kHistThreads) {
      const long long g = (static_cast<long long>(sh[e + kPlane]) << 16) + static_cast<long long>(static_cast<unsigned>(sh[e]));
      const long long h =
          (static_cast<long long>(sh[e + 3 * kPlane]) << 16) + static_cast<long long>(static_cast<unsigned>(sh[e + 2 * kPlane]));
      if ((e & 31) < lanes_used) {
        if (g != 0) atomicAdd(reinterpret_cast<unsigned long long*>(&dst[e].x), static_cast<unsigned long long>(g));
        if (h != 0) atomicAdd(reinterpret_cast<unsigned long long*>(&dst[e].y), static_cast<unsigned long long>(h));
      }
      sh[e] = 0;
      sh[e + kPlane] = 0;
      sh[e + 2 * kPlane] = 0;
      sh[e + 3 * kPlane] = 0;
    }
    __syncthreads();
  }
}

static const int kSmemQ = kPlane * 4 + kHistWarps * 32 * 8;        // 32 KB + 8 KB staging
static const int kSmemF = 4 * kPlane * 4 + kHistWarps * 32 * 16;   // 128 KB + 16 KB staging

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
    const int gx = std::max(1, (kNumSM + tiles - 1) / tiles);
    GBL_LAUNCH(hist_quant_kernel, dim3(gx, tiles), kHistThreads, kSmemQ, s, d, out, explicit_indices, explicit_count,
               flush_cap);
  } else {
    const int gx = std::max(1, (kNumSM + tiles - 1) / tiles);
    GBL_LAUNCH(hist_float_kernel, dim3(gx, tiles), kHistThreads, kSmemF, s, d, out, explicit_indices, explicit_count);
  }
}

}  // namespace gbl
