// histogram.cu -- gradient/hessian histogram construction for every node of a wave (the dominant kernel).
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
//   * float mode: gradients are exact power-of-two fixed point (|v| <= 2^30) split into 16-bit limbs, one shared
//     plane per limb (grad lo / hi, hess lo / hi): four non-returning ATOMS per cell, no carries, exact 64-bit CTA
//     partial sums at flush time (64-bit RED atomics into the node's slot).  With a constant hessian (unweighted L2)
//     the two hessian planes become one count plane: three ATOMS per cell.  Sums are order-independent, so histograms
//     are deterministic and identical for any grid shape or GPU count;
//   * memory-level parallelism: a warp requests the 32 bin sectors of a batch of 32 rows back to back and the
//     gradient words alongside, with the next batch's row ids already in flight;
//   * one launch serves a whole wave: the rows of all nodes whose histogram is due are cut into units of kHistUnit
//     rows; every CTA takes a contiguous run of units, accumulates node by node in shared memory and flushes each
//     node's partial sums to that node's slot.
#include <cstdlib>

#include "kernels.cuh"

namespace gbl {

constexpr int kHistThreads = 1024;
constexpr int kHistWarps = kHistThreads / 32;
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

// wave entry owning histogram unit u (entries without units own nothing)
__device__ inline int hist_owner(const Expand* __restrict__ w, int n, int u) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (w[mid].hist_unit_begin > u) {
      hi = mid;
    } else {
      lo = mid + 1;
    }
  }
  return lo - 1;
}

// ------------------------------------------------------------------------------------------ quantized mode
// smem: int32 hist[256 bins][32 lanes] + per-warp staging of 32 rows (row ids, packed (grad << 16) + hess words).
// Per batch of 32 rows and warp: 8 LDS.128 (row ids, broadcast) + 32 LDG.U8 (32 bins = one sector each, all in
// flight together) + 8 LDS.128 (gradient words) + 32 ATOMS.ADD.
struct QuantAcc {
  static constexpr int kPlanes = 1;
  static constexpr int kStageBytes = 8;
  int* sh;
  int* sr;  // per-warp staging: row ids of the batch
  int* sv;  // per-warp staging: packed gradient words of the batch
  const uint8_t* rp;
  unsigned rs, col;
  short* gh_q;  // int8 [hess, grad] pair of every row
  const LearnerDev* dev;
  double inv_g, inv_h;
  bool convert;

  __device__ __forceinline__ void accumulate(const int* __restrict__ indices, int r0, int r1, int warp, int lane) const {
    int base = r0 + warp * 32;
    if (base >= r1) return;
    int r_next = base + lane < r1 ? (indices != nullptr ? indices[base + lane] : base + lane) : -1;
    for (; base < r1; base += kHistWarps * 32) {
      const int r = r_next;
      {  // row ids of the warp's next batch: in flight while this batch is accumulated
        const int nx = base + kHistWarps * 32 + lane;
        r_next = nx < r1 ? (indices != nullptr ? indices[nx] : nx) : -1;
      }
      sr[lane] = max(r, 0);  // padding rows add 0 to bin(row 0)
      __syncwarp();
      int q = 0;  // int16: high byte grad (signed), low byte hess
      if (r >= 0) {
        // The int8 pairs live in their own 2-byte-per-row array (L2-resident: 2 x num_data bytes), NOT in the row
        // records: the records stay read-only, so the root pass does not write a 32-byte sector back per row
        // (measured before: 777 MB of DRAM traffic for 315 MB algorithmic on the Higgs shape).
        if (convert) {  // root pass: GradientDiscretizer::DiscretizeGradients for this row, kept for the later waves
          const short packed = discretize_row(*dev, r, inv_g, inv_h);
          gh_q[r] = packed;
          q = packed;
        } else {
          q = __ldg(gh_q + r);
        }
      }
      // all 32 bin sectors of the batch are requested before the first one is consumed (memory-level parallelism:
      // 33 loads in flight per warp instead of 16 + a dependent gradient load in front of them)
      int bins[32];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int4 rr = reinterpret_cast<const int4*>(sr)[k];
        bins[4 * k + 0] = __ldg(rp + static_cast<size_t>(static_cast<unsigned>(rr.x)) * rs);
        bins[4 * k + 1] = __ldg(rp + static_cast<size_t>(static_cast<unsigned>(rr.y)) * rs);
        bins[4 * k + 2] = __ldg(rp + static_cast<size_t>(static_cast<unsigned>(rr.z)) * rs);
        bins[4 * k + 3] = __ldg(rp + static_cast<size_t>(static_cast<unsigned>(rr.w)) * rs);
      }
      sv[lane] = r >= 0 ? ((q >> 8) << 16) + (q & 0xff) : 0;  // packed (grad << 16) + hess
      __syncwarp();
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int4 vv = reinterpret_cast<const int4*>(sv)[k];
        red_shared_add(col + bins[4 * k + 0] * (kTile * 4), vv.x);
        red_shared_add(col + bins[4 * k + 1] * (kTile * 4), vv.y);
        red_shared_add(col + bins[4 * k + 2] * (kTile * 4), vv.z);
        red_shared_add(col + bins[4 * k + 3] * (kTile * 4), vv.w);
      }
    }
  }
  // flush + re-zero: widen (g16 | h16) to the reference's (g32 | h32) packing
  __device__ __forceinline__ void flush(long long* __restrict__ dst, int lanes_used) const {
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
  }
};

// ------------------------------------------------------------------------------------------ float mode
// smem: 4 planes of int32 [256][32]: g_lo, g_hi, h_lo, h_hi (16-bit limbs of the 2^k fixed-point value, so no
// carries are needed: lo limbs sum < 2^16 * 65536 rows, hi limbs < 2^15 * 65536 rows) + per-warp staging.
// Per batch of 32 rows and warp: 8 LDS.128 (row ids) + 32 LDG.U8 in flight together + 16 LDS.128 + 128 ATOMS.ADD.
constexpr int kFloatFlushCap = 65536;

// CONST_H: ObjectiveFunction::IsConstantHessian() (unweighted L2, regression_objective.hpp:152-158).  Every row has the
// same fixed-point hessian, so the hessian planes are replaced by ONE count plane (3 ATOMS per cell instead of 4) and
// the hessian sum of a bin is count * h_const at flush time -- the same integer the four-plane path accumulates.  The
// reference takes the analogous short cut (count-only ConstructHistogram, dense_bin.hpp:124-127).
template <bool CONST_H>
struct FloatAcc {
  static constexpr int kPlanes = 4;
  static constexpr int kStageBytes = 16;
  int* sh;
  int* sr;   // per-warp staging: row ids of the batch
  int2* sv;  // per-warp staging: fixed-point (grad, hess) of the batch
  const uint8_t* rp;
  unsigned rs, col;
  uint8_t* ghp;  // record base + gradient offset
  const float* grad;
  const float* hess;
  double g_scale, h_scale;
  long long h_const;  // CONST_H: the fixed-point hessian every row carries
  bool convert;

  __device__ __forceinline__ void accumulate(const int* __restrict__ indices, int r0, int r1, int warp, int lane) const {
    int base = r0 + warp * 32;
    if (base >= r1) return;
    int r_next = base + lane < r1 ? (indices != nullptr ? indices[base + lane] : base + lane) : -1;
    for (; base < r1; base += kHistWarps * 32) {
      const int r = r_next;
      {  // row ids of the warp's next batch: in flight while this batch is accumulated
        const int nx = base + kHistWarps * 32 + lane;
        r_next = nx < r1 ? (indices != nullptr ? indices[nx] : nx) : -1;
      }
      sr[lane] = max(r, 0);  // padding rows add 0 to bin(row 0)
      __syncwarp();
      int2 v = make_int2(0, 0);
      if (r >= 0) {
        uint8_t* rec_gh = ghp + static_cast<size_t>(static_cast<unsigned>(r)) * rs;
        if (convert) {  // root pass: float gradients -> fixed point, kept in the row record for the later waves
          v.x = __double2int_rn(static_cast<double>(grad[r]) * g_scale);
          v.y = __double2int_rn(static_cast<double>(hess[r]) * h_scale);
          *reinterpret_cast<int2*>(rec_gh) = v;
        } else {
          v = *reinterpret_cast<const int2*>(rec_gh);
        }
      }
      // all 32 bin sectors of the batch are requested before the first one is consumed (memory-level parallelism:
      // 33 loads in flight per warp instead of 8 + a dependent gradient load in front of them)
      int bins[32];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int4 rr = reinterpret_cast<const int4*>(sr)[k];
        bins[4 * k + 0] = __ldg(rp + static_cast<size_t>(static_cast<unsigned>(rr.x)) * rs);
        bins[4 * k + 1] = __ldg(rp + static_cast<size_t>(static_cast<unsigned>(rr.y)) * rs);
        bins[4 * k + 2] = __ldg(rp + static_cast<size_t>(static_cast<unsigned>(rr.z)) * rs);
        bins[4 * k + 3] = __ldg(rp + static_cast<size_t>(static_cast<unsigned>(rr.w)) * rs);
      }
      if (CONST_H) v.y = r >= 0 ? 1 : 0;  // the count plane's addend
      sv[lane] = v;
      __syncwarp();
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        const int4 vv = reinterpret_cast<const int4*>(sv)[k];  // two rows: {g, h, g, h}
        const unsigned a = col + bins[2 * k] * (kTile * 4);
        red_shared_add(a, vv.x & 0xffff);
        red_shared_add(a + kPlane * 4, vv.x >> 16);
        if (CONST_H) {
          red_shared_add(a + 2 * kPlane * 4, vv.y);
        } else {
          red_shared_add(a + 2 * kPlane * 4, vv.y & 0xffff);
          red_shared_add(a + 3 * kPlane * 4, vv.y >> 16);
        }
        const unsigned b = col + bins[2 * k + 1] * (kTile * 4);
        red_shared_add(b, vv.z & 0xffff);
        red_shared_add(b + kPlane * 4, vv.z >> 16);
        if (CONST_H) {
          red_shared_add(b + 2 * kPlane * 4, vv.w);
        } else {
          red_shared_add(b + 2 * kPlane * 4, vv.w & 0xffff);
          red_shared_add(b + 3 * kPlane * 4, vv.w >> 16);
        }
      }
    }
  }
  __device__ __forceinline__ void flush(long long* __restrict__ dst_words, int lanes_used) const {
    longlong2* __restrict__ dst = reinterpret_cast<longlong2*>(dst_words);
    for (int e = threadIdx.x; e < kPlane; e += kHistThreads) {
      const long long g = (static_cast<long long>(sh[e + kPlane]) << 16) + static_cast<long long>(static_cast<unsigned>(sh[e]));
      const long long h =
          CONST_H ? static_cast<long long>(static_cast<unsigned>(sh[e + 2 * kPlane])) * h_const
                  : (static_cast<long long>(sh[e + 3 * kPlane]) << 16) +
                        static_cast<long long>(static_cast<unsigned>(sh[e + 2 * kPlane]));
      if ((e & 31) < lanes_used) {
        if (g != 0) atomicAdd(reinterpret_cast<unsigned long long*>(&dst[e].x), static_cast<unsigned long long>(g));
        if (h != 0) atomicAdd(reinterpret_cast<unsigned long long*>(&dst[e].y), static_cast<unsigned long long>(h));
      }
      sh[e] = 0;
      sh[e + kPlane] = 0;
      sh[e + 2 * kPlane] = 0;
      if (!CONST_H) sh[e + 3 * kPlane] = 0;
    }
  }
};

// Shared driver.  Tap mode (explicit_count >= 0): the given row list, split evenly over the CTAs of the tile,
// into `explicit_out`.  Wave mode: this CTA's contiguous share of the wave's work units, node by node.
template <class Acc>
__device__ __forceinline__ void hist_drive(const LearnerDev& d, const Acc& acc, long long* __restrict__ explicit_out,
                                           const int* explicit_indices, int explicit_count, int cap, int entry_words) {
  const int tile = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int lanes_used = min(kTile, d.p.num_features - tile * kTile);
  const long long tile_off = static_cast<long long>(tile) * kPlane * entry_words;
  const int gx = gridDim.x, x = blockIdx.x;
  if (explicit_count >= 0) {
    int per = (explicit_count + gx - 1) / gx;
    per = (per + kHistUnit - 1) / kHistUnit * kHistUnit;
    const long long r0 = static_cast<long long>(x) * per;
    const int r1 = static_cast<int>(min(static_cast<long long>(explicit_count), r0 + per));
    for (int p = static_cast<int>(r0); p < r1; p += cap) {
      acc.accumulate(explicit_indices, p, min(r1, p + cap), warp, lane);
      __syncthreads();
      acc.flush(explicit_out + tile_off, lanes_used);
      __syncthreads();
    }
    return;
  }
  const TreeState& t = *d.ts;
  const int U = t.hist_units, C = t.num_cand;
  const int u0 = static_cast<int>(static_cast<long long>(U) * x / gx);
  const int u1 = static_cast<int>(static_cast<long long>(U) * (x + 1) / gx);
  if (u0 >= u1) return;
  int ci = hist_owner(d.wave, C, u0);
  int u = u0;
  while (u < u1) {
    const Expand w = d.wave[ci];
    const int next_begin = d.wave[ci + 1].hist_unit_begin;
    if (next_begin <= u) {  // entry without units
      ++ci;
      continue;
    }
    const int ue = min(u1, next_begin);
    const Node& nd = d.nodes[w.hist_node];
    const int* __restrict__ indices = d.idx[nd.buf] + nd.begin;
    const int r0 = (u - w.hist_unit_begin) * kHistUnit;
    const int r1 = min(nd.count, (ue - w.hist_unit_begin) * kHistUnit);
    long long* __restrict__ dst = d.hist_pool + static_cast<long long>(w.hist_slot) * d.slot_stride + tile_off;
    for (int p = r0; p < r1; p += cap) {
      acc.accumulate(indices, p, min(r1, p + cap), warp, lane);
      __syncthreads();
      acc.flush(dst, lanes_used);
      __syncthreads();
    }
    u = ue;
    ++ci;
  }
}

__global__ void __launch_bounds__(kHistThreads, 1) hist_quant_kernel(LearnerDev d, long long* __restrict__ out,
                                                                     const int* explicit_indices, int explicit_count,
                                                                     int flush_cap) {
  extern __shared__ int sh[];
  if (explicit_count < 0 && (d.ts->done || d.ts->hist_units == 0)) return;
  for (int e = threadIdx.x; e < kPlane; e += kHistThreads) sh[e] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  QuantAcc acc;
  acc.sh = sh;
  acc.sr = sh + kPlane + warp * 64;
  acc.sv = acc.sr + 32;
  acc.rp = opaque_ptr(d.rec + blockIdx.y * d.tile_stride + lane);
  acc.rs = static_cast<unsigned>(d.rec_stride);
  acc.col = opaque_u32(static_cast<unsigned>(__cvta_generic_to_shared(sh + lane)));
  acc.gh_q = d.gh_q;
  acc.dev = &d;
  acc.inv_g = d.ts->q_inv_grad_scale;
  acc.inv_h = d.ts->q_inv_hess_scale;
  acc.convert = explicit_count < 0 && d.ts->waves == 0;
  hist_drive(d, acc, out, explicit_indices, explicit_count, flush_cap, 1);
}

template <bool CONST_H>
__global__ void __launch_bounds__(kHistThreads, 1) hist_float_kernel(LearnerDev d, long long* __restrict__ out,
                                                                     const int* explicit_indices, int explicit_count) {
  extern __shared__ int sh[];
  if (explicit_count < 0 && (d.ts->done || d.ts->hist_units == 0)) return;
  for (int e = threadIdx.x; e < 4 * kPlane; e += kHistThreads) sh[e] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  FloatAcc<CONST_H> acc;
  acc.sh = sh;
  acc.sr = sh + 4 * kPlane + warp * 128;
  acc.sv = reinterpret_cast<int2*>(acc.sr + 32);
  acc.rp = opaque_ptr(d.rec + blockIdx.y * d.tile_stride + lane);
  acc.rs = static_cast<unsigned>(d.rec_stride);
  acc.col = opaque_u32(static_cast<unsigned>(__cvta_generic_to_shared(sh + lane)));
  acc.ghp = d.rec + blockIdx.y * d.tile_stride + d.gh_off;
  acc.grad = d.grad;
  acc.hess = d.hess;
  acc.g_scale = d.ts->g_scale;
  acc.h_scale = d.ts->h_scale;
  // every row's hessian is hess[0] (the caller's IsConstantHessian contract, serial_tree_learner.cpp:63,
  // dataset.cpp:1403-1406 multiplies the counts by hessians[0] the same way)
  acc.h_const = CONST_H ? static_cast<long long>(__double2int_rn(static_cast<double>(d.hess[0]) * acc.h_scale)) : 0;
  // the first pass of a tree also converts the gradients (to_fixed: exact power-of-two fixed point, |value| <= 2^30);
  // with several tiles every tile block of a record carries its own copy of the pair, written by that tile's CTAs
  acc.convert = explicit_count < 0 && d.ts->waves == 0;
  hist_drive(d, acc, out, explicit_indices, explicit_count, kFloatFlushCap, 2);
}

static const int kSmemQ = kPlane * 4 + kHistWarps * 32 * 8;        // 32 KB + 8 KB staging
static const int kSmemF = 4 * kPlane * 4 + kHistWarps * 32 * 16;   // 128 KB + 16 KB staging

void histogram_setup() {
  GBL_CUDA(cudaFuncSetAttribute(hist_float_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemF));
  GBL_CUDA(cudaFuncSetAttribute(hist_float_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemF));
  GBL_CUDA(cudaFuncSetAttribute(hist_quant_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemQ));
}

void launch_histogram(const LearnerDev& d, const int* explicit_indices, int explicit_count, long long* explicit_out,
                      cudaStream_t s) {
  const int tiles = d.p.num_tiles;
  // one CTA per SM at most, all tiles resident at once: a ceil here (e.g. 5 tiles x 30 = 150 CTAs on 148 SMs) leaves
  // two CTAs for a second round that doubles the launch time
  const int gx = std::max(1, kNumSM / tiles);
  if (d.p.quantized) {
    // 16-bit halves hold sums of at most flush_cap rows: |g| <= bins/2, h <= bins (gradient_discretizer.cpp:105-127)
    int flush_cap = (65535 / std::max(1, d.p.num_grad_quant_bins)) & ~255;
    flush_cap = std::max(256, std::min(flush_cap, 1 << 20));
    GBL_LAUNCH(hist_quant_kernel, dim3(gx, tiles), kHistThreads, kSmemQ, s, d, explicit_out, explicit_indices,
               explicit_count, flush_cap);
  } else if (d.p.is_constant_hessian) {
    GBL_LAUNCH(hist_float_kernel<true>, dim3(gx, tiles), kHistThreads, kSmemF, s, d, explicit_out, explicit_indices,
               explicit_count);
  } else {
    GBL_LAUNCH(hist_float_kernel<false>, dim3(gx, tiles), kHistThreads, kSmemF, s, d, explicit_out, explicit_indices,
               explicit_count);
  }
}

}  // namespace gbl
