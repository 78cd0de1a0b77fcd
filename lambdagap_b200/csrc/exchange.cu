// exchange.cu -- the data-parallel learner's histogram all-reduce as kernels over NVLink peer memory.
//
// Reference: DataParallelTreeLearner::FindBestSplits -> Network::ReduceScatter of the smaller-leaf histograms
// (src/treelearner/data_parallel_tree_learner.cpp:285-297, src/network/network.cpp:227-323: host sockets / MPI,
// recursive halving over serialized byte blocks) and GlobalSyncUpByMax (gradient_discretizer.cpp:101-104).
//
// B200 design.  Every rank maps every other rank's histogram pool (CUDA IPC; one process per GPU).  After a wave's
// local histograms are built, the wave's fresh slots -- one contiguous range of the pool, the same on every rank --
// are summed in place inside ONE kernel per rank and wave: rank k owns the k-th share of the range, loads that
// share from all R pools over NVLink (16-byte loads), adds the exact integers, and stores the sum back into all R
// pools (reduce-scatter and all-gather of a two-shot all-reduce, with no staging buffer and no atomics).  The same kernel
// meets the other ranks at a flag barrier before (all local histograms complete) and after (all sums stored).  The range is read from the
// device-resident tree state, so nothing has to come back to the host to size a collective: the whole tree -- waves,
// exchanges and barriers -- is one CUDA graph per rank, as on a single GPU.  The NCCL path it replaces needed the host
// to read the tree state and enqueue one ncclAllReduce per wave.
//
// Barrier: rank a stores the barrier's sequence number into slot a of every rank's flag array (st.release.sys) and
// waits until all R slots of its own array have reached that number (ld.acquire.sys).  Every wait is bounded
// (60 s of SM clocks by default, GBL_PEER_TIMEOUT_S: ranks may reach the first barrier of a tree seconds apart when
// their hosts do different amounts of work between iterations): a rank that died makes the others fail with an error
// instead of spinning forever.
#include <cstdlib>

#include "kernels.cuh"

namespace gbl {

__device__ inline unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ inline void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ---- three-launch form (barrier / reduce / barrier): the default, see launch_xchg_wave
// one warp; lane r talks to rank r
__global__ void __launch_bounds__(32) xchg_barrier_kernel(LearnerDev d) {
  const PeerX& x = d.px;
  const int R = d.p.num_ranks;
  unsigned e = 0;
  if (threadIdx.x == 0) {
    e = *x.epoch + 1;
    *x.epoch = e;
  }
  e = __shfl_sync(0xffffffffu, e, 0);
  if (*reinterpret_cast<volatile unsigned*>(x.dead) != 0u) {  // sticky: every later tree reports the failure too
    if (threadIdx.x == 0) d.ts->error = 9;
    return;
  }
  if (threadIdx.x < R) {
    __threadfence_system();
    st_release_sys(x.flags[threadIdx.x] + x.rank, e);
    const unsigned* mine = x.flags[x.rank] + threadIdx.x;
    const long long t0 = clock64();
    for (;;) {
      const unsigned v = ld_acquire_sys(mine);
      if (static_cast<int>(v - e) >= 0) break;
      if (clock64() - t0 > x.timeout_cycles) {
        *x.dead = 1u;
        d.ts->error = 9;
        break;
      }
    }
    __threadfence_system();
  }
}

// in-place sum over ranks of the current wave's fresh slots (is_root: slot 0)
__global__ void __launch_bounds__(256) xchg_reduce_kernel(LearnerDev d, int is_root) {
  const TreeState& t = *d.ts;
  if (!is_root && (t.done || t.last_wave)) return;
  const PeerX& x = d.px;
  if (*reinterpret_cast<volatile unsigned*>(x.dead) != 0u) return;
  const int R = d.p.num_ranks, me = x.rank;
  const long long slot0 = is_root ? 0 : d.wave[0].hist_slot;
  const long long C = is_root ? 1 : t.num_cand;
  const long long units = C * (d.slot_stride >> 1);  // 16-byte units (slot_stride is even)
  const long long chunk = (units + R - 1) / R;
  const long long u0 = me * chunk, u1 = min(units, u0 + chunk);
  const long long base = slot0 * (d.slot_stride >> 1);
  const long long gsize = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long u = u0 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; u < u1; u += gsize) {
    longlong2 acc = make_longlong2(0, 0);
    for (int r0 = 0; r0 < R; r0 += 8) {
      longlong2 v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {  // up to 8 loads in flight
        v[k] = make_longlong2(0, 0);
        if (r0 + k < R) v[k] = reinterpret_cast<const longlong2*>(x.pool[r0 + k])[base + u];
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        acc.x += v[k].x;
        acc.y += v[k].y;
      }
    }
    for (int r = 0; r < R; ++r) reinterpret_cast<longlong2*>(x.pool[r])[base + u] = acc;
  }
}

// Arrive at barrier number `e` (lane r tells rank r) and wait until every rank has arrived.  Called by the first R
// lanes of one warp.  Returns false when the wait timed out (the learner then reports error 9 instead of hanging).
__device__ inline bool peer_barrier(const LearnerDev& d, unsigned e, int lane, bool arrive) {
  const PeerX& x = d.px;
  if (arrive) {
    __threadfence_system();
    st_release_sys(x.flags[lane] + x.rank, e);
  }
  const unsigned* mine = x.flags[x.rank] + lane;
  const long long t0 = clock64();
  bool ok = true;
  for (;;) {
    const unsigned v = ld_acquire_sys(mine);
    if (static_cast<int>(v - e) >= 0) break;
    if (clock64() - t0 > x.timeout_cycles) {
      *x.dead = 1u;
      d.ts->error = 9;
      ok = false;
      break;
    }
  }
  __threadfence_system();
  return ok;
}

// The exchange of one wave as ONE kernel per rank (it used to be barrier / reduce / barrier: three launches):
//   1. CTA 0 announces "my local histograms are complete" to every rank; every CTA waits until all ranks have announced
//      (CTAs only ever wait for OTHER GPUs, never for each other, so no co-residency is assumed);
//   2. the in-place sum over ranks of this rank's share of the wave's fresh slots (is_root: slot 0): 16-byte loads from
//      all R pools over NVLink, exact integer adds, the sum stored back into all R pools;
//   3. the CTA that finishes last announces "my share is stored everywhere" and waits for every rank's announcement --
//      the kernel (and with it the stream) does not complete before all sums are in this rank's pool.
// Barrier numbers: the epoch word holds the last completed number; this exchange uses epoch + 1 and epoch + 2.
__global__ void __launch_bounds__(256) xchg_wave_kernel(LearnerDev d, int is_root) {
  const TreeState& t = *d.ts;
  if (!is_root && (t.done || t.last_wave)) return;
  const PeerX& x = d.px;
  if (*reinterpret_cast<volatile unsigned*>(x.dead) != 0u) {  // sticky: every later tree reports the failure too
    if (blockIdx.x == 0 && threadIdx.x == 0) d.ts->error = 9;
    return;
  }
  const int R = d.p.num_ranks, me = x.rank;
  const unsigned e0 = *reinterpret_cast<volatile unsigned*>(x.epoch);
  __shared__ int s_flag;
  if (blockIdx.x == 0) {
    // only this CTA talks to the other GPUs (148 CTAs polling system-scope flags slowed the exchange down: 4.45 ms per
    // tree against 4.25 ms for three separate launches); the rest wait for its word in local memory
    if (threadIdx.x < 32) {
      bool ok = true;
      if (threadIdx.x < R) ok = peer_barrier(d, e0 + 1u, threadIdx.x, true);
      ok = __all_sync(0xffffffffu, ok);
      if (threadIdx.x == 0) {
        s_flag = ok ? 1 : 0;
        __threadfence();
        asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(x.go), "r"(ok ? e0 + 1u : 0xffffffffu) : "memory");
      }
    }
  } else if (threadIdx.x == 0) {
    const long long t0 = clock64();
    unsigned v;
    for (;;) {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(x.go) : "memory");
      if (v == 0xffffffffu || static_cast<int>(v - (e0 + 1u)) >= 0) break;
      if (clock64() - t0 > x.timeout_cycles) {
        v = 0xffffffffu;
        break;
      }
      __nanosleep(20);
    }
    s_flag = v == 0xffffffffu ? 0 : 1;
  }
  __syncthreads();
  if (s_flag) {
    const long long slot0 = is_root ? 0 : d.wave[0].hist_slot;
    const long long C = is_root ? 1 : t.num_cand;
    const long long units = C * (d.slot_stride >> 1);  // 16-byte units (slot_stride is even)
    const long long chunk = (units + R - 1) / R;
    const long long u0 = me * chunk, u1 = min(units, u0 + chunk);
    const long long base = slot0 * (d.slot_stride >> 1);
    const long long gsize = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long u = u0 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; u < u1; u += gsize) {
      longlong2 acc = make_longlong2(0, 0);
      for (int r0 = 0; r0 < R; r0 += 8) {
        longlong2 v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {  // up to 8 loads in flight
          v[k] = make_longlong2(0, 0);
          if (r0 + k < R) v[k] = reinterpret_cast<const longlong2*>(x.pool[r0 + k])[base + u];
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          acc.x += v[k].x;
          acc.y += v[k].y;
        }
      }
      for (int r = 0; r < R; ++r) reinterpret_cast<longlong2*>(x.pool[r])[base + u] = acc;
    }
  }
  // ---- completion: the last CTA of this rank tells every rank and waits for theirs
  // every thread orders its stores into the peers' pools before the CTA's count at GPU scope; the one thread that sees the
  // last count then fences at system scope before it raises the flag the peers wait for (release is cumulative over the
  // counts it has observed), instead of 37 000 threads fencing at system scope each
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(x.done, 1u);
    __threadfence_system();
    s_flag = prev == gridDim.x - 1u ? 1 : 0;
  }
  __syncthreads();
  if (!s_flag) return;
  if (threadIdx.x < 32) {
    if (threadIdx.x < R) peer_barrier(d, e0 + 2u, threadIdx.x, true);
    __syncwarp();
    if (threadIdx.x == 0) {
      *x.epoch = e0 + 2u;
      *x.done = 0u;
      if (*x.go == 0xffffffffu) *x.go = 0u;  // a failed exchange must not poison the next tree's first wait
    }
  }
}

// Two-launch form: the barrier kernel in front (all local histograms complete), then the reduction with the SECOND barrier
// in its tail -- the CTA that finishes last raises this rank's flag and waits for the others.  Nothing is parked in front of
// a barrier (every other CTA has retired by then); one launch per wave less than the three-launch form.  GBL_XCHG_FUSED=2;
// measured on two B200s (Higgs shape, weak scaling): 4.31-4.32 ms per tree against 4.25-4.27 ms for three launches --
// inside a CUDA graph a 32-thread kernel between two others costs less than the fence + count every CTA pays here.
__global__ void __launch_bounds__(256) xchg_reduce_tail_kernel(LearnerDev d, int is_root) {
  const TreeState& t = *d.ts;
  if (!is_root && (t.done || t.last_wave)) return;
  const PeerX& x = d.px;
  if (*reinterpret_cast<volatile unsigned*>(x.dead) != 0u) return;
  const int R = d.p.num_ranks, me = x.rank;
  const long long slot0 = is_root ? 0 : d.wave[0].hist_slot;
  const long long C = is_root ? 1 : t.num_cand;
  const long long units = C * (d.slot_stride >> 1);  // 16-byte units (slot_stride is even)
  const long long chunk = (units + R - 1) / R;
  const long long u0 = me * chunk, u1 = min(units, u0 + chunk);
  const long long base = slot0 * (d.slot_stride >> 1);
  const long long gsize = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long u = u0 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; u < u1; u += gsize) {
    longlong2 acc = make_longlong2(0, 0);
    for (int r0 = 0; r0 < R; r0 += 8) {
      longlong2 v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {  // up to 8 loads in flight
        v[k] = make_longlong2(0, 0);
        if (r0 + k < R) v[k] = reinterpret_cast<const longlong2*>(x.pool[r0 + k])[base + u];
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        acc.x += v[k].x;
        acc.y += v[k].y;
      }
    }
    for (int r = 0; r < R; ++r) reinterpret_cast<longlong2*>(x.pool[r])[base + u] = acc;
  }
  // completion: GPU-scope fence per thread, the CTA's count, and the last CTA meets the other ranks
  __shared__ int s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(x.done, 1u);
    __threadfence_system();
    s_last = prev == gridDim.x - 1u ? 1 : 0;
  }
  __syncthreads();
  if (!s_last || threadIdx.x >= 32) return;
  const unsigned e = *reinterpret_cast<volatile unsigned*>(x.epoch) + 1u;
  if (threadIdx.x < R) peer_barrier(d, e, threadIdx.x, true);
  __syncwarp();
  if (threadIdx.x == 0) {
    *x.epoch = e;
    *x.done = 0u;
  }
}

// max |gradient|, max |hessian| over ranks in one kernel: every rank deposits its pair with every rank, meets the others
// at a barrier and takes the maximum of the R deposits (GradientDiscretizer's GlobalSyncUpByMax, also the float scale)
__global__ void __launch_bounds__(32) xchg_max_kernel(LearnerDev d) {
  const PeerX& x = d.px;
  const int R = d.p.num_ranks;
  if (*reinterpret_cast<volatile unsigned*>(x.dead) != 0u) {
    if (threadIdx.x == 0) d.ts->error = 9;
    return;
  }
  const unsigned e0 = *reinterpret_cast<volatile unsigned*>(x.epoch);
  if (threadIdx.x < R) {
    float* dst = x.maxbuf[threadIdx.x] + 2 * x.rank;
    dst[0] = d.ts->max_abs_g;
    dst[1] = d.ts->max_abs_h;
    peer_barrier(d, e0 + 1u, threadIdx.x, true);
  }
  __syncwarp();
  if (threadIdx.x == 0) {
    const volatile float* mine = x.maxbuf[x.rank];
    float g = 0.f, h = 0.f;
    for (int r = 0; r < R; ++r) {
      g = fmaxf(g, mine[2 * r]);
      h = fmaxf(h, mine[2 * r + 1]);
    }
    d.ts->max_abs_g = g;
    d.ts->max_abs_h = h;
  }
  __syncwarp();
  // second barrier: nobody may overwrite a deposit (next tree) before every rank has read its R values
  if (threadIdx.x < R) peer_barrier(d, e0 + 2u, threadIdx.x, true);
  __syncwarp();
  if (threadIdx.x == 0) *x.epoch = e0 + 2u;
}

// Default: three launches.  The one-kernel form (GBL_XCHG_FUSED=1) was measured slower on two B200s, Higgs shape,
// weak scaling: 4.38 ms per tree against 4.21 ms (4.45 ms when every CTA polled the peers' flags itself); with GPU-scope
// fences per thread and one system-scope fence before the flag, 4.32 against 4.26 ms -- the launches it saves still cost
// less than the 148 CTAs it parks in front of the barrier.
static int xchg_mode() {
  static const int mode = [] {
    const char* e = std::getenv("GBL_XCHG_FUSED");
    return e == nullptr ? 0 : std::atoi(e);
  }();
  return mode;
}
bool xchg_fused_enabled() { return xchg_mode() == 1; }
int xchg_launches_per_wave() { return xchg_mode() == 1 ? 1 : (xchg_mode() == 2 ? 2 : 3); }

void launch_xchg_wave(const LearnerDev& d, int is_root, cudaStream_t s) {
  if (xchg_mode() == 1) {
    GBL_LAUNCH(xchg_wave_kernel, kNumSM, 256, 0, s, d, is_root);
  } else if (xchg_mode() == 2) {
    GBL_LAUNCH(xchg_barrier_kernel, 1, 32, 0, s, d);
    GBL_LAUNCH(xchg_reduce_tail_kernel, kNumSM, 256, 0, s, d, is_root);
  } else {
    GBL_LAUNCH(xchg_barrier_kernel, 1, 32, 0, s, d);
    GBL_LAUNCH(xchg_reduce_kernel, kNumSM, 256, 0, s, d, is_root);
    GBL_LAUNCH(xchg_barrier_kernel, 1, 32, 0, s, d);
  }
}

void launch_xchg_max(const LearnerDev& d, cudaStream_t s) { GBL_LAUNCH(xchg_max_kernel, 1, 32, 0, s, d); }

}  // namespace gbl
