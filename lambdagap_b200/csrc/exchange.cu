// exchange.cu -- the data-parallel learner's histogram all-reduce as kernels over NVLink peer memory.
//
// Reference: DataParallelTreeLearner::FindBestSplits -> Network::ReduceScatter of the smaller-leaf histograms
// (src/treelearner/data_parallel_tree_learner.cpp:285-297, src/network/network.cpp:227-323: host sockets / MPI,
// recursive halving over serialized byte blocks) and GlobalSyncUpByMax (gradient_discretizer.cpp:101-104).
//
// B200 design.  Every rank maps every other rank's histogram pool (CUDA IPC; one process per GPU).  After a wave's
// local histograms are built, the wave's fresh slots -- one contiguous range of the pool, the same on every rank --
// are summed in place in two steps inside ONE kernel per rank: rank k owns the k-th share of the range, loads that
// share from all R pools over NVLink (16-byte loads), adds the exact integers, and stores the sum back into all R
// pools (reduce-scatter and all-gather of a two-shot all-reduce, with no staging buffer and no atomics).  Ranks meet
// at a flag barrier before (all local histograms complete) and after (all sums stored).  The range is read from the
// device-resident tree state, so nothing has to come back to the host to size a collective: the whole tree -- waves,
// exchanges and barriers -- is one CUDA graph per rank, as on a single GPU.  The NCCL path it replaces needed the host
// to read the tree state and enqueue one ncclAllReduce per wave.
//
// Barrier: rank a stores the barrier's sequence number into slot a of every rank's flag array (st.release.sys) and
// waits until all R slots of its own array have reached that number (ld.acquire.sys).  Every wait is bounded
// (60 s of SM clocks by default, GBL_PEER_TIMEOUT_S: ranks may reach the first barrier of a tree seconds apart when
// their hosts do different amounts of work between iterations): a rank that died makes the others fail with an error
// instead of spinning forever.
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

// max |gradient|, max |hessian| over ranks: every rank deposits its pair with every rank, then takes the maximum
__global__ void __launch_bounds__(32) xchg_max_push_kernel(LearnerDev d) {
  const PeerX& x = d.px;
  if (threadIdx.x < d.p.num_ranks) {
    float* dst = x.maxbuf[threadIdx.x] + 2 * x.rank;
    dst[0] = d.ts->max_abs_g;
    dst[1] = d.ts->max_abs_h;
  }
}

__global__ void __launch_bounds__(32) xchg_max_apply_kernel(LearnerDev d) {
  const PeerX& x = d.px;
  if (threadIdx.x == 0) {
    const float* mine = x.maxbuf[x.rank];
    float g = 0.f, h = 0.f;
    for (int r = 0; r < d.p.num_ranks; ++r) {
      g = fmaxf(g, mine[2 * r]);
      h = fmaxf(h, mine[2 * r + 1]);
    }
    d.ts->max_abs_g = g;
    d.ts->max_abs_h = h;
  }
}

void launch_xchg_barrier(const LearnerDev& d, cudaStream_t s) { GBL_LAUNCH(xchg_barrier_kernel, 1, 32, 0, s, d); }

void launch_xchg_reduce(const LearnerDev& d, int is_root, cudaStream_t s) {
  GBL_LAUNCH(xchg_reduce_kernel, kNumSM, 256, 0, s, d, is_root);
}

void launch_xchg_max(const LearnerDev& d, cudaStream_t s) {
  GBL_LAUNCH(xchg_max_push_kernel, 1, 32, 0, s, d);
  launch_xchg_barrier(d, s);
  GBL_LAUNCH(xchg_max_apply_kernel, 1, 32, 0, s, d);
}

}  // namespace gbl
