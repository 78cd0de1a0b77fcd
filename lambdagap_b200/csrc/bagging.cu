// bagging.cu -- the reference's row-level bagging drawn on the device.
//
// Reference: BaggingSampleStrategy::Bagging / BaggingHelper (src/boosting/bagging.hpp:31-139, :224-243): every row
// draws NextFloat() from the generator of its block of 1024 rows (bagging_rand_block_; seeded bagging_seed + block,
// state carried from iteration to iteration; LightGBM's LCG, include/LightGBM/utils/random.h:20-120) and is in the bag
// when the draw is below bagging_fraction; in-bag rows are handed to the learner in ascending order
// (ParallelPartitionRunner::Run<true>).  The block-wise generators make the bag independent of the thread count, so
// it can be drawn by one device thread per block with the same integers: the bag is bit-identical to the
// reference's.  The host version of the same loop costs ~10 ns per row per re-bagging plus an index upload, which
// at bagging_freq=1 is far more than a tree takes on the device.
//
// Three launches: draw (one thread per block of 1024 rows: LCG steps -> 1 bit per row + the block's count),
// scan (one CTA: exclusive prefix of the block counts, totals), write (one warp per block: in-bag and out-of-bag
// row lists, both ascending).
#include "kernels.cuh"

namespace gbl {

constexpr int kBagBlock = 1024;  // bagging_rand_block_ (bagging.hpp:22)

__global__ void __launch_bounds__(128) bag_seed_kernel(unsigned* __restrict__ state, int nblk, int seed) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < nblk) state[b] = static_cast<unsigned>(seed + b);  // bagging_rands_.emplace_back(bagging_seed + i)
}

// label == nullptr: BaggingHelper (every row against `fraction`); otherwise BalancedBaggingHelper (bagging.hpp:245-270):
// rows with label > 0 against pos_fraction, the others against neg_fraction -- one draw per row either way
__global__ void __launch_bounds__(128) bag_draw_kernel(unsigned* __restrict__ state, int nblk, int n, double fraction,
                                                       const float* __restrict__ label, double pos_fraction,
                                                       double neg_fraction, unsigned* __restrict__ bits,
                                                       int* __restrict__ blk_cnt) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nblk) return;
  unsigned x = state[b];
  const int begin = b * kBagBlock;
  int cnt = 0;
  for (int w = 0; w < kBagBlock / 32; ++w) {
    unsigned word = 0u;
    for (int k = 0; k < 32; ++k) {
      const int row = begin + w * 32 + k;
      if (row < n) {
        x = 214013u * x + 2531011u;  // Random::RandInt16
        const float f = static_cast<float>(static_cast<int>((x >> 16) & 0x7FFF)) / 32768.0f;  // NextFloat
        const double limit = label == nullptr ? fraction : (label[row] > 0 ? pos_fraction : neg_fraction);
        if (static_cast<double>(f) < limit) word |= 1u << k;  // bagging.hpp:233 / :255-262
      }
    }
    bits[static_cast<size_t>(b) * (kBagBlock / 32) + w] = word;
    cnt += __popc(word);
  }
  state[b] = x;
  blk_cnt[b] = cnt;
}

// one CTA: blk_off[b] = in-bag rows before block b; counts[0] = bag size, counts[1] = rows outside the bag
__global__ void __launch_bounds__(1024) bag_scan_kernel(const int* __restrict__ blk_cnt, int nblk, int n,
                                                        int* __restrict__ blk_off, int* __restrict__ counts) {
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < nblk; base += 1024) {
    const int b = base + threadIdx.x;
    const int v = b < nblk ? blk_cnt[b] : 0;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int t = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, t, o);
        if (lane >= o) t += y;
      }
      s_warp[lane] = t;  // inclusive over warps
    }
    __syncthreads();
    const int before = s_carry + (warp > 0 ? s_warp[warp - 1] : 0) + x - v;
    if (b < nblk) blk_off[b] = before;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    counts[0] = s_carry;
    counts[1] = n - s_carry;
  }
}

// one warp per block of 1024 rows, word by word: lane k owns row begin + 32 w + k
__global__ void __launch_bounds__(256) bag_write_kernel(const unsigned* __restrict__ bits, const int* __restrict__ blk_off,
                                                        int nblk, int n, int* __restrict__ bag, int* __restrict__ oob) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= nblk) return;
  const int begin = b * kBagBlock;
  int in_base = blk_off[b];
  int out_base = begin - in_base;
  const unsigned my_word = bits[static_cast<size_t>(b) * (kBagBlock / 32) + lane];
  const unsigned below = (1u << lane) - 1u;
  for (int w = 0; w < kBagBlock / 32; ++w) {
    const unsigned word = __shfl_sync(0xffffffffu, my_word, w);
    const int row = begin + w * 32 + lane;
    const int pos = __popc(word & below);
    if (row < n) {
      if ((word >> lane) & 1u) {
        bag[in_base + pos] = row;
      } else {
        oob[out_base + lane - pos] = row;
      }
    }
    const int c = __popc(word);
    in_base += c;
    out_base += 32 - c;  // rows past n come last in the last block: their slots are never written
  }
}

// bagging_by_query (bagging.hpp:54-99): the draw runs over QUERIES (qbits: 1 bit per query); a row is in the bag when its
// query is.  One warp per block of 1024 rows builds the row-level bits and the block's count from row_query[].
__global__ void __launch_bounds__(256) bag_rows_from_queries_kernel(const unsigned* __restrict__ qbits,
                                                                    const int* __restrict__ row_query, int nblk, int n,
                                                                    unsigned* __restrict__ bits, int* __restrict__ blk_cnt) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= nblk) return;
  const int begin = b * kBagBlock + lane * 32;
  unsigned word = 0u;
  for (int k = 0; k < 32; ++k) {
    const int row = begin + k;
    if (row < n) {
      const int q = row_query[row];
      word |= ((qbits[q >> 5] >> (q & 31)) & 1u) << k;
    }
  }
  bits[static_cast<size_t>(b) * (kBagBlock / 32) + lane] = word;
  int cnt = __popc(word);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_down_sync(0xffffffffu, cnt, o);
  if (lane == 0) blk_cnt[b] = cnt;
}

void launch_bag_draw_by_query(unsigned* state, int nblk_q, int nq, double fraction, unsigned* qbits, int* qblk_cnt,
                              const int* row_query, int nblk, int n, unsigned* bits, int* blk_cnt, int* blk_off, int* counts,
                              int* bag, int* oob, cudaStream_t s) {
  GBL_LAUNCH(bag_draw_kernel, (nblk_q + 127) / 128, 128, 0, s, state, nblk_q, nq, fraction, static_cast<const float*>(nullptr),
             1.0, 1.0, qbits, qblk_cnt);
  GBL_LAUNCH(bag_rows_from_queries_kernel, (nblk + 7) / 8, 256, 0, s, qbits, row_query, nblk, n, bits, blk_cnt);
  GBL_LAUNCH(bag_scan_kernel, 1, 1024, 0, s, blk_cnt, nblk, n, blk_off, counts);
  GBL_LAUNCH(bag_write_kernel, (nblk + 7) / 8, 256, 0, s, bits, blk_off, nblk, n, bag, oob);
}

void launch_bag_seed(unsigned* state, int nblk, int seed, cudaStream_t s) {
  GBL_LAUNCH(bag_seed_kernel, (nblk + 127) / 128, 128, 0, s, state, nblk, seed);
}

void launch_bag_draw(unsigned* state, int nblk, int n, double fraction, const float* label, double pos_fraction,
                     double neg_fraction, unsigned* bits, int* blk_cnt, int* blk_off, int* counts, int* bag, int* oob,
                     cudaStream_t s) {
  GBL_LAUNCH(bag_draw_kernel, (nblk + 127) / 128, 128, 0, s, state, nblk, n, fraction, label, pos_fraction, neg_fraction, bits,
             blk_cnt);
  GBL_LAUNCH(bag_scan_kernel, 1, 1024, 0, s, blk_cnt, nblk, n, blk_off, counts);
  GBL_LAUNCH(bag_write_kernel, (nblk + 7) / 8, 256, 0, s, bits, blk_off, nblk, n, bag, oob);
}

}  // namespace gbl
