// Device-wide exclusive prefix sum and LSD radix sort over W-word records (hand-written; no CUB/Thrust).
//
// Replaces qsort(elem, n_elem, 40 B, compare_hopo_element_decreasing) (reference src/hopo_counter.c:351, :28-38)
// and the two qsorts inside new_empfreq_from_int_weighted (empirical_frequency.c:202-228).  Records are sorted
// ASCENDING on the digit list given by the caller (least significant digit first); the reduce step reads the
// result back to front to obtain the reference's descending order.
#pragma once
#include "tjb_common.cuh"

// ------------------------------------------------------------------ exclusive scan (u32, wrap-around) ----
#define PSCAN_THREADS 256
#define PSCAN_ITEMS 8
#define PSCAN_TILE (PSCAN_THREADS * PSCAN_ITEMS)

__device__ __forceinline__ u32 warp_incl_scan (u32 v, int lane)
{
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { u32 t = __shfl_up_sync (0xffffffffu, v, d); if (lane >= d) v += t; }
  return v;
}

// block-wide exclusive scan of one value per thread; returns the exclusive prefix, *total gets the block sum
__device__ __forceinline__ u32 block_excl_scan (u32 v, u32 *s_warp /*[32]*/, u32 *total)
{
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  u32 inc = warp_incl_scan (v, lane);
  if (lane == 31) s_warp[wid] = inc;
  __syncthreads ();
  if (wid == 0) {
    u32 w = lane < nw ? s_warp[lane] : 0;
    u32 winc = warp_incl_scan (w, lane);
    s_warp[lane] = winc - w; // exclusive
    if (lane == 31) s_warp[32] = winc;
  }
  __syncthreads ();
  u32 r = s_warp[wid] + inc - v;
  *total = s_warp[32];
  __syncthreads ();
  return r;
}

// pass 1: per-tile sums
__global__ void __launch_bounds__ (PSCAN_THREADS) pscan_reduce_kernel (const u32 *__restrict__ in, u32 *__restrict__ sums, size_t n)
{
  __shared__ u32 s_warp[33];
  size_t base = (size_t) blockIdx.x * PSCAN_TILE + (size_t) threadIdx.x * PSCAN_ITEMS;
  u32 v = 0;
#pragma unroll
  for (int j = 0; j < PSCAN_ITEMS; j++) if (base + j < n) v += in[base + j];
  u32 total;
  block_excl_scan (v, s_warp, &total);
  if (threadIdx.x == 0) sums[blockIdx.x] = total;
}
// pass 2: scan inside tiles with the (already scanned) tile offsets; n <= PSCAN_TILE handled with offs == nullptr
__global__ void __launch_bounds__ (PSCAN_THREADS) pscan_scan_kernel (const u32 *__restrict__ in, u32 *__restrict__ out, const u32 *__restrict__ offs, size_t n, u32 *__restrict__ total_out)
{
  __shared__ u32 s_warp[33];
  size_t base = (size_t) blockIdx.x * PSCAN_TILE + (size_t) threadIdx.x * PSCAN_ITEMS;
  u32 x[PSCAN_ITEMS], v = 0;
#pragma unroll
  for (int j = 0; j < PSCAN_ITEMS; j++) { x[j] = (base + j < n) ? in[base + j] : 0; v += x[j]; }
  u32 total;
  u32 ex = block_excl_scan (v, s_warp, &total) + (offs ? offs[blockIdx.x] : 0);
#pragma unroll
  for (int j = 0; j < PSCAN_ITEMS; j++) { if (base + j < n) out[base + j] = ex; ex += x[j]; }
  if (total_out && blockIdx.x == gridDim.x - 1 && threadIdx.x == PSCAN_THREADS - 1) *total_out = ex; // grand total
}

struct Launcher; // defined in tjb_gpu.cu: counts launches / records per-kernel events

// Exclusive scan of n u32 values, in -> out (may alias).  tmp must hold pscan_tmp_elems(n) u32.
// total_out (device pointer, may be null) receives the sum of all n inputs.
inline size_t pscan_tmp_elems (size_t n)
{
  size_t t = 0;
  while (n > PSCAN_TILE) { n = (n + PSCAN_TILE - 1) / PSCAN_TILE; t += n; }
  return t + 1;
}

// ------------------------------------------------------------------ LSD radix sort -----------------------
#define RSORT_THREADS 512
#define RSORT_WARPS (RSORT_THREADS / 32)
#define RSORT_ITEMS 8
#define RSORT_TILE (RSORT_THREADS * RSORT_ITEMS)
#define RSORT_RADIX_BITS 8
#define RSORT_RADIX (1 << RSORT_RADIX_BITS)

// per-tile digit histogram, stored digit-major: hist[d * n_tiles + tile]
template <int W>
__global__ void __launch_bounds__ (RSORT_THREADS)
rsort_hist_kernel (const u64 *__restrict__ in, size_t n, DigitDesc dd, u32 *__restrict__ hist, u32 n_tiles)
{
  __shared__ u32 s_h[RSORT_RADIX];
  for (int i = threadIdx.x; i < RSORT_RADIX; i += RSORT_THREADS) s_h[i] = 0;
  __syncthreads ();
  const u32 mask = (1u << dd.nbits) - 1;
  size_t base = (size_t) blockIdx.x * RSORT_TILE;
#pragma unroll
  for (int r = 0; r < RSORT_ITEMS; r++) {
    size_t i = base + (size_t) r * RSORT_THREADS + threadIdx.x;
    if (i < n) {
      u32 d = (u32) (in[(size_t) W * i + dd.word] >> dd.shift) & mask;
      atomicAdd (&s_h[d], 1u);
    }
  }
  __syncthreads ();
  for (int i = threadIdx.x; i < RSORT_RADIX; i += RSORT_THREADS) hist[(size_t) i * n_tiles + blockIdx.x] = s_h[i];
}

// stable scatter: warp w of a tile owns RSORT_ITEMS*32 consecutive records; ranks among equal digits come from
// __match_any_sync + per-warp digit counters (no shared-memory atomics, order preserved)
template <int W>
__global__ void __launch_bounds__ (RSORT_THREADS)
rsort_scatter_kernel (const u64 *__restrict__ in, u64 *__restrict__ out, size_t n, DigitDesc dd,
                      const u32 *__restrict__ offs, u32 n_tiles)
{
  __shared__ u32 s_cnt[RSORT_WARPS][RSORT_RADIX];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < RSORT_WARPS * RSORT_RADIX; i += RSORT_THREADS) (&s_cnt[0][0])[i] = 0;
  __syncthreads ();
  const u32 mask = (1u << dd.nbits) - 1;
  const size_t wbase = (size_t) blockIdx.x * RSORT_TILE + (size_t) wid * (RSORT_ITEMS * 32);
  Rec<W> rec[RSORT_ITEMS];
  u32 dig[RSORT_ITEMS], rank[RSORT_ITEMS];
#pragma unroll
  for (int r = 0; r < RSORT_ITEMS; r++) {
    size_t i = wbase + (size_t) r * 32 + lane;
    u32 d = RSORT_RADIX; // sentinel for lanes past the end
    if (i < n) { rec[r] = load_rec<W> (in, i); d = (u32) (rec[r].w[dd.word] >> dd.shift) & mask; }
    dig[r] = d;
    u32 peers = __match_any_sync (0xffffffffu, d);
    int leader = __ffs (peers) - 1;
    u32 before = __popc (peers & ((1u << lane) - 1));
    u32 old = 0;
    if (lane == leader && d < RSORT_RADIX) { old = s_cnt[wid][d]; s_cnt[wid][d] = old + __popc (peers); }
    old = __shfl_sync (0xffffffffu, old, leader);
    rank[r] = old + before;
    __syncwarp ();
  }
  __syncthreads ();
  // turn per-warp counts into global positions: offs[d][tile] + counts of earlier warps
  for (int d = threadIdx.x; d < RSORT_RADIX; d += RSORT_THREADS) {
    u32 run = offs[(size_t) d * n_tiles + blockIdx.x];
#pragma unroll
    for (int w = 0; w < RSORT_WARPS; w++) { u32 c = s_cnt[w][d]; s_cnt[w][d] = run; run += c; }
  }
  __syncthreads ();
#pragma unroll
  for (int r = 0; r < RSORT_ITEMS; r++) {
    if (dig[r] < RSORT_RADIX) store_rec<W> (out, (size_t) s_cnt[wid][dig[r]] + rank[r], rec[r]);
  }
}
