// Segmented run-length reduction of sorted tract observations into per-context tract-length histograms.
//
// Global multi-pass fallback for ONE bucket that does not fit in shared memory (tjb_gpu.cu: reduce_oversize_bucket):
// steps 1-4 of finalise_hopo_counter() (reference src/hopo_counter.c:352-413), sequential loops over the sorted
// element array in the reference.
// Input: records sorted ASCENDING on (base, ctx0, ctx1, length, flag).  All outputs are written back to front,
// i.e. in the reference's DESCENDING order.
#pragma once
#include "tjb_common.cuh"
#include "tjb_bucket.cuh" // QC_* modes

#define RED_THREADS 256

// key (first W words) of row i of an array of S-word rows
template <int W, int S> __device__ __forceinline__ Rec<W> load_key (const u64 *__restrict__ a, size_t i)
{
  Rec<W> r;
#pragma unroll
  for (int j = 0; j < W; j++) r.w[j] = a[(size_t) S * i + j];
  return r;
}

// head[i] = 1 when row i starts a new segment.  MODE 0: (base,ctx0,ctx1,length) changes; MODE 1: (base,ctx0,ctx1)
// changes.  S = words per row (W key words, then the count word of a weighted record when S = W + 1)
template <int W, int MODE, int S>
__global__ void __launch_bounds__ (RED_THREADS) head_kernel (const u64 *__restrict__ a, u32 n, u32 *__restrict__ head)
{
  u32 i = blockIdx.x * RED_THREADS + threadIdx.x;
  if (i >= n) return;
  u32 h = 1;
  if (i) {
    Rec<W> x = load_key<W, S> (a, i), y = load_key<W, S> (a, i - 1);
    if (MODE == 0) h = !same_tract<W> (x, y);
    else h = !same_context<W> (x, y);
  }
  head[i] = h;
}

// weighted records of one bucket (keys and count words in separate arrays) -> rows of W + 1 words for the global sort
template <int W>
__global__ void __launch_bounds__ (RED_THREADS)
wgather_kernel (const u64 *__restrict__ wkey, const unsigned long long *__restrict__ wcnt, u32 n, u64 *__restrict__ rows)
{
  u32 i = blockIdx.x * RED_THREADS + threadIdx.x;
  if (i >= n) return;
  Rec<W> r = load_rec<W> (wkey, i);
#pragma unroll
  for (int j = 0; j < W; j++) rows[(size_t) (W + 1) * i + j] = r.w[j];
  rows[(size_t) (W + 1) * i + W] = wcnt[i];
}

// seg_start[pos[i]] = i for every head
__global__ void __launch_bounds__ (RED_THREADS) seg_start_kernel (const u32 *__restrict__ head, const u32 *__restrict__ pos, u32 n, u32 *__restrict__ seg_start)
{
  u32 i = blockIdx.x * RED_THREADS + threadIdx.x;
  if (i < n && head[i]) seg_start[pos[i]] = i;
}

// one thread per distinct (base,ctx0,ctx1,length): the count words of its run of equal keys added up (the run-length
// count and canon_flag OR of hopo_counter.c:361-362 happened in the aggregation table; a key can still arrive in
// several rows), QC predicate (:367-374); output index reversed -> descending order
template <int W>
__global__ void __launch_bounds__ (RED_THREADS)
wrle_emit_kernel (const u64 *__restrict__ sorted, u32 n, const u32 *__restrict__ seg_start, u32 n_dist, int qc_mode,
                  u64 *__restrict__ d_rec, int *__restrict__ d_count, u32 *__restrict__ keep)
{
  u32 d = blockIdx.x * RED_THREADS + threadIdx.x;
  if (d >= n_dist) return;
  u32 s = seg_start[d], e = (d + 1 < n_dist) ? seg_start[d + 1] : n;
  Rec<W> first = load_key<W, W + 1> (sorted, s);
  u32 cf = 0, cr = 0;
  for (u32 i = s; i < e; i++) { const u64 c = sorted[(size_t) (W + 1) * i + W]; cf += (u32) c; cr += (u32) (c >> 32) & 0x7fffffffu; }
  const int flag = (cf ? 1 : 0) | (cr ? 2 : 0);
  const int cnt = sext_field ((long long) cf + (long long) cr, 20); // the reference's count:20 signed bit-field
  first.w[0] = (first.w[0] & ~3ULL) | (u64) flag;
  u32 o = n_dist - 1 - d;
  store_rec<W> (d_rec, o, first);
  d_count[o] = cnt;
  keep[o] = qc_mode == QC_BOTH_STRANDS ? (flag == 3) : (qc_mode == QC_SEEN_TWICE ? (cnt > 1) : 1);
}

template <int W>
__global__ void __launch_bounds__ (RED_THREADS)
compact_kernel (const u64 *__restrict__ d_rec, const int *__restrict__ d_count, const u32 *__restrict__ keep,
                const u32 *__restrict__ kpos, u32 n_dist, u64 *__restrict__ k_rec, int *__restrict__ k_count)
{
  u32 i = blockIdx.x * RED_THREADS + threadIdx.x;
  if (i >= n_dist || !keep[i]) return;
  u32 o = kpos[i];
  store_rec<W> (k_rec, o, load_rec<W> (d_rec, i));
  k_count[o] = d_count[i];
}

// context groups (hopo_counter.c:388-404): group g = [gstart[g], gstart[g+1]); depth = sum of counts via the
// exclusive prefix sum csum (u32 wrap-around = the reference's int arithmetic)
__global__ void __launch_bounds__ (RED_THREADS)
group_keep_kernel (const u32 *__restrict__ gstart, u32 n_groups, u32 n_keep, const u32 *__restrict__ csum, const u32 *__restrict__ csum_total,
                   int min_coverage, u32 *__restrict__ gkeep)
{
  u32 g = blockIdx.x * RED_THREADS + threadIdx.x;
  if (g >= n_groups) return;
  u32 s = gstart[g];
  u32 hi = (g + 1 < n_groups) ? csum[gstart[g + 1]] : *csum_total;
  int depth = (int) (hi - csum[s]);
  gkeep[g] = depth >= min_coverage;
  (void) n_keep;
}
__global__ void __launch_bounds__ (RED_THREADS)
group_emit_kernel (const u32 *__restrict__ gstart, const u32 *__restrict__ gkeep, const u32 *__restrict__ gkpos, u32 n_groups, u32 n_keep,
                   int *__restrict__ idx_initial, int *__restrict__ idx_final)
{
  u32 g = blockIdx.x * RED_THREADS + threadIdx.x;
  if (g >= n_groups || !gkeep[g]) return;
  u32 o = gkpos[g];
  idx_initial[o] = (int) gstart[g];
  idx_final[o] = (int) ((g + 1 < n_groups) ? gstart[g + 1] : n_keep);
}

__global__ void __launch_bounds__ (RED_THREADS) icopy_kernel (const int *__restrict__ a, u32 n, u32 *__restrict__ o)
{
  u32 i = blockIdx.x * RED_THREADS + threadIdx.x;
  if (i < n) o[i] = (u32) a[i];
}
