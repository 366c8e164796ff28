// Segmented run-length reduction of sorted tract observations into per-context tract-length histograms.
//
// Global multi-pass fallback for ONE bucket that does not fit in shared memory (tjb_gpu.cu: reduce_oversize_bucket):
// steps 1-4 of finalise_hopo_counter() (reference src/hopo_counter.c:352-413), sequential loops over the sorted
// element array in the reference.
// Input: records sorted ASCENDING on (base, ctx0, ctx1, length, flag).  All outputs are written back to front,
// i.e. in the reference's DESCENDING order.
#pragma once
#include "tjb_common.cuh"

#define RED_THREADS 256

// head[i] = 1 when record i starts a new segment.  MODE 0: (base,ctx0,ctx1,length) changes; MODE 1: (base,ctx0,ctx1)
// changes
template <int W, int MODE>
__global__ void __launch_bounds__ (RED_THREADS) head_kernel (const u64 *__restrict__ a, u32 n, u32 *__restrict__ head)
{
  u32 i = blockIdx.x * RED_THREADS + threadIdx.x;
  if (i >= n) return;
  u32 h = 1;
  if (i) {
    Rec<W> x = load_rec<W> (a, i), y = load_rec<W> (a, i - 1);
    if (MODE == 0) h = !same_tract<W> (x, y);
    else h = !same_context<W> (x, y);
  }
  head[i] = h;
}

// seg_start[pos[i]] = i for every head
__global__ void __launch_bounds__ (RED_THREADS) seg_start_kernel (const u32 *__restrict__ head, const u32 *__restrict__ pos, u32 n, u32 *__restrict__ seg_start)
{
  u32 i = blockIdx.x * RED_THREADS + threadIdx.x;
  if (i < n && head[i]) seg_start[pos[i]] = i;
}

// one thread per distinct (base,ctx0,ctx1,length): count = run length (hopo_counter.c:361), canon_flag = OR of the
// run's flags (:362) = first | last because the flag is the least significant sort digit; QC predicate (:367-374);
// output index reversed -> descending order
template <int W>
__global__ void __launch_bounds__ (RED_THREADS)
rle_emit_kernel (const u64 *__restrict__ sorted, u32 n, const u32 *__restrict__ seg_start, u32 n_dist, int remove_biased,
                 u64 *__restrict__ d_rec, int *__restrict__ d_count, u32 *__restrict__ keep)
{
  u32 d = blockIdx.x * RED_THREADS + threadIdx.x;
  if (d >= n_dist) return;
  u32 s = seg_start[d], e = (d + 1 < n_dist) ? seg_start[d + 1] : n;
  Rec<W> first = load_rec<W> (sorted, s), last = load_rec<W> (sorted, e - 1);
  int flag = rec_flag<W> (first) | rec_flag<W> (last);
  int cnt = sext_field ((long long) (e - s), 20); // the reference's count:20 signed bit-field
  first.w[0] = (first.w[0] & ~3ULL) | (u64) flag;
  u32 o = n_dist - 1 - d;
  store_rec<W> (d_rec, o, first);
  d_count[o] = cnt;
  keep[o] = remove_biased ? (flag == 3) : (cnt > 1);
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
