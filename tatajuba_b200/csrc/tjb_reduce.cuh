// Segmented run-length reduction of sorted tract observations into per-context tract-length histograms.
//
// Replaces steps 1-4 of finalise_hopo_counter() and estimate_coverage_hopo_counter() (reference
// src/hopo_counter.c:352-413, 420-438), all of which are sequential loops over the sorted element array.
// Input: records sorted ASCENDING on (base, ctx0, ctx1, length, flag).  All outputs are written back to front,
// i.e. in the reference's DESCENDING order.
#pragma once
#include "tjb_common.cuh"

#define RED_THREADS 256

// head[i] = 1 when record i starts a new segment.  MODE 0: (base,ctx0,ctx1,length) changes; MODE 1: (base,ctx0,ctx1)
// changes; MODE 2 (W = 1, coverage pairs): bits 32..63 change
template <int W, int MODE>
__global__ void __launch_bounds__ (RED_THREADS) head_kernel (const u64 *__restrict__ a, u32 n, u32 *__restrict__ head)
{
  u32 i = blockIdx.x * RED_THREADS + threadIdx.x;
  if (i >= n) return;
  u32 h = 1;
  if (i) {
    Rec<W> x = load_rec<W> (a, i), y = load_rec<W> (a, i - 1);
    if (MODE == 0) h = !same_tract<W> (x, y);
    else if (MODE == 1) h = !same_context<W> (x, y);
    else h = (x.w[0] >> 32) != (y.w[0] >> 32);
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

// coverage (hopo_counter.c:427-432): two (key, weight) pairs per kept record, key = low 31 bits of each flank
template <int W>
__global__ void __launch_bounds__ (RED_THREADS)
cov_pairs_kernel (const u64 *__restrict__ k_rec, const int *__restrict__ k_count, u32 n_keep, int k, u64 *__restrict__ pairs)
{
  u32 i = blockIdx.x * RED_THREADS + threadIdx.x;
  if (i >= n_keep) return;
  Rec<W> r = load_rec<W> (k_rec, i);
  u64 w = (u64) (u32) k_count[i];
  pairs[i] = ((rec_ctx0<W> (r, k) & 0x7fffffffULL) << 32) | w;
  pairs[n_keep + i] = ((rec_ctx1<W> (r, k) & 0x7fffffffULL) << 32) | w;
}
__global__ void __launch_bounds__ (RED_THREADS) lo32_kernel (const u64 *__restrict__ a, u32 n, u32 *__restrict__ lo)
{
  u32 i = blockIdx.x * RED_THREADS + threadIdx.x;
  if (i < n) lo[i] = (u32) a[i];
}
__global__ void __launch_bounds__ (RED_THREADS) icopy_kernel (const int *__restrict__ a, u32 n, u32 *__restrict__ o)
{
  u32 i = blockIdx.x * RED_THREADS + threadIdx.x;
  if (i < n) o[i] = (u32) a[i];
}
// coverage = max over distinct keys of the summed weights (empirical_frequency.c:254-277, hopo_counter.c:434)
__global__ void __launch_bounds__ (RED_THREADS)
cov_max_kernel (const u32 *__restrict__ seg_start, u32 n_seg, u32 n, const u32 *__restrict__ wsum, const u32 *__restrict__ wtotal, int *__restrict__ cov)
{
  u32 g = blockIdx.x * RED_THREADS + threadIdx.x;
  if (g >= n_seg) return;
  u32 hi = (g + 1 < n_seg) ? wsum[seg_start[g + 1]] : *wtotal;
  int sum = (int) (hi - wsum[seg_start[g]]);
  atomicMax (cov, sum);
  (void) n;
}

// final table -> reference-layout 40-byte hopo_element (src/hopo_counter.h:34-49) as it stands at the BWA boundary
template <int W>
__global__ void __launch_bounds__ (RED_THREADS)
pack_hopo_kernel (const u64 *__restrict__ k_rec, const int *__restrict__ k_count, u32 n_keep, int k, u64 *__restrict__ out40)
{
  u32 i = blockIdx.x * RED_THREADS + threadIdx.x;
  if (i >= n_keep) return;
  Rec<W> r = load_rec<W> (k_rec, i);
  // gcc x86-64 bit-field order from the LSB: base:2 length:10 count:20 mismatches:12 multi:3 neg_strand:2 canon_flag:3
  u64 bits = ((u64) rec_base<W> (r, k) & 3) | (((u64) rec_len_field<W> (r) & 0x3ff) << 2) |
             (((u64) k_count[i] & 0xfffff) << 12) | (0xffeULL << 32) | (((u64) rec_flag<W> (r) & 7) << 49);
  u64 *o = out40 + 5 * (size_t) i;
  o[0] = rec_ctx0<W> (r, k);
  o[1] = rec_ctx1<W> (r, k);
  o[2] = bits;
  o[3] = ~0ULL; // read_offset = -1 (hopo_counter.c:511), loc_ref_id = -1 (:297)
  o[4] = ~0ULL; // loc_pos = loc_last = -1
}
