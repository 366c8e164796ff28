// Cross-sample union of mapper queries (SURVEY.md section 8 e / f-2).
//
// The reference builds ONE mapper query per indexed context group of ONE sample, from the group's first = longest
// record (src/hopo_counter.c:513-520: left flank + base x length + right flank), and maps every sample's queries
// separately (BWA is ~97 % of its run time).  Samples of one population share most contexts, so after the per-sample
// phase the ranks exchange their compact query tables (NCCL allgather, tatajuba_b200/genome_set.py) and this file
// computes the set of DISTINCT queries over all samples plus, for every input row, its index in that set, so that
// each distinct query is mapped once and the result scattered back.  A query's mapping does not depend on the batch
// it is mapped in, so results are unchanged.
//
// A query row is 3 x u64: {ctx0, ctx1, base | length_field << 8} (length_field = the reference's signed 10-bit
// field value).  The union is returned in the reference's own order (base, ctx0, ctx1, length descending,
// hopo_counter.c:28-38).
#pragma once
#include "tjb_common.cuh"

#define UNI_THREADS 256

template <int W>
__global__ void __launch_bounds__ (UNI_THREADS) rows_to_rec_kernel (const u64 *__restrict__ rows, u32 n, int k, u64 *__restrict__ recs)
{
  const u32 i = blockIdx.x * UNI_THREADS + threadIdx.x;
  if (i >= n) return;
  const u64 meta = rows[3 * (size_t) i + 2];
  const int len_field = sext_field ((long long) ((meta >> 8) & 0x3ff), 10);
  store_rec<W> (recs, i, pack_rec<W> (rows[3 * (size_t) i], rows[3 * (size_t) i + 1], (int) (meta & 1), len_field, 0, k));
}

// lexicographic "a < b" on (base, ctx0, ctx1, length), the order of the LSD sort (flag bits are zero here)
template <int W> __device__ __forceinline__ bool rec_less (const Rec<W> &a, const Rec<W> &b)
{
  if (W == 2) return a.w[1] != b.w[1] ? a.w[1] < b.w[1] : a.w[0] < b.w[0];
  const u64 ab = (a.w[0] >> 12) & 1, bb = (b.w[0] >> 12) & 1;
  if (ab != bb) return ab < bb;
  if (a.w[W - 1] != b.w[W - 1]) return a.w[W - 1] < b.w[W - 1];
  if (a.w[1] != b.w[1]) return a.w[1] < b.w[1];
  return (a.w[0] & 0xfff) < (b.w[0] & 0xfff);
}

// distinct records (heads of the ascending sorted array) -> union rows, written back to front (descending order)
template <int W>
__global__ void __launch_bounds__ (UNI_THREADS)
union_emit_kernel (const u64 *__restrict__ sorted, const u32 *__restrict__ head, const u32 *__restrict__ pos, u32 n, const u32 *__restrict__ n_union,
                   int k, u64 *__restrict__ out_rows)
{
  const u32 i = blockIdx.x * UNI_THREADS + threadIdx.x;
  if (i >= n || !head[i]) return;
  const Rec<W> r = load_rec<W> (sorted, i);
  const size_t o = (size_t) (*n_union - 1 - pos[i]);
  out_rows[3 * o] = rec_ctx0<W> (r, k);
  out_rows[3 * o + 1] = rec_ctx1<W> (r, k);
  out_rows[3 * o + 2] = (u64) rec_base<W> (r, k) | ((u64) (rec_len_field<W> (r) & 0x3ff) << 8);
}

// index of every input row in the union: lower bound in the sorted array, then the segment number of that position
template <int W>
__global__ void __launch_bounds__ (UNI_THREADS)
union_index_kernel (const u64 *__restrict__ rows, u32 n, int k, const u64 *__restrict__ sorted, const u32 *__restrict__ head,
                    const u32 *__restrict__ pos, const u32 *__restrict__ n_union, int *__restrict__ index)
{
  const u32 i = blockIdx.x * UNI_THREADS + threadIdx.x;
  if (i >= n) return;
  const u64 meta = rows[3 * (size_t) i + 2];
  const Rec<W> key = pack_rec<W> (rows[3 * (size_t) i], rows[3 * (size_t) i + 1], (int) (meta & 1),
                                  sext_field ((long long) ((meta >> 8) & 0x3ff), 10), 0, k);
  u32 lo = 0, hi = n; // first position whose record is not less than key
  while (lo < hi) {
    const u32 mid = (lo + hi) >> 1;
    if (rec_less<W> (load_rec<W> (sorted, mid), key)) lo = mid + 1; else hi = mid;
  }
  // pos[] is the exclusive scan of head[]: the segment of position lo is pos[lo] + head[lo] - 1
  index[i] = (int) (*n_union - 1 - (pos[lo] + head[lo] - 1));
}
