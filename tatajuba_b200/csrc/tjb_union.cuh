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
#include "tjb_table.cuh"
#include "tjb_bucket.cuh"

#define UNI_THREADS 256

// The union runs through the same machinery as a sample: every row is added to the aggregation table (equal rows of
// different samples meet in one entry), the entries are partitioned, sorted per bucket and emitted in the reference's
// descending order with nothing filtered (QC_KEEP_ALL).
template <int W> __device__ __forceinline__ Rec<W> row_key (const u64 *__restrict__ rows, size_t i, int k)
{
  const u64 meta = rows[3 * i + 2];
  const int len_field = sext_field ((long long) ((meta >> 8) & 0x3ff), 10);
  return pack_rec<W> (rows[3 * i], rows[3 * i + 1], (int) (meta & 1), len_field, 0, k);
}

template <int W>
__global__ void __launch_bounds__ (UNI_THREADS)
rows_insert_kernel (const u64 *__restrict__ rows, u32 n, int k, AggTable tab, u64 *__restrict__ ovf, unsigned long long ovf_capacity,
                    unsigned long long *__restrict__ counters, u32 *__restrict__ bin_hist, int bin_bits)
{
  const u32 i = blockIdx.x * UNI_THREADS + threadIdx.x;
  if (i >= n) return;
  const Rec<W> key = row_key<W> (rows, i, k);
  const int res = table_add<W> (tab, key, 1ULL);
  if (res == TAB_NEW) { atomicAdd (&counters[CTR_ENTRIES], 1ULL); atomicAdd (&bin_hist[rec_bin<W> (key, k, bin_bits)], 1u); }
  else if (res == TAB_DEFER) {
    const unsigned long long pos = atomicAdd (&counters[CTR_OVERFLOW], 1ULL);
    Rec<W> r = key;
    r.w[0] |= 1; // a raw record of weight 1 (flag bit 0)
    if (pos < ovf_capacity) store_rec<W> (ovf, pos, r);
    atomicAdd (&bin_hist[rec_bin<W> (key, k, bin_bits)], 1u);
  }
}

// one warp per bucket: the bucket kernel's compact output -> union rows at the bucket's global offset
template <int W>
__global__ void __launch_bounds__ (UNI_THREADS)
union_rows_kernel (const u64 *__restrict__ out_rec, const u32 *__restrict__ bkt_start, const uint4 *__restrict__ bstat, const u32 *__restrict__ koff,
                   const u32 *__restrict__ meta, int k, u64 *__restrict__ out_rows, u32 max_rows, unsigned long long capacity)
{
  const u32 nb = meta[META_NBUCKETS];
  const int lane = threadIdx.x & 31;
  const u32 warps = gridDim.x * (UNI_THREADS / 32);
  for (u32 b = blockIdx.x * (UNI_THREADS / 32) + (threadIdx.x >> 5); b < nb; b += warps) {
    const uint4 s = bstat[b];
    const u32 start = bkt_start[b], ko = koff[b];
    if ((unsigned long long) start + s.y > capacity) continue;
    for (u32 j = lane; j < s.y; j += 32) {
      if (ko + j >= max_rows) break;
      const Rec<W> r = load_rec<W> (out_rec, (size_t) start + j);
      u64 *o = out_rows + 3 * (size_t) (ko + j);
      o[0] = rec_ctx0<W> (r, k);
      o[1] = rec_ctx1<W> (r, k);
      o[2] = (u64) rec_base<W> (r, k) | ((u64) (rec_len_field<W> (r) & 0x3ff) << 8);
    }
  }
}

// a before b in the reference's descending order (base, ctx0, ctx1, length; the length is a signed 10-bit field)
__device__ __forceinline__ bool row_before (u64 a0, u64 a1, u64 a2, u64 b0, u64 b1, u64 b2)
{
  const u64 ab = a2 & 1, bb = b2 & 1;
  if (ab != bb) return ab > bb;
  if (a0 != b0) return a0 > b0;
  if (a1 != b1) return a1 > b1;
  return sext_field ((long long) ((a2 >> 8) & 0x3ff), 10) > sext_field ((long long) ((b2 >> 8) & 0x3ff), 10);
}

// index of every input row in the union: first union row that is not before it
__global__ void __launch_bounds__ (UNI_THREADS)
union_index_kernel (const u64 *__restrict__ rows, u32 n, const u64 *__restrict__ uni, const u32 *__restrict__ meta, u32 max_rows, int *__restrict__ index)
{
  const u32 i = blockIdx.x * UNI_THREADS + threadIdx.x;
  if (i >= n) return;
  const u64 r0 = rows[3 * (size_t) i], r1 = rows[3 * (size_t) i + 1], r2 = rows[3 * (size_t) i + 2] & 0x3ff01ULL;
  u32 lo = 0, hi = meta[META_NKEEP] < max_rows ? meta[META_NKEEP] : max_rows;
  while (lo < hi) {
    const u32 mid = (lo + hi) >> 1;
    if (row_before (uni[3 * (size_t) mid], uni[3 * (size_t) mid + 1], uni[3 * (size_t) mid + 2], r0, r1, r2)) lo = mid + 1; else hi = mid;
  }
  index[i] = (int) lo;
}

// query rows of a counted sample, on the device: the first = longest record of every indexed context group
// (src/hopo_counter.c:513-520), straight from the bucket kernel's output (one warp per bucket, like pack_kernel)
template <int W>
__global__ void __launch_bounds__ (UNI_THREADS)
query_rows_kernel (const u64 *__restrict__ out_rec, const u32 *__restrict__ out_idx_i, const u32 *__restrict__ bkt_start, const uint4 *__restrict__ bstat,
                   const u32 *__restrict__ ioff, const u32 *__restrict__ meta, int k, u64 *__restrict__ rows, u32 max_rows, unsigned long long capacity)
{
  const u32 nb = meta[META_NBUCKETS];
  const int lane = threadIdx.x & 31;
  const u32 warps = gridDim.x * (UNI_THREADS / 32);
  for (u32 b = blockIdx.x * (UNI_THREADS / 32) + (threadIdx.x >> 5); b < nb; b += warps) {
    const uint4 s = bstat[b];
    const u32 start = bkt_start[b], io = ioff[b];
    if ((unsigned long long) start + s.y > capacity || (unsigned long long) start + s.z > capacity) continue;
    for (u32 t = lane; t < s.z; t += 32) {
      if (io + t >= max_rows) break;
      const Rec<W> r = load_rec<W> (out_rec, (size_t) start + out_idx_i[start + t]);
      u64 *o = rows + 3 * (size_t) (io + t);
      o[0] = rec_ctx0<W> (r, k);
      o[1] = rec_ctx1<W> (r, k);
      o[2] = (u64) rec_base<W> (r, k) | ((u64) (rec_len_field<W> (r) & 0x3ff) << 8);
    }
  }
}
