// Aggregation table: tract observations are counted where they are found.
//
// The reference appends one 40-byte element per observed tract (src/hopo_counter.c:285-307), qsorts them all and only
// then counts equal neighbours (:351-365).  At sequencing depth D every (base, ctx0, ctx1, length) arrives ~D times, so
// the scan kernel instead adds each observation to an open-addressing table in global memory that stays L2 resident
// (keys: the W record words without the flag bits; value: forward-strand count | reverse-strand count << 32, which is
// the run-length count and the canon_flag OR of :361-362 in one 64-bit add).  Everything after the scan works on the
// DISTINCT observations ("weighted records": key + count word), 58 K instead of 3.75 M at BASELINE config 2.
//
// Slot protocol (no thread ever waits for another):
//   cnt == 0                      empty.  CAS 0 -> BUSY claims the slot; the claimer stores the key, fences, and adds
//                                 (weight + BUSY), which clears the busy bit (2^63 + 2^63 wraps) and publishes the slot
//   cnt & BUSY                    being written by its claimer: the key cannot be compared yet -> the observation is
//                                 DEFERRED to the overflow list (a few records per sample)
//   otherwise                     ready: compare the key, add the weight on a match, else probe on (linear)
// After TAB_MAXPROBE probes the observation is deferred as well.  Deferred observations keep their raw form (one record
// each, flag bits set); the partition kernel turns them into weighted records of weight 1, so a key may reach the
// bucket kernel in several records: it adds the weights of equal keys after sorting.  Keys are never cleared: a
// slot's key is only read once its count says it has been written for THIS sample (the count array is zeroed per sample).
#pragma once
#include "tjb_common.cuh"

#define TAB_BUSY (1ULL << 63)
#ifndef TAB_MAXPROBE
#define TAB_MAXPROBE 48
#endif

struct AggTable {
  u64 *keys;                 // [slots][W]
  unsigned long long *cnt;   // [slots] forward count | reverse count << 32 (bit 63: busy)
  u32 mask;                  // slots - 1 (power of two)
};

__device__ __forceinline__ unsigned long long ld_relaxed_u64 (const unsigned long long *p) // from L2, never from a stale L1 line
{
  unsigned long long v;
  asm volatile ("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_acquire_u64 (const unsigned long long *p)
{
  unsigned long long v;
  asm volatile ("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// weight of one raw observation: flag 1 = seen on the forward strand, 2 = reverse complement (hopo_counter.c:237, 245)
__device__ __forceinline__ unsigned long long flag_weight (u64 w0) { return (w0 & 1) ? 1ULL : (1ULL << 32); }
// canon_flag of a count word (:362) and the reference's signed 20-bit count field (hopo_counter.h:40)
__device__ __forceinline__ u32 weight_flag (unsigned long long c) { return ((u32) c ? 1u : 0u) | ((u32) (c >> 32) & 0x7fffffffu ? 2u : 0u); }
__device__ __forceinline__ int weight_count20 (unsigned long long c) { return sext_field ((long long) (u32) c + (long long) ((u32) (c >> 32) & 0x7fffffffu), 20); }

template <int W> __device__ __forceinline__ u32 key_hash (const Rec<W> &r) // flag bits do not take part
{
  u32 h = (u32) (r.w[0] >> 2) * 0x9E3779B1u ^ (u32) (r.w[0] >> 34) * 0x85EBCA6Bu ^ (u32) r.w[1] * 0xC2B2AE35u ^ (u32) (r.w[1] >> 32) * 0x27D4EB2Fu;
  if (W == 3) h ^= (u32) r.w[W - 1] * 0x165667B1u ^ (u32) (r.w[W - 1] >> 32) * 0x9E3779B1u;
  h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 13;
  return h;
}

template <int W> __device__ __forceinline__ Rec<W> table_key_at (const AggTable &t, u32 slot)
{
  Rec<W> y;
  if (W == 2) {
    const ulonglong2 v = __ldcg (reinterpret_cast<const ulonglong2*> (t.keys) + slot);
    y.w[0] = v.x; y.w[1] = v.y;
  } else {
#pragma unroll
    for (int j = 0; j < W; j++) y.w[j] = __ldcg (t.keys + (size_t) W * slot + j);
  }
  return y;
}

// result of table_add
enum { TAB_ADDED = 0, TAB_NEW = 1, TAB_DEFER = 2 };

// key: record words with the flag bits cleared.  Returns TAB_NEW when this call created the entry (the caller counts
// entries and bins them), TAB_DEFER when the observation has to go to the overflow list.
template <int W> __device__ __forceinline__ int table_add (const AggTable &t, const Rec<W> &key, unsigned long long weight)
{
  u32 slot = key_hash<W> (key) & t.mask;
  for (int p = 0; p < TAB_MAXPROBE; p++, slot = (slot + 1) & t.mask) {
    unsigned long long c = ld_acquire_u64 (t.cnt + slot);
    if (c == 0) {
      c = atomicCAS (t.cnt + slot, 0ULL, TAB_BUSY);
      if (c == 0) {
        if (W == 2) __stcg (reinterpret_cast<ulonglong2*> (t.keys) + slot, make_ulonglong2 (key.w[0], key.w[1]));
        else {
#pragma unroll
          for (int j = 0; j < W; j++) __stcg (t.keys + (size_t) W * slot + j, key.w[j]);
        }
        __threadfence (); // the key is visible before the count says so
        atomicAdd (t.cnt + slot, weight + TAB_BUSY);
        return TAB_NEW;
      }
    }
    if (c & TAB_BUSY) return TAB_DEFER;
    const Rec<W> y = table_key_at<W> (t, slot);
    bool eq = true;
#pragma unroll
    for (int j = 0; j < W; j++) eq = eq && (y.w[j] == key.w[j]);
    if (eq) { atomicAdd (t.cnt + slot, weight); return TAB_ADDED; }
  }
  return TAB_DEFER;
}
