// Aggregation table: tract observations are counted where they are found.
//
// The reference appends one 40-byte element per observed tract (src/hopo_counter.c:285-307), qsorts them all and only
// then counts equal neighbours (:351-365).  At sequencing depth D every (base, ctx0, ctx1, length) arrives ~D times, so
// the scan kernel instead adds each observation to an open-addressing table in global memory that stays L2 resident
// (keys: the W record words without the flag bits; value: forward-strand count | reverse-strand count << 32, which is
// the run-length count and the canon_flag OR of :361-362 in one 64-bit add).  Everything after the scan works on the
// DISTINCT observations ("weighted records": key + count word), 58 K instead of 3.75 M at BASELINE config 2.
//
// Slot protocol (no thread ever waits for another; keys AND counts are zeroed per sample, and no real key is zero: the
// length field is biased by 512):
//   key == mine                   add my weight to the slot's count word (one load + one RED: the common case)
//   key == 0                      empty, or claimed a moment ago: CAS count 0 -> my weight claims the slot, the claimer
//                                 then stores the key.  A thread that loses the CAS reads the key a few more times:
//                                 mine -> add; still 0 -> the observation is DEFERRED to the overflow list; else probe on
//   otherwise                     another key: probe on (linear)
// After TAB_MAXPROBE probes the observation is deferred as well.  Deferred observations keep their raw form (one record
// each, flag bits set); the partition kernel turns them into weighted records of weight 1, so a key may reach the
// bucket kernel in several records: it adds the weights of equal keys after sorting.  (A key can also end up in two
// slots when a reader sees a slot half written and probes on: same remedy.)  W = 3 keys take two 16-byte halves
// {w0, w1} {w2, w0}: the copy of the never-zero w0 in the second half tells a reader that both halves are there.
#pragma once
#include "tjb_common.cuh"

#ifndef TAB_MAXPROBE
#define TAB_MAXPROBE 48
#endif
#define TAB_RELOOK 32 // looks at a slot whose claim was lost before the observation is deferred

struct AggTable {
  unsigned long long *cnt;   // [slots] forward count | reverse count << 32
  u64 *keys;                 // [slots][2] (W = 2) or [slots][4] (W = 3), 16-byte aligned
  u32 mask;                  // slots - 1 (power of two)
};
__host__ __device__ inline int table_key_words (int W) { return W == 2 ? 2 : 4; }
// bytes of a table of that many slots: counts first, keys behind them (one allocation, one memset per sample)
__host__ __device__ inline size_t table_bytes (size_t slots, int W) { return slots * 8 * (1 + (size_t) table_key_words (W)); }

// the table lives in global memory: say so, or a pointer that went through a function call is treated as generic
// (ATOM instead of RED, and a shared-memory path next to every atomic)
__device__ __forceinline__ void red_add_global (unsigned long long *p, unsigned long long v)
{ asm volatile ("red.global.add.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ unsigned long long cas_global (unsigned long long *p, unsigned long long cmp, unsigned long long val)
{
  unsigned long long old;
  asm volatile ("atom.global.cas.b64 %0, [%1], %2, %3;" : "=l"(old) : "l"(p), "l"(cmp), "l"(val) : "memory");
  return old;
}

// weight of one raw observation: flag 1 = seen on the forward strand, 2 = reverse complement (hopo_counter.c:237, 245)
__device__ __forceinline__ unsigned long long flag_weight (u64 w0) { return 1ULL << (((u32) w0 & 2u) << 4); }
__device__ __forceinline__ void red_add_global32 (u32 *p, u32 v)
{ asm volatile ("red.global.add.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory"); }
// one observation onto a slot that holds its key: +1 on the strand's half of the count word
__device__ __forceinline__ void table_count_one (const AggTable &t, u32 slot, u64 w0)
{ red_add_global32 (reinterpret_cast<u32*> (t.cnt + slot) + (((u32) w0 >> 1) & 1u), 1u); }

template <int W> __device__ __forceinline__ u32 key_hash (const Rec<W> &r) // flag bits do not take part
{
  const u64 a = (r.w[0] >> 2) * 0x9E3779B97F4A7C15ULL, b = r.w[1] * 0xC2B2AE3D27D4EB4FULL;
  u32 h = (u32) (a >> 32) ^ (u32) (b >> 32) ^ (u32) b;
  if (W == 3) { const u64 c = r.w[W - 1] * 0x165667B19E3779F9ULL; h ^= (u32) (c >> 32) ^ (u32) c; }
  return h ^ (h >> 15);
}

// what a slot's key reads as: 0 = nothing there (yet), 1 = `key`, 2 = something else (another key, or a key half written)
template <int W> __device__ __forceinline__ int table_probe (const AggTable &t, u32 slot, const Rec<W> &key)
{
  if (W == 2) {
    const ulonglong2 v = __ldcg (reinterpret_cast<const ulonglong2*> (t.keys) + slot); // from L2: never a stale L1 line
    if (v.x == key.w[0] && v.y == key.w[1]) return 1;
    return v.x == 0 ? 0 : 2;
  }
  const ulonglong2 a = __ldcg (reinterpret_cast<const ulonglong2*> (t.keys) + 2 * (size_t) slot);
  const ulonglong2 b = __ldcg (reinterpret_cast<const ulonglong2*> (t.keys) + 2 * (size_t) slot + 1);
  if (a.x == key.w[0] && a.y == key.w[1] && b.x == key.w[2] && b.y == key.w[0]) return 1;
  return (a.x == 0 && b.y == 0) ? 0 : 2;
}

// key of an occupied slot (after the kernel that filled the table has finished)
template <int W> __device__ __forceinline__ Rec<W> table_key_at (const AggTable &t, u32 slot)
{
  Rec<W> y;
  const ulonglong2 a = __ldcg (reinterpret_cast<const ulonglong2*> (t.keys) + (W == 2 ? 1 : 2) * (size_t) slot);
  y.w[0] = a.x; y.w[1] = a.y;
  if (W == 3) y.w[W - 1] = __ldcg (t.keys + 4 * (size_t) slot + 2);
  return y;
}

// result of table_add
enum { TAB_ADDED = 0, TAB_NEW = 1, TAB_DEFER = 2 };

// key: record words with the flag bits cleared.  Returns TAB_NEW when this call created the entry (the caller counts
// entries and bins them), TAB_DEFER when the observation has to go to the overflow list.  One copy of the code per
// kernel (it is called from several places of the scan, and the probe loop must not be unrolled into each of them).
template <int W> __device__ __noinline__ int table_add_from (AggTable t, Rec<W> key, unsigned long long weight, u32 slot, int seen)
{
  // slot / seen: where the caller's own look at the home slot left off (seen < 0: nothing looked at yet)
#pragma unroll 1
  for (int p = 0; p < TAB_MAXPROBE; p++, slot = (slot + 1) & t.mask) {
    if (seen < 0) seen = table_probe<W> (t, slot, key);
    if (seen == 0) {
      if (cas_global (t.cnt + slot, 0ULL, weight) == 0ULL) {
        if (W == 2) __stcg (reinterpret_cast<ulonglong2*> (t.keys) + slot, make_ulonglong2 (key.w[0], key.w[1]));
        else {
          __stcg (reinterpret_cast<ulonglong2*> (t.keys) + 2 * (size_t) slot, make_ulonglong2 (key.w[0], key.w[1]));
          __stcg (reinterpret_cast<ulonglong2*> (t.keys) + 2 * (size_t) slot + 1, make_ulonglong2 (key.w[W - 1], key.w[0]));
        }
        return TAB_NEW;
      }
      // claimed by somebody else a moment ago: its key is on its way (the claimer stores it right after its CAS).  Look
      // again a few times (bounded: nobody is waited for indefinitely) before giving the observation to the overflow list;
      // when a kernel adds all records of a sample at once, the ~depth copies of a new key arrive together
      seen = table_probe<W> (t, slot, key);
      for (int again = 0; again < TAB_RELOOK && seen == 0; again++) { __nanosleep (64); seen = table_probe<W> (t, slot, key); }
      if (seen == 0) return TAB_DEFER;
    }
    if (seen == 1) { red_add_global (t.cnt + slot, weight); return TAB_ADDED; }
    seen = -1;
  }
  return TAB_DEFER;
}
template <int W> __device__ __forceinline__ int table_add (const AggTable &t, const Rec<W> &key, unsigned long long weight)
{ return table_add_from<W> (t, key, weight, key_hash<W> (key) & t.mask, -1); }
