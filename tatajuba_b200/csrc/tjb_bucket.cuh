// Bucketed sort + reduction of the DISTINCT tract observations (the reduce path after the aggregation table).
//
// Replaces qsort + the sequential loops of finalise_hopo_counter() and estimate_coverage_hopo_counter()
// (reference src/hopo_counter.c:351-413, 420-438).  The run-length count of :352-365 has already happened in the
// aggregation table the scan fills (tjb_table.cuh); what arrives here are weighted records: key (base, ctx0, ctx1,
// length) + count word (forward | reverse << 32).  They are moved ONCE (an MSD partition on the top F bits of
// (base, ctx0[, ctx1]), F <= 20, splitters chosen from the histogram the scan accumulates when it creates a table
// entry), and every bucket (<= BK_CAP records, sized to stay in shared memory) is sorted on chip (comparison counting
// for <= 128 records, LSD radix over the digits that differ otherwise) and reduced in the same CTA: weights of equal
// keys added (a key can arrive in several records, see tjb_table.cuh), strand-bias / singleton QC (:367-374) and the
// context groups above the minimum coverage (:388-404).  A context (base, ctx0, ctx1) never straddles two buckets
// because the partition looks at context bits only.  Everything is produced in the reference's DESCENDING order.
//
//   plan_kernel           1 CTA     bin histogram -> splitters (buckets of < BK_CAP records), bin -> bucket LUT
//   wpartition_kernel     grid      table entries + overflow list -> weighted records, bucket-contiguous
//   wbucket_reduce_kernel grid      per bucket: smem sort + merge + QC + groups; compact output at the bucket start
//   totals_kernel         1 CTA     exclusive scans of the per-bucket kept / group counts
//   pack_kernel           grid      reference-layout 40-byte hopo_element rows + idx arrays + coverage hash inserts
//   cov_max_kernel        grid      coverage = max over the hash table (only when a count wrapped; pack_kernel keeps the max otherwise)
//
// Buckets with more than BK_CAP records (one bin holding > BK_CAP/2 distinct observations: pathological samples) are
// flagged and reduced by the multi-pass global path in tjb_sort.cuh / tjb_reduce.cuh instead.
#pragma once
#include "tjb_common.cuh"
#include "tjb_table.cuh"

#define BK_THREADS 256
#define BK_WARPS (BK_THREADS / 32)
#ifndef BK_IPT
#define BK_IPT 8
#endif
#ifndef BK_CTAS2
#define BK_CTAS2 3 // bucket CTAs per SM, 2-word records (shared memory and registers allow 3 at BK_IPT 8)
#endif
#define BK_CAP (BK_THREADS * BK_IPT) // records per on-chip bucket
#define BK_CAPH (BK_CAP / 2)
#define PLAN_THREADS 1024
#define TJB_MAX_BIN_BITS 20
#define PLAN_REG_BIN_BITS 16 // up to this many bits the planning CTA keeps its bins in registers (plan_kernel), above: plan_big_kernel
#define TJB_MAX_BINS (1 << TJB_MAX_BIN_BITS)
#ifndef PART_GRID
#define PART_GRID 8   // partition CTAs (of 256 threads) per SM
#endif
#ifndef PART_UNROLL
#define PART_UNROLL 4 // records in flight per thread of the partition kernel
#endif
#define BK_CURSOR_STRIDE 32 // one partition cursor per 128-byte line: L2 atomics serialise per sector, not per address

// meta[] slots (device u32)
enum { META_NBUCKETS = 0, META_NOVERSIZE = 1, META_TICKET = 2, META_N = 3, META_NDIST = 4, META_NKEEP = 5, META_NIDX = 6,
       META_COVSCAN = 7, // a kept record has a count <= 0 (its 20-bit field wrapped): the coverage needs the table scan
       META_WORDS = 8 };

struct DigitList { int n; unsigned char word[28], shift[28], nbits[28]; };

// bin = top F bits of the context (base, ctx0[, ctx1]); F <= TJB_MAX_BIN_BITS
template <int W> __device__ __forceinline__ u32 rec_bin (const Rec<W> &r, int k, int F);
template <> __device__ __forceinline__ u32 rec_bin<2> (const Rec<2> &r, int k, int F)
{
  const int top = 2 * k + 1; // significant bits of w1 = base:ctx0
  if (F <= top) return (u32) (r.w[1] >> (top - F));
  const int extra = F - top; // spill into the top bits of ctx1 (w0 bits 12 .. 12+2k)
  return ((u32) r.w[1] << extra) | (u32) ((r.w[0] >> (12 + 2 * k - extra)) & ((1u << extra) - 1));
}
template <> __device__ __forceinline__ u32 rec_bin<3> (const Rec<3> &r, int k, int F)
{
  return ((u32) ((r.w[0] >> 12) & 1) << (F - 1)) | (u32) (r.w[2] >> (2 * k - (F - 1)));
}

// ------------------------------------------------------------------------------- block scan helpers ------
// exclusive scan of one value per thread over a block of NT threads; *total = block sum.  ONE barrier: every warp
// scans the warp totals itself, and consecutive calls alternate between the two halves of s_warp (phase), so a warp
// that runs ahead into the next call writes the other half; the call after that is behind the next call's barrier,
// which no warp passes before every warp has read this call's totals.  All threads must pass the same phase.
template <int NT>
__device__ __forceinline__ u32 blk_excl_scan (u32 v, u32 *s_warp /*[2 * NT / 32]*/, int &phase, u32 *total)
{
  constexpr int NW = NT / 32;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  u32 *buf = s_warp + phase * NW;
  phase ^= 1;
  u32 inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { u32 t = __shfl_up_sync (0xffffffffu, inc, d); if (lane >= d) inc += t; }
  if (lane == 31) buf[wid] = inc;
  __syncthreads ();
  const u32 w = lane < NW ? buf[lane] : 0;
  u32 winc = w;
#pragma unroll
  for (int d = 1; d < NW; d <<= 1) { u32 t = __shfl_up_sync (0xffffffffu, winc, d); if (lane >= d) winc += t; }
  const u32 before = __shfl_sync (0xffffffffu, winc - w, wid);
  *total = __shfl_sync (0xffffffffu, winc, NW - 1);
  return before + inc - v;
}

// --------------------------------------------------------------------------------------- prologue --------
// clears the device scalars of a sample in one launch: result block (counters, meta; coverage = INT_MIN), the window
// ticket of the first scan launch and the bins the scan is going to count
__global__ void __launch_bounds__ (256)
sample_prologue_kernel (u32 *__restrict__ res_words, int n_res_words, int cov_word, u32 *__restrict__ tile_ctr, u32 *__restrict__ bin_hist, u32 n_bins)
{
  const u32 t = blockIdx.x * 256 + threadIdx.x;
  if (t < (u32) n_res_words) res_words[t] = t == (u32) cov_word ? 0x80000000u : 0u;
  if (t == 0) *tile_ctr = 0;
  for (u32 i = t; i < n_bins; i += gridDim.x * 256) bin_hist[i] = 0;
}

// ------------------------------------------------------------------------------------------- plan --------
// Bins are walked from the highest to the lowest (descending key order).  P[j] = number of records in bins before
// j in that order; bins whose P falls into the same window of 2^bshift records form one bucket, so a bucket holds
// fewer than 2^bshift + (its last bin) records: at most BK_CAP unless a single bin exceeds BK_CAP - 2^bshift.  The
// host picks bshift <= log2 (BK_CAPH) so that a sample's records spread over all SMs (a small sample in buckets of
// ~128 records, which the bucket kernel orders by comparison counting instead of radix passes).
// One CTA; thread t owns IPT consecutive bins, loaded up front with 16-byte loads so that every global load of the
// kernel is in flight at once (the first version walked them with dependent scalar loads: 195 us for 32 Ki bins).
template <int IPT>
__global__ void __launch_bounds__ (PLAN_THREADS)
plan_kernel (const u32 *__restrict__ bin_hist, int F, int bshift, unsigned long long capacity,
             u32 *__restrict__ bkt_start, u32 *__restrict__ lut, u32 *__restrict__ cursor, u32 *__restrict__ meta, u32 max_buckets)
{
  __shared__ u32 s_warp[2 * (PLAN_THREADS / 32)];
  int scan_phase = 0;
  const u32 NB = 1u << F;
  const u32 t = threadIdx.x;
  const bool active = t * IPT < NB;       // NB < PLAN_THREADS * IPT only when IPT == 1
  const u32 lo = NB - (t + 1) * IPT;      // my bins, ascending in memory = descending walk order reversed
  u32 h[IPT];
  if (active) {
    if (IPT >= 4) {
#pragma unroll
      for (int i = 0; i < IPT / 4; i++) {
        const uint4 v = *reinterpret_cast<const uint4*> (bin_hist + lo + 4 * i);
        h[4 * i] = v.x; h[4 * i + 1] = v.y; h[4 * i + 2] = v.z; h[4 * i + 3] = v.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < IPT; i++) h[i] = bin_hist[lo + i];
    }
  } else {
#pragma unroll
    for (int i = 0; i < IPT; i++) h[i] = 0;
  }
  u32 sum = 0;
#pragma unroll
  for (int i = 0; i < IPT; i++) sum += h[i];
  u32 total;
  const u32 P0 = blk_excl_scan<PLAN_THREADS> (sum, s_warp, scan_phase, &total);
  // records of the bin walked just before my first one (bin lo + IPT)
  const u32 hprev = (active && t > 0) ? bin_hist[lo + IPT] : 0;
  u32 nheads = 0;
  {
    u32 p = P0, hp = hprev;
#pragma unroll
    for (int i = IPT - 1; i >= 0; i--) {
      const bool head = active && ((t == 0 && i == IPT - 1) || ((p >> bshift) != ((p - hp) >> bshift)));
      nheads += head;
      hp = h[i]; p += h[i];
    }
  }
  u32 nb_total;
  u32 b = blk_excl_scan<PLAN_THREADS> (nheads, s_warp, scan_phase, &nb_total); // index of my first head
  if (active) {
    u32 p = P0, hp = hprev;
#pragma unroll
    for (int i = IPT - 1; i >= 0; i--) {
      const bool head = (t == 0 && i == IPT - 1) || ((p >> bshift) != ((p - hp) >> bshift));
      // (more heads than max_buckets only after a record-array overflow: the sample is redone anyway)
      if (head) { if (b < max_buckets) { bkt_start[b] = p; cursor[(size_t) b * BK_CURSOR_STRIDE] = p; } b++; }
      hp = h[i]; p += h[i];
      h[i] = b - 1 < max_buckets ? b - 1 : max_buckets - 1; // bucket of bin lo + i (a range that does not start with a head continues bucket b - 1)
    }
    if (IPT >= 4) { // 16-byte stores of the bin -> bucket table
#pragma unroll
      for (int i = 0; i < IPT / 4; i++)
        *reinterpret_cast<uint4*> (lut + lo + 4 * i) = make_uint4 (h[4 * i], h[4 * i + 1], h[4 * i + 2], h[4 * i + 3]);
    } else {
#pragma unroll
      for (int i = 0; i < IPT; i++) lut[lo + i] = h[i];
    }
  }
  if (t == 0) {
    if (nb_total > max_buckets) nb_total = max_buckets;
    bkt_start[nb_total] = total;
    meta[META_NBUCKETS] = nb_total;
    meta[META_NOVERSIZE] = 0;
    meta[META_TICKET] = 0;
    meta[META_N] = (u32) (total < capacity ? total : capacity); // weighted records (more than the capacity: the sample is redone)
  }
}

// The same plan for more bins than fit in the registers of one CTA (F > PLAN_REG_BIN_BITS: deep or very large
// samples, where 2^16 bins would hold more than BK_CAPH records each): every thread walks its 2^F / 1024 bins three
// times from L2 (sum, count heads, emit) instead of keeping them in registers.
__global__ void __launch_bounds__ (PLAN_THREADS)
plan_big_kernel (const u32 *__restrict__ bin_hist, int F, int bshift, unsigned long long capacity,
                 u32 *__restrict__ bkt_start, u32 *__restrict__ lut, u32 *__restrict__ cursor, u32 *__restrict__ meta, u32 max_buckets)
{
  __shared__ u32 s_warp[2 * (PLAN_THREADS / 32)];
  int scan_phase = 0;
  const u32 NB = 1u << F, ipt = NB / PLAN_THREADS; // F > 10: every thread owns ipt >= 64 bins
  const u32 t = threadIdx.x;
  const u32 lo = NB - (t + 1) * ipt;
  u32 sum = 0;
  for (u32 i = 0; i < ipt; i += 4) { const uint4 v = *reinterpret_cast<const uint4*> (bin_hist + lo + i); sum += v.x + v.y + v.z + v.w; }
  u32 total;
  const u32 P0 = blk_excl_scan<PLAN_THREADS> (sum, s_warp, scan_phase, &total);
  const u32 hprev = t > 0 ? bin_hist[lo + ipt] : 0;
  u32 nheads = 0;
  {
    u32 p = P0, hp = hprev;
    for (u32 i = ipt; i > 0; i -= 4) { // 16-byte loads, bins walked from the highest to the lowest
      const uint4 v = *reinterpret_cast<const uint4*> (bin_hist + lo + i - 4);
      const u32 h4[4] = {v.w, v.z, v.y, v.x};
#pragma unroll
      for (int j = 0; j < 4; j++) {
        nheads += (t == 0 && i == ipt && j == 0) || ((p >> bshift) != ((p - hp) >> bshift));
        hp = h4[j]; p += h4[j];
      }
    }
  }
  u32 nb_total;
  u32 b = blk_excl_scan<PLAN_THREADS> (nheads, s_warp, scan_phase, &nb_total);
  {
    u32 p = P0, hp = hprev;
    for (u32 i = ipt; i > 0; i -= 4) {
      const uint4 v = *reinterpret_cast<const uint4*> (bin_hist + lo + i - 4);
      const u32 h4[4] = {v.w, v.z, v.y, v.x};
      u32 l4[4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const bool head = (t == 0 && i == ipt && j == 0) || ((p >> bshift) != ((p - hp) >> bshift));
        if (head) { if (b < max_buckets) { bkt_start[b] = p; cursor[(size_t) b * BK_CURSOR_STRIDE] = p; } b++; }
        hp = h4[j]; p += h4[j];
        l4[j] = b - 1 < max_buckets ? b - 1 : max_buckets - 1;
      }
      *reinterpret_cast<uint4*> (lut + lo + i - 4) = make_uint4 (l4[3], l4[2], l4[1], l4[0]);
    }
  }
  if (t == 0) {
    if (nb_total > max_buckets) nb_total = max_buckets;
    bkt_start[nb_total] = total;
    meta[META_NBUCKETS] = nb_total;
    meta[META_NOVERSIZE] = 0;
    meta[META_TICKET] = 0;
    meta[META_N] = (u32) (total < capacity ? total : capacity); // weighted records (more than the capacity: the sample is redone)
  }
}

// --------------------------------------------------------------------------------------- aggregate -------
// The scan's records (one per observed tract) are added to the aggregation table (tjb_table.cuh): the run-length count
// of hopo_counter.c:352-365 without a sort.  One record per thread and round, AGG_UNROLL rounds in flight; a thread
// that creates an entry also counts it for the partition plan.  (Measured, config 2 / config 5 (k=32, 39 M records on
// 234 K keys): 1 round in flight 57 / 623 us, 2: 55 / 628, 4: 62 / 1023 — more probes in flight only add reductions
// queueing on the same count words.)
#ifndef AGG_UNROLL
#define AGG_UNROLL 2
#endif
template <int W>
__global__ void __launch_bounds__ (256)
aggregate_kernel (const u64 *__restrict__ recs, unsigned long long *__restrict__ counters, unsigned long long capacity, AggTable tab,
                  u64 *__restrict__ ovf, unsigned long long ovf_capacity, int k, u32 *__restrict__ bin_hist, int F)
{
  const unsigned long long n = counters[CTR_RECORDS] < capacity ? counters[CTR_RECORDS] : capacity;
  const size_t stride = (size_t) gridDim.x * 256;
  const int lane = threadIdx.x & 31;
  u32 n_new = 0;
  for (size_t i0 = (size_t) blockIdx.x * 256 + threadIdx.x; i0 < n; i0 += AGG_UNROLL * stride) {
    // the kernel lives on L2 latency (a random 16-byte read per record): the records of AGG_UNROLL rounds are loaded
    // together, then their home slots are read together, and only what did not hit its key there takes the probe loop
    Rec<W> r[AGG_UNROLL];
    u32 slot[AGG_UNROLL];
    int seen[AGG_UNROLL];
#pragma unroll
    for (int u = 0; u < AGG_UNROLL; u++) { const size_t i = i0 + u * stride; if (i < n) r[u] = load_rec<W> (recs, i); }
#pragma unroll
    for (int u = 0; u < AGG_UNROLL; u++) {
      const size_t i = i0 + u * stride;
      seen[u] = -1;
      if (i < n && r[u].w[0] != 0) { // (all zero: the unused rest of a warp's block of positions, tjb_scan.cuh)
        r[u] = canon_obs<W> (r[u], k); // the scan's fast path leaves the strand canonicalisation to this kernel
        Rec<W> key = r[u];
        key.w[0] &= ~3ULL;
        slot[u] = key_hash<W> (key) & tab.mask;
        seen[u] = table_probe<W> (tab, slot[u], key);
      }
    }
#pragma unroll
    for (int u = 0; u < AGG_UNROLL; u++) {
      if (seen[u] < 0) continue;
      if (seen[u] == 1) { table_count_one (tab, slot[u], r[u].w[0]); continue; } // the common case at depth
      Rec<W> key = r[u];
      key.w[0] &= ~3ULL;
      const int res = seen[u] == 2 ? table_add_from<W> (tab, key, flag_weight (r[u].w[0]), (slot[u] + 1) & tab.mask, -1)
                                   : table_add_from<W> (tab, key, flag_weight (r[u].w[0]), slot[u], 0);
      if (res != TAB_ADDED) atomicAdd (&bin_hist[rec_bin<W> (key, k, F)], 1u); // a new entry, or a record of the overflow list
      n_new += res == TAB_NEW;
      if (res == TAB_DEFER) {
        const unsigned long long pos = atomicAdd (&counters[CTR_OVERFLOW], 1ULL);
        if (pos < ovf_capacity) store_rec<W> (ovf, pos, r[u]);
      }
    }
  }
#pragma unroll
  for (int d = 16; d; d >>= 1) n_new += __shfl_xor_sync (0xffffffffu, n_new, d);
  if (lane == 0 && n_new) atomicAdd (&counters[CTR_ENTRIES], (unsigned long long) n_new);
}

// --------------------------------------------------------------------------------------- partition -------
// Walks the aggregation table (and the overflow list) and moves every entry to its bucket as a weighted record.
// Also sums the weights: the number of tract observations of the sample (hc->n_elem before finalise_hopo_counter).
template <int W>
__global__ void __launch_bounds__ (256)
wpartition_kernel (AggTable tab, const u64 *__restrict__ ovf, unsigned long long *__restrict__ counters, unsigned long long ovf_capacity,
                   int k, int F, const u32 *__restrict__ lut, u32 *__restrict__ cursor, u64 *__restrict__ wkey,
                   unsigned long long *__restrict__ wcnt, unsigned long long list_cap)
{
  const size_t stride = (size_t) gridDim.x * 256, n_slots = (size_t) tab.mask + 1;
  unsigned long long n_obs = 0;
  for (size_t s0 = (size_t) blockIdx.x * 256 + threadIdx.x; s0 < n_slots; s0 += PART_UNROLL * stride) {
    unsigned long long c[PART_UNROLL];
#pragma unroll
    for (int u = 0; u < PART_UNROLL; u++) { const size_t sl = s0 + u * stride; c[u] = sl < n_slots ? tab.cnt[sl] : 0ULL; }
    // every step of an entry (key, bucket of its bin, cursor, stores) waits for the one before it: the steps of the
    // PART_UNROLL entries of a thread are issued together, not entry after entry
    Rec<W> key[PART_UNROLL];
    u32 bkt[PART_UNROLL], pos[PART_UNROLL];
#pragma unroll
    for (int u = 0; u < PART_UNROLL; u++) if (c[u]) key[u] = table_key_at<W> (tab, (u32) (s0 + u * stride));
#pragma unroll
    for (int u = 0; u < PART_UNROLL; u++) if (c[u]) bkt[u] = lut[rec_bin<W> (key[u], k, F)];
#pragma unroll
    for (int u = 0; u < PART_UNROLL; u++) if (c[u]) pos[u] = atomicAdd (&cursor[(size_t) bkt[u] * BK_CURSOR_STRIDE], 1u);
#pragma unroll
    for (int u = 0; u < PART_UNROLL; u++) {
      if (!c[u]) continue;
      if (pos[u] < list_cap) { store_rec<W> (wkey, pos[u], key[u]); wcnt[pos[u]] = c[u]; }
      n_obs += (u32) c[u] + ((u32) (c[u] >> 32) & 0x7fffffffu);
    }
  }
  const unsigned long long n_ovf = counters[CTR_OVERFLOW] < ovf_capacity ? counters[CTR_OVERFLOW] : ovf_capacity;
  for (size_t i = (size_t) blockIdx.x * 256 + threadIdx.x; i < n_ovf; i += stride) {
    Rec<W> r = load_rec<W> (ovf, i);
    const unsigned long long c = flag_weight (r.w[0]);
    r.w[0] &= ~3ULL;
    const u32 pos = atomicAdd (&cursor[(size_t) lut[rec_bin<W> (r, k, F)] * BK_CURSOR_STRIDE], 1u);
    if (pos < list_cap) { store_rec<W> (wkey, pos, r); wcnt[pos] = c; }
    n_obs++;
  }
  // one atomic per CTA on the sample's counter (9 K warps adding to one address took longer than the walk itself)
  __shared__ unsigned long long s_obs[256 / 32];
#pragma unroll
  for (int d = 16; d; d >>= 1) n_obs += __shfl_xor_sync (0xffffffffu, n_obs, d);
  if ((threadIdx.x & 31) == 0) s_obs[threadIdx.x >> 5] = n_obs;
  __syncthreads ();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
#pragma unroll
    for (int w = 0; w < 256 / 32; w++) t += s_obs[w];
    if (t) atomicAdd (&counters[CTR_NRAW], t);
  }
}

// ---------------------------------------------------------------------------------- bucket reduce --------
template <int W> constexpr size_t wbucket_smem_bytes ()
{
  return (size_t) W * BK_CAP * 8 /* s_w */ + 2 * BK_CAP * 2 /* perm */ + BK_WARPS * 256 * 2 /* counts */ + BK_CAP * 2 /* aux */ +
         BK_CAP * 4 /* val */ + BK_CAP * 4 /* kept counts */ + BK_CAP * 4 /* sum */;
}
#define BK_RANKSORT_MAX BK_THREADS // up to this many records a bucket is ordered by comparison counting instead of radix passes
enum { QC_SEEN_TWICE = 0,   // "-b": keep count > 1                 (hopo_counter.c:371-373)
       QC_BOTH_STRANDS = 1, // default: keep canon_flag == 3        (:368-370)
       QC_KEEP_ALL = 2 };   // union of mapper queries (tjb_union.cuh): nothing is filtered

// a > b in the reference's order (base, ctx0, ctx1, length), hopo_counter.c:28-38; the flag bits do not take part
template <int W> __device__ __forceinline__ bool rec_greater (const Rec<W> &a, const Rec<W> &b);
template <> __device__ __forceinline__ bool rec_greater<2> (const Rec<2> &a, const Rec<2> &b)
{ return a.w[1] != b.w[1] ? a.w[1] > b.w[1] : (a.w[0] >> 2) > (b.w[0] >> 2); }
template <> __device__ __forceinline__ bool rec_greater<3> (const Rec<3> &a, const Rec<3> &b)
{
  const u64 ba = (a.w[0] >> 12) & 1, bb = (b.w[0] >> 12) & 1;
  if (ba != bb) return ba > bb;
  if (a.w[2] != b.w[2]) return a.w[2] > b.w[2];
  if (a.w[1] != b.w[1]) return a.w[1] > b.w[1];
  return ((a.w[0] >> 2) & 0x3ff) > ((b.w[0] >> 2) & 0x3ff);
}

template <int W>
__global__ void __launch_bounds__ (BK_THREADS, W == 2 ? BK_CTAS2 : 2)
wbucket_reduce_kernel (const u64 *__restrict__ wkey, const unsigned long long *__restrict__ wcnt, const u32 *__restrict__ bkt_start,
                       u32 *__restrict__ meta, unsigned long long capacity, DigitList dl, int qc_mode, int min_coverage,
                       u64 *__restrict__ out_rec, u32 *__restrict__ out_cnt, u32 *__restrict__ out_idx_i, u32 *__restrict__ out_idx_f,
                       uint4 *__restrict__ bstat)
{
  extern __shared__ __align__ (16) unsigned char smem[];
  u64 *s_w = reinterpret_cast<u64*> (smem);                                  // [W][BK_CAP]
  const u32 *s_w32 = reinterpret_cast<const u32*> (smem);
  unsigned short *s_perm0 = reinterpret_cast<unsigned short*> (s_w + W * BK_CAP);
  unsigned short *s_perm1 = s_perm0 + BK_CAP;
  unsigned short *s_cnt = s_perm1 + BK_CAP;                                  // [BK_WARPS][256] radix counters
  unsigned short *s_aux = s_cnt + BK_WARPS * 256;                            // [BK_CAP]
  u32 *s_val = reinterpret_cast<u32*> (s_aux + BK_CAP);                      // [BK_CAP] kept j -> descending position | flag << 28
  u32 *s_kcnt = s_val + BK_CAP;                                              // [BK_CAP] kept j -> count (signed 20-bit field value)
  u32 *s_sum = s_kcnt + BK_CAP;                                              // [BK_CAP] exclusive prefix sums of the kept counts
  __shared__ u32 s_warp[2 * BK_WARPS];
  int scan_phase = 0;
  __shared__ u32 s_bucket;
  __shared__ u32 s_diff[2 * W];

  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const u32 lt_mask = (1u << lane) - 1;

  // ticket of the next bucket + the per-bucket words; published by the barrier that ends a round (one barrier serves
  // both the reuse of the shared arrays and the hand-over)
  auto next_bucket = [&] () {
    if (tid == 0) s_bucket = atomicAdd (&meta[META_TICKET], 1u);
    if (tid < 2 * W) s_diff[tid] = 0;
  };
  const u32 n_buckets = meta[META_NBUCKETS];
  next_bucket ();
  __syncthreads ();
  for (;;) {
    const u32 b = s_bucket;
    if (b >= n_buckets) break;
    unsigned long long st = bkt_start[b], en = bkt_start[b + 1];
    if (st > capacity) st = capacity;
    if (en > capacity) en = capacity;
    const u32 start = (u32) st, n = (u32) (en - st);
    if (n > BK_CAP) { // reduced by the global multi-pass path (host reads the flag after the synchronisation)
      if (tid == 0) { bstat[b] = make_uint4 (0, 0, 0, 1); atomicAdd (&meta[META_NOVERSIZE], 1u); }
      __syncthreads (); // every thread has read s_bucket
      next_bucket ();
      __syncthreads ();
      continue;
    }
    // ---- load the bucket; remember which key bits differ at all ------------------------------------------
    {
      u64 d[W];
#pragma unroll
      for (int j = 0; j < W; j++) d[j] = 0;
      Rec<W> f;
      if (n) f = load_rec<W> (wkey, start);
      for (u32 i = tid; i < n; i += BK_THREADS) {
        Rec<W> r = load_rec<W> (wkey, (size_t) start + i);
#pragma unroll
        for (int j = 0; j < W; j++) { s_w[j * BK_CAP + i] = r.w[j]; d[j] |= r.w[j] ^ f.w[j]; }
        s_perm0[i] = (unsigned short) i;
        s_val[i] = 0; // rank counters of the comparison sort (free until the QC step)
      }
#pragma unroll
      for (int j = 0; j < W; j++) {
        u32 lo = __reduce_or_sync (0xffffffffu, (u32) d[j]), hi = __reduce_or_sync (0xffffffffu, (u32) (d[j] >> 32));
        if (lane == 0) { if (lo) atomicOr (&s_diff[2 * j], lo); if (hi) atomicOr (&s_diff[2 * j + 1], hi); }
      }
    }
    __syncthreads ();
    // ---- sort the records, ascending in `cur` (equal keys end up next to each other) ---------------------------
    unsigned short *cur = s_perm0, *nxt = s_perm1;
    if (n <= BK_RANKSORT_MAX) {
      // few records (the usual case at high depth): rank by comparison, one record per thread, broadcast reads.  The
      // comparison is a 128-bit subtraction whose final borrow says "y < x" (W = 2: the flag bits of a weighted key are
      // zero, so (w1, w0) as one number is the reference's order); equal keys (a key that arrived in two records) get
      // consecutive ranks from a counter at their common rank, in any order: their weights are added below
      if ((u32) tid < n) {
        u32 less = 0;
        if (W == 2) {
          const u64 x0 = s_w[tid], x1 = s_w[BK_CAP + tid];
          const u32 x0l = (u32) x0, x0h = (u32) (x0 >> 32), x1l = (u32) x1, x1h = (u32) (x1 >> 32);
#pragma unroll 4
          for (u32 o = 0; o < n; o++) {
            const u64 y0 = s_w[o], y1 = s_w[BK_CAP + o];
            u32 borrow;
            asm ("{\n .reg .u32 t;\n sub.cc.u32 t, %1, %5;\n subc.cc.u32 t, %2, %6;\n subc.cc.u32 t, %3, %7;\n subc.cc.u32 t, %4, %8;\n subc.u32 %0, 0, 0;\n}"
                 : "=r"(borrow) : "r"((u32) y0), "r"((u32) (y0 >> 32)), "r"((u32) y1), "r"((u32) (y1 >> 32)), "r"(x0l), "r"(x0h), "r"(x1l), "r"(x1h));
            less -= borrow; // 0xffffffff when y < x
          }
        } else {
          Rec<W> x;
#pragma unroll
          for (int j = 0; j < W; j++) x.w[j] = s_w[j * BK_CAP + tid];
          for (u32 o = 0; o < n; o++) {
            Rec<W> y;
#pragma unroll
            for (int j = 0; j < W; j++) y.w[j] = s_w[j * BK_CAP + o];
            less += rec_greater<W> (x, y);
          }
        }
        nxt[less + atomicAdd (&s_val[less], 1u)] = (unsigned short) tid; // s_val[0 .. n) was zeroed while the bucket was loaded
      }
      __syncthreads ();
      unsigned short *t = cur; cur = nxt; nxt = t;
    } else {
      // LSD radix sort of the index permutation; digits on which the whole bucket agrees are skipped
      const u32 ipt = (n + BK_THREADS - 1) / BK_THREADS; // items per thread in a pass
      for (int q = 0; q < dl.n; q++) {
        const int word = dl.word[q], shift = dl.shift[q];
        const u32 mask = (1u << dl.nbits[q]) - 1;
        const int half = shift >> 5, sh = shift & 31;
        if (((s_diff[2 * word + half] >> sh) & mask) == 0) continue;
        for (int i = tid; i < BK_WARPS * 256 / 2; i += BK_THREADS) reinterpret_cast<u32*> (s_cnt)[i] = 0;
        __syncthreads ();
#pragma unroll 1
        for (u32 r = 0; r < ipt; r++) { // warp w owns items [w * 32 * ipt, (w + 1) * 32 * ipt): stable
          const u32 i = (u32) wid * (32 * ipt) + r * 32 + lane;
          u32 d = 256;
          if (i < n) d = (s_w32[2 * ((size_t) word * BK_CAP + cur[i]) + half] >> sh) & mask;
          const u32 peers = __match_any_sync (0xffffffffu, d);
          const int leader = __ffs (peers) - 1;
          u32 old = 0;
          if (lane == leader && d < 256) { old = s_cnt[wid * 256 + d]; s_cnt[wid * 256 + d] = (unsigned short) (old + __popc (peers)); }
          old = __shfl_sync (0xffffffffu, old, leader);
          if (i < n) s_aux[i] = (unsigned short) (old + __popc (peers & lt_mask)); // rank among the warp's equal digits
          __syncwarp ();
        }
        __syncthreads ();
        { // thread t owns digit t: offsets in (digit-major, warp-minor) order
          u32 c[BK_WARPS], sum = 0;
#pragma unroll
          for (int w = 0; w < BK_WARPS; w++) { c[w] = s_cnt[w * 256 + tid]; sum += c[w]; }
          u32 tot;
          u32 ex = blk_excl_scan<BK_THREADS> (sum, s_warp, scan_phase, &tot);
#pragma unroll
          for (int w = 0; w < BK_WARPS; w++) { s_cnt[w * 256 + tid] = (unsigned short) ex; ex += c[w]; }
        }
        __syncthreads ();
#pragma unroll 1
        for (u32 r = 0; r < ipt; r++) {
          const u32 i = (u32) wid * (32 * ipt) + r * 32 + lane;
          if (i < n) {
            const u32 id = cur[i], d = (s_w32[2 * ((size_t) word * BK_CAP + id) + half] >> sh) & mask;
            nxt[s_cnt[wid * 256 + d] + s_aux[i]] = (unsigned short) id;
          }
        }
        __syncthreads ();
        unsigned short *t = cur; cur = nxt; nxt = t;
      }
    }
    // record at DESCENDING position p
    auto id_at = [&] (u32 p) -> u32 { return cur[n - 1 - p]; };
    auto rec_at = [&] (u32 p) {
      const u32 id = id_at (p);
      Rec<W> r;
#pragma unroll
      for (int j = 0; j < W; j++) r.w[j] = s_w[j * BK_CAP + id];
      return r;
    };
    u32 n_dist, n_keep, n_groups, n_idx, ctotal;
    // ---- S2: weights of equal keys added, count, canon_flag, QC filter (hopo_counter.c:355-374) -----------------
    {
      const u32 p0 = tid * BK_IPT;
      unsigned long long cw[BK_IPT];
#pragma unroll
      for (int r = 0; r < BK_IPT; r++) cw[r] = p0 + r < n ? wcnt[(size_t) start + id_at (p0 + r)] : 0ULL; // all loads in flight
      u32 hmask = 0, kmask = 0, packed[BK_IPT], kc[BK_IPT];
      if (p0 < n) {
        Rec<W> prev;
        if (p0) prev = rec_at (p0 - 1);
#pragma unroll
        for (int r = 0; r < BK_IPT; r++) {
          const u32 p = p0 + r;
          if (p < n) {
            const Rec<W> x = rec_at (p);
            if (p == 0 || !same_tract<W> (x, prev)) { // first record of its key: add the weights of the rest of the run
              u32 cf = (u32) cw[r], cr = (u32) (cw[r] >> 32) & 0x7fffffffu;
              for (u32 q = p + 1; q < n && same_tract<W> (rec_at (q), x); q++) {
                const unsigned long long c = wcnt[(size_t) start + id_at (q)];
                cf += (u32) c; cr += (u32) (c >> 32) & 0x7fffffffu;
              }
              const int cnt = sext_field ((long long) cf + (long long) cr, 20); // the reference's count:20 signed bit-field
              const u32 flag = (cf ? 1u : 0u) | (cr ? 2u : 0u);
              const bool keep = qc_mode == QC_BOTH_STRANDS ? flag == 3 : (qc_mode == QC_SEEN_TWICE ? cnt > 1 : true);
              hmask |= 1u << r;
              if (keep) { kmask |= 1u << r; packed[r] = p | (flag << 28); kc[r] = (u32) cnt; }
            }
            prev = x;
          }
        }
      }
      static_assert (BK_CAP < (1 << 16), "two 16-bit sums per scanned word");
      u32 both_total;
      const u32 both = blk_excl_scan<BK_THREADS> (((u32) __popc (hmask) << 16) | (u32) __popc (kmask), s_warp, scan_phase, &both_total);
      n_dist = both_total >> 16; n_keep = both_total & 0xffffu;
      u32 kpos = both & 0xffffu;
#pragma unroll
      for (int r = 0; r < BK_IPT; r++) if ((kmask >> r) & 1) { s_val[kpos] = packed[r]; s_kcnt[kpos] = kc[r]; kpos++; }
    }
    __syncthreads ();
    // ---- S3: context groups over the kept list + prefix sums of their counts (:388-404) -------------------
    {
      const u32 j0 = tid * BK_IPT;
      u32 gmask = 0, csum = 0, c[BK_IPT];
      if (j0 < n_keep) {
        Rec<W> prev;
        if (j0) prev = rec_at (s_val[j0 - 1] & 0xfff);
#pragma unroll
        for (int r = 0; r < BK_IPT; r++) {
          const u32 j = j0 + r;
          c[r] = 0;
          if (j < n_keep) {
            c[r] = s_kcnt[j];
            csum += c[r];
            Rec<W> x = rec_at (s_val[j] & 0xfff);
            if (j == 0 || !same_context<W> (x, prev)) gmask |= 1u << r;
            prev = x;
          }
        }
      }
      u32 gpos = blk_excl_scan<BK_THREADS> ((u32) __popc (gmask), s_warp, scan_phase, &n_groups);
      u32 cpos = blk_excl_scan<BK_THREADS> (csum, s_warp, scan_phase, &ctotal); // u32 wrap-around = the reference's int sums
      // (s_aux and s_sum, written next, were last read before the barrier that ended the sort)
#pragma unroll
      for (int r = 0; r < BK_IPT; r++) {
        const u32 j = j0 + r;
        if (j < n_keep) {
          s_sum[j] = cpos; cpos += c[r];
          if ((gmask >> r) & 1) s_aux[gpos++] = (unsigned short) j;
        }
      }
    }
    __syncthreads ();
    // ---- S4: groups whose summed count reaches min_coverage -> (idx_initial, idx_final), bucket-local -------
    {
      const u32 g0 = tid * BK_IPT;
      u32 imask = 0, js[BK_IPT], je[BK_IPT];
#pragma unroll
      for (int r = 0; r < BK_IPT; r++) {
        const u32 g = g0 + r;
        if (g < n_groups) {
          js[r] = s_aux[g];
          je[r] = (g + 1 < n_groups) ? s_aux[g + 1] : n_keep;
          const u32 hi = (je[r] < n_keep) ? s_sum[je[r]] : ctotal;
          if ((int) (hi - s_sum[js[r]]) >= min_coverage) imask |= 1u << r;
        }
      }
      u32 ipos = blk_excl_scan<BK_THREADS> (__popc (imask), s_warp, scan_phase, &n_idx);
#pragma unroll
      for (int r = 0; r < BK_IPT; r++)
        if ((imask >> r) & 1) { out_idx_i[start + ipos] = js[r]; out_idx_f[start + ipos] = je[r]; ipos++; }
    }
    // ---- S5: kept records, compact, at the bucket's own offset ----------------------------------------------
    for (u32 j = tid; j < n_keep; j += BK_THREADS) {
      const u32 v = s_val[j];
      Rec<W> x = rec_at (v & 0xfff);
      x.w[0] = (x.w[0] & ~3ULL) | (u64) (v >> 28);
      store_rec<W> (out_rec, (size_t) start + j, x);
      out_cnt[start + j] = s_kcnt[j];
    }
    if (tid == 0) bstat[b] = make_uint4 (n_dist, n_keep, n_idx, 0);
    next_bucket ();
    __syncthreads ();
  }
}

// ------------------------------------------------------------------------------------------ totals -------
__global__ void __launch_bounds__ (PLAN_THREADS)
totals_kernel (const uint4 *__restrict__ bstat, u32 *__restrict__ meta, u32 *__restrict__ koff, u32 *__restrict__ ioff)
{
  __shared__ u32 s_warp[2 * (PLAN_THREADS / 32)];
  int scan_phase = 0;
  const u32 nb = meta[META_NBUCKETS];
  const u32 ipt = (nb + PLAN_THREADS - 1) / PLAN_THREADS;
  const u32 b0 = threadIdx.x * ipt, b1 = (b0 + ipt < nb) ? b0 + ipt : nb;
  u32 sd = 0, sk = 0, si = 0;
  for (u32 b = b0; b < b1; b++) { uint4 s = bstat[b]; sd += s.x; sk += s.y; si += s.z; }
  u32 td, tk, ti;
  blk_excl_scan<PLAN_THREADS> (sd, s_warp, scan_phase, &td);
  u32 ek = blk_excl_scan<PLAN_THREADS> (sk, s_warp, scan_phase, &tk);
  u32 ei = blk_excl_scan<PLAN_THREADS> (si, s_warp, scan_phase, &ti);
  for (u32 b = b0; b < b1; b++) { uint4 s = bstat[b]; koff[b] = ek; ioff[b] = ei; ek += s.y; ei += s.z; }
  if (threadIdx.x == 0) { meta[META_NDIST] = td; meta[META_NKEEP] = tk; meta[META_NIDX] = ti; }
}

// ------------------------------------------------------------------------------- coverage hash + pack ----
// coverage table: slot key = (flank & 0x7fffffff) + 1, 0 = empty (one memset clears keys and sums).  Returns the key's
// sum after this weight: with positive weights the largest value any insert returns is the largest final sum.
__device__ __forceinline__ int cov_insert (u32 *__restrict__ keys, u32 *__restrict__ vals, u32 mask, u32 key, u32 w)
{
  u32 slot = (key * 0x9E3779B1u) >> 7;
  key += 1;
  for (;;) {
    slot &= mask;
    u32 prev = keys[slot];
    if (prev != key) {
      if (prev != 0) { slot++; continue; }
      prev = atomicCAS (&keys[slot], 0u, key);
      if (prev != 0 && prev != key) { slot++; continue; }
    }
    return (int) (atomicAdd (&vals[slot], w) + w);
  }
}

// bucket that holds global output position g: the LAST b with off[b] <= g (off = exclusive prefix sums of the per-bucket
// counts: buckets that contribute nothing share their offset with the next one)
__device__ __forceinline__ u32 bucket_of (const u32 *__restrict__ off, u32 nb, u32 g)
{
  u32 lo = 0, hi = nb - 1;
  while (lo < hi) { const u32 mid = (lo + hi + 1) >> 1; if (off[mid] <= g) lo = mid; else hi = mid - 1; }
  return lo;
}

// One thread per kept record: final 40-byte hopo_element (src/hopo_counter.h:34-49) as it stands at the BWA boundary,
// and the two (flank & 0x7fffffff, count) pairs of the record (hopo_counter.c:427-432); one thread per context group:
// global idx_initial / idx_final.  out40 may be device memory or mapped pinned host memory (the result arena): rows are
// staged per warp in shared memory so that every store instruction of the warp covers 256 contiguous bytes.
template <int W>
__global__ void __launch_bounds__ (256)
pack_kernel (const u64 *__restrict__ out_rec, const u32 *__restrict__ out_cnt, const u32 *__restrict__ out_idx_i, const u32 *__restrict__ out_idx_f,
             const u32 *__restrict__ bkt_start, const uint4 *__restrict__ bstat, const u32 *__restrict__ koff, const u32 *__restrict__ ioff,
             const u32 *__restrict__ meta, int k, u64 *__restrict__ out40, int *__restrict__ idx_initial, int *__restrict__ idx_final,
             u32 *__restrict__ cov_keys, u32 *__restrict__ cov_vals, u32 cov_mask, u32 max_keep, u32 max_idx, unsigned long long capacity,
             int *__restrict__ cov, u32 *__restrict__ meta_out)
{
  // coverage (hopo_counter.c:434: the largest summed weight of any 31-bit flank key): every insert returns its key's
  // running sum and the largest one seen is kept; exact as long as every weight is positive, which the QC filter
  // guarantees unless a count wrapped its signed 20-bit field (then META_COVSCAN asks the host for cov_max_kernel)
  int best = INT_MIN;
  bool nonpositive = false;
  // The host launches this kernel before it knows the counts, into arrays sized from the previous sample (max_keep,
  // max_idx; the coverage table holds 2 max_keep keys at load <= 0.5): nothing is written beyond them, and the host
  // repeats the launch with exact sizes when the counts turn out larger (or the sample had to be redone).
  __shared__ u64 s_stage[256 / 32][32 * 5];
  const u32 nb = meta[META_NBUCKETS];
  if (!nb) return;
  const int lane = threadIdx.x & 31;
  u64 *stage = s_stage[threadIdx.x >> 5];
  const u32 n_keep = meta[META_NKEEP] < max_keep ? meta[META_NKEEP] : max_keep;
  const u32 n_idx = meta[META_NIDX] < max_idx ? meta[META_NIDX] : max_idx;
  const u32 stride = gridDim.x * 256;
  for (u32 g0 = (blockIdx.x * 256 + (threadIdx.x & ~31u)); g0 < n_keep; g0 += stride) { // g0: first record of this warp's 32
    const u32 g = g0 + lane;
    const u32 rows = n_keep - g0 < 32u ? n_keep - g0 : 32u;
    if ((u32) lane < rows) {
      const u32 b = bucket_of (koff, nb, g);
      const u32 j = g - koff[b], start = bkt_start[b];
      u64 c0 = 0, c1 = 0, bits = 0;
      if ((unsigned long long) start + j < capacity) {
        Rec<W> r = load_rec<W> (out_rec, (size_t) start + j);
        const int cnt = sext_field ((long long) out_cnt[start + j], 20); // the reference's count:20 signed bit-field
        // gcc x86-64 bit-field order from the LSB: base:2 length:10 count:20 mismatches:12 multi:3 neg_strand:2 canon_flag:3
        bits = ((u64) rec_base<W> (r, k) & 3) | (((u64) rec_len_field<W> (r) & 0x3ff) << 2) | (((u64) cnt & 0xfffff) << 12) |
               (0xffeULL << 32) | (((u64) rec_flag<W> (r) & 7) << 49);
        c0 = rec_ctx0<W> (r, k); c1 = rec_ctx1<W> (r, k);
        const int s0 = cov_insert (cov_keys, cov_vals, cov_mask, (u32) (c0 & 0x7fffffffULL), (u32) cnt);
        const int s1 = cov_insert (cov_keys, cov_vals, cov_mask, (u32) (c1 & 0x7fffffffULL), (u32) cnt);
        best = max (best, max (s0, s1));
        nonpositive = nonpositive || cnt <= 0;
      }
      stage[5 * lane] = c0; stage[5 * lane + 1] = c1; stage[5 * lane + 2] = bits;
      stage[5 * lane + 3] = ~0ULL; // read_offset = -1 (hopo_counter.c:511), loc_ref_id = -1 (:297)
      stage[5 * lane + 4] = ~0ULL; // loc_pos = loc_last = -1
    }
    __syncwarp ();
    u64 *o = out40 + 5 * (size_t) g0;
#pragma unroll
    for (int t = 0; t < 5; t++) { const u32 w = t * 32 + lane; if (w < 5 * rows) o[w] = stage[w]; }
    __syncwarp ();
  }
  best = __reduce_max_sync (0xffffffffu, best);
  if (lane == 0 && best != INT_MIN) atomicMax (cov, best);
  if (__any_sync (0xffffffffu, nonpositive) && lane == 0) meta_out[META_COVSCAN] = 1;
  for (u32 t = blockIdx.x * 256 + threadIdx.x; t < n_idx; t += stride) {
    const u32 b = bucket_of (ioff, nb, t);
    const u32 start = bkt_start[b], ko = koff[b], u = t - ioff[b];
    if ((unsigned long long) start + u < capacity) {
      idx_initial[t] = (int) (ko + out_idx_i[start + u]);
      idx_final[t] = (int) (ko + out_idx_f[start + u]);
    }
  }
}

// coverage = largest summed weight of any key (empirical_frequency.c:254-277 sorted by frequency, hopo_counter.c:434):
// the scan of the whole table, only needed when a weight was not positive (META_COVSCAN)
__global__ void __launch_bounds__ (256)
cov_max_kernel (const u32 *__restrict__ keys, const u32 *__restrict__ vals, u32 n_slots, int *__restrict__ cov)
{
  int best = INT_MIN;
  for (u32 i = blockIdx.x * 256 + threadIdx.x; i < n_slots; i += gridDim.x * 256)
    if (keys[i] != 0) { int v = (int) vals[i]; best = v > best ? v : best; }
#pragma unroll
  for (int d = 16; d; d >>= 1) { int o = __shfl_xor_sync (0xffffffffu, best, d); best = o > best ? o : best; }
  if ((threadIdx.x & 31) == 0 && best != INT_MIN) atomicMax (cov, best);
}
