// Shared definitions for the sm_100a kernels of libtjb_gpu.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long u64;
typedef unsigned int u32;

// ---------------------------------------------------------------------------------------------------------
// Tract-observation record ("raw element"): W 64-bit words, one per homopolymer tract seen in a read.
// It carries exactly the fields the reference orders and reduces on (src/hopo_counter.c:28-38, 355-365):
//   flag (2 bits, canon_flag 1|2), length (10-bit field value, biased by +512 so that unsigned order equals
//   the reference's signed order), ctx1 (right flank, 2k bits), ctx0 (left flank, 2k bits), base (1 bit).
// Significance, low -> high:  flag < length < ctx1 < ctx0 < base.
//
//   W = 2 (k <= 26):  w0 = flag | len<<2 | ctx1<<12          w1 = ctx0 | base<<2k
//   W = 3 (k >= 27):  w0 = flag | len<<2 | base<<12          w1 = ctx1          w2 = ctx0
// ---------------------------------------------------------------------------------------------------------
#define TJB_LEN_BIAS 512

__host__ __device__ inline int tjb_rec_words (int k) { return k <= 26 ? 2 : 3; }

template <int W> struct Rec { u64 w[W]; };

template <int W> __device__ __forceinline__ Rec<W> pack_rec (u64 ctx0, u64 ctx1, int base, int len_field, int flag, int k);
template <> __device__ __forceinline__ Rec<2> pack_rec<2> (u64 ctx0, u64 ctx1, int base, int len_field, int flag, int k)
{
  Rec<2> r;
  r.w[0] = (u64) flag | ((u64) (len_field + TJB_LEN_BIAS) << 2) | (ctx1 << 12);
  r.w[1] = ctx0 | ((u64) base << (2 * k));
  return r;
}
template <> __device__ __forceinline__ Rec<3> pack_rec<3> (u64 ctx0, u64 ctx1, int base, int len_field, int flag, int k)
{
  (void) k;
  Rec<3> r;
  r.w[0] = (u64) flag | ((u64) (len_field + TJB_LEN_BIAS) << 2) | ((u64) base << 12);
  r.w[1] = ctx1;
  r.w[2] = ctx0;
  return r;
}

template <int W> __device__ __forceinline__ int rec_flag (const Rec<W> &r) { return (int) (r.w[0] & 3); }
template <int W> __device__ __forceinline__ int rec_len_field (const Rec<W> &r) { return (int) ((r.w[0] >> 2) & 0x3ff) - TJB_LEN_BIAS; }
template <int W> __device__ __forceinline__ int rec_base (const Rec<W> &r, int k);
template <> __device__ __forceinline__ int rec_base<2> (const Rec<2> &r, int k) { return (int) ((r.w[1] >> (2 * k)) & 1); }
template <> __device__ __forceinline__ int rec_base<3> (const Rec<3> &r, int k) { (void) k; return (int) ((r.w[0] >> 12) & 1); }
template <int W> __device__ __forceinline__ u64 rec_ctx0 (const Rec<W> &r, int k);
template <> __device__ __forceinline__ u64 rec_ctx0<2> (const Rec<2> &r, int k) { return r.w[1] & ((1ULL << (2 * k)) - 1); }
template <> __device__ __forceinline__ u64 rec_ctx0<3> (const Rec<3> &r, int k) { (void) k; return r.w[2]; }
template <int W> __device__ __forceinline__ u64 rec_ctx1 (const Rec<W> &r, int k);
template <> __device__ __forceinline__ u64 rec_ctx1<2> (const Rec<2> &r, int k) { (void) k; return r.w[0] >> 12; }
template <> __device__ __forceinline__ u64 rec_ctx1<3> (const Rec<3> &r, int k) { (void) k; return r.w[1]; }

// same (base, ctx0, ctx1, length)?  (flag bits excluded)
template <int W> __device__ __forceinline__ bool same_tract (const Rec<W> &a, const Rec<W> &b)
{
  bool eq = (a.w[0] >> 2) == (b.w[0] >> 2);
#pragma unroll
  for (int i = 1; i < W; i++) eq = eq && (a.w[i] == b.w[i]);
  return eq;
}
// same (base, ctx0, ctx1)?  = compare_hopo_context() == 0 (src/hopo_counter.c:48-58)
template <int W> __device__ __forceinline__ bool same_context (const Rec<W> &a, const Rec<W> &b)
{
  bool eq = (a.w[0] >> 12) == (b.w[0] >> 12);
#pragma unroll
  for (int i = 1; i < W; i++) eq = eq && (a.w[i] == b.w[i]);
  return eq;
}

template <int W> __device__ __forceinline__ Rec<W> load_rec (const u64 *p, size_t i)
{
  Rec<W> r;
  if (W == 2) {
    ulonglong2 v = *reinterpret_cast<const ulonglong2*> (p + 2 * i);
    r.w[0] = v.x; r.w[1] = v.y;
  } else {
#pragma unroll
    for (int j = 0; j < W; j++) r.w[j] = p[(size_t) W * i + j];
  }
  return r;
}
template <int W> __device__ __forceinline__ void store_rec (u64 *p, size_t i, const Rec<W> &r)
{
  if (W == 2) {
    *reinterpret_cast<ulonglong2*> (p + 2 * i) = make_ulonglong2 (r.w[0], r.w[1]);
  } else {
#pragma unroll
    for (int j = 0; j < W; j++) p[(size_t) W * i + j] = r.w[j];
  }
}

// ---------------------------------------------------------------------------------------------------------
// Raw tract observation: what the scan kernel appends for a tract found on its fast path.  Both flanks as the scan
// has them (k GRAY codes each, A0 C1 T2 G3, base j in bits 2j..2j+1, in read orientation), the tract's gray code and
// its length; bit 63 of one word says "raw".  Strand canonicalisation (hopo_counter.c:233-246) is left to the kernel
// that counts the observations (aggregate_kernel has idle issue slots, the scan has none): canon_obs() turns a raw
// observation into the record above; records the scan's slow path wrote in their final form pass through.
//   W = 2:  w0 = left flank | gray << 52 | length << 54       w1 = right flank | 1 << 63
//   W = 3:  w0 = gray | length << 2 | 1 << 63                 w1 = left flank     w2 = right flank
// ---------------------------------------------------------------------------------------------------------
#ifndef SCAN_RAW_OBS
#define SCAN_RAW_OBS 1
#endif
#define TJB_OBS_RAW (1ULL << 63)

template <int W> __device__ __forceinline__ Rec<W> pack_obs (u64 left, u64 right, int gray, int len)
{
  Rec<W> r;
  if (W == 2) { r.w[0] = left | ((u64) gray << 52) | ((u64) len << 54); r.w[1] = right | TJB_OBS_RAW; }
  else { r.w[0] = (u64) gray | ((u64) len << 2) | TJB_OBS_RAW; r.w[1] = left; r.w[W - 1] = right; }
  return r;
}

// reverse complement of a k-base word of GRAY codes (A0 C1 T2 G3, base j in bits 2j..2j+1), as BINARY codes
// (A0 C1 G2 T3).  The bit reversal also swaps the two bits of every base: a base (h, l) arrives as (l, h).  Its
// binary code is (h, h ^ l), the complement of that (~h, ~(h ^ l)): the high bit from the swapped low bit, the low
// bit from the XOR of both.
__device__ __forceinline__ u64 revcomp_gray (u64 g, int k, u64 kmask)
{
  const u64 y = __brevll (g) >> (64 - 2 * k); // k >= 2
  const u64 m5 = 0x5555555555555555ULL;
  return ~(((y & m5) << 1) | ((y ^ (y >> 1)) & m5)) & kmask;
}

// flanks in gray codes + tract code -> the canonical record: A / C tracts keep the read's strand (gray -> binary
// codes, hopo_counter.c:233-238), T / G tracts are reverse-complemented (:239-246)
template <int W> __device__ __forceinline__ Rec<W> canon_from_gray (u64 gl64, u64 gr64, int gray, int len, int k)
{
  const u64 kmask = (k == 32) ? ~0ULL : ((1ULL << (2 * k)) - 1);
  const u64 m5 = 0x5555555555555555ULL;
  u64 ctx0, ctx1;
  if (gray < 2) { ctx0 = gl64 ^ ((gl64 >> 1) & m5); ctx1 = gr64 ^ ((gr64 >> 1) & m5); }
  else { ctx0 = revcomp_gray (gr64, k, kmask); ctx1 = revcomp_gray (gl64, k, kmask); }
  return pack_rec<W> (ctx0, ctx1, gray & 1, len, gray < 2 ? 1 : 2, k);
}
template <int W> __device__ __forceinline__ Rec<W> canon_obs (const Rec<W> &r, int k)
{
  if (W == 2) {
    if (!(r.w[1] & TJB_OBS_RAW)) return r;
    return canon_from_gray<W> (r.w[0] & ((1ULL << 52) - 1), r.w[1] & ~TJB_OBS_RAW, (int) (r.w[0] >> 52) & 3, (int) (r.w[0] >> 54), k);
  }
  if (!(r.w[0] & TJB_OBS_RAW)) return r;
  return canon_from_gray<W> (r.w[1], r.w[W - 1], (int) r.w[0] & 3, (int) (r.w[0] >> 2) & 0x3ff, k);
}

// value of a signed bit-field of `bits` bits holding v (gcc semantics for the reference's int64_t bit-fields)
__host__ __device__ __forceinline__ int sext_field (long long v, int bits)
{
  long long m = 1LL << (bits - 1);
  v &= (1LL << bits) - 1;
  return (int) ((v ^ m) - m);
}

// device counters of one sample (unsigned long long each)
enum { CTR_OVERFLOW = 0, // observations deferred to the overflow list (tjb_table.cuh)
       CTR_SKIP = 1,     // fenced runs (non-ACGTU run that is the first tract of its read)
       CTR_ENTRIES = 2,  // entries of the aggregation table
       CTR_NRAW = 3,     // tract observations = sum of all weights (hc->n_elem before finalise_hopo_counter)
       CTR_RECORDS = 4,  // records the scan appended to the record array (= observations, unless the array overflowed)
       CTR_WORDS = 6 };

// one digit of an LSD radix pass: bits [shift, shift+nbits) of word `word`
struct DigitDesc { int word, shift, nbits; };
