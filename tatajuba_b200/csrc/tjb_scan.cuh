// Read scan: homopolymer-tract detection + flanking-context key packing + strand canonicalisation.
//
// Replaces update_hopo_counter_from_seq() + add_kmer_to_hopo_counter() (reference src/hopo_counter.c:219-258,
// 285-307), which walk one read at a time, byte by byte, on one CPU thread.  Here a whole chunk of reads
// ("SEQ\nSEQ\n...", '\n' = read separator, written by csrc/tjb_fastq.c) is scanned by the full grid:
//
//   phase A  every thread moves 16 bytes (one LDG.128) into shared memory; 16-bit "equal to previous byte"
//            masks are computed per 16-byte group with byte-SIMD arithmetic; tract starts are the bit pattern
//            0 1 1 .. 1 (min(max(m,2),4) - 1 ones) in that mask and are appended to a per-tile candidate list;
//   phase B  candidates are processed one per lane: run end from the mask words, length test, both flanks
//            checked for a read separator (= the reference's s >= k and e <= L-k-1 rules, :229, :232), 2-bit
//            packing of the 2k flank bytes (reverse-complemented for G/T/U tracts, :239-246), record staged in
//            shared memory; each batch is appended to the global record array with ONE atomicAdd per tile.
//
// Tiles own the tracts that START inside them; tracts or flanks reaching beyond the staged halo fall back to
// global-memory byte reads (rare: needs a run longer than ~150 bytes at the tile edge).
#pragma once
#include "tjb_common.cuh"

#define SCAN_THREADS 512
#define SCAN_TILE (SCAN_THREADS * 16) // bytes owned by one tile
#define SCAN_HALO_L 64                // >= k + 1, multiple of 16
#define SCAN_HALO_R 192
#define SCAN_SM_BYTES (SCAN_HALO_L + SCAN_TILE + SCAN_HALO_R)
#define SCAN_GROUPS (SCAN_SM_BYTES / 16)

// forward code table: A0 C1 G2 T/U3 (either case), 4 = any other byte, 5 = '\n' (read separator)
__device__ __forceinline__ int tjb_code_fwd (int c)
{
  int u = c & 0xDF;
  int code = 4;
  if (u == 'A') code = 0; else if (u == 'C') code = 1; else if (u == 'G') code = 2; else if (u == 'T' || u == 'U') code = 3;
  if (c == '\n') code = 5;
  return code;
}

// 0x80 in every byte of x that is zero (exact, no cross-byte borrow)
__device__ __forceinline__ u32 zero_bytes (u32 x)
{
  u32 t = (x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;
  return ~(t | x | 0x7F7F7F7Fu);
}
// bits 7,15,23,31 -> bits 0..3
__device__ __forceinline__ u32 gather_msb4 (u32 m) { return (((m >> 7) & 0x01010101u) * 0x01020408u) >> 24; }

// "equal to previous byte" mask of 16 bytes; prev = the byte before v.x's lowest byte (in bits 24..31 of prev_word)
__device__ __forceinline__ u32 eq_prev_mask16 (uint4 v, u32 prev_word)
{
  u32 p0 = __byte_perm (prev_word, v.x, 0x6543);
  u32 p1 = __byte_perm (v.x, v.y, 0x6543);
  u32 p2 = __byte_perm (v.y, v.z, 0x6543);
  u32 p3 = __byte_perm (v.z, v.w, 0x6543);
  return gather_msb4 (zero_bytes (v.x ^ p0)) | (gather_msb4 (zero_bytes (v.y ^ p1)) << 4) |
         (gather_msb4 (zero_bytes (v.z ^ p2)) << 8) | (gather_msb4 (zero_bytes (v.w ^ p3)) << 12);
}

template <int W>
__global__ void __launch_bounds__ (SCAN_THREADS)
tjb_scan_kernel (const uint8_t *__restrict__ buf, long long n, int k, int min_len, u64 *__restrict__ out,
                 unsigned long long capacity, unsigned long long *__restrict__ counters, long long n_tiles)
{
  __shared__ __align__ (16) uint8_t s_raw[SCAN_SM_BYTES];
  __shared__ unsigned short s_e[SCAN_GROUPS + 2];
  __shared__ unsigned short s_cand[SCAN_TILE / 2];
  __shared__ u64 s_rec[SCAN_THREADS * W];
  __shared__ uint8_t s_lutf[256], s_lutr[256];
  __shared__ unsigned int s_ncand, s_nrec, s_nskip;
  __shared__ unsigned long long s_base;

  const int tid = threadIdx.x;
  if (tid < 256) {
    int f = tjb_code_fwd (tid);
    s_lutf[tid] = (uint8_t) f;
    s_lutr[tid] = (uint8_t) (f < 4 ? 3 - f : f);
  }
  if (tid == 0) { s_ncand = 0; s_nrec = 0; s_nskip = 0; }
  if (tid < 2) s_e[SCAN_GROUPS + tid] = 0;
  const int cand_ones = (min_len < 4 ? min_len : 4) - 1; // ones that must follow the start's zero bit
  __syncthreads ();

  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long g0 = tile * (long long) SCAN_TILE - SCAN_HALO_L; // global offset of s_raw[0]
    // ---- phase A.1: stage the tile + halos -----------------------------------------------------------
    for (int i = tid; i < SCAN_GROUPS; i += SCAN_THREADS) {
      long long g = g0 + 16LL * i;
      uint4 v;
      if (g >= 0 && g + 16 <= n) v = __ldg (reinterpret_cast<const uint4*> (buf + g));
      else {
        uint8_t b[16];
#pragma unroll
        for (int j = 0; j < 16; j++) { long long p = g + j; b[j] = (p >= 0 && p < n) ? buf[p] : (uint8_t) '\n'; }
        v.x = b[0] | (b[1] << 8) | (b[2] << 16) | ((u32) b[3] << 24);
        v.y = b[4] | (b[5] << 8) | (b[6] << 16) | ((u32) b[7] << 24);
        v.z = b[8] | (b[9] << 8) | (b[10] << 16) | ((u32) b[11] << 24);
        v.w = b[12] | (b[13] << 8) | (b[14] << 16) | ((u32) b[15] << 24);
      }
      *reinterpret_cast<uint4*> (s_raw + 16 * i) = v;
    }
    __syncthreads ();
    // ---- phase A.2: equal-to-previous masks ----------------------------------------------------------
    for (int i = tid; i < SCAN_GROUPS; i += SCAN_THREADS) {
      uint4 v = *reinterpret_cast<const uint4*> (s_raw + 16 * i);
      u32 prev = i ? *reinterpret_cast<const u32*> (s_raw + 16 * i - 4) : ~v.x; // group 0: "different"
      s_e[i] = (unsigned short) eq_prev_mask16 (v, prev);
    }
    __syncthreads ();
    // ---- phase A.3: tract-start candidates inside the tile -------------------------------------------
    {
      int grp = SCAN_HALO_L / 16 + tid;
      u32 e32 = (u32) s_e[grp] | ((u32) s_e[grp + 1] << 16);
      u32 s = ~e32 & (e32 >> 1);
      if (cand_ones >= 2) s &= e32 >> 2;
      if (cand_ones >= 3) s &= e32 >> 3;
      s &= 0xFFFFu;
      while (s) {
        int b = __ffs (s) - 1;
        s &= s - 1;
        unsigned slot = atomicAdd (&s_ncand, 1u);
        s_cand[slot] = (unsigned short) (16 * grp + b);
      }
    }
    __syncthreads ();
    // ---- phase B: one candidate per lane --------------------------------------------------------------
    const unsigned ncand = s_ncand;
    for (unsigned c0 = 0; c0 < ncand; c0 += SCAN_THREADS) {
      unsigned c = c0 + tid;
      if (c < ncand) {
        const int s = s_cand[c];
        const int ch = s_raw[s];
        const int code = s_lutf[ch];
        if (code != 5) {
          // run end: first zero bit of the mask after s
          int p = s + 1;
          while (p < SCAN_SM_BYTES) {
            u32 x = ((~(u32) s_e[p >> 4]) & 0xFFFFu) >> (p & 15);
            if (x) { p += __ffs (x) - 1; break; }
            p = (p | 15) + 1;
          }
          long long e = g0 + p - 1; // global index of the last byte of the run
          if (p >= SCAN_SM_BYTES) { // run leaves the staged window: finish in global memory
            e = g0 + SCAN_SM_BYTES - 1;
            while (e + 1 < n && buf[e + 1] == (uint8_t) ch) e++;
          }
          const long long gs = g0 + s;
          const long long len = e - gs + 1;
          if (len >= min_len) {
            const int erel = (int) ((e - g0) < (long long) (1 << 30) ? (e - g0) : (1 << 30));
            const bool in_smem = (erel + k) < SCAN_SM_BYTES;
            bool ok = true;
            u64 ctx0 = 0, ctx1 = 0;
            if (in_smem) {
              if (code < 2) {        // A / C tract: forward strand, hopo_counter.c:233-238
                for (int j = 0; j < k; j++) {
                  int cl = s_lutf[s_raw[s - k + j]], cr = s_lutf[s_raw[erel + 1 + j]];
                  ok = ok && (cl != 5) && (cr != 5);
                  ctx0 |= (u64) (cl & 3) << (2 * j);
                  ctx1 |= (u64) (cr & 3) << (2 * j);
                }
              } else {               // G / T / U tract (or an invalid byte: flank test only)
                for (int j = 0; j < k; j++) {
                  int cl = s_lutr[s_raw[erel + k - j]], cr = s_lutr[s_raw[s - 1 - j]];
                  ok = ok && (cl != 5) && (cr != 5);
                  ctx0 |= (u64) (cl & 3) << (2 * j);
                  ctx1 |= (u64) (cr & 3) << (2 * j);
                }
              }
            } else {                 // flank beyond the staged window: same arithmetic from global memory
              for (int j = 0; j < k; j++) {
                long long pl = (code < 2) ? gs - k + j : e + k - j;
                long long pr = (code < 2) ? e + 1 + j : gs - 1 - j;
                int bl = (pl >= 0 && pl < n) ? buf[pl] : '\n', br = (pr >= 0 && pr < n) ? buf[pr] : '\n';
                int cl = (code < 2) ? s_lutf[bl] : s_lutr[bl], cr = (code < 2) ? s_lutf[br] : s_lutr[br];
                ok = ok && (cl != 5) && (cr != 5);
                ctx0 |= (u64) (cl & 3) << (2 * j);
                ctx1 |= (u64) (cr & 3) << (2 * j);
              }
            }
            if (ok) {
              if (code == 4) atomicAdd (&s_nskip, 1u); // non-ACGTU run: undefined in the reference, fenced
              else {
                int base = code < 2 ? code : 3 - code;
                Rec<W> r = pack_rec<W> (ctx0, ctx1, base, sext_field (len, 10), code < 2 ? 1 : 2, k);
                unsigned slot = atomicAdd (&s_nrec, 1u);
#pragma unroll
                for (int j = 0; j < W; j++) s_rec[slot * W + j] = r.w[j];
              }
            }
          }
        }
      }
      __syncthreads ();
      const unsigned nrec = s_nrec;
      if (nrec) {
        if (tid == 0) s_base = atomicAdd (&counters[0], (unsigned long long) nrec);
        __syncthreads ();
        const unsigned long long base = s_base;
        for (unsigned i = tid; i < nrec * W; i += SCAN_THREADS) {
          unsigned long long r = base + i / W;
          if (r < capacity) out[r * W + (i % W)] = s_rec[i];
        }
        __syncthreads ();
        if (tid == 0) s_nrec = 0;
      }
    }
    __syncthreads ();
    if (tid == 0) s_ncand = 0;
    // the next iteration's first __syncthreads (after staging) orders this reset before any candidate append
  }
  __syncthreads ();
  if (tid == 0 && s_nskip) atomicAdd (&counters[1], (unsigned long long) s_nskip);
}
