// Read scan: homopolymer-tract detection + flanking-context key packing + strand canonicalisation.
//
// Replaces update_hopo_counter_from_seq() + add_kmer_to_hopo_counter() (reference src/hopo_counter.c:219-258,
// 285-307), which walk one read at a time, byte by byte, on one CPU thread.  Here a whole chunk of reads
// ("SEQ\nSEQ\n...", '\n' = read separator, written by csrc/tjb_fastq.c) is scanned by the full grid.
//
// The reads are touched once, at HBM speed, so the kernel is bound by instruction issue and by the integer pipe
// (LOP3 / SHF / PRMT / ISETP issue at half rate on sm_100): the design minimises instructions per input byte, moves
// what it can to the FMA pipe (shifts as multiply-high) and keeps every WARP autonomous.  A warp owns windows of 4096
// staged bytes (4000 owned + 32 left / 64 right halo), fetched by ONE 1-D bulk async copy (cp.async.bulk -> UBLKCP)
// onto a warp-private mbarrier; there is no block barrier after set-up.  The bytes are only read in pass 1, so there is
// ONE byte buffer per warp and the copy of the next window is issued right after pass 1; the derived arrays are what
// is double buffered.  60 % of the windows are dealt to the warps up front (a contiguous range each), the rest goes by
// ticket from one global counter, two windows per ticket, requested one window ahead.
//
//   pass 1   every lane turns 16 bytes into (a) one 32-bit word of 2-bit "gray" codes (byte >> 1) & 3 = A0 C1 T2 G3
//            (AND + one multiply per 4 bytes), (b) a 16-bit mask of non-letter bytes (bit 6 clear: the read
//            separator; two DP4A per 8 bytes), (c) one "bad group" bit: some byte is neither one of "ACGT" nor the
//            separator (bytes re-derived from code << 1 | letter by PRMT look-ups, two words per look-up, and
//            compared with the bytes themselves).  The three derived arrays (codes 1024 B, masks 512 B, bad bits 32 B
//            per window) stay in shared memory; everything after pass 1 works on them;
//   pass 2   code-run starts (candidates: min(max(m,2),4) equal codes, previous code differs) by funnel shifts on the
//            code words (one copy of the rows per candidate length); every lane reserves room for its candidates in
//            the warp's queue with one shared-memory atomic and pops its bits into it; the queue is carried over
//            window boundaries, so phase B always runs on 32 candidates;
//   phase B  one candidate per lane, no byte is touched: run end, both flanks (2k bits each) and the read-boundary
//            test are funnel shifts over the derived arrays.  A separator anywhere in [s-k, e+k] rejects the tract
//            (the reference's s >= k and e <= L-k-1 rules, hopo_counter.c:229, :232); a bad group in that span (N,
//            lower case, U, IUPAC, a control byte ...), a run of 32+ codes or a non-letter that is not '\n' sends the
//            candidate to the byte-exact slow path, which restates the reference loop literally.  The fast path
//            emits RAW observations (both flanks in gray codes, read orientation; tjb_common.cuh): the strand
//            canonicalisation of :233-246 is done by the kernel that counts them (aggregate_kernel waits on L2 and
//            has the issue slots; here they cost 7 % of the run time).  Records are staged per warp and appended to
//            the global array with one atomicAdd per ~50 records.
//
// Windows own the code runs that START inside their owned bytes.
#pragma once
#include <type_traits>
#include "tjb_common.cuh"


#ifndef SCAN_WARPS
#define SCAN_WARPS 8
#endif
#define SCAN_THREADS (SCAN_WARPS * 32)
#ifndef SCAN_CTAS_PER_SM
#define SCAN_CTAS_PER_SM 3
#endif
#ifndef SW_ROWS
#define SW_ROWS 8                     // 16-byte groups per lane and window (8: the only geometry the tests cover)
#endif
#define SW_WIN (SW_ROWS * 512)        // staged bytes per window = SW_ROWS rows of 32 groups of 16 bytes
#define SW_HALO_L 32                  // >= k, multiple of 16
#define SW_HALO_R 64                  // >= 31 (longest fast-path run) + k + 1
#define SW_TILE (SW_WIN - SW_HALO_L - SW_HALO_R) // owned bytes (4000 of 4096)
#define SW_GROUPS (SW_WIN / 16)
#define SW_G0 (SW_HALO_L / 16)        // owned groups [SW_G0, SW_G1)
#define SW_G1 (SW_G0 + SW_TILE / 16)
#ifndef SW_CHUNK
#define SW_CHUNK (SW_ROWS >= 8 ? 2 : 4) // consecutive windows per ticket (power of two)
#endif
#ifndef SW_QCAP
#define SW_QCAP 512                   // candidate queue (linear; moved to the front when the next round may not fit)
#endif
#define SW_QUOTA 7                    // new candidates per lane and round
#ifndef SCAN_BLOCKS_W2
#define SCAN_BLOCKS_W2 0
#endif
#ifndef SCAN_STATIC_PCT
#define SCAN_STATIC_PCT 60            // share of the windows dealt to the warps up front (the rest goes by ticket)
#endif
#ifndef SW_RCAP2
#define SW_RCAP2 64
#endif
#define SW_RCAP(W) ((W) == 2 ? SW_RCAP2 : 48) // staged records per warp (three CTAs of 8 warps must fit in 228 KB); flushed
                                      // when more than SW_RCAP - 32 are pending
#define SCAN_GRAY_MUL 0x00820820u     // ((w & 0x06060606) * this) >> 24 = the four 2-bit codes of w's bytes
#define SCAN_GRAY_LUT 0x47544341u     // code -> byte: 'A' 'C' 'T' 'G'
// PRMT look-up on (code << 1 | letter): letters A C T G in the odd slots; even slots: only code 1 can be a
// non-letter, the separator 0x0A; the other three hold a byte whose own code differs from the slot's (never equal)
#define SCAN_LUT_A 0x430A41FFu
#define SCAN_LUT_B 0x47005400u

template <int W> struct ScanWarpSmem {
  alignas (16) uint8_t raw[SW_WIN];   // sequence bytes of the window in pass 1; refilled for the next window right after
  u32 codes[2][SW_GROUPS];            // 16 gray codes per word, byte j of the group in bits 2j..2j+1
  u32 nl[2][SW_GROUPS / 2 + 4];       // bit i of the array = byte i of the window has bit 6 clear; 4 pad words
  u32 bad[2][SW_ROWS + 4];            // bit g = group g holds a byte that is neither ACGT nor separator-like
  alignas (16) u64 rec[SW_RCAP (W) * W];
  unsigned long long bar;             // mbarrier of the raw bytes
  u32 qtail, qpad;                    // end of the candidate queue (the lanes reserve their entries with an atomic)
  unsigned short cand[SW_QCAP];       // candidates: stage << 15 | row pair << 10 | lane << 5 | bit (see pass 2)
};
static_assert (sizeof (ScanWarpSmem<2>) % 16 == 0 && sizeof (ScanWarpSmem<3>) % 16 == 0, "warp slices must stay 16-byte aligned");
template <int W> constexpr size_t scan_smem_bytes () { return SCAN_WARPS * sizeof (ScanWarpSmem<W>) + 512; }

// forward code table: A0 C1 G2 T/U3 (either case), 4 = any other byte, 5 = '\n' (read separator)
__device__ __forceinline__ int tjb_code_fwd (int c)
{
  int u = c & 0xDF;
  int code = 4;
  if (u == 'A') code = 0; else if (u == 'C') code = 1; else if (u == 'G') code = 2; else if (u == 'T' || u == 'U') code = 3;
  if (c == '\n') code = 5;
  return code;
}

// ---- mbarrier / bulk async copy (sm_90+; SASS: SYNCS.*, UBLKCP) ---------------------------------------------
__device__ __forceinline__ u32 smem_u32 (const void *p) { return (u32) __cvta_generic_to_shared (p); }
__device__ __forceinline__ void mbar_init (unsigned long long *b, int count)
{ asm volatile ("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32 (b)), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx (unsigned long long *b, u32 bytes)
{ asm volatile ("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32 (b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void bulk_g2s (void *dst, const void *src, u32 bytes, unsigned long long *b)
{
  asm volatile ("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                :: "r"(smem_u32 (dst)), "l"(src), "r"(bytes), "r"(smem_u32 (b)) : "memory");
}
__device__ __forceinline__ void mbar_wait (unsigned long long *b, u32 parity)
{
  asm volatile ("{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}"
                :: "r"(smem_u32 (b)), "r"(parity) : "memory");
}
__device__ __forceinline__ int dp4a_us (u32 a_unsigned, u32 b_signed, int c)
{
  int d;
  asm ("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a_unsigned), "r"(b_signed), "r"(c));
  return d;
}
__device__ __forceinline__ u32 mad_hi (u32 a, u32 b, u32 c) // IMAD.HI: right shifts on the FMA pipe
{
  u32 d;
  asm ("mad.hi.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
// mask of the x lowest bits, any x (<= 0 -> 0, >= 32 -> all)
__device__ __forceinline__ u32 low_mask (int x)
{
  u32 d;
  asm ("bmsk.clamp.b32 %0, %1, %2;" : "=r"(d) : "r"(0), "r"(x < 0 ? 0 : x));
  return d;
}

// 64 bits of a 2-bit code array starting at code index p (bit 2p)
__device__ __forceinline__ u64 codes_at (const u32 *codes, int p)
{
  const int q = p >> 4, sh = (p & 15) << 1;
  const u32 a0 = codes[q], a1 = codes[q + 1], a2 = codes[q + 2];
  return ((u64) __funnelshift_r (a1, a2, sh) << 32) | __funnelshift_r (a0, a1, sh);
}

__device__ __forceinline__ u32 prmt (u32 a, u32 b, u32 sel)
{
  u32 d;
  asm ("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel)); // default mode: selector bit 3 replicates the sign
  return d;
}

// ---- the byte-exact slow path: the reference loop (hopo_counter.c:219-258) restated literally over every byte run of
// the code run that starts at global position ps0 (gray = its 2-bit code).  Taken by candidates whose span holds a byte
// that is neither ACGT nor the separator, and by runs of 32+.  Phase B is expanded at several places of the
// kernel and this code is cold (SCAN_SLOW_INLINE = 0 keeps one copy behind a call).  Returns the number of fenced runs (non-ACGTU run that is the first tract of its read).
#ifndef SCAN_SLOW_INLINE
#define SCAN_SLOW_INLINE 1 // 0: one copy behind a call; measured slower (spills around the call: 177 -> 188 us at configs[1])
#endif
#if SCAN_SLOW_INLINE
#define SCAN_SLOW_ATTR __forceinline__
#else
#define SCAN_SLOW_ATTR __noinline__
#endif
template <int W>
__device__ SCAN_SLOW_ATTR u32 scan_slow_path (const uint8_t *__restrict__ buf, long long n, int k, int min_len, u64 *__restrict__ out,
                                            unsigned long long capacity, unsigned long long *__restrict__ counters,
                                            const uint8_t *s_lutf, const uint8_t *s_lutr, long long ps0, int gray)
{
  u32 nskip = 0;
  auto byte_at = [&] (long long gp) -> int { return (gp >= 0 && gp < n) ? buf[gp] : '\n'; };
  // contexts of the byte run [rs, re] of a base with forward code `code` (0..3); false when a flank leaves the read
  auto flanks = [&] (long long rs, long long re, int code, u64 &ctx0, u64 &ctx1) -> bool {
    bool ok = true;
    ctx0 = ctx1 = 0;
    for (int j = 0; j < k; j++) {
      const int bl = byte_at (code < 2 ? rs - k + j : re + k - j), br = byte_at (code < 2 ? re + 1 + j : rs - 1 - j);
      const int cl = (code < 2) ? s_lutf[bl] : s_lutr[bl], cr = (code < 2) ? s_lutf[br] : s_lutr[br];
      ok = ok && (cl != 5) && (cr != 5);
      ctx0 |= (u64) (cl & 3) << (2 * j);
      ctx1 |= (u64) (cr & 3) << (2 * j);
    }
    return ok;
  };
  auto emit_slow = [&] (u64 ctx0, u64 ctx1, int code, long long len) { // rare path: straight to global
    const Rec<W> q = pack_rec<W> (ctx0, ctx1, code < 2 ? code : 3 - code, sext_field (len, 10), code < 2 ? 1 : 2, k);
    const unsigned long long pos = atomicAdd (&counters[CTR_RECORDS], 1ULL);
    if (pos < capacity) store_rec<W> (out, pos, q);
  };
  long long ps = ps0;
  while (ps < n) {
    const int b0 = byte_at (ps);
    if (((b0 >> 1) & 3) != gray) break;
    long long pe = ps;
    while (pe + 1 < n && byte_at (pe + 1) == b0) pe++;
    const long long len = pe - ps + 1;
    const int code = s_lutf[b0];
    // cheap exits first
    if (len >= min_len && code != 5 && byte_at (ps - 1) != '\n' && byte_at (pe + 1) != '\n') {
      u64 ctx0, ctx1;
      if (flanks (ps, pe, code & 3, ctx0, ctx1)) {
        if (code < 4) emit_slow (ctx0, ctx1, code, len);
        else {
          // Run of a byte that is none of ACGTU (e.g. NNNN).  The reference takes neither strand branch and calls
          // add_kmer_to_hopo_counter() with whatever context[], hopo_base_int and reverse_forward_flag hold
          // (hopo_counter.c:233-248): those of the last tract emitted for THIS read, with the length of this run.
          // Walk back to that tract; when the read has none the reference reads uninitialised stack: fenced,
          // counted and skipped.  (The right-edge rule holds for every earlier run because it holds for this one;
          // once a run is closer than k to the read start so is every run before it.)
          bool found = false;
          long long q = ps - 1;
          for (;;) {
            const int b1 = byte_at (q);
            if (b1 == '\n') break;
            long long qs = q;
            while (byte_at (qs - 1) == b1) qs--;
            const int c1 = s_lutf[b1];
            if (q - qs + 1 >= min_len && c1 < 4) {
              bool okl = true;
              for (int j = 1; j <= k; j++) okl = okl && byte_at (qs - j) != '\n';
              if (okl) { flanks (qs, q, c1, ctx0, ctx1); emit_slow (ctx0, ctx1, c1, len); found = true; }
              break;
            }
            q = qs - 1;
          }
          if (!found) nskip++;
        }
      }
    }
    ps = pe + 1;
  }
  return nskip;
}

template <int W>
__global__ void __launch_bounds__ (SCAN_THREADS, SCAN_CTAS_PER_SM)
tjb_scan_kernel (const uint8_t *__restrict__ buf, long long n, int k, int min_len, u64 *__restrict__ out,
                 unsigned long long capacity, unsigned long long *__restrict__ counters, long long n_tiles,
                 u32 *__restrict__ tile_ctr, u32 rec_block)
{
  extern __shared__ __align__ (16) unsigned char scan_smem[];
  ScanWarpSmem<W> *S = reinterpret_cast<ScanWarpSmem<W>*> (scan_smem) + (threadIdx.x >> 5);
  uint8_t *s_lutf = scan_smem + SCAN_WARPS * sizeof (ScanWarpSmem<W>), *s_lutr = s_lutf + 256;

  const int tid = threadIdx.x, lane = tid & 31;
  for (int i = tid; i < 256; i += SCAN_THREADS) {
    const int f = tjb_code_fwd (i);
    s_lutf[i] = (uint8_t) f;
    s_lutr[i] = (uint8_t) (f < 4 ? 3 - f : f);
  }
  if (lane == 0) {
    S->qtail = 0;
    mbar_init (&S->bar, 1);
    asm volatile ("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  const int cand_len = min_len < 4 ? min_len : 4;
  // powers of two (and -1) as run-time values: n < 2^62, so kq0 is zero, but the compiler cannot fold the multiplies
  // into integer-pipe shifts / logic
  const u32 kq0 = (u32) ((unsigned long long) n >> 62);
  const u32 kq4 = 16u + kq0, kq16 = (1u << 16) + kq0, kq26 = (1u << 26) + kq0, kq30 = (1u << 30) + kq0, kq31 = (1u << 31) + kq0;
  const u64 kmask = (k == 32) ? ~0ULL : ((1ULL << (2 * k)) - 1);
  __syncthreads (); // the only block-wide barrier: LUTs and mbarriers are set up

  u32 nrec = 0, nskip = 0, parity = 0;
  u32 head = 0, tail = 0, n_old = 0; // candidate FIFO; n_old = entries that belong to the previous window

  // a window is "interior" when [g0, g0 + SW_WIN) lies inside the buffer: one bulk copy stages it.  g0 >= 0 from
  // window 1 on; the last interior window is the largest t with t * SW_TILE - SW_HALO_L + SW_WIN <= n
  const long long t_int = n >= SW_WIN - SW_HALO_L ? (n - (SW_WIN - SW_HALO_L)) / SW_TILE : 0; // 0: none (window 0 never is)
  const u32 t_last_interior = t_int > 0xffffffffLL ? 0xffffffffu : (u32) t_int;
  auto interior = [&] (u32 t) { return t >= 1u && t <= t_last_interior; };
  // the caller has made sure (__syncwarp) that every lane is done reading the previous window's bytes
  u32 run_end = 0; // end of the contiguous windows this warp holds (see the main loop)
  auto stage_tile = [&] (u32 t) {
    const long long g0 = (long long) t * SW_TILE - SW_HALO_L;
    if (interior (t)) {
      if (lane == 0) {
        asm volatile ("fence.proxy.async.shared::cta;" ::: "memory"); // generic-proxy accesses of the buffer are done
        mbar_expect_tx (&S->bar, SW_WIN);
        bulk_g2s (S->raw, buf + g0, SW_WIN, &S->bar);
      }
    } else { // first / last windows: bytes outside the buffer read as separators
      for (int i = lane; i < SW_WIN / 4; i += 32) {
        u32 w = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) { const long long p = g0 + 4 * i + j; w |= (u32) ((p >= 0 && p < n) ? buf[p] : (uint8_t) '\n') << (8 * j); }
        reinterpret_cast<u32*> (S->raw)[i] = w;
      }
    }
  };
  // staged records -> global array.  Every warp of the grid takes its positions from ONE counter, and one address
  // serves about one atomic per nanosecond: with a request per ~40 records a dense sample (configs[4]: 39 M records in
  // 1.3 ms) ran at exactly that rate, whatever the instruction count.  rec_block > 0: a warp reserves rec_block (>= 64)
  // positions at a time and fills them flush by flush (a flush may straddle two blocks); what is left of its last block
  // when the kernel ends is filled with all-zero records, which no real record is (aggregate_kernel skips them).
  // rec_block == 0: exact reservations, no holes (small inputs, and callers that read CTR_RECORDS as a count).
  unsigned long long blk_pos = 0, blk_end = 0;
  auto store_staged = [&] (u32 first, unsigned long long base, u32 cnt) { // records [first, first + cnt) of the staging area
    if (W == 2 && base + cnt <= capacity) { // one record = one 16-byte store
      const ulonglong2 *src = reinterpret_cast<const ulonglong2*> (S->rec) + first;
      ulonglong2 *dst = reinterpret_cast<ulonglong2*> (out) + base;
      static_assert (SW_RCAP (2) <= 64, "two stores per lane cover the staging area");
      if ((u32) lane < cnt) dst[lane] = src[lane];
      if ((u32) lane + 32 < cnt) dst[lane + 32] = src[lane + 32];
    } else {
      for (u32 i = lane; i < cnt * W; i += 32) {
        const unsigned long long r = base + i / W;
        if (r < capacity) out[r * W + (i % W)] = S->rec[first * W + i];
      }
    }
  };
  auto flush = [&] () {
    if ((W == 2 && !SCAN_BLOCKS_W2) || !rec_block) { // (two-word records: measured slower with blocks, 181 -> 190-195 us at configs[1]; the kernel grows from 6.4 K to 7.6 K instructions)
      unsigned long long base = 0;
      if (lane == 0) base = atomicAdd (&counters[CTR_RECORDS], (unsigned long long) nrec);
      base = __shfl_sync (0xffffffffu, base, 0);
      store_staged (0, base, nrec);
    } else {
      u32 done = 0;
      while (done < nrec) { // at most twice: rec_block >= the staging area
        if (blk_pos == blk_end) {
          unsigned long long base = 0;
          if (lane == 0) base = atomicAdd (&counters[CTR_RECORDS], (unsigned long long) rec_block);
          blk_pos = __shfl_sync (0xffffffffu, base, 0);
          blk_end = blk_pos + rec_block;
        }
        const u32 room = (u32) (blk_end - blk_pos), n1 = nrec - done < room ? nrec - done : room;
        store_staged (done, blk_pos, n1);
        blk_pos += n1; done += n1;
      }
    }
    nrec = 0;
    __syncwarp ();
  };

  // ---- phase B: one candidate per lane (a queue entry, or -1) ---------------------------------------------------
  // cur_st / cur_g0: the current window; candidates carried over from the previous one start at old_g0.  Only the
  // derived arrays are in shared memory at this point: the slow path reads its bytes from global memory (L2 hits).
  auto process_batch = [&] (int cnd, int cur_st, long long cur_g0, long long old_g0) {
    bool emit = false, slow = false;
    Rec<W> r;
    int gray = 0, s = 0, par = 0;
    if (cnd >= 0) {
      par = cnd >> 15;
      s = (cnd & 0xC00) | ((cnd & 1) << 9) | ((cnd >> 1) & 0x1FF); // 1024 * row pair + 512 * (bit & 1) + 16 * lane + bit / 2
      const u32 *codes = S->codes[par];
      const u64 win = codes_at (codes, s); // codes s .. s+31
      gray = (int) ((u32) win & 3u);
      const u32 pat = (u32) gray * 0x55555555u;
      const u32 x0 = (u32) win ^ pat, x1 = (u32) (win >> 32) ^ pat; // lowest set bit: either bit of the first code that differs
      const int len = x0 ? ((__ffs (x0) - 1) >> 1) : (x1 ? 16 + ((__ffs (x1) - 1) >> 1) : 32);
      if (len >= 32) slow = true; // long run: may leave the window
      else if (len >= min_len) {
        const int P = s - k, nspan = 2 * k + len; // [P, P + nspan) = left flank, run, right flank
        const int gl = P >> 4, ng = ((P + nspan - 1) >> 4) - gl + 1;
        const u32 *bad = S->bad[par];
        if (__funnelshift_r (bad[gl >> 5], bad[(gl >> 5) + 1], gl & 31) & low_mask (ng)) slow = true;
        else {
          // every byte of the span is one of "ACGT" or the separator: a separator anywhere in it means that the
          // tract is closer than k to a read end (hopo_counter.c:229, :232), for every byte run of this code run
          const u32 *nlw = S->nl[par];
          const int q = P >> 5, sh = P & 31;
          const u32 n0 = nlw[q], n1 = nlw[q + 1], n2 = nlw[q + 2], n3 = nlw[q + 3];
          const u32 m0 = __funnelshift_r (n0, n1, sh) & low_mask (nspan);
          const u32 m1 = __funnelshift_r (n1, n2, sh) & low_mask (nspan - 32);
          const u32 m2 = __funnelshift_r (n2, n3, sh) & low_mask (nspan - 64);
          if (!(m0 | m1 | m2)) {
            const u64 gl64 = codes_at (codes, P) & kmask, gr64 = codes_at (codes, s + len) & kmask;
            // the strand canonicalisation is left to the kernel that counts the records (measured: configs[1] 191 -> 178 us;
            // configs[4] 1.30 -> 1.17 ms once the record positions were reserved in blocks: without them the faster scan
            // ran into the rate of its one position counter, 1.75 ms)
            if (SCAN_RAW_OBS) r = pack_obs<W> (gl64, gr64, gray, len); // len < 32: the 10-bit field holds it as it is
            else r = canon_from_gray<W> (gl64, gr64, gray, len, k);
            emit = true;
          }
        }
      }
    }
    const u32 em = __ballot_sync (0xffffffffu, emit);
    if (emit) {
      const u32 slot = nrec + __popc (em & ((1u << lane) - 1));
#pragma unroll
      for (int j = 0; j < W; j++) S->rec[slot * W + j] = r.w[j];
    }
    nrec += __popc (em);
    if (slow) // rare
      nskip += scan_slow_path<W> (buf, n, k, min_len, out, capacity, counters, s_lutf, s_lutr, (par == cur_st ? cur_g0 : old_g0) + s, gray);
    __syncwarp ();
    if (nrec > SW_RCAP (W) - 32) flush ();
  };

  // ---- main loop: this warp's windows; the bytes of window t+1 are fetched while window t is in pass 2 / phase B ----
  // windows are handed out by a global ticket counter (SMs on the two dies do not run at the same speed, a static
  // deal leaves a tail); the ticket of the window after next is requested one window ahead, its latency is hidden
  // (SW_CHUNK consecutive windows per ticket: one counter served 258 K atomics in 250 us and its latency showed)
  // SCAN_STATIC_PCT: that share of the windows is dealt up front, a contiguous range per warp, and only the rest goes by
  // ticket (a quarter of the atomics on the one counter, and their answers come back sooner)
  u32 t_cur = 0, t_nxt = 0, t_fut = 0;
  bool want_ticket = false;
  const u32 n_warps = gridDim.x * SCAN_WARPS;
  const u32 n_static = (u32) (((unsigned long long) n_tiles * SCAN_STATIC_PCT / 100) / n_warps) & ~(u32) (SW_CHUNK - 1);
  const u32 t_dyn0 = n_static * n_warps; // tickets count chunks from here on
  if (n_static) {
    if (lane == 0) t_fut = atomicAdd (tile_ctr, 1u); // the first chunk after the dealt range: asked for now, needed much later
    t_cur = (blockIdx.x * SCAN_WARPS + (threadIdx.x >> 5)) * n_static;
    run_end = t_cur + n_static;
  } else {
    if (lane == 0) t_fut = atomicAdd (tile_ctr, 2u);
    t_fut = __shfl_sync (0xffffffffu, t_fut, 0);
    t_cur = t_fut * SW_CHUNK;
    run_end = t_cur + SW_CHUNK;
    t_fut++; // the second chunk of the first request
  }
  if (t_cur + 1 < run_end) t_nxt = t_cur + 1;
  else { t_nxt = t_dyn0 + __shfl_sync (0xffffffffu, t_fut, 0) * SW_CHUNK; run_end = t_nxt + SW_CHUNK; want_ticket = true; }
  long long g0 = (long long) t_cur * SW_TILE - SW_HALO_L, old_g0 = 0;
  int st = 0;
  bool in_cur = false;
  if (t_cur < n_tiles) { in_cur = interior (t_cur); stage_tile (t_cur); }
  const u32 hdr = (u32) lane << 5;
  // the queue is linear: what is left moves to the front only when the next round might not fit behind it
  auto queue_to_front = [&] () {
    const u32 cnt = tail - head; // < 32
    const unsigned short v = lane < (int) cnt ? S->cand[head + lane] : (unsigned short) 0;
    __syncwarp ();
    if (lane < (int) cnt) S->cand[lane] = v;
    head = 0; tail = cnt;
    if (lane == 0) S->qtail = cnt;
    __syncwarp ();
  };
  // one more pass of the loop after the last window when candidates are left (fewer than 32, of the last window, whose
  // derived arrays are still there): it runs the rounds below without new candidates, and they drain the queue
  for (;; st ^= 1) {
    const bool have_win = t_cur < n_tiles;
    if (!have_win && !n_old) break;
    u32 cb[SW_ROWS / 2];
#pragma unroll
    for (int j = 0; j < SW_ROWS / 2; j++) cb[j] = 0;
    if (have_win) {
      if (want_ticket) { if (lane == 0) t_fut = atomicAdd (tile_ctr, 1u); want_ticket = false; }
      if (in_cur) { mbar_wait (&S->bar, parity); parity ^= 1u; }
      else __syncwarp ();
      u32 *codes = S->codes[st];
      u32 c[SW_ROWS];
      {
        // ---- pass 1: 16 bytes per lane and row -> gray codes, non-letter mask, bad-group bit -------------------
        const uint4 *raw128 = reinterpret_cast<const uint4*> (S->raw);
        unsigned short *nl16 = reinterpret_cast<unsigned short*> (S->nl[st]);
#pragma unroll
        for (int row = 0; row < SW_ROWS; row++) {
          const int grp = row * 32 + lane;
          const uint4 v = raw128[grp];
          const u32 w0 = v.x, w1 = v.y, w2 = v.z, w3 = v.w;
          const u32 m0 = w0 & 0x06060606u, m1 = w1 & 0x06060606u, m2 = w2 & 0x06060606u, m3 = w3 & 0x06060606u;
          const u32 p0 = m0 * SCAN_GRAY_MUL, p1 = m1 * SCAN_GRAY_MUL, p2 = m2 * SCAN_GRAY_MUL, p3 = m3 * SCAN_GRAY_MUL;
          c[row] = __byte_perm (__byte_perm (p0, p1, 0x0073), __byte_perm (p2, p3, 0x0073), 0x5410);
          // letters have bit 6 set.  Look-up selectors: code << 1 | letter, the nibbles of two words interleaved in one
          // register.  The integer pipe is the bottleneck, so shifts, adds and the word comparison are multiply-adds
          // with multipliers the compiler cannot see through (kq*): they issue on the FMA pipe.
          const u32 a0 = w0 & 0x40404040u, a1 = w1 & 0x40404040u, a2 = w2 & 0x40404040u, a3 = w3 & 0x40404040u;
          const u32 z01 = mad_hi (a1, kq30, mad_hi (a0, kq26, m1 * kq4 + m0)); // m0 | m1 << 4 | a0 >> 6 | a1 >> 2
          const u32 z23 = mad_hi (a3, kq30, mad_hi (a2, kq26, m3 * kq4 + m2));
          const u32 xa = prmt (SCAN_LUT_A, SCAN_LUT_B, z01), xb = prmt (SCAN_LUT_A, SCAN_LUT_B, __umulhi (z01, kq16));
          const u32 xc = prmt (SCAN_LUT_A, SCAN_LUT_B, z23), xd = prmt (SCAN_LUT_A, SCAN_LUT_B, __umulhi (z23, kq16));
          // the bytes their codes stand for, against the bytes themselves: (expected ^ byte) | ... over the group (the
          // compiler makes it four compare-and-or instructions; a multiply-add / DP4A form on the FMA pipe took twice as many)
          const u32 e01 = (prmt (xa, xb, 0x6420) ^ w0) | (prmt (xa, xb, 0x7531) ^ w1);
          const u32 e23 = (prmt (xc, xd, 0x6420) ^ w2) | (prmt (xc, xd, 0x7531) ^ w3);
          const u32 acc = e01 | e23;
          // bit j of the mask = byte j is a non-letter: 64 * (255 - sum of the weights of the letters), bytes are 0x40 / 0
          const int lo8 = dp4a_us (a1, 0x80C0E0F0u, dp4a_us (a0, 0xF8FCFEFFu, 255 * 64));
          const int hi8 = dp4a_us (a3, 0x80C0E0F0u, dp4a_us (a2, 0xF8FCFEFFu, 255 * 64));
          nl16[grp] = (unsigned short) __umulhi ((u32) (hi8 * 256 + lo8), kq26);
          codes[grp] = c[row];
          const u32 bb = __ballot_sync (0xffffffffu, acc != 0);
          if (lane == 0) S->bad[st][row] = bb;
        }
      }
      __syncwarp (); // the derived arrays are complete and nobody reads the bytes any more:
      in_cur = false;
      if (t_nxt < n_tiles) { in_cur = interior (t_nxt); stage_tile (t_nxt); } // fetch the next window over them
      // ---- pass 2: code-run starts of the owned groups.  Rows 2j (even bits) and 2j+1 (odd bits) share cand_bits[j]:
      // bit b = byte b / 2 of group 64 * j + 32 * (b & 1) + lane
      // one copy of the rows per candidate length (2, 3, 4): no test of cand_len inside them
      auto pass2 = [&] (auto cl) {
        constexpr int CL = decltype (cl)::value;
#pragma unroll
        for (int row = 0; row < SW_ROWS; row++) {
          const int grp = row * 32 + lane;
          const u32 cc = c[row], cn = codes[(grp + 1) & (SW_GROUPS - 1)], cp = codes[(grp - 1) & (SW_GROUPS - 1)];
          u32 tt = cc ^ __funnelshift_r (cc, cn, 2);
          if (CL >= 3) tt |= cc ^ __funnelshift_r (cc, cn, 4);
          if (CL >= 4) tt |= cc ^ __funnelshift_r (cc, cn, 6);
          const u32 dp = cc ^ __funnelshift_l (cp, cc, 2);
          u32 cs = ~(tt | __umulhi (tt, kq31)) & (dp | __umulhi (dp, kq31)) & 0x55555555u; // x >> 1 on the FMA pipe
          if (row == 0 && lane < SW_G0) cs = 0;                       // halo groups are owned by the neighbouring windows
          if (row == SW_ROWS - 1 && lane >= SW_G1 - 32 * (SW_ROWS - 1)) cs = 0;
          cb[row >> 1] |= cs << (row & 1);
        }
      };
      if (cand_len >= 4) pass2 (std::integral_constant<int, 4> ());
      else if (cand_len == 3) pass2 (std::integral_constant<int, 3> ());
      else pass2 (std::integral_constant<int, 2> ());
    }
    // ---- compaction into the queue (at most SW_QUOTA per lane and round); full batches of 32 go to phase B ------
    for (;;) {
      if (tail + 32 * SW_QUOTA > SW_QCAP) queue_to_front ();
      int have = 0;
#pragma unroll
      for (int j = 0; j < SW_ROWS / 2; j++) have += __popc (cb[j]);
      const int mine = have < SW_QUOTA ? have : SW_QUOTA;
      const bool capped = __any_sync (0xffffffffu, have > SW_QUOTA); // rare: some lane keeps candidates for another round
      // every lane reserves its entries with one shared-memory atomic on the warp's tail (their order within a round
      // does not matter); the new tail is read back after the entries are written
      u32 pos = 0;
      if (mine) pos = atomicAdd (&S->qtail, (u32) mine);
      {
        unsigned short *qp = S->cand + pos;
        if (!capped) {
#pragma unroll
          for (int j = 0; j < SW_ROWS / 2; j++) {
            const u32 h = ((u32) st << 15) | ((u32) j << 10) | hdr;
            u32 m = cb[j];
            while (m) { const int b = __ffs (m) - 1; m &= m - 1; *qp++ = (unsigned short) (h + b); }
            cb[j] = 0;
          }
        } else {
          int left = mine;
#pragma unroll
          for (int j = 0; j < SW_ROWS / 2; j++) {
            const u32 h = ((u32) st << 15) | ((u32) j << 10) | hdr;
            u32 m = cb[j];
            while (m && left) { const int b = __ffs (m) - 1; m &= m - 1; *qp++ = (unsigned short) (h + b); left--; }
            cb[j] = m;
          }
        }
        __syncwarp ();
        tail = *reinterpret_cast<volatile u32*> (&S->qtail);
      }
      // full batches; then, once, whatever is left when it still holds candidates of the previous window: they must go
      // before its derived arrays are overwritten by the next window.  (One expansion of phase B for both cases.)
      for (;;) {
        const u32 left = tail - head;
        if (left < 32 && !n_old) break;
        const u32 take = left < 32 ? left : 32;
        process_batch ((u32) lane < take ? (int) S->cand[head + lane] : -1, st, g0, old_g0);
        head += take;
        n_old = n_old > take ? n_old - take : 0;
      }
      if (!capped) break; // (a capped lane has bits left: another round)
    }
    n_old = tail - head; // what is left (< 32 candidates) waits for the next window
    old_g0 = g0;
    t_cur = t_nxt;
    if (t_nxt + 1 < run_end) t_nxt++;
    else { t_nxt = t_dyn0 + __shfl_sync (0xffffffffu, t_fut, 0) * SW_CHUNK; run_end = t_nxt + SW_CHUNK; want_ticket = true; }
    g0 = (long long) t_cur * SW_TILE - SW_HALO_L;
  }
  if (nrec) flush ();
  if (W != 2 || SCAN_BLOCKS_W2) for (unsigned long long r = blk_pos + lane; r < blk_end; r += 32) // the rest of this warp's last block: empty records
    if (r < capacity) {
#pragma unroll
      for (int j = 0; j < W; j++) out[r * W + j] = 0ULL;
    }
#pragma unroll
  for (int d = 16; d; d >>= 1) nskip += __shfl_xor_sync (0xffffffffu, nskip, d);
  if (lane == 0 && nskip) atomicAdd (&counters[CTR_SKIP], (unsigned long long) nskip);
}
