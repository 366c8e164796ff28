// Read scan: homopolymer-tract detection + flanking-context key packing + strand canonicalisation.
//
// Replaces update_hopo_counter_from_seq() + add_kmer_to_hopo_counter() (reference src/hopo_counter.c:219-258,
// 285-307), which walk one read at a time, byte by byte, on one CPU thread.  Here a whole chunk of reads
// ("SEQ\nSEQ\n...", '\n' = read separator, written by csrc/tjb_fastq.c) is scanned by the full grid.
//
// The kernel is instruction-issue bound (the reads are touched once, at HBM speed), so (1) the common path works on
// 32-bit words, and (2) every WARP is autonomous: it owns 2 KiB tiles, stages them with a 1-D bulk async copy
// (cp.async.bulk -> UBLKCP, completion on a warp-private mbarrier, double buffered so that tile t+1 is in flight
// while tile t is processed) and never meets another warp at a barrier.
//
//   phase A  per 512-byte row, every lane squeezes its 16 bytes into a 32-bit word of 2-bit "gray" codes,
//            (byte >> 1) & 3 = A0 C1 T2 G3, with one AND + one multiply per 4 bytes.  Runs are detected on that word
//            (funnel shifts against the neighbouring groups): a position is a CANDIDATE when min(max(m,2),4) equal
//            codes start there and the previous code differs.  Equal bytes have equal codes, so every byte run of
//            that length lies inside exactly one code run; code runs that are not byte runs (N next to G, '\n' next
//            to C, lower case ...) are sorted out exactly in phase B.  Candidates are queued per warp (ballot
//            compaction) and processed 32 at a time, one per lane;
//   phase B  run end by word compares; both flanks are read as k bytes re-aligned with PRMT, validated against
//            "ACTG"[code] (any other byte, the read separator included, shows up as a non-zero difference) and
//            packed 4 bases per multiply; strand canonicalisation is a 64-bit bit reversal (hopo_counter.c:239-246).
//            A flank holding a separator rejects the tract (the reference's s >= k and e <= L-k-1 rules, :229,
//            :232).  Anything unusual (non-ACGT byte in the run or a flank, U, lower case, a run or flank leaving the
//            staged window) takes the byte-exact slow path, which restates the reference loop literally.  Records
//            are staged per warp and appended to the global array with one atomicAdd per ~50 records; the bin
//            histogram of the partition plan (tjb_bucket.cuh) is accumulated here too.
//
// Tiles own the code runs that START inside them.
#pragma once
#include "tjb_common.cuh"
#include "tjb_bucket.cuh" // rec_bin(): the scan also accumulates the bin histogram the partition plan is made from


#define SCAN_WARPS 8
#define SCAN_THREADS (SCAN_WARPS * 32)
#define SW_TILE 2048                  // bytes owned by one warp tile
#define SW_HALO_L 64                  // >= k + 4, multiple of 16
#define SW_HALO_R 128
#define SW_BYTES (SW_HALO_L + SW_TILE + SW_HALO_R)
#define SW_ROWS (SW_TILE / 512)
#define SW_QCAP 320                   // candidate queue: < 32 left over + at most 256 new per row
#define SW_RCAP 64                    // staged records per warp; flushed when more than SW_RCAP - 32 are pending
#define SCAN_GRAY_MUL 0x00820820u     // ((w & 0x06060606) * this) >> 24 = the four 2-bit codes of w's bytes
#define SCAN_GRAY_LUT 0x47544341u     // code -> byte: 'A' 'C' 'T' 'G'

template <int W> struct ScanWarpSmem {
  uint8_t raw[2][SW_BYTES + 16];      // two stages; 16 pad bytes read (never used) by the flank word loads
  u64 rec[SW_RCAP * W];
  unsigned long long bar[2];          // mbarriers of the two stages
  unsigned short cand[SW_QCAP];
};
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

// k bytes starting at shared-memory offset pos -> 2-bit gray codes, base j in bits 2j..2j+1.
// bad != 0 when any of the k bytes is not one of "ACGT" (upper case); sep != 0 when one of them is the read separator.
__device__ __forceinline__ u64 flank_gray (const u32 *raw32, int pos, int nw, u32 tail_mask, u32 &bad, u32 &sep)
{
  const int a = pos >> 2;
  const u32 sel = 0x3210u + 0x1111u * (u32) (pos & 3);
  u32 lo = 0, hi = 0, acc = 0, sacc = 0;
  u32 w_lo = raw32[a];
#pragma unroll
  for (int j = 0; j < 8; j++) {
    if (j >= nw) break;
    const u32 w_hi = raw32[a + j + 1];
    const u32 w = __byte_perm (w_lo, w_hi, sel);
    w_lo = w_hi;
    u32 m = w & 0x06060606u;
    const u32 y = m >> 1, t = y | (y >> 4);
    u32 d = w ^ __byte_perm (SCAN_GRAY_LUT, 0, __byte_perm (t, 0, 0x4420));
    const u32 x = w ^ 0x0A0A0A0Au;
    u32 z = ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x) & 0x80808080u;
    if (j == nw - 1) { d &= tail_mask; m &= tail_mask; z &= tail_mask; }
    acc |= d;
    sacc |= z;
    const u32 c8 = (m * SCAN_GRAY_MUL) >> 24;
    if (j < 4) lo |= c8 << (8 * j); else hi |= c8 << (8 * (j - 4));
  }
  bad = acc;
  sep = sacc;
  return ((u64) hi << 32) | lo;
}

// reverse complement of a k-base word of BINARY codes (A0 C1 G2 T3), base j in bits 2j..2j+1
__device__ __forceinline__ u64 revcomp2 (u64 x, int k)
{
  u64 y = __brevll (x) >> (64 - 2 * k); // k >= 2; reverses the base order and the two bits of every base
  y = ((y & 0x5555555555555555ULL) << 1) | ((y >> 1) & 0x5555555555555555ULL);
  const u64 mask = (k == 32) ? ~0ULL : ((1ULL << (2 * k)) - 1);
  return ~y & mask;
}

template <int W>
__global__ void __launch_bounds__ (SCAN_THREADS, 4)
tjb_scan_kernel (const uint8_t *__restrict__ buf, long long n, int k, int min_len, u64 *__restrict__ out,
                 unsigned long long capacity, unsigned long long *__restrict__ counters, long long n_tiles,
                 u32 *__restrict__ bin_hist, int bin_bits)
{
  extern __shared__ __align__ (16) unsigned char scan_smem[];
  ScanWarpSmem<W> *S = reinterpret_cast<ScanWarpSmem<W>*> (scan_smem) + (threadIdx.x >> 5);
  uint8_t *s_lutf = scan_smem + SCAN_WARPS * sizeof (ScanWarpSmem<W>), *s_lutr = s_lutf + 256;

  const int tid = threadIdx.x, lane = tid & 31;
  const u32 lt_mask = (1u << lane) - 1;
  {
    const int f = tjb_code_fwd (tid); // SCAN_THREADS == 256
    s_lutf[tid] = (uint8_t) f;
    s_lutr[tid] = (uint8_t) (f < 4 ? 3 - f : f);
  }
  if (lane < 8) reinterpret_cast<u32*> (S->raw[lane >> 2] + SW_BYTES)[lane & 3] = 0x0A0A0A0Au;
  if (lane == 0) {
    mbar_init (&S->bar[0], 1);
    mbar_init (&S->bar[1], 1);
    asm volatile ("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  const int cand_len = min_len < 4 ? min_len : 4;
  const int nw = (k + 3) >> 2;
  const u32 tail_mask = (k & 3) ? ((1u << (8 * (k & 3))) - 1) : 0xffffffffu;
  __syncthreads (); // the only block-wide barrier: LUTs and mbarriers are set up

  const long long n_warps = (long long) gridDim.x * SCAN_WARPS;
  u32 nrec = 0, nskip = 0, parity = 0;

  // a tile is "interior" when its whole window [g0, g0 + SW_BYTES) lies inside the buffer: one bulk copy stages it
  auto interior = [&] (long long t) { const long long g0 = t * SW_TILE - SW_HALO_L; return g0 >= 0 && g0 + SW_BYTES <= n; };
  auto stage_tile = [&] (long long t, int st) {
    const long long g0 = t * SW_TILE - SW_HALO_L;
    if (interior (t)) {
      if (lane == 0) {
        asm volatile ("fence.proxy.async.shared::cta;" ::: "memory"); // generic-proxy accesses of this stage are done
        mbar_expect_tx (&S->bar[st], SW_BYTES);
        bulk_g2s (S->raw[st], buf + g0, SW_BYTES, &S->bar[st]);
      }
    } else { // first / last tiles: bytes outside the buffer read as separators
      for (int i = lane; i < SW_BYTES / 4; i += 32) {
        u32 w = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) { const long long p = g0 + 4 * i + j; w |= (u32) ((p >= 0 && p < n) ? buf[p] : (uint8_t) '\n') << (8 * j); }
        reinterpret_cast<u32*> (S->raw[st])[i] = w;
      }
    }
  };
  auto flush = [&] () { // staged records -> global array, one atomicAdd
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd (&counters[0], (unsigned long long) nrec);
    base = __shfl_sync (0xffffffffu, base, 0);
    for (u32 i = lane; i < nrec * W; i += 32) {
      const unsigned long long r = base + i / W;
      if (r < capacity) out[r * W + (i % W)] = S->rec[i];
    }
    nrec = 0;
    __syncwarp ();
  };

  // ---- phase B: one candidate (window offset s of a code-run start, or -1) per lane ----------------------------
  auto process_batch = [&] (int s, const uint8_t *raw, long long g0) {
    const u32 *raw32 = reinterpret_cast<const u32*> (raw);
    bool emit = false, slow = false;
    Rec<W> r;
    int gray = 0;
    if (s >= 0) {
      const int ch = raw[s];
      gray = (ch >> 1) & 3;
      // first byte after s that differs from ch (word compares on the re-aligned bytes s+1, s+2, ...)
      int p = s + 1;
      {
        const u32 sel = 0x3210u + 0x1111u * (u32) (p & 3);
        const u32 pat = (u32) ch * 0x01010101u;
        int a = p >> 2;
        u32 w_lo = raw32[a];
        for (;;) {
          const u32 w_hi = raw32[a + 1];
          const u32 x = __byte_perm (w_lo, w_hi, sel) ^ pat;
          const u32 nz = (((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x) & 0x80808080u;
          if (nz) { p += (__ffs (nz) - 1) >> 3; break; }
          p += 4; a++; w_lo = w_hi;
          if (p >= SW_BYTES) break;
        }
      }
      // the code run must end where the byte run ends (else: N next to G, '\n' next to C, ... -> exact path)
      slow = p >= SW_BYTES - 4 || ((raw[p] >> 1) & 3) == gray || (p + k + 4 >= SW_BYTES) ||
             ch != (int) ((SCAN_GRAY_LUT >> (8 * gray)) & 0xff);
      const int e = p - 1;
      if (!slow && e - s + 1 >= min_len) {
        u32 badl, badr, sepl, sepr;
        const u64 gl = flank_gray (raw32, s - k, nw, tail_mask, badl, sepl);
        const u64 gr = flank_gray (raw32, e + 1, nw, tail_mask, badr, sepr);
        if (badl | badr) {
          // a separator inside a flank: the tract is closer than k to a read end (hopo_counter.c:229, :232);
          // any other byte (N, lower case, U, ...): byte-exact path
          slow = !(sepl | sepr);
        } else {
          const u64 m5 = 0x5555555555555555ULL;
          const u64 bl = gl ^ ((gl >> 1) & m5), br = gr ^ ((gr >> 1) & m5); // gray -> A0 C1 G2 T3
          u64 ctx0, ctx1;
          if (gray < 2) { ctx0 = bl; ctx1 = br; }                       // A / C tract: forward strand, :233-238
          else { ctx0 = revcomp2 (br, k); ctx1 = revcomp2 (bl, k); }    // T / G tract: reverse complement, :239-246
          r = pack_rec<W> (ctx0, ctx1, gray & 1, sext_field (e - s + 1, 10), gray < 2 ? 1 : 2, k);
          atomicAdd (&bin_hist[rec_bin<W> (r, k, bin_bits)], 1u);
          emit = true;
        }
      }
    }
    const u32 em = __ballot_sync (0xffffffffu, emit);
    if (emit) {
      const u32 slot = nrec + __popc (em & lt_mask);
#pragma unroll
      for (int j = 0; j < W; j++) S->rec[slot * W + j] = r.w[j];
    }
    nrec += __popc (em);
    if (slow) {
      // byte-exact restatement of the reference loop over every byte run inside this code run
      auto byte_at = [&] (long long gp) -> int {
        const long long rel = gp - g0;
        if (rel >= 0 && rel < SW_BYTES) return raw[rel];
        return (gp >= 0 && gp < n) ? buf[gp] : '\n';
      };
      long long ps = g0 + s;
      while (ps < n) {
        const int b0 = byte_at (ps);
        if (((b0 >> 1) & 3) != gray) break;
        long long pe = ps;
        while (pe + 1 < n && byte_at (pe + 1) == b0) pe++;
        const long long len = pe - ps + 1;
        const int code = s_lutf[b0];
        // cheap exits first: most code runs that are not byte runs are a C tract touching a read separator
        if (len >= min_len && code != 5 && byte_at (ps - 1) != '\n' && byte_at (pe + 1) != '\n') {
          bool ok = true;
          u64 ctx0 = 0, ctx1 = 0;
          for (int j = 0; j < k; j++) {
            const int bl = byte_at (code < 2 ? ps - k + j : pe + k - j), br = byte_at (code < 2 ? pe + 1 + j : ps - 1 - j);
            const int cl = (code < 2) ? s_lutf[bl] : s_lutr[bl], cr = (code < 2) ? s_lutf[br] : s_lutr[br];
            ok = ok && (cl != 5) && (cr != 5);
            ctx0 |= (u64) (cl & 3) << (2 * j);
            ctx1 |= (u64) (cr & 3) << (2 * j);
          }
          if (ok) {
            if (code == 4) nskip++; // non-ACGTU run: undefined in the reference, fenced
            else {
              Rec<W> q = pack_rec<W> (ctx0, ctx1, code < 2 ? code : 3 - code, sext_field (len, 10), code < 2 ? 1 : 2, k);
              atomicAdd (&bin_hist[rec_bin<W> (q, k, bin_bits)], 1u);
              const unsigned long long pos = atomicAdd (&counters[0], 1ULL); // rare path: straight to global
              if (pos < capacity) store_rec<W> (out, pos, q);
            }
          }
        }
        ps = pe + 1;
      }
    }
    __syncwarp ();
    if (nrec > SW_RCAP - 32) flush ();
  };

  // ---- main loop: this warp's tiles, tile t+n_warps in flight while tile t is processed ----------------------
  long long t = (long long) blockIdx.x * SCAN_WARPS + (tid >> 5);
  int st = 0;
  if (t < n_tiles) stage_tile (t, 0);
  for (; t < n_tiles; t += n_warps, st ^= 1) {
    if (t + n_warps < n_tiles) stage_tile (t + n_warps, st ^ 1);
    if (interior (t)) { mbar_wait (&S->bar[st], (parity >> st) & 1); parity ^= 1u << st; }
    else __syncwarp ();
    const uint8_t *raw = S->raw[st];
    const u32 *raw32 = reinterpret_cast<const u32*> (raw);
    const long long g0 = t * SW_TILE - SW_HALO_L;
    u32 nq = 0;
    for (int row = 0; row < SW_ROWS; row++) {
      // ---- phase A: gray codes of my 16 bytes, code-run starts ---------------------------------------------
      const int grp = SW_HALO_L / 16 + row * 32 + lane;
      const uint4 v = *reinterpret_cast<const uint4*> (raw + 16 * grp);
      const u32 p0 = (v.x & 0x06060606u) * SCAN_GRAY_MUL, p1 = (v.y & 0x06060606u) * SCAN_GRAY_MUL;
      const u32 p2 = (v.z & 0x06060606u) * SCAN_GRAY_MUL, p3 = (v.w & 0x06060606u) * SCAN_GRAY_MUL;
      const u32 c = __byte_perm (__byte_perm (p0, p1, 0x0073), __byte_perm (p2, p3, 0x0073), 0x5410);
      const u32 wprev = raw32[4 * grp - 1], wnext = raw32[4 * grp + 4];
      const u32 cn = ((wnext & 0x00060606u) * SCAN_GRAY_MUL) >> 24; // next three bases
      const u32 cpw = (wprev << 5) & 0xC0000000u;                    // previous base, in the top two bits
      u32 tt = c ^ __funnelshift_r (c, cn, 2);
      if (cand_len >= 3) tt |= c ^ __funnelshift_r (c, cn, 4);
      if (cand_len >= 4) tt |= c ^ __funnelshift_r (c, cn, 6);
      const u32 dp = c ^ __funnelshift_l (cpw, c, 2);
      u32 cs = ~(tt | (tt >> 1)) & (dp | (dp >> 1)) & 0x55555555u;
      // ballot compaction into the warp's queue (a lane rarely holds more than one start)
      for (;;) {
        const u32 m = __ballot_sync (0xffffffffu, cs != 0);
        if (!m) break;
        if (cs) {
          S->cand[nq + __popc (m & lt_mask)] = (unsigned short) (16 * grp + ((__ffs (cs) - 1) >> 1));
          cs &= cs - 1;
        }
        nq += __popc (m);
      }
      __syncwarp ();
      while (nq >= 32) { nq -= 32; process_batch ((int) S->cand[nq + lane], raw, g0); }
    }
    if (nq) process_batch (lane < (int) nq ? (int) S->cand[lane] : -1, raw, g0);
    __syncwarp (); // every lane is done with this stage before it is refilled
  }
  if (nrec) flush ();
#pragma unroll
  for (int d = 16; d; d >>= 1) nskip += __shfl_xor_sync (0xffffffffu, nskip, d);
  if (lane == 0 && nskip) atomicAdd (&counters[1], (unsigned long long) nskip);
}
