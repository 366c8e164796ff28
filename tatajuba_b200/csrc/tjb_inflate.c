/* Block-streaming gzip decoder for the file feed (".gz" samples).
 *
 * The reference reads its samples through zlib's gzread() under the klib parser (src/hopo_counter.c:141-153,
 * src/kseq.h): one thread per sample inflates at a few hundred MB/s, which is what bounds a GPU that counts 10^11
 * bases/s.  This is our own DEFLATE decoder (RFC 1951 / RFC 1952, written from the format description): the compressed
 * file is mapped, so the input never has to be suspended; a 64-bit bit buffer is refilled once per two or three symbols;
 * literal/length codes are looked up in one 11-bit table (sub-tables behind it for longer codes), distances in an 8-bit
 * one; matches are copied eight bytes at a time into an output buffer that has room to overshoot.  A producer thread
 * decodes into a ring of buffers, each starting with the 32 KiB history of the one before; the consumer (the FASTQ
 * parser's thread) checks CRC-32 and length of every member (zlib's crc32(): the only zlib function left on the path)
 * while the producer is already inflating the next buffer.  Concatenated members (bgzip) are supported; what follows
 * the last member is ignored like gzread() does.
 */
#include <fcntl.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>
#include "tjb_gpu.h"

#define LIT_BITS 11                 /* primary table of the literal/length code */
#define DIST_BITS 8                 /* primary table of the distance code */
#define LIT_TABLE ((1 << LIT_BITS) + 1024)
#define DIST_TABLE ((1 << DIST_BITS) + 512)
#define WINDOW 32768u
#ifndef GZ_PAYLOAD_LOG
#define GZ_PAYLOAD_LOG 22
#endif
#define GZ_PAYLOAD (1u << GZ_PAYLOAD_LOG) /* decoded bytes per buffer */
#define GZ_SLACK 320u               /* one loop iteration never writes more than this past the soft end */
#define GZ_NBUF 3
#define TAIL_BYTES 32

/* table entry: bits 0-7 bits to consume (code + extra bits; sub-table pointer: the primary bits), bits 8-11 bits of the
 * code alone (pointer: index bits of the sub-table), bit 15 literal, bit 14 end of block, bit 13 pointer, bits 16-31
 * value (literal, base length, base distance, sub-table start).  0 = no such code. */
#define E_LIT 0x8000u
#define E_EOB 0x4000u
#define E_SUB 0x2000u

static const uint16_t kLenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
static const uint8_t kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
static const uint16_t kDistBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
static const uint8_t kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
static const uint8_t kPrecodeOrder[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};

enum { ST_MEMBER = 0, ST_BLOCK, ST_STORED, ST_HUFF, ST_TRAILER, ST_END, ST_CORRUPT };

typedef struct {
  const uint8_t *in, *in_end;   /* mapped file */
  uint64_t bb;                  /* bit buffer: the next bits of the stream, lowest first */
  unsigned bn;                  /* how many of them count */
  int state, last_block, members;
  uint32_t stored_left;
  uint32_t lit[LIT_TABLE], dist[DIST_TABLE];
  uint32_t crc_expect, size_expect;
  const uint8_t *tail_src;      /* the last bytes of the file ... */
  uint8_t tail[TAIL_BYTES + 32];/* ... copied in front of zeros, so that the decoder can load 8 bytes at any position */
  int tail_ready;
} inflater;

static inline uint64_t load64 (const uint8_t *p) { uint64_t v; memcpy (&v, p, 8); return v; }
static inline void store64 (uint8_t *p, uint64_t v) { memcpy (p, &v, 8); }
#define LOW64(x, n) ((n) >= 64 ? (x) : ((x) & ((1ULL << (n)) - 1)))

/* byte-wise refill for headers and for the last bytes of the file: past the end it feeds zeros (the caller notices
 * when more bits were used than the file had: `in` runs past `in_end` by more than the bit buffer holds) */
static inline void refill_slow (inflater *s)
{
  while (s->bn <= 56) {
    const uint64_t b = s->in < s->in_end ? *s->in : 0;
    s->in++;
    s->bb |= b << s->bn;
    s->bn += 8;
  }
}
static inline uint32_t take_bits (inflater *s, unsigned n) /* n <= 32 */
{
  if (s->bn < n) refill_slow (s);
  const uint32_t v = (uint32_t) (s->bb & ((1ULL << n) - 1));
  s->bb >>= n; s->bn -= n;
  return v;
}
/* drop the rest of the current byte and give whole bytes back to the input */
static inline void byte_align (inflater *s)
{
  const unsigned drop = s->bn & 7;
  s->bb >>= drop; s->bn -= drop;
  s->in -= s->bn >> 3;
  s->bb = 0; s->bn = 0;
}
static inline int input_overrun (const inflater *s) { return s->in > s->in_end + (s->bn >> 3); }

static inline uint32_t reverse_bits (uint32_t code, int len)
{
  uint32_t r = 0;
  for (int i = 0; i < len; i++) { r = (r << 1) | (code & 1); code >>= 1; }
  return r;
}

enum { KIND_PRECODE = 0, KIND_LITLEN, KIND_DIST };
static inline uint32_t make_entry (int kind, int sym, int bits)
{
  if (kind == KIND_PRECODE) return ((uint32_t) sym << 16) | (uint32_t) bits;
  if (kind == KIND_DIST) return sym < 30 ? ((uint32_t) kDistBase[sym] << 16) | ((uint32_t) bits << 8) | (uint32_t) (bits + kDistExtra[sym]) : 0u;
  if (sym < 256) return E_LIT | ((uint32_t) sym << 16) | (uint32_t) bits;
  if (sym == 256) return E_EOB | (uint32_t) bits;
  return sym < 286 ? ((uint32_t) kLenBase[sym - 257] << 16) | ((uint32_t) bits << 8) | (uint32_t) (bits + kLenExtra[sym - 257]) : 0u;
}

/* canonical Huffman code of `n_sym` code lengths (0 = unused, <= 15) -> lookup table indexed by the next bits of the
 * stream.  Returns 0, or -1 for a code that is over-subscribed, or incomplete with more than one symbol (the rule zlib
 * applies, so the same files are accepted and rejected). */
static int build_table (uint32_t *table, int capacity, int root, const uint8_t *lens, int n_sym, int kind)
{
  int count[16] = {0}, offs[16], left_at[16];
  uint16_t sorted[288];
  for (int i = 0; i < n_sym; i++) count[lens[i]]++;
  memset (table, 0, sizeof (uint32_t) << root);
  if (count[0] == n_sym) return 0; /* no code at all: every lookup is an error */
  int left = 1, max = 0;
  for (int l = 1; l <= 15; l++) {
    left = (left << 1) - count[l];
    if (left < 0) return -1;
    if (count[l]) max = l;
  }
  if (left > 0 && max != 1) return -1;
  offs[1] = 0;
  for (int l = 1; l < 15; l++) offs[l + 1] = offs[l] + count[l];
  for (int i = 0; i < n_sym; i++) if (lens[i]) sorted[offs[lens[i]]++] = (uint16_t) i;
  for (int l = 0; l <= 15; l++) left_at[l] = count[l]; /* codes of each length not handed out yet */

  uint32_t code = 0;
  int k = 0, next_sub = 1 << root, sub_start = 0, sub_bits = 0;
  uint32_t sub_prefix = 0xffffffffu;
  for (int l = 1; l <= 15; l++, code <<= 1) {
    for (int c = 0; c < count[l]; c++, k++, code++) {
      const int sym = sorted[k];
      const uint32_t rev = reverse_bits (code, l);
      if (l <= root) {
        const uint32_t e = make_entry (kind, sym, l);
        for (uint32_t i = rev; i < (1u << root); i += 1u << l) table[i] = e;
      } else {
        const uint32_t prefix = rev & ((1u << root) - 1);
        if (prefix != sub_prefix) { /* codes that share their first `root` bits are neighbours in canonical order */
          int bits = l - root, room = 1 << bits;
          while (bits + root < 15) { /* as many index bits as the longest code below this prefix needs */
            room -= (bits + root == l) ? left_at[l] : count[bits + root];
            if (room <= 0) break;
            bits++; room <<= 1;
          }
          if (next_sub + (1 << bits) > capacity) return -1;
          sub_prefix = prefix; sub_start = next_sub; sub_bits = bits; next_sub += 1 << bits;
          memset (table + sub_start, 0, sizeof (uint32_t) << bits);
          table[prefix] = E_SUB | ((uint32_t) sub_start << 16) | ((uint32_t) bits << 8) | (uint32_t) root;
        }
        const uint32_t e = make_entry (kind, sym, l - root);
        for (uint32_t i = rev >> root; i < (1u << sub_bits); i += 1u << (l - root)) table[sub_start + i] = e;
      }
      left_at[l]--;
    }
  }
  return 0;
}

static void fixed_tables (inflater *s)
{
  uint8_t lens[288 + 32];
  int i = 0;
  for (; i < 144; i++) lens[i] = 8;
  for (; i < 256; i++) lens[i] = 9;
  for (; i < 280; i++) lens[i] = 7;
  for (; i < 288; i++) lens[i] = 8;
  build_table (s->lit, LIT_TABLE, LIT_BITS, lens, 288, KIND_LITLEN);
  for (i = 0; i < 32; i++) lens[i] = 5;
  build_table (s->dist, DIST_TABLE, DIST_BITS, lens, 32, KIND_DIST);
}

/* header of a dynamic block (RFC 1951 3.2.7): the two codes, themselves coded by the "code length" code */
static int dynamic_tables (inflater *s)
{
  const int n_lit = (int) take_bits (s, 5) + 257, n_dist = (int) take_bits (s, 5) + 1, n_pre = (int) take_bits (s, 4) + 4;
  if (n_lit > 286 || n_dist > 30) return -1;
  uint8_t pre_lens[19] = {0}, lens[286 + 30 + 138];
  for (int i = 0; i < n_pre; i++) pre_lens[kPrecodeOrder[i]] = (uint8_t) take_bits (s, 3);
  uint32_t pre[1 << 7];
  if (build_table (pre, 1 << 7, 7, pre_lens, 19, KIND_PRECODE) < 0) return -1;
  int n = 0;
  while (n < n_lit + n_dist) {
    if (s->bn < 7 + 7) refill_slow (s);
    const uint32_t e = pre[s->bb & 127];
    if (!e) return -1;
    s->bb >>= (e & 0xff); s->bn -= (e & 0xff);
    const int sym = (int) (e >> 16);
    if (sym < 16) { lens[n++] = (uint8_t) sym; continue; }
    int rep, val = 0;
    if (sym == 16) { if (!n) return -1; val = lens[n - 1]; rep = 3 + (int) take_bits (s, 2); }
    else if (sym == 17) rep = 3 + (int) take_bits (s, 3);
    else rep = 11 + (int) take_bits (s, 7);
    if (n + rep > n_lit + n_dist) return -1;
    memset (lens + n, val, (size_t) rep);
    n += rep;
  }
  if (!lens[256]) return -1; /* a block cannot end without its end-of-block code */
  if (build_table (s->lit, LIT_TABLE, LIT_BITS, lens, n_lit, KIND_LITLEN) < 0) return -1;
  if (build_table (s->dist, DIST_TABLE, DIST_BITS, lens + n_lit, n_dist, KIND_DIST) < 0) return -1;
  return input_overrun (s) ? -1 : 0;
}

/* Decode symbols of the current Huffman block into out[] until the block ends (returns 1), the output reaches
 * `soft_end` (returns 0; never more than GZ_SLACK bytes past it) or the data is wrong (-1).  `base` is the oldest byte
 * a match may reach.
 * The bit buffer is refilled by one unaligned 8-byte load (the last TAIL_BYTES of the file are decoded from a copy
 * followed by zeros, so the load is always legal); after it all 64 bits of `bb` are stream bits although only `bn` >= 56
 * are counted, so the table entry of the NEXT symbol can be fetched from `bb` right after a symbol is consumed, before
 * the refill: the serial chain per symbol is one table load and one shift (extra bits are cut out of a copy of `bb`). */
static inline __attribute__ ((always_inline)) int decode_block_body (inflater *s, const uint8_t *base, uint8_t **out_io, uint8_t *soft_end)
{
  if (!s->tail_ready) {
    const size_t n = (size_t) (s->in_end - s->in) < TAIL_BYTES ? (size_t) (s->in_end - s->in) : TAIL_BYTES;
    s->tail_src = s->in_end - n;
    memset (s->tail, 0, sizeof s->tail);
    memcpy (s->tail, s->tail_src, n);
    s->tail_ready = 1;
  }
  const uint8_t *in = s->in, *in_limit = s->in_end - 16;
  int in_tail = 0;
  uint64_t bb = s->bb;
  unsigned bn = s->bn;
  uint8_t *out = *out_io, *out_safe = *out_io;
  const uint32_t *const lit = s->lit, *const dtab = s->dist;
  int rc = 0;
#define LIT_MASK ((1u << LIT_BITS) - 1)
#define LOW(x, n) ((x) & ((1ULL << (n)) - 1))
#define REFILL() do { \
    if (in > in_limit) { \
      if (in_tail || in < s->tail_src) { rc = -1; goto done; } /* past the padding: more bits used than the file has */ \
      in_tail = 1; in = s->tail + (in - s->tail_src); in_limit = s->tail + sizeof s->tail - 8; out_safe = out; \
    } \
    bb |= load64 (in) << bn; in += (63 - bn) >> 3; bn |= 56; } while (0)
#define CONSUME(n) do { bb >>= (n); bn -= (n); } while (0)
#define TAIL_OVERRUN() ((size_t) (in - s->tail) > (size_t) (s->in_end - s->tail_src) + (bn >> 3))
  {
    REFILL ();
    uint32_t e = lit[bb & LIT_MASK];
    while (out < soft_end) {
      if (in_tail) { /* the last bytes of the file: nothing decoded from the zeros behind them is kept */
        if (TAIL_OVERRUN ()) goto done;
        out_safe = out;
      }
      REFILL (); /* off the chain: `e` is known already, the load address does not depend on it */
      if (e & E_LIT) { /* up to three literals per refill: 3 x 15 bits <= 56 */
        CONSUME (e & 0xff); *out++ = (uint8_t) (e >> 16);
        e = lit[bb & LIT_MASK];
        if (e & E_LIT) {
          CONSUME (e & 0xff); *out++ = (uint8_t) (e >> 16);
          e = lit[bb & LIT_MASK];
          if (e & E_LIT) { CONSUME (e & 0xff); *out++ = (uint8_t) (e >> 16); e = lit[bb & LIT_MASK]; continue; }
        }
        REFILL (); /* at least 26 bits were left; what follows may need 48 */
      }
      if (e & E_SUB) {
        CONSUME (e & 0xff);
        e = lit[(e >> 16) + LOW (bb, (e >> 8) & 15)];
        if (e & E_LIT) { CONSUME (e & 0xff); *out++ = (uint8_t) (e >> 16); e = lit[bb & LIT_MASK]; continue; }
      }
      if (e & E_EOB) { CONSUME (e & 0xff); rc = 1; goto done; }
      if (!e) { rc = -1; goto done; }
      const uint64_t len_bits = bb;
      CONSUME (e & 0xff);
      const unsigned len = (e >> 16) + (unsigned) (LOW (len_bits, e & 0xff) >> ((e >> 8) & 15));
      uint32_t d = dtab[bb & ((1u << DIST_BITS) - 1)];
      if (d & E_SUB) { CONSUME (d & 0xff); d = dtab[(d >> 16) + LOW (bb, (d >> 8) & 15)]; }
      if (!d) { rc = -1; goto done; }
      const uint64_t dist_bits = bb;
      CONSUME (d & 0xff);
      e = lit[bb & LIT_MASK]; /* the next symbol's entry, while this match is copied */
      const size_t dist = (d >> 16) + (size_t) (LOW (dist_bits, d & 0xff) >> ((d >> 8) & 15));
      if (dist > (size_t) (out - base)) { rc = -1; goto done; }
      const uint8_t *src = out - dist;
      uint8_t *const end = out + len;
      if (dist >= 8) { /* sixteen bytes without asking (most matches of sequence data are shorter), then the rest */
        store64 (out, load64 (src)); store64 (out + 8, load64 (src + 8));
        if (len > 16) { out += 16; src += 16; do { store64 (out, load64 (src)); out += 8; src += 8; } while (out < end); }
      } else if (dist == 1) { /* a run of one byte (the quality lines of simulated reads, poly-N stretches) */
        const uint64_t v = 0x0101010101010101ULL * src[0];
        do { store64 (out, v); out += 8; } while (out < end);
      } else {
        do { *out++ = *src++; } while (out < end);
      }
      out = end;
    }
  }
done:
  if (in_tail && TAIL_OVERRUN ()) { out = out_safe; rc = -1; }
#undef TAIL_OVERRUN
#undef REFILL
#undef CONSUME
#undef LOW
#undef LIT_MASK
  s->in = in_tail ? s->tail_src + (in - s->tail) : in;
  s->bb = LOW64 (bb, bn); s->bn = bn; /* the uncounted bits go: the byte-wise reader ORs onto zeros */
  *out_io = out;
  if (rc >= 0 && input_overrun (s)) rc = -1;
  return rc;
}
static int decode_block_generic (inflater *s, const uint8_t *base, uint8_t **out_io, uint8_t *soft_end)
{ return decode_block_body (s, base, out_io, soft_end); }
#if defined(__x86_64__) && defined(__GNUC__)
__attribute__ ((target ("bmi2"))) static int decode_block_bmi2 (inflater *s, const uint8_t *base, uint8_t **out_io, uint8_t *soft_end)
{ return decode_block_body (s, base, out_io, soft_end); }
#endif
static int decode_block (inflater *s, const uint8_t *base, uint8_t **out_io, uint8_t *soft_end)
{
#if defined(__x86_64__) && defined(__GNUC__)
  static int has_bmi2 = -1;
  if (has_bmi2 < 0) has_bmi2 = __builtin_cpu_supports ("bmi2") ? 1 : 0;
  if (has_bmi2) return decode_block_bmi2 (s, base, out_io, soft_end);
#endif
  return decode_block_generic (s, base, out_io, soft_end);
}

/* gzip member header (RFC 1952 2.3); returns 0, -1 = not a gzip member / truncated */
static int member_header (inflater *s)
{
  const uint8_t *p = s->in, *const e = s->in_end;
  if (e - p < 18 || p[0] != 0x1f || p[1] != 0x8b || p[2] != 8 || (p[3] & 0xe0)) return -1;
  const int flags = p[3];
  p += 10;
  if (flags & 4) { if (e - p < 2) return -1; const size_t n = p[0] | ((size_t) p[1] << 8); p += 2; if ((size_t) (e - p) < n) return -1; p += n; }
  for (int f = 8; f <= 16; f <<= 1) if (flags & f) { /* file name, comment: zero terminated */
    const uint8_t *z = (const uint8_t*) memchr (p, 0, (size_t) (e - p));
    if (!z) return -1;
    p = z + 1;
  }
  if (flags & 2) { if (e - p < 2) return -1; p += 2; }
  s->in = p; s->bb = 0; s->bn = 0;
  return 0;
}

/* Inflate into out[0 .. capacity) (GZ_SLACK more bytes must be writable) with out[-history .. 0) as the window so far.
 * Returns the bytes produced.  *member_end is set when a member's trailer was reached (crc_expect / size_expect hold
 * it): the call stops there.  Wrong data leaves the decoder in ST_CORRUPT (what was decoded before it is still
 * returned, as gzread() would); ST_END = nothing follows the last member. */
static size_t inflate_some (inflater *s, uint8_t *out, size_t history, size_t capacity, int *member_end)
{
  uint8_t *const out0 = out, *const soft_end = out + capacity;
  const uint8_t *const base = out - history;
  *member_end = 0;
#define CORRUPT() do { s->state = ST_CORRUPT; return (size_t) (out - out0); } while (0)
  for (;;) {
    switch (s->state) {
    case ST_MEMBER:
      if (s->in >= s->in_end || member_header (s) < 0) { /* end of file, or something that is no gzip member: ignored */
        if (!s->members) CORRUPT ();
        s->state = ST_END;
        return (size_t) (out - out0);
      }
      s->members++;
      s->state = ST_BLOCK;
      break;
    case ST_BLOCK: {
      if (s->last_block) { s->state = ST_TRAILER; break; }
      s->last_block = (int) take_bits (s, 1);
      const int type = (int) take_bits (s, 2);
      if (type == 0) {
        byte_align (s);
        if (s->in > s->in_end || s->in_end - s->in < 4) CORRUPT ();
        const uint32_t len = s->in[0] | ((uint32_t) s->in[1] << 8), nlen = s->in[2] | ((uint32_t) s->in[3] << 8);
        if ((len ^ nlen) != 0xffffu) CORRUPT ();
        s->in += 4;
        if ((size_t) (s->in_end - s->in) < len) CORRUPT ();
        s->stored_left = len; s->state = ST_STORED;
      } else if (type == 1) { fixed_tables (s); s->state = ST_HUFF; }
      else if (type == 2) { if (dynamic_tables (s) < 0) CORRUPT (); s->state = ST_HUFF; }
      else CORRUPT ();
      break;
    }
    case ST_STORED: {
      const size_t room = out < soft_end ? (size_t) (soft_end - out) : 0, n = s->stored_left < room ? s->stored_left : room;
      memcpy (out, s->in, n);
      out += n; s->in += n; s->stored_left -= (uint32_t) n;
      if (s->stored_left) return (size_t) (out - out0); /* buffer full */
      s->state = ST_BLOCK;
      break;
    }
    case ST_HUFF: {
      if (out >= soft_end) return (size_t) (out - out0);
      const int rc = decode_block (s, base, &out, soft_end);
      if (rc < 0) CORRUPT ();
      if (rc == 1) s->state = ST_BLOCK;
      break;
    }
    case ST_TRAILER:
      byte_align (s);
      if (s->in > s->in_end || s->in_end - s->in < 8) CORRUPT ();
      memcpy (&s->crc_expect, s->in, 4); memcpy (&s->size_expect, s->in + 4, 4); /* little endian, like every host here */
      s->in += 8;
      s->last_block = 0;
      s->state = ST_MEMBER;
      *member_end = 1;
      return (size_t) (out - out0);
    default:
      return (size_t) (out - out0);
    }
  }
#undef CORRUPT
}

/* ---- the stream: a producer thread inflating into a ring of buffers ------------------------------------------ */
typedef struct {
  uint8_t *mem;          /* WINDOW bytes of history, GZ_PAYLOAD of payload, GZ_SLACK */
  size_t n;              /* payload bytes */
  int full, member_end, last; /* last: nothing follows (end of file or corrupt data) */
  int corrupt;
  uint32_t crc, size;
} gz_buffer;

struct tjb_gzstream {
  inflater inf;
  void *map; size_t map_len;
  gz_buffer buf[GZ_NBUF];
  pthread_t thread;
  pthread_mutex_t mu;
  pthread_cond_t cv;
  int stop, started;
  int take;              /* next buffer the consumer gets */
  int held;              /* buffer the consumer is reading, or -1 */
  unsigned long crc_run; /* of the current member, so far */
  uint32_t size_run;
  int finished;
};

static void *gz_producer (void *arg)
{
  tjb_gzstream *g = (tjb_gzstream*) arg;
  size_t history = 0;
  gz_buffer *prev = NULL;
  for (int i = 0;; i = (i + 1) % GZ_NBUF) {
    gz_buffer *b = &g->buf[i];
    pthread_mutex_lock (&g->mu);
    while (b->full && !g->stop) pthread_cond_wait (&g->cv, &g->mu);
    const int stop = g->stop;
    pthread_mutex_unlock (&g->mu);
    if (stop) return NULL;
    if (history) memcpy (b->mem + WINDOW - history, prev->mem + WINDOW + prev->n - history, history);
    int member_end = 0;
    const size_t n = inflate_some (&g->inf, b->mem + WINDOW, history, GZ_PAYLOAD, &member_end);
    const int last = g->inf.state == ST_END || g->inf.state == ST_CORRUPT;
    history = member_end ? 0 : (history + n < WINDOW ? history + n : WINDOW);
    prev = b;
    pthread_mutex_lock (&g->mu);
    b->n = n; b->member_end = member_end; b->last = last; b->corrupt = g->inf.state == ST_CORRUPT;
    b->crc = g->inf.crc_expect; b->size = g->inf.size_expect;
    b->full = 1;
    pthread_cond_broadcast (&g->cv);
    pthread_mutex_unlock (&g->mu);
    if (last) return NULL;
  }
}

tjb_gzstream *tjb_gz_open (const char *path)
{
  const int fd = open (path, O_RDONLY);
  if (fd < 0) return NULL;
  struct stat sb;
  unsigned char magic[3];
  if (fstat (fd, &sb) != 0 || !S_ISREG (sb.st_mode) || sb.st_size < 18 || pread (fd, magic, 3, 0) != 3
      || magic[0] != 0x1f || magic[1] != 0x8b || magic[2] != 8) { close (fd); return NULL; } /* the caller falls back to zlib */
  void *map = mmap (NULL, (size_t) sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
  close (fd);
  if (map == MAP_FAILED) return NULL;
  madvise (map, (size_t) sb.st_size, MADV_SEQUENTIAL);
  tjb_gzstream *g = (tjb_gzstream*) calloc (1, sizeof (tjb_gzstream));
  if (!g) { munmap (map, (size_t) sb.st_size); return NULL; }
  g->map = map; g->map_len = (size_t) sb.st_size;
  g->inf.in = (const uint8_t*) map; g->inf.in_end = g->inf.in + g->map_len;
  g->held = -1;
  for (int i = 0; i < GZ_NBUF; i++) {
    g->buf[i].mem = (uint8_t*) malloc (WINDOW + GZ_PAYLOAD + GZ_SLACK);
    if (!g->buf[i].mem) { tjb_gz_close (g); return NULL; }
  }
  pthread_mutex_init (&g->mu, NULL);
  pthread_cond_init (&g->cv, NULL);
  if (pthread_create (&g->thread, NULL, gz_producer, g) != 0) { tjb_gz_close (g); return NULL; }
  g->started = 1;
  return g;
}

void tjb_gz_close (tjb_gzstream *g)
{
  if (!g) return;
  if (g->started) {
    pthread_mutex_lock (&g->mu);
    g->stop = 1;
    pthread_cond_broadcast (&g->cv);
    pthread_mutex_unlock (&g->mu);
    pthread_join (g->thread, NULL);
    pthread_mutex_destroy (&g->mu);
    pthread_cond_destroy (&g->cv);
  }
  for (int i = 0; i < GZ_NBUF; i++) free (g->buf[i].mem);
  if (g->map) munmap (g->map, g->map_len);
  free (g);
}

/* the next piece of decoded data (valid until the next call); 0 = end of the file, TJB_ERR_IO = corrupt or truncated
 * data, or a member whose CRC-32 / length does not match (everything decoded before it has been delivered) */
int64_t tjb_gz_next (tjb_gzstream *g, const uint8_t **data)
{
  if (!g || !data) return TJB_ERR_ARG;
  *data = NULL;
  for (;;) {
    pthread_mutex_lock (&g->mu);
    if (g->held >= 0) { g->buf[g->held].full = 0; g->held = -1; pthread_cond_broadcast (&g->cv); }
    if (g->finished) { pthread_mutex_unlock (&g->mu); return g->finished < 0 ? TJB_ERR_IO : 0; }
    gz_buffer *b = &g->buf[g->take];
    while (!b->full) pthread_cond_wait (&g->cv, &g->mu);
    g->held = g->take;
    g->take = (g->take + 1) % GZ_NBUF;
    pthread_mutex_unlock (&g->mu);
    const uint8_t *p = b->mem + WINDOW;
    if (b->n) { g->crc_run = crc32 (g->crc_run, p, (uInt) b->n); g->size_run += (uint32_t) b->n; }
    if (b->member_end) {
      if ((uint32_t) g->crc_run != b->crc || g->size_run != b->size) g->finished = -1;
      g->crc_run = 0; g->size_run = 0;
    }
    if (b->last && !g->finished) g->finished = b->corrupt ? -1 : 1;
    if (b->n) { *data = p; return (int64_t) b->n; }
    /* an empty piece (a member that ended on a buffer boundary, or an empty member): look at the next one */
  }
}
