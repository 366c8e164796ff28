/* CPU restatement of tatajuba's hot path — see tjb_oracle.h (TEST INFRASTRUCTURE ONLY). */
#include <stdlib.h>
#include <string.h>
#include "tjb_oracle.h"

/* 2-bit codes, hopo_counter.c:205-216: A0 C1 G2 T/U3 (either case), anything else 4 */
static int code_fwd (uint8_t c)
{
  switch (c) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': case 'U': case 'u': return 3;
    default: return 4;
  }
}
static int code_rev (uint8_t c) { int f = code_fwd (c); return f < 4 ? 3 - f : 4; }

static int32_t sext (int64_t v, int bits) /* value of a signed bit-field of that width */
{
  int64_t m = 1LL << (bits - 1);
  v &= (1LL << bits) - 1;
  return (int32_t) ((v ^ m) - m);
}

typedef struct { tjbo_rec *v; int64_t n, cap; } recvec;

static void push (recvec *r, tjbo_rec x)
{
  if (r->n == r->cap) { r->cap = r->cap ? 2 * r->cap : 1024; r->v = (tjbo_rec*) realloc (r->v, (size_t) r->cap * sizeof (tjbo_rec)); }
  r->v[r->n++] = x;
}

/* one read; hopo_counter.c:219-258 restated as "maximal byte runs [s,e] with len >= max(m,2), s >= k, e <= L-k-1" */
static void scan_read (const uint8_t *seq, int64_t L, int k, int m, recvec *out, int64_t *skipped)
{
  int min_len = m < 2 ? 2 : m; /* the length test only happens in the "same as previous" branch (:227-229) */
  int64_t s = 0;
  tjbo_rec last; /* context[], hopo_base_int and reverse_forward_flag of the last tract emitted for this read (:222) */
  int have_last = 0;
  while (s < L) {
    int64_t e = s;
    while (e + 1 < L && seq[e + 1] == seq[s]) e++;
    int64_t len = e - s + 1;
    if (len >= min_len && s >= k && e <= L - k - 1) {
      int f = code_fwd (seq[s]), r = code_rev (seq[s]);
      tjbo_rec x;
      x.ctx[0] = x.ctx[1] = 0; x.count = 1; x.length = sext (len, 10);
      if (f < r) {        /* A or C: forward (:233-238) */
        for (int j = 0; j < k; j++) {
          x.ctx[0] |= (uint64_t) (code_fwd (seq[s - k + j]) & 3) << (2 * j);
          x.ctx[1] |= (uint64_t) (code_fwd (seq[e + 1 + j]) & 3) << (2 * j);
        }
        x.base = f; x.flag = 1;
        push (out, x);
        last = x; have_last = 1;
      } else if (f > r) { /* G or T/U: reverse complement (:239-246) */
        for (int j = 0; j < k; j++) {
          x.ctx[0] |= (uint64_t) (code_rev (seq[e + k - j]) & 3) << (2 * j);
          x.ctx[1] |= (uint64_t) (code_rev (seq[s - 1 - j]) & 3) << (2 * j);
        }
        x.base = r; x.flag = 2;
        push (out, x);
        last = x; have_last = 1;
      } else if (have_last) {
        /* run of a byte that is none of ACGTU (e.g. NNNN): neither branch is taken and add_kmer_to_hopo_counter() is
         * called with the variables of the previous tract of this read (:233-248): its context, base and flag, with
         * the length of THIS run */
        x = last; x.length = sext (len, 10);
        push (out, x);
      } else (*skipped)++; /* no earlier tract in this read: the reference reads an uninitialised context[] and
                              hopo_base_int (undefined behaviour); fenced: counted and skipped */
    }
    s = e + 1;
  }
}

int64_t
tjbo_scan_buffer (const uint8_t *buf, size_t n_bytes, int k, int m, tjbo_rec **recs, int64_t *n_skipped)
{
  recvec r = {NULL, 0, 0};
  int64_t skipped = 0;
  size_t start = 0;
  while (start < n_bytes) {
    const uint8_t *nl = (const uint8_t*) memchr (buf + start, '\n', n_bytes - start);
    size_t end = nl ? (size_t) (nl - buf) : n_bytes;
    if (end > start) scan_read (buf + start, (int64_t) (end - start), k, m, &r, &skipped);
    start = end + 1;
  }
  *recs = r.v;
  if (n_skipped) *n_skipped = skipped;
  return r.n;
}

/* hopo_counter.c:28-38, descending */
static int cmp_rec_desc (const void *pa, const void *pb)
{
  const tjbo_rec *a = (const tjbo_rec*) pa, *b = (const tjbo_rec*) pb;
  if (a->base != b->base) return b->base - a->base;
  if (a->ctx[0] != b->ctx[0]) return b->ctx[0] > a->ctx[0] ? 1 : -1;
  if (a->ctx[1] != b->ctx[1]) return b->ctx[1] > a->ctx[1] ? 1 : -1;
  return b->length - a->length;
}
static int same_context (const tjbo_rec *a, const tjbo_rec *b)
{ return a->base == b->base && a->ctx[0] == b->ctx[0] && a->ctx[1] == b->ctx[1]; }

typedef struct { int32_t key, w; } kw;
static int cmp_kw (const void *a, const void *b)
{ int32_t x = ((const kw*) a)->key, y = ((const kw*) b)->key; return (x > y) - (x < y); }

int
tjbo_count_buffer (const uint8_t *buf, size_t n_bytes, int k, int m, int min_coverage, int remove_biased, tjbo_result *out)
{
  memset (out, 0, sizeof (*out));
  tjbo_rec *raw = NULL;
  int64_t n = tjbo_scan_buffer (buf, n_bytes, k, m, &raw, &out->n_skipped_nonacgt);
  out->n_raw = n;
  if (!n) { free (raw); return out->status = 1; }                       /* :345-349 */

  qsort (raw, (size_t) n, sizeof (tjbo_rec), cmp_rec_desc);             /* :351 */
  int64_t nd = 0;                                                       /* :352-365 */
  for (int64_t i = 1; i < n; i++) {
    if (cmp_rec_desc (&raw[nd], &raw[i])) raw[++nd] = raw[i];
    else { raw[nd].count = sext ((int64_t) raw[nd].count + 1, 20); raw[nd].flag |= raw[i].flag; }
  }
  nd++;
  out->n_distinct = nd;

  int64_t nk = 0;                                                       /* :367-374 */
  for (int64_t i = 0; i < nd; i++)
    if (remove_biased ? raw[i].flag == 3 : raw[i].count > 1) raw[nk++] = raw[i];
  if (!nk) { free (raw); return out->status = 2; }                      /* :376-381 */

  out->n_elem = (int32_t) nk;
  out->elem = (tjbo_rec*) realloc (raw, (size_t) nk * sizeof (tjbo_rec));
  out->idx_initial = (int32_t*) malloc ((size_t) nk * sizeof (int32_t));
  out->idx_final   = (int32_t*) malloc ((size_t) nk * sizeof (int32_t));
  int32_t ng = 0;                                                       /* :388-404 */
  for (int64_t g0 = 0; g0 < nk; ) {
    int64_t g1 = g0 + 1, depth = out->elem[g0].count;
    while (g1 < nk && same_context (&out->elem[g0], &out->elem[g1])) depth += out->elem[g1++].count;
    if ((int32_t) depth >= min_coverage) { out->idx_initial[ng] = (int32_t) g0; out->idx_final[ng] = (int32_t) g1; ng++; }
    g0 = g1;
  }
  out->n_idx = ng;
  if (!ng) { /* :406-411; like the reference, nothing of the counter is used afterwards */
    free (out->elem); free (out->idx_initial); free (out->idx_final);
    out->elem = NULL; out->idx_initial = out->idx_final = NULL; out->n_elem = 0;
    return out->status = 3;
  }

  /* coverage, :420-438: pool the low 31 bits of both flanks of ALL kept records, weight = count */
  kw *p = (kw*) malloc ((size_t) (2 * nk) * sizeof (kw));
  for (int64_t i = 0; i < nk; i++) {
    p[i].key      = (int32_t) (out->elem[i].ctx[0] & 0x7fffffffULL); p[i].w      = out->elem[i].count;
    p[nk + i].key = (int32_t) (out->elem[i].ctx[1] & 0x7fffffffULL); p[nk + i].w = out->elem[i].count;
  }
  qsort (p, (size_t) (2 * nk), sizeof (kw), cmp_kw);
  int32_t best = 0, have = 0;
  for (int64_t i = 0; i < 2 * nk; ) {
    int32_t sum = 0; int64_t j = i;
    while (j < 2 * nk && p[j].key == p[i].key) sum += p[j++].w;
    if (!have || sum > best) { best = sum; have = 1; }
    i = j;
  }
  free (p);
  out->coverage = best;
  return out->status = 0;
}

void tjbo_free_result (tjbo_result *r)
{
  if (!r) return;
  free (r->elem); free (r->idx_initial); free (r->idx_final);
  r->elem = NULL; r->idx_initial = r->idx_final = NULL;
}

void tjbo_free (void *p) { free (p); }

void
tjbo_pack_hopo_elements (const tjbo_rec *recs, int64_t n, void *out40)
{
  uint8_t *o = (uint8_t*) out40;
  for (int64_t i = 0; i < n; i++, o += 40) {
    /* gcc/x86-64 bit-field allocation from the LSB: base:2 length:10 count:20 mismatches:12 multi:3 neg_strand:2 canon_flag:3 */
    uint64_t bits = ((uint64_t) recs[i].base & 3) | (((uint64_t) recs[i].length & 0x3ff) << 2) |
                    (((uint64_t) recs[i].count & 0xfffff) << 12) | (0xffeULL << 32) | (((uint64_t) recs[i].flag & 7) << 49);
    int32_t tail[4] = {-1, -1, -1, -1}; /* read_offset (reset at :511), loc_ref_id, loc_pos, loc_last */
    memcpy (o, recs[i].ctx, 16); memcpy (o + 16, &bits, 8); memcpy (o + 24, tail, 16);
  }
}
