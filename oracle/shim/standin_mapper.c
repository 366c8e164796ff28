/* Deterministic stand-in for the missing mapper (submodules/bwa, fork leomrtns/bwa @ 661e331: not in this checkout and
 * no libbwa on this image), implementing exactly the part of wrapper_bwa.h the reference's src/ uses.
 *
 * TEST INFRASTRUCTURE ONLY.  It is linked into BOTH command-line builds that oracle/Makefile makes from the reference
 * sources — the stock one and the one patched to call libtjb_gpu (integration/tatajuba-libtjb_gpu.patch) — so that
 * their output files can be compared byte for byte: "bit-exact with the reference given an identical mapper".  It is
 * not BWA: no seeds with gaps, no quality-aware scores.  What it finds, for a query = left flank (k) + tract + right
 * flank (k) and for its reverse complement, anchored on the exact first or last min(k, 12) bases:
 *   1. exact occurrences                                               nm = 0
 *   2. both flanks exact, tract of a different length in the reference   nm = | length difference |
 *   3. same length, at most 2 substitutions                            nm = substitutions
 * and reports, per query, the two best hits ordered by (nm, contig, position) — enough to exercise the reference's
 * best / leftmost / multi logic (src/hopo_counter.c:531-555).  position = 0-based leftmost reference base of the hit,
 * ref_length = reference bases covered, as the reference reads them. */
#include <pthread.h>
#include "wrapper_bwa.h"

#define SM_ANCHOR_MAX 12
#define SM_MAX_SUBST 2
#define SM_HITS_PER_QUERY 2

typedef struct {
  char *fasta;
  int n_seqs;
  char **name; int *len; long *off; /* contigs, concatenated in seq[] (upper case), off[i] = start of contig i */
  char *seq; long total;
  int a;                            /* anchor length of the index */
  uint32_t *key; int32_t *pos;      /* positions of the concatenation sorted by their a-mer (2 bits per base) */
  long n_pos;
} sm_index;

static sm_index *g_index = NULL;
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;

static int code_of (char c) { switch (c) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; default: return -1; } }

static void read_fasta (sm_index *ix, const char *fasta)
{
  FILE *f = fopen (fasta, "r");
  if (!f) biomcmc_error ("stand-in mapper: cannot open %s", fasta);
  char *line = NULL; size_t cap = 0; ssize_t n;
  long alloc = 1 << 20;
  ix->seq = (char*) malloc ((size_t) alloc); ix->total = 0; ix->n_seqs = 0;
  ix->name = NULL; ix->len = NULL; ix->off = NULL;
  while ((n = getline (&line, &cap, f)) != -1) {
    while (n > 0 && (line[n - 1] == '\n' || line[n - 1] == '\r')) line[--n] = '\0';
    if (!n) continue;
    if (line[0] == '>') {
      ix->name = (char**) realloc (ix->name, (size_t) (ix->n_seqs + 1) * sizeof (char*));
      ix->len = (int*) realloc (ix->len, (size_t) (ix->n_seqs + 1) * sizeof (int));
      ix->off = (long*) realloc (ix->off, (size_t) (ix->n_seqs + 1) * sizeof (long));
      char *sp = line + 1; size_t l = strcspn (sp, " \t");
      ix->name[ix->n_seqs] = strndup (sp, l);
      ix->len[ix->n_seqs] = 0; ix->off[ix->n_seqs] = ix->total;
      ix->n_seqs++;
    } else if (ix->n_seqs) {
      if (ix->total + n + 1 > alloc) { while (ix->total + n + 1 > alloc) alloc *= 2; ix->seq = (char*) realloc (ix->seq, (size_t) alloc); }
      for (ssize_t i = 0; i < n; i++) if (line[i] != ' ' && line[i] != '\t') { ix->seq[ix->total++] = (char) toupper ((unsigned char) line[i]); ix->len[ix->n_seqs - 1]++; }
    }
  }
  free (line); fclose (f);
  if (!ix->n_seqs) biomcmc_error ("stand-in mapper: no sequence in %s", fasta);
}

typedef struct { uint32_t key; int32_t pos; } kp;
static int cmp_kp (const void *a, const void *b)
{
  const kp *x = (const kp*) a, *y = (const kp*) b;
  if (x->key != y->key) return x->key < y->key ? -1 : 1;
  return (x->pos > y->pos) - (x->pos < y->pos);
}

static void build_index (sm_index *ix, int a)
{
  ix->a = a;
  kp *v = (kp*) malloc ((size_t) (ix->total + 1) * sizeof (kp));
  long m = 0;
  for (int c = 0; c < ix->n_seqs; c++) { /* a-mers never straddle two contigs */
    const char *s = ix->seq + ix->off[c];
    for (long p = 0; p + a <= ix->len[c]; p++) {
      uint32_t key = 0; int ok = 1;
      for (int j = 0; j < a; j++) { int cd = code_of (s[p + j]); if (cd < 0) { ok = 0; break; } key = (key << 2) | (uint32_t) cd; }
      if (ok) { v[m].key = key; v[m].pos = (int32_t) (ix->off[c] + p); m++; }
    }
  }
  qsort (v, (size_t) m, sizeof (kp), cmp_kp);
  ix->key = (uint32_t*) malloc ((size_t) (m + 1) * sizeof (uint32_t));
  ix->pos = (int32_t*) malloc ((size_t) (m + 1) * sizeof (int32_t));
  for (long i = 0; i < m; i++) { ix->key[i] = v[i].key; ix->pos[i] = v[i].pos; }
  ix->n_pos = m;
  free (v);
}

static sm_index *get_index (const char *fasta, int a)
{
  pthread_mutex_lock (&g_mu);
  if (g_index && (strcmp (g_index->fasta, fasta) != 0 || (a > 0 && g_index->a != a))) { /* another reference or k: rebuild */
    sm_index *o = g_index;
    for (int i = 0; i < o->n_seqs; i++) free (o->name[i]);
    free (o->name); free (o->len); free (o->off); free (o->seq); free (o->key); free (o->pos); free (o->fasta); free (o);
    g_index = NULL;
  }
  if (!g_index) {
    sm_index *ix = (sm_index*) calloc (1, sizeof (sm_index));
    ix->fasta = strdup (fasta);
    read_fasta (ix, fasta);
    if (a > 0) build_index (ix, a);
    g_index = ix;
  } else if (a > 0 && !g_index->key) build_index (g_index, a);
  sm_index *r = g_index;
  pthread_mutex_unlock (&g_mu);
  return r;
}

static int contig_of (const sm_index *ix, long gpos)
{
  int lo = 0, hi = ix->n_seqs - 1;
  while (lo < hi) { int mid = (lo + hi + 1) / 2; if (ix->off[mid] <= gpos) lo = mid; else hi = mid - 1; }
  return lo;
}

typedef struct { int ref_id, position, ref_length, nm, neg; } hit;
static int cmp_hit (const void *a, const void *b)
{
  const hit *x = (const hit*) a, *y = (const hit*) b;
  if (x->nm != y->nm) return x->nm - y->nm;
  if (x->ref_id != y->ref_id) return x->ref_id - y->ref_id;
  if (x->position != y->position) return x->position - y->position;
  return x->neg - y->neg;
}

static void add_hit (hit **h, int *n, int *cap, hit x)
{
  for (int i = 0; i < *n; i++) if ((*h)[i].ref_id == x.ref_id && (*h)[i].position == x.position && (*h)[i].neg == x.neg) { if (x.nm < (*h)[i].nm) (*h)[i] = x; return; }
  if (*n == *cap) { *cap = *cap ? 2 * *cap : 8; *h = (hit*) realloc (*h, (size_t) *cap * sizeof (hit)); }
  (*h)[(*n)++] = x;
}

/* candidates = reference positions whose a-mer equals x[at .. at+a) */
static void lookup (const sm_index *ix, const char *x, int at, long *first, long *last)
{
  uint32_t key = 0;
  *first = 0; *last = -1;
  for (int j = 0; j < ix->a; j++) { int cd = code_of (x[at + j]); if (cd < 0) return; key = (key << 2) | (uint32_t) cd; }
  long lo = 0, hi = ix->n_pos;
  while (lo < hi) { long mid = (lo + hi) / 2; if (ix->key[mid] < key) lo = mid + 1; else hi = mid; }
  *first = lo;
  while (lo < ix->n_pos && ix->key[lo] == key) lo++;
  *last = lo - 1;
}

static void try_position (const sm_index *ix, const char *x, int n, int k, long gp, int neg, hit **h, int *nh, int *cap)
{
  if (gp < 0) return;
  const int c = contig_of (ix, gp);
  const long cend = ix->off[c] + ix->len[c];
  const char *r = ix->seq + gp;
  hit y; y.ref_id = c; y.position = (int) (gp - ix->off[c]); y.neg = neg;
  if (gp + n <= cend) { /* same length: substitutions only */
    int mm = 0;
    for (int i = 0; i < n && mm <= SM_MAX_SUBST; i++) mm += (r[i] != x[i]);
    if (mm <= SM_MAX_SUBST) { y.nm = mm; y.ref_length = n; add_hit (h, nh, cap, y); if (!mm) return; }
  }
  /* flanks exact, tract of another length */
  const int t = n - 2 * k;
  if (t < 1 || gp + k >= cend || memcmp (r, x, (size_t) k) != 0) return;
  const char b = x[k];
  long tr = 0;
  while (gp + k + tr < cend && r[k + tr] == b) tr++;
  if (tr < 1 || tr == t || gp + k + tr + k > cend) return;
  if (memcmp (r + k + tr, x + k + t, (size_t) k) != 0) return;
  y.nm = (int) (tr > t ? tr - t : t - tr); y.ref_length = (int) (2 * k + tr);
  add_hit (h, nh, cap, y);
}

static void revcomp_into (const char *s, int n, char *o)
{
  for (int i = 0; i < n; i++) {
    char c = s[n - 1 - i];
    o[i] = c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : 'N';
  }
  o[n] = '\0';
}

static int flank_length (const char *x, int n)
{
  if (n < 3) return 0;
  int lo = n / 2, hi = n / 2;
  while (lo > 0 && x[lo - 1] == x[n / 2]) lo--;
  while (hi + 1 < n && x[hi + 1] == x[n / 2]) hi++;
  return (n - (hi - lo + 1)) / 2;
}

/* ---- the wrapper_bwa.h interface ---- */
bwase_options_t new_bwase_options_t (int level) { bwase_options_t o; o.dummy = level; return o; }

static bwase_match_t new_match_with_names (const sm_index *ix)
{
  bwase_match_t match = (bwase_match_t) malloc (sizeof (struct bwase_match_struct));
  match->m = NULL; match->n_m = 0;
  match->bns = (shim_bns_t*) malloc (sizeof (shim_bns_t));
  match->bns->n_seqs = ix->n_seqs;
  match->bns->anns = (shim_ann_t*) malloc ((size_t) ix->n_seqs * sizeof (shim_ann_t));
  for (int i = 0; i < ix->n_seqs; i++) { match->bns->anns[i].name = strdup (ix->name[i]); match->bns->anns[i].len = ix->len[i]; }
  return match;
}

bwase_match_t
new_bwase_match_from_bwa_and_char_vector (const char *fasta, char_vector seqname, char_vector seqs, int n_occ, bwase_options_t opt)
{
  (void) seqname; (void) n_occ; (void) opt;
  /* The queries do not carry the flank length k, but query = k + tract + k where the tract is a maximal run of one
   * base, so the run that covers the middle of the query is the tract and k = (length - run) / 2. */
  int a = SM_ANCHOR_MAX;
  for (int q = 0; q < seqs->nstrings; q++) { int kq = flank_length (seqs->string[q], (int) strlen (seqs->string[q])); if (kq >= 1 && kq < a) a = kq; }
  const sm_index *ix = get_index (fasta, a);
  bwase_match_t match = new_match_with_names (ix);
  int cap_m = 0;
  char *rc = NULL; int rc_cap = 0;
  hit *h = NULL; int nh = 0, hcap = 0;
  for (int q = 0; q < seqs->nstrings; q++) {
    const char *x = seqs->string[q];
    const int n = (int) strlen (x);
    const int k = flank_length (x, n);
    if (k < a) continue;
    if (n + 1 > rc_cap) { rc_cap = 2 * n + 2; rc = (char*) realloc (rc, (size_t) rc_cap); }
    revcomp_into (x, n, rc);
    nh = 0;
    for (int neg = 0; neg < 2; neg++) {
      const char *y = neg ? rc : x;
      long f, l;
      lookup (ix, y, 0, &f, &l);                       /* anchored on the first a bases */
      for (long i = f; i <= l; i++) try_position (ix, y, n, k, ix->pos[i], neg, &h, &nh, &hcap);
      lookup (ix, y, n - a, &f, &l);                   /* anchored on the last a bases (same length only) */
      for (long i = f; i <= l; i++) try_position (ix, y, n, k, (long) ix->pos[i] - (n - a), neg, &h, &nh, &hcap);
    }
    if (!nh) continue;
    qsort (h, (size_t) nh, sizeof (hit), cmp_hit);
    for (int i = 0; i < nh && i < SM_HITS_PER_QUERY; i++) {
      if (match->n_m == cap_m) { cap_m = cap_m ? 2 * cap_m : 1024; match->m = (bwase_elem_t*) realloc (match->m, (size_t) cap_m * sizeof (bwase_elem_t)); }
      bwase_elem_t *e = &match->m[match->n_m++];
      e->query_id = q; e->ref_id = h[i].ref_id; e->position = h[i].position; e->ref_length = h[i].ref_length;
      e->nm = h[i].nm; e->mm = h[i].nm; e->gape = 0; e->gapo = 0; e->neg_strand = h[i].neg;
    }
  }
  free (rc); free (h);
  return match;
}

bwase_match_t new_bwase_match_t (const char *fasta) { return new_match_with_names (get_index (fasta, 0)); }

void del_bwase_match_t (bwase_match_t match)
{
  if (!match) return;
  if (match->bns) {
    for (int i = 0; i < match->bns->n_seqs; i++) free (match->bns->anns[i].name);
    free (match->bns->anns); free (match->bns);
  }
  free (match->m); free (match);
}

char *save_bwa_index (const char *genome_filename, const char *suffix, bool overwrite)
{ (void) suffix; (void) overwrite; get_index (genome_filename, 0); return strdup (genome_filename); }
