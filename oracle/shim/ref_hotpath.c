/* Driver + stub mapper linked with the UNMODIFIED reference sources (see oracle/Makefile).
 * The reference calls new_bwase_match_from_bwa_and_char_vector() exactly at the BWA boundary
 * (src/hopo_counter.c:523); the stub below copies the counter state there and returns "no matches".
 * TEST INFRASTRUCTURE ONLY (see ref_hotpath.h). */
#include <time.h>
#include "hopo_counter.h"
#include "ref_hotpath.h"

static __thread hopo_counter tl_hc = NULL;
static __thread ref_dump_t *tl_out = NULL;
static __thread double tl_t_finalise_start = 0.;

static double now_secs (void)
{
  struct timespec ts;
  clock_gettime (CLOCK_MONOTONIC, &ts);
  return (double) ts.tv_sec + 1e-9 * (double) ts.tv_nsec;
}

/* ---- stubbed wrapper_bwa.h ---- */
bwase_options_t new_bwase_options_t (int level) { bwase_options_t o; o.dummy = level; return o; }

static bwase_match_t new_empty_match (void)
{
  bwase_match_t match = (bwase_match_t) malloc (sizeof (struct bwase_match_struct));
  match->m = NULL; match->n_m = 0;
  match->bns = (shim_bns_t*) malloc (sizeof (shim_bns_t));
  match->bns->n_seqs = 1;
  match->bns->anns = (shim_ann_t*) malloc (sizeof (shim_ann_t));
  match->bns->anns[0].name = strdup ("stub_contig");
  match->bns->anns[0].len = 1;
  return match;
}

bwase_match_t
new_bwase_match_from_bwa_and_char_vector (const char *fasta, char_vector seqname, char_vector seqs, int n_occ, bwase_options_t opt)
{
  (void) fasta; (void) seqname; (void) seqs; (void) n_occ; (void) opt;
  if (tl_hc && tl_out) { /* capture the state at the boundary */
    hopo_counter hc = tl_hc;
    ref_dump_t *d = tl_out;
    d->secs_finalise = now_secs () - tl_t_finalise_start;
    d->status = 0;
    d->n_elem = hc->n_elem; d->n_idx = hc->n_idx; d->coverage = hc->coverage;
    d->elem = malloc ((size_t) hc->n_elem * sizeof (hopo_element));
    memcpy (d->elem, hc->elem, (size_t) hc->n_elem * sizeof (hopo_element));
    d->idx_initial = (int*) malloc ((size_t) hc->n_idx * sizeof (int));
    d->idx_final   = (int*) malloc ((size_t) hc->n_idx * sizeof (int));
    memcpy (d->idx_initial, hc->idx_initial, (size_t) hc->n_idx * sizeof (int));
    memcpy (d->idx_final,   hc->idx_final,   (size_t) hc->n_idx * sizeof (int));
  }
  return new_empty_match ();
}

bwase_match_t new_bwase_match_t (const char *fasta) { (void) fasta; return new_empty_match (); }

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
{ (void) suffix; (void) overwrite; return strdup (genome_filename); }

/* ---- driver ---- */
static tatajuba_options_t make_opt (int k, int m, int min_cov, int remove_biased, int paired)
{
  tatajuba_options_t opt;
  memset (&opt, 0, sizeof (opt));
  opt.reference_fasta_filename = (char*) "stub.fasta";
  opt.outdir = (char*) ".";
  opt.paired_end = paired; opt.remove_biased = remove_biased; opt.save_vcf = false;
  opt.gff = NULL;
  opt.max_distance_per_flank = 1; opt.levenshtein_distance = 2;
  opt.kmer_size = k; opt.min_tract_size = m; opt.min_coverage = min_cov;
  opt.n_samples = 1; opt.n_threads = 1;
  return opt;
}

static void finalise_and_classify (hopo_counter hc, ref_dump_t *out)
{
  out->n_raw = hc->n_elem;
  out->status = -1;
  tl_hc = hc; tl_out = out;
  tl_t_finalise_start = now_secs ();
  finalise_hopo_counter (hc);
  tl_hc = NULL; tl_out = NULL;
  if (out->status != 0) { /* the mapper was never reached: one of the three "empty counter" paths */
    out->secs_finalise = now_secs () - tl_t_finalise_start;
    if (!out->n_raw) out->status = 1;
    else if (!hc->idx_initial) out->status = 2;
    else out->status = 3;
  }
}

static void reset_dump (ref_dump_t *out)
{
  memset (out, 0, sizeof (*out));
}

int
ref_hotpath_run_files (const char *const *files, int n_files, int kmer_size, int min_tract_size,
                       int min_coverage, int remove_biased, ref_dump_t *out)
{
  tatajuba_options_t opt = make_opt (kmer_size, min_tract_size, min_coverage, remove_biased, n_files > 1);
  hopo_counter hc = NULL;
  reset_dump (out);
  double t0 = now_secs ();
  for (int i = 0; i < n_files; i++) hc = new_or_append_hopo_counter_from_file (hc, files[i], opt);
  out->secs_scan = now_secs () - t0;
  finalise_and_classify (hc, out);
  del_hopo_counter (hc);
  return out->status;
}

int
ref_hotpath_run_buffer (const char *buf, size_t n_bytes, int kmer_size, int min_tract_size,
                        int min_coverage, int remove_biased, ref_dump_t *out)
{
  tatajuba_options_t opt = make_opt (kmer_size, min_tract_size, min_coverage, remove_biased, 0);
  hopo_counter hc = new_hopo_counter (kmer_size);
  reset_dump (out);
  hc->name = strdup ("buffer");
  hc->opt = opt;
  double t0 = now_secs ();
  size_t start = 0;
  while (start < n_bytes) {
    const char *nl = (const char*) memchr (buf + start, '\n', n_bytes - start);
    size_t end = nl ? (size_t) (nl - buf) : n_bytes;
    if (end > start) update_hopo_counter_from_seq (hc, (char*) (buf + start), (int) (end - start), min_tract_size);
    start = end + 1;
  }
  out->secs_scan = now_secs () - t0;
  finalise_and_classify (hc, out);
  del_hopo_counter (hc);
  return out->status;
}

void ref_dump_free (ref_dump_t *d)
{
  if (!d) return;
  free (d->elem); free (d->idx_initial); free (d->idx_final);
  d->elem = NULL; d->idx_initial = d->idx_final = NULL;
}

size_t ref_sizeof_hopo_element (void) { return sizeof (hopo_element); }
