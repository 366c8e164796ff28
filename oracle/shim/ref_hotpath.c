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

#ifdef TJB_PATCHED
/* the same with the sources patched by integration/tatajuba-libtjb_gpu.patch: new_hopo_counter_from_files_gpu() counts
 * on the GPU, finalise_hopo_counter() goes straight to the mapper call, where the stub above captures the state */
int
ref_hotpath_run_files_gpu (const char *const *files, int n_files, int kmer_size, int min_tract_size,
                           int min_coverage, int remove_biased, int device, ref_dump_t *out)
{
  tatajuba_options_t opt = make_opt (kmer_size, min_tract_size, min_coverage, remove_biased, n_files > 1);
  reset_dump (out);
  double t0 = now_secs ();
  hopo_counter hc = new_hopo_counter_from_files_gpu (files[0], n_files > 1 ? files[1] : NULL, opt, device);
  out->secs_scan = now_secs () - t0;
  out->status = -1;
  tl_hc = hc; tl_out = out;
  tl_t_finalise_start = now_secs ();
  finalise_hopo_counter (hc);
  tl_hc = NULL; tl_out = NULL;
  if (out->status != 0) out->status = hc->gpu_status; /* the mapper was never reached: the sample is excluded */
  out->n_raw = -1; /* not kept by the binding (hc->n_elem before finalise is the GPU library's diagnostic n_raw) */
  del_hopo_counter (hc);
  return out->status;
}
#endif

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

/* ---- cross-sample merge (genome_set.c:195-289): the reference's own g_tract_vector_concatenate_tracts() and
 * context_histograms_overlap() run on context_histograms built from plain arrays.  The tract_id loop below follows
 * genome_set.c:208-222 line by line; the summary statistics that loop also feeds (update_g_tract_summary_...) need
 * complete mapped histograms and are not part of the merge order / tract ids checked here. */
#include "genome_set.h"
int
ref_merge_tracts (int n_genome, const int *n_hist, const int *location, const int *base, const int *mode_len,
                  const char *names, int name_stride, int lev_distance, int *out_genome, int *out_index, int *out_tract_id)
{
  tatajuba_options_t opt = make_opt (25, 4, 5, 1, 0);
  opt.levenshtein_distance = lev_distance;
  genomic_context_list_t *genome = (genomic_context_list_t*) malloc ((size_t) n_genome * sizeof (genomic_context_list_t));
  int off = 0, n_total = 0;
  for (int g = 0; g < n_genome; g++) { /* samples are already compacted (remove_empty_genomes, genome_set.c:97) */
    genome[g] = (genomic_context_list_t) calloc (1, sizeof (struct genomic_context_list_struct));
    genome[g]->opt = opt;
    genome[g]->n_hist = n_hist[g];
    genome[g]->hist = (context_histogram_t*) malloc ((size_t) (n_hist[g] ? n_hist[g] : 1) * sizeof (context_histogram_t));
    for (int j = 0; j < n_hist[g]; j++, off++) {
      context_histogram_t ch = (context_histogram_t) calloc (1, sizeof (struct context_histogram_struct));
      ch->location = location[off];
      ch->base = base[off];
      ch->mode_context_length = mode_len[off];
      ch->name = strndup (names + (size_t) off * name_stride, name_stride);
      ch->index = g;        /* genome_set.c:100 */
      ch->n_tracts = j;     /* stash: index inside the sample (untouched by the functions under test) */
      genome[g]->hist[j] = ch;
    }
    n_total += n_hist[g];
  }
  g_tract_vector_t tract = (g_tract_vector_t) calloc (1, sizeof (struct g_tract_vector_struct));
  g_tract_vector_concatenate_tracts (tract, genome, n_genome);
  if (tract->n_concat > 0) {
    int i, prev, lev = 0;
    tract->concat[0]->tract_id = 0;
    for (prev = 0, i = 1; i < tract->n_concat; i++) {
      bool is_same = context_histograms_overlap (tract->concat[i], tract->concat[prev], &lev, opt);
      if (!is_same) { tract->concat[i]->tract_id = tract->concat[prev]->tract_id + 1; prev = i; }
      else tract->concat[i]->tract_id = tract->concat[prev]->tract_id;
    }
  }
  for (int i = 0; i < tract->n_concat; i++) {
    out_genome[i] = tract->concat[i]->index;
    out_index[i] = tract->concat[i]->n_tracts;
    out_tract_id[i] = tract->concat[i]->tract_id;
  }
  int n_concat = tract->n_concat;
  free (tract->concat); free (tract);
  for (int g = 0; g < n_genome; g++) {
    for (int j = 0; j < genome[g]->n_hist; j++) { free (genome[g]->hist[j]->name); free (genome[g]->hist[j]); }
    free (genome[g]->hist); free (genome[g]);
  }
  free (genome);
  (void) n_total;
  return n_concat;
}
