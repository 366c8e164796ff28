/* The UNMODIFIED reference from reads to the per-sample context histograms, with the deterministic stand-in mapper
 * (standin_mapper.c) behind wrapper_bwa.h: new_genomic_context_list() runs exactly as in the reference
 * (src/context_histogram.c:224-272: finalise_hopo_counter, mapper, location sort, context merge, empfreq histograms,
 * merge at the same location), then its two ends are copied out as flat arrays:
 *   - the mapped elements hc->elem[ref_start .. n_elem) in the order the location sort left them in (the INPUT of the
 *     product's tjb_context_histograms), and
 *   - every context_histogram_t of the resulting genomic_context_list (what that function must reproduce).
 * TEST INFRASTRUCTURE ONLY (see ref_hotpath.h). */
#include "context_histogram.h"
#include "tjb_gpu.h" /* only for the layout of tjb_context_table, so that both sides fill the same struct */

/* The reference's read_gff3_from_file() (biomcmc-lib gff3_format.c:197-279) deletes its seq_region vector a second time:
 * gff3_finalise() takes the pointer BY VALUE, reorder_seqid_charvector_from_pragma() (:531-535, :560-563) deletes the
 * vector and replaces the callee's copy, and :273 then deletes the caller's stale pointer (--ref_counter on freed memory).
 * The command line does that once at start-up; a test process that builds many samples does it per call and
 * eventually decrements a live object that reused the chunk.  The features of the GFF3 are not what this shim pins, so
 * the file is read with the reference's own building blocks (same calls, same order as :213-271) minus that last delete.
 * Only what synth.write_reference() writes is handled: pragmas, feature lines, no ##FASTA section. */
gff3_t new_gff3_t (const char *filename);
gff3_fields gff3_fields_from_char_line (char *line);
void add_fields_to_gff3_t (gff3_t g3, gff3_fields gfield);
void gff3_finalise (gff3_t g3, char_vector seq_region);
void gff3_generate_seq_vectors (gff3_t g3, hashtable hgs);

static gff3_t shim_read_gff3 (const char *path)
{
  FILE *f = fopen (path, "r");
  if (!f) biomcmc_error ("shim: cannot open %s", path);
  gff3_t g3 = new_gff3_t (path);
  char_vector seq_region = new_char_vector (1);
  int *reg_size = NULL, n_reg = 0;
  char *line = NULL, name[1024]; size_t cap = 0; ssize_t n;
  while ((n = getline (&line, &cap, f)) != -1) {
    while (n > 0 && (line[n - 1] == '\n' || line[n - 1] == '\r')) line[--n] = '\0';
    if (!n) continue;
    if (line[0] == '#') {
      int i1, i2;
      if (sscanf (line, "##sequence-region %1023s %d %d", name, &i1, &i2) == 3) {
        char_vector_add_string (seq_region, name);
        reg_size = (int*) realloc (reg_size, (size_t) (n_reg + 1) * sizeof (int));
        reg_size[n_reg++] = i2 - 1;
      }
      continue;
    }
    gff3_fields gf = gff3_fields_from_char_line (line);
    if (gff3_fields_is_valid (gf)) add_fields_to_gff3_t (g3, gf);
  }
  free (line); fclose (f);
  char_vector_finalise_big (g3->sequence);
  hashtable hgs = new_hashtable (n_reg);
  for (int i = 0; i < n_reg; i++) insert_hashtable (hgs, seq_region->string[i], reg_size[i]);
  gff3_finalise (g3, seq_region); /* releases seq_region (see above) */
  gff3_generate_seq_vectors (g3, hgs);
  del_hashtable (hgs);
  free (reg_size);
  return g3;
}

typedef struct {
  int n_mapped;            /* hc->n_elem - hc->ref_start */
  void *elem;              /* malloc'ed copy, 40 bytes each */
  tjb_context_table hist;  /* n_hist = -1: the sample was excluded (new_genomic_context_list returned NULL) */
} ref_pipeline_dump_t;

int
ref_pipeline_run (const char *buf, size_t n_bytes, const char *fasta, const char *gff, int kmer_size, int min_tract_size,
                  int min_coverage, int remove_biased, int max_distance_per_flank, int levenshtein_distance, ref_pipeline_dump_t *out)
{
  tatajuba_options_t opt;
  memset (&opt, 0, sizeof (opt));
  memset (out, 0, sizeof (*out));
  opt.gff = shim_read_gff3 (gff);
  /* (main.c:133-137 also attaches the fasta's DNA to the GFF3; nothing on this path reads it) */
  opt.reference_fasta_filename = (char*) fasta;
  opt.outdir = (char*) ".";
  opt.paired_end = false; opt.remove_biased = remove_biased; opt.save_vcf = false;
  opt.max_distance_per_flank = max_distance_per_flank; opt.levenshtein_distance = levenshtein_distance;
  opt.kmer_size = kmer_size; opt.min_tract_size = min_tract_size; opt.min_coverage = min_coverage;
  opt.n_samples = 1; opt.n_threads = 1;
  char *s = save_bwa_index (fasta, NULL, false);
  if (s) free (s);

  hopo_counter hc = new_hopo_counter (kmer_size);
  hc->name = strdup ("buffer");
  hc->opt = opt;
  size_t start = 0;
  while (start < n_bytes) {
    const char *nl = (const char*) memchr (buf + start, '\n', n_bytes - start);
    size_t end = nl ? (size_t) (nl - buf) : n_bytes;
    if (end > start) update_hopo_counter_from_seq (hc, (char*) (buf + start), (int) (end - start), min_tract_size);
    start = end + 1;
  }
  genomic_context_list_t g = new_genomic_context_list (hc);
  if (!g) { out->hist.n_hist = -1; del_hopo_counter (hc); del_gff3_t (opt.gff); return -1; }

  out->n_mapped = hc->n_elem - hc->ref_start;
  out->elem = malloc ((size_t) (out->n_mapped ? out->n_mapped : 1) * sizeof (hopo_element));
  memcpy (out->elem, hc->elem + hc->ref_start, (size_t) out->n_mapped * sizeof (hopo_element));

  tjb_context_table *t = &out->hist;
  const int n = g->n_hist, stride = 2 * kmer_size + 4;
  size_t n_ctx = 0, n_bins = 0;
  for (int i = 0; i < n; i++) { n_ctx += (size_t) g->hist[i]->n_context; n_bins += (size_t) g->hist[i]->h->n; }
  t->n_hist = n; t->name_stride = stride;
#define ARR(T, cnt) ((T*) calloc ((cnt) ? (cnt) : 1, sizeof (T)))
  t->location = ARR (int32_t, n); t->loc2d = ARR (int32_t, 3 * (size_t) n); t->base = ARR (int32_t, n); t->neg_strand = ARR (int32_t, n);
  t->mismatches = ARR (int32_t, n); t->multi = ARR (int32_t, n); t->indel = ARR (int32_t, n); t->integral = ARR (int32_t, n);
  t->n_context = ARR (int32_t, n); t->mode_context_count = ARR (int32_t, n); t->mode_context_length = ARR (int32_t, n);
  t->mode_context_id = ARR (int32_t, n); t->h_min = ARR (int32_t, n); t->h_max = ARR (int32_t, n);
  t->names = ARR (char, (size_t) n * stride); t->ctx_first = ARR (int32_t, (size_t) n + 1); t->contexts = ARR (uint64_t, 2 * n_ctx);
  t->h_first = ARR (int32_t, (size_t) n + 1); t->h_length = ARR (int32_t, n_bins); t->h_freq = ARR (int32_t, n_bins);
#undef ARR
  size_t c = 0, b = 0;
  for (int i = 0; i < n; i++) {
    context_histogram_t h = g->hist[i];
    t->location[i] = h->location; t->base[i] = h->base; t->neg_strand[i] = h->neg_strand ? 1 : 0; t->mismatches[i] = h->mismatches;
    for (int j = 0; j < 3; j++) t->loc2d[3 * i + j] = h->loc2d[j];
    t->multi[i] = h->multi; t->indel[i] = h->indel; t->integral[i] = h->integral; t->n_context[i] = h->n_context;
    t->mode_context_count[i] = h->mode_context_count; t->mode_context_length[i] = h->mode_context_length;
    t->mode_context_id[i] = h->mode_context_id; t->h_min[i] = h->h->min; t->h_max[i] = h->h->max;
    strncpy (t->names + (size_t) i * stride, h->name, (size_t) stride - 1);
    t->ctx_first[i] = (int32_t) c;
    memcpy (t->contexts + 2 * c, h->context, 2 * (size_t) h->n_context * sizeof (uint64_t));
    c += (size_t) h->n_context;
    t->h_first[i] = (int32_t) b;
    for (int j = 0; j < h->h->n; j++, b++) { t->h_length[b] = h->h->i[j].idx; t->h_freq[b] = h->h->i[j].freq; }
  }
  t->ctx_first[n] = (int32_t) c; t->h_first[n] = (int32_t) b;
  del_genomic_context_list (g);
  del_hopo_counter (hc);
  del_gff3_t (opt.gff);
  return n;
}

void ref_pipeline_free (ref_pipeline_dump_t *d)
{
  if (!d) return;
  free (d->elem);
  tjb_context_table *t = &d->hist;
  void *p[] = {t->location, t->loc2d, t->base, t->neg_strand, t->mismatches, t->multi, t->indel, t->integral, t->n_context,
               t->mode_context_count, t->mode_context_length, t->mode_context_id, t->h_min, t->h_max, t->names, t->ctx_first,
               t->contexts, t->h_first, t->h_length, t->h_freq};
  for (size_t i = 0; i < sizeof (p) / sizeof (p[0]); i++) free (p[i]);
  memset (d, 0, sizeof (*d));
}
