/* C interface of oracle/_ref/libref_hotpath.so: the UNMODIFIED reference sources (src/hopo_counter.c and
 * biomcmc-lib, compiled where they lie under /root/reference) driven up to the BWA boundary
 * (hopo_counter.c:523), where the stubbed mapper captures the counter state.
 * TEST INFRASTRUCTURE ONLY: may be used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs, never by the product path. */
#ifndef TJB_REF_HOTPATH_H
#define TJB_REF_HOTPATH_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int status;       /* 0 = mapper boundary reached; 1 = no HT at all (hopo_counter.c:345); 2 = none after QC (:376);
                       3 = no context above min coverage (:406) */
  int n_raw;        /* hc->n_elem before finalise_hopo_counter() */
  int n_elem, n_idx, coverage;
  void *elem;       /* malloc'ed copy of hc->elem[0..n_elem) = reference-layout 40-byte hopo_element */
  int *idx_initial, *idx_final; /* malloc'ed copies, n_idx each */
  double secs_scan;     /* read + update_hopo_counter_from_seq (the reference's secs[0] span, genome_set.c:71-74) */
  double secs_finalise; /* finalise_hopo_counter() entry -> mapper call (hopo_counter.c:339-523) */
} ref_dump_t;

/* files: 1 (single-end) or 2 (paired: R2 appended to the same counter, genome_set.c:72-73) FASTQ/FASTA(.gz) paths */
int ref_hotpath_run_files (const char *const *files, int n_files, int kmer_size, int min_tract_size,
                           int min_coverage, int remove_biased, ref_dump_t *out);
/* only in oracle/_ref/libref_hotpath_gpu.so (the sources patched by integration/tatajuba-libtjb_gpu.patch): the binding's
 * new_hopo_counter_from_files_gpu() + finalise_hopo_counter(); n_raw is not reported (-1) */
int ref_hotpath_run_files_gpu (const char *const *files, int n_files, int kmer_size, int min_tract_size,
                               int min_coverage, int remove_biased, int device, ref_dump_t *out);
/* buf: reads separated by '\n' (a trailing '\n' is optional); each read goes through update_hopo_counter_from_seq */
int ref_hotpath_run_buffer (const char *buf, size_t n_bytes, int kmer_size, int min_tract_size,
                            int min_coverage, int remove_biased, ref_dump_t *out);
void ref_dump_free (ref_dump_t *d);
size_t ref_sizeof_hopo_element (void);

/* Cross-sample merge: the reference's g_tract_vector_concatenate_tracts() (genome_set.c:250-289) and
 * context_histograms_overlap() (context_histogram.c:87-110) on per-sample histogram tables given as flat arrays
 * (samples concatenated, each sorted by location; names = name_stride bytes per histogram).  Outputs, per element of
 * the merged vector: sample, index inside the sample, tract_id (genome_set.c:208-222).  Returns n_concat. */
int ref_merge_tracts (int n_genome, const int *n_hist, const int *location, const int *base, const int *mode_len,
                      const char *names, int name_stride, int lev_distance, int *out_genome, int *out_index,
                      int *out_tract_id);

#ifdef __cplusplus
}
#endif
#endif
