/* libtjb_gpu — C ABI of the B200-native (sm_100a) replacement for tatajuba's hot path.
 *
 * tatajuba has no plugin/FFI interface; the seam is its internal C API (reference src/hopo_counter.h:61-81)
 * as used by the per-sample OpenMP loop body (src/genome_set.c:69-78 and 84-93).  One call here stands in for
 *
 *     hc = new_or_append_hopo_counter_from_file (NULL, R1, opt);          src/hopo_counter.c:135-157
 *     hc = new_or_append_hopo_counter_from_file (hc,   R2, opt);          (paired end only)
 *     finalise_hopo_counter (hc);   -- steps 0-4 + coverage,              src/hopo_counter.c:339-415
 *                                      i.e. everything BEFORE the mapper call at :416
 *
 * and returns the counter state exactly as the reference holds it when it calls BWA (src/hopo_counter.c:523):
 * elem[] (reference-layout 40-byte hopo_element, reference order), n_elem, idx_initial[], idx_final[], n_idx,
 * coverage.  INTEGRATION.md shows the ~30-line patch that makes the reference call this.
 *
 * Plain pointers and sizes only; no torch / CUDA types in any signature (streams and device pointers travel
 * as void* / integers).  There is NO CPU fallback: every compute entry point returns TJB_ERR_CUDA when no
 * sm_100-class device is usable.
 */
#ifndef TJB_GPU_H
#define TJB_GPU_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* status of a counted sample; 1..3 mirror the reference's three "sample will be excluded" paths */
enum {
  TJB_OK = 0,
  TJB_NO_HT = 1,                 /* hopo_counter.c:345-349  no tract at all                              */
  TJB_NO_HT_AFTER_QC = 2,        /* hopo_counter.c:376-381  none left after strand-bias / singleton QC   */
  TJB_NO_CTX_ABOVE_COVERAGE = 3, /* hopo_counter.c:406-411  no context reaches min_coverage              */
  /* errors (negative): the reference would biomcmc_error() -> exit(EXIT_FAILURE) (lowlevel.c:39-53) */
  TJB_ERR_CUDA = -1,
  TJB_ERR_ARG = -2,
  TJB_ERR_IO = -3,
  TJB_ERR_NOMEM = -4,
  TJB_ERR_TOO_LARGE = -5,        /* more than INT32_MAX tract observations (the reference's int n_elem) */
  TJB_ERR_NOALLOC_LIMIT = -6     /* tjb_tract_rows_noalloc: more reads than its fixed scratch is sized for */
};

/* the four tatajuba_options_t fields the path reads (src/hopo_counter.h:20-32); clamps as src/main.c:184-192 */
typedef struct {
  int32_t kmer_size;      /* -k, 2..32  flank length in bases                        */
  int32_t min_tract_size; /* -m, 1..32  (effective minimum is max(m,2), hopo_counter.c:227-229) */
  int32_t min_coverage;   /* -i, 0..65535                                            */
  int32_t remove_biased;  /* 1 = default (keep canon_flag==3), 0 = "-b" (keep count>1) */
  int32_t paired_end;     /* only echoed in warnings by the reference; not used by the arithmetic */
} tjb_opts;

/* counter state at the BWA boundary; replaces the fields of struct hopo_counter_struct (src/hopo_counter.h:51-59).
 * elem/idx_initial/idx_final/name are malloc()'ed host memory owned by the caller (the reference free()s them,
 * hopo_counter.c:180-183, 568-571); release with tjb_counter_out_free() or plain free(). */
typedef struct {
  int32_t status;            /* TJB_OK or one of the three drop codes; then n_elem = n_idx = 0 like the reference */
  int32_t n_elem, n_idx, coverage;
  void    *elem;             /* n_elem x 40-byte hopo_element (hopo_counter.h:34-49); mapping fields as initialised
                                at hopo_counter.c:296-302 and read_offset = -1 as after :511 */
  int32_t *idx_initial, *idx_final;
  char    *name;             /* copy of the first file name (hopo_counter.c:145-149), NULL for buffer inputs */
  /* diagnostics (not part of the reference state) */
  int64_t n_reads, n_bases;  /* sequences / sequence bytes consumed (0 for device buffers: not counted there) */
  int64_t n_raw;             /* tract observations before the reduce (hc->n_elem before finalise)       */
  int64_t n_distinct;        /* distinct (base, contexts, length) before the QC filter                   */
  int64_t n_skipped_nonacgt; /* qualifying runs of a non-ACGTU byte (undefined behaviour in the reference,
                                hopo_counter.c:233-248): counted and skipped, see DESIGN.md                 */
  double  ms_decode, ms_device; /* host FASTQ decode wall time; device time (CUDA events) for scan+reduce, profiling level 2 */
  int32_t arena;             /* non-zero: elem / idx_initial / idx_final point into the context's result arena
                                (TJB_RESULT_ARENA): not to be free()d, valid until the context counts its next sample */
  int32_t reserved;
} tjb_counter_out;

typedef struct tjb_ctx tjb_ctx; /* one per (host thread, device): stream, pinned double buffers, device workspace */

/* per-kernel timing, filled when profiling is on (CUDA events on the context's stream) */
typedef struct {
  char    name[48];
  int64_t launches;
  double  ms_total;
  int64_t bytes_algorithmic; /* algorithmic bytes summed over those launches (DESIGN.md states the per-unit figure) */
} tjb_kernel_stat;

const char *tjb_version (void);
const char *tjb_strerror (int code);
/* last error message of this context (CUDA error string, file name, ...) */
const char *tjb_ctx_last_error (const tjb_ctx *ctx);

/* device = CUDA ordinal.  Fails with TJB_ERR_CUDA when there is no usable GPU (no CPU fallback exists). */
int  tjb_ctx_create (int device, tjb_ctx **ctx);
void tjb_ctx_destroy (tjb_ctx *ctx);
/* the context's stream as a cudaStream_t cast to void* (for callers that time with their own events) */
void *tjb_ctx_stream (tjb_ctx *ctx);
/* 0 = off; 1 = the scan kernel's launches are timed (two events per launch, nothing else: cheap enough for a timed
 * region); 2 = every kernel, and ms_device of tjb_counter_out */
int  tjb_ctx_set_profiling (tjb_ctx *ctx, int on);
/* Where the result arrays live.  TJB_RESULT_MALLOC (default): malloc()ed copies owned by the caller, as the reference
 * needs them (it realloc()s and free()s hc->elem, hopo_counter.c:180-183, 562).  TJB_RESULT_ARENA: pointers into the
 * context's pinned result arena, which the device writes directly: no host copy, no page faults; the arrays are valid
 * until this context counts its next sample (callers that only read the boundary state, e.g. to build mapper queries). */
enum { TJB_RESULT_MALLOC = 0, TJB_RESULT_ARENA = 1 };
int  tjb_ctx_set_result_mode (tjb_ctx *ctx, int mode);
/* usable CUDA devices (0 without a driver); the binding deals samples to devices by thread number */
int  tjb_device_count (void);
/* returns the number of kernels with at least one launch since the last reset; fills up to max_stats entries */
int  tjb_ctx_kernel_stats (tjb_ctx *ctx, tjb_kernel_stat *stats, int max_stats);
void tjb_ctx_reset_kernel_stats (tjb_ctx *ctx);
/* kernels launched by this context since creation (or since reset_kernel_stats) */
int64_t tjb_ctx_launch_count (const tjb_ctx *ctx);
/* bytes tjb_count_sample_host() has sent host -> device since the last tjb_ctx_reset_kernel_stats().  Large host
 * buffers cross PCIe as 2-bit codes plus (position, byte) pairs for every byte that is not one of "ACGT" and are
 * rebuilt byte for byte on the device, so this is less than the bytes handed in (environment TJB_HOST_PACK=0 sends
 * them as they are, TJB_HOST_PACK=n packs with n threads; default one per hardware thread, at most 32). */
int64_t tjb_ctx_htod_bytes (const tjb_ctx *ctx);

/* Whole replaced span from FASTQ/FASTA(.gz) files: n_files = 1 (single end) or 2 (R1 then R2 into one counter,
 * genome_set.c:72-73).  Host C decode (kseq semantics, src/kseq.h:172-212) into pinned double-buffered chunks,
 * scan per chunk overlapped with the next chunk's decode + H2D, then sort + reduce. */
int tjb_count_sample_files (tjb_ctx *ctx, const char *const *files, int n_files, const tjb_opts *opt,
                            tjb_counter_out *out);
/* Same from sequences already decoded on the host: reads separated by '\n' (pageable or pinned memory). */
int tjb_count_sample_host (tjb_ctx *ctx, const uint8_t *seqs, size_t n_bytes, const tjb_opts *opt,
                           tjb_counter_out *out);
/* Same from sequences resident in device memory (16-byte aligned, reads separated by '\n'). */
int tjb_count_sample_device (tjb_ctx *ctx, const void *d_seqs, size_t n_bytes, const tjb_opts *opt,
                             tjb_counter_out *out);

/* One-shot, re-entrant form of tjb_count_sample_files for the reference's OpenMP loop body: uses a context
 * cached per (calling thread, device). */
int tjb_count_sample (const char *const *files, int n_files, const tjb_opts *opt, int device,
                      tjb_counter_out *out);

void tjb_counter_out_free (tjb_counter_out *out);

/* ---- one sample in pieces (SURVEY.md section 8 e, "by read chunk within a sample"): every GPU scans its share of the
 * reads into its own aggregation table (distinct (base, contexts, length) -> count per strand); the tables are exported
 * as rows, exchanged (NCCL allgather, tatajuba_b200/genome_set.py) and imported (counts added per strand, which is what
 * one scan of all the reads would have counted: + and | are associative); QC, context groups and coverage
 * (hopo_counter.c:367-438) run once, on the complete table, in tjb_sample_finalise().
 *   tjb_sample_begin -> tjb_sample_scan_device (any number of chunks) -> [tjb_sample_export -> exchange ->
 *   tjb_sample_import (the other ranks' rows)] -> tjb_sample_finalise
 * tjb_count_sample_device() is begin + scan + finalise.  A row is row_words = W + 1 uint64: the W key words of the
 * internal record layout (csrc/tjb_common.cuh; W = 2 for k <= 26, 3 above) and forward count | reverse count << 32.
 * TJB_ERR_TOO_LARGE from export / finalise: the table was sized too small for what arrived; begin the sample again
 * (the next tjb_sample_begin sizes it from what was seen). */
int tjb_sample_begin (tjb_ctx *ctx, const tjb_opts *opt, size_t n_bytes_hint);
/* distinct observations the table of the current sample can take, and a way to size the NEXT tjb_sample_begin for a
 * known number (ranks agree on the sum of their exported rows before they import, so that none of them overflows) */
int64_t tjb_ctx_table_capacity (const tjb_ctx *ctx);
int tjb_ctx_reserve_entries (tjb_ctx *ctx, int64_t n_entries);
int tjb_sample_scan_device (tjb_ctx *ctx, const void *d_seqs, size_t n_bytes);
/* table -> rows (once per sample; synchronous).  d_rows = NULL: only *n_rows and *row_words are returned, so that the
 * caller can size its device buffer; otherwise up to max_rows rows are copied to d_rows (device memory). */
int tjb_sample_export (tjb_ctx *ctx, void *d_rows, int64_t max_rows, int64_t *n_rows, int32_t *row_words);
int tjb_sample_import (tjb_ctx *ctx, const void *d_rows, int64_t n_rows);
int tjb_sample_finalise (tjb_ctx *ctx, tjb_counter_out *out);

/* ---- cross-sample exchange (SURVEY.md section 8 e): the step between finalise_hopo_counter() steps 0-4 and the
 * mapper call.  The reference builds one mapper query per indexed context group from the group's first = longest
 * record (src/hopo_counter.c:513-520) and maps every sample separately (src/genome_set.c:75, 90).  A "query row"
 * is that record as 3 x uint64 {context[0], context[1], base | length_field << 8}.  Ranks exchange their rows
 * (NCCL allgather, tatajuba_b200/genome_set.py); tjb_union_queries_device() then returns the DISTINCT rows over
 * all samples in the reference's order (hopo_counter.c:28-38, descending) and, per input row, its index in that
 * union, so that each distinct query is mapped once and the result scattered back to every sample's groups. */
/* rows: 3 * in->n_idx uint64 (host).  Returns n_idx or a negative error. */
int tjb_query_rows (const tjb_counter_out *in, uint64_t *rows);
/* the same rows for the sample this context counted LAST, written on the device from the device-resident result
 * (d_rows: device buffer of 3 * max_rows uint64); nothing crosses PCIe.  Returns n_idx (0 for a dropped sample). */
int tjb_query_rows_device (tjb_ctx *ctx, void *d_rows, int64_t max_rows);
/* d_rows: 3 * n_rows uint64 in DEVICE memory; d_union: device buffer of 3 * n_rows uint64 (the first 3 * *n_union
 * are written); d_index: device int32[n_rows] or NULL.  Synchronous. */
int tjb_union_queries_device (tjb_ctx *ctx, const void *d_rows, int64_t n_rows, int kmer_size, void *d_union,
                              int64_t *n_union, void *d_index);

/* ---- cross-sample merge after the mapper (reference src/genome_set.c:195-289).  Inputs: the per-sample context
 * tables of the SURVIVING samples (remove_empty_genomes, genome_set.c:132-142, already applied; sample ids are the
 * compacted ones, :100), concatenated sample after sample, each sorted by location: location (1-D flattened genome
 * position, context_histogram.h location), base, mode_context_length, and the histogram name ("LEFT-B-RIGHT",
 * name_stride bytes per row, NUL padded).  Outputs, per element of the merged vector (g_tract_vector concat[],
 * genome_set.c:250-289): sample, row index inside the sample, tract_id (:208-222, rows compared with the first row of
 * the current tract by context_histograms_overlap, context_histogram.c:87-110, with -l = levenshtein_distance).
 * Host-side sequential merge (no device work); returns the number of merged rows or a negative error. */
int tjb_merge_tracts (int32_t n_samples, const int32_t *n_hist, const int32_t *location, const int32_t *base,
                      const int32_t *mode_len, const char *names, int32_t name_stride, int32_t levenshtein_distance,
                      int32_t *out_sample, int32_t *out_index, int32_t *out_tract_id);

/* ---- per-sample context histograms after the mapper (reference src/context_histogram.c:224-305, 364-421): what
 * new_genomic_context_list() builds from a hopo_counter once finalise_hopo_counter() has returned.  Input: the mapped
 * elements hc->elem[hc->ref_start .. hc->n_elem) in the order the location sort left them in (hopo_counter.c:562),
 * reference layout (40-byte hopo_element with the mapper's fields filled in).  Output: the genomic_context_list's
 * histograms as flat malloc()ed arrays (release with tjb_context_table_free): one row per context_histogram_t with the
 * fields of src/context_histogram.h:17-43, the context pairs and the (tract length, reads) bins of its empfreq, most
 * frequent first.  The reference's quirks are kept (the last histogram of a sample is dropped, :379-382; the context
 * union of :411-418).  GFF3 features (:331-351) are not looked up here.  Sequential host code; returns n_hist. */
typedef struct {
  int32_t kmer_size;              /* -k */
  int32_t min_tract_size;         /* -m: also the largest location difference inside one histogram (:262) */
  int32_t max_distance_per_flank; /* -d */
  int32_t levenshtein_distance;   /* -l (after the defaulting of src/main.c:192) */
} tjb_hist_opts;

typedef struct {
  int32_t n_hist, name_stride;    /* name_stride = 2 k + 4 bytes per name ("LEFT.B.RIGHT", NUL padded) */
  int32_t *location, *loc2d /* 3 per row */, *base, *neg_strand, *mismatches, *multi, *indel, *integral, *n_context,
          *mode_context_count, *mode_context_length, *mode_context_id, *h_min, *h_max;
  char *names;
  int32_t *ctx_first;             /* n_hist + 1: contexts of row i = contexts[2 * ctx_first[i] .. 2 * ctx_first[i + 1]) */
  uint64_t *contexts;
  int32_t *h_first;               /* n_hist + 1: bins of row i = h_length / h_freq [h_first[i] .. h_first[i + 1]) */
  int32_t *h_length, *h_freq;
} tjb_context_table;

int tjb_context_histograms (const tjb_hist_opts *opt, const void *elem, int32_t n_elem, tjb_context_table *out);
void tjb_context_table_free (tjb_context_table *t);

/* Scan + count of one small batch of reads WITHOUT any allocation.  This entry point has no counterpart in the
 * reference: it was added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  d_seqs: n_bytes of reads separated by '\n' in
 * device memory, n_bytes <= TJB_NOALLOC_MAX_BYTES (else TJB_ERR_NOALLOC_LIMIT, nothing is enqueued).  d_rows (device,
 * room for max_rows rows) receives the sample's aggregation table in the row format of tjb_sample_export(): W + 1
 * 64-bit words per row, W = 2 for kmer_size <= 26 and 3 above (tjb_sample_export's row_words - 1), {key words without
 * the flag bits, forward count | reverse count << 32}, in no particular order; a key may appear in more than one row
 * (add their counts), as in tjb_sample_export().  d_status (device, 3 x int64): [0] rows written, or -1 when the batch
 * did not fit (more than 2^18 tract observations, more distinct ones than the 2^17-slot table takes, or more rows than
 * max_rows); [1] tract observations; [2] skipped runs of non-ACGT bytes.  Everything is enqueued on `stream` (a
 * cudaStream_t cast to void*, NULL = the default stream) and NOTHING is synchronised: read d_status after the stream.
 * The scratch (records, table, counters: 11.9 MB) is one static device array that every GPU which loads the library
 * holds until exit; all calls share it and it keeps the last call's contents (each call clears what must start at
 * zero), so calls that overlap on the GPU (different streams, several host threads) race: one call at a time per GPU.
 * Calls no cudaMalloc / cudaFree / cudaHostAlloc and no allocating library routine. */
#define TJB_NOALLOC_MAX_BYTES ((size_t) 4 << 20)
int tjb_tract_rows_noalloc (int device, const void *d_seqs, size_t n_bytes, const tjb_opts *opt, void *d_rows, int64_t max_rows,
                            int64_t *d_status, void *stream);
size_t tjb_noalloc_max_bytes (void);

/* host FASTQ/FASTA(.gz) decoder on its own (kseq semantics): appends "SEQ\n" per record to dst (capacity cap);
 * returns bytes written (0 = end of file), <0 on error.  Exposed for tests and for callers that feed
 * tjb_count_sample_host themselves. */
typedef struct tjb_reader tjb_reader;
tjb_reader *tjb_reader_open (const char *path);
int64_t tjb_reader_fill (tjb_reader *r, uint8_t *dst, size_t cap, int64_t *n_reads, int64_t *n_bases);
/* a read longer than `cap`: tjb_reader_fill() first returns what the chunk holds, then TJB_ERR_TOO_LARGE; this call
 * hands out the waiting read in the reader's own buffer (no '\n' behind it; valid until the next call on the reader) and
 * returns its length (0: none waiting).  src/kseq.h:172-212 accepts reads of any length, and so does the file path of
 * tjb_count_sample_files(): such a read is copied and scanned as a chunk of its own. */
int64_t tjb_reader_take_long (tjb_reader *r, const uint8_t **seq, int64_t *n_reads, int64_t *n_bases);
void tjb_reader_close (tjb_reader *r);

/* the gzip stream under tjb_reader for regular .gz files (replaces gzopen / gzread of the reference's reader,
 * src/hopo_counter.c:141-153): our own inflate on a producer thread.  tjb_gz_open() returns NULL for anything that is
 * not a regular file starting with a gzip member (the reader then uses zlib, which also reads plain text);
 * tjb_gz_next() hands out the next decoded piece (valid until the following call): bytes, 0 at the end of the file,
 * TJB_ERR_IO for corrupt or truncated data or a CRC-32 / length mismatch, after everything decodable was delivered. */
typedef struct tjb_gzstream tjb_gzstream;
tjb_gzstream *tjb_gz_open (const char *path);
int64_t tjb_gz_next (tjb_gzstream *g, const uint8_t **data);
void tjb_gz_close (tjb_gzstream *g);

#ifdef __cplusplus
}
#endif
#endif
