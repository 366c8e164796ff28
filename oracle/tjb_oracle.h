/* CPU restatement (plain C) of tatajuba's hot path up to the BWA boundary.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this; the product path (tatajuba_b200/, libtjb_gpu.so) never does.
 *
 * Parity status: PINNED — the reference ships no golden vectors for this path (its tests/ only cover the BWA
 * bindings), so this restatement is pinned against the reference's own code compiled here
 * (oracle/_ref/libref_hotpath.so, see oracle/Makefile) on the reference's fixture
 * tests/files/err1750956.fastq.gz and on seeded synthetic/adversarial inputs (tests/test_oracle_vs_reference.py),
 * and against tests/golden/ vectors generated from that compiled reference (tests/golden/make_golden.py).
 *
 * Follows: src/hopo_counter.c:205-258 (scan), :285-307 (key packing), :28-38 + :351 (order),
 * :352-365 (run-length count), :367-374 (QC filter), :388-413 (context groups), :420-438 +
 * submodules/biomcmc-lib/lib/empirical_frequency.c:202-228,254-277 (coverage). */
#ifndef TJB_ORACLE_H
#define TJB_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  uint64_t ctx[2]; /* left / right flanking k-mer, 2 bits per base, first base in the LSBs (hopo_counter.c:304-305) */
  int32_t base;    /* 0 = A/T tract, 1 = C/G tract */
  int32_t length;  /* value of the reference's signed 10-bit field (hopo_counter.h:39) */
  int32_t count;   /* value of the reference's signed 20-bit field (hopo_counter.h:40) */
  int32_t flag;    /* canon_flag: 1 = forward, 2 = reverse-complemented, 3 = both */
} tjbo_rec;

typedef struct {
  int status;                /* same codes as ref_dump_t.status / tjb_counter_out.status */
  int64_t n_raw;             /* emitted tract observations */
  int64_t n_skipped_nonacgt; /* qualifying runs of a non-ACGTU byte with NO earlier tract in the same read: the
                                reference emits uninitialised stack there (hopo_counter.c:222, 233-248), counted and
                                skipped here.  With an earlier tract the reference re-emits that tract's context,
                                base and flag with this run's length (well defined): replicated */
  int64_t n_distinct;        /* after the run-length count, before the QC filter */
  int32_t n_elem, n_idx, coverage;
  tjbo_rec *elem;            /* n_elem records in the reference's order */
  int32_t *idx_initial, *idx_final;
} tjbo_result;

/* buf: reads separated by '\n' */
int tjbo_count_buffer (const uint8_t *buf, size_t n_bytes, int kmer_size, int min_tract_size, int min_coverage,
                       int remove_biased, tjbo_result *out);
/* raw scan only: returns malloc'ed records in emission order (count = 1) */
int64_t tjbo_scan_buffer (const uint8_t *buf, size_t n_bytes, int kmer_size, int min_tract_size,
                          tjbo_rec **recs, int64_t *n_skipped_nonacgt);
void tjbo_free_result (tjbo_result *r);
void tjbo_free (void *p);
/* packs records into the reference's 40-byte hopo_element layout (hopo_counter.h:34-49) as it stands at the
 * BWA boundary: mismatches = 0xffe, multi = neg_strand = 0, read_offset = loc_* = -1 */
void tjbo_pack_hopo_elements (const tjbo_rec *recs, int64_t n, void *out40);

#ifdef __cplusplus
}
#endif
#endif
