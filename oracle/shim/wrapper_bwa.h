/* Stub of the missing submodules/bwa/wrapper_bwa.h (fork leomrtns/bwa @ 661e331, not in this checkout).
 * Declares only what the reference's src/ touches (hopo_counter.c:505-555, context_histogram.c:336-350,
 * genome_set.c:467-470, main.c:148).  Test infrastructure only. */
#ifndef TJB_ORACLE_SHIM_WRAPPER_BWA_H
#define TJB_ORACLE_SHIM_WRAPPER_BWA_H
#include <biomcmc.h>

typedef struct { int dummy; } bwase_options_t;
typedef struct { int query_id, ref_id, position, ref_length, nm, mm, gape, gapo, neg_strand; } bwase_elem_t;
typedef struct { char *name; int len; } shim_ann_t;
typedef struct { int n_seqs; shim_ann_t *anns; } shim_bns_t;
typedef struct bwase_match_struct { bwase_elem_t *m; int n_m; shim_bns_t *bns; } *bwase_match_t;

bwase_options_t new_bwase_options_t (int level);
bwase_match_t new_bwase_match_from_bwa_and_char_vector (const char *fasta, char_vector seqname, char_vector seqs, int n_occ, bwase_options_t opt);
bwase_match_t new_bwase_match_t (const char *fasta);
void del_bwase_match_t (bwase_match_t match);
char *save_bwa_index (const char *genome_filename, const char *suffix, bool overwrite);
#endif
