/* fem_oracle.h -- CPU restatement of the reference's `FEM map` hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under fem_b200/ may include, link or call this.  Only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg use it, as the checker.
 *
 * Parity status: PINNED against the reference itself.  The reference holds no golden vectors or
 * tests for this path, so the pin is (a) oracle/_ref/FEM_ref_{asis,hdr}, the
 * reference compiled from /root/reference by oracle/Makefile, run on seeded synthetic inputs
 * (tests/test_oracle_vs_ref.py, executed wherever oracle/_ref exists), and (b) the SAM files those
 * binaries produced, committed under tests/golden/ with the generating script.
 *
 * Every function cites the reference file:line it restates (paths relative to /root/reference).
 */
#ifndef FEM_ORACLE_H_
#define FEM_ORACLE_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fo_ctx fo_ctx;

/* Mapping record, src/utils.h:44-49 */
typedef struct {
  uint64_t candidate_position; /* (seq_index<<32 | pos) of the band's left edge          */
  int32_t end_position_offset; /* end = candidate + offset                                */
  uint8_t direction;           /* 0 forward, 1 reverse complement                         */
  uint8_t edit_distance;
  uint8_t pad_[2];
} fo_mapping;

/* counters printed at src/FEM_map.c:214-218 (+ one of ours: reads the reference cannot handle) */
typedef struct {
  uint64_t num_reads, num_mapped_reads, num_candidates_before_filter, num_candidates, num_mappings;
  uint64_t num_infeasible_reads;
} fo_stats;

/* ---- context: reference records + index ---------------------------------------------------- */
/* Load a FASTA (plain or gz) the way src/sequence_batch.c:82-121 does. */
fo_ctx *fo_open_reference(const char *fasta_path);
/* Build from memory: n records, seqs[i] of lens[i] chars, names[i] NUL-terminated. */
fo_ctx *fo_create_reference(uint32_t n, const char *const *names, const char *const *seqs,
                            const uint32_t *lens);
void fo_close(fo_ctx *ctx);
uint32_t fo_num_records(const fo_ctx *ctx);
uint32_t fo_record_length(const fo_ctx *ctx, uint32_t i);
const char *fo_record_name(const fo_ctx *ctx, uint32_t i);
const char *fo_record_bases(const fo_ctx *ctx, uint32_t i);

/* src/index.c:57-98 (construct), :100-131 (load), :133-168 (save) */
int fo_build_index(fo_ctx *ctx, int k, int step);
int fo_load_index(fo_ctx *ctx, const char *index_path);
int fo_save_index(const fo_ctx *ctx, const char *index_path);
int fo_index_k(const fo_ctx *ctx);
int fo_index_step(const fo_ctx *ctx);
size_t fo_index_num_occurrences(const fo_ctx *ctx);
const uint32_t *fo_index_lookup_table(const fo_ctx *ctx);
const uint64_t *fo_index_occurrences(const fo_ctx *ctx);

/* seeding parameters used by map (src/FEM_map.c:67-72).  k/step <= 0 -> take the index header's
 * ("hdr" oracle); pass 12/3 to model the unmodified binary. */
int fo_set_params(fo_ctx *ctx, int k, int step, int e, int a);

/* ---- stage-level functions ------------------------------------------------------------------ */
/* src/sequence_batch.h:90-98 */
void fo_reverse_complement(const char *seq, int l, char *out);
/* src/utils.h:101-117; returns the ambiguous-seed count */
int fo_hash_all_seeds(const char *text, int l, int k, uint32_t *hashes);
/* src/filter.c:146-223.  `text` is the strand-oriented read.  Returns the candidate count, or -1
 * when the reference would run into undefined behaviour -- no candidates are written.
 * *n_before receives the "candidates before additional q-gram filter" contribution. */
long fo_generate_candidates(const fo_ctx *ctx, const char *text, int l, uint64_t **cands,
                            size_t *cap, uint32_t *n_before);
/* src/align.c:4-51 (8x16-bit lanes for full groups of 8, 32-bit scalar for the tail) */
uint32_t fo_verify_candidates(const fo_ctx *ctx, const char *text, int l, int direction,
                              const uint64_t *cands, uint32_t n, fo_mapping *out);
/* single-candidate kernels: src/align.c:102-147 (width 32) and one lane of :149-277 (width 16,
 * without the all-lanes early exit) */
int fo_banded_edit_distance(const fo_ctx *ctx, uint64_t cand, const char *text, int l, int width,
                            int *end_offset);
/* src/align.c:53-57 + src/ksort.h:98-151 */
void fo_sort_mappings(fo_mapping *m, size_t n);
/* src/align.c:279-544.  cigar (BAM encoding) must hold l+1 words, md 4*l+16 chars.  Returns the
 * 0-based start relative to the candidate (src/align.c:78). */
int fo_generate_alignment(const fo_ctx *ctx, uint64_t cand, const char *text, int l,
                          int edit_distance, int end_offset, uint32_t *cigar, int *n_cigar,
                          char *md);

/* ---- whole reads ---------------------------------------------------------------------------- */
/* src/map.c:27-56 for one read: returns the number of mappings (sorted, src/align.c:57) written to
 * *maps (grown with realloc), updates stats. */
size_t fo_map_read(const fo_ctx *ctx, const char *seq, int l, fo_mapping **maps, size_t *cap,
                   fo_stats *stats);
/* src/align.c:56-92 + htslib sam.c:3004-3061: SAM text of one read's sorted mappings, appended
 * to *buf (realloc-grown, len and cap in bytes). */
void fo_format_sam(const fo_ctx *ctx, const char *name, const char *seq, const char *qual, int l,
                   const fo_mapping *maps, size_t n, char **buf, size_t *len, size_t *cap);
/* src/output_queue.c:93-116 */
void fo_format_header(const fo_ctx *ctx, char **buf, size_t *len, size_t *cap);
/* `FEM map` end to end on files, -t 1 order (src/FEM_map.c:57-227).  Returns 0 on success. */
int fo_map_file(const fo_ctx *ctx, const char *fastq_path, const char *sam_path, fo_stats *stats);

#ifdef __cplusplus
}
#endif
#endif
