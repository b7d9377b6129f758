/* femhost.h -- host side (plain C) of the B200 `FEM map` path: everything the reference does around
 * the mapping kernel -- sequence files, the index file, per-read mapping order, alignment
 * traceback, SAM text -- plus the batch pipeline that feeds the GPUs through include/femgpu.h.
 *
 * Each entry point names the reference interface it replaces (file:line relative to the reference
 * tree).  Error convention as in femgpu.h: 0 on success, negative on failure, text from
 * femhost_last_error(); nothing here calls exit().
 */
#ifndef FEMHOST_H_
#define FEMHOST_H_

#include <stddef.h>
#include <stdint.h>

#include "femgpu.h"

#ifdef __cplusplus
extern "C" {
#endif

const char *femhost_last_error(void);

/* ---- sequence records ------------------------------------------------------------------------ */
/* A SequenceBatch (src/sequence_batch.h:15-24) flattened: record i has
 *   bases  [base_off[i], base_off[i+1])      raw characters as they stand in the file
 *   quals  same range (only meaningful when has_qual[i])
 *   name   names + name_off[i], NUL terminated (text up to the first white space, src/kseq.h:196)
 * Buffers are grow-only and reused from batch to batch. */
typedef struct {
  uint32_t n_records;
  uint64_t n_bases;
  char *bases;
  char *quals;
  uint64_t *base_off; /* n_records + 1 */
  uint8_t *has_qual;
  char *names;
  uint64_t *name_off; /* n_records + 1 */
  /* private */
  size_t cap_records, cap_bases, cap_names;
  int pinned; /* bases/base_off come from femgpu_alloc_pinned */
} femhost_records;

void femhost_records_init(femhost_records *r, int pinned);
void femhost_records_free(femhost_records *r);
static inline uint32_t femhost_record_length(const femhost_records *r, uint32_t i) {
  return (uint32_t)(r->base_off[i + 1] - r->base_off[i]);
}
static inline const char *femhost_record_name(const femhost_records *r, uint32_t i) {
  return r->names + r->name_off[i];
}

/* FASTA/FASTQ reader, plain or gzip (src/sequence_batch.c:30-37 + src/kseq.h:185-226 record
 * semantics: zero-length records are skipped, a truncated quality string is an error). */
typedef struct femhost_reader femhost_reader;
femhost_reader *femhost_reader_open(const char *path);
void femhost_reader_close(femhost_reader *rd);
/* src/sequence_batch.c:44-80: up to max_records records replace the content of *out.
 * Returns the number loaded (0 at end of file) or a negative error. */
int64_t femhost_reader_next(femhost_reader *rd, uint32_t max_records, femhost_records *out);
/* src/sequence_batch.c:82-121 */
int femhost_load_all(const char *path, femhost_records *out);
/* The same records for a plain four-line FASTQ file, read the way femhost_map_file's parallel
 * reader does: independent byte ranges of chunk_bytes that resynchronise on record starts.
 * Returns 1 (and callers fall back to femhost_load_all) when the file is not of that shape. */
int femhost_load_all_ranges(const char *path, uint64_t chunk_bytes, femhost_records *out);
/* The records femhost_map_file maps for byte-range shard shard_index of shard_count (see
 * femhost_map_options); returns 1 when the file is not plain four-line FASTQ. */
int femhost_load_shard(const char *path, uint32_t shard_index, uint32_t shard_count, femhost_records *out);
/* Number of records the readers deliver for a file of any accepted format (FASTA/FASTQ, multi-line,
 * gzip; zero-length records are skipped as in src/sequence_batch.c:50-52), or a negative error. */
int64_t femhost_count_records(const char *path);

/* ---- index ------------------------------------------------------------------------------------ */
/* src/index.h:7-14 */
typedef struct {
  int32_t k, step;
  uint32_t *lut; /* 4^k + 1 prefix sums */
  size_t lut_n;
  uint64_t *occ; /* (record << 32) | position, ascending inside a bucket */
  size_t n_occ;
} femhost_index;
/* src/index.c:57-98 (`FEM index`): same arrays, built by counting sort */
int femhost_index_build(const femhost_records *ref, int k, int step, femhost_index *out);
/* the same file written from a GPU (femgpu_build_index + femgpu_save_index_file); seconds (may be
 * NULL) receives the upload, build and save times */
int femhost_index_build_device(const femhost_records *ref, int k, int step, int device, const char *out_path,
                               double seconds[3]);
/* src/index.c:133-168 and :100-131; the file layout is frozen */
int femhost_index_save(const femhost_index *idx, const char *path);
int femhost_index_load(const char *path, femhost_index *out);
void femhost_index_free(femhost_index *idx);

/* ---- SAM ---------------------------------------------------------------------------------------- */
/* growable text buffer */
typedef struct {
  char *s;
  size_t l, m;
} femhost_str;
void femhost_str_free(femhost_str *s);

/* src/output_queue.c:93-116: "@SQ\tSN:<name>\tLN:<len>\n" per record, nothing else */
int femhost_sam_header(const femhost_records *ref, femhost_str *out);
/* src/align.c:279-544 for one mapping: CIGAR in BAM encoding (cigar must hold l+2 words), MD text
 * (md must hold 4*l+64 chars).  `window` = reference characters from the candidate position on,
 * `text` = the read in the mapping's orientation.  Returns the 0-based start relative to the
 * window (src/align.c:78), or FEMHOST_ALIGN_ERROR where the reference would assert. */
#define FEMHOST_ALIGN_ERROR (-0x40000000)
int femhost_alignment(const char *window, const char *text, int l, int e, int edit_distance,
                      int end_offset, uint32_t *cigar, int *n_cigar, char *md);
/* src/align.c:56-92 + htslib sam.c:3004-3061 for one read: sorts maps[0..n) in place with the
 * reference's own order (src/align.c:57) and appends one SAM line per mapping.  alns (may be
 * NULL) are the device's start/CIGAR per mapping (femgpu_result.alns), permuted along; entries
 * marked FEMGPU_ALN_HOST and all entries of a NULL array are computed here by femhost_alignment. */
int femhost_sam_read(const femhost_records *ref, const char *name, const char *seq,
                     const char *qual, int l, int e, femgpu_mapping *maps, femgpu_alignment *alns,
                     size_t n, femhost_str *out);

/* ---- the `FEM map` engine ---------------------------------------------------------------------- */
/* src/FEM_map.c:136-143: reference records + index, resident on n_gpus devices. */
typedef struct femhost_mapper femhost_mapper;
femhost_mapper *femhost_mapper_create(const char *ref_path, const char *index_path, int n_gpus,
                                      const int *gpu_ids);
void femhost_mapper_destroy(femhost_mapper *m);
const femhost_records *femhost_mapper_reference(const femhost_mapper *m);
femgpu_ctx *femhost_mapper_context(femhost_mapper *m, int i);

typedef struct {
  femgpu_params params;   /* src/FEM_map.c:67-70; k/step <= 0: the index header's            */
  int32_t n_threads;      /* -t: host threads that build SAM text                            */
  uint32_t batch_reads;   /* reads per GPU batch (0 = default)                                */
  uint64_t skip_reads;    /* one process per GPU: this process maps reads                      */
  uint64_t max_reads;     /*   [skip_reads, skip_reads + max_reads) of the file (0 = to the end)*/
  int32_t no_header;      /* and only the first shard prints the @SQ lines                     */
  int32_t no_reads;       /* this process maps nothing (more shards than reads); max_reads = 0 means "to the end" */
  uint32_t shard_index;   /* shard_count > 1: this process maps the records that START in byte range            */
  uint32_t shard_count;   /*   [size*index/count, size*(index+1)/count) of a plain four-line FASTQ file -- no    */
                          /*   counting pass, no skipped reads; other inputs return FEMHOST_ENOTFASTQ4          */
} femhost_map_options;
/* femhost_map_file with shard_count > 1 on input that is not plain four-line FASTQ (FASTA, gzip, CRLF,
 * multi-line records): every rank must then fall back to read-count shards (skip_reads / max_reads
 * from femhost_count_records), see fem_b200/shard.py */
#define FEMHOST_ENOTFASTQ4 (-5)

/* per-phase host seconds of the last run (summed over threads where parallel) */
typedef struct {
  double wall, parse, gpu_wait, format, write;
  uint64_t h2d_bytes, d2h_bytes, sam_bytes;
  uint32_t n_batches, n_launches;
  uint64_t device_sam_reads; /* mapped reads whose records the device formatted (k_sam) */
} femhost_map_timing;

/* src/FEM_map.c:172-219, the reference's "Time:" window: read the FASTQ, map, write the SAM file
 * in input order (the reference's -t 1 order).  sam_path NULL: build the text but write nothing. */
int femhost_map_file(femhost_mapper *m, const femhost_map_options *opt, const char *reads_path,
                     const char *sam_path, femgpu_stats *stats, femhost_map_timing *timing);

#ifdef __cplusplus
}
#endif
#endif /* FEMHOST_H_ */
