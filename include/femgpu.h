/* femgpu.h -- C ABI of the B200 (sm_100a) implementation of the `FEM map` hot path.
 *
 * This is the drop-in boundary: plain C types only, one context per GPU.  A maintainer of the
 * reference replaces the per-read calls inside the mapper thread (src/map.c:27-56) by one
 * femgpu_map_batch() per SequenceBatch; INTEGRATION.md shows the exact C stub.  Every entry point
 * names the reference interface it stands in for (file:line relative to the reference tree).
 *
 * Error convention: functions return 0 on success and a negative femgpu_status otherwise; the text
 * is available from femgpu_last_error() (thread local).  Nothing here calls exit() or abort().
 * There is no CPU fallback: without a CUDA device every call fails with FEMGPU_ENODEV.
 */
#ifndef FEMGPU_H_
#define FEMGPU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FEMGPU_ABI_VERSION 5

typedef enum {
  FEMGPU_OK = 0,
  FEMGPU_EINVAL = -1,   /* bad argument (e out of 0..7, a out of 0..2, k/step mismatch ...)   */
  FEMGPU_ENODEV = -2,   /* no usable CUDA device                                               */
  FEMGPU_ENOMEM = -3,   /* device or pinned-host allocation failed                             */
  FEMGPU_ECUDA = -4,    /* a CUDA call or kernel failed                                        */
  FEMGPU_ESTATE = -5,   /* index / reference not uploaded yet                                  */
  FEMGPU_ELIMIT = -6    /* a read needs more scratch than the configured limit                 */
} femgpu_status;

typedef struct femgpu_ctx femgpu_ctx;

/* src/FEM_map.c:63-72 (FEMArgs).  k/step <= 0: use the uploaded index header's values (the
 * evidently intended behaviour); k=12, step=3 reproduces the unmodified reference binary, which
 * never copies the header. */
typedef struct {
  int32_t k;    /* q-gram (window) size, 1..15   */
  int32_t step; /* index step size               */
  int32_t e;    /* error threshold 0..7          */
  int32_t a;    /* additional q-grams 0..2       */
} femgpu_params;

/* src/utils.h:44-49 (Mapping).  16 bytes. */
typedef struct {
  uint64_t candidate_position;  /* (seq_index << 32) | band left edge                          */
  int32_t end_position_offset;  /* mapping end = candidate_position + end_position_offset       */
  uint8_t direction;            /* 0 forward, 1 reverse complement                              */
  uint8_t edit_distance;
  uint8_t reserved[2];
} femgpu_mapping;

/* Start, CIGAR and MD text of one mapping as generate_alignment and generate_MD_tag
 * (src/align.c:279-544) produce them, computed on the device.  88 bytes.  cigar[i] = (length << 2) | op with op 0 M, 1 I, 2 D, in
 * SAM order.  n_cigar == FEMGPU_ALN_HOST: not computed (a character outside "ACGTN" in the read
 * or the reference -- the reference compares raw characters there); the host's
 * femhost_alignment handles those. */
#define FEMGPU_ALN_HOST 0xFF
#define FEMGPU_MD_MAX 47
typedef struct {
  int16_t start;      /* 0-based, relative to candidate_position (src/align.c:78)           */
  uint8_t n_cigar;
  uint8_t md_len;     /* characters of md; FEMGPU_ALN_HOST: not computed (longer than FEMGPU_MD_MAX) */
  uint16_t cigar[18];
  char md[48];        /* the MD text of generate_MD_tag (src/align.c:501-544), not NUL terminated   */
} femgpu_alignment;

/* One batch of reads: what a SequenceBatch (src/sequence_batch.h:15-24) holds, flattened.
 * bases: the reads' sequence characters back to back (any case, non-ACGT = ambiguous);
 * offsets[i]..offsets[i+1] delimit read i.  Host memory, pinned or pageable. */
typedef struct {
  const char *bases;
  const uint64_t *offsets; /* n_reads + 1 entries */
  uint32_t n_reads;
} femgpu_batch;

/* src/utils.h:55-61 (MappingStats) plus the count of reads the reference cannot process
 * (undefined behaviour there, reported and skipped here -- SURVEY.md F2). */
typedef struct {
  uint64_t num_reads;
  uint64_t num_mapped_reads;
  uint64_t num_candidates_before_filter;
  uint64_t num_candidates;
  uint64_t num_mappings;
  uint64_t num_infeasible_reads;
} femgpu_stats;

/* Result of one batch.  All pointers are owned by the context (pinned host memory) and stay valid
 * until the next call that uses the same context slot.  mappings of read i are
 * maps[first[i] .. first[i+1]) in the reference's emission order *before* process_mappings'
 * sort (src/map.c:36,47: forward-strand hits in candidate order, then reverse-strand hits);
 * femgpu_sort_mappings() applies src/align.c:57. */
typedef struct {
  const uint64_t *first;      /* n_reads + 1 */
  const femgpu_mapping *maps; /* first[n_reads] records */
  femgpu_stats stats;         /* this batch only */
  const femgpu_alignment *alns; /* parallel to maps; NULL when "device_traceback" is off */
  /* SAM text formatted on the device (femgpu_set_sam_inputs before the submit): the records of read i,
   * in the reference's order (src/align.c:56-92), are sam_text[sam_off[i] .. sam_off[i+1]).  Reads with
   * sam_host[i] != 0 were left to the host (more than 64 mappings: src/ksort.h's unstable radix order;
   * a character outside "ACGTN"; an MD text longer than FEMGPU_MD_MAX) and own no bytes.  NULL
   * when no inputs were set. */
  const char *sam_text;
  const uint64_t *sam_off;  /* n_reads + 1 */
  const uint8_t *sam_host;  /* n_reads */
} femgpu_result;

/* What the SAM records need beyond the bases (src/align.c:546-632 reads them from the kseq record):
 * name i = names[name_off[i] .. name_off[i+1] - 1) followed by one NUL (back to back, as kseq's name.s
 * copied with its terminator); quals share the batch's offsets, has_qual[i] == 0 prints '*'.
 * Host memory, valid until the submit returns. */
typedef struct {
  const char *names;
  const uint64_t *name_off; /* n_reads + 1 */
  const char *quals;
  const uint8_t *has_qual;  /* n_reads */
  /* Optional (all NULL/0: the library's own buffers, valid until the slot's next submit): pinned host
   * memory of the caller that femgpu_wait_batch copies the text, the n_reads + 1 offsets and the
   * n_reads flags into, so that the result outlives the slot.  A text larger than text_cap lands in the
   * library's buffer instead (result.sam_text != text_out tells). */
  char *text_out;
  uint64_t text_cap;
  uint64_t *off_out;
  uint8_t *host_out;
} femgpu_sam_inputs;

/* ---- life cycle ------------------------------------------------------------------------------ */
int femgpu_abi_version(void);
int femgpu_device_count(void);
/* One context per GPU; owns its streams, device copies of index + reference, staging buffers. */
femgpu_ctx *femgpu_create(int device);
void femgpu_destroy(femgpu_ctx *ctx);
const char *femgpu_last_error(void);

/* Page-locked host memory for batch staging (replaces the kstring buffers a SequenceBatch owns,
 * src/sequence_batch.h:15-24).  Returns NULL on failure. */
void *femgpu_alloc_pinned(size_t bytes);
void femgpu_free_pinned(void *p);

/* Tunables (all optional): "blocking_sync" (1: host waits sleep on an event instead of spinning; for hosts
 * with few cores per GPU), "scratch_limit_bytes", "arena_candidates", "sort_capacity",
 * "acc_capacity", "l2_persist_lut", "seed_dp_mode" (0 automatic, 1 lane per group, 2 warp scan, 3 lane per group with several read-strands per warp),
 * "cand_mode" (0 automatic: near-pair path when a >= 1, hash path when a = 0; 1 sort path only;
 * 2 hash path first; 3 near-pair path first; 4 all-groups-at-once path first, the automatic
 * choice on sparse indexes; 5 the near-pair path's register form (k_cand_reg) whatever the index density;
 * 6 the near-pair path without its register form), "small_shape" (k_cand_small: 0 automatic, 1 one seed per lane
 * when step x Q <= 32, 4 up to 128 seeds), "hash_slots", "pair_bits", "pair_values" (0 automatic),
 * "verify_planes" (1 default: k_verify computes candidates without ambiguous bases from bit planes; 0: the general
 * form for all), "device_traceback" (1 default: k_trace fills femgpu_result.alns).  Returns FEMGPU_EINVAL for unknown keys. */
int femgpu_set_option(femgpu_ctx *ctx, const char *key, int64_t value);

/* ---- resident data --------------------------------------------------------------------------- */
/* Stands in for load_index (src/index.h:19, src/index.c:100-131): the arrays of the index file
 * (`int k; int step; uint32 L[4^k+1]; size_t n; uint64 O[n]`, src/index.c:133-168) go to HBM
 * verbatim.  Host pointers may be pageable. */
int femgpu_upload_index(femgpu_ctx *ctx, int32_t k, int32_t step, const uint32_t *lookup_table,
                        size_t lookup_table_size, const uint64_t *occurrence_table,
                        size_t occurrence_table_size);
/* Reads the index file itself (same layout) and uploads it in chunks without holding a second
 * full host copy. */
int femgpu_upload_index_file(femgpu_ctx *ctx, const char *index_path);
/* Stands in for the reference SequenceBatch handed to every call (src/map.h:59): n records of
 * raw characters; packed on the device to 2 bits per base plus an ambiguity mask. */
int femgpu_upload_reference(femgpu_ctx *ctx, uint32_t n_records, const char *const *bases,
                            const uint32_t *lengths);
/* Stands in for construct_index (src/index.c:57-98) behind `FEM index`: builds the lookup table and
 * the occurrence table in HBM from the uploaded reference -- the same arrays, entry for entry, as the
 * reference's two radix sorts produce.  The result is resident exactly as after femgpu_upload_index. */
int femgpu_build_index(femgpu_ctx *ctx, int32_t k, int32_t step);
/* Sizes of the resident index: 4^k + 1 lookup-table words, n occurrence words. */
int femgpu_index_sizes(const femgpu_ctx *ctx, size_t *lookup_table_size, size_t *occurrence_table_size);
/* The resident arrays back to host memory (sizes must match femgpu_index_sizes). */
int femgpu_download_index(femgpu_ctx *ctx, uint32_t *lookup_table, size_t lookup_table_size,
                          uint64_t *occurrence_table, size_t occurrence_table_size);
/* Stands in for save_index (src/index.c:133-168): the resident index in the frozen file layout. */
int femgpu_save_index_file(femgpu_ctx *ctx, const char *index_path);
int femgpu_index_k(const femgpu_ctx *ctx);
int femgpu_index_step(const femgpu_ctx *ctx);

/* ---- the hot path ---------------------------------------------------------------------------- */
/* Whole path for a batch = src/map.c:27-49 for every read: generate_group_seeding_candidates
 * (src/filter.h:9) and verify_candidates (src/align.h:13) on both strands.  Synchronous:
 * H2D of the batch, kernels, D2H of the compacted mappings all happen inside the call. */
int femgpu_map_batch(femgpu_ctx *ctx, const femgpu_params *params, const femgpu_batch *batch,
                     femgpu_result *result);

/* Asynchronous form for double buffering (replaces the InputQueue/OutputQueue hand-off,
 * src/input_queue.c:34-51, src/output_queue.c:46-58): `slot` is 0 or 1; submit returns once the
 * batch has been staged, wait blocks until its result is in host memory. */
int femgpu_submit_batch(femgpu_ctx *ctx, int slot, const femgpu_params *params,
                        const femgpu_batch *batch);
int femgpu_wait_batch(femgpu_ctx *ctx, int slot, femgpu_result *result);
/* SAM text on the device (stands in for process_mappings + generate_bam1_t + sam_write1,
 * src/align.c:56-92,546-632, htslib sam.c:3004-3061).  femgpu_set_reference_names once after
 * femgpu_upload_reference (names as femhost_sam_header prints them); femgpu_set_sam_inputs applies to
 * the NEXT submit on that slot only. */
int femgpu_set_reference_names(femgpu_ctx *ctx, uint32_t n_records, const char *const *names);
/* Several GPUs in one process: the index, the packed reference and the record names resident in `src`
 * copied device to device into `dst` (peer copies over NVLink where the devices are peers), instead of
 * a second load_index / load_all_sequences (src/FEM_map.c:136-143, src/index.c:100-131) per device. */
int femgpu_clone_resident(femgpu_ctx *dst, const femgpu_ctx *src);
int femgpu_set_sam_inputs(femgpu_ctx *ctx, int slot, const femgpu_sam_inputs *inputs);

/* Stage-level entry points (tests, profiling, and callers that keep their own verification).
 * Candidates: generate_group_seeding_candidates (src/filter.c:146-223) for both strands of every
 * read.  Strand s of read i is "read-strand" 2*i+s; its candidates are
 * cands[first[2i+s] .. first[2i+s+1]); before[2i+s] is its contribution to the "candidates before
 * additional q-gram filter" counter; status[2i+s]: 0 ok, 1 reference returns 0 early
 * (src/filter.c:168-172,180-182), 2 infeasible (reference UB).  Pointers: context-owned pinned
 * memory, valid until the next call on the context. */
typedef struct {
  const uint64_t *first;   /* 2*n_reads + 1 */
  const uint64_t *cands;
  const uint32_t *before;  /* 2*n_reads */
  const uint8_t *status;   /* 2*n_reads */
} femgpu_candidates;
int femgpu_candidates_batch(femgpu_ctx *ctx, const femgpu_params *params,
                            const femgpu_batch *batch, femgpu_candidates *out);

/* src/align.c:53-57 + src/ksort.h:98-151: the reference's per-read mapping order (host side,
 * used by the SAM writer; exported so callers get the same permutation on ties). */
void femgpu_sort_mappings(femgpu_mapping *maps, size_t n);
/* the same, carrying the parallel array of device alignments along (alns may be NULL) */
void femgpu_sort_mappings_aln(femgpu_mapping *maps, femgpu_alignment *alns, size_t n);

/* ---- an entry point that allocates nothing ---------------------------------------------------- */
/* No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The lookup table of the index (the prefix sums of src/index.c:57-98 that femgpu_build_index computes
 * with a cudaMalloc'ed buffer of tile sums) from per-bucket counts: d_lut[0] = 0,
 * d_lut[h + 1] = d_counts[0] + ... + d_counts[h] (uint32 sums; the total must fit, as it does for any
 * index the reference can write).  d_counts (n_buckets entries) and d_lut (n_buckets + 1 entries) are
 * DEVICE pointers of the caller on `device`; stream is a cudaStream_t (NULL: the default stream).
 * The three kernels are queued on that stream and the call returns without waiting for them.
 * The tile sums live in a static device array inside the library (32 KB per GPU, held until exit),
 * which bounds the size: 1 <= n_buckets <= FEMGPU_NOALLOC_MAX_BUCKETS (4^12, the reference's default
 * window), FEMGPU_EINVAL with a message beyond.  The call makes no cudaMalloc, cudaFree or other
 * allocating call; the first call per device looks the array's address up (cudaGetSymbolAddress),
 * so make one call before a graph capture starts.  Every call on a device shares the array: calls
 * that overlap on the GPU (two streams, two threads) race -- order them.  Nothing in it has to
 * start at zero (each call overwrites the part it reads). */
#define FEMGPU_NOALLOC_MAX_BUCKETS 16777216ull
int femgpu_lookup_table_noalloc(int device, const uint32_t *d_counts, uint64_t n_buckets, uint32_t *d_lut,
                                void *stream);

/* ---- measurement hooks ------------------------------------------------------------------------ */
/* Device time (CUDA events on the context's stream) of the kernels of the last finished batch, in
 * milliseconds: [0] pack+seed, [1] candidates, [2] verify, [3] compaction, [4] traceback,
 * [5] whole device pipeline including H2D/D2H.  n_launches receives the number of kernel launches. */
int femgpu_last_timings(femgpu_ctx *ctx, float ms[6], uint32_t *n_launches);
/* Work counters of the last finished batch, the numerators of the roofline figures: [0] banded
 * Myers columns actually computed by the verification kernel, [1] candidates verified,
 * [2] occurrence-table entries of the selected seeds ("candidates before additional q-gram
 * filter", src/filter.c:202), [3] mappings emitted. */
int femgpu_last_work(femgpu_ctx *ctx, uint64_t work[4]);
/* Run the device pipeline of the last submitted batch again with inputs already resident in HBM
 * (no H2D, no D2H of mappings); returns device milliseconds.  bench.py's `value`.  ms_stage: seeds,
 * candidates, verification, compaction, traceback, SAM text (0 unless the batch had SAM inputs). */
int femgpu_rerun_resident(femgpu_ctx *ctx, float *ms_total, float ms_stage[6]);

/* INT32 pipe peak of the device in 10^9 instructions per second, measured with dependency-chain
 * free kernels: [0] LOP3 (logic), [1] IADD3 (add), [2] shift+logic pairs as in the Myers column.
 * The denominator of the verification kernel's roofline. */
int femgpu_measure_int32_peak(femgpu_ctx *ctx, double giops[3]);

#ifdef __cplusplus
}
#endif
#endif /* FEMGPU_H_ */
