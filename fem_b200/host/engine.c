/* engine.c -- the `FEM map` engine: reference + index resident on the GPUs, and the batch pipeline
 * that replaces the reference's reader thread / mapper threads / writer thread and their two
 * mutex-protected rings (src/input_queue.c, src/map.c, src/output_queue.c, src/FEM_map.c:146-198):
 *
 *   reader thread      FASTQ -> pinned flat batches                      (fastx.c)
 *   2 threads per GPU  one femgpu slot each: H2D, kernels, D2H, so that one batch's copies and
 *                      host hand-shakes overlap the other batch's kernels on the same device
 *   -t worker threads  per-read mapping order, traceback, SAM text, chunk by chunk (sam.c)
 *   writer thread      batches in file order -> output: byte-identical to the reference's -t 1
 */
#define _GNU_SOURCE /* fallocate, MADV_POPULATE_WRITE */
#include <fcntl.h>
#include <inttypes.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/vfs.h>
#include <unistd.h>

#include "femhost.h"
#include "host_internal.h"

/* reads per GPU batch: small enough that the first SAM text is on its way a few milliseconds into the
 * run and the last batch's tail is short, large enough to fill the device (measured on cfg2: 65536
 * beats 131072 by 10-20 %, 262144 loses 30 %) */
#define DEFAULT_BATCH_READS 65536u

#define PREALLOC_AHEAD (48ull << 20) /* prealloc_main: bytes of output pages kept allocated ahead of the text */
#define PREALLOC_STEP (4ull << 20)

/* ---- blocking FIFO of pointers ---------------------------------------------------------------- */
typedef struct {
  void **item;
  size_t cap, head, count;
  int closed;
  pthread_mutex_t mu;
  pthread_cond_t not_empty, not_full;
} fifo;

static void fifo_init(fifo *q, size_t cap) {
  memset(q, 0, sizeof *q);
  q->item = (void **)calloc(cap, sizeof(void *));
  q->cap = cap;
  pthread_mutex_init(&q->mu, NULL);
  pthread_cond_init(&q->not_empty, NULL);
  pthread_cond_init(&q->not_full, NULL);
}
static void fifo_destroy(fifo *q) {
  free(q->item);
  pthread_mutex_destroy(&q->mu);
  pthread_cond_destroy(&q->not_empty);
  pthread_cond_destroy(&q->not_full);
}
static void fifo_push(fifo *q, void *p) {
  pthread_mutex_lock(&q->mu);
  while (q->count == q->cap) pthread_cond_wait(&q->not_full, &q->mu);
  q->item[(q->head + q->count++) % q->cap] = p;
  pthread_cond_signal(&q->not_empty);
  pthread_mutex_unlock(&q->mu);
}
/* NULL once the queue is closed and drained */
static void *fifo_pop(fifo *q) {
  pthread_mutex_lock(&q->mu);
  while (q->count == 0 && !q->closed) pthread_cond_wait(&q->not_empty, &q->mu);
  void *p = NULL;
  if (q->count) {
    p = q->item[q->head];
    q->head = (q->head + 1) % q->cap;
    --q->count;
    pthread_cond_signal(&q->not_full);
  }
  pthread_mutex_unlock(&q->mu);
  return p;
}
static void fifo_close(fifo *q) {
  pthread_mutex_lock(&q->mu);
  q->closed = 1;
  pthread_cond_broadcast(&q->not_empty);
  pthread_mutex_unlock(&q->mu);
}

/* ---- mapper: resident data ---------------------------------------------------------------------- */
#define FEM_MAX_GPUS 16
struct batch;
struct femhost_mapper {
  femhost_records ref;
  int n_gpus;
  femgpu_ctx *ctx[FEM_MAX_GPUS];
  /* batch objects (pinned staging, result copies, SAM text buffers) live as long as the mapper,
   * so a second run pays for none of their allocations */
  struct batch *pool;
  size_t n_pool;
};
static void pool_free(femhost_mapper *m);

femhost_mapper *femhost_mapper_create(const char *ref_path, const char *index_path, int n_gpus, const int *gpu_ids) {
  if (!ref_path || !index_path || n_gpus < 1 || n_gpus > FEM_MAX_GPUS) {
    femhost_set_error(-1, "mapper: bad arguments");
    return NULL;
  }
  femhost_mapper *m = (femhost_mapper *)calloc(1, sizeof *m);
  femhost_records_init(&m->ref, 0);
  femhost_index idx;
  memset(&idx, 0, sizeof idx);
  if (femhost_load_all(ref_path, &m->ref)) goto fail;
  if (m->ref.n_records == 0) { femhost_set_error(-1, "reference %s holds no sequence", ref_path); goto fail; }
  const char **seqs = (const char **)malloc(m->ref.n_records * sizeof(char *));
  uint32_t *lens = (uint32_t *)malloc(m->ref.n_records * sizeof(uint32_t));
  for (uint32_t i = 0; i < m->ref.n_records; ++i) {
    seqs[i] = m->ref.bases + m->ref.base_off[i];
    lens[i] = femhost_record_length(&m->ref, i);
  }
  int rc = 0;
  /* the file is read and the reference packed once, on the first device; the others receive copies
   * device to device (FEM_CLONE_RESIDENT=0: every device loads for itself) */
  const int clone = !(getenv("FEM_CLONE_RESIDENT") && atoi(getenv("FEM_CLONE_RESIDENT")) == 0);
  if (n_gpus > 1 && !clone && femhost_index_load(index_path, &idx)) rc = -1;
  for (int g = 0; g < n_gpus && !rc; ++g) {
    m->ctx[g] = femgpu_create(gpu_ids ? gpu_ids[g] : g);
    if (!m->ctx[g]) { femhost_set_error(-2, "%s", femgpu_last_error()); rc = -2; break; }
    m->n_gpus = g + 1;
    if (g > 0 && clone) rc = femgpu_clone_resident(m->ctx[g], m->ctx[0]);
    else {
      rc = n_gpus > 1 && !clone ? femgpu_upload_index(m->ctx[g], idx.k, idx.step, idx.lut, idx.lut_n, idx.occ, idx.n_occ)
                                : femgpu_upload_index_file(m->ctx[g], index_path);
      if (!rc) rc = femgpu_upload_reference(m->ctx[g], m->ref.n_records, seqs, lens);
      if (!rc) { /* record names for the device's SAM text (k_sam) */
        const char **rn = (const char **)malloc(m->ref.n_records * sizeof(char *));
        for (uint32_t i = 0; i < m->ref.n_records; ++i) rn[i] = femhost_record_name(&m->ref, i);
        rc = femgpu_set_reference_names(m->ctx[g], m->ref.n_records, rn);
        free(rn);
      }
    }
    const char *tb = getenv("FEM_DEVICE_TRACEBACK"); /* "0": start/CIGAR of every mapping on the host */
    if (!rc && tb) rc = femgpu_set_option(m->ctx[g], "device_traceback", atoi(tb));
    if (rc) femhost_set_error(rc, "%s", femgpu_last_error());
  }
  free(seqs);
  free(lens);
  femhost_index_free(&idx);
  if (rc) goto fail;
  return m;
fail:
  femhost_mapper_destroy(m);
  return NULL;
}

/* `FEM index` on a GPU: the records go to HBM 2-bit packed, femgpu_build_index produces the arrays of
 * src/index.c:57-98 there, femgpu_save_index_file streams them into the frozen file layout. */
int femhost_index_build_device(const femhost_records *ref, int k, int step, int device, const char *out_path,
                               double seconds[3]) {
  if (!ref || !out_path || ref->n_records == 0) return femhost_set_error(-1, "index: bad arguments");
  femgpu_ctx *ctx = femgpu_create(device);
  if (!ctx) return femhost_set_error(-2, "%s", femgpu_last_error());
  const char **seqs = (const char **)malloc(ref->n_records * sizeof(char *));
  uint32_t *lens = (uint32_t *)malloc(ref->n_records * sizeof(uint32_t));
  for (uint32_t i = 0; i < ref->n_records; ++i) {
    seqs[i] = ref->bases + ref->base_off[i];
    lens[i] = femhost_record_length(ref, i);
  }
  double t0 = femhost_now();
  int rc = femgpu_upload_reference(ctx, ref->n_records, seqs, lens);
  if (seconds) seconds[0] = femhost_now() - t0;
  t0 = femhost_now();
  if (!rc) rc = femgpu_build_index(ctx, k, step);
  if (seconds) seconds[1] = femhost_now() - t0;
  t0 = femhost_now();
  if (!rc) rc = femgpu_save_index_file(ctx, out_path);
  if (seconds) seconds[2] = femhost_now() - t0;
  if (rc) femhost_set_error(rc, "%s", femgpu_last_error());
  free(seqs);
  free(lens);
  femgpu_destroy(ctx);
  return rc;
}

void femhost_mapper_destroy(femhost_mapper *m) {
  if (!m) return;
  pool_free(m);
  for (int g = 0; g < m->n_gpus; ++g) femgpu_destroy(m->ctx[g]);
  femhost_records_free(&m->ref);
  free(m);
}
const femhost_records *femhost_mapper_reference(const femhost_mapper *m) { return &m->ref; }
femgpu_ctx *femhost_mapper_context(femhost_mapper *m, int i) { return i >= 0 && i < m->n_gpus ? m->ctx[i] : NULL; }

/* ---- one batch travelling through the pipeline -------------------------------------------------- */
#define CHUNK_READS 2048
typedef struct batch batch;
typedef struct {
  batch *b;
  uint32_t chunk;
  int write;                /* 0: build the chunk's SAM text; 1: pwrite it at file_off */
  uint64_t file_off;
  const char *ext;          /* the chunk's text as the device formatted it (NULL: b->text[chunk]) */
  size_t ext_len;
} work_item;
struct batch {
  uint64_t seq;             /* position in file order */
  femhost_records reads;    /* pinned: the device copies straight out of it */
  uint64_t *first;          /* copy of the device result (the slot is reused at once) */
  femgpu_mapping *maps;
  femgpu_alignment *alns;   /* NULL-able: device start/CIGAR per mapping */
  int has_alns;
  char *sam_text;           /* copy of the device's SAM text (k_sam), NULL-able */
  uint64_t *sam_off;        /* n_reads + 1 */
  uint8_t *sam_host;        /* n_reads: 1 = this read is formatted here */
  int has_sam;
  double t_trace[6];        /* FEM_TRACE: parsed, gpu start, gpu end, text ready, offsets assigned, written */
  uint64_t sam_dev_reads;   /* mapped reads of the batch the device formatted */
  size_t cap_sam_text, cap_sam_reads;
  size_t cap_first, cap_maps, cap_alns;
  femgpu_stats stats;
  uint32_t n_launches;
  uint32_t n_chunks;
  femhost_str *text;        /* one SAM text buffer per chunk */
  work_item *items;
  size_t cap_chunks;
  int chunks_left;          /* guarded by run->mu */
  int writes_left;          /* guarded by run->mu */
  char *map_base;           /* the batch's span of the output file, mapped (NULL: pwrite) */
  size_t map_len;
  uint64_t map_file_off;    /* file offset of map_base[0] */
};

static void pool_free(femhost_mapper *m) {
  for (size_t i = 0; i < m->n_pool; ++i) {
    batch *b = &m->pool[i];
    femhost_records_free(&b->reads);
    free(b->first);
    free(b->maps);
    free(b->alns);
    femgpu_free_pinned(b->sam_text); femgpu_free_pinned(b->sam_off); femgpu_free_pinned(b->sam_host);
    for (size_t c = 0; c < b->cap_chunks; ++c) femhost_str_free(&b->text[c]);
    free(b->text);
    free(b->items);
  }
  free(m->pool);
  m->pool = NULL;
  m->n_pool = 0;
}

typedef struct {
  femhost_mapper *m;
  const femhost_map_options *opt;
  femhost_reader *rd;
  int out_fd;               /* -1: build the text, write nothing */
  int out_seekable;         /* regular file: chunks are written in parallel with pwrite */
  uint64_t out_off;         /* next file offset (writer thread only) */
  int trace;                /* FEM_TRACE=1: one line per batch with its times on stderr */
  double t_start, t_enter;
  int device_sam;           /* SAM text formatted by k_sam (FEM_DEVICE_SAM=0: on the host) */
  int out_populate;         /* madvise(MADV_POPULATE_WRITE) before the copy (FEM_SAM_POPULATE=0 turns it off) */
  int out_mmap;             /* tmpfs: chunks are copied into a shared mapping (pwrite holds the inode lock) */
  uint64_t out_alloc;       /* size the file has been extended to (writer thread only) */
  /* page-cache pages of the output allocated ahead of the copies (prealloc_main) */
  int prealloc, prealloc_stop;
  uint64_t prealloc_goal;   /* guarded by mu: allocate up to this file offset */
  pthread_cond_t prealloc_cv;
  int writes_outstanding;   /* guarded by mu; writer waits for 0 before the file is closed */
  pthread_cond_t writes_done;
  fifo q_free, q_parsed, q_work, q_write;
  /* finished batches wait here for the writer, indexed by seq % n_batches */
  batch **done;
  size_t n_batches;
  pthread_mutex_t mu;
  pthread_cond_t batch_done;
  uint64_t n_parsed;        /* batches handed to the GPUs; valid once parsing_over */
  int parsing_over;
  /* parallel reader (plain four-line FASTQ): the file is cut into byte ranges, batch = range */
  int fd;
  uint64_t file_size, chunk_bytes, n_chunks, next_chunk;
  uint64_t range_lo, range_hi; /* this run's byte range of the file (the whole file unless sharded) */
  int readers_left;
  int restart_sequential;   /* some range did not parse as four-line FASTQ: run again with the kseq reader */
  int failed;
  char error[512];
  /* accounting */
  femgpu_stats stats;
  femhost_map_timing tm;
} run_state;

static void run_fail(run_state *r, const char *msg) {
  pthread_mutex_lock(&r->mu);
  if (!r->failed) {
    r->failed = 1;
    snprintf(r->error, sizeof r->error, "%s", msg);
  }
  pthread_mutex_unlock(&r->mu);
}

static void *reader_main(void *arg) {
  run_state *r = (run_state *)arg;
  const uint32_t per_batch = r->opt->batch_reads ? r->opt->batch_reads : DEFAULT_BATCH_READS;
  uint64_t to_skip = r->opt->skip_reads, left = r->opt->no_reads ? 0 : r->opt->max_reads ? r->opt->max_reads : UINT64_MAX, seq = 0;
  double busy = 0;
  batch *b = NULL;
  while (left && !r->failed) {
    if (!b) b = (batch *)fifo_pop(&r->q_free);
    const double t0 = femhost_now();
    uint32_t want = per_batch;
    if (to_skip) want = (uint32_t)(to_skip < per_batch ? to_skip : per_batch);
    else if (left < want) want = (uint32_t)left;
    const int64_t n = femhost_reader_next(r->rd, want, &b->reads);
    busy += femhost_now() - t0;
    if (n < 0) { run_fail(r, femhost_last_error()); break; }
    if (n == 0) break;
    if (to_skip) { to_skip -= (uint64_t)n; continue; } /* another shard's reads: parsed and dropped */
    left -= (uint64_t)n;
    b->seq = seq++;
    b->t_trace[0] = femhost_now();
    fifo_push(&r->q_parsed, b);
    b = NULL;
  }
  if (b) fifo_push(&r->q_free, b);
  pthread_mutex_lock(&r->mu);
  r->n_parsed = seq;
  r->parsing_over = 1;
  r->tm.parse = busy;
  pthread_cond_broadcast(&r->batch_done);
  pthread_mutex_unlock(&r->mu);
  fifo_close(&r->q_parsed);
  return NULL;
}

/* ---- parallel reader --------------------------------------------------------------------------- */
/* Range i covers the records that start in [sync(i*chunk), sync((i+1)*chunk)), where sync(a) is the
 * first record start after the first line break at or after a (femhost_fastq4_sync).  Every range
 * must parse exactly, record by record, with the four-line fast path; since range i ends where
 * range i+1 starts, the ranges together are the sequential parse of the file.  Anything else
 * (multi-line records, FASTA, CRLF, junk) aborts the run, which map_file repeats with the
 * sequential kseq-compatible reader. */
/* byte where range i starts.  The first four ranges are 1/8, 1/8, 1/4 and 1/2 of the others: the readers
 * start together, so the first SAM text is on its way to the file after an eighth of a batch's parse and
 * device time instead of a whole one (the run is bound by the copies into the page cache, which can
 * only start then) */
static uint64_t par_range_start(const run_state *r, uint64_t i) {
  static const uint64_t eighths[4] = {0, 1, 2, 4};
  return r->range_lo + (i < 4 ? r->chunk_bytes * eighths[i] / 8 : (i - 3) * r->chunk_bytes);
}

static int64_t par_sync(run_state *r, uint64_t a, unsigned char *win, size_t win_cap) {
  if (a == 0) return 0;
  if (a >= r->file_size) return (int64_t)r->file_size;
  size_t n = win_cap;
  if (a + n > r->file_size) n = (size_t)(r->file_size - a);
  if (pread(r->fd, win, n, (off_t)a) != (ssize_t)n) return -1;
  const long o = femhost_fastq4_sync(win, n, 0, a + n == r->file_size);
  return o < 0 ? -1 : (int64_t)(a + (uint64_t)o);
}

typedef struct {
  run_state *r;
  double busy;
} par_arg;

static void *par_reader_main(void *arg) {
  par_arg *pa = (par_arg *)arg;
  run_state *r = pa->r;
  const size_t win_cap = 1 << 18;
  unsigned char *win = (unsigned char *)malloc(win_cap), *raw = NULL;
  size_t raw_cap = 0;
  for (;;) {
    /* a batch object first, a range number second: the lowest range in flight always owns one */
    batch *b = (batch *)fifo_pop(&r->q_free);
    pthread_mutex_lock(&r->mu);
    const uint64_t i = r->next_chunk < r->n_chunks && !r->failed && !r->restart_sequential ? r->next_chunk++ : UINT64_MAX;
    pthread_mutex_unlock(&r->mu);
    if (i == UINT64_MAX) { fifo_push(&r->q_free, b); break; }
    const double t0 = femhost_now();
    const int64_t lo = par_sync(r, par_range_start(r, i), win, win_cap);
    const int64_t hi = par_sync(r, i + 1 == r->n_chunks ? r->range_hi : par_range_start(r, i + 1), win, win_cap);
    int rc = lo < 0 || hi < lo ? 1 : 0;
    if (!rc) {
      const size_t n = (size_t)(hi - lo);
      if (n > raw_cap) { raw_cap = n + n / 8 + 4096; free(raw); raw = (unsigned char *)malloc(raw_cap); }
      if (n && pread(r->fd, raw, n, (off_t)lo) != (ssize_t)n) rc = 1;
      else rc = femhost_parse_fastq4(raw, raw + n, &b->reads);
    }
    pa->busy += femhost_now() - t0;
    b->t_trace[0] = femhost_now();
    b->seq = i;
    if (rc) {
      pthread_mutex_lock(&r->mu);
      if (rc < 0) { if (!r->failed) { r->failed = 1; snprintf(r->error, sizeof r->error, "%s", femhost_last_error()); } }
      else r->restart_sequential = 1;
      pthread_mutex_unlock(&r->mu);
      b->reads.n_records = 0;
    }
    fifo_push(&r->q_parsed, b);
  }
  free(win);
  free(raw);
  pthread_mutex_lock(&r->mu);
  if (--r->readers_left == 0) {
    r->n_parsed = r->next_chunk;
    r->parsing_over = 1;
    pthread_cond_broadcast(&r->batch_done);
    pthread_mutex_unlock(&r->mu);
    fifo_close(&r->q_parsed);
  } else pthread_mutex_unlock(&r->mu);
  return NULL;
}

typedef struct {
  run_state *r;
  int gpu, slot;
  double busy;
} gpu_arg;

/* pinned room of a batch for the device's SAM text (text_bytes) or offsets + flags (n_off); contents are not kept */
static void batch_sam_room(batch *b, size_t text_bytes, size_t n_off) {
  if (text_bytes) {
    femgpu_free_pinned(b->sam_text);
    b->sam_text = (char *)femgpu_alloc_pinned(text_bytes);
    b->cap_sam_text = b->sam_text ? text_bytes : 0;
  }
  if (n_off) {
    femgpu_free_pinned(b->sam_off);
    femgpu_free_pinned(b->sam_host);
    b->sam_off = (uint64_t *)femgpu_alloc_pinned(n_off * 8);
    b->sam_host = (uint8_t *)femgpu_alloc_pinned(n_off);
    b->cap_sam_reads = b->sam_off && b->sam_host ? n_off : 0;
  }
}

/* grow-only array of a batch; 0 on success (the old block stays valid on failure) */
static int batch_grow(void **p, size_t bytes) {
  void *q = realloc(*p, bytes ? bytes : 1);
  if (!q) return -1;
  *p = q;
  return 0;
}

/* marks a batch complete for the writer */
static void batch_finished(run_state *r, batch *b) {
  b->t_trace[3] = femhost_now();
  pthread_mutex_lock(&r->mu);
  r->done[b->seq % r->n_batches] = b;
  pthread_cond_broadcast(&r->batch_done);
  pthread_mutex_unlock(&r->mu);
}

static void *gpu_main(void *arg) {
  gpu_arg *ga = (gpu_arg *)arg;
  run_state *r = ga->r;
  femgpu_ctx *ctx = r->m->ctx[ga->gpu];
  batch *b;
  while ((b = (batch *)fifo_pop(&r->q_parsed)) != NULL) {
    const uint32_t n = b->reads.n_records;
    memset(&b->stats, 0, sizeof b->stats);
    b->n_chunks = 0;
    int ok = !r->failed && !r->restart_sequential;
    if (ok) {
      const double t0 = femhost_now();
      b->t_trace[1] = t0;
      femgpu_batch gb = {b->reads.bases, b->reads.base_off, n};
      femgpu_result res;
      int rc = 0;
      if (r->device_sam && n && b->reads.quals && b->reads.has_qual) {
        /* the text comes back into pinned buffers of the batch: one record per read is the first guess,
         * a batch that needs more lands in the library's buffer once and the room grows */
        const uint64_t guess = 2 * b->reads.n_bases + b->reads.name_off[n] + 160ull * n;
        if (guess > b->cap_sam_text) batch_sam_room(b, guess + guess / 4, 0);
        if ((size_t)n + 1 > b->cap_sam_reads) batch_sam_room(b, 0, (size_t)n + 1);
        femgpu_sam_inputs si = {b->reads.names, b->reads.name_off, b->reads.quals, b->reads.has_qual,
                                b->sam_text, b->sam_text ? b->cap_sam_text : 0, b->sam_off, b->sam_host};
        if (!b->sam_off || !b->sam_host) rc = femhost_set_error(-3, "out of pinned memory for the SAM offsets");
        else rc = femgpu_set_sam_inputs(ctx, ga->slot, &si);
      }
      if (!rc) rc = femgpu_submit_batch(ctx, ga->slot, &r->opt->params, &gb);
      if (!rc) rc = femgpu_wait_batch(ctx, ga->slot, &res);
      if (rc) { run_fail(r, femgpu_last_error()); ok = 0; }
      else {
        const uint64_t n_maps = res.first[n];
        int oom = 0;
        if ((size_t)n + 1 > b->cap_first) { if (batch_grow((void **)&b->first, ((size_t)n + 1) * 8)) oom = 1; else b->cap_first = (size_t)n + 1; }
        if (!oom && n_maps > b->cap_maps) {
          const size_t want = n_maps + n_maps / 4 + 1024;
          if (batch_grow((void **)&b->maps, want * sizeof(femgpu_mapping))) oom = 1; else b->cap_maps = want;
        }
        if (!oom && res.alns && n_maps > b->cap_alns) {
          const size_t want = n_maps + n_maps / 4 + 1024;
          if (batch_grow((void **)&b->alns, want * sizeof(femgpu_alignment))) oom = 1; else b->cap_alns = want;
        }
        if (oom) { run_fail(r, "out of memory for a batch's results"); batch_finished(r, b); continue; }
        memcpy(b->first, res.first, ((size_t)n + 1) * 8);
        memcpy(b->maps, res.maps, n_maps * sizeof(femgpu_mapping));
        b->has_alns = res.alns != NULL;
        if (res.alns) memcpy(b->alns, res.alns, n_maps * sizeof(femgpu_alignment));
        b->has_sam = res.sam_off != NULL;
        if (b->has_sam) {
          const uint64_t nb = res.sam_off[n];
          if (res.sam_text != b->sam_text && nb) { /* did not fit: copy out of the slot's buffer, more room next time */
            batch_sam_room(b, nb + nb / 4 + 4096, 0);
            if (!b->sam_text) { run_fail(r, "out of pinned memory for the SAM text"); b->has_sam = 0; }
            else memcpy(b->sam_text, res.sam_text, nb);
          }
        }
        if (b->has_sam) {
          b->sam_dev_reads = 0;
          for (uint32_t i = 0; i < n; ++i) b->sam_dev_reads += b->sam_off[i + 1] > b->sam_off[i];
        }
        b->stats = res.stats;
        float ms[6];
        femgpu_last_timings(ctx, ms, &b->n_launches);
      }
      b->t_trace[2] = femhost_now();
      ga->busy += b->t_trace[2] - t0;
    }
    if (!ok) { batch_finished(r, b); continue; }
    /* hand the batch to the SAM workers, CHUNK_READS reads at a time */
    const uint32_t n_chunks = (n + CHUNK_READS - 1) / CHUNK_READS;
    if (n_chunks > b->cap_chunks) {
      if (batch_grow((void **)&b->text, n_chunks * sizeof(femhost_str)) || batch_grow((void **)&b->items, n_chunks * sizeof(work_item))) {
        run_fail(r, "out of memory for a batch's SAM chunks");
        b->n_chunks = 0;
        batch_finished(r, b);
        continue;
      }
      memset(b->text + b->cap_chunks, 0, (n_chunks - b->cap_chunks) * sizeof(femhost_str));
      b->cap_chunks = n_chunks;
    }
    b->n_chunks = n_chunks;
    /* chunks the device formatted completely are ready as they are; the others go to the workers */
    int to_format = 0;
    for (uint32_t c = 0; c < n_chunks; ++c) {
      work_item *it = &b->items[c];
      const uint32_t lo = c * CHUNK_READS, hi = lo + CHUNK_READS < n ? lo + CHUNK_READS : n;
      it->b = b;
      it->chunk = c;
      it->write = 0;
      it->ext = NULL;
      it->ext_len = 0;
      if (b->has_sam && !memchr(b->sam_host + lo, 1, hi - lo)) {
        it->ext = b->sam_text + b->sam_off[lo];
        it->ext_len = (size_t)(b->sam_off[hi] - b->sam_off[lo]);
        b->text[c].l = 0;
      } else ++to_format;
    }
    b->chunks_left = to_format;
    if (to_format == 0) { batch_finished(r, b); continue; }
    for (uint32_t c = 0; c < n_chunks; ++c)
      if (!b->items[c].ext) fifo_push(&r->q_work, &b->items[c]);
  }
  return NULL;
}

typedef struct {
  run_state *r;
  fifo *q;                  /* q_work: SAM text of chunks; q_write: copies into the output file */
  double busy, busy_write;
  double t_populate, t_copy, t_unmap; /* FEM_TRACE: parts of busy_write */
} worker_arg;

static int write_all(int fd, const char *p, size_t n, int positional, uint64_t off) {
  while (n) {
    const ssize_t w = positional ? pwrite(fd, p, n, (off_t)off) : write(fd, p, n);
    if (w <= 0) return -1;
    p += w; n -= (size_t)w; off += (uint64_t)w;
  }
  return 0;
}

static void *worker_main(void *arg) {
  worker_arg *wa = (worker_arg *)arg;
  run_state *r = wa->r;
  const femhost_records *ref = &r->m->ref;
  const int e = r->opt->params.e;
  static const char no_qual[2] = {' ', 0}; /* prints as '*' (htslib sam.c:3047) */
  work_item *it;
  while ((it = (work_item *)fifo_pop(wa->q)) != NULL) {
    const double t0 = femhost_now();
    batch *b = it->b;
    femhost_str *out = &b->text[it->chunk];
    if (it->write) { /* the writer thread fixed the offset; the copy into the page cache happens here */
      if (b->map_base) {
        char *dst = b->map_base + (it->file_off - b->map_file_off);
#ifndef MADV_POPULATE_WRITE
#define MADV_POPULATE_WRITE 23
#endif
        const size_t wl = it->ext ? it->ext_len : out->l;
        const double tp0 = r->trace ? femhost_now() : 0;
        if (r->out_populate && wl) { /* all pages of the chunk in one call instead of one fault each */
          const uintptr_t pg = (uintptr_t)sysconf(_SC_PAGESIZE), p0 = (uintptr_t)dst & ~(pg - 1);
          madvise((void *)p0, (size_t)((uintptr_t)dst + wl - p0), MADV_POPULATE_WRITE);
        }
        const double tp1 = r->trace ? femhost_now() : 0;
        memcpy(dst, it->ext ? it->ext : out->s, it->ext ? it->ext_len : out->l);
        if (r->trace) { wa->t_populate += tp1 - tp0; wa->t_copy += femhost_now() - tp1; }
      }
      else if (write_all(r->out_fd, it->ext ? it->ext : out->s, it->ext ? it->ext_len : out->l, 1, it->file_off)) run_fail(r, "short write to the SAM file");
      pthread_mutex_lock(&r->mu);
      const int left = --b->writes_left;
      pthread_mutex_unlock(&r->mu);
      if (left == 0 && b->map_base) {
        const double tu0 = r->trace ? femhost_now() : 0;
        munmap(b->map_base, b->map_len);
        b->map_base = NULL;
        if (r->trace) wa->t_unmap += femhost_now() - tu0;
      }
      wa->busy_write += femhost_now() - t0;
      if (left == 0 && r->trace)
        fprintf(stderr, "[trace] batch %" PRIu64 " (%u reads): parsed %.1f, gpu %.1f-%.1f, text %.1f, offsets %.1f, written %.1f ms\n", b->seq,
                b->reads.n_records, 1e3 * (b->t_trace[0] - r->t_start), 1e3 * (b->t_trace[1] - r->t_start), 1e3 * (b->t_trace[2] - r->t_start),
                1e3 * (b->t_trace[3] - r->t_start), 1e3 * (b->t_trace[4] - r->t_start), 1e3 * (femhost_now() - r->t_start));
      pthread_mutex_lock(&r->mu);
      if (--r->writes_outstanding == 0) pthread_cond_broadcast(&r->writes_done);
      pthread_mutex_unlock(&r->mu);
      if (left == 0) fifo_push(&r->q_free, b);
      continue;
    }
    out->l = 0;
    const uint32_t lo = it->chunk * CHUNK_READS;
    const uint32_t hi = lo + CHUNK_READS < b->reads.n_records ? lo + CHUNK_READS : b->reads.n_records;
    for (uint32_t i = lo; i < hi && !r->failed; ++i) {
      const size_t n = (size_t)(b->first[i + 1] - b->first[i]);
      if (!n) continue;
      if (b->has_sam && !b->sam_host[i]) { /* this read's records as the device wrote them */
        const size_t tl = (size_t)(b->sam_off[i + 1] - b->sam_off[i]);
        if (femhost_str_reserve(out, tl)) { run_fail(r, femhost_last_error()); break; }
        memcpy(out->s + out->l, b->sam_text + b->sam_off[i], tl);
        out->l += tl;
        continue;
      }
      const uint64_t off = b->reads.base_off[i];
      if (femhost_sam_read(ref, femhost_record_name(&b->reads, i), b->reads.bases + off,
                           b->reads.has_qual[i] ? b->reads.quals + off : no_qual, (int)femhost_record_length(&b->reads, i), e,
                           b->maps + b->first[i], b->has_alns ? b->alns + b->first[i] : NULL, n, out))
        run_fail(r, femhost_last_error());
    }
    wa->busy += femhost_now() - t0;
    pthread_mutex_lock(&r->mu);
    const int left = --b->chunks_left;
    pthread_mutex_unlock(&r->mu);
    if (left == 0) batch_finished(r, b);
  }
  return NULL;
}

static void *writer_main(void *arg) {
  run_state *r = (run_state *)arg;
  uint64_t next = 0;
  double busy = 0;
  for (;;) {
    pthread_mutex_lock(&r->mu);
    batch *b;
    while ((b = r->done[next % r->n_batches]) == NULL || b->seq != next) {
      if (r->parsing_over && next >= r->n_parsed) { b = NULL; break; }
      pthread_cond_wait(&r->batch_done, &r->mu);
    }
    if (b) r->done[next % r->n_batches] = NULL;
    pthread_mutex_unlock(&r->mu);
    if (!b) break;
    const double t0 = femhost_now();
    int n_writes = 0;
    for (uint32_t c = 0; c < b->n_chunks; ++c) {
      const size_t len = b->items[c].ext ? b->items[c].ext_len : b->text[c].l;
      r->tm.sam_bytes += len;
      if (r->out_fd < 0 || !len || r->failed || r->restart_sequential) continue;
      if (!r->out_seekable) { /* pipe: in order, from this thread */
        if (write_all(r->out_fd, b->items[c].ext ? b->items[c].ext : b->text[c].s, len, 0, 0)) run_fail(r, "short write to the SAM file");
        continue;
      }
      b->items[c].write = 1;
      b->items[c].file_off = r->out_off;
      r->out_off += len;
      ++n_writes;
    }
    b->map_base = NULL;
    if (n_writes && r->prealloc) {
      pthread_mutex_lock(&r->mu);
      if (r->out_off + PREALLOC_AHEAD > r->prealloc_goal) r->prealloc_goal = r->out_off + PREALLOC_AHEAD;
      pthread_cond_signal(&r->prealloc_cv);
      pthread_mutex_unlock(&r->mu);
    }
    if (n_writes && r->out_mmap) { /* extend the file in big steps and map this batch's span of it */
      const uint64_t hi = r->out_off;
      uint64_t first = hi;
      for (uint32_t c = 0; c < b->n_chunks; ++c) if (b->items[c].write && b->items[c].file_off < first) first = b->items[c].file_off;
      if (hi > r->out_alloc) {
        const uint64_t want = hi + (256ull << 20);
        struct statfs fs;
        if (fstatfs(r->out_fd, &fs) == 0 && (uint64_t)fs.f_bavail * (uint64_t)fs.f_bsize > (want - r->out_alloc) + (64ull << 20) &&
            ftruncate(r->out_fd, (off_t)want) == 0) r->out_alloc = want;
      }
      if (hi <= r->out_alloc) {
        const uint64_t page = (uint64_t)sysconf(_SC_PAGESIZE), m0 = first & ~(page - 1);
        void *mp = mmap(NULL, (size_t)(hi - m0), PROT_READ | PROT_WRITE, MAP_SHARED, r->out_fd, (off_t)m0);
        if (mp != MAP_FAILED) { b->map_base = (char *)mp; b->map_len = (size_t)(hi - m0); b->map_file_off = m0; }
      }
    }
    b->t_trace[4] = femhost_now();
    if (n_writes) { /* hand the copies to the workers; the last one recycles the batch */
      pthread_mutex_lock(&r->mu);
      b->writes_left = n_writes;
      r->writes_outstanding += n_writes;
      pthread_mutex_unlock(&r->mu);
      for (uint32_t c = 0; c < b->n_chunks; ++c)
        if (b->items[c].write) fifo_push(&r->q_write, &b->items[c]);
    }
    busy += femhost_now() - t0;
    r->stats.num_reads += b->stats.num_reads;
    r->stats.num_mapped_reads += b->stats.num_mapped_reads;
    r->stats.num_candidates_before_filter += b->stats.num_candidates_before_filter;
    r->stats.num_candidates += b->stats.num_candidates;
    r->stats.num_mappings += b->stats.num_mappings;
    r->stats.num_infeasible_reads += b->stats.num_infeasible_reads;
    r->tm.h2d_bytes += b->reads.n_bases + ((uint64_t)b->reads.n_records + 1) * 8 +
                       (b->has_sam ? b->reads.n_bases + b->reads.name_off[b->reads.n_records] + 9ull * b->reads.n_records : 0);
    r->tm.d2h_bytes += ((uint64_t)b->reads.n_records + 1) * 8 + b->stats.num_mappings * (sizeof(femgpu_mapping) + (b->has_alns ? sizeof(femgpu_alignment) : 0)) + 2ull * b->reads.n_records +
                       (b->has_sam ? b->sam_off[b->reads.n_records] + 9ull * b->reads.n_records : 0);
    r->tm.n_launches += b->n_launches;
    if (b->has_sam) r->tm.device_sam_reads += b->sam_dev_reads;
    ++r->tm.n_batches;
    ++next;
    if (!n_writes) fifo_push(&r->q_free, b);
  }
  pthread_mutex_lock(&r->mu);
  while (r->writes_outstanding) pthread_cond_wait(&r->writes_done, &r->mu);
  pthread_mutex_unlock(&r->mu);
  r->tm.write = busy; /* map_file_run cuts the file to its size once the allocator has stopped */
  return NULL;
}

/* Allocates the output's page-cache pages ahead of the copies.  On tmpfs a first write to a page costs a
 * page allocation under locks shared by every thread that writes the file (measured: 8-9 GB/s from six
 * or sixteen copying threads alike); fallocate allocates them from one thread at 15 GB/s without
 * zeroing, and copies into allocated pages scale with the threads (14 GB/s from six).  The pages past
 * the end of the text go away with the final ftruncate. */
static void *prealloc_main(void *arg) {
  run_state *r = (run_state *)arg;
  uint64_t done = 0;
  pthread_mutex_lock(&r->mu);
  for (;;) {
    while (!r->prealloc_stop && r->prealloc_goal <= done) pthread_cond_wait(&r->prealloc_cv, &r->mu);
    if (r->prealloc_stop) break;
    const uint64_t goal = r->prealloc_goal;
    pthread_mutex_unlock(&r->mu);
    while (done < goal) {
      if (fallocate(r->out_fd, FALLOC_FL_KEEP_SIZE, (off_t)done, (off_t)PREALLOC_STEP) != 0) { done = UINT64_MAX; break; } /* no room or not supported: page faults do it */
      done += PREALLOC_STEP;
      pthread_mutex_lock(&r->mu);
      const int stop = r->prealloc_stop;
      pthread_mutex_unlock(&r->mu);
      if (stop) break;
    }
    pthread_mutex_lock(&r->mu);
  }
  pthread_mutex_unlock(&r->mu);
  return NULL;
}

/* plain (not gzip) regular file that starts with '@': candidate for the parallel reader */
static int parallel_readable(const char *path, uint64_t *size) {
  struct stat st;
  if (stat(path, &st) != 0 || !S_ISREG(st.st_mode) || st.st_size < 4) return 0;
  const int fd = open(path, O_RDONLY);
  if (fd < 0) return 0;
  unsigned char b[2] = {0, 0};
  const ssize_t n = read(fd, b, 2);
  close(fd);
  *size = (uint64_t)st.st_size;
  return n == 2 && b[0] == '@';
}

#define N_PAR_READERS 4 /* at most; -t of them on small hosts (one process per GPU leaves each a few cores) */
static int par_readers(int n_threads) { const char *e = getenv("FEM_READERS"); const int n = e ? atoi(e) : n_threads; return n < 1 ? 1 : n > N_PAR_READERS ? N_PAR_READERS : n; }
/* threads that copy finished text into the output file: the page cache's per-file locks stop scaling
 * beyond about six (measured on tmpfs: 7.5 GB/s from six threads, the same from sixteen at three times the CPU time) */
#define MAX_COPIERS 6
static int max_copiers(void) { const char *e = getenv("FEM_SAM_COPIERS"); const int n = e ? atoi(e) : MAX_COPIERS; return n < 1 ? 1 : n > 64 ? 64 : n; }

static int map_file_run(femhost_mapper *m, const femhost_map_options *opt, const char *reads_path, const char *sam_path,
                        int parallel, femgpu_stats *stats, femhost_map_timing *timing, int *restart) {
  run_state r;
  memset(&r, 0, sizeof r);
  r.m = m;
  r.opt = opt;
  r.fd = -1;
  r.t_enter = femhost_now();
  if (parallel) {
    r.fd = open(reads_path, O_RDONLY);
    if (r.fd < 0) return femhost_set_error(-1, "Cannot find sequence file!");
    struct stat st;
    fstat(r.fd, &st);
    r.file_size = (uint64_t)st.st_size;
    const uint64_t per_batch = opt->batch_reads ? opt->batch_reads : DEFAULT_BATCH_READS;
    r.chunk_bytes = per_batch * 224; /* about per_batch reads of 100 bp; only the batch granularity depends on it */
    if (r.chunk_bytes < 4096) r.chunk_bytes = 4096;
    /* one process per GPU: shard i of n takes the records that start in [sync(size*i/n), sync(size*(i+1)/n)) */
    r.range_lo = 0;
    r.range_hi = r.file_size;
    if (opt->shard_count > 1) {
      r.range_lo = (uint64_t)((__uint128_t)r.file_size * opt->shard_index / opt->shard_count);
      r.range_hi = (uint64_t)((__uint128_t)r.file_size * (opt->shard_index + 1) / opt->shard_count);
    }
    r.n_chunks = 0;
    while (par_range_start(&r, r.n_chunks) < r.range_hi) ++r.n_chunks;
    r.readers_left = par_readers(opt->n_threads);
  } else {
    r.rd = femhost_reader_open(reads_path);
    if (!r.rd) return -1;
  }
  r.device_sam = !(getenv("FEM_DEVICE_SAM") && atoi(getenv("FEM_DEVICE_SAM")) == 0);
  /* few host threads (one process per GPU on a small host): the two GPU threads sleep while they wait instead of
   * spinning on cores the readers and copiers need (four cores, -t 2: 13.7 -> 14.4 M reads/s on cfg2) */
  if (!getenv("FEM_BLOCKING_SYNC"))
    for (int g = 0; g < m->n_gpus; ++g) femgpu_set_option(m->ctx[g], "blocking_sync", opt->n_threads <= 3);

  r.out_fd = -1;
  if (sam_path) {
    r.out_fd = open(sam_path, O_RDWR | O_CREAT | O_TRUNC, 0644);
    if (r.out_fd < 0) {
      if (r.rd) femhost_reader_close(r.rd);
      if (r.fd >= 0) close(r.fd);
      return femhost_set_error(-1, "Cannot open output file %s", sam_path);
    }
    struct stat ost;
    r.out_seekable = fstat(r.out_fd, &ost) == 0 && S_ISREG(ost.st_mode);
    /* FEM_SAM_MMAP=1/0 forces the mapped / the pwrite form; default: mapped on tmpfs, where every
     * pwrite to one file serialises on the inode lock while page faults of a shared mapping do not */
    struct statfs fs;
    const char *mm = getenv("FEM_SAM_MMAP");
    r.out_populate = !(getenv("FEM_SAM_POPULATE") && atoi(getenv("FEM_SAM_POPULATE")) == 0);
    r.out_mmap = r.out_seekable && (mm ? atoi(mm) != 0 : (fstatfs(r.out_fd, &fs) == 0 && fs.f_type == 0x01021994));
  }
  pthread_cond_init(&r.writes_done, NULL);
  const double t_start = femhost_now();
  r.t_start = t_start;
  r.trace = getenv("FEM_TRACE") && atoi(getenv("FEM_TRACE")) != 0;
  const int n_gpu_threads = 2 * m->n_gpus, n_format = opt->n_threads, n_readers = parallel ? par_readers(opt->n_threads) : 1;
  const int n_copiers = opt->n_threads < max_copiers() ? opt->n_threads : max_copiers(), n_workers = n_format + n_copiers;
  r.n_batches = (size_t)(2 * n_gpu_threads + 6 + (parallel ? N_PAR_READERS : 0));
  if (m->n_pool != r.n_batches) {
    pool_free(m);
    m->pool = (batch *)calloc(r.n_batches, sizeof(batch));
    m->n_pool = r.n_batches;
    for (size_t i = 0; i < r.n_batches; ++i) femhost_records_init(&m->pool[i].reads, 1);
  }
  batch *pool = m->pool;
  r.done = (batch **)calloc(r.n_batches, sizeof(batch *));
  fifo_init(&r.q_free, r.n_batches + 1);
  fifo_init(&r.q_parsed, r.n_batches + 1);
  fifo_init(&r.q_work, 1 << 16);
  fifo_init(&r.q_write, 1 << 16);
  pthread_mutex_init(&r.mu, NULL);
  pthread_cond_init(&r.batch_done, NULL);
  for (size_t i = 0; i < r.n_batches; ++i) fifo_push(&r.q_free, &pool[i]);
  if (!opt->no_header) { /* src/output_queue.c:93-116 */
    femhost_str hdr = {0, 0, 0};
    if (femhost_sam_header(&m->ref, &hdr) == 0) {
      r.tm.sam_bytes += hdr.l;
      if (r.out_fd >= 0 && write_all(r.out_fd, hdr.s, hdr.l, 0, 0)) run_fail(&r, "short write to the SAM file");
      r.out_off = hdr.l;
    }
    femhost_str_free(&hdr);
  }
  pthread_t th_reader[N_PAR_READERS], th_writer, th_gpu[2 * FEM_MAX_GPUS], *th_worker = (pthread_t *)calloc((size_t)n_workers, sizeof(pthread_t));
  gpu_arg ga[2 * FEM_MAX_GPUS];
  par_arg pa[N_PAR_READERS];
  worker_arg *wa = (worker_arg *)calloc((size_t)n_workers, sizeof(worker_arg));
  for (int i = 0; i < n_readers; ++i) {
    pa[i].r = &r; pa[i].busy = 0;
    if (parallel) pthread_create(&th_reader[i], NULL, par_reader_main, &pa[i]);
    else pthread_create(&th_reader[i], NULL, reader_main, &r);
  }
  pthread_t th_prealloc;
  pthread_cond_init(&r.prealloc_cv, NULL);
  r.prealloc = r.out_mmap && !(getenv("FEM_SAM_PREALLOC") && atoi(getenv("FEM_SAM_PREALLOC")) == 0);
  if (r.prealloc) {
    struct statfs fs; /* only with room to spare: the pages ahead of the text are real memory until the end */
    if (fstatfs(r.out_fd, &fs) != 0 || (uint64_t)fs.f_bavail * (uint64_t)fs.f_bsize < (1ull << 30)) r.prealloc = 0;
  }
  if (r.prealloc) {
    /* a file of four-line FASTQ maps to about its own size of SAM text when most reads map once: those
     * pages are allocated while the first batches are still being parsed and mapped, before the copies
     * compete for the same locks; beyond that the allocator keeps PREALLOC_AHEAD in front of the text */
    const char *pe = getenv("FEM_SAM_PREALLOC_PERCENT");
    const uint64_t guess = parallel ? r.file_size / 100 * (uint64_t)(pe ? atoi(pe) : 110) : 0;
    r.prealloc_goal = r.out_off + (guess > PREALLOC_AHEAD ? guess : PREALLOC_AHEAD);
    struct statfs fs;
    if (fstatfs(r.out_fd, &fs) == 0 && r.prealloc_goal > (uint64_t)fs.f_bavail * (uint64_t)fs.f_bsize / 4)
      r.prealloc_goal = r.out_off + PREALLOC_AHEAD;
    pthread_create(&th_prealloc, NULL, prealloc_main, &r);
  }
  pthread_create(&th_writer, NULL, writer_main, &r);
  for (int i = 0; i < n_gpu_threads; ++i) {
    ga[i].r = &r; ga[i].gpu = i / 2; ga[i].slot = i % 2; ga[i].busy = 0;
    pthread_create(&th_gpu[i], NULL, gpu_main, &ga[i]);
  }
  for (int i = 0; i < n_workers; ++i) {
    wa[i].r = &r;
    wa[i].q = i < n_format ? &r.q_work : &r.q_write;
    pthread_create(&th_worker[i], NULL, worker_main, &wa[i]);
  }
  for (int i = 0; i < n_readers; ++i) { pthread_join(th_reader[i], NULL); if (parallel) r.tm.parse += pa[i].busy; }
  for (int i = 0; i < n_gpu_threads; ++i) { pthread_join(th_gpu[i], NULL); r.tm.gpu_wait += ga[i].busy; }
  pthread_join(th_writer, NULL); /* returns once every chunk is formatted, ordered and written */
  fifo_close(&r.q_work);
  fifo_close(&r.q_write);
  for (int i = 0; i < n_workers; ++i) { pthread_join(th_worker[i], NULL); r.tm.format += wa[i].busy; r.tm.write += wa[i].busy_write; }
  if (r.prealloc) {
    pthread_mutex_lock(&r.mu);
    r.prealloc_stop = 1;
    pthread_cond_signal(&r.prealloc_cv);
    pthread_mutex_unlock(&r.mu);
    pthread_join(th_prealloc, NULL);
  }
  pthread_cond_destroy(&r.prealloc_cv);
  if ((r.out_alloc || r.prealloc) && ftruncate(r.out_fd, (off_t)r.out_off) != 0) run_fail(&r, "truncating the SAM file to its size failed");
  if (r.out_fd >= 0 && close(r.out_fd) != 0) run_fail(&r, "closing the SAM file failed");
  pthread_cond_destroy(&r.writes_done);
  r.tm.wall = femhost_now() - t_start;
  if (r.trace) {
    double tp = 0, tc = 0, tu = 0;
    for (int i = 0; i < n_workers; ++i) { tp += wa[i].t_populate; tc += wa[i].t_copy; tu += wa[i].t_unmap; }
    fprintf(stderr, "[trace] copiers: populate %.3f s, copy %.3f s, unmap %.3f s of write %.3f s\n", tp, tc, tu, r.tm.write);
  }
  if (r.trace) fprintf(stderr, "[trace] entry to thread start %.1f ms, threads %.1f ms\n", 1e3 * (t_start - r.t_enter), 1e3 * r.tm.wall);
  if (r.rd) femhost_reader_close(r.rd);
  if (r.fd >= 0) close(r.fd);
  free(r.done); free(th_worker); free(wa);
  fifo_destroy(&r.q_free); fifo_destroy(&r.q_parsed); fifo_destroy(&r.q_work); fifo_destroy(&r.q_write);
  pthread_mutex_destroy(&r.mu);
  pthread_cond_destroy(&r.batch_done);
  if (stats) *stats = r.stats;
  if (timing) *timing = r.tm;
  *restart = r.restart_sequential;
  if (r.failed) return femhost_set_error(-4, "%s", r.error);
  return 0;
}

/* ---- src/FEM_map.c:146-219 ------------------------------------------------------------------------ */
/* a path that exists and is not a regular file (FIFO, /dev/stdout, a socket): bytes written to it
 * cannot be taken back, so the run must not start over after it has emitted any */
static int output_is_stream(const char *sam_path) {
  struct stat st;
  return sam_path && stat(sam_path, &st) == 0 && !S_ISREG(st.st_mode);
}

int femhost_map_file(femhost_mapper *m, const femhost_map_options *opt, const char *reads_path, const char *sam_path,
                     femgpu_stats *stats, femhost_map_timing *timing) {
  if (!m || !opt || !reads_path) return femhost_set_error(-1, "map: null argument");
  if (opt->params.e < 0 || opt->params.e > 7) return femhost_set_error(-1, "Wrong error threshold.");         /* src/FEM_map.c:30 */
  if (opt->n_threads <= 0 || opt->n_threads > 1024) return femhost_set_error(-1, "Wrong number of threads."); /* :34 */
  if (opt->params.a < 0 || opt->params.a > 2) return femhost_set_error(-1, "Wrong number of additional q-grams."); /* :38 */
  const int sharded = opt->shard_count > 1;
  if (sharded && opt->shard_index >= opt->shard_count) return femhost_set_error(-1, "map: shard %u of %u", opt->shard_index, opt->shard_count);
  if (sharded && (opt->skip_reads || opt->max_reads || opt->no_reads))
    return femhost_set_error(-1, "map: byte-range shards and read-count shards exclude each other");
  uint64_t size = 0;
  const char *force = getenv("FEM_READER"); /* "kseq": always the sequential reader */
  /* The parallel reader starts over with the sequential one when some range is not four-line FASTQ; that
   * is only possible while nothing irrevocable has been written, i.e. with a regular output file. */
  int parallel = opt->skip_reads == 0 && opt->max_reads == 0 && !opt->no_reads && !(force && strcmp(force, "kseq") == 0) &&
                 !output_is_stream(sam_path) && parallel_readable(reads_path, &size);
  if (sharded && !parallel)
    return femhost_set_error(FEMHOST_ENOTFASTQ4, "map: byte-range shards need a plain four-line FASTQ file and a regular output file");
  int restart = 0;
  const double t0 = femhost_now();
  int rc = map_file_run(m, opt, reads_path, sam_path, parallel, stats, timing, &restart);
  if (rc == 0 && restart && sharded) /* the other shards cannot know: the caller repeats the run with read-count shards */
    return femhost_set_error(FEMHOST_ENOTFASTQ4, "map: %s is not plain four-line FASTQ, byte-range shards do not apply", reads_path);
  if (rc == 0 && restart) { /* not plain four-line FASTQ after all: once more, sequentially */
    rc = map_file_run(m, opt, reads_path, sam_path, 0, stats, timing, &restart);
    if (timing) timing->wall = femhost_now() - t0;
  }
  return rc;
}

/* records of a sequence file as femhost_map_file's readers see them (zero-length records skipped,
 * FASTA, multi-line and gzip input included): what read-count shards are cut from */
int64_t femhost_count_records(const char *path) {
  femhost_reader *rd = femhost_reader_open(path);
  if (!rd) return -1;
  femhost_records tmp;
  femhost_records_init(&tmp, 0);
  int64_t total = 0, n;
  while ((n = femhost_reader_next(rd, 1u << 16, &tmp)) > 0) total += n;
  femhost_records_free(&tmp);
  femhost_reader_close(rd);
  return n < 0 ? n : total;
}
