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
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "femhost.h"
#include "host_internal.h"

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
  if (n_gpus > 1 && femhost_index_load(index_path, &idx)) rc = -1; /* one read of the file for all devices */
  for (int g = 0; g < n_gpus && !rc; ++g) {
    m->ctx[g] = femgpu_create(gpu_ids ? gpu_ids[g] : g);
    if (!m->ctx[g]) { femhost_set_error(-2, "%s", femgpu_last_error()); rc = -2; break; }
    m->n_gpus = g + 1;
    rc = n_gpus > 1 ? femgpu_upload_index(m->ctx[g], idx.k, idx.step, idx.lut, idx.lut_n, idx.occ, idx.n_occ)
                    : femgpu_upload_index_file(m->ctx[g], index_path);
    if (!rc) rc = femgpu_upload_reference(m->ctx[g], m->ref.n_records, seqs, lens);
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
} work_item;
struct batch {
  uint64_t seq;             /* position in file order */
  femhost_records reads;    /* pinned: the device copies straight out of it */
  uint64_t *first;          /* copy of the device result (the slot is reused at once) */
  femgpu_mapping *maps;
  size_t cap_first, cap_maps;
  femgpu_stats stats;
  uint32_t n_launches;
  uint32_t n_chunks;
  femhost_str *text;        /* one SAM text buffer per chunk */
  work_item *items;
  size_t cap_chunks;
  int chunks_left;          /* guarded by run->mu */
};

static void pool_free(femhost_mapper *m) {
  for (size_t i = 0; i < m->n_pool; ++i) {
    batch *b = &m->pool[i];
    femhost_records_free(&b->reads);
    free(b->first);
    free(b->maps);
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
  FILE *out;
  fifo q_free, q_parsed, q_work;
  /* finished batches wait here for the writer, indexed by seq % n_batches */
  batch **done;
  size_t n_batches;
  pthread_mutex_t mu;
  pthread_cond_t batch_done;
  uint64_t n_parsed;        /* batches handed to the GPUs; valid once parsing_over */
  int parsing_over;
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
  const uint32_t per_batch = r->opt->batch_reads ? r->opt->batch_reads : 131072u;
  uint64_t to_skip = r->opt->skip_reads, left = r->opt->max_reads ? r->opt->max_reads : UINT64_MAX, seq = 0;
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

typedef struct {
  run_state *r;
  int gpu, slot;
  double busy;
} gpu_arg;

/* marks a batch complete for the writer */
static void batch_finished(run_state *r, batch *b) {
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
    int ok = !r->failed;
    if (ok) {
      const double t0 = femhost_now();
      femgpu_batch gb = {b->reads.bases, b->reads.base_off, n};
      femgpu_result res;
      int rc = femgpu_submit_batch(ctx, ga->slot, &r->opt->params, &gb);
      if (!rc) rc = femgpu_wait_batch(ctx, ga->slot, &res);
      if (rc) { run_fail(r, femgpu_last_error()); ok = 0; }
      else {
        const uint64_t n_maps = res.first[n];
        if ((size_t)n + 1 > b->cap_first) { b->cap_first = (size_t)n + 1; b->first = (uint64_t *)realloc(b->first, b->cap_first * 8); }
        if (n_maps > b->cap_maps) { b->cap_maps = n_maps + n_maps / 4 + 1024; b->maps = (femgpu_mapping *)realloc(b->maps, b->cap_maps * sizeof(femgpu_mapping)); }
        memcpy(b->first, res.first, ((size_t)n + 1) * 8);
        memcpy(b->maps, res.maps, n_maps * sizeof(femgpu_mapping));
        b->stats = res.stats;
        float ms[5];
        femgpu_last_timings(ctx, ms, &b->n_launches);
      }
      ga->busy += femhost_now() - t0;
    }
    if (!ok) { batch_finished(r, b); continue; }
    /* hand the batch to the SAM workers, CHUNK_READS reads at a time */
    const uint32_t n_chunks = (n + CHUNK_READS - 1) / CHUNK_READS;
    if (n_chunks > b->cap_chunks) {
      b->text = (femhost_str *)realloc(b->text, n_chunks * sizeof(femhost_str));
      memset(b->text + b->cap_chunks, 0, (n_chunks - b->cap_chunks) * sizeof(femhost_str));
      b->items = (work_item *)realloc(b->items, n_chunks * sizeof(work_item));
      b->cap_chunks = n_chunks;
    }
    b->n_chunks = n_chunks;
    b->chunks_left = (int)n_chunks;
    if (n_chunks == 0) { batch_finished(r, b); continue; }
    for (uint32_t c = 0; c < n_chunks; ++c) {
      b->items[c].b = b;
      b->items[c].chunk = c;
      fifo_push(&r->q_work, &b->items[c]);
    }
  }
  return NULL;
}

typedef struct {
  run_state *r;
  double busy;
} worker_arg;

static void *worker_main(void *arg) {
  worker_arg *wa = (worker_arg *)arg;
  run_state *r = wa->r;
  const femhost_records *ref = &r->m->ref;
  const int e = r->opt->params.e;
  static const char no_qual[2] = {' ', 0}; /* prints as '*' (htslib sam.c:3047) */
  work_item *it;
  while ((it = (work_item *)fifo_pop(&r->q_work)) != NULL) {
    const double t0 = femhost_now();
    batch *b = it->b;
    femhost_str *out = &b->text[it->chunk];
    out->l = 0;
    const uint32_t lo = it->chunk * CHUNK_READS;
    const uint32_t hi = lo + CHUNK_READS < b->reads.n_records ? lo + CHUNK_READS : b->reads.n_records;
    for (uint32_t i = lo; i < hi && !r->failed; ++i) {
      const size_t n = (size_t)(b->first[i + 1] - b->first[i]);
      if (!n) continue;
      const uint64_t off = b->reads.base_off[i];
      if (femhost_sam_read(ref, femhost_record_name(&b->reads, i), b->reads.bases + off,
                           b->reads.has_qual[i] ? b->reads.quals + off : no_qual, (int)femhost_record_length(&b->reads, i), e,
                           b->maps + b->first[i], n, out))
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
    for (uint32_t c = 0; c < b->n_chunks; ++c) {
      r->tm.sam_bytes += b->text[c].l;
      if (r->out && b->text[c].l && fwrite(b->text[c].s, 1, b->text[c].l, r->out) != b->text[c].l) run_fail(r, "short write to the SAM file");
    }
    busy += femhost_now() - t0;
    r->stats.num_reads += b->stats.num_reads;
    r->stats.num_mapped_reads += b->stats.num_mapped_reads;
    r->stats.num_candidates_before_filter += b->stats.num_candidates_before_filter;
    r->stats.num_candidates += b->stats.num_candidates;
    r->stats.num_mappings += b->stats.num_mappings;
    r->stats.num_infeasible_reads += b->stats.num_infeasible_reads;
    r->tm.h2d_bytes += b->reads.n_bases + ((uint64_t)b->reads.n_records + 1) * 8;
    r->tm.d2h_bytes += ((uint64_t)b->reads.n_records + 1) * 8 + b->stats.num_mappings * sizeof(femgpu_mapping) + 2ull * b->reads.n_records;
    r->tm.n_launches += b->n_launches;
    ++r->tm.n_batches;
    ++next;
    fifo_push(&r->q_free, b);
  }
  r->tm.write = busy;
  return NULL;
}

/* ---- src/FEM_map.c:146-219 ------------------------------------------------------------------------ */
int femhost_map_file(femhost_mapper *m, const femhost_map_options *opt, const char *reads_path, const char *sam_path,
                     femgpu_stats *stats, femhost_map_timing *timing) {
  if (!m || !opt || !reads_path) return femhost_set_error(-1, "map: null argument");
  if (opt->params.e < 0 || opt->params.e > 7) return femhost_set_error(-1, "Wrong error threshold.");         /* src/FEM_map.c:30 */
  if (opt->n_threads <= 0 || opt->n_threads > 1024) return femhost_set_error(-1, "Wrong number of threads."); /* :34 */
  if (opt->params.a < 0 || opt->params.a > 2) return femhost_set_error(-1, "Wrong number of additional q-grams."); /* :38 */
  run_state r;
  memset(&r, 0, sizeof r);
  r.m = m;
  r.opt = opt;
  r.rd = femhost_reader_open(reads_path);
  if (!r.rd) return -1;
  if (sam_path) {
    r.out = fopen(sam_path, "wb");
    if (!r.out) { femhost_reader_close(r.rd); return femhost_set_error(-1, "Cannot open output file %s", sam_path); }
    setvbuf(r.out, NULL, _IOFBF, 8 << 20);
  }
  const double t_start = femhost_now();
  const int n_gpu_threads = 2 * m->n_gpus, n_workers = opt->n_threads;
  r.n_batches = (size_t)(2 * n_gpu_threads + 2);
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
  pthread_mutex_init(&r.mu, NULL);
  pthread_cond_init(&r.batch_done, NULL);
  for (size_t i = 0; i < r.n_batches; ++i) fifo_push(&r.q_free, &pool[i]);
  if (!opt->no_header) { /* src/output_queue.c:93-116 */
    femhost_str hdr = {0, 0, 0};
    if (femhost_sam_header(&m->ref, &hdr) == 0) {
      r.tm.sam_bytes += hdr.l;
      if (r.out) fwrite(hdr.s, 1, hdr.l, r.out);
    }
    femhost_str_free(&hdr);
  }
  pthread_t th_reader, th_writer, th_gpu[2 * FEM_MAX_GPUS], *th_worker = (pthread_t *)calloc((size_t)n_workers, sizeof(pthread_t));
  gpu_arg ga[2 * FEM_MAX_GPUS];
  worker_arg *wa = (worker_arg *)calloc((size_t)n_workers, sizeof(worker_arg));
  pthread_create(&th_reader, NULL, reader_main, &r);
  pthread_create(&th_writer, NULL, writer_main, &r);
  for (int i = 0; i < n_gpu_threads; ++i) {
    ga[i].r = &r; ga[i].gpu = i / 2; ga[i].slot = i % 2; ga[i].busy = 0;
    pthread_create(&th_gpu[i], NULL, gpu_main, &ga[i]);
  }
  for (int i = 0; i < n_workers; ++i) {
    wa[i].r = &r;
    pthread_create(&th_worker[i], NULL, worker_main, &wa[i]);
  }
  pthread_join(th_reader, NULL);
  for (int i = 0; i < n_gpu_threads; ++i) { pthread_join(th_gpu[i], NULL); r.tm.gpu_wait += ga[i].busy; }
  fifo_close(&r.q_work);
  for (int i = 0; i < n_workers; ++i) { pthread_join(th_worker[i], NULL); r.tm.format += wa[i].busy; }
  pthread_join(th_writer, NULL);
  if (r.out && fclose(r.out) != 0) run_fail(&r, "closing the SAM file failed");
  r.tm.wall = femhost_now() - t_start;
  femhost_reader_close(r.rd);
  free(r.done); free(th_worker); free(wa);
  fifo_destroy(&r.q_free); fifo_destroy(&r.q_parsed); fifo_destroy(&r.q_work);
  pthread_mutex_destroy(&r.mu);
  pthread_cond_destroy(&r.batch_done);
  if (stats) *stats = r.stats;
  if (timing) *timing = r.tm;
  if (r.failed) return femhost_set_error(-4, "%s", r.error);
  return 0;
}
