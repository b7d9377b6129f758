/* map_b200.c -- INTEGRATION.md section B made executable (TEST INFRASTRUCTURE, built by oracle/Makefile into
 * oracle/_ref/FEM_ref_b200, never part of the product).
 *
 * A drop-in replacement for the reference's src/map.c: the reference keeps its own main, argument parsing,
 * InputQueue, OutputQueue, process_mappings and htslib writer; only the body of the mapper thread changes.
 * Where the reference calls generate_group_seeding_candidates (src/filter.h:9) and verify_candidates
 * (src/align.h:13) per read and strand (src/map.c:27-49), this thread flattens the batch the input queue handed
 * over and makes ONE femgpu_map_batch call (include/femgpu.h).  The mappings come back in the reference's
 * emission order, so process_mappings (src/align.h:14) and push_output_queue run unchanged.
 *
 * The device context is created by the first mapper thread from the Index and reference SequenceBatch that
 * map_main loaded (src/FEM_map.c:136-143) -- this file is all a maintainer has to swap.
 */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "map.h"     /* the reference's own header (REF/src) */
#include "femgpu.h"  /* include/femgpu.h */

static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_mutex_t g_slot_mu[2] = {PTHREAD_MUTEX_INITIALIZER, PTHREAD_MUTEX_INITIALIZER};
static femgpu_ctx *g_ctx = NULL;

static void die(const char *what) {
  fprintf(stderr, "%s: %s\n", what, femgpu_last_error());
  exit(EXIT_FAILURE);
}

/* load_index + the reference batch, as they stand in host memory, go to HBM once */
static femgpu_ctx *shared_context(const MappingArgs *a) {
  pthread_mutex_lock(&g_mu);
  if (!g_ctx) {
    femgpu_ctx *c = femgpu_create(0);
    if (!c) die("femgpu_create");
    const Index *ix = a->index; /* src/index.h:7-14 */
    if (femgpu_upload_index(c, ix->kmer_size, ix->step_size, ix->lookup_table, ((size_t)1 << (2 * ix->kmer_size)) + 1,
                            ix->occurrence_table, ix->occurrence_table_size))
      die("femgpu_upload_index");
    const SequenceBatch *ref = a->reference_sequence_batch;
    const uint32_t n = ref->num_loaded_sequences;
    const char **seq = (const char **)malloc(n * sizeof *seq);
    uint32_t *len = (uint32_t *)malloc(n * sizeof *len);
    for (uint32_t i = 0; i < n; ++i) { /* src/sequence_batch.h:45-55 */
      seq[i] = get_sequence_from_sequence_batch_at(ref, i);
      len[i] = get_sequence_length_from_sequence_batch_at(ref, i);
    }
    if (femgpu_upload_reference(c, n, seq, len)) die("femgpu_upload_reference");
    free(seq);
    free(len);
    g_ctx = c;
  }
  pthread_mutex_unlock(&g_mu);
  return g_ctx;
}

void *single_end_read_mapping_thread(void *mapping_args_v) {
  MappingArgs *mapping_args = (MappingArgs *)mapping_args_v;
  femgpu_ctx *gpu = shared_context(mapping_args);
  const int slot = mapping_args->thread_id & 1; /* two slots = two streams per context */
  SequenceBatch read_batch;
  initialize_sequence_batch_with_max_size(mapping_args->max_read_batch_size, &read_batch);
  kvec_t_Mapping mappings;
  kv_init(mappings.v);
  kvec_t_bam1_t_ptr sam_alignment_kvec;
  kv_init(sam_alignment_kvec.v);
  char *bases = NULL;
  uint64_t *off = NULL;
  size_t cap_bases = 0, cap_reads = 0;
  const FEMArgs *fa = mapping_args->fem_args;
  const femgpu_params p = {fa->kmer_size, fa->step_size, fa->error_threshold, fa->num_additional_qgrams};
  while (1) {
    pop_input_queue(&read_batch, mapping_args->input_queue); /* src/map.c:19 */
    const uint32_t n = read_batch.num_loaded_sequences;
    if (n == 0) break;
    mapping_args->mapping_stats.num_reads += n;
    /* flatten the SequenceBatch */
    if ((size_t)n + 1 > cap_reads) { cap_reads = (size_t)n + 1; off = (uint64_t *)realloc(off, cap_reads * sizeof *off); }
    off[0] = 0;
    for (uint32_t i = 0; i < n; ++i) off[i + 1] = off[i] + get_sequence_length_from_sequence_batch_at(&read_batch, i);
    if (off[n] > cap_bases) { cap_bases = off[n] + off[n] / 4 + 64; bases = (char *)realloc(bases, cap_bases); }
    for (uint32_t i = 0; i < n; ++i) memcpy(bases + off[i], get_sequence_from_sequence_batch_at(&read_batch, i), off[i + 1] - off[i]);
    const femgpu_batch b = {bases, off, n};
    femgpu_result r;
    pthread_mutex_lock(&g_slot_mu[slot]);
    if (femgpu_submit_batch(gpu, slot, &p, &b) || femgpu_wait_batch(gpu, slot, &r)) die("femgpu_map_batch");
    /* what src/map.c:50-55 did per read; r stays valid until this slot is submitted again */
    for (uint32_t i = 0; i < n; ++i) {
      const uint32_t m = (uint32_t)(r.first[i + 1] - r.first[i]);
      if (!m) continue;
      kv_clear(mappings.v);
      int any_reverse = 0;
      for (uint32_t j = 0; j < m; ++j) { /* femgpu_mapping == Mapping, src/utils.h:44-49 */
        const femgpu_mapping *g = &r.maps[r.first[i] + j];
        Mapping mp;
        memset(&mp, 0, sizeof mp);
        mp.direction = g->direction;
        mp.edit_distance = g->edit_distance;
        mp.candidate_position = g->candidate_position;
        mp.end_position_offset = (int16_t)g->end_position_offset;
        any_reverse |= g->direction;
        kv_push(Mapping, mappings.v, mp);
      }
      if (any_reverse) prepare_negative_sequence_at(i, &read_batch); /* src/map.c:40: process_mappings reads it */
      process_mappings(fa, &read_batch, i, mapping_args->reference_sequence_batch, mappings.v.a, m, &sam_alignment_kvec);
      push_output_queue(&sam_alignment_kvec, mapping_args->output_queue);
    }
    /* the five counters of src/FEM_map.c:214-218 (num_reads above) */
    mapping_args->mapping_stats.num_mapped_reads += r.stats.num_mapped_reads;
    mapping_args->mapping_stats.num_candidates_without_additonal_qgram_filter += r.stats.num_candidates_before_filter;
    mapping_args->mapping_stats.num_candidates += r.stats.num_candidates;
    mapping_args->mapping_stats.num_mappings += r.stats.num_mappings;
    pthread_mutex_unlock(&g_slot_mu[slot]);
  }
  pthread_mutex_lock(&(mapping_args->output_queue->queue_mutex)); /* src/map.c:59-62 */
  ++(mapping_args->output_queue->num_finished_mapping_threads);
  pthread_cond_signal(&(mapping_args->output_queue->pro_cond));
  pthread_mutex_unlock(&(mapping_args->output_queue->queue_mutex));
  destory_sequence_batch(&read_batch);
  kv_destroy(sam_alignment_kvec.v);
  kv_destroy(mappings.v);
  free(bases);
  free(off);
  return NULL;
}
