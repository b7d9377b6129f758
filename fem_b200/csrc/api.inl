// api.inl -- batch-level entry points (included by femgpu.cu)

static int scan_batch(const femgpu_batch *b, uint64_t *n_bases, int *l_max, uint32_t *n_too_long) {
  if (!b || (b->n_reads && (!b->bases || !b->offsets))) return set_error(FEMGPU_EINVAL, "null batch");
  uint64_t lm = 0;
  *n_too_long = 0;
  for (uint32_t i = 0; i < b->n_reads; ++i) {
    if (b->offsets[i + 1] < b->offsets[i]) return set_error(FEMGPU_EINVAL, "read offsets must be non-decreasing");
    const uint64_t l = b->offsets[i + 1] - b->offsets[i];
    // a read beyond the seed kernel's shared-memory plan is left unmapped and counted as infeasible (k_seed marks
    // it); the reference maps reads of any length, so the command line reports them like the too-short ones
    if (l > (uint64_t)FEM_MAX_READ_LEN) ++*n_too_long;
    else lm = std::max(lm, l);
  }
  if (b->n_reads && b->offsets[0] != 0) return set_error(FEMGPU_EINVAL, "offsets[0] must be 0");
  *n_bases = b->n_reads ? b->offsets[b->n_reads] : 0;
  if (b->n_reads > 0x7FFFFFF0u / 2) return set_error(FEMGPU_EINVAL, "too many reads in one batch");
  *l_max = (int)lm;
  return 0;
}

static int stage_batch(femgpu_ctx *c, Slot &s, const femgpu_params *p, const femgpu_batch *b, Plan *pl) {
  if (!c || !p) return set_error(FEMGPU_EINVAL, "null argument");
  CU(cudaSetDevice(c->device));
  uint64_t n_bases = 0;
  int l_max = 0;
  uint32_t n_too_long = 0;
  TRY(scan_batch(b, &n_bases, &l_max, &n_too_long));
  TRY(make_plan(c, p, l_max, pl, n_too_long != 0));
  s.n_too_long = n_too_long;
  if (pl->seed_smem > c->smem_optin || pl->cand_smem > c->smem_optin || pl->hash_smem > c->smem_optin)
    return set_error(FEMGPU_EINVAL, "shared-memory plan too large (seed %zu, cand %zu bytes)", pl->seed_smem, pl->cand_smem);
  TRY(slot_reserve(c, s, *pl, b->n_reads, n_bases));
  s.params = *p;
  s.blocking_sync = c->blocking_sync != 0;
  s.n_reads = b->n_reads;
  s.n_bases = n_bases;
  s.l_max = l_max;
  memset(&s.stats, 0, sizeof s.stats);
  s.stats.num_reads = b->n_reads;
  cudaEventRecord(s.ev[0], s.stream);
  if (b->n_reads) {
    CU(cudaMemcpyAsync(s.bases.p, b->bases, n_bases, cudaMemcpyHostToDevice, s.stream));
    CU(cudaMemcpyAsync(s.read_off.p, b->offsets, (b->n_reads + 1ull) * 8, cudaMemcpyHostToDevice, s.stream));
  }
  // SAM text inputs apply to this submit only
  s.sam_on = s.sam_pending && c->device_traceback && c->have_rnames && b->n_reads > 0;
  s.sam_pending = false;
  if (s.sam_on) {
    const femgpu_sam_inputs &in = s.sam_in;
    const uint64_t name_bytes = in.name_off[b->n_reads];
    TRY(s.names.reserve(name_bytes + 16));
    TRY(s.name_off.reserve((b->n_reads + 1ull) * 8));
    TRY(s.quals.reserve(n_bases + 16));
    TRY(s.has_qual.reserve(b->n_reads + 16ull));
    CU(cudaMemcpyAsync(s.names.p, in.names, name_bytes, cudaMemcpyHostToDevice, s.stream));
    CU(cudaMemcpyAsync(s.name_off.p, in.name_off, (b->n_reads + 1ull) * 8, cudaMemcpyHostToDevice, s.stream));
    CU(cudaMemcpyAsync(s.quals.p, in.quals, n_bases, cudaMemcpyHostToDevice, s.stream));
    CU(cudaMemcpyAsync(s.has_qual.p, in.has_qual, b->n_reads, cudaMemcpyHostToDevice, s.stream));
  }
  return 0;
}

// k_sam: lengths -> exclusive scan over reads -> text.  After run_verify_and_compact (needs k_trace).
static int run_sam(femgpu_ctx *c, Slot &s) {
  s.sam_total = 0;
  s.sam_flagged = 0;
  if (!s.sam_on) return 0;
  const uint32_t n = s.n_reads;
  Counters *dc = s.counters.as<Counters>();
  TRY(s.sam_bytes.reserve(n * 4ull + 16));
  TRY(s.sam_host.reserve(n + 16ull));
  TRY(s.sam_off.reserve((n + 1ull) * 8));
  const unsigned n_blocks = (unsigned)((n + SCAN_BLOCK - 1) / SCAN_BLOCK);
  TRY(s.sam_blocks.reserve((n_blocks + 2ull) * 8));
  SamArgs a;
  a.bases = s.bases.as<char>();
  a.read_off = s.read_off.as<uint64_t>();
  a.n_reads = n;
  a.quals = s.quals.as<char>();
  a.has_qual = s.has_qual.as<uint8_t>();
  a.names = s.names.as<char>();
  a.name_off = s.name_off.as<uint64_t>();
  a.rnames = c->rnames.as<char>();
  a.rname_off = c->rname_off.as<uint32_t>();
  a.maps = s.maps.as<ulonglong2>();
  a.aln = s.aln.as<uint32_t>();
  a.rs_map_first = s.rs_map_first.as<uint64_t>();
  a.read_bytes = s.sam_bytes.as<uint32_t>();
  a.host_flag = s.sam_host.as<uint8_t>();
  a.text_off = s.sam_off.as<uint64_t>();
  a.n_flagged = &dc->sam_flagged;
  TRY(s.sam_big_list.reserve(n * 4ull + 16));
  TRY(s.sam_big_ord.reserve(s.n_maps * 2ull + 16));
  TRY(s.sam_big_rec_off.reserve(s.n_maps * 4ull + 16));
  a.big_list = s.sam_big_list.as<uint32_t>();
  a.n_big = &dc->sam_big;
  a.big_ord = s.sam_big_ord.as<uint16_t>();
  a.big_rec_off = s.sam_big_rec_off.as<uint32_t>();
  const unsigned big_grid = (unsigned)c->sm_count * 2u;
  const size_t big_smem = 4 * (size_t)SAM_BIG_MAX * 10;
  a.text = nullptr;
  const unsigned grid = (n + 127) / 128;
  k_sam_len<<<grid, 128, 0, s.stream>>>(a);
  CU(cudaFuncSetAttribute(k_sam_big_len, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)big_smem));
  k_sam_big_len<<<big_grid, 128, big_smem, s.stream>>>(a); // reads k_sam_len listed (65..SAM_BIG_MAX mappings); idles without any
  k_scan_reduce<<<n_blocks, SCAN_BLOCK, 0, s.stream>>>(s.sam_bytes.as<uint32_t>(), n, s.sam_blocks.as<uint64_t>());
  k_scan_blocks<<<1, 32, 0, s.stream>>>(s.sam_blocks.as<uint64_t>(), n_blocks, (uint64_t *)&dc->sam_total);
  k_scan_apply<<<n_blocks, SCAN_BLOCK, 0, s.stream>>>(s.sam_bytes.as<uint32_t>(), n, s.sam_blocks.as<uint64_t>(),
                                                     s.sam_off.as<uint64_t>(), (uint64_t *)&dc->sam_total);
  s.n_launches += 5;
  CU(cudaGetLastError());
  Counters hc;
  TRY(read_counters(s, &hc));
  s.sam_total = hc.sam_total;
  const uint64_t n_big = hc.sam_big;
  s.sam_flagged = hc.sam_flagged;
  if (s.sam_total) {
    TRY(s.sam_text.reserve(s.sam_total + 16));
    a.text = s.sam_text.as<char>();
    // per-warp image: 1.25 x the mean span of 32 reads (more resident warps than with the maximum);
    // the spans that do not fit are written byte-wise
    uint64_t image = (s.sam_total / n + 1) * 40 + 256;
    image = std::min<uint64_t>(std::max<uint64_t>((image + 15) & ~15ull, 2048), SAM_IMAGE_MAX);
    CU(cudaFuncSetAttribute(k_sam_write, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(4 * SAM_IMAGE_MAX)));
    k_sam_write<<<grid, 128, 4 * image, s.stream>>>(a, (uint32_t)image);
    ++s.n_launches;
    if (n_big) {
      k_sam_big_write<<<big_grid, 128, 0, s.stream>>>(a);
      ++s.n_launches;
    }
    CU(cudaGetLastError());
  }
  return 0;
}

// per-read `first` array = every other entry of the per read-strand scan
static int fetch_result(Slot &s, femgpu_result *r) {
  const uint32_t n = s.n_reads;
  uint64_t *hf = s.h_first.as<uint64_t>();
  if (n == 0) {
    hf[0] = 0;
  } else {
    CU(cudaMemcpy2DAsync(hf, 8, s.rs_map_first.p, 16, 8, (size_t)n + 1, cudaMemcpyDeviceToHost, s.stream));
    if (s.n_maps) {
      TRY(s.h_maps.reserve(s.n_maps * 16));
      CU(cudaMemcpyAsync(s.h_maps.p, s.maps.p, s.n_maps * 16, cudaMemcpyDeviceToHost, s.stream));
      if (s.has_aln && s.sam_on && s.sam_flagged == 0) s.has_aln = false; // every record is already text
      if (s.has_aln) {
        TRY(s.h_aln.reserve(s.n_maps * sizeof(femgpu_alignment)));
        CU(cudaMemcpyAsync(s.h_aln.p, s.aln.p, s.n_maps * sizeof(femgpu_alignment), cudaMemcpyDeviceToHost, s.stream));
      }
    }
    if (s.sam_on) { // into the caller's buffers where it gave some (femgpu_sam_inputs)
      const femgpu_sam_inputs &in = s.sam_in;
      s.r_sam_off = in.off_out;
      s.r_sam_host = in.host_out;
      s.r_sam_text = in.text_out && s.sam_total <= in.text_cap ? in.text_out : nullptr;
      if (!s.r_sam_off) { TRY(s.h_sam_off.reserve((n + 1ull) * 8)); s.r_sam_off = s.h_sam_off.as<uint64_t>(); }
      if (!s.r_sam_host) { TRY(s.h_sam_host.reserve(n + 16ull)); s.r_sam_host = s.h_sam_host.as<uint8_t>(); }
      if (!s.r_sam_text) { TRY(s.h_sam_text.reserve(s.sam_total + 16)); s.r_sam_text = s.h_sam_text.as<char>(); }
      CU(cudaMemcpyAsync(s.r_sam_off, s.sam_off.p, (n + 1ull) * 8, cudaMemcpyDeviceToHost, s.stream));
      CU(cudaMemcpyAsync(s.r_sam_host, s.sam_host.p, n, cudaMemcpyDeviceToHost, s.stream));
      if (s.sam_total) CU(cudaMemcpyAsync(s.r_sam_text, s.sam_text.p, s.sam_total, cudaMemcpyDeviceToHost, s.stream));
    }
    // infeasible read-strands are rare; count them from the status bytes
    TRY(s.h_status.reserve(2ull * n));
    CU(cudaMemcpyAsync(s.h_status.p, s.rs_status.p, 2ull * n, cudaMemcpyDeviceToHost, s.stream));
  }
  cudaEventRecord(s.ev[5], s.stream);
  CU(slot_sync(s));
  uint64_t mapped = 0, infeasible = 0;
  const uint8_t *st = s.h_status.as<uint8_t>();
  for (uint32_t i = 0; i < n; ++i) {
    mapped += hf[i + 1] > hf[i];
    infeasible += st[2 * i] == RS_INFEASIBLE || st[2 * i + 1] == RS_INFEASIBLE;
  }
  s.stats.num_mapped_reads = mapped;
  s.stats.num_infeasible_reads = infeasible;
  float t;
  cudaEventElapsedTime(&t, s.ev[0], s.ev[1]); s.ms[0] = t;
  cudaEventElapsedTime(&t, s.ev[1], s.ev[2]); s.ms[1] = t;
  cudaEventElapsedTime(&t, s.ev[2], s.ev[3]); s.ms[2] = t;
  cudaEventElapsedTime(&t, s.ev[3], s.ev[4]); s.ms[3] = t;
  cudaEventElapsedTime(&t, s.ev[4], s.ev[6]); s.ms[4] = t;
  cudaEventElapsedTime(&t, s.ev[0], s.ev[5]); s.ms[5] = t;
  if (r) {
    r->first = hf;
    r->maps = s.h_maps.as<femgpu_mapping>();
    r->alns = s.has_aln && s.n_maps ? s.h_aln.as<femgpu_alignment>() : nullptr;
    r->stats = s.stats;
    const bool sam = s.sam_on && n > 0;
    r->sam_text = sam ? s.r_sam_text : nullptr;
    r->sam_off = sam ? s.r_sam_off : nullptr;
    r->sam_host = sam ? s.r_sam_host : nullptr;
  }
  return 0;
}

extern "C" int femgpu_submit_batch(femgpu_ctx *c, int slot, const femgpu_params *p, const femgpu_batch *b) {
  if (!c || slot < 0 || slot > 1) return set_error(FEMGPU_EINVAL, "bad context or slot");
  Slot &s = c->slot[slot];
  if (s.busy) return set_error(FEMGPU_ESTATE, "slot %d still holds an unfinished batch", slot);
  Plan pl;
  TRY(stage_batch(c, s, p, b, &pl));
  // The candidate stage needs two small device->host handshakes (arena size, big-path plan), so
  // the kernels of this slot are driven here; the other slot's D2H and the caller's host work
  // overlap with them because each slot owns its stream.
  TRY(run_seed_and_candidates(c, s, pl));
  TRY(run_verify_and_compact(c, s, pl));
  s.has_aln = c->device_traceback != 0;
  TRY(run_sam(c, s));
  s.busy = true;
  c->last_slot = slot;
  return 0;
}

extern "C" int femgpu_set_sam_inputs(femgpu_ctx *c, int slot, const femgpu_sam_inputs *in) {
  if (!c || slot < 0 || slot > 1) return set_error(FEMGPU_EINVAL, "bad context or slot");
  Slot &s = c->slot[slot];
  if (!in) { s.sam_pending = false; return 0; }
  if (!in->names || !in->name_off || !in->quals || !in->has_qual) return set_error(FEMGPU_EINVAL, "null SAM input");
  s.sam_in = *in;
  s.sam_pending = true;
  return 0;
}

extern "C" int femgpu_set_reference_names(femgpu_ctx *c, uint32_t n_records, const char *const *names) {
  if (!c || !names) return set_error(FEMGPU_EINVAL, "null argument");
  if (!c->n_rec || n_records != c->n_rec) return set_error(FEMGPU_ESTATE, "reference names must follow femgpu_upload_reference (%u records resident)", c->n_rec);
  CU(cudaSetDevice(c->device));
  std::vector<uint32_t> off(n_records + 1, 0);
  std::string all;
  for (uint32_t i = 0; i < n_records; ++i) {
    if (!names[i]) return set_error(FEMGPU_EINVAL, "null record name");
    all += names[i];
    off[i + 1] = (uint32_t)all.size();
  }
  TRY(c->rnames.reserve(all.size() + 16));
  TRY(c->rname_off.reserve((n_records + 1ull) * 4));
  CU(cudaMemcpy(c->rnames.p, all.data(), all.size(), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(c->rname_off.p, off.data(), (n_records + 1ull) * 4, cudaMemcpyHostToDevice));
  c->have_rnames = true;
  return 0;
}

extern "C" int femgpu_wait_batch(femgpu_ctx *c, int slot, femgpu_result *r) {
  if (!c || slot < 0 || slot > 1) return set_error(FEMGPU_EINVAL, "bad context or slot");
  Slot &s = c->slot[slot];
  if (!s.busy) return set_error(FEMGPU_ESTATE, "slot %d has no batch in flight", slot);
  CU(cudaSetDevice(c->device));
  s.busy = false;
  return fetch_result(s, r);
}

extern "C" int femgpu_map_batch(femgpu_ctx *c, const femgpu_params *p, const femgpu_batch *b, femgpu_result *r) {
  TRY(femgpu_submit_batch(c, 0, p, b));
  return femgpu_wait_batch(c, 0, r);
}

// ---- stage-level: candidates only -------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_gather_candidates(const uint64_t *cand, const uint64_t *rs_cand_first,
                                                           const uint32_t *rs_cand_count, const uint64_t *sorted_first,
                                                           uint32_t n_rs, uint64_t *out) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t rs = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (rs >= n_rs) return;
  const uint64_t src = rs_cand_first[rs], dst = sorted_first[rs];
  const uint32_t n = rs_cand_count[rs];
  for (uint32_t t = lane; t < n; t += 32) out[dst + t] = cand[src + t];
}

extern "C" int femgpu_candidates_batch(femgpu_ctx *c, const femgpu_params *p, const femgpu_batch *b, femgpu_candidates *out) {
  if (!c || !out) return set_error(FEMGPU_EINVAL, "null argument");
  Slot &s = c->slot[0];
  if (s.busy) return set_error(FEMGPU_ESTATE, "slot 0 still holds an unfinished batch");
  Plan pl;
  TRY(stage_batch(c, s, p, b, &pl));
  TRY(run_seed_and_candidates(c, s, pl));
  const uint64_t n_rs = 2ull * s.n_reads;
  TRY(s.h_cands.reserve((n_rs + 1) * 8 + s.n_cand * 8 + 16));
  TRY(s.h_before.reserve(n_rs * 4 + 16));
  TRY(s.h_status.reserve(n_rs + 16));
  uint64_t *hfirst = s.h_cands.as<uint64_t>();
  if (n_rs == 0) {
    hfirst[0] = 0;
  } else {
    Counters *dc = s.counters.as<Counters>();
    TRY(s.rs_cand_sorted_first.reserve((n_rs + 1) * 8));
    TRY(s.cand_sorted.reserve(s.n_cand * 8 + 16));
    const unsigned n_blocks = (unsigned)((n_rs + SCAN_BLOCK - 1) / SCAN_BLOCK);
    k_scan_reduce<<<n_blocks, SCAN_BLOCK, 0, s.stream>>>(s.rs_cand_count.as<uint32_t>(), n_rs, s.block_sums.as<uint64_t>());
    k_scan_blocks<<<1, 32, 0, s.stream>>>(s.block_sums.as<uint64_t>(), n_blocks, (uint64_t *)&dc->total_maps);
    k_scan_apply<<<n_blocks, SCAN_BLOCK, 0, s.stream>>>(s.rs_cand_count.as<uint32_t>(), n_rs, s.block_sums.as<uint64_t>(),
                                                       s.rs_cand_sorted_first.as<uint64_t>(), (uint64_t *)&dc->total_maps);
    k_gather_candidates<<<(unsigned)((n_rs + 7) / 8), 256, 0, s.stream>>>(s.cand.as<uint64_t>(), s.rs_cand_first.as<uint64_t>(),
                                                                        s.rs_cand_count.as<uint32_t>(), s.rs_cand_sorted_first.as<uint64_t>(),
                                                                        (uint32_t)n_rs, s.cand_sorted.as<uint64_t>());
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(hfirst, s.rs_cand_sorted_first.p, (n_rs + 1) * 8, cudaMemcpyDeviceToHost, s.stream));
    if (s.n_cand) CU(cudaMemcpyAsync(hfirst + n_rs + 1, s.cand_sorted.p, s.n_cand * 8, cudaMemcpyDeviceToHost, s.stream));
    CU(cudaMemcpyAsync(s.h_before.p, s.rs_before.p, n_rs * 4, cudaMemcpyDeviceToHost, s.stream));
    CU(cudaMemcpyAsync(s.h_status.p, s.rs_status.p, n_rs, cudaMemcpyDeviceToHost, s.stream));
  }
  CU(slot_sync(s));
  out->first = hfirst;
  out->cands = hfirst + n_rs + 1;
  out->before = s.h_before.as<uint32_t>();
  out->status = s.h_status.as<uint8_t>();
  return 0;
}

// ---- measurement hooks -----------------------------------------------------------------------------
extern "C" int femgpu_last_timings(femgpu_ctx *c, float ms[6], uint32_t *n_launches) {
  if (!c) return set_error(FEMGPU_EINVAL, "null context");
  const Slot &s = c->slot[c->last_slot];
  if (ms) memcpy(ms, s.ms, sizeof s.ms);
  if (n_launches) *n_launches = s.n_launches;
  return 0;
}

extern "C" int femgpu_last_work(femgpu_ctx *c, uint64_t work[4]) {
  if (!c || !work) return set_error(FEMGPU_EINVAL, "null argument");
  const Slot &s = c->slot[c->last_slot];
  work[0] = s.cols_executed;
  work[1] = s.n_cand;
  work[2] = s.stats.num_candidates_before_filter;
  work[3] = s.n_maps;
  return 0;
}

extern "C" int femgpu_rerun_resident(femgpu_ctx *c, float *ms_total, float ms_stage[6]) {
  if (!c) return set_error(FEMGPU_EINVAL, "null context");
  Slot &s = c->slot[c->last_slot];
  if (s.busy) return set_error(FEMGPU_ESTATE, "slot busy");
  CU(cudaSetDevice(c->device));
  Plan pl;
  TRY(make_plan(c, &s.params, s.l_max, &pl, s.n_too_long != 0));
  cudaEventRecord(s.ev[0], s.stream);
  TRY(run_seed_and_candidates(c, s, pl));
  TRY(run_verify_and_compact(c, s, pl));
  TRY(run_sam(c, s)); // the batch's names and qualities are still resident when it was submitted with SAM inputs
  cudaEventRecord(s.ev[7], s.stream);
  CU(slot_sync(s));
  float t;
  if (ms_stage)
  {
    for (int i = 0; i < 4; ++i) { cudaEventElapsedTime(&t, s.ev[i], s.ev[i + 1]); ms_stage[i] = t; }
    cudaEventElapsedTime(&t, s.ev[4], s.ev[6]); ms_stage[4] = t;
    cudaEventElapsedTime(&t, s.ev[6], s.ev[7]); ms_stage[5] = t;
  }
  cudaEventElapsedTime(&t, s.ev[0], s.ev[7]);
  if (ms_total) *ms_total = t;
  return 0;
}

// ---- host: the reference's per-read mapping order ----------------------------------------------------
// Key and algorithm of src/align.c:53-57 + klib's radix sort (src/ksort.h:98-151): <= 64 records
// insertion sort (stable), otherwise an in-place MSD byte-wise "American flag" permutation whose
// tie order for equal keys depends on the cycle-leader walk, so it is reproduced step for step.
namespace {
inline uint64_t map_key(const femgpu_mapping &m) {
  return ((uint64_t)m.edit_distance << 60) | ((uint64_t)m.direction << 59) |
         (m.candidate_position + (uint64_t)(int64_t)m.end_position_offset);
}
void map_insertion_sort(femgpu_mapping *b, femgpu_mapping *e) {
  for (femgpu_mapping *i = b + 1; i < e; ++i) {
    if (map_key(*i) >= map_key(*(i - 1))) continue;
    const femgpu_mapping t = *i;
    const uint64_t kt = map_key(t);
    femgpu_mapping *j = i;
    for (; j > b && kt < map_key(*(j - 1)); --j) *j = *(j - 1);
    *j = t;
  }
}
void map_flag_sort(femgpu_mapping *beg, femgpu_mapping *end, int shift) {
  struct Bucket { femgpu_mapping *head, *tail; } bk[256];
  size_t hist[256] = {0};
  for (femgpu_mapping *i = beg; i != end; ++i) ++hist[(map_key(*i) >> shift) & 255];
  femgpu_mapping *p = beg;
  for (int d = 0; d < 256; ++d) { bk[d].head = p; p += hist[d]; bk[d].tail = p; }
  for (int d = 0; d < 256;) {
    Bucket &cur = bk[d];
    if (cur.head == cur.tail) { ++d; continue; }
    int dest = (int)((map_key(*cur.head) >> shift) & 255);
    if (dest == d) { ++cur.head; continue; }
    femgpu_mapping moving = *cur.head;
    while (dest != d) { // push `moving` into its bucket, pick up the record it displaces
      std::swap(moving, *bk[dest].head);
      ++bk[dest].head;
      dest = (int)((map_key(moving) >> shift) & 255);
    }
    *cur.head++ = moving;
  }
  if (shift == 0) return;
  const int next = shift > 8 ? shift - 8 : 0;
  femgpu_mapping *b = beg;
  for (int d = 0; d < 256; ++d) {
    femgpu_mapping *e = bk[d].tail;
    if (e - b > 64) map_flag_sort(b, e, next);
    else if (e - b > 1) map_insertion_sort(b, e);
    b = e;
  }
}
} // namespace

// The same permutation applied to a parallel array of alignments: the records are sorted with
// their original index in the two reserved bytes' place (a wider copy), then the alignments follow.
extern "C" void femgpu_sort_mappings(femgpu_mapping *maps, size_t n);
extern "C" void femgpu_sort_mappings_aln(femgpu_mapping *maps, femgpu_alignment *alns, size_t n) {
  if (!maps || n < 2 || !alns) { femgpu_sort_mappings(maps, n); return; }
  // equal sort keys are the only source of ambiguity; tag every record with its position so that
  // the permutation can be read back.  The tag lives in bytes the key never looks at.
  std::vector<uint32_t> idx(n);
  std::vector<femgpu_mapping> tagged(maps, maps + n);
  const bool small = n <= 0xFFFF;
  std::vector<femgpu_alignment> src(alns, alns + n);
  if (small) {
    for (size_t i = 0; i < n; ++i) { tagged[i].reserved[0] = (uint8_t)(i & 255); tagged[i].reserved[1] = (uint8_t)(i >> 8); }
    femgpu_sort_mappings(tagged.data(), n);
    for (size_t i = 0; i < n; ++i) {
      const size_t from = (size_t)tagged[i].reserved[0] | ((size_t)tagged[i].reserved[1] << 8);
      alns[i] = src[from];
      maps[i] = tagged[i];
      maps[i].reserved[0] = maps[i].reserved[1] = 0;
    }
    return;
  }
  // more than 65535 mappings of one read: 32-bit tag in the upper half of end_position_offset's
  // slot is not available, so sort (record, index) pairs with the same algorithm
  struct Pair { femgpu_mapping m; uint32_t i; };
  std::vector<Pair> pr(n);
  for (size_t i = 0; i < n; ++i) { pr[i].m = maps[i]; pr[i].i = (uint32_t)i; }
  struct S {
    static uint64_t key(const Pair &p) { return map_key(p.m); }
    static void insertion(Pair *b, Pair *e) {
      for (Pair *i = b + 1; i < e; ++i) {
        if (key(*i) >= key(*(i - 1))) continue;
        const Pair t = *i;
        Pair *j = i;
        for (; j > b && key(t) < key(*(j - 1)); --j) *j = *(j - 1);
        *j = t;
      }
    }
    static void flag(Pair *beg, Pair *end, int shift) {
      struct Bucket { Pair *head, *tail; } bk[256];
      size_t hist[256] = {0};
      for (Pair *i = beg; i != end; ++i) ++hist[(key(*i) >> shift) & 255];
      Pair *p = beg;
      for (int d = 0; d < 256; ++d) { bk[d].head = p; p += hist[d]; bk[d].tail = p; }
      for (int d = 0; d < 256;) {
        Bucket &cur = bk[d];
        if (cur.head == cur.tail) { ++d; continue; }
        int dest = (int)((key(*cur.head) >> shift) & 255);
        if (dest == d) { ++cur.head; continue; }
        Pair moving = *cur.head;
        while (dest != d) {
          std::swap(moving, *bk[dest].head);
          ++bk[dest].head;
          dest = (int)((key(moving) >> shift) & 255);
        }
        *cur.head++ = moving;
      }
      if (shift == 0) return;
      const int next = shift > 8 ? shift - 8 : 0;
      Pair *b = beg;
      for (int d = 0; d < 256; ++d) {
        Pair *e = bk[d].tail;
        if (e - b > 64) flag(b, e, next);
        else if (e - b > 1) insertion(b, e);
        b = e;
      }
    }
  };
  S::flag(pr.data(), pr.data() + n, 56);
  for (size_t i = 0; i < n; ++i) { maps[i] = pr[i].m; alns[i] = src[pr[i].i]; }
}

extern "C" void femgpu_sort_mappings(femgpu_mapping *maps, size_t n) {
  if (!maps || n < 2) return;
  if (n <= 64) map_insertion_sort(maps, maps + n);
  else map_flag_sort(maps, maps + n, 56);
}
