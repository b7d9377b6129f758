// upload.inl -- index and reference upload (included by femgpu.cu)

// Packs raw reference characters into 2 bits per base + ambiguity mask.  Each thread packs 32
// bases: two ref2 words and one refn word.  `g0` (multiple of 32) is the position of chunk[0] in
// the concatenated reference.
__global__ void __launch_bounds__(256) k_pack_reference(const unsigned char *chunk, uint64_t n, uint64_t g0,
                                                        uint32_t *ref2, uint32_t *refn, int *any_n) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t i0 = t * 32;
  if (i0 >= n) return;
  uint32_t w0 = 0, w1 = 0, wn = 0;
  bool noncanon = false;
  const int cnt = (int)min((uint64_t)32, n - i0);
  for (int j = 0; j < cnt; ++j) {
    const unsigned char ch = chunk[i0 + j];
    const uint32_t c = fem_base_code(ch);
    if (!(ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T' || ch == 'N')) noncanon = true;
    const uint32_t two = c & 3u; // ambiguous packs as 0; the mask decides
    if (j < 16) w0 |= (c < 4 ? two : 0u) << (2 * j);
    else w1 |= (c < 4 ? two : 0u) << (2 * (j - 16));
    if (c == 4) wn |= 1u << j;
  }
  const uint64_t g = g0 + i0;
  ref2[g >> 4] = w0;
  ref2[(g >> 4) + 1] = w1;
  refn[g >> 5] = wn;
  if (wn) atomicOr(any_n, 1);
  if (noncanon) atomicOr(any_n, 2); // a character outside "ACGTN": k_trace leaves such input to the host
}

extern "C" int femgpu_upload_index(femgpu_ctx *c, int32_t k, int32_t step, const uint32_t *lut, size_t lut_n,
                                   const uint64_t *occ, size_t occ_n) {
  if (!c || !lut || (!occ && occ_n)) return set_error(FEMGPU_EINVAL, "null argument");
  if (k < 1 || k > 15 || step < 1) return set_error(FEMGPU_EINVAL, "bad index header: k=%d step=%d", k, step);
  if (lut_n != ((size_t)1 << (2 * k)) + 1) return set_error(FEMGPU_EINVAL, "lookup table size %zu does not match k=%d", lut_n, k);
  if (lut[lut_n - 1] != (uint32_t)occ_n) return set_error(FEMGPU_EINVAL, "lookup table total %u != occurrence table size %zu", lut[lut_n - 1], occ_n);
  CU(cudaSetDevice(c->device));
  TRY(c->lut.reserve(lut_n * sizeof(uint32_t)));
  TRY(c->occ.reserve(std::max<size_t>(occ_n, 1) * sizeof(uint64_t) + 64));
  CU(cudaMemcpy(c->lut.p, lut, lut_n * sizeof(uint32_t), cudaMemcpyHostToDevice));
  if (occ_n) CU(cudaMemcpy(c->occ.p, occ, occ_n * sizeof(uint64_t), cudaMemcpyHostToDevice));
  c->n_occ = occ_n;
  c->idx_k = k;
  c->idx_step = step;
  return 0;
}

extern "C" int femgpu_upload_index_file(femgpu_ctx *c, const char *path) {
  if (!c || !path) return set_error(FEMGPU_EINVAL, "null argument");
  FILE *f = fopen(path, "rb");
  if (!f) return set_error(FEMGPU_EINVAL, "Failed to open index file %s", path);
  int k = 0, step = 0;
  if (fread(&k, sizeof(int), 1, f) != 1 || fread(&step, sizeof(int), 1, f) != 1 || k < 1 || k > 15 || step < 1) {
    fclose(f);
    return set_error(FEMGPU_EINVAL, "bad index header in %s", path);
  }
  CU(cudaSetDevice(c->device));
  const size_t lut_n = ((size_t)1 << (2 * k)) + 1;
  const size_t chunk = 64u << 20;
  HBuf stage[2];
  int rc = 0;
  auto fail = [&](int code) { fclose(f); stage[0].release(); stage[1].release(); return code; };
  if ((rc = stage[0].reserve(chunk)) || (rc = stage[1].reserve(chunk))) return fail(rc);
  if ((rc = c->lut.reserve(lut_n * sizeof(uint32_t)))) return fail(rc);
  cudaStream_t st = c->slot[0].stream;
  cudaEvent_t ev[2];
  cudaEventCreate(&ev[0]); cudaEventCreate(&ev[1]);
  auto stream_in = [&](void *dst, size_t bytes) -> int { // file -> pinned -> device, double buffered
    size_t done = 0;
    int b = 0;
    while (done < bytes) {
      const size_t n = std::min(chunk, bytes - done);
      cudaEventSynchronize(ev[b]);
      if (fread(stage[b].p, 1, n, f) != n) return set_error(FEMGPU_EINVAL, "index file %s is truncated", path);
      if (cudaMemcpyAsync((char *)dst + done, stage[b].p, n, cudaMemcpyHostToDevice, st) != cudaSuccess)
        return set_error(FEMGPU_ECUDA, "H2D of index failed: %s", cudaGetErrorString(cudaGetLastError()));
      cudaEventRecord(ev[b], st);
      done += n;
      b ^= 1;
    }
    cudaStreamSynchronize(st);
    return 0;
  };
  rc = stream_in(c->lut.p, lut_n * sizeof(uint32_t));
  size_t occ_n = 0;
  if (!rc && fread(&occ_n, sizeof(size_t), 1, f) != 1) rc = set_error(FEMGPU_EINVAL, "index file %s is truncated", path);
  if (!rc) rc = c->occ.reserve(std::max<size_t>(occ_n, 1) * sizeof(uint64_t) + 64);
  if (!rc) rc = stream_in(c->occ.p, occ_n * sizeof(uint64_t));
  cudaEventDestroy(ev[0]); cudaEventDestroy(ev[1]);
  if (rc) return fail(rc);
  uint32_t total = 0;
  cudaMemcpy(&total, c->lut.as<uint32_t>() + lut_n - 1, sizeof total, cudaMemcpyDeviceToHost);
  if (total != (uint32_t)occ_n) return fail(set_error(FEMGPU_EINVAL, "index file %s: lookup table total %u != occurrence count %zu", path, total, occ_n));
  fail(0);
  c->n_occ = occ_n;
  c->idx_k = k;
  c->idx_step = step;
  return 0;
}

extern "C" int femgpu_upload_reference(femgpu_ctx *c, uint32_t n_rec, const char *const *bases, const uint32_t *lens) {
  if (!c || !bases || !lens || n_rec == 0) return set_error(FEMGPU_EINVAL, "null or empty reference");
  CU(cudaSetDevice(c->device));
  std::vector<uint64_t> base(n_rec + 1, 0);
  for (uint32_t i = 0; i < n_rec; ++i) base[i + 1] = base[i] + lens[i];
  const uint64_t total = base[n_rec];
  const size_t n_w2 = (size_t)((total + 31) / 32) * 2 + 4, n_wn = (size_t)((total + 31) / 32) + 2;
  TRY(c->ref2.reserve(n_w2 * 4));
  TRY(c->refn.reserve(n_wn * 4));
  TRY(c->rec_base.reserve((n_rec + 1) * 8));
  TRY(c->rec_len.reserve(n_rec * 4));
  CU(cudaMemset(c->ref2.p, 0, n_w2 * 4));
  CU(cudaMemset(c->refn.p, 0, n_wn * 4));
  CU(cudaMemcpy(c->rec_base.p, base.data(), (n_rec + 1) * 8, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(c->rec_len.p, lens, n_rec * 4, cudaMemcpyHostToDevice));
  const size_t chunk = 64u << 20; // multiple of 32
  HBuf stage[2];
  DBuf dchunk[2], flag;
  int rc = 0;
  auto cleanup = [&](int code) { stage[0].release(); stage[1].release(); dchunk[0].release(); dchunk[1].release(); flag.release(); return code; };
  if ((rc = stage[0].reserve(chunk)) || (rc = stage[1].reserve(chunk)) || (rc = dchunk[0].reserve(chunk)) ||
      (rc = dchunk[1].reserve(chunk)) || (rc = flag.reserve(4)))
    return cleanup(rc);
  cudaStream_t st = c->slot[0].stream;
  cudaMemsetAsync(flag.p, 0, 4, st);
  cudaEvent_t ev[2];
  cudaEventCreate(&ev[0]); cudaEventCreate(&ev[1]);
  uint32_t rec = 0;
  uint64_t in_rec = 0;
  int b = 0;
  for (uint64_t g0 = 0; g0 < total; g0 += chunk, b ^= 1) {
    const size_t n = (size_t)std::min<uint64_t>(chunk, total - g0);
    cudaEventSynchronize(ev[b]);
    size_t filled = 0;
    while (filled < n) { // concatenate records into the staging chunk
      while (in_rec == lens[rec]) { ++rec; in_rec = 0; }
      const size_t take = (size_t)std::min<uint64_t>(n - filled, lens[rec] - in_rec);
      memcpy(stage[b].as<char>() + filled, bases[rec] + in_rec, take);
      filled += take;
      in_rec += take;
    }
    cudaMemcpyAsync(dchunk[b].p, stage[b].p, n, cudaMemcpyHostToDevice, st);
    const uint64_t threads = (n + 31) / 32;
    k_pack_reference<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(dchunk[b].as<unsigned char>(), n, g0,
                                                                      c->ref2.as<uint32_t>(), c->refn.as<uint32_t>(), flag.as<int>());
    cudaEventRecord(ev[b], st);
  }
  int any_n = 0;
  cudaMemcpyAsync(&any_n, flag.p, 4, cudaMemcpyDeviceToHost, st);
  cudaError_t err = cudaStreamSynchronize(st);
  cudaEventDestroy(ev[0]); cudaEventDestroy(ev[1]);
  cleanup(0);
  if (err != cudaSuccess) return set_error(FEMGPU_ECUDA, "reference upload failed: %s", cudaGetErrorString(err));
  c->n_rec = n_rec;
  c->have_rnames = false; // names belong to the previous reference
  c->ref_total = total;
  c->ref_has_n = (any_n & 1) != 0;
  c->ref_canonical = (any_n & 2) == 0;
  return derive_ref_planes(c);
}

// The resident index, reference and record names of `src` copied device to device (NVLink peer copies
// where the devices are peers, through the host otherwise): one read of the index file and one packing
// of the reference serve every GPU of the node.  Stands in for the per-process load_index +
// load_all_sequences of src/FEM_map.c:136-143 on the second and later devices.
extern "C" int femgpu_clone_resident(femgpu_ctx *dst, const femgpu_ctx *src) {
  if (!dst || !src || dst == src) return set_error(FEMGPU_EINVAL, "clone: bad contexts");
  if (!src->lut.p || !src->n_rec) return set_error(FEMGPU_ESTATE, "clone: the source holds no index or no reference");
  CU(cudaSetDevice(dst->device));
  if (dst->device != src->device) {
    int can = 0;
    if (cudaDeviceCanAccessPeer(&can, dst->device, src->device) == cudaSuccess && can) {
      const cudaError_t e = cudaDeviceEnablePeerAccess(src->device, 0);
      if (e != cudaSuccess) cudaGetLastError(); // already enabled, or not possible: cudaMemcpyPeer stages through the host
    }
  }
  cudaStream_t st = dst->slot[0].stream;
  const uint64_t k_lut = (((uint64_t)1 << (2 * src->idx_k)) + 1) * sizeof(uint32_t);
  const uint64_t n_w2 = (uint64_t)((src->ref_total + 31) / 32) * 2 + 4, n_wn = (uint64_t)((src->ref_total + 31) / 32) + 2;
  struct { DBuf *d; const DBuf *s; uint64_t bytes; } parts[] = {
      {&dst->lut, &src->lut, k_lut},
      {&dst->occ, &src->occ, std::max<uint64_t>(src->n_occ, 1) * sizeof(uint64_t)},
      {&dst->ref2, &src->ref2, n_w2 * 4},
      {&dst->refn, &src->refn, n_wn * 4},
      {&dst->rec_base, &src->rec_base, ((uint64_t)src->n_rec + 1) * 8},
      {&dst->rec_len, &src->rec_len, (uint64_t)src->n_rec * 4},
      {&dst->rnames, &src->rnames, src->have_rnames ? src->rnames.bytes : 0},
      {&dst->rname_off, &src->rname_off, src->have_rnames ? ((uint64_t)src->n_rec + 1) * 4 : 0},
  };
  for (auto &pt : parts) {
    if (!pt.bytes) continue;
    if (pt.bytes > pt.s->bytes) return set_error(FEMGPU_ESTATE, "clone: source buffer smaller than its contents");
    TRY(pt.d->reserve(pt.bytes + 64));
    CU(cudaMemcpyPeerAsync(pt.d->p, dst->device, pt.s->p, src->device, pt.bytes, st));
  }
  CU(cudaStreamSynchronize(st));
  dst->n_occ = src->n_occ;
  dst->idx_k = src->idx_k;
  dst->idx_step = src->idx_step;
  dst->n_rec = src->n_rec;
  dst->ref_total = src->ref_total;
  dst->ref_has_n = src->ref_has_n;
  dst->ref_canonical = src->ref_canonical;
  dst->have_rnames = src->have_rnames;
  return derive_ref_planes(dst);
}

// Bit planes of the packed reference for k_verify's plane form (cheap to derive: one pass over ref2 on the device).
static int derive_ref_planes(femgpu_ctx *c) {
  const uint64_t n_words = (c->ref_total + 31) / 32;
  TRY(c->refp0.reserve((n_words + 4) * 4));
  TRY(c->refp1.reserve((n_words + 4) * 4));
  cudaStream_t st = c->slot[0].stream;
  CU(cudaMemsetAsync(c->refp0.p, 0, (n_words + 4) * 4, st));
  CU(cudaMemsetAsync(c->refp1.p, 0, (n_words + 4) * 4, st));
  if (n_words) k_ref_planes<<<(unsigned)((n_words + 255) / 256), 256, 0, st>>>(c->ref2.as<uint32_t>(), n_words, c->refp0.as<uint32_t>(), c->refp1.as<uint32_t>());
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(st));
  return 0;
}
