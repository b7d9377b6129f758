// index_build.inl -- femgpu_build_index / femgpu_download_index / femgpu_save_index_file
// (included by femgpu.cu after pipeline.inl)

extern "C" int femgpu_build_index(femgpu_ctx *c, int32_t k, int32_t step) {
  if (!c) return set_error(FEMGPU_EINVAL, "null argument");
  if (k < 1 || k > 15 || step < 1) return set_error(FEMGPU_EINVAL, "bad index parameters k=%d step=%d", k, step);
  if (!c->n_rec) return set_error(FEMGPU_ESTATE, "reference not uploaded");
  CU(cudaSetDevice(c->device));
  cudaStream_t st = c->slot[0].stream;
  const uint32_t n_rec = c->n_rec;
  // samples per record: positions 0, step, ... while pos + k <= len (src/index.c:65)
  std::vector<uint32_t> lens(n_rec);
  CU(cudaMemcpy(lens.data(), c->rec_len.p, (size_t)n_rec * 4, cudaMemcpyDeviceToHost));
  std::vector<uint64_t> sbase(n_rec + 1, 0);
  for (uint32_t r = 0; r < n_rec; ++r)
    sbase[r + 1] = sbase[r] + (lens[r] >= (uint32_t)k ? (uint64_t)(lens[r] - (uint32_t)k) / (uint32_t)step + 1 : 0);
  const uint64_t n = sbase[n_rec];
  if (n > 0xFFFFFFFFull) return set_error(FEMGPU_EINVAL, "more than 2^32-1 sampled q-grams");
  const uint64_t n_buckets = 1ull << (2 * k);
  const size_t lut_n = (size_t)n_buckets + 1;
  const uint32_t n_tiles = (uint32_t)((n_buckets + INDEX_SCAN_TILE - 1) / INDEX_SCAN_TILE);
  const uint64_t big_cap = n / INDEX_WARP_CAP + 1;
  DBuf d_sbase, tile_sum, big_list, big_count, big_b, big_e, scratch, scratch_off;
  auto cleanup = [&](int code) {
    d_sbase.release(); tile_sum.release(); big_list.release(); big_count.release(); big_b.release(); big_e.release();
    scratch.release(); scratch_off.release();
    return code;
  };
#define TRYC(x) do { int rc__ = (x); if (rc__) return cleanup(rc__); } while (0)
#define CUC(call) do { cudaError_t err__ = (call); if (err__ != cudaSuccess) return cleanup(set_error(err__ == cudaErrorMemoryAllocation ? FEMGPU_ENOMEM : FEMGPU_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(err__), __FILE__, __LINE__)); } while (0)
  c->n_occ = 0; c->idx_k = 0; c->idx_step = 0; // no index until the build has finished
  TRYC(c->lut.reserve(lut_n * sizeof(uint32_t)));
  TRYC(c->occ.reserve(std::max<uint64_t>(n, 1) * sizeof(uint64_t) + 64));
  TRYC(d_sbase.reserve((n_rec + 1) * 8));
  TRYC(tile_sum.reserve((size_t)n_tiles * 4 + 16));
  TRYC(big_list.reserve(big_cap * 4));
  TRYC(big_count.reserve(16));
  CUC(cudaMemcpyAsync(d_sbase.p, sbase.data(), (n_rec + 1) * 8, cudaMemcpyHostToDevice, st));
  CUC(cudaMemsetAsync(c->lut.p, 0, lut_n * sizeof(uint32_t), st));
  CUC(cudaMemsetAsync(big_count.p, 0, 16, st));
  IndexArgs ia;
  ia.ref2 = c->ref2.as<uint32_t>();
  ia.rec_base = c->rec_base.as<uint64_t>();
  ia.sample_base = d_sbase.as<uint64_t>();
  ia.n_rec = n_rec;
  ia.n_samples = n;
  ia.k = k; ia.step = step;
  ia.lut = c->lut.as<uint32_t>();
  ia.occ = c->occ.as<uint64_t>();
  const unsigned pass_grid = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>((n + 255) / 256, (uint64_t)c->sm_count * 64));
  if (n) k_index_pass<false><<<pass_grid, 256, 0, st>>>(ia);
  k_index_scan_reduce<<<n_tiles, 256, 0, st>>>(ia.lut + 1, n_buckets, tile_sum.as<uint32_t>());
  k_index_scan_tiles<<<1, 256, 0, st>>>(tile_sum.as<uint32_t>(), n_tiles);
  k_index_scan_apply<<<n_tiles, 256, 0, st>>>(ia.lut + 1, n_buckets, tile_sum.as<uint32_t>());
  if (n) k_index_pass<true><<<pass_grid, 256, 0, st>>>(ia);
  const size_t warp_smem = 4 * (size_t)INDEX_WARP_CAP * 8;
  const unsigned sort_grid = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>((n_buckets + 3) / 4, (uint64_t)c->sm_count * 48));
  k_index_sort_warp<<<sort_grid, 128, warp_smem, st>>>(ia.lut, n_buckets, ia.occ, big_list.as<uint32_t>(), big_count.as<uint32_t>());
  CUC(cudaGetLastError());
  uint32_t n_big = 0;
  CUC(cudaMemcpyAsync(&n_big, big_count.p, 4, cudaMemcpyDeviceToHost, st));
  CUC(cudaStreamSynchronize(st));
  if (n_big) {
    // sizes of the big buckets -> largest first, global scratch slices for those beyond shared memory
    std::vector<uint32_t> list(n_big), hb(n_big), he(n_big), next(n_big);
    TRYC(big_b.reserve((size_t)n_big * 4));
    TRYC(big_e.reserve((size_t)n_big * 4));
    TRYC(scratch_off.reserve((size_t)n_big * 8));
    CUC(cudaMemcpyAsync(list.data(), big_list.p, (size_t)n_big * 4, cudaMemcpyDeviceToHost, st));
    CUC(cudaStreamSynchronize(st));
    for (uint32_t i = 0; i < n_big; ++i) next[i] = list[i] + 1;
    k_gather_u32<<<(n_big + 255) / 256, 256, 0, st>>>(ia.lut, big_list.as<uint32_t>(), n_big, big_b.as<uint32_t>());
    CUC(cudaMemcpyAsync(big_list.p, next.data(), (size_t)n_big * 4, cudaMemcpyHostToDevice, st));
    k_gather_u32<<<(n_big + 255) / 256, 256, 0, st>>>(ia.lut, big_list.as<uint32_t>(), n_big, big_e.as<uint32_t>());
    CUC(cudaMemcpyAsync(hb.data(), big_b.p, (size_t)n_big * 4, cudaMemcpyDeviceToHost, st));
    CUC(cudaMemcpyAsync(he.data(), big_e.p, (size_t)n_big * 4, cudaMemcpyDeviceToHost, st));
    CUC(cudaStreamSynchronize(st));
    std::vector<uint32_t> order(n_big);
    for (uint32_t i = 0; i < n_big; ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) {
      const uint32_t nx = he[x] - hb[x], ny = he[y] - hb[y];
      return nx != ny ? nx > ny : list[x] < list[y];
    });
    std::vector<uint32_t> sorted_list(n_big);
    std::vector<uint64_t> off(n_big, 0);
    uint64_t used = 0;
    for (uint32_t i = 0; i < n_big; ++i) {
      sorted_list[i] = list[order[i]];
      const uint32_t sz = he[order[i]] - hb[order[i]];
      if (sz > INDEX_BLOCK_CAP) { off[i] = used; used += pow2_ceil_u32(sz); }
    }
    if (used) TRYC(scratch.reserve(used * 8));
    CUC(cudaMemcpyAsync(big_list.p, sorted_list.data(), (size_t)n_big * 4, cudaMemcpyHostToDevice, st));
    CUC(cudaMemcpyAsync(scratch_off.p, off.data(), (size_t)n_big * 8, cudaMemcpyHostToDevice, st));
    const size_t block_smem = (size_t)INDEX_BLOCK_CAP * 8;
    CUC(cudaFuncSetAttribute(k_index_sort_block, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)block_smem));
    k_index_sort_block<<<std::min<uint32_t>(n_big, (uint32_t)c->sm_count * 2), 1024, block_smem, st>>>(
        ia.lut, ia.occ, big_list.as<uint32_t>(), n_big, scratch.as<uint64_t>(), scratch_off.as<uint64_t>());
    CUC(cudaGetLastError());
    CUC(cudaStreamSynchronize(st)); // the host vectors above must outlive the copies
  }
  uint32_t total = 0;
  CUC(cudaMemcpy(&total, c->lut.as<uint32_t>() + n_buckets, 4, cudaMemcpyDeviceToHost));
  if (total != (uint32_t)n) return cleanup(set_error(FEMGPU_ECUDA, "internal: index build counted %u of %llu samples", total, (unsigned long long)n));
#undef TRYC
#undef CUC
  cleanup(0);
  c->n_occ = n;
  c->idx_k = k;
  c->idx_step = step;
  return 0;
}

extern "C" int femgpu_index_sizes(const femgpu_ctx *c, size_t *lut_n, size_t *occ_n) {
  if (!c || !c->idx_k) return set_error(FEMGPU_ESTATE, "index not uploaded");
  if (lut_n) *lut_n = ((size_t)1 << (2 * c->idx_k)) + 1;
  if (occ_n) *occ_n = (size_t)c->n_occ;
  return 0;
}

extern "C" int femgpu_download_index(femgpu_ctx *c, uint32_t *lut, size_t lut_n, uint64_t *occ, size_t occ_n) {
  if (!c || !lut || (!occ && occ_n)) return set_error(FEMGPU_EINVAL, "null argument");
  if (!c->idx_k) return set_error(FEMGPU_ESTATE, "index not uploaded");
  if (lut_n != ((size_t)1 << (2 * c->idx_k)) + 1 || occ_n != c->n_occ) return set_error(FEMGPU_EINVAL, "array sizes do not match the resident index");
  CU(cudaSetDevice(c->device));
  CU(cudaMemcpy(lut, c->lut.p, lut_n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  if (occ_n) CU(cudaMemcpy(occ, c->occ.p, occ_n * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  return 0;
}

// Writes the resident index in the reference's file layout (src/index.c:133-168): device -> pinned
// chunks -> file, double buffered.
extern "C" int femgpu_save_index_file(femgpu_ctx *c, const char *path) {
  if (!c || !path) return set_error(FEMGPU_EINVAL, "null argument");
  if (!c->idx_k) return set_error(FEMGPU_ESTATE, "index not uploaded");
  CU(cudaSetDevice(c->device));
  FILE *f = fopen(path, "wb");
  if (!f) return set_error(FEMGPU_EINVAL, "Cannot open index file %s for writing", path);
  const size_t chunk = 64u << 20;
  HBuf stage[2];
  cudaEvent_t ev[2];
  cudaEventCreate(&ev[0]); cudaEventCreate(&ev[1]);
  cudaStream_t st = c->slot[0].stream;
  int rc = 0;
  if ((rc = stage[0].reserve(chunk)) == 0) rc = stage[1].reserve(chunk);
  auto stream_out = [&](const void *src, size_t bytes) -> int { // chunk i+1 is copied while chunk i is written
    size_t issued = 0, written = 0;
    int bi = 0, bw = 0;
    size_t len[2] = {0, 0};
    while (written < bytes) {
      while (issued < bytes && issued - written < 2 * chunk) {
        const size_t nb = std::min(chunk, bytes - issued);
        if (cudaMemcpyAsync(stage[bi].p, (const char *)src + issued, nb, cudaMemcpyDeviceToHost, st) != cudaSuccess)
          return set_error(FEMGPU_ECUDA, "D2H of index failed: %s", cudaGetErrorString(cudaGetLastError()));
        cudaEventRecord(ev[bi], st);
        len[bi] = nb;
        issued += nb;
        bi ^= 1;
      }
      if (cudaEventSynchronize(ev[bw]) != cudaSuccess) return set_error(FEMGPU_ECUDA, "D2H of index failed: %s", cudaGetErrorString(cudaGetLastError()));
      if (fwrite(stage[bw].p, 1, len[bw], f) != len[bw]) return set_error(FEMGPU_EINVAL, "short write to index file %s", path);
      written += len[bw];
      bw ^= 1;
    }
    return 0;
  };
  const int k = c->idx_k, step = c->idx_step;
  const size_t n = (size_t)c->n_occ, lut_n = ((size_t)1 << (2 * k)) + 1;
  if (!rc && (fwrite(&k, sizeof(int), 1, f) != 1 || fwrite(&step, sizeof(int), 1, f) != 1)) rc = set_error(FEMGPU_EINVAL, "short write to index file %s", path);
  if (!rc) rc = stream_out(c->lut.p, lut_n * sizeof(uint32_t));
  if (!rc && fwrite(&n, sizeof(size_t), 1, f) != 1) rc = set_error(FEMGPU_EINVAL, "short write to index file %s", path);
  if (!rc) rc = stream_out(c->occ.p, n * sizeof(uint64_t));
  cudaEventDestroy(ev[0]); cudaEventDestroy(ev[1]);
  stage[0].release(); stage[1].release();
  if (fclose(f) != 0 && !rc) rc = set_error(FEMGPU_EINVAL, "short write to index file %s", path);
  return rc;
}
