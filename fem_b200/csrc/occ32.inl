// occ32.inl -- the 32-bit copy of the occurrence table for k_cand_pair32 (included by femgpu.cu)

__global__ void __launch_bounds__(256) k_occ32(const uint64_t *occ, uint64_t n, const uint32_t *gbase, uint32_t n_rec, uint32_t *occ32) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t o = occ[i];
    const uint32_t rec = (uint32_t)(o >> 32);
    occ32[i] = (rec < n_rec ? __ldg(gbase + rec) : 0u) + (uint32_t)o; // an index of another reference maps nothing sensible either way
  }
}

// Uploads and builds only invalidate the copy; it is derived when a batch first asks for the 32-bit
// layout ("pair_layout" = 32).  Leaves c->occ32_ok = false (the 64-bit kernel is used) when the padded
// coordinate does not fit in 32 bits.
static int derive_occ32(femgpu_ctx *c) {
  c->occ32_ok = false;
  c->occ32_tried = false;
  return 0;
}

static int derive_occ32_now(femgpu_ctx *c) {
  c->occ32_ok = false;
  c->occ32_tried = true;
  if (!c->n_rec || !c->idx_k || !c->n_occ) return 0;
  std::vector<uint32_t> lens(c->n_rec);
  CU(cudaMemcpy(lens.data(), c->rec_len.p, (size_t)c->n_rec * 4, cudaMemcpyDeviceToHost));
  const uint64_t align = 1ull << PAIR32_ALIGN_LOG;
  std::vector<uint32_t> gbase(c->n_rec + 1);
  std::vector<uint32_t> flags((1u << (32 - PAIR32_ALIGN_LOG)) / 32, 0u);
  uint64_t g = align; // the first record starts one block in: its pad is the end of block 0
  for (uint32_t r = 0; r <= c->n_rec; ++r) {
    if (g >= (1ull << 32) - 2 * align) return 0; // does not fit: keep the 64-bit path
    gbase[r] = (uint32_t)g;
    if (r < c->n_rec) {
      const uint32_t blk = (uint32_t)(g >> PAIR32_ALIGN_LOG) - 1; // the block that ends at this record's start
      flags[blk >> 5] |= 1u << (blk & 31u);
      g = (g + lens[r] + PAIR32_PAD + align - 1) & ~(align - 1);
    }
  }
  TRY(c->occ32.reserve(c->n_occ * 4 + 256));
  TRY(c->gbase.reserve((size_t)(c->n_rec + 1) * 4));
  TRY(c->padflag.reserve(flags.size() * 4));
  cudaStream_t st = c->slot[0].stream;
  CU(cudaMemcpyAsync(c->gbase.p, gbase.data(), (size_t)(c->n_rec + 1) * 4, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(c->padflag.p, flags.data(), flags.size() * 4, cudaMemcpyHostToDevice, st));
  CU(cudaMemsetAsync((char *)c->occ32.p + c->n_occ * 4, 0, 256, st));
  const unsigned grid = (unsigned)std::min<uint64_t>((c->n_occ + 255) / 256, (uint64_t)c->sm_count * 32);
  k_occ32<<<grid, 256, 0, st>>>(c->occ.as<uint64_t>(), c->n_occ, c->gbase.as<uint32_t>(), c->n_rec, c->occ32.as<uint32_t>());
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(st)); // the host vectors above must outlive the copies
  c->occ32_ok = true;
  return 0;
}
