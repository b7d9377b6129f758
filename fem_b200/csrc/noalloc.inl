// noalloc.inl -- femgpu_lookup_table_noalloc: the lookup-table prefix sum of the index build for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
// (included by femgpu.cu after index_build.inl: one translation unit, so the array below and the host code that
// asks for its address belong to the same module)
//
// No counterpart in the reference.  femgpu_build_index does the same
// arithmetic (k_index_scan_reduce / _tiles / _apply over the bucket counts) with the tile sums in a cudaMalloc'ed
// buffer; here they live in a file-scope device array of a fixed, modest size that every GPU which loads the
// library holds until exit.

#define FEMGPU_NOALLOC_MAX_TILES 8192u // x INDEX_SCAN_TILE (2048) buckets = 4^12: the reference's default window
static_assert((uint64_t)FEMGPU_NOALLOC_MAX_TILES * INDEX_SCAN_TILE == FEMGPU_NOALLOC_MAX_BUCKETS, "header and kernel limits differ");

// 32 KB + one padding line per device.  In HOST code this name is a host-side stand-in, not a device address: the
// kernels get the address cudaGetSymbolAddress returns (noalloc_scratch_of below).
static __device__ __align__(256) uint32_t g_noalloc_tile_sum[FEMGPU_NOALLOC_MAX_TILES + 64];

// k_index_scan_apply writing a separate table: lut[0] = 0, lut[h + 1] = counts[0] + ... + counts[h]
__global__ void __launch_bounds__(256) k_index_scan_apply_out(const uint32_t *counts, uint64_t n, const uint32_t *tile_sum,
                                                              uint32_t *lut) {
  __shared__ uint32_t s_warp[9];
  const uint64_t base = (uint64_t)blockIdx.x * INDEX_SCAN_TILE + (uint64_t)threadIdx.x * INDEX_SCAN_ITEMS;
  uint32_t x[INDEX_SCAN_ITEMS], sum = 0;
#pragma unroll
  for (int i = 0; i < INDEX_SCAN_ITEMS; ++i) { x[i] = base + i < n ? counts[base + i] : 0u; sum += x[i]; }
  uint32_t total;
  uint32_t run = tile_sum[blockIdx.x] + block_exclusive_scan_256(sum, &total, s_warp);
  if (blockIdx.x == 0 && threadIdx.x == 0) lut[0] = 0;
#pragma unroll
  for (int i = 0; i < INDEX_SCAN_ITEMS; ++i) {
    run += x[i];
    if (base + i < n) lut[base + i + 1] = run;
  }
}

#define FEMGPU_NOALLOC_MAX_DEVICES 64
static std::mutex g_noalloc_mutex;
static uint32_t *g_noalloc_scratch[FEMGPU_NOALLOC_MAX_DEVICES]; // device address of g_noalloc_tile_sum, per device

// set-up, once per device, with that device current
static int noalloc_scratch_of(int device, uint32_t **out) {
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_scratch[device]) {
    uint32_t *p = nullptr;
    CU(cudaGetSymbolAddress((void **)&p, g_noalloc_tile_sum));
    if (!p) return set_error(FEMGPU_ECUDA, "cudaGetSymbolAddress returned a null address on device %d", device);
    g_noalloc_scratch[device] = p;
  }
  *out = g_noalloc_scratch[device];
  return 0;
}

extern "C" int femgpu_lookup_table_noalloc(int device, const uint32_t *d_counts, uint64_t n_buckets, uint32_t *d_lut,
                                           void *stream) {
  if (!d_counts || !d_lut) return set_error(FEMGPU_EINVAL, "null argument");
  if (device < 0 || device >= FEMGPU_NOALLOC_MAX_DEVICES) return set_error(FEMGPU_EINVAL, "device %d out of 0..%d", device, FEMGPU_NOALLOC_MAX_DEVICES - 1);
  if (n_buckets == 0 || n_buckets > FEMGPU_NOALLOC_MAX_BUCKETS)
    return set_error(FEMGPU_EINVAL,
                     "femgpu_lookup_table_noalloc: %llu buckets, the static scratch holds the tile sums of 1..%llu (4^12); "
                     "use femgpu_build_index for larger tables",
                     (unsigned long long)n_buckets, (unsigned long long)FEMGPU_NOALLOC_MAX_BUCKETS);
  CU(cudaSetDevice(device));
  uint32_t *tile_sum = nullptr;
  TRY(noalloc_scratch_of(device, &tile_sum));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const uint32_t n_tiles = (uint32_t)((n_buckets + INDEX_SCAN_TILE - 1) / INDEX_SCAN_TILE);
  // tile_sum[0 .. n_tiles) is written in full by the first kernel before anything reads it: nothing of the previous
  // call's contents is used, and nothing has to start at zero
  k_index_scan_reduce<<<n_tiles, 256, 0, st>>>(d_counts, n_buckets, tile_sum);
  k_index_scan_tiles<<<1, 256, 0, st>>>(tile_sum, n_tiles);
  k_index_scan_apply_out<<<n_tiles, 256, 0, st>>>(d_counts, n_buckets, tile_sum, d_lut);
  CU(cudaGetLastError());
  return 0;
}
