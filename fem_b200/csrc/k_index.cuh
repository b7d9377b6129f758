// k_index.cuh -- index construction on the device (the arrays of src/index.c:57-98, byte-identical).
//
// The reference collects one (hash, location) pair per sampled q-gram (positions 0, step, 2*step,
// ... of every record while the q-gram lies inside it, src/index.c:65), radix-sorts the pairs by
// hash and then sorts every bucket by location.  Locations are unique, so the result is fully
// determined by the SET of pairs: lut = prefix sums of the bucket sizes, occ = every bucket's
// locations ascending.  Here:
//   k_index_count   one thread per sample: hash from the 2-bit packed reference (ambiguous bases
//                   are packed as A, which is what the reference's hash does, src/utils.h:91-92),
//                   atomicAdd into lut[1 + hash]                       (L2-resident histogram)
//   k_index_scan_*  exclusive prefix sum in place: lut[1 + h] = start of bucket h
//   k_index_fill    same enumeration: occ[atomicAdd(&lut[1 + hash], 1)] = location; afterwards
//                   lut[1 + h] = end of bucket h = start of bucket h + 1, i.e. the final table
//   k_index_sort_*  every bucket sorted ascending: a warp per bucket (rank sort up to 32 entries,
//                   shared-memory bitonic network up to 1024), a block per larger bucket
//                   (shared memory up to 8192 entries, global-memory bitonic network beyond)
#pragma once
#include "common.cuh"
#include "k_cand.cuh" // warp_bitonic_sort

#define INDEX_MAX_REC_SMEM 1024 // records whose sample prefix is staged in shared memory
#define INDEX_WARP_CAP 1024u    // entries a warp sorts in shared memory
#define INDEX_BLOCK_CAP 8192u   // entries a block sorts in shared memory

struct IndexArgs {
  const uint32_t *ref2;
  const uint64_t *rec_base;    // [n_rec + 1] record starts in the packed reference
  const uint64_t *sample_base; // [n_rec + 1] prefix sums of the records' sample counts
  uint32_t n_rec;
  uint64_t n_samples;
  int k, step;
  uint32_t *lut; // [4^k + 1]
  uint64_t *occ;
};

// q-gram hash at packed position g: first base in the most significant position (src/utils.h:83-99)
__device__ __forceinline__ uint32_t packed_hash(const uint32_t *ref2, uint64_t g, int k) {
  const uint64_t w = g >> 4;
  const uint32_t sh = 2u * (uint32_t)(g & 15);
  const uint64_t lo = ref2[w], hi = ref2[w + 1];
  uint32_t x = (uint32_t)(((hi << 32) | lo) >> sh); // base j of the q-gram at bits 2j, 2j+1 (k <= 15 < 16)
  x = __brev(x);                                    // base j now at bits 31-2j, 30-2j with its two bits swapped
  x = ((x & 0xAAAAAAAAu) >> 1) | ((x & 0x55555555u) << 1);
  return x >> (32 - 2 * k);
}

// record of sample t: last r with sample_base[r] <= t (records without samples are skipped by the search)
__device__ __forceinline__ uint32_t sample_record(const uint64_t *sb, uint32_t n_rec, uint64_t t) {
  uint32_t lo = 0, hi = n_rec; // invariant: sb[lo] <= t < sb[hi]
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (sb[mid] <= t) lo = mid; else hi = mid;
  }
  return lo;
}

template <bool FILL>
__global__ void __launch_bounds__(256) k_index_pass(IndexArgs A) {
  __shared__ uint64_t s_sb[INDEX_MAX_REC_SMEM + 1];
  const bool staged = A.n_rec <= INDEX_MAX_REC_SMEM;
  if (staged) {
    for (uint32_t i = threadIdx.x; i <= A.n_rec; i += blockDim.x) s_sb[i] = A.sample_base[i];
    __syncthreads();
  }
  const uint64_t *sb = staged ? s_sb : A.sample_base;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < A.n_samples; t += stride) {
    const uint32_t r = sample_record(sb, A.n_rec, t);
    const uint64_t pos = (t - sb[r]) * (uint64_t)A.step;
    const uint32_t h = packed_hash(A.ref2, A.rec_base[r] + pos, A.k);
    if (FILL) {
      const uint32_t slot = atomicAdd(&A.lut[1 + h], 1u);
      A.occ[slot] = ((uint64_t)r << 32) | (uint64_t)(uint32_t)pos;
    } else {
      atomicAdd(&A.lut[1 + h], 1u);
    }
  }
}

// ---- exclusive scan of v[0..n) in place (uint32 sums; the total fits by construction) ----------
#define INDEX_SCAN_ITEMS 8
#define INDEX_SCAN_TILE (256 * INDEX_SCAN_ITEMS)

__device__ __forceinline__ uint32_t block_exclusive_scan_256(uint32_t v, uint32_t *total, uint32_t *s_warp) {
  const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
  uint32_t incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
    if ((int)lane >= d) incl += o;
  }
  if (lane == 31) s_warp[w] = incl;
  __syncthreads();
  if (w == 0) {
    uint32_t x = lane < 8 ? s_warp[lane] : 0u, xi = x;
#pragma unroll
    for (int d = 1; d < 8; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, xi, d);
      if ((int)lane >= d) xi += o;
    }
    if (lane < 8) s_warp[lane] = xi - x;
    if (lane == 7) s_warp[8] = xi;
  }
  __syncthreads();
  *total = s_warp[8];
  return s_warp[w] + incl - v;
}

__global__ void __launch_bounds__(256) k_index_scan_reduce(const uint32_t *v, uint64_t n, uint32_t *tile_sum) {
  __shared__ uint32_t s_warp[9];
  const uint64_t base = (uint64_t)blockIdx.x * INDEX_SCAN_TILE + (uint64_t)threadIdx.x * INDEX_SCAN_ITEMS;
  uint32_t sum = 0;
#pragma unroll
  for (int i = 0; i < INDEX_SCAN_ITEMS; ++i) sum += base + i < n ? v[base + i] : 0u;
  uint32_t total;
  block_exclusive_scan_256(sum, &total, s_warp);
  if (threadIdx.x == 0) tile_sum[blockIdx.x] = total;
}

// one block: exclusive scan of the tile sums
__global__ void __launch_bounds__(256) k_index_scan_tiles(uint32_t *tile_sum, uint32_t n_tiles) {
  __shared__ uint32_t s_warp[9];
  uint32_t carry = 0;
  for (uint32_t base = 0; base < n_tiles; base += 256) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < n_tiles ? tile_sum[i] : 0u;
    uint32_t total;
    const uint32_t ex = block_exclusive_scan_256(v, &total, s_warp);
    if (i < n_tiles) tile_sum[i] = carry + ex;
    carry += total;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256) k_index_scan_apply(uint32_t *v, uint64_t n, const uint32_t *tile_sum) {
  __shared__ uint32_t s_warp[9];
  const uint64_t base = (uint64_t)blockIdx.x * INDEX_SCAN_TILE + (uint64_t)threadIdx.x * INDEX_SCAN_ITEMS;
  uint32_t x[INDEX_SCAN_ITEMS], sum = 0;
#pragma unroll
  for (int i = 0; i < INDEX_SCAN_ITEMS; ++i) { x[i] = base + i < n ? v[base + i] : 0u; sum += x[i]; }
  uint32_t total;
  uint32_t run = tile_sum[blockIdx.x] + block_exclusive_scan_256(sum, &total, s_warp);
#pragma unroll
  for (int i = 0; i < INDEX_SCAN_ITEMS; ++i) {
    if (base + i < n) v[base + i] = run;
    run += x[i];
  }
}

// ---- bucket sorts ----------------------------------------------------------------------------------
// A warp per bucket.  Buckets above INDEX_WARP_CAP entries are appended to big_list.
__global__ void __launch_bounds__(128) k_index_sort_warp(const uint32_t *lut, uint64_t n_buckets, uint64_t *occ,
                                                         uint32_t *big_list, uint32_t *big_count) {
  extern __shared__ uint64_t smem_index[];
  const uint32_t lane = lane_id(), wib = threadIdx.x >> 5;
  uint64_t *buf = smem_index + (size_t)wib * INDEX_WARP_CAP;
  const uint64_t n_warps = (uint64_t)gridDim.x * (blockDim.x >> 5);
  for (uint64_t h = (uint64_t)blockIdx.x * (blockDim.x >> 5) + wib; h < n_buckets; h += n_warps) {
    const uint32_t b = lut[h], n = lut[h + 1] - b;
    if (n < 2) continue;
    uint64_t *o = occ + b;
    if (n <= 32) { // rank by counting through shuffles (values are distinct)
      const uint64_t x = lane < n ? o[lane] : FEM_SENTINEL;
      uint32_t rank = 0;
      for (uint32_t j = 0; j < n; ++j) rank += __shfl_sync(0xffffffffu, x, j) < x;
      if (lane < n) o[rank] = x;
    } else if (n <= INDEX_WARP_CAP) {
      const uint32_t n_pad = pow2_ceil_u32(n);
      for (uint32_t t = lane; t < n_pad; t += 32) buf[t] = t < n ? o[t] : FEM_SENTINEL;
      __syncwarp();
      warp_bitonic_sort(buf, n_pad, lane);
      for (uint32_t t = lane; t < n; t += 32) o[t] = buf[t];
      __syncwarp();
    } else if (lane == 0) {
      big_list[atomicAdd(big_count, 1u)] = (uint32_t)h;
    }
  }
}

__device__ __forceinline__ void block_bitonic_sort(uint64_t *buf, uint64_t n_pad) {
  for (uint64_t k = 2; k <= n_pad; k <<= 1) {
    for (uint64_t j = k >> 1; j > 0; j >>= 1) {
      for (uint64_t t = threadIdx.x; t < (n_pad >> 1); t += blockDim.x) {
        const uint64_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        const uint64_t p = i | j;
        const uint64_t x = buf[i], y = buf[p];
        const bool up = (i & k) == 0;
        if ((x > y) == up) { buf[i] = y; buf[p] = x; }
      }
      __syncthreads();
    }
  }
}

// A block per big bucket.  Buckets beyond INDEX_BLOCK_CAP are sorted in a global scratch slice of
// pow2_ceil(n) entries at scratch_off[i] (planned on the host).
__global__ void __launch_bounds__(1024) k_index_sort_block(const uint32_t *lut, uint64_t *occ, const uint32_t *big_list,
                                                           uint32_t n_big, uint64_t *scratch, const uint64_t *scratch_off) {
  extern __shared__ uint64_t smem_index[];
  for (uint32_t w = blockIdx.x; w < n_big; w += gridDim.x) {
    const uint32_t h = big_list[w];
    const uint32_t b = lut[h], n = lut[h + 1] - b;
    uint64_t *o = occ + b;
    uint64_t n_pad = 1;
    while (n_pad < n) n_pad <<= 1;
    uint64_t *buf = n <= INDEX_BLOCK_CAP ? smem_index : scratch + scratch_off[w];
    for (uint64_t t = threadIdx.x; t < n_pad; t += blockDim.x) buf[t] = t < n ? o[t] : FEM_SENTINEL;
    __syncthreads();
    block_bitonic_sort(buf, n_pad);
    for (uint64_t t = threadIdx.x; t < n; t += blockDim.x) o[t] = buf[t];
    __syncthreads();
  }
}
