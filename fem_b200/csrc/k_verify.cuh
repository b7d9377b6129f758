// k_verify.cuh -- K4: banded Myers bit-vector verification, one thread per candidate, and
// K5: ordered compaction of the accepted mappings.
//
// Restates src/align.c:4-51 (verify_candidates), :102-147 (32-bit scalar kernel) and :149-277
// (8 x 16-bit SSE lanes).  The reference runs candidates j < 8*floor(n/8) of a read-strand in
// 16-bit lanes and the remaining n%8 in 32-bit words; bits above the band are never masked and
// `D0 >> 1` pulls them back, so the two widths are not interchangeable.  Each thread
// therefore computes in the width class of ITS candidate: `wmask` is 0xFFFF or 0xFFFFFFFF and is
// applied after the add and the two complements, which is all that distinguishes the lane widths.
// The per-lane early exit (> 3e errors) is result-equivalent to the reference's all-lanes exit:
// the bottom-cell count never decreases and the final walk can lower it by at most 2e.
#pragma once
#include "common.cuh"

struct VerifyArgs {
  DevRef ref;
  const uint8_t *codes;        // strand-oriented read codes (k_seed)
  const uint64_t *read_off;
  const uint64_t *cand;
  const uint32_t *cand_rs;
  const uint64_t *rs_cand_first;
  const uint32_t *rs_cand_count;
  uint64_t n_cand;
  int e;
  uint32_t *result;            // out per candidate: dist<<16 | end_offset, or 0xFFFFFFFF
  uint32_t *rs_map_count;      // out per read-strand (atomic)
  unsigned long long *cols_executed; // out: Myers columns actually computed (roofline accounting)
};

template <bool HAS_N>
__device__ __forceinline__ uint32_t ref_code_at(const DevRef &R, uint64_t g, uint32_t &w2, uint32_t &wn, bool reload) {
  if (reload || (g & 15) == 0) w2 = __ldg(R.ref2 + (g >> 4));
  uint32_t c = (w2 >> (2 * (g & 15))) & 3u;
  if (HAS_N) {
    if (reload || (g & 31) == 0) wn = __ldg(R.refn + (g >> 5));
    if ((wn >> (g & 31)) & 1u) c = 4u;
  }
  return c;
}

// The same codes for a forward walk over the reference: the current words are kept shifted down, so a base
// costs an AND, a shift and a counter (the random-access form above spends more on 64-bit index arithmetic
// than on the base).  Reads one word past the walk's last base at most (the arrays are padded).
template <bool HAS_N>
struct RefCursor {
  const uint32_t *p2, *pn;
  uint32_t w2, wn;
  int left2, leftn;
  __device__ __forceinline__ void seek(const DevRef &R, uint64_t g) {
    const uint32_t s2 = (uint32_t)g & 15u;
    p2 = R.ref2 + (g >> 4);
    w2 = __ldg(p2) >> (2u * s2);
    left2 = 16 - (int)s2;
    if (HAS_N) {
      const uint32_t sn = (uint32_t)g & 31u;
      pn = R.refn + (g >> 5);
      wn = __ldg(pn) >> sn;
      leftn = 32 - (int)sn;
    }
  }
  __device__ __forceinline__ uint32_t next() {
    uint32_t c = w2 & 3u;
    w2 >>= 2;
    if (--left2 == 0) { w2 = __ldg(++p2); left2 = 16; }
    if (HAS_N) {
      if (wn & 1u) c = 4u;
      wn >>= 1;
      if (--leftn == 0) { wn = __ldg(++pn); leftn = 32; }
    }
    return c;
  }
};

template <bool HAS_N>
__global__ void __launch_bounds__(256) k_verify(VerifyArgs A) {
  const uint64_t ci = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (ci >= A.n_cand) return; // grid tail only: whole warps or the last partial one
  const uint32_t rs = A.cand_rs[ci];
  const uint32_t j = (uint32_t)(ci - A.rs_cand_first[rs]), n = A.rs_cand_count[rs];
  const uint32_t wmask = j < (n & ~7u) ? 0xFFFFu : 0xFFFFFFFFu; // align.c:12-13
  const uint32_t read = rs >> 1, dir = rs & 1;
  const uint64_t off = A.read_off[read];
  const int l = (int)(A.read_off[read + 1] - off);
  const uint8_t *text = A.codes + 2 * off + (uint64_t)dir * (uint64_t)l;
  const uint64_t cand = A.cand[ci];
  const int e = A.e;
  RefCursor<HAS_N> ref;
  ref.seek(A.ref, A.ref.rec_base[cand >> 32] + (uint32_t)cand);

  uint32_t p0 = 0, p1 = 0, p2 = 0, p3 = 0, p4 = 0; // Peq rows
  for (int i = 0; i < 2 * e; ++i) {                // align.c:104-107
    const uint32_t c = ref.next();
    const uint32_t bit = 1u << i;
    p0 |= c == 0 ? bit : 0u; p1 |= c == 1 ? bit : 0u; p2 |= c == 2 ? bit : 0u;
    p3 |= c == 3 ? bit : 0u; p4 |= c == 4 ? bit : 0u;
  }
  const uint32_t hi = 1u << (2 * e);
  uint32_t VP = 0, VN = 0;
  int err = 0, cols = l;
  bool dead = false;
  for (int i = 0; i < l; ++i) {                    // align.c:117-134
    const uint32_t c = ref.next();
    p0 |= c == 0 ? hi : 0u; p1 |= c == 1 ? hi : 0u; p2 |= c == 2 ? hi : 0u;
    p3 |= c == 3 ? hi : 0u; p4 |= c == 4 ? hi : 0u;
    const uint32_t t = text[i];
    const uint32_t eq = t == 0 ? p0 : t == 1 ? p1 : t == 2 ? p2 : t == 3 ? p3 : p4;
    uint32_t X = eq | VN;
    const uint32_t D0 = ((((VP + (X & VP)) & wmask) ^ VP) | X);
    const uint32_t HN = VP & D0;
    const uint32_t HP = VN | (~(VP | D0) & wmask);
    X = D0 >> 1;
    VN = X & HP;
    VP = HN | (~(X | HP) & wmask);
    err += 1 - (int)(D0 & 1u);
    if (err > 3 * e) { dead = true; cols = i + 1; break; }
    p0 >>= 1; p1 >>= 1; p2 >>= 1; p3 >>= 1; p4 >>= 1;
  }
  uint32_t out = 0xFFFFFFFFu;
  if (!dead) {
    int best = err, end_off = l - 1;               // align.c:135-146
    for (int i = 0; i < 2 * e; ++i) {
      err += (int)((VP >> i) & 1u) - (int)((VN >> i) & 1u);
      if (err < best) { best = err; end_off = l + i; }
    }
    if (best <= e) {
      out = ((uint32_t)best << 16) | (uint32_t)end_off;
      atomicAdd(A.rs_map_count + rs, 1u);
    }
  }
  A.result[ci] = out;
  const uint32_t live = __activemask();
  const uint32_t warp_cols = __reduce_add_sync(live, (uint32_t)cols);
  if (lane_id() == (uint32_t)(__ffs(live) - 1)) atomicAdd(A.cols_executed, (unsigned long long)warp_cols);
}

// ---- exclusive scan of uint32 counts into uint64 offsets (3 small kernels) ----------------------
#define SCAN_BLOCK 1024
__global__ void __launch_bounds__(SCAN_BLOCK) k_scan_reduce(const uint32_t *in, uint64_t n, uint64_t *block_sums) {
  __shared__ uint64_t sm[32];
  const uint64_t i = (uint64_t)blockIdx.x * SCAN_BLOCK + threadIdx.x;
  uint64_t v = i < n ? in[i] : 0;
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    v = sm[threadIdx.x];
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = v;
  }
}
__global__ void k_scan_blocks(uint64_t *block_sums, uint32_t n_blocks, uint64_t *total) {
  // single thread block, sequential over chunks: n_blocks is at most a few thousand
  __shared__ uint64_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < n_blocks; base += 32) {
    if (threadIdx.x < 32) {
      const uint32_t i = base + threadIdx.x;
      const uint64_t v = i < n_blocks ? block_sums[i] : 0;
      uint64_t incl = v;
      for (int d = 1; d < 32; d <<= 1) {
        const uint64_t o = __shfl_up_sync(0xffffffffu, incl, d);
        if ((int)threadIdx.x >= d) incl += o;
      }
      if (i < n_blocks) block_sums[i] = carry + incl - v;
      const uint64_t tot = __shfl_sync(0xffffffffu, incl, 31);
      __syncwarp();
      if (threadIdx.x == 0) carry += tot;
      __syncwarp();
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) *total = carry;
}
__global__ void __launch_bounds__(SCAN_BLOCK) k_scan_apply(const uint32_t *in, uint64_t n, const uint64_t *block_sums,
                                                           uint64_t *out /* n+1 */, const uint64_t *total) {
  __shared__ uint64_t sm[32];
  const uint64_t i = (uint64_t)blockIdx.x * SCAN_BLOCK + threadIdx.x;
  const uint64_t v = i < n ? in[i] : 0;
  uint64_t incl = v;
  const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int d = 1; d < 32; d <<= 1) {
    const uint64_t o = __shfl_up_sync(0xffffffffu, incl, d);
    if ((int)lane >= d) incl += o;
  }
  if (lane == 31) sm[w] = incl;
  __syncthreads();
  if (w == 0) {
    uint64_t s = sm[lane], si = s;
    for (int d = 1; d < 32; d <<= 1) {
      const uint64_t o = __shfl_up_sync(0xffffffffu, si, d);
      if ((int)lane >= d) si += o;
    }
    sm[lane] = si - s;
  }
  __syncthreads();
  if (i < n) out[i] = block_sums[blockIdx.x] + sm[w] + incl - v;
  if (i == 0) out[n] = *total;
}

// ---- K5: ordered compaction, one warp per read-strand ---------------------------------------------
struct CompactArgs {
  const uint64_t *cand;
  const uint32_t *result;
  const uint64_t *rs_cand_first;
  const uint32_t *rs_cand_count;
  const uint64_t *rs_map_first; // exclusive scan of rs_map_count
  const uint32_t *rs_map_count;
  uint32_t n_rs;
  ulonglong2 *maps;             // 16-byte femgpu_mapping records
  uint32_t *map_rs;             // owning read-strand of each record (k_trace)
};

// Eight lanes per read-strand, four read-strands per warp: most read-strands hold one or two candidates (a
// warp each was 2 M nearly empty warps per 1 M reads, 384 us), the repeat-rich ones thousands (eight at a time).
#define COMPACT_LPR 8u
__global__ void __launch_bounds__(256) k_compact(CompactArgs A) {
  const uint32_t lane = lane_id(), sub = lane & (COMPACT_LPR - 1), grp = lane / COMPACT_LPR;
  const uint64_t rs = ((uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * (32u / COMPACT_LPR) + grp;
  const bool live = rs < A.n_rs && A.rs_map_count[rs] != 0;
  uint64_t cf = 0, mf = 0;
  uint32_t n = 0;
  if (live) { cf = A.rs_cand_first[rs]; mf = A.rs_map_first[rs]; n = A.rs_cand_count[rs]; }
  const uint32_t n_max = __reduce_max_sync(0xffffffffu, n);
  const uint32_t below = (1u << sub) - 1u;
  uint32_t w = 0;
  for (uint32_t base = 0; base < n_max; base += COMPACT_LPR) {
    const uint32_t t = base + sub;
    uint32_t r = 0xFFFFFFFFu;
    if (t < n) r = A.result[cf + t];
    const uint32_t m = (__ballot_sync(0xffffffffu, r != 0xFFFFFFFFu) >> (COMPACT_LPR * grp)) & ((1u << COMPACT_LPR) - 1u);
    if (r != 0xFFFFFFFFu) {
      ulonglong2 rec;
      rec.x = A.cand[cf + t];
      // {int32 end_offset, uint8 direction, uint8 edit_distance, 2 reserved} little endian
      rec.y = (uint64_t)(r & 0xFFFFu) | ((uint64_t)(rs & 1) << 32) | ((uint64_t)(r >> 16) << 40);
      const uint64_t o = mf + w + __popc(m & below);
      A.maps[o] = rec;
      A.map_rs[o] = (uint32_t)rs;
    }
    w += __popc(m);
  }
}
