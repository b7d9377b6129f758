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
  // bit planes (k_planes / k_ref_planes): bit 0 and bit 1 of the base codes, one bit per base, 32 bases per word
  const uint32_t *refp0, *refp1;     // reference (nullptr: not derived -> general form)
  const uint32_t *planes;            // reads: per read-strand [3][plane_words]: bit 0, bit 1, "is N"
  uint32_t plane_words;              // words per plane = ceil(l_max / 32)
  const uint8_t *rs_has_n;           // per read-strand: some base is ambiguous (general form)
};

// ---- bit planes ----------------------------------------------------------------------------------------
// Reads: one thread per plane word (32 bases) of a read-strand with candidates.  The 32 code bytes are read as nine
// aligned words, realigned by funnel shifts; bit b of four codes at once is gathered by a multiply.  rs_has_n is
// zeroed before the launch.
__device__ __forceinline__ uint32_t bits_of_bytes(uint32_t y) { // y: one bit per byte (at bit 0) -> those four bits
  return (y * 0x01020408u) >> 24;
}
__global__ void __launch_bounds__(256) k_planes(const uint8_t *codes, const uint64_t *read_off, const uint32_t *rs_cand_count, uint32_t n_rs,
                                                uint32_t plane_words, uint32_t *planes, uint8_t *rs_has_n) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t rs = t / plane_words;
  const uint32_t w = (uint32_t)(t - rs * plane_words);
  if (rs >= n_rs || rs_cand_count[rs] == 0) return; // only read-strands with candidates are looked at again
  const uint32_t read = (uint32_t)(rs >> 1), dir = (uint32_t)(rs & 1);
  const uint64_t off = read_off[read];
  const uint32_t l = (uint32_t)(read_off[read + 1] - off);
  const uint8_t *p = codes + 2 * off + (uint64_t)dir * l + 32u * w;
  const uint32_t *pa = reinterpret_cast<const uint32_t *>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)3);
  const uint32_t sh = 8u * (uint32_t)(reinterpret_cast<uintptr_t>(p) & 3u);
  uint32_t b0 = 0, b1 = 0, bn = 0;
  if (32u * w < l) { // the codes array is padded: the words past the read's end are readable, their bits masked below
    uint32_t wd[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) wd[k] = __ldg(pa + k);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t x = __funnelshift_r(wd[k], wd[k + 1], sh);
      b0 |= bits_of_bytes(x & 0x01010101u) << (4 * k);
      b1 |= bits_of_bytes((x >> 1) & 0x01010101u) << (4 * k);
      bn |= bits_of_bytes((x >> 2) & 0x01010101u) << (4 * k);
    }
    const uint32_t n = l - 32u * w;
    if (n < 32u) { const uint32_t m = (1u << n) - 1u; b0 &= m; b1 &= m; bn &= m; }
  }
  uint32_t *out = planes + rs * 3ull * plane_words + w;
  out[0] = b0; out[plane_words] = b1; out[2 * plane_words] = bn;
  if (bn) rs_has_n[rs] = 1;
}

// Reference: plane word j (bases 32 j .. 32 j + 31) from the two 2-bit words ref2[2 j], ref2[2 j + 1].
__device__ __forceinline__ uint32_t even_bits16(uint32_t x) { // bits 0, 2, 4, ... of x gathered into the low 16 bits
  x &= 0x55555555u;
  x = (x | (x >> 1)) & 0x33333333u;
  x = (x | (x >> 2)) & 0x0F0F0F0Fu;
  x = (x | (x >> 4)) & 0x00FF00FFu;
  return (x | (x >> 8)) & 0xFFFFu;
}
__global__ void __launch_bounds__(256) k_ref_planes(const uint32_t *ref2, uint64_t n_words, uint32_t *p0, uint32_t *p1) {
  const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_words) return;
  const uint32_t a = ref2[2 * j], b = ref2[2 * j + 1];
  p0[j] = even_bits16(a) | (even_bits16(b) << 16);
  p1[j] = even_bits16(a >> 1) | (even_bits16(b >> 1) << 16);
}

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

// One candidate in the general form: base codes one at a time, five Peq rows (ambiguous bases in the read or the
// reference).  Leaves VP, VN and the bottom-cell error count of the last column, or dead.
template <bool HAS_N>
__device__ __forceinline__ void verify_columns_general(const VerifyArgs &A, uint64_t g0, const uint8_t *text, int l, int e, uint32_t wmask,
                                                       uint32_t &VP_out, uint32_t &VN_out, int &err_out, int &cols_out, bool &dead_out) {
  RefCursor<HAS_N> ref;
  ref.seek(A.ref, g0);

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
  VP_out = VP; VN_out = VN; err_out = err; cols_out = cols; dead_out = dead;
}

// The same columns for a candidate without ambiguous bases on either side, from bit planes: at column i the Peq
// row of read base t, restricted to the band, is the set of j in [0, 2e] with reference[i + j] == t, i.e.
//   ~((R0 >> i) ^ T0) & ~((R1 >> i) ^ T1) & band      R0/R1: planes of the window, T0/T1: bit 0 / bit 1 of t, replicated
// -- two funnel shifts of the window planes, two predicated complements, one LOP3, instead of five conditional ORs, five
// shifts, a byte load and a four-way select per column.  32 columns per round (one word of each read plane); the
// bottom-cell bits of D0 are collected by a funnel shift and counted once per eight columns (the early exit of
// align.c:128-130 only ever turns a candidate into a rejected one, and the count never decreases, so testing it
// later changes nothing but the number of columns computed).
__device__ __forceinline__ void verify_columns_planes(const VerifyArgs &A, uint64_t g0, const uint32_t *q, int l, int e, uint32_t wmask,
                                                      uint32_t &VP_out, uint32_t &VN_out, int &err_out, int &cols_out, bool &dead_out) {
  const uint32_t band = (2u << (2 * e)) - 1u;
  const uint32_t *r0p = A.refp0 + (g0 >> 5), *r1p = A.refp1 + (g0 >> 5);
  const uint32_t s0 = (uint32_t)g0 & 31u;
  uint32_t VP = 0, VN = 0;
  int err = 0, cols = l;
  bool dead = false;
  // window planes of the current round: bits [32 r, 32 r + 64) of the window, refilled word by word
  uint32_t a0 = __funnelshift_r(__ldg(r0p), __ldg(r0p + 1), s0), a1 = __funnelshift_r(__ldg(r1p), __ldg(r1p + 1), s0);
  const int three_e = 3 * e;
#define FEM_COLUMN(ii)                                                                                   \
  {                                                                                                        \
    uint32_t u0 = __funnelshift_r(a0, b0, ii), u1 = __funnelshift_r(a1, b1, ii);                           \
    if (t0 & (1u << (ii))) u0 = ~u0;                                                                       \
    if (t1 & (1u << (ii))) u1 = ~u1;                                                                       \
    uint32_t X = (~(u0 | u1) & band) | VN;                                                                 \
    const uint32_t D0 = ((((VP + (X & VP)) & wmask) ^ VP) | X);                                            \
    const uint32_t HP = VN | (~(VP | D0) & wmask);                                                         \
    X = D0 >> 1;                                                                                           \
    VP = (VP & D0) | (~(X | HP) & wmask);                                                                  \
    VN = X & HP;                                                                                           \
    d0bits = __funnelshift_r(d0bits, D0, 1);                                                               \
  }
  int i0 = 0;
  for (; i0 + 32 <= l; i0 += 32) {
    const int w = i0 >> 5;
    const uint32_t b0 = __funnelshift_r(__ldg(r0p + w + 1), __ldg(r0p + w + 2), s0), b1 = __funnelshift_r(__ldg(r1p + w + 1), __ldg(r1p + w + 2), s0);
    const uint32_t t0 = q[w], t1 = q[A.plane_words + w];
    uint32_t d0bits = 0;
#pragma unroll
    for (int oct = 0; oct < 4; ++oct) {
#pragma unroll
      for (int k = 0; k < 8; ++k) FEM_COLUMN(8 * oct + k)
      // d0bits holds the matches of the round's columns so far in its top 8 (oct + 1) bits
      if (err + 8 * (oct + 1) - __popc(d0bits) > three_e) { dead = true; cols = i0 + 8 * (oct + 1); break; }
    }
    if (dead) break;
    err += 32 - __popc(d0bits);
    a0 = b0; a1 = b1;
  }
  if (!dead && i0 < l) { // the last, partial round: columns one at a time
    const int w = i0 >> 5;
    const uint32_t b0 = __funnelshift_r(__ldg(r0p + w + 1), __ldg(r0p + w + 2), s0), b1 = __funnelshift_r(__ldg(r1p + w + 1), __ldg(r1p + w + 2), s0);
    const uint32_t t0 = q[w], t1 = q[A.plane_words + w];
    for (int ii = 0; i0 + ii < l; ++ii) {
      uint32_t d0bits = 0;
      FEM_COLUMN(ii)
      err += 1 - (int)(d0bits >> 31);
      if (err > three_e) { dead = true; cols = i0 + ii + 1; break; }
    }
  }
#undef FEM_COLUMN
  VP_out = VP; VN_out = VN; err_out = err; cols_out = cols; dead_out = dead;
}

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
  const uint64_t cand = A.cand[ci];
  const int e = A.e;
  const uint64_t g0 = A.ref.rec_base[cand >> 32] + (uint32_t)cand;
  uint32_t VP, VN;
  int err, cols;
  bool dead;
  if (!HAS_N && A.refp0 && !A.rs_has_n[rs])
    verify_columns_planes(A, g0, A.planes + (uint64_t)rs * 3ull * A.plane_words, l, e, wmask, VP, VN, err, cols, dead);
  else
    verify_columns_general<HAS_N>(A, g0, A.codes + 2 * off + (uint64_t)dir * (uint64_t)l, l, e, wmask, VP, VN, err, cols, dead);
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
