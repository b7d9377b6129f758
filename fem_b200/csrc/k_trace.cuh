// k_trace.cuh -- K6: alignment traceback of the accepted mappings on the device, one thread per
// mapping: start position and CIGAR of src/align.c:279-499 (generate_alignment).
//
// The reference recomputes the band in 32-bit words keeping D0 and HP of every column
// (src/align.c:303-338), then walks back from the end cell consuming exactly `edit_distance`
// errors (:340-440) and emits the run-length operations reversed (:441-498).  The same walk is
// done here per thread with the two per-column words in local memory, followed by the MD text of
// src/align.c:501-544 (generate_MD_tag); the host (sam.c) keeps only the text formatting.  The reference compares RAW characters in this function
// while the band uses base codes; codes and characters agree exactly when every
// character is one of "ACGTN", so mappings of reads (or references) with any other character are
// flagged FEMGPU_ALN_HOST and the host runs its own femhost_alignment on them.
#pragma once
#include "common.cuh"
#include "k_verify.cuh" // RefCursor

#define ALN_MAX_OPS 18
#define ALN_HOST 0xFFu // n_cigar value: not computed on the device

struct TraceArgs {
  DevRef ref;
  const uint8_t *codes;        // strand-oriented read codes (k_seed)
  const uint64_t *read_off;
  const uint8_t *read_flags;   // bit 0: read holds a character outside "ACGTN"
  const ulonglong2 *maps;      // compacted 16-byte mapping records (k_compact)
  const uint32_t *map_rs;      // owning read-strand of each record
  uint64_t n_maps;
  int e;
  int ref_canonical;
  uint32_t *aln;               // out: 22 words per mapping (femgpu_alignment, 88 bytes)
  // bit planes (k_ref_planes / k_planes, see k_verify.cuh); refp0 == nullptr: not available
  const uint32_t *refp0, *refp1, *planes;
  uint32_t plane_words;
  const uint8_t *rs_has_n;
};

template <bool HAS_N>
__device__ __forceinline__ uint32_t ref_code_fetch(const DevRef &R, uint64_t g) {
  uint32_t c = (__ldg(R.ref2 + (g >> 4)) >> (2 * (g & 15))) & 3u;
  if (HAS_N && ((__ldg(R.refn + (g >> 5)) >> (g & 31)) & 1u)) c = 4u;
  return c;
}

#define ALN_WORDS 22
#define ALN_MD_MAX 47

__device__ __forceinline__ void aln_store(uint32_t *dst, int start, uint32_t n_cigar, const uint16_t *cig, uint32_t md_len = ALN_HOST) {
  dst[0] = ((uint32_t)(uint16_t)(int16_t)start) | (n_cigar << 16) | (md_len << 24);
#pragma unroll
  for (int i = 0; i < ALN_MAX_OPS / 2; ++i) dst[1 + i] = (uint32_t)cig[2 * i] | ((uint32_t)cig[2 * i + 1] << 16);
}

// decimal digits of v appended at md[o..]; returns the new length (may exceed ALN_MD_MAX: caller checks)
__device__ __forceinline__ int md_put_uint(char *md, int o, uint32_t v) {
  char tmp[10];
  int n = 0;
  do { tmp[n++] = (char)('0' + v % 10u); v /= 10u; } while (v);
  while (n) { if (o < 48) md[o] = tmp[n - 1]; ++o; --n; }
  return o;
}

// generate_MD_tag (src/align.c:501-544) for a CIGAR in femgpu_alignment encoding; the reference starts at
// gref, the read codes at text.  Characters are those of a canonical ("ACGTN") reference.  Writes the
// twelve MD words of the record and returns the length, or ALN_HOST when the text does not fit.
template <bool HAS_N>
__device__ uint32_t md_build_store(const DevRef &R, uint64_t gref, const uint8_t *text, const uint16_t *cig, int nc, uint32_t *dst) {
  char md[48];
  int o = 0, tp = 0;
  uint32_t run = 0;
  RefCursor<HAS_N> ref; // the reference is walked forward
  ref.seek(R, gref);
  for (int ci = 0; ci < nc; ++ci) {
    const uint32_t op = cig[ci] & 3u, len = cig[ci] >> 2;
    if (op == 0) {
      for (uint32_t j = 0; j < len; ++j, ++tp) {
        const uint32_t c = ref.next();
        if (c == text[tp]) { ++run; continue; }
        if (run) { o = md_put_uint(md, o, run); run = 0; }
        if (o < 48) md[o] = "ACGTN"[c];
        ++o;
      }
    } else if (op == 1) {
      tp += (int)len;
    } else {
      if (run) { o = md_put_uint(md, o, run); run = 0; }
      if (o < 48) md[o] = '^';
      ++o;
      for (uint32_t j = 0; j < len; ++j) {
        const uint32_t c = ref.next();
        if (o < 48) md[o] = "ACGTN"[c];
        ++o;
      }
    }
    if (o > ALN_MD_MAX) return ALN_HOST;
  }
  if (run) o = md_put_uint(md, o, run);
  if (o > ALN_MD_MAX) return ALN_HOST;
  for (int i = o; i < 48; ++i) md[i] = 0;
#pragma unroll
  for (int w = 0; w < 12; ++w)
    dst[w] = (uint32_t)(uint8_t)md[4 * w] | ((uint32_t)(uint8_t)md[4 * w + 1] << 8) | ((uint32_t)(uint8_t)md[4 * w + 2] << 16) |
             ((uint32_t)(uint8_t)md[4 * w + 3] << 24);
  return (uint32_t)o;
}

// ---- reads and references without N: bit planes, 32 bases per word --------------------------------------------
// 32 bits of a plane from bit position g on
__device__ __forceinline__ uint32_t plane_word_at(const uint32_t *p, uint64_t g) {
  const uint32_t *w = p + (g >> 5);
  return __funnelshift_r(__ldg(w), __ldg(w + 1), (uint32_t)g & 31u);
}

// md_build_store for an N-free reference and read: the mismatches of a match run are the set bits of
// (R0 ^ T0) | (R1 ^ T1), 32 bases at a time, so the cost follows the number of differences, not the read length
__device__ uint32_t md_build_store_planes(const uint32_t *rp0, const uint32_t *rp1, uint64_t gref, const uint32_t *q0, const uint32_t *q1,
                                          const uint16_t *cig, int nc, uint32_t *dst) {
  char md[48];
  int o = 0, tp = 0;
  uint32_t run = 0;
  for (int ci = 0; ci < nc; ++ci) {
    const uint32_t op = cig[ci] & 3u, len = cig[ci] >> 2;
    if (op == 0) {
      uint32_t pos = 0; // bases of this run already accounted for
      for (uint32_t j0 = 0; j0 < len; j0 += 32) {
        const uint32_t r0 = plane_word_at(rp0, gref + j0), r1 = plane_word_at(rp1, gref + j0), n = min(32u, len - j0);
        uint32_t mm = (r0 ^ plane_word_at(q0, (uint64_t)tp + j0)) | (r1 ^ plane_word_at(q1, (uint64_t)tp + j0));
        if (n < 32u) mm &= (1u << n) - 1u;
        while (mm) {
          const uint32_t b = (uint32_t)__ffs((int)mm) - 1u, j = j0 + b;
          mm &= mm - 1u;
          run += j - pos;
          pos = j + 1;
          if (run) { o = md_put_uint(md, o, run); run = 0; }
          if (o < 48) md[o] = "ACGT"[((r0 >> b) & 1u) | (((r1 >> b) & 1u) << 1)];
          ++o;
        }
      }
      run += len - pos;
      gref += len;
      tp += (int)len;
    } else if (op == 1) {
      tp += (int)len;
    } else {
      if (run) { o = md_put_uint(md, o, run); run = 0; }
      if (o < 48) md[o] = '^';
      ++o;
      for (uint32_t j = 0; j < len; ++j) {
        if (o < 48) md[o] = "ACGT"[(plane_word_at(rp0, gref + j) & 1u) | ((plane_word_at(rp1, gref + j) & 1u) << 1)];
        ++o;
      }
      gref += len;
    }
    if (o > ALN_MD_MAX) return ALN_HOST;
  }
  if (run) o = md_put_uint(md, o, run);
  if (o > ALN_MD_MAX) return ALN_HOST;
  for (int i = o; i < 48; ++i) md[i] = 0;
#pragma unroll
  for (int w = 0; w < 12; ++w)
    dst[w] = (uint32_t)(uint8_t)md[4 * w] | ((uint32_t)(uint8_t)md[4 * w + 1] << 8) | ((uint32_t)(uint8_t)md[4 * w + 2] << 16) |
             ((uint32_t)(uint8_t)md[4 * w + 3] << 24);
  return (uint32_t)o;
}

template <bool HAS_N, int LCAP>
// 12 blocks per SM (40 registers, 20 bytes of spills): the kernel waits on its history words in local memory (ncu: 12 stall
// cycles per instruction on loads at 48 registers and 10 blocks), so more resident warps pay -- cfg3 1.19 -> 1.04 ms per 1 M
// reads, cfg2 0.73 -> 0.65, cfg5 1.82 -> 1.57; 16 blocks (32 registers, 64 bytes of spills) 1.21.
__global__ void __launch_bounds__(128, 12) k_trace(TraceArgs A) {
  const uint64_t mi = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (mi >= A.n_maps) return;
  uint32_t *out = A.aln + mi * ALN_WORDS;
  uint16_t cig[ALN_MAX_OPS];
#pragma unroll
  for (int i = 0; i < ALN_MAX_OPS; ++i) cig[i] = 0;
  const uint32_t rs = A.map_rs[mi];
  const uint32_t read = rs >> 1, dir = rs & 1;
  const uint64_t off = A.read_off[read];
  const int l = (int)(A.read_off[read + 1] - off);
  const ulonglong2 rec = A.maps[mi];
  const int end_off = (int)(int32_t)(uint32_t)rec.y, dist = (int)((rec.y >> 40) & 0xFFu), e = A.e;
  int start = end_off - l + 1;
  if (!A.ref_canonical || (A.read_flags[read] & 1u) || l > LCAP || l <= 0 || start < 0) { aln_store(out, 0, ALN_HOST, cig); return; }
  const uint8_t *text = A.codes + 2 * off + (uint64_t)dir * (uint64_t)l;
  const uint64_t g0 = A.ref.rec_base[rec.x >> 32] + (uint32_t)rec.x; // window[0]
  RefCursor<HAS_N> ref;

  // packed = neither the read nor the reference holds an N: every comparison below is a word-wise XOR of bit planes
  // (k_planes wrote the read's for k_verify)
  const bool packed = !HAS_N && A.refp0 && !A.rs_has_n[rs];
  const uint32_t *q0 = A.planes + (uint64_t)rs * 3ull * A.plane_words, *q1 = q0 + A.plane_words;

  // :289-300 plain comparison at the reported end; no difference -> lM
  int diff = 0;
  if (packed) {
    const uint64_t gs = g0 + (uint64_t)start;
    for (int t = 0; t < l; t += 32) {
      uint32_t x = (plane_word_at(A.refp0, gs + (uint64_t)t) ^ q0[t >> 5]) | (plane_word_at(A.refp1, gs + (uint64_t)t) ^ q1[t >> 5]);
      if (l - t < 32) x &= (1u << (l - t)) - 1u;
      diff |= x != 0;
    }
  } else {
    ref.seek(A.ref, g0 + (uint64_t)start);
    for (int i = 0; i < l; ++i) diff += text[i] != ref.next();
  }
  if (diff == 0) {
    cig[0] = (uint16_t)(l << 2);
    char md[48]; // no difference: the MD text is the read length
    const int o = md_put_uint(md, 0, (uint32_t)l);
    for (int i = o; i < 48; ++i) md[i] = 0;
#pragma unroll
    for (int w = 0; w < 12; ++w)
      out[10 + w] = (uint32_t)(uint8_t)md[4 * w] | ((uint32_t)(uint8_t)md[4 * w + 1] << 8) | ((uint32_t)(uint8_t)md[4 * w + 2] << 16) |
                    ((uint32_t)(uint8_t)md[4 * w + 3] << 24);
    aln_store(out, start, 1, cig, (uint32_t)o);
    return;
  }
  // :303-338 the band again, 32-bit words; the low 16 bits of D0 and HP of every column kept in one
  // word (the band is 2e + 1 <= 15 bits wide; the walk below only looks at bit positions 0..15 and leaves
  // anything else to the host): half the local-memory traffic of two 32-bit histories
  uint32_t hist[LCAP];
  if (packed) { // the columns from bit planes, as in k_verify's verify_columns_planes (32-bit words: no width mask)
    const uint32_t band = (2u << (2 * e)) - 1u;
    const uint32_t *r0p = A.refp0 + (g0 >> 5), *r1p = A.refp1 + (g0 >> 5);
    const uint32_t s0 = (uint32_t)g0 & 31u;
    uint32_t VP = 0, VN = 0;
    uint32_t a0 = __funnelshift_r(__ldg(r0p), __ldg(r0p + 1), s0), a1 = __funnelshift_r(__ldg(r1p), __ldg(r1p + 1), s0);
    for (int i0 = 0; i0 < l; i0 += 32) {
      const int w = i0 >> 5;
      const uint32_t b0 = __funnelshift_r(__ldg(r0p + w + 1), __ldg(r0p + w + 2), s0), b1 = __funnelshift_r(__ldg(r1p + w + 1), __ldg(r1p + w + 2), s0);
      const uint32_t t0 = q0[w], t1 = q1[w];
      const int n = min(32, l - i0);
#pragma unroll 8
      for (int ii = 0; ii < n; ++ii) {
        uint32_t u0 = __funnelshift_r(a0, b0, ii), u1 = __funnelshift_r(a1, b1, ii);
        if (t0 & (1u << ii)) u0 = ~u0;
        if (t1 & (1u << ii)) u1 = ~u1;
        uint32_t X = (~(u0 | u1) & band) | VN;
        const uint32_t D0 = ((VP + (X & VP)) ^ VP) | X;
        const uint32_t HP = VN | ~(VP | D0);
        X = D0 >> 1;
        VP = (VP & D0) | ~(X | HP);
        VN = X & HP;
        hist[i0 + ii] = (D0 & 0xFFFFu) | (HP << 16);
      }
      a0 = b0; a1 = b1;
    }
  } else {
    uint32_t p0 = 0, p1 = 0, p2 = 0, p3 = 0, p4 = 0, VP = 0, VN = 0;
    ref.seek(A.ref, g0);
    for (int i = 0; i < 2 * e; ++i) {
      const uint32_t c = ref.next(), bit = 1u << i;
      p0 |= c == 0 ? bit : 0u; p1 |= c == 1 ? bit : 0u; p2 |= c == 2 ? bit : 0u; p3 |= c == 3 ? bit : 0u; p4 |= c == 4 ? bit : 0u;
    }
    const uint32_t hi = 1u << (2 * e);
    for (int i = 0; i < l; ++i) {
      const uint32_t c = ref.next();
      p0 |= c == 0 ? hi : 0u; p1 |= c == 1 ? hi : 0u; p2 |= c == 2 ? hi : 0u; p3 |= c == 3 ? hi : 0u; p4 |= c == 4 ? hi : 0u;
      const uint32_t t = text[i];
      const uint32_t eq = t == 0 ? p0 : t == 1 ? p1 : t == 2 ? p2 : t == 3 ? p3 : p4;
      uint32_t X = eq | VN;
      const uint32_t D0 = ((VP + (X & VP)) ^ VP) | X;
      const uint32_t HN = VP & D0, HP = VN | ~(VP | D0);
      X = D0 >> 1;
      VN = X & HP;
      VP = HN | ~(X | HP);
      hist[i] = (D0 & 0xFFFFu) | (HP << 16);
      p0 >>= 1; p1 >>= 1; p2 >>= 1; p3 >>= 1; p4 >>= 1;
    }
  }
  // :340-459 the walk.  Runs are collected from the read's 3' end; errors before the first match
  // sit in a pseudo run 'S' that is folded into the run after it (:466-469).  Starting in an
  // empty 'S' run makes the reference's special first step the general step (see sam.c).
  enum { OP_M = 0, OP_I = 1, OP_D = 2, OP_S = 3 };
  uint8_t ops[ALN_MAX_OPS + 2];
  uint16_t cnt[ALN_MAX_OPS + 2];
  int n_ops = 0, bit = start, tp = l - 1, rp = end_off, used = 0, run = OP_S, run_n = 0;
  bool fail = false;
#define CLOSE_RUN(next)                                                                            \
  do {                                                                                             \
    if (n_ops >= ALN_MAX_OPS + 1) { fail = true; break; }                                          \
    ops[n_ops] = (uint8_t)run; cnt[n_ops++] = (uint16_t)run_n; run = (next); run_n = 1;            \
  } while (0)
  // The walk is one long dependency chain, so its loads are kept off it: the history word of the next column is
  // requested while this one is used, and (packed) "reference base == read base" comes from a 32-column match mask of
  // the current diagonal rp - tp, rebuilt from the bit planes when the walk enters another 32-column block or an
  // indel changes the diagonal.
  uint32_t hcur = hist[tp], eqm = 0;
  int eq_blk = -1, eq_delta = 0;
  while (tp >= 0 && (used != dist || tp == l - 1) && !fail) {
    const uint32_t sh = (uint32_t)bit & 31u; // the reference shifts by `bit`; out-of-range shifts are UB there
    if (bit < 0 || bit > 15) { fail = true; break; }
    const uint32_t hnext = tp > 0 ? hist[tp - 1] : 0u;
    const uint32_t h = hcur >> sh, d0 = h & 1u;
    const int tp_was = tp;
    bool same;
    if (packed && rp >= tp) {
      if ((tp >> 5) != eq_blk || rp - tp != eq_delta) {
        eq_blk = tp >> 5; eq_delta = rp - tp;
        const uint64_t gb = g0 + (uint64_t)(32 * eq_blk + eq_delta);
        eqm = ~((plane_word_at(A.refp0, gb) ^ q0[eq_blk]) | (plane_word_at(A.refp1, gb) ^ q1[eq_blk]));
      }
      same = (eqm >> (tp & 31)) & 1u;
    } else same = rp >= 0 && ref_code_fetch<HAS_N>(A.ref, g0 + (uint64_t)rp) == text[tp];
    if (d0 && same) { // match
      --tp; --rp;
      if (run != OP_M) CLOSE_RUN(OP_M); else ++run_n;
    } else if (!d0) { // mismatch
      --tp; --rp; ++used;
      if (run == OP_S || run == OP_M) ++run_n; else CLOSE_RUN(OP_M);
    } else if ((h >> 16) & 1u) { // insertion: a read base only
      --tp; ++bit; ++used; ++start;
      if (run == OP_S || run == OP_I) ++run_n; else CLOSE_RUN(OP_I);
    } else { // deletion: a reference base only
      if (tp == l - 1 && n_ops == 0 && run == OP_S && run_n == 0) { fail = true; break; } // reference asserts (:367)
      --bit; --rp; ++used; --start;
      if (run != OP_D) CLOSE_RUN(OP_D); else ++run_n;
    }
    if (tp != tp_was) hcur = hnext;
  }
  if (!fail && tp >= 0) { // :441-459 the rest of the read matches
    if (run != OP_M) { CLOSE_RUN(OP_M); run_n = tp + 1; }
    else run_n += tp + 1;
  }
#undef CLOSE_RUN
  if (!fail) { ops[n_ops] = (uint8_t)run; cnt[n_ops] = (uint16_t)run_n; }
  int first = 0;
  if (!fail && ops[0] == OP_S) { // :466-469
    if (n_ops == 0) fail = true;
    else { cnt[1] = (uint16_t)(cnt[1] + cnt[0]); first = 1; }
  }
  if (fail || n_ops - first + 1 > ALN_MAX_OPS) { aln_store(out, 0, ALN_HOST, cig); return; }
  int nc = 0;
  for (int i = n_ops; i >= first; --i) cig[nc++] = (uint16_t)(((uint32_t)cnt[i] << 2) | ops[i]);
  const uint32_t md_len = packed ? md_build_store_planes(A.refp0, A.refp1, g0 + (uint64_t)start, q0, q1, cig, nc, out + 10)
                                 : md_build_store<HAS_N>(A.ref, g0 + (uint64_t)start, text, cig, nc, out + 10);
  aln_store(out, start, (uint32_t)nc, cig, md_len);
}
