// k_seed_multi.cuh -- K1 in the form the BASELINE configs run: several read-strands per warp.
//
// Same arithmetic as k_seed.cuh's lane form (src/filter.c:3-43,162-204): one lane walks one group's
// small DP table.  There a warp owns one read-strand, so only `step` of its 32 lanes work during the DP
// (4 of 32 on cfg2) and the ncu profile showed that part as 38 % of the kernel's instructions.  Here a warp
// owns R = 32 / step consecutive read-strands: base coding and lookup-table gathers run read-strand
// after read-strand with all lanes, the DP runs once with lane = (read-strand, group), the frequency sort
// runs over the flattened (read-strand, group, seed) entries.  Divisions by the run-time constants step,
// Q and step * Q are multiplications by 2^32 / d (exact for the 16-bit dividends here).
#pragma once
#include "common.cuh"
#include "k_seed.cuh"

struct SeedMultiArgs {
  SeedArgs a;
  int R;                 // read-strands per warp
  uint32_t words_per_rs; // shared-memory words of one read-strand's region (even)
  uint32_t inv_step, inv_Q, inv_stepQ; // floor(2^32 / d) + 1
};

// lut[h], lut[h + 1]: one 8-byte load when h is even
__device__ __forceinline__ uint2 seed_lut_pair(const uint32_t *lut, uint32_t h) {
  if (!(h & 1u)) return __ldg(reinterpret_cast<const uint2 *>(lut + h));
  return make_uint2(__ldg(lut + h), __ldg(lut + h + 1));
}

__device__ __forceinline__ uint32_t div_small(uint32_t x, uint32_t inv) { return __umulhi(x, inv); } // x * d < 2^32

// region of one read-strand, in words: meta[2] (status) | packed read | freq | M[step][w_max + 1] |
// arg-min bytes [step][Q][w_max] | selected positions (uint16) [step][Q].  The first occurrence index of
// a selected seed is read again from the lookup table (an L2 hit) instead of being kept for every q-gram.
__host__ __device__ inline uint32_t seed_multi_words_per_rs(int n_words_max, int ns_max, int w_max, int Q, int step) {
  const uint32_t w = 2u + (uint32_t)n_words_max + (uint32_t)ns_max + (uint32_t)step * (uint32_t)(w_max + 1) +
                     ((uint32_t)step * Q * w_max + 3u) / 4u + ((uint32_t)step * Q + 1u) / 2u;
  return (w + 1u) & ~1u;
}
#define SEED_MULTI_SCRATCH 64u // words per warp after the regions: per-lane DP minimum and failure flag

__global__ void __launch_bounds__(128, 10) k_seed_multi(SeedMultiArgs B) {
  extern __shared__ __align__(16) uint32_t smem_seed_multi[];
  const SeedArgs &A = B.a;
  const uint32_t lane = lane_id(), warp_in_block = threadIdx.x >> 5;
  const int R = B.R, k = A.k, step = A.step, Q = A.Q, s = A.s, w_max = A.w_max;
  const uint64_t n_rs = 2ull * A.n_reads;
  const uint64_t rs0 = ((uint64_t)blockIdx.x * (blockDim.x >> 5) + warp_in_block) * (uint64_t)R;
  if (rs0 >= n_rs) return;
  const int n_here = (int)min((uint64_t)R, n_rs - rs0);
  uint32_t *warp_base = smem_seed_multi + (size_t)warp_in_block * ((size_t)R * B.words_per_rs + SEED_MULTI_SCRATCH);
  uint32_t *scratch = warp_base + (size_t)R * B.words_per_rs;
  const uint32_t o_pw = 2u, o_freq = o_pw + (uint32_t)A.n_words_max,
                 o_M = o_freq + (uint32_t)A.ns_max, o_amin = o_M + (uint32_t)step * (uint32_t)(w_max + 1),
                 o_sel = o_amin + ((uint32_t)step * Q * w_max + 3u) / 4u;

  // ---- phase 1: base codes, 2-bit packing, feasibility; lane r keeps read-strand r's status and n_s ----
  uint32_t my_status = RS_OK;
  int my_ns = 0, my_l = 0;
  uint64_t my_off = 0;
  if ((int)lane < n_here) { // lane r reads read-strand r's offsets once
    const uint32_t read = (uint32_t)((rs0 + lane) >> 1);
    my_off = A.read_off[read];
    my_l = (int)(A.read_off[read + 1] - my_off);
  }
  // characters of a read-strand, four per lane (positions lane, lane + 32, ...; reads longer than 128
  // take further rounds); the next read-strand's are requested before this one's are used
  uint32_t cur[4], nxt[4] = {0, 0, 0, 0};
  auto fetch = [&](int r, int base, uint32_t (&dst)[4]) {
    const uint64_t off = __shfl_sync(0xffffffffu, my_off, r);
    const int l = __shfl_sync(0xffffffffu, my_l, r);
    const uint32_t dir = (uint32_t)((rs0 + (uint64_t)r) & 1);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = base + u * 32 + (int)lane;
      dst[u] = i < l ? (uint32_t)(unsigned char)A.bases[off + (uint64_t)(dir ? l - 1 - i : i)] : 0u;
    }
  };
  fetch(0, 0, cur);
  if (!(R & 1)) {
    // Both strands of a read at once (R even: a warp's read-strands are whole reads).  The characters are
    // read and decoded once; the reverse strand's codes are stored mirrored, its ambiguity count comes from
    // the same ballot, and its packed words are the forward words read backwards: a 16-base window of the
    // forward stream, 2-bit groups reversed (brev + swap of neighbouring bits), complemented.  Ambiguous
    // bases (hashed as A on both strands, utils.h:95-100) are cleared afterwards where a read has any.
    for (int q = 0; 2 * q < n_here; ++q) {
      const int r0 = 2 * q;
      const uint32_t read = (uint32_t)((rs0 + (uint64_t)r0) >> 1);
      const uint64_t off = __shfl_sync(0xffffffffu, my_off, r0);
      const int l = __shfl_sync(0xffffffffu, my_l, r0);
      const int n_s = l - k + 1;
      uint8_t *codes_f = A.codes + 2 * off, *codes_r = codes_f + l;
      uint32_t *pw_f = warp_base + (size_t)r0 * B.words_per_rs + o_pw, *pw_r = pw_f + B.words_per_rs;
      int amb_f = 0, amb_r = 0;
      bool noncanon = false;
      uint32_t any_amb = 0;
      const int n_words = (l + 15) / 16 + 1;
      for (int base0 = 0; base0 < n_words * 16; base0 += 128) {
        if (base0 + 128 < n_words * 16) fetch(r0, base0 + 128, nxt);
        else if (r0 + 2 < n_here) fetch(r0 + 2, 0, nxt);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i0 = base0 + u * 32, i = i0 + (int)lane;
          if (i0 >= n_words * 16) break;
          const bool valid = i < l;
          uint32_t c = 0;
          if (valid) {
            const uint32_t ch = cur[u];
            const uint32_t two = ((ch >> 1) & 3u) ^ ((ch >> 2) & 1u);
            const bool acgt = ((0x54474341u >> (8u * two)) & 0xFFu) == ch;
            noncanon |= !(acgt || ch == 'N');
            c = acgt ? two : 4u;
            codes_f[i] = (uint8_t)c;
            codes_r[l - 1 - i] = (uint8_t)(acgt ? 3u - two : 4u); // ambiguous stays ambiguous on the reverse strand
          }
          const uint32_t amb4 = __ballot_sync(0xffffffffu, valid && c == 4u);
          if (amb4) { // only incoming bases count (utils.h:113): forward i >= k, reverse l - 1 - i >= k
            const int nf = k - i0, nr = l - k - i0;
            const uint32_t mask_f = nf <= 0 ? 0xFFFFFFFFu : nf >= 32 ? 0u : 0xFFFFFFFFu << nf;
            const uint32_t mask_r = nr <= 0 ? 0u : nr >= 32 ? 0xFFFFFFFFu : (1u << nr) - 1u;
            amb_f += __popc(amb4 & mask_f);
            amb_r += __popc(amb4 & mask_r);
            any_amb = 1;
          }
          uint32_t v = (valid && c < 4u) ? c << (30 - 2 * (i & 15)) : 0u;          // ambiguous hashed as A
          v |= __shfl_xor_sync(0xffffffffu, v, 8);
          v |= __shfl_xor_sync(0xffffffffu, v, 4);
          v |= __shfl_xor_sync(0xffffffffu, v, 2);
          v |= __shfl_xor_sync(0xffffffffu, v, 1);
          if ((lane & 15u) == 0 && (i >> 4) < n_words) pw_f[i >> 4] = v;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) cur[u] = nxt[u];
      }
      __syncwarp();
      for (int w = (int)lane; w < n_words; w += 32) {
        const int s0 = l - 16 - 16 * w;            // first forward base of the window, may lie before the read
        const int a0 = s0 >> 4, sh = 2 * (s0 & 15); // floor division: s0 = 16 * a0 + (s0 & 15)
        const uint32_t hi = a0 >= 0 && a0 < n_words ? pw_f[a0] : 0u;
        const uint32_t lo = a0 + 1 >= 0 && a0 + 1 < n_words ? pw_f[a0 + 1] : 0u;
        uint32_t x = __brev(__funnelshift_l(lo, hi, sh));
        x = ((x & 0x55555555u) << 1) | ((x >> 1) & 0x55555555u);
        pw_r[w] = ~x;
      }
      if (any_amb) {
        __syncwarp();
        for (int i = (int)lane; i < l; i += 32)
          if (codes_f[i] == 4) {
            const int j = l - 1 - i;
            atomicAnd(&pw_r[j >> 4], ~(3u << (30 - 2 * (j & 15))));
          }
      }
      const bool any_noncanon = __any_sync(0xffffffffu, noncanon);
      if (lane == 0) A.read_flags[read] = any_noncanon ? 1 : 0;
#pragma unroll
      for (int d = 0; d < 2; ++d) {
        const int amb = d ? amb_r : amb_f;
        uint32_t status = RS_OK;
        if (n_s <= 0) status = RS_INFEASIBLE;                                              // reference: assert, filter.c:167
        else if ((uint32_t)Q > div_small((uint32_t)n_s, B.inv_step)) status = RS_EMPTY;     // filter.c:168-172
        else if (amb > A.e) status = RS_EMPTY;                                              // filter.c:180-182
        else if (n_s - (step - 1) < 0 || div_small((uint32_t)(n_s - (step - 1)), B.inv_step) < (uint32_t)(Q * s)) status = RS_INFEASIBLE;
        if ((int)lane == r0 + d) { my_status = status; my_ns = n_s; }
      }
    }
  } else
  for (int r = 0; r < n_here; ++r) {
    const uint64_t rs = rs0 + (uint64_t)r;
    const uint32_t read = (uint32_t)(rs >> 1), dir = (uint32_t)(rs & 1);
    const uint64_t off = __shfl_sync(0xffffffffu, my_off, r);
    const int l = __shfl_sync(0xffffffffu, my_l, r);
    const int n_s = l - k + 1;
    uint8_t *codes = A.codes + 2 * off + (uint64_t)dir * (uint64_t)l;
    uint32_t *pw = warp_base + (size_t)r * B.words_per_rs + o_pw;
    int amb = 0;
    bool noncanon = false;
    const int n_words = (l + 15) / 16 + 1;
    for (int base0 = 0; base0 < n_words * 16; base0 += 128) {
      if (base0 + 128 < n_words * 16) fetch(r, base0 + 128, nxt);
      else if (r + 1 < n_here) fetch(r + 1, 0, nxt);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = base0 + u * 32 + (int)lane;
        if (base0 + u * 32 >= n_words * 16) break;
        const bool valid = i < l;
        uint32_t c = 0;
        if (valid) {
          const uint32_t ch = cur[u];
          // A 0x41, C 0x43, G 0x47, T 0x54: bits 1..2 give 0, 1, 3, 2; bit 2 swaps the last two
          const uint32_t two = ((ch >> 1) & 3u) ^ ((ch >> 2) & 1u);
          const bool acgt = ((0x54474341u >> (8u * two)) & 0xFFu) == ch;
          noncanon |= !(acgt || ch == 'N');
          c = acgt ? (dir ? 3u - two : two) : 4u; // ambiguous stays ambiguous on the reverse strand
          codes[i] = (uint8_t)c;
        }
        amb += __popc(__ballot_sync(0xffffffffu, valid && c == 4u && i >= k)); // only incoming bases, utils.h:113
        uint32_t v = (valid && c < 4u) ? c << (30 - 2 * (i & 15)) : 0u;          // ambiguous hashed as A
        v |= __shfl_xor_sync(0xffffffffu, v, 8);
        v |= __shfl_xor_sync(0xffffffffu, v, 4);
        v |= __shfl_xor_sync(0xffffffffu, v, 2);
        v |= __shfl_xor_sync(0xffffffffu, v, 1);
        if ((lane & 15u) == 0 && (i >> 4) < n_words) pw[i >> 4] = v;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) cur[u] = nxt[u];
    }
    if (dir == 0) {
      const bool any_noncanon = __any_sync(0xffffffffu, noncanon);
      if (lane == 0) A.read_flags[read] = any_noncanon ? 1 : 0;
    }
    uint32_t status = RS_OK;
    if (n_s <= 0) status = RS_INFEASIBLE;                                              // reference: assert, filter.c:167
    else if ((uint32_t)Q > div_small((uint32_t)n_s, B.inv_step)) status = RS_EMPTY;     // filter.c:168-172
    else if (amb > A.e) status = RS_EMPTY;                                              // filter.c:180-182
    else if (n_s - (step - 1) < 0 || div_small((uint32_t)(n_s - (step - 1)), B.inv_step) < (uint32_t)(Q * s)) status = RS_INFEASIBLE;
    if ((int)lane == r) { my_status = status; my_ns = n_s; }
  }
  __syncwarp();

  // ---- phase 2: hashes and lookup-table gathers (two adjacent uint32 per q-gram) ----------------------
  const uint32_t mask = (uint32_t)((1ull << (2 * k)) - 1);
  for (int r = 0; r < n_here; ++r) {
    const uint32_t st = __shfl_sync(0xffffffffu, my_status, r);
    const int n_s = __shfl_sync(0xffffffffu, my_ns, r);
    if (st != RS_OK) continue;
    uint32_t *reg = warp_base + (size_t)r * B.words_per_rs;
    const uint32_t *pw = reg + o_pw;
    for (int p0 = 0; p0 < n_s; p0 += 128) { // four gathers per lane in flight
      uint32_t b[4], e2[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int p = p0 + u * 32 + (int)lane;
        b[u] = e2[u] = 0;
        if (p < n_s) {
          const uint64_t x = ((uint64_t)pw[p >> 4] << 32) | pw[(p >> 4) + 1];
          const uint32_t h = (uint32_t)(x >> (64 - 2 * (p & 15) - 2 * k)) & mask;
          const uint2 be = seed_lut_pair(A.lut, h);
          b[u] = be.x;
          e2[u] = be.y;
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int p = p0 + u * 32 + (int)lane;
        if (p < n_s) reg[o_freq + p] = e2[u] - b[u];
      }
    }
  }
  __syncwarp();

  // ---- phase 3: DP, lane = (read-strand, group) --------------------------------------------------------
  {
    const uint32_t rl = div_small(lane, B.inv_step);
    const int g = (int)(lane - rl * (uint32_t)step);
    const uint32_t src = rl < (uint32_t)n_here ? rl : 0u;
    const uint32_t st = __shfl_sync(0xffffffffu, my_status, src);
    const int n_s = __shfl_sync(0xffffffffu, my_ns, src);
    uint32_t before = 0, bad = 0;
    if (rl < (uint32_t)n_here && st == RS_OK) {
      uint32_t *reg = warp_base + (size_t)rl * B.words_per_rs;
      const uint32_t *sm_freq = reg + o_freq;
      const int n_g = (int)div_small((uint32_t)(n_s - g), B.inv_step);
      const int W = n_g - Q * s + 1; // real columns 1..W (>= 1 by the feasibility check)
      uint32_t *M = reg + o_M + (size_t)g * (w_max + 1);
      uint8_t *am = reinterpret_cast<uint8_t *>(reg + o_amin) + (size_t)g * Q * w_max;
      uint16_t *sel = reinterpret_cast<uint16_t *>(reg + o_sel);
      for (int c = 0; c <= W; ++c) M[c] = 0; // row 0 (column 0 unused)
      for (int r = 1; r <= Q; ++r) {
        uint32_t cv = A.inf, ci = 0; // M[r][0] = "infinity", stop marker
        const uint32_t *f = sm_freq + g + ((r - 1) * s - 1) * step;
        uint8_t *amr = am + (r - 1) * w_max - 1;
        for (int c = 1; c <= W; ++c) {
          uint32_t v = M[c] + f[c * step], idx = (uint32_t)c;
          if (cv <= v) { v = cv; idx = ci; }
          M[c] = v;
          amr[c] = (uint8_t)idx;
          cv = v; ci = idx;
        }
      }
      before = M[W];
      int c = W; // traceback (filter.c:30-41): right-most seed first
      for (int r = Q; r >= 1; --r) {
        c = am[(r - 1) * w_max + (c - 1)];
        if (c == 0) { bad = 1; break; } // fewer than Q seeds found: the reference reads uninitialised memory
        sel[g * Q + (Q - r)] = (uint16_t)(g + (c + (r - 1) * s - 1) * step);
      }
    }
    scratch[lane] = before;
    scratch[32 + lane] = bad;
  }
  __syncwarp();
  uint32_t my_before = 0;
  if ((int)lane < n_here) { // lane r closes read-strand r: sums over its groups, meta for phase 4
    uint32_t any_bad = 0;
    for (int g = 0; g < step; ++g) {
      my_before += scratch[lane * step + g];
      any_bad |= scratch[32 + lane * step + g];
    }
    if (my_status == RS_OK && any_bad) { my_status = RS_INFEASIBLE; my_before = 0; }
    if (my_status != RS_OK) my_before = 0;
    warp_base[(size_t)lane * B.words_per_rs] = my_status;
  }
  __syncwarp();
  scratch[lane] = 0; // from here: total frequency of the selected seeds per (read-strand, group)
  __syncwarp();

  // ---- phase 4: stable ascending sort by frequency inside each group (filter.c:204) and the seed tables
  //      of the candidate kernels: one lane per (read-strand, group, seed) ------------------------------
  const int stepQ = step * Q, n_entries = n_here * stepQ;
  for (int base = 0; base < n_entries; base += 32) {
    const int idx = base + (int)lane;
    if (idx >= n_entries) continue;
    const uint32_t rl = div_small((uint32_t)idx, B.inv_stepQ);
    uint32_t *reg = warp_base + (size_t)rl * B.words_per_rs;
    if (reg[0] != RS_OK) continue;
    const int rem = idx - (int)rl * stepQ;
    const int g = (int)div_small((uint32_t)rem, B.inv_Q), i = rem - g * Q;
    const uint32_t *sm_freq = reg + o_freq;
    const uint16_t *sg = reinterpret_cast<const uint16_t *>(reg + o_sel) + g * Q;
    const uint32_t my_pos = sg[i], f = sm_freq[my_pos];
    const uint32_t *pw = reg + o_pw;
    const uint64_t x = ((uint64_t)pw[my_pos >> 4] << 32) | pw[(my_pos >> 4) + 1];
    const uint32_t begin = __ldg(A.lut + ((uint32_t)(x >> (64 - 2 * (my_pos & 15) - 2 * k)) & mask));
    int rank = 0;
    uint32_t tot = 0;
    for (int j = 0; j < Q; ++j) {
      const uint32_t fj = sm_freq[sg[j]];
      rank += (fj < f) || (fj == f && j < i);
      tot += fj;
    }
    const size_t o = ((size_t)(rs0 + rl) * step + g) * Q + rank;
    A.seed_begin[o] = begin;
    A.seed_freq[o] = f;
    A.seed_start[o] = my_pos;
    if (i == 0) scratch[rl * (uint32_t)step + (uint32_t)g] = tot; // one lane per group carries the group's total
  }
  __syncwarp();
  if ((int)lane < n_here) {
    uint32_t max_f = 0;
    uint64_t sum64 = 0;
    for (int g = 0; g < step; ++g) {
      const uint32_t tot = scratch[lane * step + g];
      max_f = max(max_f, tot);
      sum64 += tot;
    }
    const uint64_t rs = rs0 + lane;
    A.rs_status[rs] = (uint8_t)my_status;
    A.rs_before[rs] = my_before;
    A.rs_max_f[rs] = max_f;
    A.rs_sum_f[rs] = sum64 > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)sum64;
  }
  // batch-wide sum of the DP minima (filter.c:202): one atomic per warp
  unsigned long long tb = my_before;
#pragma unroll
  for (int d = 16; d; d >>= 1) tb += __shfl_xor_sync(0xffffffffu, tb, d);
  if (lane == 0 && tb) atomicAdd(A.total_before, tb);
}
