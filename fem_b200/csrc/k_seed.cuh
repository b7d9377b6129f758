// k_seed.cuh -- K1: per read-strand base coding, q-gram hashing, lookup-table frequency gather
// and per-group optimal seed selection.  One warp per read-strand.
//
// Restates, in a warp-parallel form, src/sequence_batch.h:90-98 (reverse complement),
// src/utils.h:83-117 (hashing), src/filter.c:162-204 (feasibility checks, per-group seed tables,
// DP of src/filter.c:3-43 and the frequency sort).
//
// The DP  M[r][c] = min(M[r][c-1], M[r-1][c] + f[c+(r-1)s-1])  (seed taken only when strictly
// cheaper) is a prefix minimum along c with "left-most achiever wins" tie breaking.  Only the
// arg-min column per cell is kept (uint16) for the traceback of src/filter.c:30-41.  Two forms:
//   lane_dp = 1  one lane per seed group walks its own small table serially -- the tables of the
//                BASELINE configs are 1..6 columns wide (W = n_g - Q*s + 1), so all `step` groups
//                finish in Q*W short steps at once;
//   lane_dp = 0  wide tables (short Q, long reads): the warp scans one row of one group at a time.
// The ncu profile of round 1 showed the all-warp-scan form issue bound at ~4700 instructions per
// read-strand; the lane form needs a few hundred.
#pragma once
#include "common.cuh"

struct SeedArgs {
  const char *bases;          // batch bases, reads back to back
  const uint64_t *read_off;   // n_reads + 1
  uint32_t n_reads;
  const uint32_t *lut;
  uint32_t inf;               // (uint32) occurrence_table_size: the DP's "infinity" (filter.c:9)
  int k, step, e, a, Q, s;
  int n_words_max, ns_max, w_max; // shared-memory plan for the longest read of the batch
  int lane_dp;
  uint8_t *codes;             // out: 2 * total_bases; strand d of read i at 2*off[i] + d*len
  uint8_t *read_flags;        // out: per read, bit 0 = a character outside "ACGTN" (k_trace leaves it to the host)
  uint8_t *rs_status;         // out: per read-strand
  uint32_t *rs_before;        // out: DP minimum summed over groups (filter.c:202)
  uint32_t *rs_max_f;         // out: max over groups of the selected seeds' total frequency
  uint32_t *rs_sum_f;         // out: sum over groups of the same (saturating)
  uint32_t *seed_begin;       // out: [rs*step+g][Q] first occurrence index, frequency order
  uint32_t *seed_freq;
  uint32_t *seed_start;       // read offset of the seed
  unsigned long long *total_before; // out: batch-wide sum of rs_before
};

// words of shared memory one warp needs: packed read + freq + begin + DP rows + arg-min tables
// (uint16) + selected seed positions (uint16); the DP parts exist once per group in the lane form
__host__ __device__ inline size_t seed_smem_words_per_warp(int n_words_max, int ns_max, int w_max, int Q, int step,
                                                           int lane_dp) {
  const size_t groups = lane_dp ? (size_t)step : 1;
  return (size_t)n_words_max + 2 * (size_t)ns_max + groups * (size_t)(w_max + 1) + (groups * (size_t)Q * w_max + 1) / 2 +
         ((size_t)step * Q + 1) / 2;
}

__global__ void __launch_bounds__(128) k_seed(SeedArgs A) {
  extern __shared__ uint32_t smem_seed[];
  const uint32_t lane = lane_id();
  const uint32_t warp_in_block = threadIdx.x >> 5;
  const uint64_t rs = (uint64_t)blockIdx.x * (blockDim.x >> 5) + warp_in_block;
  if (rs >= 2ull * A.n_reads) return;
  const int k = A.k, step = A.step, Q = A.Q, s = A.s, w_max = A.w_max;
  const size_t groups = A.lane_dp ? (size_t)step : 1;
  const size_t per_warp = seed_smem_words_per_warp(A.n_words_max, A.ns_max, w_max, Q, step, A.lane_dp);
  uint32_t *pw = smem_seed + warp_in_block * per_warp;      // 2-bit packed read, big-endian in word
  uint32_t *sm_freq = pw + A.n_words_max;
  uint32_t *sm_begin = sm_freq + A.ns_max;
  uint32_t *m_rows = sm_begin + A.ns_max;                   // [groups][w_max + 1], columns 0..W
  uint16_t *amin = reinterpret_cast<uint16_t *>(m_rows + groups * (w_max + 1)); // [groups][Q][w_max]
  uint16_t *sel = reinterpret_cast<uint16_t *>(m_rows + groups * (w_max + 1) + (groups * Q * w_max + 1) / 2); // [step][Q]

  const uint32_t read = (uint32_t)(rs >> 1), dir = (uint32_t)(rs & 1);
  const uint64_t off = A.read_off[read];
  const int l = (int)(A.read_off[read + 1] - off);
  const int n_s = l - k + 1;
  uint8_t *codes = A.codes + 2 * off + (uint64_t)dir * (uint64_t)l;
  if (l > FEM_MAX_READ_LEN) { // beyond the shared-memory plan: left unmapped, counted as infeasible
    if (lane == 0) {
      A.rs_status[rs] = RS_INFEASIBLE; A.rs_before[rs] = 0; A.rs_max_f[rs] = 0; A.rs_sum_f[rs] = 0;
      if (dir == 0) A.read_flags[read] = 0;
    }
    return;
  }

  // ---- base codes (strand oriented), 2-bit packing, ambiguous-base count --------------------
  int amb = 0;
  bool noncanon = false;
  const int n_words = (l + 15) / 16 + 1;
  for (int base = 0; base < n_words * 16; base += 32) {
    const int i = base + (int)lane;
    uint32_t c = 0, valid = i < l;
    if (valid) {
      const unsigned char ch = (unsigned char)(dir ? A.bases[off + (uint64_t)(l - 1 - i)] : A.bases[off + (uint64_t)i]);
      c = fem_base_code(ch);
      noncanon |= !(ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T' || ch == 'N');
      if (dir && c < 4) c = 3u - c; // 3 ^ code; ambiguous stays ambiguous (-> 'N')
      codes[i] = (uint8_t)c;
    }
    amb += __popc(__ballot_sync(0xffffffffu, valid && c == 4 && i >= k)); // only incoming bases, utils.h:113
    uint32_t v = (valid && c < 4) ? c << (30 - 2 * (i & 15)) : 0u;         // ambiguous hashed as A
    v |= __shfl_xor_sync(0xffffffffu, v, 8);
    v |= __shfl_xor_sync(0xffffffffu, v, 4);
    v |= __shfl_xor_sync(0xffffffffu, v, 2);
    v |= __shfl_xor_sync(0xffffffffu, v, 1);
    if ((lane & 15) == 0 && (i >> 4) < n_words) pw[i >> 4] = v;
  }
  __syncwarp();
  if (dir == 0) {
    const bool any_noncanon = __any_sync(0xffffffffu, noncanon);
    if (lane == 0) A.read_flags[read] = any_noncanon ? 1 : 0;
  }

  uint8_t status = RS_OK;
  if (n_s <= 0) status = RS_INFEASIBLE;                  // reference: assert, filter.c:167
  else if (Q > n_s / step) status = RS_EMPTY;            // filter.c:168-172
  else if (amb > A.e) status = RS_EMPTY;                 // filter.c:180-182
  else if ((n_s - (step - 1)) / step < Q * s) status = RS_INFEASIBLE; // DP table would be < 2 wide
  uint32_t before = 0, max_f = 0, sum_f = 0;

  if (status == RS_OK) {
    // ---- hashes and lookup-table gathers (two adjacent uint32 per q-gram) --------------------
    const uint32_t mask = (uint32_t)((1ull << (2 * k)) - 1);
    for (int p = (int)lane; p < n_s; p += 32) {
      const uint64_t x = ((uint64_t)pw[p >> 4] << 32) | pw[(p >> 4) + 1];
      const uint32_t h = (uint32_t)(x >> (64 - 2 * (p & 15) - 2 * k)) & mask;
      const uint32_t b = __ldg(A.lut + h), e2 = __ldg(A.lut + h + 1);
      sm_begin[p] = b;
      sm_freq[p] = e2 - b;
    }
    __syncwarp();
    bool bad = false; // fewer than Q seeds found: the reference reads uninitialised memory
    if (A.lane_dp) {
      // ---- one lane per group: serial DP over its own Q x W table -------------------------------
      for (int g = (int)lane; g < step; g += 32) {
        const int n_g = (n_s - g) / step;
        const int W = n_g - Q * s + 1; // real columns 1..W (>= 1 by the feasibility check)
        if (W == 1) {
          // zero slack (cfg3: 27 seeds per group for 9 x 3): the table is one column wide and the selection is forced --
          // seed r is the group's q-gram (r - 1) * s; a row whose sum reaches "infinity" ends the reference's traceback
          // (column 0), exactly as in the general form below
          uint32_t sum = 0;
          for (int r = 1; r <= Q; ++r) {
            const int q = g + (r - 1) * s * step;
            uint32_t v = sum + sm_freq[q];
            if (A.inf <= v) { v = A.inf; bad = true; }
            sum = v;
            sel[g * Q + (Q - r)] = (uint16_t)q;
          }
          before += sum;
          continue;
        }
        uint32_t *M = m_rows + (size_t)g * (w_max + 1);
        uint16_t *am = amin + (size_t)g * Q * w_max;
        for (int c = 0; c <= W; ++c) M[c] = 0; // row 0 (column 0 unused)
        for (int r = 1; r <= Q; ++r) {
          uint32_t cv = A.inf, ci = 0; // M[r][0] = "infinity", stop marker
          const uint32_t *f = sm_freq + g + ((r - 1) * s - 1) * step;
          for (int c = 1; c <= W; ++c) {
            uint32_t v = M[c] + f[c * step], idx = (uint32_t)c;
            if (cv <= v) { v = cv; idx = ci; }
            M[c] = v;
            am[(r - 1) * w_max + (c - 1)] = (uint16_t)idx;
            cv = v; ci = idx;
          }
        }
        before += M[W];
        int c = W; // traceback (filter.c:30-41): right-most seed first
        for (int r = Q; r >= 1; --r) {
          c = am[(r - 1) * w_max + (c - 1)];
          if (c == 0) { bad = true; break; }
          sel[g * Q + (Q - r)] = (uint16_t)(g + (c + (r - 1) * s - 1) * step);
        }
      }
      before = __reduce_add_sync(0xffffffffu, before);
    } else {
      // ---- wide tables: the warp scans one row of one group at a time ----------------------------
      uint32_t *m_row = m_rows;
      for (int g = 0; g < step && !bad; ++g) {
        const int n_g = (n_s - g) / step;
        const int W = n_g - Q * s + 1;
        for (int c = (int)lane; c <= W; c += 32) m_row[c] = 0;
        __syncwarp();
        for (int r = 1; r <= Q; ++r) {
          uint32_t carry_v = A.inf, carry_i = 0;
          for (int c0 = 1; c0 <= W; c0 += 32) {
            const int c = c0 + (int)lane;
            const bool valid = c <= W;
            uint32_t v = 0xFFFFFFFFu, idx = (uint32_t)c;
            if (valid) v = m_row[c] + sm_freq[g + (c + (r - 1) * s - 1) * step];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
              const uint32_t ov = __shfl_up_sync(0xffffffffu, v, d), oi = __shfl_up_sync(0xffffffffu, idx, d);
              if ((int)lane >= d && ov <= v) { v = ov; idx = oi; }
            }
            if (carry_v <= v) { v = carry_v; idx = carry_i; }
            if (valid) { m_row[c] = v; amin[(r - 1) * w_max + (c - 1)] = (uint16_t)idx; }
            const int last = min(31, W - c0);
            carry_v = __shfl_sync(0xffffffffu, v, last);
            carry_i = __shfl_sync(0xffffffffu, idx, last);
          }
          __syncwarp();
        }
        before += m_row[W];
        int c = W;
        for (int r = Q; r >= 1; --r) {
          c = amin[(r - 1) * w_max + (c - 1)];
          if (c == 0) { bad = true; break; }
          if (lane == 0) sel[g * Q + (Q - r)] = (uint16_t)(g + (c + (r - 1) * s - 1) * step);
        }
        __syncwarp();
      }
    }
    bad = __any_sync(0xffffffffu, bad);
    __syncwarp();
    if (bad) {
      status = RS_INFEASIBLE;
      before = 0;
    } else {
      // ---- stable ascending sort by frequency inside each group (filter.c:204) and the
      //      seed tables of the candidate kernel: one lane per (group, seed) ----------------------
      uint64_t sum64 = 0;
      for (int base = 0; base < step * Q; base += 32) {
        const int idx = base + (int)lane;
        uint32_t tot = 0;
        if (idx < step * Q) {
          const int g = idx / Q, i = idx - g * Q;
          const uint16_t *sg = sel + g * Q;
          const uint32_t my_pos = sg[i], f = sm_freq[my_pos];
          int rank = 0;
          for (int j = 0; j < Q; ++j) {
            const uint32_t fj = sm_freq[sg[j]];
            rank += (fj < f) || (fj == f && j < i);
            tot += fj;
          }
          const size_t o = ((size_t)rs * step + g) * Q + rank;
          A.seed_begin[o] = sm_begin[my_pos];
          A.seed_freq[o] = f;
          A.seed_start[o] = my_pos;
          if (i != 0) tot = 0; // one lane per group carries the group's total
        }
        max_f = max(max_f, __reduce_max_sync(0xffffffffu, tot));
        sum64 += (uint64_t)__reduce_add_sync(0xffffffffu, tot & 0xFFFFu) + ((uint64_t)__reduce_add_sync(0xffffffffu, tot >> 16) << 16);
      }
      sum_f = sum64 > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)sum64;
    }
  }
  if (lane == 0) {
    A.rs_status[rs] = status;
    A.rs_before[rs] = before;
    A.rs_max_f[rs] = max_f;
    A.rs_sum_f[rs] = sum_f;
    if (before) atomicAdd(A.total_before, (unsigned long long)before);
  }
}
