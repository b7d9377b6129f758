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
  uint32_t smem_words;        // seed_smem_words_per_warp of this plan
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

__global__ void __launch_bounds__(128, 12) k_seed(SeedArgs A) {
  extern __shared__ uint32_t smem_seed[];
  const uint32_t lane = lane_id();
  const uint32_t warp_in_block = threadIdx.x >> 5;
  const uint64_t rs = (uint64_t)blockIdx.x * (blockDim.x >> 5) + warp_in_block;
  if (rs >= 2ull * A.n_reads) return;
  const int k = A.k, step = A.step, Q = A.Q, s = A.s, w_max = A.w_max;
  const uint32_t groups = A.lane_dp ? (uint32_t)step : 1u;
  uint32_t *pw = smem_seed + warp_in_block * A.smem_words;  // 2-bit packed read, big-endian in word
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
  // Four consecutive bases per lane and trip (128 bases per trip: a 150 bp read takes two).  A byte ch is a letter of
  // "ACGT" in either case iff x = ch ^ 0x40 without its case bit is < 32 and bit x of ACGT_MASK is set (A 1, C 3, G 7,
  // T 20); its code (src/utils.h:72) is then ((ch >> 1) ^ (ch >> 2)) & 3 -- A 0, C 1, G 2, T 3 -- and 4 otherwise.
  constexpr uint32_t ACGT_MASK = (1u << 1) | (1u << 3) | (1u << 7) | (1u << 20), ACGTN_MASK = ACGT_MASK | (1u << 14);
  uint32_t amb_lane = 0;
  bool noncanon = false;
  const int n_words = (l + 15) / 16 + 1;
  const char *rb = A.bases + off;
  for (int base = 0; base < n_words * 16; base += 128) {
    const int i0 = base + 4 * (int)lane;
    uint32_t v = 0;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int i = i0 + t;
      if (i < l) {
        const uint32_t ch = (unsigned char)(dir ? rb[l - 1 - i] : rb[i]);
        const uint32_t x = ch ^ 0x40u;
        const bool acgt = (x & 0xDFu) < 32u && ((ACGT_MASK >> (x & 31u)) & 1u);
        noncanon |= !(x < 32u && ((ACGTN_MASK >> (x & 31u)) & 1u)); // anything but upper-case "ACGTN"
        uint32_t c = ((ch >> 1) ^ (ch >> 2)) & 3u;
        if (dir) c ^= 3u;                        // complement; ambiguous stays ambiguous (-> 'N')
        if (!acgt) { c = 4u; amb_lane += i >= k; } // only incoming bases count, utils.h:113
        codes[i] = (uint8_t)c;
        v |= (c & 3u) << (30 - 8 * (int)(lane & 3u) - 2 * t); // ambiguous (4) hashed as A
      }
    }
    v |= __shfl_xor_sync(0xffffffffu, v, 1);
    v |= __shfl_xor_sync(0xffffffffu, v, 2);
    if ((lane & 3u) == 0 && (i0 >> 4) < n_words) pw[i0 >> 4] = v;
  }
  const int amb = (int)__reduce_add_sync(0xffffffffu, amb_lane);
  __syncwarp();
  if (dir == 0) {
    const bool any_noncanon = __any_sync(0xffffffffu, noncanon);
    if (lane == 0) A.read_flags[read] = any_noncanon ? 1 : 0;
  }

  uint8_t status = RS_OK;
  if (n_s <= 0) status = RS_INFEASIBLE;                  // reference: assert, filter.c:167
  else if (Q * step > n_s) status = RS_EMPTY;            // Q > n_s / step, filter.c:168-172
  else if (amb > A.e) status = RS_EMPTY;                 // filter.c:180-182
  else if (n_s - (step - 1) < Q * s * step) status = RS_INFEASIBLE; // (n_s - (step - 1)) / step < Q * s: DP table would be < 2 wide
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
    // Zero slack (cfg3: 27 seeds per group for 9 x 3): group 0 has the most seeds, so when its table is one column wide
    // all are, and every selection is forced -- seed r of group g is the group's q-gram (r - 1) * s, read offset
    // g + (r - 1) * s * step.  No table, no traceback: the sort stage below works out the offsets and the group totals.
    const bool forced = n_s / step - Q * s + 1 == 1;
    if (forced) {
    } else if (A.lane_dp) {
      // ---- one lane per group: serial DP over its own Q x W table -------------------------------
      for (int g = (int)lane; g < step; g += 32) {
        const int n_g = (n_s - g) / step;
        const int W = n_g - Q * s + 1; // real columns 1..W (>= 1 by the feasibility check)
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
      //      seed tables of the candidate kernel: one lane per (group, seed), whole groups per trip (32 / Q of them) so
      //      that a lane finds its group's frequencies in its neighbours' registers ----------------------
      uint64_t sum64 = 0;
      uint32_t forced_before = 0;
      const uint32_t gl = lane / (uint32_t)Q, il = lane - gl * (uint32_t)Q, gpt = 32u / (uint32_t)Q; // groups per trip
      for (uint32_t g0 = 0; g0 < (uint32_t)step; g0 += gpt) {
        const uint32_t g = g0 + gl;
        const bool active = gl < gpt && g < (uint32_t)step;
        uint32_t my_pos = 0, f = 0;
        if (active) {
          my_pos = forced ? g + ((uint32_t)Q - 1u - il) * (uint32_t)(s * step) : sel[g * Q + il];
          f = sm_freq[my_pos];
        }
        uint32_t rank = 0, tot = 0;
        bool wrapped = false;
        const uint32_t lane0 = gl * (uint32_t)Q; // the group's first lane
#pragma unroll
        for (uint32_t j = 0; j < FEM_MAX_Q; ++j) {
          if (j >= (uint32_t)Q) break;
          const uint32_t fj = __shfl_sync(0xffffffffu, f, lane0 + j);
          rank += (fj < f || (fj == f && j < il)) ? 1u : 0u;
          tot += fj;
          wrapped |= tot < fj;
        }
        if (forced) {
          // the single column of the reference's table: running sums, "infinity" once one of them reaches it.  They
          // ascend, so only the total needs the test -- unless a sum wrapped around 2^32 (a q-gram that fills most of a
          // four-billion-entry index): then the lane walks the column as the reference does, wrap-around included
          bool bad_g = A.inf <= tot;
          if (__any_sync(0xffffffffu, active && wrapped)) {
            if (active && il == 0) {
              uint32_t sum = 0;
              bad_g = false;
              for (int r = 1; r <= Q; ++r) {
                uint32_t v = sum + sm_freq[g + (uint32_t)(r - 1) * (uint32_t)(s * step)];
                if (A.inf <= v) { v = A.inf; bad_g = true; }
                sum = v;
              }
              forced_before += sum;
            }
          } else if (active && il == 0) forced_before += tot;
          bad |= active && il == 0 && bad_g;
        }
        if (active) {
          const size_t o = ((size_t)rs * step + g) * Q + rank;
          A.seed_begin[o] = sm_begin[my_pos];
          A.seed_freq[o] = f;
          A.seed_start[o] = my_pos;
        }
        if (!active || il != 0) tot = 0; // one lane per group carries the group's total
        max_f = max(max_f, __reduce_max_sync(0xffffffffu, tot));
        sum64 += (uint64_t)__reduce_add_sync(0xffffffffu, tot & 0xFFFFu) + ((uint64_t)__reduce_add_sync(0xffffffffu, tot >> 16) << 16);
      }
      if (forced) {
        before = __reduce_add_sync(0xffffffffu, forced_before);
        if (__any_sync(0xffffffffu, bad)) { status = RS_INFEASIBLE; before = 0; max_f = 0; sum64 = 0; }
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
