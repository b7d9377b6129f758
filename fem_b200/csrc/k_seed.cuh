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
  const uint8_t *base_table;  // [strand][byte]: code (0..4, complemented on strand 1) | 8 if the byte is none of "ACGTN" (seed_base_table)
  uint32_t inf;               // (uint32) occurrence_table_size: the DP's "infinity" (filter.c:9)
  int k, step, e, a, Q, s;
  int n_words_max, ns_max, w_max; // shared-memory plan for the longest read of the batch
  int lane_dp;
  uint32_t smem_words;        // seed_smem_words_per_warp of this plan
  uint32_t amin_word, sel_word; // word offsets of the arg-min tables and the selected positions in a warp's area
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

#define SEED_TABLE_BYTES 512 // shared-memory copy of base_table, after the four warps' areas

// src/utils.h:72 (codes) and src/sequence_batch.h:96 (complement of strand 1) for every byte value, plus the "not one
// of ACGTN in upper case" flag that sends a read's SAM text to the host
inline void seed_base_table(uint8_t out[SEED_TABLE_BYTES]) {
  for (int d = 0; d < 2; ++d)
    for (int ch = 0; ch < 256; ++ch) {
      uint32_t c = fem_base_code((unsigned char)ch);
      if (d && c < 4) c = 3u - c;
      const bool canon = ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T' || ch == 'N';
      out[d * 256 + ch] = (uint8_t)(c | (canon ? 0u : 8u));
    }
}

template <int QT>
__global__ void __launch_bounds__(128, 12) k_seed(SeedArgs A) {
  extern __shared__ uint32_t smem_seed[];
  const uint32_t lane = lane_id();
  const uint32_t warp_in_block = threadIdx.x >> 5;
  uint8_t *tbl = reinterpret_cast<uint8_t *>(smem_seed + 4u * A.smem_words);
  reinterpret_cast<uint32_t *>(tbl)[threadIdx.x] = __ldg(reinterpret_cast<const uint32_t *>(A.base_table) + threadIdx.x); // 128 x 4 bytes
  __syncthreads();
  const uint64_t rs = (uint64_t)blockIdx.x * (blockDim.x >> 5) + warp_in_block;
  if (rs >= 2ull * A.n_reads) return;
  const int k = A.k, step = A.step, s = A.s, w_max = A.w_max;
  constexpr int Q = QT;
  uint32_t *pw = smem_seed + warp_in_block * A.smem_words;  // 2-bit packed read, big-endian in word
  uint32_t *sm_freq = pw + A.n_words_max;
  uint32_t *sm_begin = sm_freq + A.ns_max;
  uint32_t *m_rows = sm_begin + A.ns_max;                   // [groups][w_max + 1], columns 0..W
  uint16_t *amin = reinterpret_cast<uint16_t *>(pw + A.amin_word); // [groups][Q][w_max]
  uint16_t *sel = reinterpret_cast<uint16_t *>(pw + A.sel_word);   // [step][Q]

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
  // Four consecutive characters per lane and trip (128 per trip: a 150 bp read takes two), fetched as the two aligned
  // words that hold them; on strand 1 they are the four that END at l - 1 - i0, byte-reversed.  Each goes through the
  // 256-entry table once; the four entries side by side in one register give the codes (stored as one word when the
  // destination is aligned), the 2-bit packing, the ambiguous count and the "not ACGTN" flag without per-character work.
  uint32_t amb_lane = 0, flags = 0;
  const int n_words = (l + 15) / 16 + 1;
  const uint32_t *bw = reinterpret_cast<const uint32_t *>(A.bases);
  const uint8_t *tb = tbl + 256u * dir;
  const uint32_t store_align = (uint32_t)(2 * off + (uint64_t)dir * (uint64_t)l) & 3u; // of codes + i0: the same for every lane
  for (int base = 0; base < n_words * 16; base += 128) {
    const int i0 = base + 4 * (int)lane;
    uint32_t E = 0;
    const int nvalid = min(max(l - i0, 0), 4);
    if (nvalid > 0) {
      const int64_t sb = (int64_t)off + (dir ? l - 4 - i0 : i0); // first byte of the four; >= -3
      const int64_t wi = sb >> 2;
      const uint32_t w0 = wi >= 0 ? __ldg(bw + wi) : 0u, w1 = __ldg(bw + wi + 1); // the buffer ends 16 bytes after the last base
      uint32_t x = __funnelshift_r(w0, w1, ((uint32_t)sb & 3u) * 8u);
      if (dir) x = __byte_perm(x, 0u, 0x0123);
      if (nvalid < 4) { const uint32_t m = (1u << (8 * nvalid)) - 1u; x = (x & m) | (0x41414141u & ~m); } // past the end: 'A'
      E = (uint32_t)tb[x & 0xFFu] | ((uint32_t)tb[(x >> 8) & 0xFFu] << 8) | ((uint32_t)tb[(x >> 16) & 0xFFu] << 16) | ((uint32_t)tb[x >> 24] << 24);
      flags |= E;
      // ambiguous incoming bases (utils.h:113): code 4 = bit 2 of the entry, characters i >= k only
      uint32_t a4 = E & 0x04040404u;
      if (i0 < k) a4 = k - i0 >= 4 ? 0u : a4 & ~((1u << (8 * (k - i0))) - 1u);
      amb_lane += __popc(a4);
      const uint32_t cw = E & 0x07070707u;
      if (nvalid == 4 && store_align == 0) *reinterpret_cast<uint32_t *>(codes + i0) = cw;
      else if (nvalid == 4 && store_align == 2) {
        *reinterpret_cast<uint16_t *>(codes + i0) = (uint16_t)cw;
        *reinterpret_cast<uint16_t *>(codes + i0 + 2) = (uint16_t)(cw >> 16);
      } else if (nvalid == 4) {
        codes[i0] = (uint8_t)cw;
        *reinterpret_cast<uint16_t *>(codes + i0 + 1) = (uint16_t)(cw >> 8);
        codes[i0 + 3] = (uint8_t)(cw >> 24);
      } else
        for (int t = 0; t < nvalid; ++t) codes[i0 + t] = (uint8_t)(cw >> (8 * t));
    }
    const uint32_t t1 = E & 0x03030303u; // ambiguous (4) hashed as A
    uint32_t v = (((t1 << 6) | (t1 >> 4) | (t1 >> 14) | (t1 >> 24)) & 0xFFu) << (24 - 8 * (int)(lane & 3u));
    v |= __shfl_xor_sync(0xffffffffu, v, 1);
    v |= __shfl_xor_sync(0xffffffffu, v, 2);
    if ((lane & 3u) == 0 && (i0 >> 4) < n_words) pw[i0 >> 4] = v;
  }
  const bool noncanon = (flags & 0x08080808u) != 0;
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
    for (int p0 = (int)lane; p0 < n_s; p0 += 128) { // four q-grams per lane with their eight loads in flight together
      uint32_t b[4], e2[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int p = p0 + 32 * u;
        b[u] = e2[u] = 0;
        if (p < n_s) {
          const uint64_t x = ((uint64_t)pw[p >> 4] << 32) | pw[(p >> 4) + 1];
          const uint32_t h = (uint32_t)(x >> (64 - 2 * (p & 15) - 2 * k)) & mask;
          b[u] = __ldg(A.lut + h);
          e2[u] = __ldg(A.lut + h + 1);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int p = p0 + 32 * u;
        if (p < n_s) { sm_begin[p] = b[u]; sm_freq[p] = e2[u] - b[u]; }
      }
    }
    __syncwarp();
    bool bad = false; // fewer than Q seeds found: the reference reads uninitialised memory
    // Zero slack (cfg3: 27 seeds per group for 9 x 3): group 0 has the most seeds, so when its table is one column wide
    // all are, and every selection is forced -- seed r of group g is the group's q-gram (r - 1) * s, read offset
    // g + (r - 1) * s * step.  No table, no traceback: the sort stage below works out the offsets and the group totals.
    const bool forced = Q * s * step <= n_s && n_s < (Q * s + 1) * step; // n_s / step - Q * s + 1 == 1
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
        uint32_t rank = 0, tot = 0, fmx = 0;
        const uint32_t lane0 = gl * (uint32_t)Q; // the group's first lane
#pragma unroll
        for (uint32_t j = 0; j < (uint32_t)Q; ++j) {
          const uint32_t fj = __shfl_sync(0xffffffffu, f, lane0 + j);
          rank += (fj < f || (fj == f && j < il)) ? 1u : 0u;
          tot += fj;
          fmx = max(fmx, fj);
        }
        const bool wrapped = fmx > 0xFFFFFFFFu / (uint32_t)Q; // else the sums of Q frequencies cannot pass 2^32
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
