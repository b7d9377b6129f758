// k_cand_small.cuh -- K3 path for sparse indexes (one or two occurrences per seed: cfg1, cfg2, cfg5):
// all groups of a read-strand at once, one warp per read-strand.
//
// Same result as k_cand / k_cand_hash / k_cand_pair and therefore as src/filter.c:45-144.  When the
// location lists of ALL step x Q selected seeds of a read-strand together hold at most CAP entries,
// the per-group machinery of the other paths (tables, bitmaps, one dependent round of loads per
// group) costs more than the work itself.  Here
//   gather   every lane takes seeds lane, lane + 32, ... and copies their shifted occurrences into one
//            shared-memory array (all groups' loads in flight together: one round trip to HBM);
//   F5       the lane that owns a group's last (most frequent) seed drops that seed's entries above
//            the maximum of the group's other entries (src/filter.c:85), or all of them when the
//            other seeds contribute nothing;
//   exact    one lane per entry counts the entries of the same group in [x, x+e] (itself included):
//            x survives additional_qgram_filter (src/filter.c:118-131) iff that count exceeds a;
//            repeated values keep their first copy only.  The count runs on the low 32 bits first
//            (an upper bound of the exact count); only entries that pass it are counted in 64 bits;
//   merge    survivors are rank-sorted by (group, value) and folded group by group into the running
//            list with the greedy "> last + e" rule (src/filter.c:45-78) by one lane;
//   emit     remove_out_ranged_candidates (src/filter.c:133-144) in one ballot (the list holds <= 32).
// Read-strands with more entries, seeds or survivors than the buffers hold go to k_cand (sort based)
// through the overflow list.
//
// Two shapes: R = 1 (step x Q <= 32 seeds: one seed per lane, cfg1 and cfg2; 2.1 KB of shared memory
// per warp) and R = 4 (up to 128 seeds, cfg5).
#pragma once
#include "common.cuh"
#include "k_cand.cuh"

template <int R> struct SmallShape {
  static constexpr uint32_t SEEDS = 32u * R;              // step * Q
  static constexpr uint32_t CAP = R == 1 ? 128u : 256u;   // entries per read-strand
  static constexpr uint32_t SC = R == 1 ? 32u : 64u;      // survivors per read-strand
  static constexpr uint32_t AC = R == 1 ? 16u : 32u;      // running candidate list (<= 32: emit is one ballot)
  __host__ __device__ static constexpr size_t bytes_per_warp() {
    return ((size_t)CAP * 8 + (SC + 2 * AC) * 8 + SEEDS * 8 + (SEEDS + 1) * 4 + SC + CAP + 15) & ~(size_t)15;
  }
};
#define SMALL_SEEDS_MAX 128u
#define SMALL_CAP_NARROW 128u

__device__ __forceinline__ void small_hand_on(const CandArgs &A, uint64_t rs, uint32_t lane) {
  if (lane == 0) {
    A.overflow_list[atomicAdd(A.overflow_count, 1u)] = (uint32_t)rs;
    A.rs_cand_first[rs] = 0;
    A.rs_cand_count[rs] = 0;
  }
}

template <int R> __global__ void __launch_bounds__(128, R == 1 ? 12 : 8) k_cand_small(CandArgs A, uint32_t inv_Q) {
  typedef SmallShape<R> Sh;
  extern __shared__ uint64_t smem_small[];
  const uint32_t lane = lane_id(), wib = threadIdx.x >> 5;
  const uint64_t rs = (uint64_t)blockIdx.x * (blockDim.x >> 5) + wib;
  if (rs >= A.n_rs) return;
  if (A.rs_status[rs] != RS_OK) {
    if (lane == 0) { A.rs_cand_first[rs] = 0; A.rs_cand_count[rs] = 0; }
    return;
  }
  // read length for the range filter: requested now, used at the very end
  const uint32_t read = (uint32_t)(rs >> 1);
  const uint64_t ro0 = A.read_off[read], ro1 = A.read_off[read + 1];
  unsigned char *base = reinterpret_cast<unsigned char *>(smem_small) + (size_t)wib * Sh::bytes_per_warp();
  uint64_t *vals = reinterpret_cast<uint64_t *>(base);       // [CAP] shifted occurrences, FEM_SENTINEL = dropped
  uint64_t *sv = vals + Sh::CAP;                              // [SC] survivors
  uint64_t *acc = sv + Sh::SC, *acc2 = acc + Sh::AC;
  uint64_t *smax = acc2 + Sh::AC;                             // [SEEDS] largest kept entry per seed + 1 (0: none); shape 4 only
  uint32_t *eo = reinterpret_cast<uint32_t *>(smax + Sh::SEEDS); // [SEEDS + 1] entry offsets of the seeds
  uint8_t *sg = reinterpret_cast<uint8_t *>(eo + Sh::SEEDS + 1); // [SC] group of each survivor
  uint8_t *eg = sg + Sh::SC;                                  // [CAP] group of each entry
  const uint32_t Q = (uint32_t)A.Q, S = Q * (uint32_t)A.step;
  const uint32_t a = (uint32_t)A.a;
  const uint32_t e32 = (uint32_t)A.e;
  const uint64_t e = (uint64_t)A.e;
  if (S > Sh::SEEDS) { small_hand_on(A, rs, lane); return; }

  // ---- gather: seeds lane, lane + 32, ... ; exclusive prefix of the frequencies over all seeds ------
  uint32_t T = 0;
  const size_t so = (size_t)rs * S;
  uint32_t sb[R], sf[R], sst[R], off[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const uint32_t s = lane + 32u * r;
    sb[r] = 0; sf[r] = 0; sst[r] = 0;
    if (s < S) { sb[r] = A.seed_begin[so + s]; sf[r] = A.seed_freq[so + s]; sst[r] = A.seed_start[so + s]; }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    off[r] = T;
    if (32u * r < S) { // uniform
      uint32_t incl = sf[r];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
        if ((int)lane >= d) incl += o;
      }
      off[r] = T + incl - sf[r];
      T += __shfl_sync(0xffffffffu, incl, 31);
    }
  }
  if (T > Sh::CAP) { small_hand_on(A, rs, lane); return; }
  // largest kept entry of each seed plus one (0: none), for F5
  uint64_t mxp[R];
  {
    // the first two entries of every seed are requested before any is used: one round trip for
    // almost all of them on a sparse index
    uint64_t o0[R], o1[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      o0[r] = 0; o1[r] = 0;
      if (sf[r] > 0) o0[r] = A.occ[(uint64_t)sb[r]];
      if (sf[r] > 1) o1[r] = A.occ[(uint64_t)sb[r] + 1];
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint32_t s = lane + 32u * r;
      mxp[r] = 0;
      if (s < S) {
        eo[s] = off[r];
        const uint8_t g = (uint8_t)(Q == 1 ? s : __umulhi(s, inv_Q)); // s / Q
        const uint32_t p = off[r];
        if (sf[r] > 0) {
          uint64_t v = FEM_SENTINEL;
          if ((uint32_t)o0[r] >= sst[r]) { v = o0[r] - sst[r]; mxp[r] = v + 1; } // filter.c:89,106; the run ascends
          vals[p] = v;
          eg[p] = g;
        }
        if (sf[r] > 1) {
          uint64_t v = FEM_SENTINEL;
          if ((uint32_t)o1[r] >= sst[r]) { v = o1[r] - sst[r]; mxp[r] = v + 1; }
          vals[p + 1] = v;
          eg[p + 1] = g;
        }
        for (uint32_t k = 2; k < sf[r]; ++k) {
          const uint64_t o = A.occ[(uint64_t)sb[r] + k];
          uint64_t v = FEM_SENTINEL;
          if ((uint32_t)o >= sst[r]) { v = o - sst[r]; mxp[r] = v + 1; }
          vals[p + k] = v;
          eg[p + k] = g;
        }
        if (R != 1) smax[s] = mxp[r];
      }
    }
    if (lane == 0) eo[S] = T;
  }
  // ---- F5: the last seed of a group only contributes entries up to the largest entry of the others; the
  //      lane that holds the last seed drops its own entries above that bound (the run ascends: walk back
  //      from its end to the first entry that stays)
  if (R == 1) {
    // one seed per lane: maximum over the Q - 1 lanes below by a sliding window of shuffles (windows of the
    // lanes that hold a group's last seed lie inside their group)
    const uint32_t W = Q - 1;
    uint64_t w = 0;
    if (W) {
      w = __shfl_up_sync(0xffffffffu, mxp[0], 1);
      if (lane < 1) w = 0;
      uint32_t win = 1;
      while (2 * win <= W) {
        uint64_t t = __shfl_up_sync(0xffffffffu, w, win);
        if (lane < win) t = 0;
        w = max(w, t);
        win *= 2;
      }
      if (win < W) {
        const uint32_t d = W - win;
        uint64_t t = __shfl_up_sync(0xffffffffu, w, d);
        if (lane < d) t = 0;
        w = max(w, t);
      }
    }
    const uint32_t s = lane;
    if (s < S && sf[0]) {
      const uint32_t g = Q == 1 ? s : __umulhi(s, inv_Q);
      if (s == g * Q + Q - 1)
        for (uint32_t k = sf[0]; k-- > 0;) {
          const uint64_t v = vals[off[0] + k];
          if (v == FEM_SENTINEL) continue;
          if (v < w) break; // v <= bound, w = bound + 1 (0: no bound, nothing stays)
          vals[off[0] + k] = FEM_SENTINEL;
        }
    }
  } else {
    __syncwarp();
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint32_t s = lane + 32u * r;
      if (s < S && sf[r]) {
        const uint32_t g = Q == 1 ? s : __umulhi(s, inv_Q);
        if (s == g * Q + Q - 1) {
          uint64_t w = 0;
          for (uint32_t i = s + 1 - Q; i < s; ++i) w = max(w, smax[i]);
          for (uint32_t k = sf[r]; k-- > 0;) {
            const uint64_t v = vals[off[r] + k];
            if (v == FEM_SENTINEL) continue;
            if (v < w) break;
            vals[off[r] + k] = FEM_SENTINEL;
          }
        }
      }
    }
  }
  __syncwarp();
  // ---- exact: one lane per entry ---------------------------------------------------------------------
  uint32_t ns = 0;
  const uint32_t *vals_lo = reinterpret_cast<const uint32_t *>(vals); // little endian: word 2j = low half of vals[j]
  for (uint32_t t0 = 0; t0 < T; t0 += 32) {
    const uint32_t t = t0 + lane;
    bool keep = false;
    uint64_t x = FEM_SENTINEL;
    uint32_t g = 0;
    if (t < T) {
      x = vals[t];
      g = eg[t];
      if (x != FEM_SENTINEL) {
        const uint32_t gs = eo[g * Q], ge = eo[g * Q + Q];
        // low halves first: y - x <= e in 64 bits implies it in 32 bits, so this count bounds the exact one
        const uint32_t xl = (uint32_t)x;
        uint32_t c32 = 0;
#pragma unroll 2
        for (uint32_t j = gs; j < ge; ++j) c32 += (vals_lo[2 * j] - xl <= e32) ? 1u : 0u;
        if (c32 > a) {
          uint32_t cnt = 0;
          bool dup = false;
          for (uint32_t j = gs; j < ge; ++j) {
            const uint64_t y = vals[j];
            cnt += (y - x <= e) ? 1u : 0u;        // FEM_SENTINEL - x is far above e
            dup = dup || (y == x && j < t);
          }
          keep = cnt > a && !dup;
        }
      }
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, keep);
    if (keep) {
      const uint32_t k = ns + __popc(bal & lanemask_lt());
      if (k < Sh::SC) { sv[k] = x; sg[k] = (uint8_t)g; }
    }
    ns += __popc(bal);
  }
  if (ns > Sh::SC) { small_hand_on(A, rs, lane); return; }
  if (ns == 0) { // nothing survives: no candidates (the usual case for the strand the read did not come from)
    if (lane == 0) { A.rs_cand_first[rs] = 0; A.rs_cand_count[rs] = 0; }
    return;
  }
  __syncwarp();
  // ---- survivors in (group, value) order, then the groups folded one after the other -----------------
  uint32_t na = 0;
  if (ns == 1) {
    if (lane == 0) acc[0] = sv[0];
    na = 1;
  } else {
    bool too_long = false;
    uint64_t x[Sh::SC / 32];
    uint32_t g[Sh::SC / 32], rank[Sh::SC / 32];
#pragma unroll
    for (int r = 0; r < (int)(Sh::SC / 32); ++r) {
      const uint32_t i = lane + 32u * r;
      x[r] = 0; g[r] = 0; rank[r] = 0;
      if (i < ns) {
        x[r] = sv[i]; g[r] = sg[i];
        for (uint32_t j = 0; j < ns; ++j) {
          const uint32_t gj = sg[j];
          rank[r] += (gj < g[r] || (gj == g[r] && sv[j] < x[r])) ? 1u : 0u; // (group, value) pairs are distinct
        }
      }
    }
    __syncwarp();
#pragma unroll
    for (int r = 0; r < (int)(Sh::SC / 32); ++r)
      if (lane + 32u * r < ns) { sv[rank[r]] = x[r]; sg[rank[r]] = (uint8_t)g[r]; }
    __syncwarp();
    if (lane == 0) {
      uint32_t i2 = 0;
      while (i2 < ns && !too_long) {
        const uint32_t gcur = sg[i2];
        uint32_t end = i2;
        while (end < ns && sg[end] == gcur) ++end;
        uint32_t i1 = 0, no = 0;
        uint64_t last = 0;
        while (i1 < na || i2 < end) { // filter.c:45-78
          uint64_t v;
          if (i1 < na && (i2 >= end || acc[i1] < sv[i2])) v = acc[i1++];
          else v = sv[i2++];
          if (no == 0 || v > last + e) {
            if (no < Sh::AC) acc2[no] = v;
            ++no;
            last = v;
          }
        }
        if (no > Sh::AC) { too_long = true; break; }
        na = no;
        uint64_t *sw = acc; acc = acc2; acc2 = sw;
      }
    }
    too_long = __shfl_sync(0xffffffffu, (int)too_long, 0) != 0;
    na = __shfl_sync(0xffffffffu, na, 0);
    // lane 0 may have swapped the two lists an odd number of times
    const uint32_t swapped = __shfl_sync(0xffffffffu, (uint32_t)(acc != sv + Sh::SC), 0);
    acc = swapped ? sv + Sh::SC + Sh::AC : sv + Sh::SC;
    if (too_long) { small_hand_on(A, rs, lane); return; }
  }
  __syncwarp();
  // ---- emit: the range filter and the arena slice (same outputs as rs_emit_candidates; na <= 32) ------
  {
    const uint32_t l = (uint32_t)(ro1 - ro0);
    bool keep = false;
    uint64_t v = 0;
    if (lane < na) {
      v = acc[lane];
      const uint32_t pos = (uint32_t)v, len = A.rec_len[v >> 32];
      keep = pos >= e32 && pos + l + e32 < len;
    }
    const uint32_t m = __ballot_sync(0xffffffffu, keep);
    const uint32_t n_keep = __popc(m);
    unsigned long long first = 0;
    if (lane == 0 && n_keep) first = atomicAdd(A.cand_count, (unsigned long long)n_keep);
    first = __shfl_sync(0xffffffffu, first, 0);
    if (n_keep && first + n_keep > A.cand_cap) {
      if (lane == 0) { *A.error_flag = 1; A.rs_cand_first[rs] = 0; A.rs_cand_count[rs] = 0; }
      return;
    }
    if (keep) {
      const uint64_t o = first + __popc(m & lanemask_lt());
      A.cand[o] = v - e; // the stored candidate is the band's left edge
      A.cand_rs[o] = (uint32_t)rs;
    }
    if (lane == 0) { A.rs_cand_first[rs] = first; A.rs_cand_count[rs] = n_keep; }
  }
}
