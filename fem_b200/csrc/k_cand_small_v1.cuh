// k_cand_small_v1.cuh -- previous form of k_cand_small, kept for one A/B run on the box (small_shape = 9).
// K3 path for sparse indexes (one or two occurrences per seed: cfg1, cfg2, cfg5):
// all groups of a read-strand at once, one warp per read-strand.
//
// Same result as k_cand / k_cand_hash / k_cand_pair and therefore as src/filter.c:45-144.  When the
// location lists of ALL step x Q selected seeds of a read-strand together hold at most V1_SMALL_CAP
// entries, the per-group machinery of the other paths (tables, bitmaps, one dependent round of loads
// per group) costs more than the work itself.  Here
//   gather   every lane takes seeds lane, lane + 32, ... and copies their shifted occurrences into one
//            shared-memory array (all groups' loads in flight together: one round trip to HBM);
//   F5       the lane that owns a group's last (most frequent) seed drops that seed's entries above
//            the maximum of the group's other entries (src/filter.c:85), or the whole group when the
//            other seeds contribute nothing;
//   exact    one lane per entry counts the entries of the same group in [x, x+e] (itself included):
//            x survives additional_qgram_filter (src/filter.c:118-131) iff that count exceeds a;
//            repeated values keep their first copy only;
//   merge    survivors are rank-sorted by (group, value) and folded group by group into the running
//            list with the greedy "> last + e" rule (src/filter.c:45-78) by one lane.
// Read-strands with more entries, seeds or survivors than the buffers hold go to k_cand (sort based)
// through the overflow list.
#pragma once
#include "common.cuh"
#include "k_cand.cuh"

#define V1_SMALL_CAP 256u   // entries per read-strand
#define V1_SMALL_SEEDS 128u // step * Q
#define V1_SMALL_SC 64u     // survivors per read-strand
#define V1_SMALL_AC 32u     // running candidate list

__host__ __device__ inline size_t cand_small_v1_smem_bytes_per_warp() {
  return (size_t)V1_SMALL_CAP * 8 + (V1_SMALL_SC + 2 * V1_SMALL_AC) * 8 + (V1_SMALL_SEEDS + 1) * 4 + V1_SMALL_SEEDS * 8 + V1_SMALL_SC + 64 + V1_SMALL_CAP + 12;
}

__global__ void __launch_bounds__(128) k_cand_small_v1(CandArgs A) {
  extern __shared__ uint64_t smem_small_v1[];
  const uint32_t lane = lane_id(), wib = threadIdx.x >> 5;
  const uint64_t rs = (uint64_t)blockIdx.x * (blockDim.x >> 5) + wib;
  if (rs >= A.n_rs) return;
  if (A.rs_status[rs] != RS_OK) {
    if (lane == 0) { A.rs_cand_first[rs] = 0; A.rs_cand_count[rs] = 0; }
    return;
  }
  unsigned char *base = reinterpret_cast<unsigned char *>(smem_small_v1) + (size_t)wib * ((cand_small_v1_smem_bytes_per_warp() + 15) & ~(size_t)15);
  uint64_t *vals = reinterpret_cast<uint64_t *>(base);       // [V1_SMALL_CAP] shifted occurrences, FEM_SENTINEL = dropped
  uint64_t *sv = vals + V1_SMALL_CAP;                            // [V1_SMALL_SC] survivors
  uint64_t *acc = sv + V1_SMALL_SC, *acc2 = acc + V1_SMALL_AC;
  uint64_t *smax = acc2 + V1_SMALL_AC;                           // [V1_SMALL_SEEDS] largest kept entry per seed (FEM_SENTINEL: none)
  uint32_t *eo = reinterpret_cast<uint32_t *>(smax + V1_SMALL_SEEDS); // [V1_SMALL_SEEDS + 1] entry offsets of the seeds
  uint8_t *sg = reinterpret_cast<uint8_t *>(eo + V1_SMALL_SEEDS + 1); // [V1_SMALL_SC] group of each survivor
  uint8_t *dead = sg + V1_SMALL_SC;                              // [64] group contributes nothing
  uint8_t *eg = dead + 64;                                    // [V1_SMALL_CAP] group of each entry
  const uint32_t Q = (uint32_t)A.Q, step = (uint32_t)A.step, S = Q * step;
  const uint32_t a = (uint32_t)A.a;
  const uint64_t e = (uint64_t)A.e;
  bool overflow = S > V1_SMALL_SEEDS || step > 64;

  // ---- gather: seeds lane, lane + 32, ... ; exclusive prefix of the frequencies over all seeds ------
  uint32_t T = 0;
  if (!overflow) {
    const size_t so = (size_t)rs * S;
    uint32_t sb[4], sf[4], sst[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const uint32_t s = lane + 32u * r;
      sb[r] = 0; sf[r] = 0; sst[r] = 0;
      if (s < S) { sb[r] = A.seed_begin[so + s]; sf[r] = A.seed_freq[so + s]; sst[r] = A.seed_start[so + s]; }
    }
    uint32_t off[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      uint32_t incl = sf[r];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
        if ((int)lane >= d) incl += o;
      }
      off[r] = T + incl - sf[r];
      T += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (T > V1_SMALL_CAP) overflow = true;
    else {
      // the first two entries of every seed are requested before any is used: one round trip for
      // almost all of them on a sparse index
      uint64_t o0[4], o1[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        o0[r] = 0; o1[r] = 0;
        if (sf[r] > 0) o0[r] = A.occ[(uint64_t)sb[r]];
        if (sf[r] > 1) o1[r] = A.occ[(uint64_t)sb[r] + 1];
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const uint32_t s = lane + 32u * r;
        if (s < S) {
          eo[s] = off[r];
          const uint8_t g = (uint8_t)(s / Q);
          uint64_t mx = FEM_SENTINEL;
          for (uint32_t k = 0; k < sf[r]; ++k) {
            const uint64_t o = k == 0 ? o0[r] : k == 1 ? o1[r] : A.occ[(uint64_t)sb[r] + k];
            uint64_t v = FEM_SENTINEL;
            if ((uint32_t)o >= sst[r]) { v = o - sst[r]; mx = v; } // filter.c:89,106; the run ascends
            vals[off[r] + k] = v;
            eg[off[r] + k] = g;
          }
          smax[s] = mx;
        }
      }
      if (lane == 0) eo[S] = T;
    }
  }
  if (overflow) {
    if (lane == 0) {
      A.overflow_list[atomicAdd(A.overflow_count, 1u)] = (uint32_t)rs;
      A.rs_cand_first[rs] = 0;
      A.rs_cand_count[rs] = 0;
    }
    return;
  }
  __syncwarp();
  // ---- F5: the last seed of a group only contributes entries up to the largest entry of the others: one
  //      lane per group finds that bound (kept in sv, free until the count), one lane per entry applies it
  for (uint32_t g = lane; g < step; g += 32) {
    uint64_t mv = 0;
    bool has = false;
    for (uint32_t i = 0; i + 1 < Q; ++i) {
      const uint64_t m = smax[g * Q + i];
      if (m != FEM_SENTINEL && (!has || m > mv)) { mv = m; has = true; }
    }
    dead[g] = has ? 0 : 1;
    sv[g] = mv;
  }
  __syncwarp();
  for (uint32_t t = lane; t < T; t += 32) {
    const uint32_t g = eg[t];
    if (t >= eo[g * Q + Q - 1] && (dead[g] || (vals[t] != FEM_SENTINEL && vals[t] > sv[g]))) vals[t] = FEM_SENTINEL;
  }
  __syncwarp();
  // ---- exact: one lane per entry ---------------------------------------------------------------------
  uint32_t ns = 0;
  for (uint32_t t0 = 0; t0 < T; t0 += 32) {
    const uint32_t t = t0 + lane;
    bool keep = false;
    uint64_t x = FEM_SENTINEL;
    uint32_t g = 0;
    if (t < T) {
      x = vals[t];
      g = eg[t];
      if (x != FEM_SENTINEL && !dead[g]) {
        const uint32_t gs = eo[g * Q], ge = eo[g * Q + Q];
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
    const uint32_t bal = __ballot_sync(0xffffffffu, keep);
    if (keep) {
      const uint32_t k = ns + __popc(bal & lanemask_lt());
      if (k < V1_SMALL_SC) { sv[k] = x; sg[k] = (uint8_t)g; }
    }
    ns += __popc(bal);
  }
  if (ns > V1_SMALL_SC) {
    if (lane == 0) {
      A.overflow_list[atomicAdd(A.overflow_count, 1u)] = (uint32_t)rs;
      A.rs_cand_first[rs] = 0;
      A.rs_cand_count[rs] = 0;
    }
    return;
  }
  __syncwarp();
  // ---- survivors in (group, value) order, then the groups folded one after the other -----------------
  uint32_t na = 0;
  bool too_long = false;
  if (ns) {
    uint64_t x[2];
    uint32_t g[2], rank[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
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
    for (int r = 0; r < 2; ++r)
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
            if (no < V1_SMALL_AC) acc2[no] = v;
            ++no;
            last = v;
          }
        }
        if (no > V1_SMALL_AC) { too_long = true; break; }
        na = no;
        uint64_t *sw = acc; acc = acc2; acc2 = sw;
      }
    }
    too_long = __shfl_sync(0xffffffffu, (int)too_long, 0) != 0;
    na = __shfl_sync(0xffffffffu, na, 0);
    // lane 0 may have swapped the two lists an odd number of times
    const uint32_t swapped = __shfl_sync(0xffffffffu, (uint32_t)(acc != sv + V1_SMALL_SC), 0);
    acc = swapped ? sv + V1_SMALL_SC + V1_SMALL_AC : sv + V1_SMALL_SC;
  }
  if (too_long) {
    if (lane == 0) {
      A.overflow_list[atomicAdd(A.overflow_count, 1u)] = (uint32_t)rs;
      A.rs_cand_first[rs] = 0;
      A.rs_cand_count[rs] = 0;
    }
    return;
  }
  __syncwarp();
  rs_emit_candidates(A, rs, lane, acc, na);
}
