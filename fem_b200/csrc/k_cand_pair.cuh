// k_cand_pair.cuh -- K3 main path for a >= 1: candidate generation by near-pair detection, one warp
// per read-strand, no sorting and no value table.
//
// Same result as k_cand / k_cand_hash and therefore as src/filter.c:80-144.  With a >= 1 additional
// q-grams a shifted location x of a group survives additional_qgram_filter (src/filter.c:118-131)
// iff the group's multiset holds at least a other entries in [x, x+e]; every such witness lies
// within e of x.  So only entries that have SOME other entry within e can matter, as survivor or
// as witness, and on a non-repetitive reference those are the handful of entries at the read's true
// location out of hundreds.  The warp therefore
//   pass 1  gathers the group's occurrence runs (all loads of a 128-entry chunk in flight together),
//           keeps the shifted values in shared memory and marks bit hash(x >> 3) of bitmap A
//           (bitmap B: the bit was already set -- two entries in one 8-wide bucket);
//   pass 2  flags an entry iff B holds its own bucket or A holds bucket-1 or bucket+1: entries within
//           e <= 7 of each other sit in the same or in adjacent buckets, so every entry with a
//           neighbour is flagged; hash collisions only add a few false flags;
//   exact   counts, among the flagged entries only, the others in [x, x+e] and drops duplicates;
// and the survivors go through the same greedy cross-group merge as in k_cand_hash.  The last
// (most frequent) seed's run counts only up to the maximum of the other runs.
// Read-strands whose group exceeds the value buffer, the flagged list, the survivor list or the
// running list are handed to k_cand (sort based) through the overflow list.
#pragma once
#include "common.cuh"
#include "k_cand.cuh"
#include "k_cand_hash.cuh" // warp_max_u64

#define PAIR_FC 64u // flagged entries per group
#define PAIR_SC 32u // survivors per group
#define PAIR_AC 32u // running candidate list

// M bits of bitmap A (power of two >= 1024), M/4 bits of bitmap B, VC values
__host__ __device__ inline size_t cand_pair_smem_bytes_per_warp(uint32_t M, uint32_t VC) {
  return (size_t)VC * 8 + (PAIR_FC + PAIR_SC + 2 * PAIR_AC) * 8 + M / 8 + M / 32 + 32 * 4;
}

__device__ __forceinline__ uint32_t pair_hash_base(uint64_t v) {
  return (((uint32_t)v >> 3) + (uint32_t)(v >> 32) * 0x9E3779B1u) * 0x85EBCA6Bu;
}

__global__ void __launch_bounds__(128) k_cand_pair(CandArgs A, uint32_t M, uint32_t shift, uint32_t VC) {
  extern __shared__ uint64_t smem_pair[];
  const uint32_t lane = lane_id(), wib = threadIdx.x >> 5;
  const uint64_t rs = (uint64_t)blockIdx.x * (blockDim.x >> 5) + wib;
  if (rs >= A.n_rs) return;
  if (A.rs_status[rs] != RS_OK) {
    if (lane == 0) { A.rs_cand_first[rs] = 0; A.rs_cand_count[rs] = 0; }
    return;
  }
  unsigned char *base = reinterpret_cast<unsigned char *>(smem_pair) + (size_t)wib * cand_pair_smem_bytes_per_warp(M, VC);
  uint64_t *vals = reinterpret_cast<uint64_t *>(base);
  uint64_t *flist = vals + VC;
  uint64_t *surv = flist + PAIR_FC;
  uint64_t *acc = surv + PAIR_SC, *acc2 = acc + PAIR_AC;
  uint32_t *bmA = reinterpret_cast<uint32_t *>(acc2 + PAIR_AC);
  uint32_t *bmB = bmA + M / 32;
  uint32_t *s_b = bmB + M / 128, *s_st = s_b + 10, *s_incl = s_st + 10;
  const int Q = A.Q;
  const uint32_t a = (uint32_t)A.a;
  const uint64_t e = (uint64_t)A.e;
  const uint32_t K2 = 0x85EBCA6Bu;
  uint32_t na = 0;
  bool overflow = false;

  for (int g = 0; g < A.step && !overflow; ++g) {
    const size_t so = ((size_t)rs * A.step + g) * Q;
    uint32_t b = 0, f = 0, st = 0;
    if ((int)lane < Q) { b = A.seed_begin[so + lane]; f = A.seed_freq[so + lane]; st = A.seed_start[so + lane]; }
    uint32_t incl = f;
#pragma unroll
    for (int d = 1; d < 16; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
      if ((int)lane >= d) incl += o;
    }
    const uint32_t n_tot = __shfl_sync(0xffffffffu, incl, Q - 1);
    const uint32_t n_first = Q > 1 ? __shfl_sync(0xffffffffu, incl, Q - 2) : 0u; // all runs but the last
    if (n_first == 0) continue;                 // the last run alone contributes nothing (F5)
    if (n_tot > VC) { overflow = true; break; }
    __syncwarp();
    if ((int)lane < Q) { s_b[lane] = b; s_st[lane] = st; s_incl[lane] = incl; }
    {
      uint4 *z = reinterpret_cast<uint4 *>(bmA); // A and B are contiguous: (M/32 + M/128) words
      for (uint32_t t = lane; t < (M / 32 + M / 128) / 4; t += 32) z[t] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncwarp();
    // ---- pass 1, runs 0..Q-2: one flat index space, four rounds of loads in flight ---------------
    uint64_t mv = 0;
    bool has = false;
    uint32_t i0 = 0; // run that holds flat index j0 (warp uniform)
    for (uint32_t j0 = 0; j0 < n_first; j0 += 128) {
      while (s_incl[i0] <= j0) ++i0;
      uint64_t o[4];
      uint32_t sti[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t j = j0 + 32u * u + lane;
        o[u] = 0; sti[u] = 1; // (uint32)0 < 1: dropped
        if (j < n_first) {
          uint32_t i = i0;
          while (s_incl[i] <= j) ++i;
          const uint32_t first = i ? s_incl[i - 1] : 0u;
          sti[u] = s_st[i];
          o[u] = A.occ[(uint64_t)s_b[i] + (j - first)];
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t j = j0 + 32u * u + lane;
        if (j < n_first) {
          uint64_t v = FEM_SENTINEL;
          if ((uint32_t)o[u] >= sti[u]) { // filter.c:89,106
            v = o[u] - sti[u];
            const uint32_t h = pair_hash_base(v) >> shift, bit = 1u << (h & 31u);
            if (atomicOr(&bmA[h >> 5], bit) & bit) atomicOr(&bmB[h >> 7], 1u << ((h >> 2) & 31u));
            if (!has || v > mv) mv = v;
            has = true;
          }
          vals[j] = v;
        }
      }
    }
    const bool any = __any_sync(0xffffffffu, has);
    if (!any) continue; // nothing to merge the last run into
    mv = warp_max_u64(has ? mv : 0ull);
    // ---- pass 1, last run: entries up to that maximum (ascending, so stop at the first larger one) --
    uint32_t n_all = n_first;
    {
      const uint32_t bl = s_b[Q - 1], stl = s_st[Q - 1], fl = n_tot - n_first;
      for (uint32_t j0 = 0; j0 < fl; j0 += 32) {
        const uint32_t j = j0 + lane;
        bool beyond = false;
        if (j < fl) {
          const uint64_t o = A.occ[(uint64_t)bl + j];
          uint64_t v = FEM_SENTINEL;
          if ((uint32_t)o >= stl) {
            if (o - stl <= mv) {
              v = o - stl;
              const uint32_t h = pair_hash_base(v) >> shift, bit = 1u << (h & 31u);
              if (atomicOr(&bmA[h >> 5], bit) & bit) atomicOr(&bmB[h >> 7], 1u << ((h >> 2) & 31u));
            } else beyond = true;
          }
          vals[n_first + j] = v;
        }
        n_all = n_first + min(fl, j0 + 32u);
        if (__any_sync(0xffffffffu, beyond)) break;
      }
    }
    __syncwarp();
    // ---- pass 2: entries whose own bucket holds two entries or whose neighbour bucket holds one -----
    uint32_t nf = 0;
    for (uint32_t j0 = 0; j0 < n_all; j0 += 32) {
      const uint32_t j = j0 + lane;
      const uint64_t v = j < n_all ? vals[j] : FEM_SENTINEL;
      bool flag = false;
      if (v != FEM_SENTINEL) {
        const uint32_t m = pair_hash_base(v);
        const uint32_t h0 = m >> shift, hm = (m - K2) >> shift, hp = (m + K2) >> shift;
        flag = (((bmB[h0 >> 7] >> ((h0 >> 2) & 31u)) | (bmA[hm >> 5] >> (hm & 31u)) | (bmA[hp >> 5] >> (hp & 31u))) & 1u) != 0;
      }
      const uint32_t bal = __ballot_sync(0xffffffffu, flag);
      if (flag) {
        const uint32_t idx = nf + __popc(bal & lanemask_lt());
        if (idx < PAIR_FC) flist[idx] = v;
      }
      nf += __popc(bal);
    }
    if (nf > PAIR_FC) { overflow = true; break; }
    if (nf < a + 1) continue; // a survivor needs a other flagged entries
    __syncwarp();
    // ---- exact: distinct flagged values with at least a other flagged entries in [x, x+e] -----------
    uint32_t ns = 0;
    for (uint32_t t0 = 0; t0 < nf; t0 += 32) {
      const uint32_t idx = t0 + lane;
      const uint64_t x = idx < nf ? flist[idx] : 0ull;
      uint32_t cnt = 0;
      bool dup = false;
      for (uint32_t j = 0; j < nf; ++j) {
        const uint64_t y = flist[j];
        if (j != idx) {
          cnt += (y - x <= e) ? 1u : 0u;
          dup = dup || (y == x && j < idx);
        }
      }
      const bool keep = idx < nf && cnt >= a && !dup;
      const uint32_t bal = __ballot_sync(0xffffffffu, keep);
      if (keep) {
        const uint32_t k = ns + __popc(bal & lanemask_lt());
        if (k < PAIR_SC) surv[k] = x;
      }
      ns += __popc(bal);
    }
    if (ns > PAIR_SC) { overflow = true; break; }
    if (ns == 0) continue;
    __syncwarp();
    // ---- sort the survivors (rank by counting), then the greedy "> last + e" merge with the
    //      running list (filter.c:45-78), walked by one lane: both lists are a handful long ----
    uint64_t x = 0;
    uint32_t rank = 0;
    if (lane < ns) {
      x = surv[lane];
      for (uint32_t j = 0; j < ns; ++j) rank += surv[j] < x; // values are distinct
    }
    __syncwarp();
    if (lane < ns) surv[rank] = x;
    __syncwarp();
    uint32_t no = 0;
    if (lane == 0) {
      uint32_t i1 = 0, i2 = 0;
      uint64_t last = 0;
      while (i1 < na || i2 < ns) {
        uint64_t v;
        if (i1 < na && (i2 >= ns || acc[i1] < surv[i2])) v = acc[i1++];
        else v = surv[i2++];
        if (no == 0 || v > last + e) {
          if (no < PAIR_AC) acc2[no] = v;
          ++no;
          last = v;
        }
      }
    }
    no = __shfl_sync(0xffffffffu, no, 0);
    if (no > PAIR_AC) { overflow = true; break; }
    na = no;
    uint64_t *sw = acc; acc = acc2; acc2 = sw;
    __syncwarp();
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
  rs_emit_candidates(A, rs, lane, acc, na);
}
