// k_cand_hash.cuh -- K3 fast path: candidate generation for read-strands whose location lists are
// short (the common case on non-repetitive sequence), one warp per read-strand, no sorting.
//
// Same result as k_cand (k_cand.cuh) and therefore as src/filter.c:80-144, from the observation
// that only the SET of surviving values of a group matters downstream:
//   * additional_qgram_filter keeps v[i] iff v[i+a] <= v[i]+e in the sorted multiset; for a value
//     x that holds for its first copy iff the multiset has at least a+1 elements in [x, x+e], and
//     later copies add nothing because
//   * the cross-group pass (merge_kvec_t_uint64_t) keeps a value only if it exceeds the last kept
//     value by more than e, which drops every duplicate.
// So the warp builds a small open-addressing table (value -> multiplicity) of the group's shifted
// locations in shared memory, asks for every distinct value how many entries lie in [x, x+e]
// (its own 8-wide bucket and the next one), and sorts only the few survivors.  The last (most
// frequent) seed's entries count only up to the maximum of the other seeds' entries.
// Read-strands that exceed the table, the survivor list or the running list go to k_cand (sort
// based) through the overflow list, and from there to k_cand_big if needed.
#pragma once
#include "common.cuh"
#include "k_cand.cuh"

#define HASH_SC 32u // survivors per group
#define HASH_AC 32u // running candidate list

__host__ __device__ inline size_t cand_hash_smem_bytes_per_warp(uint32_t T) {
  return (size_t)T * 8 + (size_t)T * 4 + (HASH_SC + 2 * HASH_AC) * 8 + 32 * 4;
}

__device__ __forceinline__ uint32_t hash_bucket(uint64_t bucket, uint32_t shift) {
  return (uint32_t)((bucket * 0x9E3779B97F4A7C15ull) >> shift);
}

__device__ __forceinline__ void table_insert(uint64_t *tab, uint32_t *cnt, uint32_t mask, uint32_t shift, uint64_t x) {
  uint32_t slot = hash_bucket(x >> 3, shift);
  for (;;) {
    const unsigned long long old = atomicCAS((unsigned long long *)&tab[slot], (unsigned long long)FEM_SENTINEL, (unsigned long long)x);
    if (old == FEM_SENTINEL || old == x) { atomicAdd(&cnt[slot], 1u); return; }
    slot = (slot + 1) & mask;
  }
}

// entries with value in [x, x+e] whose bucket is `bucket`
__device__ __forceinline__ uint32_t table_count(const uint64_t *tab, const uint32_t *cnt, uint32_t mask, uint32_t shift,
                                                uint64_t bucket, uint64_t x, uint64_t e) {
  uint32_t slot = hash_bucket(bucket, shift), n = 0;
  for (;;) {
    const uint64_t y = tab[slot];
    if (y == FEM_SENTINEL) return n;
    if ((y >> 3) == bucket && y - x <= e) n += cnt[slot];
    slot = (slot + 1) & mask;
  }
}

__device__ __forceinline__ uint64_t warp_max_u64(uint64_t v) {
  const uint32_t hi = (uint32_t)(v >> 32);
  const uint32_t mhi = __reduce_max_sync(0xffffffffu, hi);
  const uint32_t lo = hi == mhi ? (uint32_t)v : 0u;
  return ((uint64_t)mhi << 32) | __reduce_max_sync(0xffffffffu, lo);
}

__global__ void __launch_bounds__(128) k_cand_hash(CandArgs A, uint32_t T, uint32_t shift) {
  extern __shared__ uint64_t smem_hash[];
  const uint32_t lane = lane_id(), wib = threadIdx.x >> 5;
  const uint64_t rs = (uint64_t)blockIdx.x * (blockDim.x >> 5) + wib;
  if (rs >= A.n_rs) return;
  if (A.rs_status[rs] != RS_OK) {
    if (lane == 0) { A.rs_cand_first[rs] = 0; A.rs_cand_count[rs] = 0; }
    return;
  }
  unsigned char *base = reinterpret_cast<unsigned char *>(smem_hash) + (size_t)wib * cand_hash_smem_bytes_per_warp(T);
  uint64_t *tab = reinterpret_cast<uint64_t *>(base);
  uint64_t *surv = tab + T;
  uint64_t *acc = surv + HASH_SC, *acc2 = acc + HASH_AC;
  uint32_t *cnt = reinterpret_cast<uint32_t *>(acc2 + HASH_AC);
  uint32_t *s_b = cnt + T, *s_st = s_b + 10, *s_incl = s_st + 10; // per-seed run begin, read offset, inclusive prefix
  const uint32_t mask = T - 1;
  const int Q = A.Q, a = A.a;
  const uint64_t e = (uint64_t)A.e;
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
    if (n_tot > T / 2) { overflow = true; break; }
    if ((int)lane < Q) { s_b[lane] = b; s_st[lane] = st; s_incl[lane] = incl; }
    for (uint32_t t = lane; t < T; t += 32) { tab[t] = FEM_SENTINEL; cnt[t] = 0; }
    __syncwarp();
    // ---- runs 0..Q-2: one flat index space, all loads of a round in flight together ------------
    uint64_t mv = 0;
    bool has = false;
    for (uint32_t j0 = 0; j0 < n_first; j0 += 32) {
      const uint32_t j = j0 + lane;
      if (j < n_first) {
        int i = 0;
        while (s_incl[i] <= j) ++i;
        const uint32_t first = i ? s_incl[i - 1] : 0u, sti = s_st[i];
        const uint64_t o = A.occ[(uint64_t)s_b[i] + (j - first)];
        if ((uint32_t)o >= sti) { // filter.c:89,106
          const uint64_t v = o - sti;
          table_insert(tab, cnt, mask, shift, v);
          if (!has || v > mv) mv = v;
          has = true;
        }
      }
    }
    const bool any = __any_sync(0xffffffffu, has);
    if (!any) continue; // nothing to merge the last run into
    mv = warp_max_u64(has ? mv : 0ull);
    // ---- last run: entries up to that maximum ----------------------------------------------------
    {
      const uint32_t bl = s_b[Q - 1], stl = s_st[Q - 1], fl = n_tot - n_first;
      for (uint32_t j = lane; j < fl; j += 32) {
        const uint64_t o = A.occ[(uint64_t)bl + j];
        if ((uint32_t)o >= stl && o - stl <= mv) table_insert(tab, cnt, mask, shift, o - stl);
      }
    }
    __syncwarp();
    // ---- distinct values with at least a+1 entries in [x, x+e] -----------------------------------
    uint32_t ns = 0;
    for (uint32_t t0 = 0; t0 < T; t0 += 32) {
      const uint64_t x = tab[t0 + lane];
      bool keep = false;
      if (x != FEM_SENTINEL) {
        uint32_t n = table_count(tab, cnt, mask, shift, x >> 3, x, e);
        if (n <= (uint32_t)a && ((x + e) >> 3) != (x >> 3)) n += table_count(tab, cnt, mask, shift, (x >> 3) + 1, x, e);
        keep = n > (uint32_t)a;
      }
      const uint32_t m = __ballot_sync(0xffffffffu, keep);
      if (keep) {
        const uint32_t idx = ns + __popc(m & lanemask_lt());
        if (idx < HASH_SC) surv[idx] = x;
      }
      ns += __popc(m);
    }
    if (ns > HASH_SC) { overflow = true; break; }
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
          if (no < HASH_AC) acc2[no] = v;
          ++no;
          last = v;
        }
      }
    }
    no = __shfl_sync(0xffffffffu, no, 0);
    if (no > HASH_AC) { overflow = true; break; }
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
