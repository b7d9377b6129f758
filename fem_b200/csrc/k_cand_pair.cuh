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
//           marks the entry's 8-wide bucket in bitmap A (bitmap B: the bit was already set -- two
//           entries in one bucket) and keeps only the 16-bit bitmap position per entry;
//   flags   F = B | (A & (A << 1 | A >> 1)), word-wise: entries within e <= 7 of each other sit in
//           the same or in adjacent buckets, and adjacent buckets are adjacent bits, so every entry
//           with a neighbour has its bucket in F; aliasing only adds a few false flags;
//   pass 2  collects the entries whose bucket is in F and loads their values again (cache hits);
//   exact   counts, among the flagged entries only, those in [x, x+e];
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
#define PAIR_DROPPED 0xFFFFu

// M bits per bitmap (power of two, 1024..32768), VC entries per group (multiple of 64)
__host__ __device__ inline size_t cand_pair_smem_bytes_per_warp(uint32_t M, uint32_t VC) {
  return (size_t)M / 8 * 2 + (PAIR_SC + 2 * PAIR_AC) * 8 + PAIR_FC * 8 + 48 * 4 + (size_t)VC * 2 + PAIR_FC * 2;
}

// Bitmap position of the 8-wide bucket of shifted location v.  Adjacent buckets of one record map
// to adjacent positions (mod M), so "bucket and a neighbour bucket both occupied" is a shift-and-AND
// over bitmap words; the record number only offsets the positions.
__device__ __forceinline__ uint32_t pair_bucket(uint64_t v, uint32_t mask) {
  return (((uint32_t)v >> 3) + (uint32_t)(v >> 32) * 0x9E3779B1u) & mask;
}

__global__ void __launch_bounds__(128) k_cand_pair(CandArgs A, uint32_t M, uint32_t VC) {
  extern __shared__ uint64_t smem_pair[];
  const uint32_t lane = lane_id(), wib = threadIdx.x >> 5;
  uint64_t rs = (uint64_t)blockIdx.x * (blockDim.x >> 5) + wib;
  if (A.rs_list) { // list mode: the read-strands k_cand_reg handed on; the list's length may still stand on the device
    rs += A.list_first;
    if (rs >= A.n_list || (A.n_list_dev && rs >= *A.n_list_dev)) return;
    rs = A.rs_list[rs];
  } else if (rs >= A.n_rs) return;
  if (A.rs_status[rs] != RS_OK) {
    if (lane == 0) { A.rs_cand_first[rs] = 0; A.rs_cand_count[rs] = 0; }
    return;
  }
  unsigned char *base = reinterpret_cast<unsigned char *>(smem_pair) + (size_t)wib * cand_pair_smem_bytes_per_warp(M, VC);
  uint32_t *bmA = reinterpret_cast<uint32_t *>(base);         // occupied buckets
  uint32_t *bmB = bmA + M / 32;                                // buckets with two entries, then the flag map F
  uint64_t *surv = reinterpret_cast<uint64_t *>(bmB + M / 32);
  uint64_t *acc = surv + PAIR_SC, *acc2 = acc + PAIR_AC;
  uint64_t *fv = acc2 + PAIR_AC;                               // values of the flagged entries
  uint32_t *s_b = reinterpret_cast<uint32_t *>(fv + PAIR_FC), *s_st = s_b + 16, *s_incl = s_st + 16; // 16 slots each; s_incl padded with ~0
  uint16_t *hst = reinterpret_cast<uint16_t *>(s_incl + 16);            // bitmap position per entry
  uint16_t *fj = hst + VC;                                              // flat indices of the flagged entries
  const int Q = A.Q;
  const uint32_t a = (uint32_t)A.a, e32 = (uint32_t)A.e, mask = M - 1, wmask = M / 32 - 1;
  const uint64_t e = (uint64_t)A.e;
  uint32_t na = 0;
  bool overflow = false;
  if (lane < 16) s_incl[lane] = 0xFFFFFFFFu;
  // pass 1 decomposition: the Q-1 first runs side by side, lpr lanes each (lane -> run myrun, entry
  // sub, sub + lpr, ...): every lane keeps its run's begin / length / shift in registers
  const uint32_t R = (uint32_t)Q - 1u, lpr = R ? 32u / R : 32u;
  const uint32_t myrun = lane / lpr, sub = lane - myrun * lpr;
  const bool runner = myrun < R;
  // seeds of the next group are fetched while the current one is processed
  uint32_t nb = 0, nfreq = 0, nst = 0;
  if ((int)lane < Q) {
    const size_t so = (size_t)rs * A.step * Q;
    nb = A.seed_begin[so + lane]; nfreq = A.seed_freq[so + lane]; nst = A.seed_start[so + lane];
  }

  for (int g = 0; g < A.step && !overflow; ++g) {
    const uint32_t b = nb, f = nfreq, st = nst;
    if (g + 1 < A.step && (int)lane < Q) {
      const size_t so = ((size_t)rs * A.step + g + 1) * Q;
      nb = A.seed_begin[so + lane]; nfreq = A.seed_freq[so + lane]; nst = A.seed_start[so + lane];
    }
    uint32_t incl = (int)lane < Q ? f : 0u;
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
      uint4 *z = reinterpret_cast<uint4 *>(bmA); // A and B are contiguous: 2 * M / 32 words
      for (uint32_t t = lane; t < M / 64; t += 32) z[t] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncwarp();
    // ---- pass 1, runs 0..Q-2 side by side, four rounds of loads in flight ---------------------------
    // run registers of this lane; with lpr a multiple of 4 every lane group reads one aligned 32-byte
    // sector per round (the first round of a run starts below its begin by begin mod 4 entries)
    const uint32_t rb = __shfl_sync(0xffffffffu, b, myrun), rf = __shfl_sync(0xffffffffu, f, myrun);
    const uint32_t rst = __shfl_sync(0xffffffffu, st, myrun), rex = __shfl_sync(0xffffffffu, incl - f, myrun);
    const uint32_t skew = (lpr & 3u) == 0 ? (rb & 3u) : 0u;
    uint32_t rounds = runner ? (rf + skew + lpr - 1) / lpr : 0u;
    rounds = __reduce_max_sync(0xffffffffu, rounds);
    // the last run's first 64 entries are requested now, so that they arrive with the other runs'
    const uint32_t bl = __shfl_sync(0xffffffffu, b, Q - 1), stl = __shfl_sync(0xffffffffu, st, Q - 1), fl = n_tot - n_first;
    uint64_t ol0 = 0, ol1 = 0;
    if (lane < fl) ol0 = A.occ[(uint64_t)bl + lane];
    if (lane + 32u < fl) ol1 = A.occ[(uint64_t)bl + 32u + lane];
    uint64_t mv = 0;
    bool has = false;
    {
      const uint64_t *rp = A.occ + rb;      // entry k of this lane's run
      uint16_t *hp = hst + rex;
      uint32_t k = sub - skew;              // wraps below zero for the skewed first lanes: k >= rf
      const uint32_t lpr4 = 4 * lpr;
      for (uint32_t it0 = 0; it0 < rounds; it0 += 4, k += lpr4) {
        uint64_t o[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t ku = k + u * lpr;
          o[u] = 0;
          if (runner && ku < rf) o[u] = rp[ku];
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t ku = k + u * lpr;
          if (runner && ku < rf) {
            uint32_t h = PAIR_DROPPED;
            if ((uint32_t)o[u] >= rst) { // filter.c:89,106
              const uint64_t v = o[u] - rst;
              h = pair_bucket(v, mask);
              const uint32_t bit = 1u << (h & 31u);
              if (atomicOr(&bmA[h >> 5], bit) & bit) atomicOr(&bmB[h >> 5], bit);
              mv = v; // the run ascends: the last kept entry is this lane's largest
              has = true;
            }
            hp[ku] = (uint16_t)h;
          }
        }
      }
    }
    // the next group's seeds have arrived by now: pull its occurrence runs into L2 (128-byte lines,
    // lpr lines per run, four of the last run) while this group is finished
    if (g + 1 < A.step) {
      const uint32_t pb = __shfl_sync(0xffffffffu, nb, myrun), pf = __shfl_sync(0xffffffffu, nfreq, myrun);
      const uint32_t lb = __shfl_sync(0xffffffffu, nb, Q - 1), lf = __shfl_sync(0xffffffffu, nfreq, Q - 1);
      if (runner && 16u * sub < pf + 15u) asm volatile("prefetch.global.L2 [%0];" ::"l"(A.occ + pb + 16u * sub));
      if (lane < 4 && 16u * lane < lf + 15u) asm volatile("prefetch.global.L2 [%0];" ::"l"(A.occ + lb + 16u * lane));
    }
    const bool any = __any_sync(0xffffffffu, has);
    if (!any) continue; // nothing to merge the last run into
    mv = warp_max_u64(has ? mv : 0ull);
    // ---- pass 1, last run: entries up to that maximum (ascending, so stop at the first larger one) --
    uint32_t n_all = n_first;
    for (uint32_t j0 = 0; j0 < fl; j0 += 32) {
      const uint32_t j = j0 + lane;
      bool beyond = false;
      if (j < fl) {
        const uint64_t o = j0 == 0 ? ol0 : j0 == 32 ? ol1 : A.occ[(uint64_t)bl + j];
        uint32_t h = PAIR_DROPPED;
        if ((uint32_t)o >= stl) {
          if (o - stl <= mv) {
            h = pair_bucket(o - stl, mask);
            const uint32_t bit = 1u << (h & 31u);
            if (atomicOr(&bmA[h >> 5], bit) & bit) atomicOr(&bmB[h >> 5], bit);
          } else beyond = true;
        }
        hst[n_first + j] = (uint16_t)h;
      }
      n_all = n_first + min(fl, j0 + 32u);
      if (__any_sync(0xffffffffu, beyond)) break;
    }
    __syncwarp();
    // ---- flag map F = B | (A & (A rotated by one bucket either way)), four words per lane and round --
    for (uint32_t w0 = 4 * lane; w0 < M / 32; w0 += 128) {
      const uint4 x = *reinterpret_cast<const uint4 *>(bmA + w0);
      uint4 y = *reinterpret_cast<const uint4 *>(bmB + w0);
      const uint32_t lft = bmA[(w0 - 1) & wmask], rgt = bmA[(w0 + 4) & wmask];
      y.x |= x.x & (__funnelshift_l(lft, x.x, 1) | __funnelshift_r(x.x, x.y, 1));
      y.y |= x.y & (__funnelshift_l(x.x, x.y, 1) | __funnelshift_r(x.y, x.z, 1));
      y.z |= x.z & (__funnelshift_l(x.y, x.z, 1) | __funnelshift_r(x.z, x.w, 1));
      y.w |= x.w & (__funnelshift_l(x.z, x.w, 1) | __funnelshift_r(x.w, rgt, 1));
      *reinterpret_cast<uint4 *>(bmB + w0) = y;
    }
    __syncwarp();
    // ---- pass 2: flat indices of the entries whose bucket is flagged, four entries per lane and round;
    //      a lane's count of flagged entries (0..4) is prefix-summed over the warp from three ballots ----
    uint32_t nf = 0;
    for (uint32_t j0 = 0; j0 < n_all; j0 += 128) {
      const uint32_t j = j0 + 4 * lane;
      uint32_t fl4 = 0; // bit u: entry j + u is flagged
      if (j < n_all) {
        const uint2 hh = *reinterpret_cast<const uint2 *>(hst + j); // entries j .. j + 3
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t h = ((u < 2 ? hh.x : hh.y) >> (16 * (u & 1))) & 0xFFFFu;
          if (j + u < n_all && h != PAIR_DROPPED) fl4 |= ((bmB[h >> 5] >> (h & 31u)) & 1u) << u;
        }
      }
      const uint32_t any4 = __ballot_sync(0xffffffffu, fl4 != 0);
      if (any4) {
        const uint32_t c = __popc(fl4), lt = lanemask_lt();
        const uint32_t c0 = __ballot_sync(0xffffffffu, c & 1u), c1 = __ballot_sync(0xffffffffu, c & 2u), c2 = __ballot_sync(0xffffffffu, c & 4u);
        uint32_t idx = nf + __popc(c0 & lt) + 2 * __popc(c1 & lt) + 4 * __popc(c2 & lt);
        for (uint32_t m = fl4; m; m &= m - 1, ++idx)
          if (idx < PAIR_FC) fj[idx] = (uint16_t)(j + __ffs(m) - 1);
        nf += __popc(c0) + 2 * __popc(c1) + 4 * __popc(c2);
      }
    }
    if (nf > PAIR_FC) { overflow = true; break; }
    if (nf < a + 1) continue; // a survivor needs a other flagged entries
    __syncwarp();
    // ---- the flagged entries' values again (L1/L2 hits) --------------------------------------------
    for (uint32_t t = lane; t < nf; t += 32) {
      const uint32_t j = fj[t];
      uint32_t i = 0;
      while (s_incl[i] <= j) ++i;
      const uint32_t first = i ? s_incl[i - 1] : 0u;
      const uint64_t v = A.occ[(uint64_t)s_b[i] + (j - first)] - s_st[i];
      fv[t] = v;
    }
    __syncwarp();
    // ---- exact: flagged entries with at least a other flagged entries in [x, x+e] ----------------------
    uint32_t ns = 0;
    for (uint32_t t0 = 0; t0 < nf; t0 += 32) {
      const uint32_t idx = t0 + lane;
      const uint64_t xv = idx < nf ? fv[idx] : FEM_SENTINEL;
      const uint32_t xlo = (uint32_t)xv, xhi = (uint32_t)(xv >> 32);
      uint32_t cnt = 0;
#pragma unroll 4
      for (uint32_t j = 0; j < nf; ++j) {
        const uint64_t y = fv[j];
        cnt += ((uint32_t)y - xlo <= e32 && (uint32_t)(y >> 32) == xhi) ? 1u : 0u;
      }
      const bool keep = idx < nf && cnt >= a + 1; // cnt includes the entry itself
      const uint32_t bal = __ballot_sync(0xffffffffu, keep);
      if (keep) {
        const uint32_t k = ns + __popc(bal & lanemask_lt());
        if (k < PAIR_SC) surv[k] = ((uint64_t)xhi << 32) | xlo;
      }
      ns += __popc(bal);
    }
    if (ns > PAIR_SC) { overflow = true; break; }
    if (ns == 0) continue;
    __syncwarp();
    // ---- sort the survivors (stable rank by counting), drop repeated values, then the greedy
    //      "> last + e" merge with the running list (filter.c:45-78), walked by one lane ----
    uint64_t x = FEM_SENTINEL;
    uint32_t rank = 0;
    if (lane < ns) {
      x = surv[lane];
      for (uint32_t j = 0; j < ns; ++j) { const uint64_t y = surv[j]; rank += (y < x || (y == x && j < lane)) ? 1u : 0u; }
    }
    __syncwarp();
    if (lane < ns) surv[rank] = x;
    __syncwarp();
    {
      const uint64_t s = lane < ns ? surv[lane] : FEM_SENTINEL;
      const bool fresh = lane < ns && (lane == 0 || surv[lane - 1] != s);
      const uint32_t bal = __ballot_sync(0xffffffffu, fresh);
      __syncwarp();
      if (fresh) surv[__popc(bal & lanemask_lt())] = s;
      ns = __popc(bal);
    }
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
