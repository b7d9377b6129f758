// k_cand_pair32.cuh -- k_cand_pair (near-pair candidate generation, a >= 1) over a 32-bit copy of the
// occurrence table: half the HBM bytes per entry, four entries per 128-bit load, 32-bit arithmetic.
//
// HBM layout (derived on the device when the index is uploaded or built, femgpu.cu):
//   occ32[i] = gbase[record] + position      one word per occurrence, same order as occ
//   gbase[r]  start of record r in a padded global coordinate: multiples of 64 Ki, and at least
//             PAIR32_PAD positions after the end of record r-1
//   padflag   one bit per 64 Ki block: the block ends at a record start
// The reference's arithmetic on (record << 32 | position) words carries over exactly:
//   * order: gbase ascends with the record, so every run stays ascending;
//   * "o - start, entries with position < start dropped" (src/filter.c:89-92,106): o32 - start lands in
//     the PAIR32_PAD (> longest read) positions before the record start exactly when the reference
//     drops the entry, and nothing else ever lies there: test the low 16 bits, then the block's flag;
//   * distances: inside a record they are unchanged, across records they exceed PAIR32_PAD > e, so
//     every "within e" / "> last + e" decision (src/filter.c:53-75,118-131) is the same.
// The surviving candidates are converted back to (record << 32 | position) words by a binary search
// over gbase before the range filter.  Indexes whose padded coordinate does not fit in 32 bits keep
// the 64-bit kernel.  Algorithm and buffers: see k_cand_pair.cuh.
#pragma once
#include "common.cuh"
#include "k_cand.cuh"

#define PAIR32_PAD 2048u        // > FEM_MAX_READ_LEN + 7
#define PAIR32_ALIGN_LOG 16
#define PAIR32_DROPPED 0xFFFFFFFFu
#define PAIR32_FC 64u
#define PAIR32_SC 32u
#define PAIR32_AC 32u

__host__ __device__ inline size_t cand_pair32_smem_bytes_per_warp(uint32_t M, uint32_t VC) {
  return (size_t)M / 8 * 2 + (size_t)VC * 4 + PAIR32_AC * 8 + (PAIR32_FC + PAIR32_SC + 2 * PAIR32_AC) * 4 + 48 * 4;
}

// value in the pad before a record start <=> the reference drops the entry (position < seed offset)
__device__ __forceinline__ bool pair32_in_pad(const uint32_t *padflag, uint32_t v) {
  if ((v & 0xFFFFu) < 0x10000u - PAIR32_PAD) return false;
  const uint32_t blk = v >> PAIR32_ALIGN_LOG;
  return (__ldg(padflag + (blk >> 5)) >> (blk & 31u)) & 1u;
}

__device__ __forceinline__ void pair32_insert(uint32_t *bmA, uint32_t *bmB, uint32_t mask, uint32_t v) {
  const uint32_t h = (v >> 3) & mask, bit = 1u << (h & 31u);
  if (atomicOr(&bmA[h >> 5], bit) & bit) atomicOr(&bmB[h >> 5], bit);
}

__global__ void __launch_bounds__(128) k_cand_pair32(CandArgs A, uint32_t M, uint32_t VC) {
  extern __shared__ uint64_t smem_pair32[];
  const uint32_t lane = lane_id(), wib = threadIdx.x >> 5;
  const uint64_t rs = (uint64_t)blockIdx.x * (blockDim.x >> 5) + wib;
  if (rs >= A.n_rs) return;
  if (A.rs_status[rs] != RS_OK) {
    if (lane == 0) { A.rs_cand_first[rs] = 0; A.rs_cand_count[rs] = 0; }
    return;
  }
  unsigned char *base = reinterpret_cast<unsigned char *>(smem_pair32) + (size_t)wib * cand_pair32_smem_bytes_per_warp(M, VC);
  uint32_t *bmA = reinterpret_cast<uint32_t *>(base);      // occupied buckets
  uint32_t *bmB = bmA + M / 32;                             // buckets with two entries, then the flag map F
  uint32_t *vals = bmB + M / 32;                            // [VC] shifted values of the group, PAIR32_DROPPED = none
  uint64_t *out64 = reinterpret_cast<uint64_t *>(vals + VC); // [AC] final list as (record << 32 | position)
  uint32_t *fv = reinterpret_cast<uint32_t *>(out64 + PAIR32_AC); // [FC] flagged values
  uint32_t *surv = fv + PAIR32_FC;                          // [SC]
  uint32_t *acc = surv + PAIR32_SC, *acc2 = acc + PAIR32_AC;
  const int Q = A.Q;
  const uint32_t a = (uint32_t)A.a, e = (uint32_t)A.e, mask = M - 1, wmask = M / 32 - 1;
  uint32_t na = 0;
  bool overflow = false;
  // pass 1 decomposition: the Q-1 first runs side by side, lpr lanes each, four entries per lane and round
  const uint32_t R = (uint32_t)Q - 1u, lpr = R ? 32u / R : 32u;
  const uint32_t myrun = lane / lpr, sub = lane - myrun * lpr;
  const bool runner = myrun < R;
  uint32_t nb = 0, nfreq = 0, nst = 0; // seeds of the next group, fetched one group ahead
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
    if (n_tot + 4 > VC) { overflow = true; break; }
    __syncwarp();
    {
      uint4 *z = reinterpret_cast<uint4 *>(bmA); // A and B are contiguous: 2 * M / 32 words
      for (uint32_t t = lane; t < M / 64; t += 32) z[t] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncwarp();
    // ---- pass 1: this lane's run in registers; every lane group reads aligned 16-byte quads ---------
    const uint32_t rb = __shfl_sync(0xffffffffu, b, myrun), rf = __shfl_sync(0xffffffffu, f, myrun);
    const uint32_t rst = __shfl_sync(0xffffffffu, st, myrun), rex = __shfl_sync(0xffffffffu, incl - f, myrun);
    const uint32_t skew = rb & 3u;
    uint32_t rounds = runner ? (rf + skew + 4 * lpr - 1) / (4 * lpr) : 0u;
    rounds = __reduce_max_sync(0xffffffffu, rounds);
    // the last run's first 128 entries are requested now, so that they arrive with the other runs'
    const uint32_t bl = __shfl_sync(0xffffffffu, b, Q - 1), stl = __shfl_sync(0xffffffffu, st, Q - 1), fl = n_tot - n_first;
    const uint32_t lskew = bl & 3u;
    uint4 ol = make_uint4(0u, 0u, 0u, 0u);
    if (4 * lane < fl + lskew) ol = *reinterpret_cast<const uint4 *>(A.occ32 + (bl - lskew) + 4 * lane);
    uint32_t mv = 0;
    bool has = false;
    {
      const uint32_t *rp = A.occ32 + (rb - skew);
      uint32_t *vp = vals + rex;
      for (uint32_t it0 = 0; it0 < rounds; it0 += 2) {
        uint4 q[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const uint32_t k4 = ((it0 + u) * lpr + sub) * 4; // quad start, in entries from rb - skew
          q[u] = make_uint4(0u, 0u, 0u, 0u);
          if (runner && k4 < rf + skew) q[u] = *reinterpret_cast<const uint4 *>(rp + k4);
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const uint32_t k4 = ((it0 + u) * lpr + sub) * 4;
          if (runner && k4 < rf + skew) {
            const uint32_t o4[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
#pragma unroll
            for (int w = 0; w < 4; ++w) {
              const uint32_t k = k4 + w - skew; // entry of the run; wraps below zero before its begin
              if (k < rf) {
                uint32_t v = o4[w] - rst;
                if (pair32_in_pad(A.padflag, v)) v = PAIR32_DROPPED; // filter.c:89,106
                else { pair32_insert(bmA, bmB, mask, v); mv = v; has = true; } // the run ascends
                vp[k] = v;
              }
            }
          }
        }
      }
    }
    // the next group's runs are pulled into L2 while this group is finished
    if (g + 1 < A.step) {
      const uint32_t pb = __shfl_sync(0xffffffffu, nb, myrun), pf = __shfl_sync(0xffffffffu, nfreq, myrun);
      const uint32_t lb = __shfl_sync(0xffffffffu, nb, Q - 1), lf = __shfl_sync(0xffffffffu, nfreq, Q - 1);
      if (runner && 32u * sub < pf + 31u) asm volatile("prefetch.global.L2 [%0];" ::"l"(A.occ32 + pb + 32u * sub));
      if (lane < 4 && 32u * lane < lf + 31u) asm volatile("prefetch.global.L2 [%0];" ::"l"(A.occ32 + lb + 32u * lane));
    }
    const bool any = __any_sync(0xffffffffu, has);
    if (!any) continue; // nothing to merge the last run into
    mv = __reduce_max_sync(0xffffffffu, has ? mv : 0u);
    // ---- pass 1, last run: entries up to that maximum (ascending, so stop at the first larger one) --
    uint32_t n_all = n_first;
    {
      const uint32_t *lp = A.occ32 + (bl - lskew);
      uint32_t *vp = vals + n_first;
      for (uint32_t j0 = 0; j0 < fl + lskew; j0 += 128) {
        const uint32_t k4 = j0 + 4 * lane;
        bool beyond = false;
        if (k4 < fl + lskew) {
          const uint4 q = j0 == 0 ? ol : *reinterpret_cast<const uint4 *>(lp + k4);
          const uint32_t o4[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            const uint32_t k = k4 + w - lskew;
            if (k < fl) {
              uint32_t v = o4[w] - stl;
              if (pair32_in_pad(A.padflag, v)) v = PAIR32_DROPPED;
              else if (v <= mv) pair32_insert(bmA, bmB, mask, v);
              else { v = PAIR32_DROPPED; beyond = true; }
              vp[k] = v;
            }
          }
        }
        n_all = n_first + min(fl, j0 + 128u - lskew);
        if (__any_sync(0xffffffffu, beyond)) break;
      }
    }
    if (lane < 4) vals[n_all + lane] = PAIR32_DROPPED; // pass 2 reads whole quads
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
    // ---- pass 2: the entries whose bucket is flagged, four per lane and round; a lane's count of
    //      flagged entries (0..4) is prefix-summed over the warp from three ballots ----------------------
    uint32_t nf = 0;
    for (uint32_t j0 = 0; j0 < n_all; j0 += 128) {
      const uint32_t j = j0 + 4 * lane;
      uint32_t fl4 = 0; // bit w: entry j + w is flagged
      uint32_t v4[4] = {0u, 0u, 0u, 0u};
      if (j < n_all) {
        const uint4 q = *reinterpret_cast<const uint4 *>(vals + j); // entries past n_all are PAIR32_DROPPED
        v4[0] = q.x; v4[1] = q.y; v4[2] = q.z; v4[3] = q.w;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          const uint32_t h = (v4[w] >> 3) & mask;
          if (v4[w] != PAIR32_DROPPED) fl4 |= ((bmB[h >> 5] >> (h & 31u)) & 1u) << w;
        }
      }
      const uint32_t any4 = __ballot_sync(0xffffffffu, fl4 != 0);
      if (any4) {
        const uint32_t c = __popc(fl4), lt = lanemask_lt();
        const uint32_t c0 = __ballot_sync(0xffffffffu, c & 1u), c1 = __ballot_sync(0xffffffffu, c & 2u), c2 = __ballot_sync(0xffffffffu, c & 4u);
        uint32_t idx = nf + __popc(c0 & lt) + 2 * __popc(c1 & lt) + 4 * __popc(c2 & lt);
#pragma unroll
        for (int w = 0; w < 4; ++w)
          if ((fl4 >> w) & 1u) { if (idx < PAIR32_FC) fv[idx] = v4[w]; ++idx; }
        nf += __popc(c0) + 2 * __popc(c1) + 4 * __popc(c2);
      }
    }
    if (nf > PAIR32_FC) { overflow = true; break; }
    if (nf < a + 1) continue; // a survivor needs a other flagged entries
    __syncwarp();
    // ---- exact: flagged entries with at least a other flagged entries in [x, x+e] ----------------------
    uint32_t ns = 0;
    for (uint32_t t0 = 0; t0 < nf; t0 += 32) {
      const uint32_t idx = t0 + lane;
      const uint32_t x = idx < nf ? fv[idx] : 0u;
      uint32_t cnt = 0;
#pragma unroll 4
      for (uint32_t j = 0; j < nf; ++j) cnt += (fv[j] - x <= e) ? 1u : 0u;
      const bool keep = idx < nf && cnt >= a + 1; // cnt includes the entry itself
      const uint32_t bal = __ballot_sync(0xffffffffu, keep);
      if (keep) {
        const uint32_t k = ns + __popc(bal & lanemask_lt());
        if (k < PAIR32_SC) surv[k] = x;
      }
      ns += __popc(bal);
    }
    if (ns > PAIR32_SC) { overflow = true; break; }
    if (ns == 0) continue;
    __syncwarp();
    // ---- sort the survivors (stable rank by counting), drop repeated values, then the greedy
    //      "> last + e" merge with the running list (filter.c:45-78), walked by one lane ----
    uint32_t x = PAIR32_DROPPED, rank = 0;
    if (lane < ns) {
      x = surv[lane];
      for (uint32_t j = 0; j < ns; ++j) { const uint32_t y = surv[j]; rank += (y < x || (y == x && j < lane)) ? 1u : 0u; }
    }
    __syncwarp();
    if (lane < ns) surv[rank] = x;
    __syncwarp();
    {
      const uint32_t s = lane < ns ? surv[lane] : PAIR32_DROPPED;
      const bool fresh = lane < ns && (lane == 0 || surv[lane - 1] != s);
      const uint32_t bal = __ballot_sync(0xffffffffu, fresh);
      __syncwarp();
      if (fresh) surv[__popc(bal & lanemask_lt())] = s;
      ns = __popc(bal);
    }
    __syncwarp();
    uint32_t no = 0;
    if (lane == 0) {
      uint32_t i1 = 0, i2 = 0, last = 0;
      while (i1 < na || i2 < ns) {
        uint32_t v;
        if (i1 < na && (i2 >= ns || acc[i1] < surv[i2])) v = acc[i1++];
        else v = surv[i2++];
        if (no == 0 || v > last + e) {
          if (no < PAIR32_AC) acc2[no] = v;
          ++no;
          last = v;
        }
      }
    }
    no = __shfl_sync(0xffffffffu, no, 0);
    if (no > PAIR32_AC) { overflow = true; break; }
    na = no;
    uint32_t *sw = acc; acc = acc2; acc2 = sw;
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
  // ---- back to (record << 32 | position): last record whose start is <= the value ------------------
  if (lane < na) {
    const uint32_t v = acc[lane];
    uint32_t lo = 0, hi = A.n_rec; // gbase[lo] <= v < gbase[hi] (gbase[n_rec] = end of the coordinate)
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (__ldg(A.gbase + mid) <= v) lo = mid; else hi = mid;
    }
    out64[lane] = ((uint64_t)lo << 32) | (uint64_t)(v - __ldg(A.gbase + lo));
  }
  __syncwarp();
  rs_emit_candidates(A, rs, lane, out64, na);
}
