// k_cand_reg.cuh -- K3 main path on dense indexes (a >= 1, cfg3-like: ~37 occurrences per seed): the near-pair
// detection of k_cand_pair.cuh with a group's entries kept in REGISTERS between its two passes.
//
// Same result as k_cand / k_cand_hash / k_cand_pair and therefore as src/filter.c:80-144 (+ the group
// loop of :190-213).  k_cand_pair spends most of its instructions moving per-entry state through shared
// memory (a 16-bit bitmap position per entry, a flat index per flagged entry, the search for the run a
// flat index belongs to).  Here the shape is fixed at compile time -- Q seeds per group, UA load slots
// per lane -- so that every entry has a NAME (lane, slot):
//   loads   all occurrence words of the group are requested at once: runs 0..Q-2 side by side, LPR = 32/(Q-1)
//           lanes per run and one aligned 32-byte sector per run and slot (when LPR is a multiple of 4), the
//           last run 32 entries per slot; the loads of a whole group (~2.7 KB at cfg3) are in flight together;
//   pass 1  every entry marks its 8-wide bucket in the bitmap (and the next bucket too when its own was taken: see
//           preg_mark) and keeps the bitmap position, two to a register.  Entries the reference drops (position < seed offset,
//           filter.c:89,106) are marked all the same: they can only add false flags, and the exact test
//           follows below;
//   flags   F = A & (A << 1 | A >> 1), in place;
//   pass 2  one bit test per slot gives the lane's mask of flagged slots; only flagged entries (~30 of ~330
//           per group at cfg3) are loaded again (cache hits), filtered exactly and written to the flagged list;
//   exact   as in k_cand_pair: count in [x, x+e] among the flagged entries, rank sort, greedy merge.
// The last (most frequent) seed's run counts only up to the maximum of the other runs; that
// maximum comes from one lane per run reading its run's last kept entry.
// A group that does not fit the slots (a run of more than ~UA*LPR entries, a last run of more than 96) hands
// the read-strand to k_cand_pair in list mode, and from there to k_cand as before.
#pragma once
#include "common.cuh"
#include "k_cand.cuh"
#include "k_cand_hash.cuh" // warp_max_u64
#include "k_cand_pair.cuh" // PAIR_FC, PAIR_SC, PAIR_AC

#define PREG_UB 3 // slots of the last run (32 entries each)

// M bits of the bitmap (power of two, 2048..32768: positions are kept in 16 bits)
__host__ __device__ inline size_t cand_reg_smem_bytes_per_warp(uint32_t M) {
  return (size_t)M / 8 + (PAIR_SC + 2 * PAIR_AC) * 8 + PAIR_FC * 8;
}

// Bitmap marks of NS slots at once: the entry's 8-wide bucket h is bit h & 31 of the word at shared address
// bm_s + 4 * (h >> 5).  An entry that finds its bit set already (two entries in one bucket) also marks the NEXT bucket:
// "bucket and a neighbour both occupied" then flags the bucket of both, and no second bitmap is needed for buckets that
// hold two entries -- which lets the one bitmap be twice as long in the same shared memory (half the false flags).
// All NS first atomics are issued before the first result is looked at (a shared atomic returns after ~30 cycles; one
// at a time, the kernel waited on the short scoreboard for a third of its cycles).  A slot without an entry ORs 0 into
// the word its filler value points at -- the fillers differ from lane to lane, so these no-ops do not pile up on one
// address -- and so needs no predicate: ptxas turns a predicated shared atomic into a divergence region.
template <int NS>
__device__ __forceinline__ void preg_mark(uint32_t bm_s, uint32_t mask, const uint32_t (&h)[NS], const uint32_t (&bit)[NS]) {
  uint32_t old[NS];
#pragma unroll
  for (int i = 0; i < NS; ++i) {
    uint32_t addr;
    asm("{\n\t.reg .u32 t;\n\tshr.u32 t, %1, 5;\n\tmad.lo.u32 %0, t, 4, %2;\n\t}" : "=r"(addr) : "r"(h[i]), "r"(bm_s));
    asm volatile("atom.shared.or.b32 %0, [%1], %2;" : "=r"(old[i]) : "r"(addr), "r"(bit[i]));
  }
#pragma unroll
  for (int i = 0; i < NS; ++i) {
    if (old[i] & bit[i]) {
      const uint32_t h1 = (h[i] + 1u) & mask;
      asm volatile("red.shared.or.b32 [%0], %1;" ::"r"(bm_s + (h1 >> 5) * 4u), "r"(1u << (h1 & 31u)));
    }
  }
}

template <int Q, int UA, int MINB>
__global__ void __launch_bounds__(128, MINB) k_cand_reg(CandArgs A, uint32_t M) {
  constexpr uint32_t R = Q - 1, LPR = 32u / R, UB = PREG_UB;
  static_assert(Q >= 2 && Q <= FEM_MAX_Q && UA % 4 == 0 && UA + UB <= 32, "shape");
  extern __shared__ uint64_t smem_reg[];
  const uint32_t lane = lane_id(), wib = threadIdx.x >> 5;
  const uint64_t rs = (uint64_t)blockIdx.x * (blockDim.x >> 5) + wib;
  if (rs >= A.n_rs) return;
  if (A.rs_status[rs] != RS_OK) {
    if (lane == 0) { A.rs_cand_first[rs] = 0; A.rs_cand_count[rs] = 0; }
    return;
  }
  unsigned char *base = reinterpret_cast<unsigned char *>(smem_reg) + (size_t)wib * cand_reg_smem_bytes_per_warp(M);
  uint32_t *bm = reinterpret_cast<uint32_t *>(base);           // occupied buckets, then (in place) the flagged buckets
  const uint32_t bm_s = (uint32_t)__cvta_generic_to_shared(bm);
  uint64_t *surv = reinterpret_cast<uint64_t *>(bm + M / 32);
  uint64_t *acc = surv + PAIR_SC, *acc2 = acc + PAIR_AC;
  uint64_t *fv = acc2 + PAIR_AC;                               // values of the flagged entries
  const uint32_t a = (uint32_t)A.a, e32 = (uint32_t)A.e, mask = M - 1, nw = M / 32;
  const uint64_t e = (uint64_t)A.e;
  const uint32_t myrun = lane / LPR, sub = lane - myrun * LPR;
  const uint64_t filler = (uint64_t)(lane * 0x9E3779B1u); // lane-specific: the no-op atomics of empty slots spread over the words
  const bool runner = myrun < R;
  uint32_t na = 0;
  bool overflow = false;
  // seeds of the next group are fetched while the current one is processed
  uint32_t nb = 0, nfreq = 0, nst = 0;
  size_t so = (size_t)rs * A.step * Q + (lane < (uint32_t)Q ? lane : 0u); // this lane's seed of the next group
  if (lane < (uint32_t)Q) { nb = A.seed_begin[so]; nfreq = A.seed_freq[so]; nst = A.seed_start[so]; }

  for (int g = 0; g < A.step; ++g) {
    const uint32_t b = nb, f = nfreq, st = nst;
    so += Q;
    if (g + 1 < A.step && lane < (uint32_t)Q) { nb = A.seed_begin[so]; nfreq = A.seed_freq[so]; nst = A.seed_start[so]; }
    const uint32_t n_first = __reduce_add_sync(0xffffffffu, lane < R ? f : 0u); // all runs but the last
    if (n_first == 0) continue;                 // the last run alone contributes nothing (F5)
    // run registers of this lane: entry k0 + u * LPR of run myrun in slot u; with LPR a multiple of 4 every
    // lane group reads one aligned 32-byte sector per slot (slot 0 starts below the run's begin by begin mod 4)
    const uint32_t rb = __shfl_sync(0xffffffffu, b, myrun), rf = __shfl_sync(0xffffffffu, f, myrun) & (runner ? 0xFFFFFFFFu : 0u);
    const uint32_t rst = __shfl_sync(0xffffffffu, st, myrun);
    const uint32_t bl = __shfl_sync(0xffffffffu, b, Q - 1), stl = __shfl_sync(0xffffffffu, st, Q - 1), fl = __shfl_sync(0xffffffffu, f, Q - 1);
    const uint32_t skew = (LPR & 3u) == 0 ? (rb & 3u) : 0u;
    const uint32_t rounds = __reduce_max_sync(0xffffffffu, (rf + skew + LPR - 1) / LPR);
    if (rounds > (uint32_t)UA || fl > 32u * UB) { overflow = true; break; }
    const uint64_t *rp = A.occ + rb;
    // slot u of this lane holds entry k0 + u * LPR of its run; k0 < 0 for the skewed first lanes of slot 0.
    // One 64-bit address per lane and group; the slots are constant offsets from it.
    const int32_t k0 = (int32_t)sub - (int32_t)skew;
    const uint64_t *p0 = rp + k0;
    const int32_t lim = (int32_t)rf - k0;   // slot u >= 1 holds an entry iff u * LPR < lim; slot 0 iff 0 <= k0 < rf
    const uint64_t *pb = A.occ + bl + lane; // last run: entry lane + 32 j in slot UA + j
    const int32_t limb = (int32_t)fl - (int32_t)lane;
    // ---- every load of the group at once -------------------------------------------------------------
    uint64_t o[UA], ob[UB];
#pragma unroll
    for (int u = 0; u < UA; ++u) {
      o[u] = filler; // a slot without an entry: see preg_mark
      if ((int32_t)(u * LPR) < lim && (u > 0 || k0 >= 0)) o[u] = p0[u * LPR];
    }
#pragma unroll
    for (int j = 0; j < (int)UB; ++j) {
      ob[j] = filler;
      if (32 * j < limb) ob[j] = pb[32 * j];
    }
    // valid slots of this lane, as a mask: slots [k0 < 0, ceil(lim / LPR)) of runs 0..Q-2 (rounds <= UA was checked)
    uint32_t vm = 0;
    if (lim > 0) vm = ((1u << (((uint32_t)lim + LPR - 1) / LPR)) - 1u) & ~(k0 < 0 ? 1u : 0u);
    // the largest kept entry of runs 0..Q-2 (they ascend): the last entry of each run, or the one before ...
    uint64_t mv = 0;
    bool has = false;
    if (sub == 0 && rf)
      for (uint32_t t = rf; t > 0;) {
        const uint64_t x = rp[--t];
        if ((uint32_t)x >= rst) { mv = x - rst; has = true; break; } // filter.c:89,106
      }
    {
      uint4 *z = reinterpret_cast<uint4 *>(bm);
      for (uint32_t t = lane; t < M / 128; t += 32) z[t] = make_uint4(0u, 0u, 0u, 0u);
    }
    if (!__any_sync(0xffffffffu, has)) continue; // nothing to merge the last run into
    mv = warp_max_u64(has ? mv : 0ull);
    __syncwarp();
    // ---- pass 1, runs 0..Q-2 ---------------------------------------------------------------------------
    // Entries the reference drops (position < seed offset) are marked like the others: a false flag at worst.
    uint32_t hs[UA / 2 + 2]; // bitmap positions, two to a register: slots 0..UA-1, then the last run's
#pragma unroll
    for (int c = 0; c < UA / 4; ++c) {
      hs[2 * c] = hs[2 * c + 1] = 0;
      if (4u * c < rounds) {
        uint32_t h[4], bit[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int u = 4 * c + i;
          h[i] = ((((uint32_t)o[u] - rst) >> 3) + (uint32_t)(o[u] >> 32) * 0x9E3779B1u) & mask;
          bit[i] = ((vm >> u) & 1u) << (h[i] & 31u);
          hs[u >> 1] |= h[i] << (16 * (u & 1));
        }
        preg_mark<4>(bm_s, mask, h, bit);
      }
    }
    // ---- pass 1, last run: entries up to the maximum of the others (ascending: the rest lies beyond) ----------
    hs[UA / 2] = hs[UA / 2 + 1] = 0;
    if (fl) {
      uint32_t h[UB], bit[UB];
#pragma unroll
      for (int j = 0; j < (int)UB; ++j) {
        const uint32_t lo = (uint32_t)ob[j];
        const uint64_t v = ob[j] - stl;
        h[j] = (((uint32_t)v >> 3) + (uint32_t)(v >> 32) * 0x9E3779B1u) & mask;
        const bool keep = 32 * j < limb && lo >= stl && v <= mv;
        bit[j] = keep ? 1u << (h[j] & 31u) : 0u;
        vm |= keep ? 1u << (UA + j) : 0u;
        hs[(UA + j) >> 1] |= h[j] << (16 * ((UA + j) & 1));
      }
      preg_mark<UB>(bm_s, mask, h, bit);
    }
    // the next group's seeds have arrived by now: pull its occurrence runs into L2 (128-byte lines, one per lane of a
    // run, six for the last run) while this group is finished
    if (g + 1 < A.step) {
      const uint32_t qb = __shfl_sync(0xffffffffu, nb, myrun), qf = __shfl_sync(0xffffffffu, nfreq, myrun);
      const uint32_t lb = __shfl_sync(0xffffffffu, nb, Q - 1), lf = __shfl_sync(0xffffffffu, nfreq, Q - 1);
      if (runner && 16u * sub < qf + 15u) asm volatile("prefetch.global.L2 [%0];" ::"l"(A.occ + qb + 16u * sub));
      if (lane < 2u * UB && 16u * lane < lf + 15u) asm volatile("prefetch.global.L2 [%0];" ::"l"(A.occ + lb + 16u * lane));
    }
    __syncwarp();
    // ---- flagged buckets F = A & (A rotated by one bucket either way), in place, four words per lane and round.  The
    //      neighbour words come from the neighbour lanes' registers; the words across a round's edge are read before
    //      they are overwritten (the last word of the previous round is carried, word 0 is kept for the wrap-around). ----
    if (nw >= 128) { // four words per lane and round (eight: 2-way bank conflicts of the 16-byte accesses, 15.4 against 14.4 ms)
      const uint32_t first = bm[0];
      uint32_t carry = bm[nw - 1];
      for (uint32_t r0 = 0; r0 < nw; r0 += 128) {
        uint32_t *p = bm + r0 + 4 * lane;
        const uint4 x = *reinterpret_cast<const uint4 *>(p);
        uint32_t lft = __shfl_up_sync(0xffffffffu, x.w, 1), rgt = __shfl_down_sync(0xffffffffu, x.x, 1);
        const uint32_t next = r0 + 128 < nw ? bm[r0 + 128] : first; // the word after this round's last, still unchanged
        if (lane == 0) lft = carry;
        if (lane == 31) rgt = next;
        carry = __shfl_sync(0xffffffffu, x.w, 31);
        uint4 y;
        y.x = x.x & (__funnelshift_l(lft, x.x, 1) | __funnelshift_r(x.x, x.y, 1));
        y.y = x.y & (__funnelshift_l(x.x, x.y, 1) | __funnelshift_r(x.y, x.z, 1));
        y.z = x.z & (__funnelshift_l(x.y, x.z, 1) | __funnelshift_r(x.z, x.w, 1));
        y.w = x.w & (__funnelshift_l(x.z, x.w, 1) | __funnelshift_r(x.w, rgt, 1));
        __syncwarp(); // every read of this round (and of the next round's first word) is done before the stores
        *reinterpret_cast<uint4 *>(p) = y;
      }
    } else { // short bitmaps (tests): four words per lane, lanes past the end idle
      const uint32_t first = bm[0];
      uint32_t carry = bm[nw - 1];
      for (uint32_t r0 = 0; r0 < nw; r0 += 128) {
        const uint32_t w0 = r0 + 4 * lane;
        const bool in = w0 < nw;
        uint4 x = make_uint4(0u, 0u, 0u, 0u);
        if (in) x = *reinterpret_cast<const uint4 *>(bm + w0);
        uint32_t lft = __shfl_up_sync(0xffffffffu, x.w, 1), rgt = __shfl_down_sync(0xffffffffu, x.x, 1);
        if (lane == 0) lft = carry;
        if (w0 + 4 >= nw) rgt = first;
        else if (lane == 31) rgt = bm[w0 + 4];
        carry = __shfl_sync(0xffffffffu, x.w, 31);
        uint4 y;
        y.x = x.x & (__funnelshift_l(lft, x.x, 1) | __funnelshift_r(x.x, x.y, 1));
        y.y = x.y & (__funnelshift_l(x.x, x.y, 1) | __funnelshift_r(x.y, x.z, 1));
        y.z = x.z & (__funnelshift_l(x.y, x.z, 1) | __funnelshift_r(x.z, x.w, 1));
        y.w = x.w & (__funnelshift_l(x.z, x.w, 1) | __funnelshift_r(x.w, rgt, 1));
        __syncwarp();
        if (in) *reinterpret_cast<uint4 *>(bm + w0) = y;
      }
    }
    __syncwarp();
    // ---- pass 2: this lane's flagged slots ---------------------------------------------------------------
    uint32_t fm = 0;
#pragma unroll
    for (int c = 0; c < UA / 4; ++c)
      if (4u * c < rounds) {
#pragma unroll
        for (int u = 4 * c; u < 4 * c + 4; ++u) {
          const uint32_t h = (hs[u >> 1] >> (16 * (u & 1))) & 0xFFFFu;
          fm |= ((bm[h >> 5] >> (h & 31u)) & 1u) << u;
        }
      }
    if (fl) {
#pragma unroll
      for (int j = 0; j < (int)UB; ++j) {
        const uint32_t h = (hs[(UA + j) >> 1] >> (16 * ((UA + j) & 1))) & 0xFFFFu;
        fm |= ((bm[h >> 5] >> (h & 31u)) & 1u) << (UA + j);
      }
    }
    fm &= vm;
    // ---- the flagged entries' values again (L1/L2 hits), filtered exactly ---------------------------------
    uint32_t nf = 0;
    while (__any_sync(0xffffffffu, fm != 0)) {
      bool keep = false;
      uint64_t v = 0;
      if (fm) {
        const uint32_t u = (uint32_t)__ffs(fm) - 1u;
        fm &= fm - 1;
        if (u < (uint32_t)UA) {
          const uint64_t x = p0[u * LPR];
          keep = (uint32_t)x >= rst; // filter.c:89,106
          v = x - rst;
        } else {
          v = pb[32u * (u - UA)] - stl; // kept in pass 1 already
          keep = true;
        }
      }
      const uint32_t bal = __ballot_sync(0xffffffffu, keep);
      if (keep) {
        const uint32_t idx = nf + __popc(bal & lanemask_lt());
        if (idx < PAIR_FC) fv[idx] = v;
      }
      nf += __popc(bal);
    }
    if (nf > PAIR_FC) { overflow = true; break; }
    if (nf < a + 1) continue; // a survivor needs a other flagged entries
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
