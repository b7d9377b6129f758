// k_cand.cuh -- K3: candidate generation for one read-strand per warp.
//
// Restates src/filter.c:80-144 + the group loop of :190-213 in a set-based form:
//   * merge_candidate_locations (:80-116) yields the sorted multiset union of the first Q-1
//     occurrence runs (each shifted by its seed's read offset, entries with pos < offset dropped)
//     plus those entries of the LAST (most frequent) run that are <= the maximum of that union
//     (the reference stops folding the last run when the accumulator is exhausted).
//     Order inside a multiset is irrelevant, so the runs are gathered with coalesced loads and
//     sorted by a warp bitonic network instead of being folded one at a time.
//   * additional_qgram_filter (:118-131): keep v[i] iff v[i+a] <= v[i]+e.
//   * merge_kvec_t_uint64_t (:45-78): merge with the running list and keep x iff
//     x > last kept + e.  This greedy pass is order dependent across groups, so groups are
//     processed in the reference's order; inside a pass the warp resolves 32 elements at a time
//     with ballots.
//   * remove_out_ranged_candidates (:133-144).
// Buffers are passed as generic pointers: shared memory on the fast path, a slice of a global
// scratch arena for read-strands whose lists do not fit (k_cand_big).
#pragma once
#include "common.cuh"

struct CandArgs {
  const uint64_t *occ;
  uint32_t n_rec;
  const uint32_t *seed_begin, *seed_freq, *seed_start; // [rs*step+g][Q]
  const uint8_t *rs_status;
  const uint64_t *read_off;
  const uint32_t *rec_len;
  uint32_t n_rs;
  int step, e, a, Q;
  // outputs
  uint64_t *cand;                  // arena of candidates
  uint32_t *cand_rs;               // owning read-strand of each arena entry
  unsigned long long *cand_count;  // arena bump pointer
  uint64_t cand_cap;
  uint64_t *rs_cand_first;         // per read-strand: arena offset
  uint32_t *rs_cand_count;
  uint32_t *overflow_list;         // read-strands that need the big path
  uint32_t *overflow_count;
  int *error_flag;                 // 1: arena full
  // fast-path shared-memory plan (elements per warp)
  uint32_t S, Acap;
  // list mode of k_cand / k_cand_pair: the read-strands an earlier path handed on
  const uint32_t *rs_list;
  uint32_t n_list;
  const uint32_t *n_list_dev;      // k_cand_pair: the list's length where it stands on the device (nullptr: n_list);
  uint32_t list_first;             //   the launch covers entries [list_first, min(n_list, *n_list_dev))
  // big path
  const uint32_t *big_list;        // read-strand ids
  const uint64_t *big_off;         // scratch offset (elements) per entry
  const uint32_t *big_S, *big_A;
  uint64_t *scratch;
  uint32_t n_big;
};

__device__ __forceinline__ void warp_bitonic_sort(uint64_t *buf, uint32_t n, uint32_t lane) {
  for (uint32_t k = 2; k <= n; k <<= 1) {
    for (uint32_t j = k >> 1; j > 0; j >>= 1) {
      for (uint32_t t = lane; t < (n >> 1); t += 32) {
        const uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        const uint32_t p = i | j;
        const uint64_t x = buf[i], y = buf[p];
        const bool up = (i & k) == 0;
        if ((x > y) == up) { buf[i] = y; buf[p] = x; }
      }
      __syncwarp();
    }
  }
}

// number of elements < x (lower) in sorted v[0..n)
__device__ __forceinline__ uint32_t lower_bound_u64(const uint64_t *v, uint32_t n, uint64_t x) {
  uint32_t lo = 0, hi = n;
  while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (v[mid] < x) lo = mid + 1; else hi = mid; }
  return lo;
}
// number of elements <= x
__device__ __forceinline__ uint32_t upper_bound_u64(const uint64_t *v, uint32_t n, uint64_t x) {
  uint32_t lo = 0, hi = n;
  while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (v[mid] <= x) lo = mid + 1; else hi = mid; }
  return lo;
}

// Returns 0 and the final list in acc[0..*n_out) (before the range filter), or 1 on overflow.
__device__ int rs_candidate_list(const CandArgs &A, uint64_t rs, uint32_t lane, uint64_t *buf, uint32_t S,
                                 uint64_t *acc, uint32_t Acap, uint32_t *n_out) {
  const int Q = A.Q, e = A.e, a = A.a;
  uint32_t na = 0;
  for (int g = 0; g < A.step; ++g) {
    const size_t so = ((size_t)rs * A.step + g) * Q;
    uint32_t b = 0, f = 0, st = 0;
    if ((int)lane < Q) { b = A.seed_begin[so + lane]; f = A.seed_freq[so + lane]; st = A.seed_start[so + lane]; }
    // maximum of the first Q-1 shifted runs = the largest surviving entry of any of them
    uint64_t mv = 0;
    bool has = false;
    if ((int)lane < Q - 1)
      for (int64_t t = (int64_t)f - 1; t >= 0; --t) {
        const uint64_t o = A.occ[(uint64_t)b + (uint64_t)t];
        if ((uint32_t)o >= st) { mv = o - st; has = true; break; }
      }
    const uint32_t any = __ballot_sync(0xffffffffu, has);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) mv = max(mv, __shfl_xor_sync(0xffffffffu, mv, d));
    if ((int)lane == Q - 1) { // F5: the last run contributes only entries <= that maximum
      if (!any) f = 0;
      else f = upper_bound_u64(A.occ + b, f, mv + st);
    }
    // exclusive prefix of the run lengths over lanes 0..Q-1
    uint32_t incl = f;
#pragma unroll
    for (int d = 1; d < 16; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
      if ((int)lane >= d) incl += o;
    }
    const uint32_t n_tot = __shfl_sync(0xffffffffu, incl, Q - 1);
    const uint32_t my_off = incl - f;
    if (n_tot == 0) continue; // nothing new: the greedy pass leaves the running list unchanged
    const uint32_t n_pad = max(32u, pow2_ceil_u32(n_tot));
    if (n_pad > S) return 1;
    uint32_t dropped = 0;
    for (int i = 0; i < Q; ++i) {
      const uint32_t bi = __shfl_sync(0xffffffffu, b, i), fi = __shfl_sync(0xffffffffu, f, i);
      const uint32_t sti = __shfl_sync(0xffffffffu, st, i), oi = __shfl_sync(0xffffffffu, my_off, i);
      for (uint32_t t0 = 0; t0 < fi; t0 += 32) {
        const uint32_t t = t0 + lane;
        bool drop = false;
        if (t < fi) {
          const uint64_t o = A.occ[(uint64_t)bi + t];
          drop = (uint32_t)o < sti; // filter.c:89,106
          buf[oi + t] = drop ? FEM_SENTINEL : o - sti;
        }
        dropped += __popc(__ballot_sync(0xffffffffu, drop));
      }
    }
    for (uint32_t t = n_tot + lane; t < n_pad; t += 32) buf[t] = FEM_SENTINEL;
    __syncwarp();
    warp_bitonic_sort(buf, n_pad, lane);
    const uint32_t n_valid = n_tot - dropped;
    // ---- additional q-gram filter, compacted in place (writes never pass the read cursor) ----
    uint32_t ns = 0;
    if (a == 0) ns = n_valid;
    else
      for (uint32_t base = 0; base < n_valid; base += 32) {
        const uint32_t t = base + lane;
        uint64_t v = 0;
        bool keep = false;
        if (t < n_valid) { v = buf[t]; keep = t + a < n_valid && buf[t + a] <= v + (uint64_t)e; }
        __syncwarp();
        const uint32_t m = __ballot_sync(0xffffffffu, keep);
        if (keep) buf[ns + __popc(m & lanemask_lt())] = v;
        ns += __popc(m);
        __syncwarp();
      }
    if (ns == 0) continue;
    // ---- merge the survivors with the running list --------------------------------------------
    const uint64_t *T = buf;
    const uint32_t nt = ns + na;
    if (na) {
      if ((uint64_t)ns + nt > S) return 1;
      uint64_t *Tm = buf + ns;
      for (uint32_t i = lane; i < na; i += 32) { const uint64_t x = acc[i]; Tm[i + lower_bound_u64(buf, ns, x)] = x; }
      for (uint32_t j = lane; j < ns; j += 32) { const uint64_t y = buf[j]; Tm[j + upper_bound_u64(acc, na, y)] = y; }
      __syncwarp();
      T = Tm;
    }
    // ---- greedy "> last + e" pass (filter.c:53-75) ---------------------------------------------
    uint32_t no = 0;
    uint64_t last = 0;
    bool has_last = false;
    for (uint32_t base = 0; base < nt; base += 32) {
      const uint32_t t = base + lane;
      const uint64_t v = t < nt ? T[t] : 0;
      uint32_t remaining = __ballot_sync(0xffffffffu, t < nt), keepmask = 0;
      for (;;) {
        const bool c = ((remaining >> lane) & 1u) && (!has_last || v > last + (uint64_t)e);
        const uint32_t m = __ballot_sync(0xffffffffu, c);
        if (!m) break;
        const int fl = __ffs(m) - 1;
        last = __shfl_sync(0xffffffffu, v, fl);
        has_last = true;
        keepmask |= 1u << fl;
        remaining &= ~((2u << fl) - 1u);
      }
      if ((keepmask >> lane) & 1u) {
        const uint32_t idx = no + __popc(keepmask & lanemask_lt());
        if (idx < Acap) acc[idx] = v;
      }
      no += __popc(keepmask);
    }
    __syncwarp();
    if (no > Acap) return 1;
    na = no;
  }
  *n_out = na;
  return 0;
}

// range filter (filter.c:133-144) + arena write
__device__ void rs_emit_candidates(const CandArgs &A, uint64_t rs, uint32_t lane, const uint64_t *acc, uint32_t na) {
  const uint32_t read = (uint32_t)(rs >> 1);
  const uint32_t l = (uint32_t)(A.read_off[read + 1] - A.read_off[read]);
  const uint32_t e = (uint32_t)A.e;
  // pass 1: count
  uint32_t n_keep = 0;
  for (uint32_t base = 0; base < na; base += 32) {
    const uint32_t t = base + lane;
    bool keep = false;
    if (t < na) {
      const uint64_t v = acc[t];
      const uint32_t pos = (uint32_t)v, len = A.rec_len[v >> 32];
      keep = pos >= e && pos + l + e < len;
    }
    n_keep += __popc(__ballot_sync(0xffffffffu, keep));
  }
  unsigned long long first = 0;
  if (lane == 0 && n_keep) first = atomicAdd(A.cand_count, (unsigned long long)n_keep);
  first = __shfl_sync(0xffffffffu, first, 0);
  if (n_keep && first + n_keep > A.cand_cap) {
    if (lane == 0) { *A.error_flag = 1; A.rs_cand_first[rs] = 0; A.rs_cand_count[rs] = 0; }
    return;
  }
  uint32_t w = 0;
  for (uint32_t base = 0; base < na && n_keep; base += 32) {
    const uint32_t t = base + lane;
    bool keep = false;
    uint64_t v = 0;
    if (t < na) {
      v = acc[t];
      const uint32_t pos = (uint32_t)v, len = A.rec_len[v >> 32];
      keep = pos >= e && pos + l + e < len;
    }
    const uint32_t m = __ballot_sync(0xffffffffu, keep);
    if (keep) {
      const uint64_t o = first + w + __popc(m & lanemask_lt());
      A.cand[o] = v - e; // the stored candidate is the band's left edge
      A.cand_rs[o] = (uint32_t)rs;
    }
    w += __popc(m);
  }
  if (lane == 0) { A.rs_cand_first[rs] = first; A.rs_cand_count[rs] = n_keep; }
}

__global__ void __launch_bounds__(128) k_cand(CandArgs A) {
  extern __shared__ uint64_t smem_cand[];
  const uint32_t lane = lane_id(), wib = threadIdx.x >> 5;
  uint64_t rs = (uint64_t)blockIdx.x * (blockDim.x >> 5) + wib;
  if (A.rs_list) {
    if (rs >= A.n_list) return;
    rs = A.rs_list[rs];
  } else if (rs >= A.n_rs) return;
  if (A.rs_status[rs] != RS_OK) {
    if (lane == 0) { A.rs_cand_first[rs] = 0; A.rs_cand_count[rs] = 0; }
    return;
  }
  uint64_t *buf = smem_cand + (size_t)wib * (A.S + A.Acap);
  uint64_t *acc = buf + A.S;
  uint32_t na = 0;
  if (rs_candidate_list(A, rs, lane, buf, A.S, acc, A.Acap, &na)) {
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

// big path: same algorithm, buffers in the global scratch arena (sized so that it cannot overflow)
__global__ void __launch_bounds__(128) k_cand_big(CandArgs A) {
  const uint32_t lane = lane_id();
  const uint32_t w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (w >= A.n_big) return;
  const uint64_t rs = A.big_list[w];
  uint64_t *buf = A.scratch + A.big_off[w];
  uint64_t *acc = buf + A.big_S[w];
  uint32_t na = 0;
  if (rs_candidate_list(A, rs, lane, buf, A.big_S[w], acc, A.big_A[w], &na)) {
    if (lane == 0) *A.error_flag = 2; // cannot happen by construction
    return;
  }
  __syncwarp();
  rs_emit_candidates(A, rs, lane, acc, na);
}
