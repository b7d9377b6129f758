// pipeline.inl -- the per-batch device pipeline (included by femgpu.cu)

#define REG_UA 16          // load slots per lane of k_cand_reg (runs of up to 16 * LPR - 3 entries)
#define REG_CHAIN 16384ull // warps of the k_cand_pair launch chained behind k_cand_reg on the device

struct Plan { // derived from params + the batch
  int k, step, e, a, Q, s;
  int l_max, n_words_max, ns_max, w_max, lane_dp;
  size_t seed_smem;   // bytes per block of k_seed
  int seed_R;         // > 0: k_seed_multi with this many read-strands per warp
  uint32_t seed_words_per_rs;
  size_t cand_smem;   // bytes per block of k_cand
  uint32_t hash_T, hash_shift; // k_cand_hash: table slots per warp, 64 - log2(slots)
  size_t hash_smem;
  uint32_t pair_M, pair_shift, pair_VC; // k_cand_pair: bitmap bits, 32 - log2(bits), values per warp
  size_t pair_smem, reg_smem;
  uint32_t reg_M;     // k_cand_reg: bits of its one bitmap
  bool pair_reg; // k_cand_reg first (shape instantiated for this Q and runs expected to fit its slots)
  int first_path; // 0 sort, 1 hash, 2 pair, 3 small
  size_t small_smem;
  int small_R;          // k_cand_small shape: 1 (one seed per lane) or 4
  uint32_t small_inv_Q; // multiply-high reciprocal of Q for seed -> group
};

static int make_plan(femgpu_ctx *c, const femgpu_params *p, int l_max, Plan *pl, bool over_long_reads = false) {
  if (!c->n_occ && !c->lut.p) return set_error(FEMGPU_ESTATE, "index not uploaded");
  if (!c->n_rec) return set_error(FEMGPU_ESTATE, "reference not uploaded");
  pl->k = p->k > 0 ? p->k : c->idx_k;
  pl->step = p->step > 0 ? p->step : c->idx_step;
  pl->e = p->e;
  pl->a = p->a;
  if (pl->e < 0 || pl->e > 7) return set_error(FEMGPU_EINVAL, "Wrong error threshold.");          // src/FEM_map.c:30
  if (pl->a < 0 || pl->a > 2) return set_error(FEMGPU_EINVAL, "Wrong number of additional q-grams."); // :38
  // the lookup table is addressed by 2k-bit hashes: a seeding k other than the index's would
  // read out of bounds when larger, and is only meaningful when equal
  if (pl->k != c->idx_k) return set_error(FEMGPU_EINVAL, "seeding k=%d differs from the index's k=%d", pl->k, c->idx_k);
  if (pl->step < 1 || pl->step > 64) return set_error(FEMGPU_EINVAL, "bad step %d", pl->step);
  if (l_max > FEM_MAX_READ_LEN) return set_error(FEMGPU_EINVAL, "read of %d bases exceeds the supported maximum %d", l_max, FEM_MAX_READ_LEN);
  pl->Q = pl->e + 1 + pl->a;
  pl->s = pl->k / pl->step + (pl->k % pl->step > 0);
  pl->l_max = std::max(l_max, 1);
  pl->n_words_max = (pl->l_max + 15) / 16 + 1;
  pl->ns_max = std::max(pl->l_max - pl->k + 1, 1);
  // widest DP table of the batch: W = n_g - Q*s + 1 columns for the longest read's group 0
  pl->w_max = std::max(pl->ns_max / pl->step - pl->Q * pl->s + 1, 1);
  // one lane per group pays Q*W serial steps, the warp scan pays step*Q*ceil(W/32) rounds of ~7x the cost
  pl->lane_dp = c->seed_dp_mode == 1 ? 1 : c->seed_dp_mode == 2 ? 0 : pl->w_max <= 8 * pl->step;
  pl->seed_smem = 4 * seed_smem_words_per_warp(pl->n_words_max, pl->ns_max, pl->w_max, pl->Q, pl->step, pl->lane_dp) * 4 + SEED_TABLE_BYTES; // 4 warps + the base table
  // several read-strands per warp where the lane form leaves most lanes idle (k_seed_multi.cuh).  Automatic
  // choice: only while the lookup table stays in L2 (k <= 12, 67 MB; with k = 14 its gathers go to HBM and the
  // one-read-strand form's larger number of warps in flight wins: cfg5 10 ms against 22 ms) and while the
  // regions of 32/step read-strands are small enough for ten blocks per SM (each warp works through its
  // read-strands one after the other, so it needs many neighbours), and only for the shapes it was measured
  // to win on -- reads of up to 128 bases (one round of base coding per read-strand) with at most 7 seeds per
  // group: cfg2 2.1 against 2.9 ms, cfg1 0.20 against 0.29 ms per 1 M / 0.1 M reads; on cfg3 (150 bases, 9 seeds per
  // group) it is bound by the shared-memory pipe (ncu: short-scoreboard and MIO-throttle stalls) at 6.2 against
  // 3.7 ms, so the one-read-strand form stays there.
  pl->seed_R = 0;
  // over-long reads are marked by k_seed alone
  if (!over_long_reads && (c->seed_dp_mode == 0 || c->seed_dp_mode == 3) && pl->lane_dp && pl->w_max <= 254 && pl->step <= 32 && pl->ns_max < 65536) {
    pl->seed_words_per_rs = seed_multi_words_per_rs(pl->n_words_max, pl->ns_max, pl->w_max, pl->Q, pl->step);
    // read-strands per warp: as many as fill the DP's lanes, but at most four -- smaller regions mean more warps
    // beside the L1 half of the SM (cfg2 per 1 M reads: 8 per warp 1.78 ms, 6: 1.66, 4: 1.65, 2: 2.03)
    int R = std::min(32 / pl->step, 4);
    if (const char *rv = getenv("FEM_SEED_R")) { const int v = atoi(rv); if (v >= 1 && v <= 32 / pl->step) R = v; } // experiments
    size_t bytes = 16ull * ((size_t)R * pl->seed_words_per_rs + SEED_MULTI_SCRATCH);
    bool use = c->seed_dp_mode == 3 || (pl->k <= 12 && bytes <= 24 * 1024 && pl->l_max <= 128 && pl->Q <= 7);
    if (c->seed_dp_mode == 3)
      while (R > 1 && bytes > 56 * 1024) { --R; bytes = 16ull * ((size_t)R * pl->seed_words_per_rs + SEED_MULTI_SCRATCH); }
    if (use && bytes <= c->smem_optin) { pl->seed_R = R; pl->seed_smem = bytes; }
  }
  pl->cand_smem = (size_t)4 * (c->sort_capacity + c->acc_capacity) * 8;
  // hash path: room for 4x the expected location-list total of a group on a uniform index
  uint32_t T = c->hash_slots;
  if (!T) {
    const double mean = (double)c->n_occ / (double)(1ull << (2 * pl->k)) * pl->Q;
    T = pow2_ceil_u32((uint32_t)std::min(1024.0, std::max(128.0, 4.0 * mean)));
  }
  pl->hash_T = T;
  pl->hash_shift = 64;
  for (uint32_t t = T; t > 1; t >>= 1) --pl->hash_shift;
  pl->hash_smem = 4 * cand_hash_smem_bytes_per_warp(T);
  // pair path: value buffer for a group's expected location-list total plus six standard deviations,
  // 32 bitmap bits per expected entry (a few false flags per group)
  {
    const double mean = (double)c->n_occ / (double)(1ull << (2 * pl->k)) * pl->Q;
    uint32_t VC = c->pair_values;
    if (!VC) VC = (uint32_t)std::min(2048.0, std::max(64.0, mean + 6.0 * std::sqrt(mean) + 32.0));
    VC = (VC + 63u) & ~63u;
    uint32_t M = c->pair_bits;
    if (!M) M = pow2_ceil_u32((uint32_t)std::min(32768.0, std::max(1024.0, 32.0 * mean)));
    M = std::min<uint32_t>(M, 32768u); // bitmap positions are kept in 16 bits, 0xFFFF = dropped
    pl->pair_M = M;
    pl->pair_VC = VC;
    pl->pair_shift = 32;
    for (uint32_t t = M; t > 1; t >>= 1) --pl->pair_shift;
    pl->pair_smem = 4 * cand_pair_smem_bytes_per_warp(M, VC);
  }
  pl->first_path = c->cand_mode == 1 ? 0 : c->cand_mode == 2 ? 1 : (pl->a >= 1 ? 2 : 1);
  {
    // register form of the pair path: its slots hold runs of up to UA * LPR - 3 entries (and a last run of 96); a
    // run that does not fit sends its read-strand to k_cand_pair at ~2.5 x the cost, so it is chosen while the runs
    // of a uniform index stay below that by 3.3 standard deviations (~2 % of the read-strands of a batch hand on)
    const double mean = (double)c->n_occ / (double)(1ull << (2 * pl->k));
    const int lpr = pl->Q > 1 ? 32 / (pl->Q - 1) : 32;
    const bool shape = pl->Q == 5 || pl->Q == 7 || pl->Q == 9;
    const bool fits = mean + 3.3 * std::sqrt(mean) + 1.0 <= (double)(REG_UA * lpr - 3) && mean + 6.0 * std::sqrt(mean) <= 32.0 * PREG_UB;
    // one bitmap of twice k_cand_pair's bits in the same shared memory (positions stay below 2^15)
    pl->reg_M = std::max<uint32_t>(2048u, std::min<uint32_t>(32768u, 2u * pl->pair_M));
    pl->reg_smem = 4 * cand_reg_smem_bytes_per_warp(pl->reg_M);
    pl->pair_reg = pl->a >= 1 && shape && c->cand_mode != 6 && (c->cand_mode == 5 || ((c->cand_mode == 0 || c->cand_mode == 3) && fits)) &&
                   pl->reg_smem <= c->smem_optin;
    if (c->cand_mode == 5 || c->cand_mode == 6) pl->first_path = pl->a >= 1 ? 2 : 1;
  }
  {
    // sparse index: all groups of a read-strand at once when their lists are expected to hold <= 128 entries
    // (expectation over random q-grams; the DP picks the rarest, cfg5: 104 expected, 47 found)
    const double per_rs = (double)c->n_occ / (double)(1ull << (2 * pl->k)) * pl->Q * pl->step;
    const bool fits = (uint32_t)(pl->Q * pl->step) <= SMALL_SEEDS_MAX;
    const bool narrow = (uint32_t)(pl->Q * pl->step) <= 32u && per_rs <= 0.5 * SMALL_CAP_NARROW;
    pl->small_R = c->small_shape == 4 ? 4 : c->small_shape == 1 ? ((uint32_t)(pl->Q * pl->step) <= 32u ? 1 : 4) : (narrow ? 1 : 4);
    pl->small_smem = 4 * (pl->small_R == 1 ? SmallShape<1>::bytes_per_warp() : SmallShape<4>::bytes_per_warp());
    pl->small_inv_Q = pl->Q > 1 ? (uint32_t)(0x100000000ull / (uint64_t)pl->Q) + 1u : 0u;
    if (c->cand_mode == 4 || (c->cand_mode == 0 && fits && per_rs <= 128.0)) pl->first_path = 3;
    if (c->cand_mode == 3 && pl->a < 1) pl->first_path = 1;
  }
  return 0;
}

static int slot_reserve(femgpu_ctx *c, Slot &s, const Plan &pl, uint32_t n_reads, uint64_t n_bases) {
  const uint64_t n_rs = 2ull * n_reads;
  const uint64_t n_task = n_rs * pl.step;
  TRY(s.bases.reserve(n_bases + 16));
  TRY(s.read_off.reserve((n_reads + 1ull) * 8));
  TRY(s.codes.reserve(2 * n_bases + 64)); // k_planes reads nine aligned words from a plane word's first code on
  TRY(s.rs_status.reserve(n_rs + 16));
  TRY(s.read_flags.reserve(n_reads + 16));
  TRY(s.rs_before.reserve(n_rs * 4 + 16));
  TRY(s.rs_max_f.reserve(n_rs * 4 + 16));
  TRY(s.rs_sum_f.reserve(n_rs * 4 + 16));
  TRY(s.seed_begin.reserve(n_task * pl.Q * 4 + 16));
  TRY(s.seed_freq.reserve(n_task * pl.Q * 4 + 16));
  TRY(s.seed_start.reserve(n_task * pl.Q * 4 + 16));
  TRY(s.rs_cand_first.reserve(n_rs * 8 + 16));
  TRY(s.rs_cand_count.reserve(n_rs * 4 + 16));
  TRY(s.rs_map_count.reserve(n_rs * 4 + 16));
  TRY(s.rs_map_first.reserve((n_rs + 1) * 8));
  TRY(s.block_sums.reserve((n_rs / SCAN_BLOCK + 2) * 8));
  TRY(s.overflow_list.reserve(n_rs * 4 + 16));
  TRY(s.overflow_list2.reserve(n_rs * 4 + 16));
  TRY(s.counters.reserve(sizeof(Counters)));
  TRY(s.h_counters.reserve(sizeof(Counters)));
  TRY(s.h_first.reserve((n_reads + 1ull) * 8));
  uint64_t want = c->arena_candidates > 0 ? (uint64_t)c->arena_candidates : std::max<uint64_t>(1u << 20, 16ull * n_reads);
  if (want > s.cand_cap) {
    TRY(s.cand.reserve(want * 8));
    TRY(s.cand_rs.reserve(want * 4));
    TRY(s.result.reserve(want * 4));
    s.cand_cap = want;
  }
  return 0;
}

static int grow_arena(Slot &s) {
  const uint64_t want = s.cand_cap * 2;
  s.cand_cap = 0;
  TRY(s.cand.reserve(want * 8));
  TRY(s.cand_rs.reserve(want * 4));
  TRY(s.result.reserve(want * 4));
  s.cand_cap = want;
  return 0;
}

__global__ void k_gather_u32(const uint32_t *src, const uint32_t *idx, uint32_t n, uint32_t *dst) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[idx[i]];
}

static int read_counters(Slot &s, Counters *out) {
  CU(cudaMemcpyAsync(s.h_counters.p, s.counters.p, sizeof(Counters), cudaMemcpyDeviceToHost, s.stream));
  CU(slot_sync(s));
  *out = *s.h_counters.as<Counters>();
  return 0;
}

// k_seed .. candidates resident in the arena.  Records events 1 (after seed) and 2 (after cand).
static int run_seed_and_candidates(femgpu_ctx *c, Slot &s, const Plan &pl) {
  const uint32_t n_reads = s.n_reads;
  const uint64_t n_rs = 2ull * n_reads;
  s.n_launches = 0;
  CU(cudaMemsetAsync(s.counters.p, 0, sizeof(Counters), s.stream));
  if (n_reads == 0) {
    s.n_cand = 0;
    cudaEventRecord(s.ev[1], s.stream);
    cudaEventRecord(s.ev[2], s.stream);
    return 0;
  }
  Counters *dc = s.counters.as<Counters>();
  SeedArgs sa;
  sa.bases = s.bases.as<char>();
  sa.read_off = s.read_off.as<uint64_t>();
  sa.n_reads = n_reads;
  sa.lut = c->lut.as<uint32_t>();
  sa.base_table = c->base_table.as<uint8_t>();
  sa.inf = (uint32_t)c->n_occ;
  sa.k = pl.k; sa.step = pl.step; sa.e = pl.e; sa.a = pl.a; sa.Q = pl.Q; sa.s = pl.s;
  sa.n_words_max = pl.n_words_max; sa.ns_max = pl.ns_max; sa.w_max = pl.w_max;
  sa.lane_dp = pl.lane_dp;
  sa.smem_words = (uint32_t)seed_smem_words_per_warp(pl.n_words_max, pl.ns_max, pl.w_max, pl.Q, pl.step, pl.lane_dp);
  {
    const uint32_t groups = pl.lane_dp ? (uint32_t)pl.step : 1u;
    sa.amin_word = (uint32_t)pl.n_words_max + 2u * (uint32_t)pl.ns_max + groups * (uint32_t)(pl.w_max + 1);
    sa.sel_word = sa.amin_word + (groups * (uint32_t)pl.Q * (uint32_t)pl.w_max + 1u) / 2u;
  }
  sa.codes = s.codes.as<uint8_t>();
  sa.read_flags = s.read_flags.as<uint8_t>();
  sa.rs_status = s.rs_status.as<uint8_t>();
  sa.rs_before = s.rs_before.as<uint32_t>();
  sa.rs_max_f = s.rs_max_f.as<uint32_t>();
  sa.rs_sum_f = s.rs_sum_f.as<uint32_t>();
  sa.seed_begin = s.seed_begin.as<uint32_t>();
  sa.seed_freq = s.seed_freq.as<uint32_t>();
  sa.seed_start = s.seed_start.as<uint32_t>();
  sa.total_before = &dc->total_before;
  if (pl.seed_R > 0) {
    SeedMultiArgs sm;
    sm.a = sa;
    sm.R = pl.seed_R;
    sm.words_per_rs = pl.seed_words_per_rs;
    sm.inv_step = (uint32_t)(0x100000000ull / (uint64_t)pl.step) + 1u;
    sm.inv_Q = (uint32_t)(0x100000000ull / (uint64_t)pl.Q) + 1u;
    sm.inv_stepQ = (uint32_t)(0x100000000ull / (uint64_t)(pl.step * pl.Q)) + 1u;
    if (pl.seed_smem > 48 * 1024) CU(cudaFuncSetAttribute(k_seed_multi, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.seed_smem));
    // L1 / shared-memory split: the kernel's gathers and byte stores live on L1 hits, so half of the 256 KB goes to
    // L1 (six blocks of ~20 KB per SM instead of ten).  Measured on cfg2 per 1 M reads: default split (ten blocks)
    // 2.05 ms, 80 % shared 2.12, 70 % 1.81, 50 % 1.78, 40 % / 30 % 1.88, 100 % 4.0.  FEM_SEED_CARVEOUT overrides.
    {
      int carve = 50;
      if (const char *cv = getenv("FEM_SEED_CARVEOUT")) carve = atoi(cv);
      if (carve >= 0 && carve <= 100) cudaFuncSetAttribute(k_seed_multi, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
    }
    const uint64_t per_block = 4ull * (uint64_t)pl.seed_R;
    k_seed_multi<<<(unsigned)((n_rs + per_block - 1) / per_block), 128, pl.seed_smem, s.stream>>>(sm);
  } else {
#define FEM_SEED_LAUNCH(QT)                                                                                                          \
  case QT:                                                                                                                         \
    if (pl.seed_smem > 48 * 1024) CU(cudaFuncSetAttribute(k_seed<QT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.seed_smem)); \
    k_seed<QT><<<(unsigned)((n_rs + 3) / 4), 128, pl.seed_smem, s.stream>>>(sa);                                                  \
    break;
    switch (pl.Q) { // Q = e + 1 + a seeds per group, a compile-time constant of the kernel
      FEM_SEED_LAUNCH(1) FEM_SEED_LAUNCH(2) FEM_SEED_LAUNCH(3) FEM_SEED_LAUNCH(4) FEM_SEED_LAUNCH(5)
      FEM_SEED_LAUNCH(6) FEM_SEED_LAUNCH(7) FEM_SEED_LAUNCH(8) FEM_SEED_LAUNCH(9) FEM_SEED_LAUNCH(10)
      default: return set_error(FEMGPU_EINVAL, "Q = %d seeds per group outside 1..%d", pl.Q, FEM_MAX_Q);
    }
#undef FEM_SEED_LAUNCH
  }
  ++s.n_launches;
  CU(cudaGetLastError());
  cudaEventRecord(s.ev[1], s.stream);

  CandArgs ca;
  memset(&ca, 0, sizeof ca);
  ca.occ = c->occ.as<uint64_t>();
  ca.seed_begin = sa.seed_begin; ca.seed_freq = sa.seed_freq; ca.seed_start = sa.seed_start;
  ca.rs_status = sa.rs_status;
  ca.read_off = sa.read_off;
  ca.rec_len = c->rec_len.as<uint32_t>();
  ca.n_rec = c->n_rec;
  ca.n_rs = (uint32_t)n_rs;
  ca.step = pl.step; ca.e = pl.e; ca.a = pl.a; ca.Q = pl.Q;
  ca.cand_count = &dc->cand_count;
  ca.rs_cand_first = s.rs_cand_first.as<uint64_t>();
  ca.rs_cand_count = s.rs_cand_count.as<uint32_t>();
  ca.overflow_list = s.overflow_list.as<uint32_t>();
  ca.overflow_count = &dc->overflow_count;
  ca.error_flag = &dc->error_flag;
  ca.S = c->sort_capacity; ca.Acap = c->acc_capacity;
  if (pl.cand_smem > 48 * 1024) CU(cudaFuncSetAttribute(k_cand, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.cand_smem));
  if (pl.hash_smem > 48 * 1024) CU(cudaFuncSetAttribute(k_cand_hash, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.hash_smem));
  if (pl.pair_smem > 48 * 1024) CU(cudaFuncSetAttribute(k_cand_pair, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.pair_smem));
  if (pl.pair_reg && pl.reg_smem > 48 * 1024) {
    CU(cudaFuncSetAttribute(k_cand_reg<9, REG_UA, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.reg_smem));
    CU(cudaFuncSetAttribute(k_cand_reg<9, REG_UA, 7>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.reg_smem));
    CU(cudaFuncSetAttribute(k_cand_reg<9, REG_UA, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.reg_smem));
    CU(cudaFuncSetAttribute(k_cand_reg<7, REG_UA, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.reg_smem));
    CU(cudaFuncSetAttribute(k_cand_reg<5, REG_UA, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.reg_smem));
  }

  Counters hc;
  for (int attempt = 0;; ++attempt) {
    if (attempt > 12) return set_error(FEMGPU_ELIMIT, "candidate arena keeps overflowing");
    ca.cand = s.cand.as<uint64_t>();
    ca.cand_rs = s.cand_rs.as<uint32_t>();
    ca.cand_cap = s.cand_cap;
    ca.rs_list = nullptr;
    ca.n_list = 0;
    const bool reg_first = pl.first_path == 2 && pl.pair_reg;
    if (pl.first_path == 3 && pl.small_R == 1) k_cand_small<1><<<(unsigned)((n_rs + 3) / 4), 128, pl.small_smem, s.stream>>>(ca, pl.small_inv_Q);
    else if (pl.first_path == 3) k_cand_small<4><<<(unsigned)((n_rs + 3) / 4), 128, pl.small_smem, s.stream>>>(ca, pl.small_inv_Q);
    else if (reg_first) {
      const unsigned grid = (unsigned)((n_rs + 3) / 4);
      static const int variant = getenv("FEM_REG_VARIANT") ? atoi(getenv("FEM_REG_VARIANT")) : 1; // experiments: resident blocks per SM (1: 8 blocks of 64 registers, measured best: 14.5 against 15.0 ms)
      if (pl.Q == 9 && variant == 1) k_cand_reg<9, REG_UA, 8><<<grid, 128, pl.reg_smem, s.stream>>>(ca, pl.reg_M);
      else if (pl.Q == 9 && variant == 2) k_cand_reg<9, REG_UA, 7><<<grid, 128, pl.reg_smem, s.stream>>>(ca, pl.reg_M);
      else if (pl.Q == 9) k_cand_reg<9, REG_UA, 6><<<grid, 128, pl.reg_smem, s.stream>>>(ca, pl.reg_M);
      else if (pl.Q == 7) k_cand_reg<7, REG_UA, 6><<<grid, 128, pl.reg_smem, s.stream>>>(ca, pl.reg_M);
      else k_cand_reg<5, REG_UA, 6><<<grid, 128, pl.reg_smem, s.stream>>>(ca, pl.reg_M);
      ++s.n_launches;
      // what it hands on goes to k_cand_pair without a trip to the host: the list's length is read on the device;
      // REG_CHAIN warps are launched, a longer list (seen below) gets a second launch for its tail
      CandArgs c2 = ca;
      c2.rs_list = s.overflow_list.as<uint32_t>();
      c2.n_list = 0xFFFFFFFFu;
      c2.n_list_dev = &dc->overflow_count;
      c2.list_first = 0;
      c2.overflow_list = s.overflow_list2.as<uint32_t>();
      c2.overflow_count = &dc->overflow_count2;
      const uint32_t chain = (uint32_t)std::min<uint64_t>(n_rs, REG_CHAIN);
      k_cand_pair<<<(chain + 3) / 4, 128, pl.pair_smem, s.stream>>>(c2, pl.pair_M, pl.pair_VC);
    }
    else if (pl.first_path == 2) k_cand_pair<<<(unsigned)((n_rs + 3) / 4), 128, pl.pair_smem, s.stream>>>(ca, pl.pair_M, pl.pair_VC);
    else if (pl.first_path == 1) k_cand_hash<<<(unsigned)((n_rs + 3) / 4), 128, pl.hash_smem, s.stream>>>(ca, pl.hash_T, pl.hash_shift);
    else k_cand<<<(unsigned)((n_rs + 3) / 4), 128, pl.cand_smem, s.stream>>>(ca);
    ++s.n_launches;
    CU(cudaGetLastError());
    TRY(read_counters(s, &hc));
    bool arena_full = hc.error_flag == 1;
    uint32_t n1 = hc.overflow_count;                         // read-strands handed on by the first path ...
    const uint32_t *list1 = s.overflow_list.as<uint32_t>();
    if (reg_first) {
      if (!arena_full && hc.overflow_count > REG_CHAIN) { // the tail of a list longer than the chained launch covered
        CandArgs c2 = ca;
        c2.rs_list = s.overflow_list.as<uint32_t>();
        c2.n_list = hc.overflow_count;
        c2.list_first = (uint32_t)REG_CHAIN;
        c2.overflow_list = s.overflow_list2.as<uint32_t>();
        c2.overflow_count = &dc->overflow_count2;
        k_cand_pair<<<(hc.overflow_count - (uint32_t)REG_CHAIN + 3) / 4, 128, pl.pair_smem, s.stream>>>(c2, pl.pair_M, pl.pair_VC);
        ++s.n_launches;
        CU(cudaGetLastError());
        TRY(read_counters(s, &hc));
        arena_full = hc.error_flag == 1;
      }
      n1 = hc.overflow_count2;                               // ... here: by k_cand_pair after k_cand_reg
      list1 = s.overflow_list2.as<uint32_t>();
    }
    if (pl.first_path != 0 && !arena_full && n1) {
      // ---- sort path for the read-strands the first path handed on ----------------------------------
      TRY(s.hash_list.reserve((size_t)n1 * 4));
      CU(cudaMemcpyAsync(s.hash_list.p, list1, (size_t)n1 * 4, cudaMemcpyDeviceToDevice, s.stream));
      CU(cudaMemsetAsync(&dc->overflow_count, 0, 2 * sizeof(uint32_t), s.stream)); // both lists' counters (adjacent)
      ca.rs_list = s.hash_list.as<uint32_t>();
      ca.n_list = n1;
      k_cand<<<(n1 + 3) / 4, 128, pl.cand_smem, s.stream>>>(ca);
      ++s.n_launches;
      CU(cudaGetLastError());
      TRY(read_counters(s, &hc));
      arena_full = hc.error_flag == 1;
    } else if (reg_first) hc.overflow_count = 0; // everything handed on was resolved by k_cand_pair
    if (!arena_full && hc.overflow_count) {
      // ---- big path: plan scratch slices on the host, run in rounds within the scratch limit ----
      const uint32_t n_over = hc.overflow_count;
      TRY(s.h_overflow.reserve((size_t)n_over * 4));
      CU(cudaMemcpyAsync(s.h_overflow.p, s.overflow_list.p, (size_t)n_over * 4, cudaMemcpyDeviceToHost, s.stream));
      CU(slot_sync(s));
      const uint32_t *ol = s.h_overflow.as<uint32_t>();
      std::vector<uint32_t> order(ol, ol + n_over);
      std::sort(order.begin(), order.end()); // deterministic processing order
      // frequencies of exactly those read-strands (small gather kernel, one D2H)
      TRY(s.big_list.reserve((size_t)n_over * 4));
      TRY(s.big_S.reserve((size_t)n_over * 4));
      TRY(s.big_A.reserve((size_t)n_over * 4));
      TRY(s.h_plan.reserve((size_t)n_over * 8));
      CU(cudaMemcpyAsync(s.big_list.p, order.data(), (size_t)n_over * 4, cudaMemcpyHostToDevice, s.stream));
      k_gather_u32<<<(n_over + 255) / 256, 256, 0, s.stream>>>(s.rs_max_f.as<uint32_t>(), s.big_list.as<uint32_t>(), n_over, s.big_S.as<uint32_t>());
      k_gather_u32<<<(n_over + 255) / 256, 256, 0, s.stream>>>(s.rs_sum_f.as<uint32_t>(), s.big_list.as<uint32_t>(), n_over, s.big_A.as<uint32_t>());
      s.n_launches += 2;
      CU(cudaMemcpyAsync(s.h_plan.p, s.big_S.p, (size_t)n_over * 4, cudaMemcpyDeviceToHost, s.stream));
      CU(cudaMemcpyAsync(s.h_plan.as<uint32_t>() + n_over, s.big_A.p, (size_t)n_over * 4, cudaMemcpyDeviceToHost, s.stream));
      CU(slot_sync(s));
      const uint32_t *mx = s.h_plan.as<uint32_t>(), *sm = mx + n_over;
      const uint64_t limit_elems = (uint64_t)c->scratch_limit_bytes / 8;
      uint32_t done = 0;
      while (done < n_over && !arena_full) {
        std::vector<uint32_t> list, bS, bA;
        std::vector<uint64_t> boff;
        uint64_t used = 0;
        while (done < n_over) {
          if (sm[done] == 0xFFFFFFFFu) return set_error(FEMGPU_ELIMIT, "read-strand %u: location lists too long", order[done]);
          const uint64_t S = std::max<uint64_t>(64, 2ull * mx[done] + sm[done]);
          const uint64_t Acap = std::max<uint64_t>(1, sm[done]);
          if (S + Acap > limit_elems || S > 0xFFFFFFFFull)
            return set_error(FEMGPU_ELIMIT, "read-strand %u needs %llu bytes of scratch (limit %lld)", order[done],
                             (unsigned long long)((S + Acap) * 8), (long long)c->scratch_limit_bytes);
          if (used + S + Acap > limit_elems && !list.empty()) break;
          list.push_back(order[done]); boff.push_back(used); bS.push_back((uint32_t)S); bA.push_back((uint32_t)Acap);
          used += S + Acap;
          ++done;
        }
        const size_t nb = list.size();
        TRY(s.scratch.reserve(used * 8));
        TRY(s.big_list.reserve(nb * 4)); TRY(s.big_off.reserve(nb * 8)); TRY(s.big_S.reserve(nb * 4)); TRY(s.big_A.reserve(nb * 4));
        CU(cudaMemcpyAsync(s.big_list.p, list.data(), nb * 4, cudaMemcpyHostToDevice, s.stream));
        CU(cudaMemcpyAsync(s.big_off.p, boff.data(), nb * 8, cudaMemcpyHostToDevice, s.stream));
        CU(cudaMemcpyAsync(s.big_S.p, bS.data(), nb * 4, cudaMemcpyHostToDevice, s.stream));
        CU(cudaMemcpyAsync(s.big_A.p, bA.data(), nb * 4, cudaMemcpyHostToDevice, s.stream));
        ca.big_list = s.big_list.as<uint32_t>(); ca.big_off = s.big_off.as<uint64_t>();
        ca.big_S = s.big_S.as<uint32_t>(); ca.big_A = s.big_A.as<uint32_t>();
        ca.scratch = s.scratch.as<uint64_t>();
        ca.n_big = (uint32_t)nb;
        k_cand_big<<<(unsigned)((nb + 3) / 4), 128, 0, s.stream>>>(ca);
        ++s.n_launches;
        CU(cudaGetLastError());
        TRY(read_counters(s, &hc)); // also keeps the host vectors alive until the copies are done
        if (hc.error_flag == 2) return set_error(FEMGPU_ECUDA, "internal: big-path scratch plan too small");
        arena_full = hc.error_flag == 1;
      }
    }
    if (!arena_full) break;
    TRY(grow_arena(s));
    CU(cudaMemsetAsync(&dc->cand_count, 0, sizeof(unsigned long long), s.stream));
    CU(cudaMemsetAsync(&dc->overflow_count, 0, 2 * sizeof(uint32_t), s.stream)); // and overflow_count2
    CU(cudaMemsetAsync(&dc->error_flag, 0, sizeof(int), s.stream));
  }
  s.n_cand = hc.cand_count;
  s.stats.num_candidates_before_filter = hc.total_before;
  s.stats.num_candidates = hc.cand_count;
  if (s.n_cand) {
    // bit planes of the reads that have candidates (k_verify's and k_trace's plane forms); timed with this stage
    const uint32_t plane_words = (uint32_t)((std::max(s.l_max, 1) + 31) / 32);
    TRY(s.planes.reserve(n_rs * 3ull * plane_words * 4 + 16));
    TRY(s.rs_has_n.reserve(n_rs + 16));
    CU(cudaMemsetAsync(s.rs_has_n.p, 0, n_rs, s.stream));
    k_planes<<<(unsigned)((n_rs * plane_words + 255) / 256), 256, 0, s.stream>>>(s.codes.as<uint8_t>(), s.read_off.as<uint64_t>(), s.rs_cand_count.as<uint32_t>(),
                                                                             (uint32_t)n_rs, plane_words, s.planes.as<uint32_t>(), s.rs_has_n.as<uint8_t>());
    ++s.n_launches;
    CU(cudaGetLastError());
  }
  cudaEventRecord(s.ev[2], s.stream);
  return 0;
}

// k_verify -> scan -> k_compact.  Records events 3 (after verify) and 4 (after compaction); leaves
// s.n_maps and the device-side `maps` / `rs_map_first` arrays.
static int run_verify_and_compact(femgpu_ctx *c, Slot &s, const Plan &pl) {
  const uint64_t n_rs = 2ull * s.n_reads;
  Counters *dc = s.counters.as<Counters>();
  s.n_maps = 0;
  if (s.n_reads == 0) {
    cudaEventRecord(s.ev[3], s.stream);
    cudaEventRecord(s.ev[4], s.stream);
    cudaEventRecord(s.ev[6], s.stream);
    return 0;
  }
  CU(cudaMemsetAsync(s.rs_map_count.p, 0, n_rs * 4, s.stream));
  if (s.n_cand) {
    VerifyArgs va;
    va.ref.ref2 = c->ref2.as<uint32_t>();
    va.ref.refn = c->ref_has_n ? c->refn.as<uint32_t>() : nullptr;
    va.ref.rec_base = c->rec_base.as<uint64_t>();
    va.ref.rec_len = c->rec_len.as<uint32_t>();
    va.ref.n_rec = c->n_rec;
    va.codes = s.codes.as<uint8_t>();
    va.read_off = s.read_off.as<uint64_t>();
    va.cand = s.cand.as<uint64_t>();
    va.cand_rs = s.cand_rs.as<uint32_t>();
    va.rs_cand_first = s.rs_cand_first.as<uint64_t>();
    va.rs_cand_count = s.rs_cand_count.as<uint32_t>();
    va.n_cand = s.n_cand;
    va.e = pl.e;
    va.result = s.result.as<uint32_t>();
    va.rs_map_count = s.rs_map_count.as<uint32_t>();
    va.cols_executed = &dc->cols_executed;
    const uint32_t plane_words = (uint32_t)((std::max(s.l_max, 1) + 31) / 32);
    va.refp0 = c->refp0.as<uint32_t>(); va.refp1 = c->refp1.as<uint32_t>();
    va.planes = s.planes.as<uint32_t>();
    va.plane_words = plane_words;
    va.rs_has_n = s.rs_has_n.as<uint8_t>();
    if (!c->verify_planes) va.refp0 = nullptr; // tests / experiments: the general form for every candidate
    const unsigned grid = (unsigned)((s.n_cand + 255) / 256);
    if (c->ref_has_n) k_verify<true><<<grid, 256, 0, s.stream>>>(va);
    else k_verify<false><<<grid, 256, 0, s.stream>>>(va);
    ++s.n_launches;
    CU(cudaGetLastError());
  }
  cudaEventRecord(s.ev[3], s.stream);
  const unsigned n_blocks = (unsigned)((n_rs + SCAN_BLOCK - 1) / SCAN_BLOCK);
  k_scan_reduce<<<n_blocks, SCAN_BLOCK, 0, s.stream>>>(s.rs_map_count.as<uint32_t>(), n_rs, s.block_sums.as<uint64_t>());
  k_scan_blocks<<<1, 32, 0, s.stream>>>(s.block_sums.as<uint64_t>(), n_blocks, (uint64_t *)&dc->total_maps);
  k_scan_apply<<<n_blocks, SCAN_BLOCK, 0, s.stream>>>(s.rs_map_count.as<uint32_t>(), n_rs, s.block_sums.as<uint64_t>(),
                                                     s.rs_map_first.as<uint64_t>(), (uint64_t *)&dc->total_maps);
  s.n_launches += 3;
  CU(cudaGetLastError());
  Counters hc;
  TRY(read_counters(s, &hc));
  s.n_maps = hc.total_maps;
  s.cols_executed = hc.cols_executed;
  s.stats.num_mappings = hc.total_maps;
  if (s.n_maps) {
    TRY(s.maps.reserve(s.n_maps * 16));
    TRY(s.map_rs.reserve(s.n_maps * 4));
    CompactArgs pa;
    pa.cand = s.cand.as<uint64_t>();
    pa.result = s.result.as<uint32_t>();
    pa.rs_cand_first = s.rs_cand_first.as<uint64_t>();
    pa.rs_cand_count = s.rs_cand_count.as<uint32_t>();
    pa.rs_map_first = s.rs_map_first.as<uint64_t>();
    pa.rs_map_count = s.rs_map_count.as<uint32_t>();
    pa.n_rs = (uint32_t)n_rs;
    pa.maps = s.maps.as<ulonglong2>();
    pa.map_rs = s.map_rs.as<uint32_t>();
    k_compact<<<(unsigned)((n_rs + 31) / 32), 256, 0, s.stream>>>(pa); // 8 warps x (32 / COMPACT_LPR) read-strands
    ++s.n_launches;
    CU(cudaGetLastError());
  }
  cudaEventRecord(s.ev[4], s.stream);
  // ---- K6: start + CIGAR of every mapping (src/align.c:279-499) ----------------------------------
  if (s.n_maps && c->device_traceback) {
    TRY(s.aln.reserve(s.n_maps * sizeof(femgpu_alignment)));
    TraceArgs ta;
    ta.ref.ref2 = c->ref2.as<uint32_t>();
    ta.ref.refn = c->ref_has_n ? c->refn.as<uint32_t>() : nullptr;
    ta.ref.rec_base = c->rec_base.as<uint64_t>();
    ta.ref.rec_len = c->rec_len.as<uint32_t>();
    ta.ref.n_rec = c->n_rec;
    ta.codes = s.codes.as<uint8_t>();
    ta.read_off = s.read_off.as<uint64_t>();
    ta.read_flags = s.read_flags.as<uint8_t>();
    ta.maps = s.maps.as<ulonglong2>();
    ta.map_rs = s.map_rs.as<uint32_t>();
    ta.n_maps = s.n_maps;
    ta.e = pl.e;
    ta.ref_canonical = c->ref_canonical ? 1 : 0;
    ta.aln = s.aln.as<uint32_t>();
    ta.refp0 = c->verify_planes ? c->refp0.as<uint32_t>() : nullptr; ta.refp1 = c->refp1.as<uint32_t>();
    ta.planes = s.planes.as<uint32_t>();
    ta.plane_words = (uint32_t)((std::max(s.l_max, 1) + 31) / 32);
    ta.rs_has_n = s.rs_has_n.as<uint8_t>();
    const unsigned grid = (unsigned)((s.n_maps + 127) / 128);
    const int cap = s.l_max <= 128 ? 128 : s.l_max <= 160 ? 160 : s.l_max <= 256 ? 256 : 1024; // columns of per-thread history (local memory)
#define LAUNCH_TRACE(N, L) k_trace<N, L><<<grid, 128, 0, s.stream>>>(ta)
    if (c->ref_has_n) { if (cap == 128) LAUNCH_TRACE(true, 128); else if (cap == 160) LAUNCH_TRACE(true, 160); else if (cap == 256) LAUNCH_TRACE(true, 256); else LAUNCH_TRACE(true, 1024); }
    else { if (cap == 128) LAUNCH_TRACE(false, 128); else if (cap == 160) LAUNCH_TRACE(false, 160); else if (cap == 256) LAUNCH_TRACE(false, 256); else LAUNCH_TRACE(false, 1024); }
#undef LAUNCH_TRACE
    ++s.n_launches;
    CU(cudaGetLastError());
  }
  cudaEventRecord(s.ev[6], s.stream);
  return 0;
}
