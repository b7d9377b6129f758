// femgpu.cu -- C-ABI implementation (include/femgpu.h): context, resident data, batch pipeline.
//
// One context per GPU.  Per batch ("slot", two of them for double buffering) the pipeline is
//   H2D(bases, offsets) -> k_seed -> k_cand [-> k_cand_big] -> k_verify -> scan -> k_compact -> D2H
// on the slot's own stream.  All device buffers are grow-only and reused across batches.
#include "../../include/femgpu.h"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"
#include "k_cand.cuh"
#include "k_cand_hash.cuh"
#include "k_cand_pair.cuh"
#include "k_cand_reg.cuh"
#include "k_cand_small.cuh"
#include "k_index.cuh"
#include "k_peak.cuh"
#include "k_seed.cuh"
#include "k_seed_multi.cuh"
#include "k_trace.cuh"
#include "k_sam.cuh"
#include "k_verify.cuh"

// ------------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------------
static thread_local std::string g_last_error;
static int set_error(int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_last_error = buf;
  return code;
}
#define CU(call)                                                                                   \
  do {                                                                                             \
    cudaError_t err__ = (call);                                                                    \
    if (err__ != cudaSuccess)                                                                      \
      return set_error(err__ == cudaErrorMemoryAllocation ? FEMGPU_ENOMEM : FEMGPU_ECUDA,          \
                       "%s failed: %s (%s:%d)", #call, cudaGetErrorString(err__), __FILE__, __LINE__); \
  } while (0)

// ------------------------------------------------------------------------------------------------
// grow-only buffers
// ------------------------------------------------------------------------------------------------
struct DBuf {
  void *p = nullptr;
  size_t bytes = 0;
  int reserve(size_t need) {
    if (need <= bytes) return 0;
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
    size_t want = need + need / 4 + 256;
    if (cudaMalloc(&p, want) != cudaSuccess) {
      cudaGetLastError();
      want = need;
      if (cudaMalloc(&p, want) != cudaSuccess) {
        cudaGetLastError();
        p = nullptr;
        return set_error(FEMGPU_ENOMEM, "cudaMalloc of %zu bytes failed", need);
      }
    }
    bytes = want;
    return 0;
  }
  void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
  template <class T> T *as() const { return static_cast<T *>(p); }
};
struct HBuf { // pinned host memory
  void *p = nullptr;
  size_t bytes = 0;
  int reserve(size_t need) {
    if (need <= bytes) return 0;
    if (p) cudaFreeHost(p);
    p = nullptr;
    bytes = 0;
    size_t want = need + need / 4 + 256;
    if (cudaMallocHost(&p, want) != cudaSuccess) {
      cudaGetLastError();
      p = nullptr;
      return set_error(FEMGPU_ENOMEM, "cudaMallocHost of %zu bytes failed", want);
    }
    bytes = want;
    return 0;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; bytes = 0; }
  template <class T> T *as() const { return static_cast<T *>(p); }
};
#define TRY(x) do { int rc__ = (x); if (rc__) return rc__; } while (0)

// counters shared between kernels and host (one small device block per slot)
struct Counters {
  unsigned long long cand_count;
  unsigned long long total_before;
  unsigned long long total_maps;
  unsigned long long cols_executed;
  unsigned long long sam_total;
  unsigned long long sam_flagged; // reads k_sam left to the host
  unsigned long long sam_big;     // reads with 65..SAM_BIG_MAX mappings (k_sam_big_*)
  uint32_t overflow_count;
  uint32_t overflow_count2; // second list: what k_cand_pair (list mode) hands on after k_cand_reg
  int error_flag;
};

struct Slot {
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[8] = {};
  cudaEvent_t ev_block = nullptr; // cudaEventBlockingSync: slot_sync
  bool blocking_sync = false;     // copied from the context with every batch
  // inputs / per read-strand state
  DBuf bases, read_off, codes, rs_status, rs_before, rs_max_f, rs_sum_f;
  DBuf seed_begin, seed_freq, seed_start;
  DBuf rs_cand_first, rs_cand_count, rs_map_count, rs_map_first, block_sums;
  DBuf cand, cand_rs, result, maps, overflow_list, overflow_list2, counters;
  DBuf read_flags, map_rs, aln, planes, rs_has_n;
  DBuf big_list, big_off, big_S, big_A, scratch, hash_list;
  DBuf cand_sorted, rs_cand_sorted_first; // femgpu_candidates_batch only
  // SAM text on the device (k_sam): inputs of the batch, per-read sizes / offsets / flags, the text
  DBuf names, name_off, quals, has_qual, sam_bytes, sam_host, sam_off, sam_blocks, sam_text;
  DBuf sam_big_list, sam_big_ord, sam_big_rec_off;
  HBuf h_sam_text, h_sam_off, h_sam_host;
  femgpu_sam_inputs sam_in{};
  bool sam_pending = false, sam_on = false;
  uint64_t sam_total = 0;
  HBuf h_aln;
  HBuf h_counters, h_first, h_maps, h_overflow, h_plan, h_cands, h_before, h_status;
  uint64_t cand_cap = 0;
  // state of the batch in flight
  bool busy = false;
  bool has_aln = false;
  uint64_t sam_flagged = 0;
  char *r_sam_text = nullptr;     // where this batch's SAM text went on the host (caller's or ours)
  uint64_t *r_sam_off = nullptr;
  uint8_t *r_sam_host = nullptr;
  femgpu_params params{};
  uint32_t n_reads = 0;
  uint64_t n_bases = 0;
  int l_max = 0;          // longest read within FEM_MAX_READ_LEN
  uint32_t n_too_long = 0; // reads beyond it (left unmapped, counted as infeasible)
  uint64_t n_cand = 0, n_maps = 0;
  femgpu_stats stats{};
  float ms[6] = {0, 0, 0, 0, 0, 0};
  uint32_t n_launches = 0;
  uint64_t cols_executed = 0;
};

// wait for the slot's stream: spinning (lowest latency) or asleep on an event (leaves the core to the other threads)
static inline cudaError_t slot_sync(Slot &s) {
  if (!s.blocking_sync || !s.ev_block) return cudaStreamSynchronize(s.stream);
  cudaError_t e = cudaEventRecord(s.ev_block, s.stream);
  return e != cudaSuccess ? e : cudaEventSynchronize(s.ev_block);
}

struct femgpu_ctx {
  int device = 0;
  int sm_count = 0;
  size_t smem_optin = 0;
  // index
  DBuf lut, occ;
  DBuf base_table;     // k_seed: code and flags of every byte value, both strands (seed_base_table)
  uint64_t n_occ = 0;
  int idx_k = 0, idx_step = 0;
  // reference
  DBuf ref2, refn, rec_base, rec_len;
  DBuf refp0, refp1;          // bit planes of ref2 (k_ref_planes): bit 0 / bit 1 of the base codes, 32 bases per word
  DBuf rnames, rname_off;     // record names for k_sam (femgpu_set_reference_names)
  bool have_rnames = false;
  uint32_t n_rec = 0;
  uint64_t ref_total = 0;
  bool ref_has_n = false;
  bool ref_canonical = true; // every reference character is one of "ACGTN"
  // options
  int64_t scratch_limit_bytes = 8ll << 30;
  int64_t arena_candidates = 0; // 0 = automatic
  uint32_t sort_capacity = 1024, acc_capacity = 256;
  int l2_persist_lut = 1;
  int small_shape = 0;  // k_cand_small: 0 automatic, 1 one seed per lane (step x Q <= 32), 4 up to 128 seeds
  int cand_mode = 0;    // 0: pair path (a >= 1) or hash path (a = 0) first, then sort path, then global scratch;
                        // 1: sort path only; 2: hash path first; 3: pair path first (a >= 1); 4: small path first;
                        // 5: register form of the pair path first (k_cand_reg) whatever the index density;
                        // 6: pair path first, never its register form
  uint32_t pair_bits = 0, pair_values = 0; // 0 automatic (from the index density)
  uint32_t hash_slots = 0; // 0 automatic (from the index density), else a power of two 32..4096
  int verify_planes = 1; // k_verify: bit-plane form of the columns for candidates without ambiguous bases (0: general form only)
  int device_traceback = 1; // k_trace after compaction (0: the host computes every alignment)
  int blocking_sync = 0; // host waits sleep on an event instead of spinning in cudaStreamSynchronize (small hosts, one process per GPU)
  int seed_dp_mode = 0; // 0 automatic, 1 one lane per group, 2 warp scan, 3 several read-strands per warp (tests force each)
  Slot slot[2];
  int last_slot = 0;
};

// ------------------------------------------------------------------------------------------------
// life cycle
// ------------------------------------------------------------------------------------------------
extern "C" int femgpu_abi_version(void) { return FEMGPU_ABI_VERSION; }

extern "C" int femgpu_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

extern "C" const char *femgpu_last_error(void) { return g_last_error.c_str(); }

extern "C" femgpu_ctx *femgpu_create(int device) {
  int n = femgpu_device_count();
  if (n <= 0) { set_error(FEMGPU_ENODEV, "no CUDA device available (this library has no CPU fallback)"); return nullptr; }
  if (device < 0 || device >= n) { set_error(FEMGPU_EINVAL, "device %d out of range (0..%d)", device, n - 1); return nullptr; }
  if (cudaSetDevice(device) != cudaSuccess) { set_error(FEMGPU_ECUDA, "cudaSetDevice(%d) failed", device); return nullptr; }
  femgpu_ctx *c = new femgpu_ctx();
  c->device = device;
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  c->sm_count = prop.multiProcessorCount;
  c->smem_optin = prop.sharedMemPerBlockOptin;
  if (const char *b = getenv("FEM_BLOCKING_SYNC")) c->blocking_sync = atoi(b) != 0;
  if (const char *m = getenv("FEM_CAND_MODE")) { const int v = atoi(m); if (v >= 0 && v <= 6) c->cand_mode = v; } // experiments: see "cand_mode"
  if (const char *m = getenv("FEM_VERIFY_PLANES")) c->verify_planes = atoi(m) != 0; // experiments: see "verify_planes"
  if (const char *m = getenv("FEM_PAIR_BITS")) { const int v = atoi(m); if (v >= 1024 && v <= 32768 && (v & (v - 1)) == 0) c->pair_bits = (uint32_t)v; } // experiments: see "pair_bits"
  if (const char *m = getenv("FEM_SMALL_SHAPE")) { const int v = atoi(m); if (v == 1 || v == 4) c->small_shape = v; } // experiments: see "small_shape"
  if (const char *m = getenv("FEM_SEED_DP_MODE")) c->seed_dp_mode = atoi(m) >= 0 && atoi(m) <= 3 ? atoi(m) : 0; // experiments: see "seed_dp_mode"
  for (auto &s : c->slot) {
    if (cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) != cudaSuccess) {
      set_error(FEMGPU_ECUDA, "cudaStreamCreate failed: %s", cudaGetErrorString(cudaGetLastError()));
      delete c;
      return nullptr;
    }
    for (auto &e : s.ev) cudaEventCreate(&e);
    cudaEventCreateWithFlags(&s.ev_block, cudaEventBlockingSync | cudaEventDisableTiming);
  }
  {
    uint8_t table[SEED_TABLE_BYTES];
    seed_base_table(table);
    if (c->base_table.reserve(SEED_TABLE_BYTES) || cudaMemcpy(c->base_table.p, table, SEED_TABLE_BYTES, cudaMemcpyHostToDevice) != cudaSuccess) {
      set_error(FEMGPU_ECUDA, "base table upload failed: %s", cudaGetErrorString(cudaGetLastError()));
      femgpu_destroy(c);
      return nullptr;
    }
  }
  return c;
}

static void slot_release(Slot &s) {
  DBuf *d[] = {&s.bases, &s.read_off, &s.codes, &s.rs_status, &s.rs_before, &s.rs_max_f, &s.rs_sum_f, &s.seed_begin,
               &s.seed_freq, &s.seed_start, &s.rs_cand_first, &s.rs_cand_count, &s.rs_map_count, &s.rs_map_first,
               &s.block_sums, &s.read_flags, &s.map_rs, &s.aln, &s.planes, &s.rs_has_n, &s.cand, &s.cand_rs, &s.result, &s.maps, &s.overflow_list, &s.overflow_list2, &s.counters, &s.big_list,
               &s.big_off, &s.big_S, &s.big_A, &s.scratch, &s.hash_list, &s.cand_sorted, &s.rs_cand_sorted_first,
               &s.names, &s.name_off, &s.quals, &s.has_qual, &s.sam_bytes, &s.sam_host, &s.sam_off, &s.sam_blocks, &s.sam_text,
               &s.sam_big_list, &s.sam_big_ord, &s.sam_big_rec_off};
  for (auto *b : d) b->release();
  HBuf *h[] = {&s.h_sam_text, &s.h_sam_off, &s.h_sam_host, &s.h_aln, &s.h_counters, &s.h_first, &s.h_maps, &s.h_overflow, &s.h_plan, &s.h_cands, &s.h_before, &s.h_status};
  for (auto *b : h) b->release();
  for (auto &e : s.ev) if (e) cudaEventDestroy(e);
  if (s.ev_block) cudaEventDestroy(s.ev_block);
  if (s.stream) cudaStreamDestroy(s.stream);
}

extern "C" void femgpu_destroy(femgpu_ctx *c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  for (auto &s : c->slot) slot_release(s);
  c->lut.release(); c->occ.release(); c->base_table.release(); c->ref2.release(); c->refn.release(); c->refp0.release(); c->refp1.release(); c->rec_base.release(); c->rec_len.release(); c->rnames.release(); c->rname_off.release();
  delete c;
}

extern "C" int femgpu_set_option(femgpu_ctx *c, const char *key, int64_t value) {
  if (!c || !key) return set_error(FEMGPU_EINVAL, "null argument");
  std::string k(key);
  if (k == "scratch_limit_bytes" && value > 0) c->scratch_limit_bytes = value;
  else if (k == "arena_candidates" && value >= 0) c->arena_candidates = value;
  else if (k == "sort_capacity" && value >= 32 && value <= 8192 && (value & (value - 1)) == 0) c->sort_capacity = (uint32_t)value;
  else if (k == "acc_capacity" && value >= 1 && value <= 8192) c->acc_capacity = (uint32_t)value;
  else if (k == "l2_persist_lut") c->l2_persist_lut = value != 0;
  else if (k == "cand_mode" && value >= 0 && value <= 6) c->cand_mode = (int)value;
  else if (k == "small_shape" && (value == 0 || value == 1 || value == 4)) c->small_shape = (int)value;
  else if (k == "pair_bits" && (value == 0 || (value >= 1024 && value <= 32768 && (value & (value - 1)) == 0))) c->pair_bits = (uint32_t)value;
  else if (k == "pair_values" && (value == 0 || (value >= 32 && value <= 4096 && value % 2 == 0))) c->pair_values = (uint32_t)value;
  else if (k == "hash_slots" && (value == 0 || (value >= 32 && value <= 4096 && (value & (value - 1)) == 0))) c->hash_slots = (uint32_t)value;
  else if (k == "device_traceback") c->device_traceback = value != 0;
  else if (k == "verify_planes") c->verify_planes = value != 0;
  else if (k == "blocking_sync" && (value == 0 || value == 1)) c->blocking_sync = (int)value;
  else if (k == "seed_dp_mode" && value >= 0 && value <= 3) c->seed_dp_mode = (int)value;
  else return set_error(FEMGPU_EINVAL, "unknown option or bad value: %s=%lld", key, (long long)value);
  return 0;
}

extern "C" void *femgpu_alloc_pinned(size_t bytes) {
  void *p = nullptr;
  if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) {
    cudaGetLastError();
    set_error(FEMGPU_ENOMEM, "cudaMallocHost of %zu bytes failed", bytes);
    return nullptr;
  }
  return p;
}
extern "C" void femgpu_free_pinned(void *p) { if (p) cudaFreeHost(p); }

extern "C" int femgpu_index_k(const femgpu_ctx *c) { return c ? c->idx_k : 0; }
extern "C" int femgpu_index_step(const femgpu_ctx *c) { return c ? c->idx_step : 0; }

// INT32 pipe peak: instructions per second of the three kinds, 148 SMs x 8 CTAs x 256 threads
extern "C" int femgpu_measure_int32_peak(femgpu_ctx *c, double giops[3]) {
  if (!c || !giops) return set_error(FEMGPU_EINVAL, "null argument");
  CU(cudaSetDevice(c->device));
  uint32_t *d = nullptr;
  CU(cudaMalloc(&d, 64));
  cudaStream_t st = c->slot[0].stream;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 4096, blocks = c->sm_count * 8;
  const double instr = (double)blocks * 256 * iters * 16 * 8;
  for (int kind = 0; kind < 3; ++kind) {
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
      cudaEventRecord(e0, st);
      if (kind == 0) k_int_peak<0><<<blocks, 256, 0, st>>>(d, iters, 12345u + rep);
      else if (kind == 1) k_int_peak<1><<<blocks, 256, 0, st>>>(d, iters, 12345u + rep);
      else k_int_peak<2><<<blocks, 256, 0, st>>>(d, iters, 12345u + rep);
      cudaEventRecord(e1, st);
      cudaEventSynchronize(e1);
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      if (rep && ms < best) best = ms;
    }
    giops[kind] = instr * (kind == 2 ? 2.0 : 1.0) / (best * 1e-3) / 1e9;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(d);
  CU(cudaGetLastError());
  return 0;
}

static int derive_ref_planes(femgpu_ctx *c);
#include "upload.inl"
#include "pipeline.inl"
#include "index_build.inl"
#include "noalloc.inl"
#include "api.inl"
