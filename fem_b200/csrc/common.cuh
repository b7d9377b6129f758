// common.cuh -- shared device/host definitions of the FEM map path on B200 (sm_100a).
//
// Data layout in HBM (one copy per GPU, uploaded once):
//   lut    uint32[4^k+1]   prefix sums of bucket sizes, verbatim from the index file
//                          (reference: src/index.h:7-14, src/index.c:133-168)
//   occ    uint64[n]       (seq_index<<32 | pos) words, verbatim (src/index.c:65-67,93)
//   ref2   uint32[]        reference bases, all records back to back, 2 bits per base,
//                          16 bases per word, base g at bits 2*(g&15) of word g>>4
//   refn   uint32[]        ambiguity mask (1 = not ACGT), 32 bases per word; absent when the
//                          reference has no such base
//   rec_base uint64[n_rec+1], rec_len uint32[n_rec]   record starts in the concatenation
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define FEM_SENTINEL 0xFFFFFFFFFFFFFFFFull
#define FEM_MAX_Q 10          // e + 1 + a <= 7 + 1 + 2
#define FEM_MAX_READ_LEN 1024 // longest read the seed kernel's shared-memory plan accepts

// read-strand status written by the seed kernel
#define RS_OK 0
#define RS_EMPTY 1      // the reference returns 0 candidates early (src/filter.c:168-172,180-182)
#define RS_INFEASIBLE 2 // the reference runs into undefined behaviour

struct DevIndex {
  const uint32_t *lut;
  const uint64_t *occ;
  uint64_t n_occ;
  int k, step;
};

struct DevRef {
  const uint32_t *ref2;
  const uint32_t *refn; // nullptr when the reference is pure ACGT
  const uint64_t *rec_base;
  const uint32_t *rec_len;
  uint32_t n_rec;
};

// base code of the reference (src/utils.h:72): A/a 0, C/c 1, G/g 2, T/t 3, everything else 4
__host__ __device__ __forceinline__ uint32_t fem_base_code(unsigned char c) {
  c &= 0xDF; // fold case: 'a'..'z' -> 'A'..'Z' (other bytes may alias, checked below)
  return c == 'A' ? 0u : c == 'C' ? 1u : c == 'G' ? 2u : c == 'T' ? 3u : 4u;
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}
__host__ __device__ __forceinline__ uint32_t pow2_ceil_u32(uint32_t x) {
  uint32_t p = 1;
  while (p < x) p <<= 1;
  return p;
}
