// k_peak.cuh -- INT32 pipe peak microbenchmark (the roofline denominator of k_verify; the driver's
// MEASURED_PEAKS.json only holds HBM and bf16 numbers).  Each thread runs 8 independent
// dependency chains of one instruction kind so that issue rate, not latency, is measured.
#pragma once
#include "common.cuh"

template <int KIND> // 0: LOP3 (a ^ b & c), 1: IADD3, 2: SHF/shift + LOP3 mix as in the Myers column
__global__ void __launch_bounds__(256) k_int_peak(uint32_t *out, int iters, uint32_t seed) {
  uint32_t a0 = seed + threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, a4 = a0 * 11, a5 = a0 * 13, a6 = a0 * 17, a7 = a0 * 19;
  const uint32_t b = seed * 0x9E3779B9u + blockIdx.x, c = ~seed;
#pragma unroll 1
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      if (KIND == 0) {
        a0 = (a0 ^ b) & (a0 | c); a1 = (a1 ^ b) & (a1 | c); a2 = (a2 ^ b) & (a2 | c); a3 = (a3 ^ b) & (a3 | c);
        a4 = (a4 ^ b) & (a4 | c); a5 = (a5 ^ b) & (a5 | c); a6 = (a6 ^ b) & (a6 | c); a7 = (a7 ^ b) & (a7 | c);
      } else if (KIND == 1) {
        a0 = a0 + b + c; a1 = a1 + b + c; a2 = a2 + b + c; a3 = a3 + b + c;
        a4 = a4 + b + c; a5 = a5 + b + c; a6 = a6 + b + c; a7 = a7 + b + c;
      } else {
        a0 = (a0 >> 1) ^ b; a1 = (a1 >> 1) ^ b; a2 = (a2 >> 1) ^ b; a3 = (a3 >> 1) ^ b;
        a4 = (a4 >> 1) ^ b; a5 = (a5 >> 1) ^ b; a6 = (a6 >> 1) ^ b; a7 = (a7 >> 1) ^ b;
      }
    }
  }
  const uint32_t r = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7;
  if (r == 0x12345678u) out[0] = r; // keeps the chains alive
}
