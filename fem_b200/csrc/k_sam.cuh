// k_sam.cuh -- K7: the SAM records of a batch as text, on the device.
//
// One thread per read restates what the reference does per mapped read on the host:
//   * process_mappings (src/align.c:56-92): mapping order = radix_sort_mapping on
//     (distance << 60) | (direction << 59) | (candidate + end offset); for <= 64 records klib's sort is
//     an insertion sort (src/ksort.h:98-111), i.e. a stable sort by that key, reproduced here by rank
//     counting.  Reads with more records (unstable in-place radix passes) are left to the host;
//   * generate_bam1_t + htslib's sam_format1 (src/align.c:546-632, extern/htslib/sam.c:3004-3061):
//     QNAME FLAG RNAME POS 255 CIGAR * 0 0 SEQ QUAL NM:i MD:Z, the first record primary with SEQ/QUAL in
//     the file's orientation, the others flag | 256 with "*".
// Start, CIGAR and MD come from k_trace; reads with an alignment k_trace left to the host
// (FEMGPU_ALN_HOST) are left to the host here as well.  Two passes of the same emitter: lengths (then
// an exclusive scan over reads), text.
#pragma once
#include "common.cuh"
#include "k_trace.cuh"

struct SamArgs {
  const char *bases;          // the batch's reads as uploaded (file orientation)
  const uint64_t *read_off;
  uint32_t n_reads;
  const char *quals;          // same offsets as bases
  const uint8_t *has_qual;
  const char *names;
  const uint64_t *name_off;
  const char *rnames;         // reference record names back to back
  const uint32_t *rname_off;  // n_rec + 1
  const ulonglong2 *maps;     // compacted mappings (k_compact), emission order
  const uint32_t *aln;        // ALN_WORDS per mapping (k_trace)
  const uint64_t *rs_map_first; // per read-strand; read i owns [rs_map_first[2i], rs_map_first[2i+2])
  uint32_t *read_bytes;       // out (length pass): bytes of text of each read
  uint8_t *host_flag;         // out (length pass): 1 = left to the host
  unsigned long long *n_flagged; // out (length pass): how many
  const uint64_t *text_off;   // in (text pass): exclusive scan of read_bytes
  char *text;                 // out (text pass)
};

__device__ __forceinline__ uint64_t sam_key(const ulonglong2 &m) { // src/align.c:53-55
  const uint64_t dist = (m.y >> 40) & 0xFFu, dir = (m.y >> 32) & 0xFFu;
  return (dist << 60) | (dir << 59) | (m.x + (uint64_t)(int64_t)(int32_t)(uint32_t)m.y);
}

template <bool WRITE>
struct SamOut {
  char *p;
  uint32_t n;
  __device__ __forceinline__ void ch(char c) { if (WRITE) p[n] = c; ++n; }
  __device__ __forceinline__ void mem(const char *s, uint32_t len) {
    if (WRITE) for (uint32_t i = 0; i < len; ++i) p[n + i] = s[i];
    n += len;
  }
  __device__ __forceinline__ void lit(const char *s) { while (*s) ch(*s++); }
  __device__ __forceinline__ void num(uint64_t v) {
    char tmp[20];
    int k = 0;
    do { tmp[k++] = (char)('0' + (int)(v % 10u)); v /= 10u; } while (v);
    while (k) ch(tmp[--k]);
  }
};

template <bool WRITE>
__global__ void __launch_bounds__(128) k_sam(SamArgs A) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.n_reads) return;
  const uint64_t m0 = A.rs_map_first[2ull * i], m1 = A.rs_map_first[2ull * i + 2];
  const uint32_t n = (uint32_t)(m1 - m0);
  if (!WRITE) { A.read_bytes[i] = 0; A.host_flag[i] = 0; }
  if (n == 0) return;
  if (WRITE && A.host_flag[i]) return;
  if (!WRITE) {
    bool host = n > 64;
    for (uint32_t j = 0; j < n && !host; ++j) {
      const uint32_t w0 = A.aln[(m0 + j) * ALN_WORDS];
      host = ((w0 >> 16) & 0xFFu) == ALN_HOST || (w0 >> 24) == ALN_HOST;
    }
    if (host) { A.host_flag[i] = 1; atomicAdd(A.n_flagged, 1ull); return; }
  }
  // mapping order: stable by key (n <= 64)
  uint8_t ord[64];
  if (n == 1) ord[0] = 0;
  else
    for (uint32_t a = 0; a < n; ++a) {
      const uint64_t ka = sam_key(A.maps[m0 + a]);
      uint32_t rank = 0;
      for (uint32_t b = 0; b < n; ++b) {
        const uint64_t kb = sam_key(A.maps[m0 + b]);
        rank += (kb < ka || (kb == ka && b < a)) ? 1u : 0u;
      }
      ord[rank] = (uint8_t)a;
    }
  const uint64_t off = A.read_off[i];
  const uint32_t l = (uint32_t)(A.read_off[i + 1] - off);
  const char *name = A.names + A.name_off[i];
  uint32_t name_l = 0;
  while (name[name_l]) ++name_l;
  SamOut<WRITE> o;
  o.p = WRITE ? A.text + A.text_off[i] : nullptr;
  o.n = 0;
  for (uint32_t mi = 0; mi < n; ++mi) {
    const uint64_t m = m0 + ord[mi];
    const ulonglong2 rec = A.maps[m];
    const uint32_t *al = A.aln + m * ALN_WORDS;
    const uint32_t w0 = al[0];
    const int start = (int)(int16_t)(uint16_t)w0;
    const uint32_t n_cigar = (w0 >> 16) & 0xFFu, md_len = w0 >> 24;
    const uint32_t dir = (uint32_t)(rec.y >> 32) & 0xFFu, dist = (uint32_t)(rec.y >> 40) & 0xFFu;
    const uint32_t r = (uint32_t)(rec.x >> 32), pos = (uint32_t)start + (uint32_t)rec.x; // src/align.c:80
    o.mem(name, name_l);
    o.ch('\t');
    o.num((dir ? 16u : 0u) | (mi ? 256u : 0u));
    o.ch('\t');
    o.mem(A.rnames + A.rname_off[r], A.rname_off[r + 1] - A.rname_off[r]);
    o.ch('\t');
    o.num((uint64_t)pos + 1);
    o.lit("\t255\t");
    if (n_cigar == 0) o.ch('*');
    for (uint32_t ci = 0; ci < n_cigar; ++ci) {
      const uint32_t c = (al[1 + (ci >> 1)] >> (16 * (ci & 1))) & 0xFFFFu;
      o.num(c >> 2);
      o.ch("MID"[c & 3u]);
    }
    o.lit("\t*\t0\t0\t");
    if (mi == 0) { // the primary record carries SEQ/QUAL in the file's orientation (src/align.c:79,87)
      o.mem(A.bases + off, l); // canonical reads only (others are left to the host): printed as they are
      o.ch('\t');
      if (!A.has_qual[i] || A.quals[off] == ' ') o.ch('*'); // htslib sam.c:3047
      else o.mem(A.quals + off, l);
    } else {
      o.lit("*\t*");
    }
    o.lit("\tNM:i:");
    o.num(dist);
    o.lit("\tMD:Z:");
    if (WRITE) for (uint32_t k = 0; k < md_len; ++k) o.p[o.n + k] = (char)((al[10 + (k >> 2)] >> (8 * (k & 3))) & 0xFFu);
    o.n += md_len;
    o.ch('\n');
  }
  if (!WRITE) A.read_bytes[i] = o.n;
}
