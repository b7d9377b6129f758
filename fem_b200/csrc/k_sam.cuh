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
// (FEMGPU_ALN_HOST) are left to the host here as well.  Two passes of the same emitter: lengths
// (k_sam_len, then an exclusive scan over reads), text (k_sam_write).
#pragma once
#include "common.cuh"
#include "k_trace.cuh"

#define SAM_BIG_MAX 2048u    // most mappings of a read the device still formats (k_sam_big_*)
#define SAM_IMAGE_MAX 24576u // most bytes of shared memory per warp in k_sam_write (four warps per block)

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
  // reads with 65..SAM_BIG_MAX mappings: a warp per read (k_sam_big_len / k_sam_big_write)
  uint32_t *big_list;         // read numbers, in no particular order
  unsigned long long *n_big;
  uint16_t *big_ord;          // per mapping slot: which of the read's mappings is printed r-th
  uint32_t *big_rec_off;      // per mapping slot: offset of the r-th record within the read's text
  const uint64_t *text_off;   // in (text pass): exclusive scan of read_bytes
  char *text;                 // out (text pass)
};

__device__ __forceinline__ uint64_t sam_key(const ulonglong2 &m) { // src/align.c:53-55
  const uint64_t dist = (m.y >> 40) & 0xFFu, dir = (m.y >> 32) & 0xFFu;
  return (dist << 60) | (dir << 59) | (m.x + (uint64_t)(int64_t)(int32_t)(uint32_t)m.y);
}

// MODE 0: lengths only; 1: every byte to p; 2: every byte but the primary record's SEQ and QUAL, whose
// room is left free and whose position is returned (the warp copies them, k_sam_write)
template <int MODE>
struct SamOut {
  char *p;
  uint32_t n;
  __device__ __forceinline__ void ch(char c) { if (MODE) p[n] = c; ++n; }
  __device__ __forceinline__ void mem(const char *s, uint32_t len) {
    if (MODE) for (uint32_t i = 0; i < len; ++i) p[n + i] = s[i];
    n += len;
  }
  __device__ __forceinline__ void lit(const char *s) { while (*s) ch(*s++); }
  __device__ __forceinline__ void num(uint32_t v) { // every number of a record fits 32 bits (positions are < 2^31)
    uint32_t digits = 1;
    for (uint32_t t = v; t >= 10u; t /= 10u) ++digits;
    if (MODE) for (uint32_t k = digits, t = v; k--; t /= 10u) p[n + k] = (char)('0' + t % 10u);
    n += digits;
  }
};

// one record: mapping m of read i (bases at off, l long), primary (SEQ/QUAL present) or not
template <int MODE>
__device__ __forceinline__ void sam_emit_record(const SamArgs &A, SamOut<MODE> &o, uint32_t i, uint64_t m, bool primary,
                                                const char *name, uint32_t name_l, uint64_t off, uint32_t l,
                                                uint32_t *seq_at, bool *with_qual) {
  const ulonglong2 rec = A.maps[m];
  const uint32_t *al = A.aln + m * ALN_WORDS;
  const uint32_t w0 = al[0];
  const int start = (int)(int16_t)(uint16_t)w0;
  const uint32_t n_cigar = (w0 >> 16) & 0xFFu, md_len = w0 >> 24;
  const uint32_t dir = (uint32_t)(rec.y >> 32) & 0xFFu, dist = (uint32_t)(rec.y >> 40) & 0xFFu;
  const uint32_t r = (uint32_t)(rec.x >> 32), pos = (uint32_t)start + (uint32_t)rec.x; // src/align.c:80
  o.mem(name, name_l);
  o.ch('\t');
  o.num((dir ? 16u : 0u) | (primary ? 0u : 256u));
  o.ch('\t');
  o.mem(A.rnames + A.rname_off[r], A.rname_off[r + 1] - A.rname_off[r]);
  o.ch('\t');
  o.num(pos + 1u);
  o.lit("\t255\t");
  if (n_cigar == 0) o.ch('*');
  for (uint32_t ci = 0; ci < n_cigar; ++ci) {
    const uint32_t c = (al[1 + (ci >> 1)] >> (16 * (ci & 1))) & 0xFFFFu;
    o.num(c >> 2);
    o.ch("MID"[c & 3u]);
  }
  o.lit("\t*\t0\t0\t");
  if (primary) { // the primary record carries SEQ/QUAL in the file's orientation (src/align.c:79,87)
    const bool q = A.has_qual[i] && A.quals[off] != ' '; // htslib sam.c:3047 prints '*' otherwise
    if (MODE == 2) {
      *seq_at = o.n;
      *with_qual = q;
      o.n += l;
      o.ch('\t');
      if (q) o.n += l;
      else o.ch('*');
    } else {
      o.mem(A.bases + off, l); // canonical reads only (others are left to the host): printed as they are
      o.ch('\t');
      if (q) o.mem(A.quals + off, l);
      else o.ch('*');
    }
  } else {
    o.lit("*\t*");
  }
  o.lit("\tNM:i:");
  o.num(dist);
  o.lit("\tMD:Z:");
  if (MODE) for (uint32_t k = 0; k < md_len; ++k) o.p[o.n + k] = (char)((al[10 + (k >> 2)] >> (8 * (k & 3))) & 0xFFu);
  o.n += md_len;
  o.ch('\n');
}

// the records of read i (1 <= n <= 64 mappings, not left to the host); returns the bytes.  seq_at: MODE 2,
// offset of SEQ within the text; with_qual: MODE 2, QUAL follows SEQ + tab (else '*' was written)
template <int MODE>
__device__ __forceinline__ uint32_t sam_emit_read(const SamArgs &A, uint32_t i, uint64_t m0, uint32_t n, char *p,
                                                  uint32_t *seq_at, bool *with_qual) {
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
  const uint64_t name_at = A.name_off[i];
  const char *name = A.names + name_at;
  const uint32_t name_l = (uint32_t)(A.name_off[i + 1] - name_at) - 1u; // names lie back to back, one NUL after each
  SamOut<MODE> o;
  o.p = p;
  o.n = 0;
  for (uint32_t mi = 0; mi < n; ++mi) sam_emit_record<MODE>(A, o, i, m0 + ord[mi], mi == 0, name, name_l, off, l, seq_at, with_qual);
  return o.n;
}

// pass 1: bytes of every read's text, host flags
__global__ void __launch_bounds__(128) k_sam_len(SamArgs A) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.n_reads) return;
  const uint64_t m0 = A.rs_map_first[2ull * i], m1 = A.rs_map_first[2ull * i + 2];
  const uint32_t n = (uint32_t)(m1 - m0);
  A.read_bytes[i] = 0;
  A.host_flag[i] = 0;
  if (n == 0) return;
  if (n > 64 && n <= SAM_BIG_MAX) { // a warp sorts and measures this one
    A.host_flag[i] = 2;
    A.big_list[atomicAdd(A.n_big, 1ull)] = i;
    return;
  }
  bool host = n > 64;
  for (uint32_t j = 0; j < n && !host; ++j) {
    const uint32_t w0 = A.aln[(m0 + j) * ALN_WORDS];
    host = ((w0 >> 16) & 0xFFu) == ALN_HOST || (w0 >> 24) == ALN_HOST;
  }
  if (host) { A.host_flag[i] = 1; atomicAdd(A.n_flagged, 1ull); return; }
  A.read_bytes[i] = sam_emit_read<0>(A, i, m0, n, nullptr, nullptr, nullptr);
}

// len bytes of src[at..] into shared memory by one lane: aligned words, shifted into place (the buffers
// are padded past their last byte, slot_reserve).  Every lane of the warp copies its own read's field, so
// all lanes work (a warp-wide copy of one 100-byte field keeps 25 of them busy for one round).
__device__ __forceinline__ void sam_copy_own(char *dst, const char *src, uint64_t at, uint32_t len) {
  const uint32_t *w = reinterpret_cast<const uint32_t *>(src + (at & ~3ull));
  const uint32_t sh = (uint32_t)(at & 3u) * 8u;
  uint32_t prev = w[0];
  uint32_t k = 0;
  for (; k + 4 <= len; k += 4) {
    const uint32_t next = w[(k >> 2) + 1];
    const uint32_t v = __funnelshift_r(prev, next, sh);
    prev = next;
    dst[k] = (char)v; dst[k + 1] = (char)(v >> 8); dst[k + 2] = (char)(v >> 16); dst[k + 3] = (char)(v >> 24);
  }
  if (k < len) {
    const uint32_t v = __funnelshift_r(prev, w[(k >> 2) + 1], sh);
    for (uint32_t j = 0; k + j < len; ++j) dst[k + j] = (char)(v >> (8 * j));
  }
}

// pass 2: the text.  The 32 reads of a warp own one contiguous span of the output (text_off is the scan
// of their sizes): each lane builds its read's records in the warp's shared-memory image of that span,
// SEQ and QUAL included (word-wise copies), and the image goes out as
// 16-byte stores (the image starts at the span's offset within its 16-byte line, so both sides are
// aligned alike).  Spans larger than `image` bytes are written byte-wise straight to global memory.
__global__ void __launch_bounds__(128) k_sam_write(SamArgs A, uint32_t image) {
  extern __shared__ __align__(16) char sam_smem[];
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x, i0 = i - lane;
  if (i0 >= A.n_reads) return;
  const uint32_t i1 = min(i0 + 32u, A.n_reads);
  const uint64_t t0 = A.text_off[i0], t1 = A.text_off[i1];
  if (t1 == t0) return;
  const bool live = i < A.n_reads;
  uint64_t m0 = 0;
  uint32_t n = 0;
  if (live && !A.host_flag[i]) {
    m0 = A.rs_map_first[2ull * i];
    n = (uint32_t)(A.rs_map_first[2ull * i + 2] - m0);
  }
  const uint32_t skew = (uint32_t)(t0 & 15u);
  const uint64_t total = (t1 - t0) + skew;
  // a read of k_sam_big_write inside the span: its bytes are not in the image, so no image for this warp
  const bool big_here = __any_sync(0xffffffffu, live && n > 64);
  if (n > 64) n = 0;
  if (total > image || big_here) {
    if (n) sam_emit_read<1>(A, i, m0, n, A.text + A.text_off[i], nullptr, nullptr);
    return;
  }
  char *img = sam_smem + (size_t)warp * image;
  uint32_t seq_at = 0xFFFFFFFFu, l = 0;
  uint64_t roff = 0;
  bool with_qual = false;
  if (n) {
    const uint32_t mine = (uint32_t)(A.text_off[i] - t0) + skew;
    sam_emit_read<2>(A, i, m0, n, img + mine, &seq_at, &with_qual);
    seq_at += mine;
    roff = A.read_off[i];
    l = (uint32_t)(A.read_off[i + 1] - roff);
  }
  __syncwarp();
  if (seq_at != 0xFFFFFFFFu) { // SEQ and QUAL of this lane's read
    sam_copy_own(img + seq_at, A.bases, roff, l);
    if (with_qual) sam_copy_own(img + seq_at + l + 1, A.quals, roff, l);
  }
  __syncwarp();
  char *dst = A.text + (t0 - skew); // 16-byte aligned (the text buffer is)
  const uint32_t tot = (uint32_t)total;
  for (uint32_t c = lane * 16u; c < tot; c += 512u) {
    if (c >= skew && c + 16u <= tot) *reinterpret_cast<uint4 *>(dst + c) = *reinterpret_cast<const uint4 *>(img + c);
    else
      for (uint32_t k = max(c, skew); k < min(c + 16u, tot); ++k) dst[k] = img[k];
  }
}

// ---- reads with 65..SAM_BIG_MAX mappings ---------------------------------------------------------------
// klib's radix sort is not stable beyond 64 records (src/ksort.h:113-151), but with pairwise different
// keys every sort gives the same order: a warp sorts (key, index) by a bitonic network in shared memory,
// hands reads with two equal keys (or an alignment left to the host) to the host, and otherwise measures
// the records in their order.  One lane per record then writes it (k_sam_big_write).
__global__ void __launch_bounds__(128) k_sam_big_len(SamArgs A) {
  extern __shared__ __align__(16) unsigned char sam_big_smem[];
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, warps = gridDim.x * (blockDim.x >> 5);
  uint64_t *key = reinterpret_cast<uint64_t *>(sam_big_smem) + (size_t)warp * SAM_BIG_MAX;
  uint16_t *idx = reinterpret_cast<uint16_t *>(sam_big_smem + (size_t)(blockDim.x >> 5) * SAM_BIG_MAX * 8) + (size_t)warp * SAM_BIG_MAX;
  const uint64_t n_big = *A.n_big;
  for (uint64_t b = (uint64_t)blockIdx.x * (blockDim.x >> 5) + warp; b < n_big; b += warps) {
    const uint32_t i = A.big_list[b];
    const uint64_t m0 = A.rs_map_first[2ull * i];
    const uint32_t n = (uint32_t)(A.rs_map_first[2ull * i + 2] - m0);
    uint32_t P = 128;
    while (P < n) P <<= 1;
    bool host = false;
    for (uint32_t j = lane; j < P; j += 32) {
      uint64_t kj = ~0ull;
      if (j < n) {
        kj = sam_key(A.maps[m0 + j]);
        const uint32_t w0 = A.aln[(m0 + j) * ALN_WORDS];
        host |= ((w0 >> 16) & 0xFFu) == ALN_HOST || (w0 >> 24) == ALN_HOST;
      }
      key[j] = kj;
      idx[j] = (uint16_t)j;
    }
    host = __any_sync(0xffffffffu, host);
    __syncwarp();
    if (!host) {
      for (uint32_t k = 2; k <= P; k <<= 1)
        for (uint32_t d = k >> 1; d > 0; d >>= 1) {
          for (uint32_t t = lane; t < (P >> 1); t += 32) {
            const uint32_t lo = ((t & ~(d - 1)) << 1) | (t & (d - 1)), hi = lo | d; // the pair t of this stage
            const bool up = (lo & k) == 0;
            const uint64_t a = key[lo], c = key[hi];
            if ((a > c) == up) {
              key[lo] = c; key[hi] = a;
              const uint16_t x = idx[lo]; idx[lo] = idx[hi]; idx[hi] = x;
            }
          }
          __syncwarp();
        }
      for (uint32_t j = lane; j + 1 < n; j += 32) host |= key[j] == key[j + 1];
      host = __any_sync(0xffffffffu, host);
    }
    if (host) {
      if (lane == 0) { A.host_flag[i] = 1; A.read_bytes[i] = 0; atomicAdd(A.n_flagged, 1ull); }
      __syncwarp();
      continue;
    }
    const uint64_t off = A.read_off[i];
    const uint32_t l = (uint32_t)(A.read_off[i + 1] - off);
    const uint64_t name_at = A.name_off[i];
    const char *name = A.names + name_at;
    const uint32_t name_l = (uint32_t)(A.name_off[i + 1] - name_at) - 1u; // names lie back to back, one NUL after each
    uint32_t running = 0;
    for (uint32_t r0 = 0; r0 < n; r0 += 32) {
      const uint32_t r = r0 + lane;
      uint32_t len = 0;
      if (r < n) {
        SamOut<0> o;
        o.p = nullptr;
        o.n = 0;
        sam_emit_record<0>(A, o, i, m0 + idx[r], r == 0, name, name_l, off, l, nullptr, nullptr);
        len = o.n;
      }
      uint32_t incl = len;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o2 = __shfl_up_sync(0xffffffffu, incl, d);
        if ((int)lane >= d) incl += o2;
      }
      if (r < n) {
        A.big_ord[m0 + r] = idx[r];
        A.big_rec_off[m0 + r] = running + incl - len;
      }
      running += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 0) { A.read_bytes[i] = running; A.host_flag[i] = 0; }
    __syncwarp();
  }
}

#define SAM_BIG_IMAGE 8192u // bytes of shared memory per warp in k_sam_big_write

// 32 consecutive records of a read own one contiguous span of its text: built in a shared-memory image
// by one lane per record, written out as 16-byte stores (as in k_sam_write); longer spans byte-wise.
__global__ void __launch_bounds__(128) k_sam_big_write(SamArgs A) {
  __shared__ __align__(16) char sam_big_img[4 * SAM_BIG_IMAGE];
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, warps = gridDim.x * (blockDim.x >> 5);
  char *img = sam_big_img + (size_t)warp * SAM_BIG_IMAGE;
  const uint64_t n_big = *A.n_big;
  for (uint64_t b = (uint64_t)blockIdx.x * (blockDim.x >> 5) + warp; b < n_big; b += warps) {
    const uint32_t i = A.big_list[b];
    if (A.host_flag[i]) continue;
    const uint64_t m0 = A.rs_map_first[2ull * i];
    const uint32_t n = (uint32_t)(A.rs_map_first[2ull * i + 2] - m0);
    const uint64_t off = A.read_off[i];
    const uint32_t l = (uint32_t)(A.read_off[i + 1] - off);
    const uint64_t name_at = A.name_off[i];
    const char *name = A.names + name_at;
    const uint32_t name_l = (uint32_t)(A.name_off[i + 1] - name_at) - 1u; // names lie back to back, one NUL after each
    const uint64_t t_read = A.text_off[i];
    const uint32_t read_len = (uint32_t)(A.text_off[i + 1] - t_read);
    for (uint32_t r0 = 0; r0 < n; r0 += 32) {
      const uint32_t r = r0 + lane;
      const uint32_t mine = r < n ? A.big_rec_off[m0 + r] : 0u;
      const uint32_t s0 = __shfl_sync(0xffffffffu, mine, 0);
      const uint32_t s1 = r0 + 32 < n ? A.big_rec_off[m0 + r0 + 32] : read_len;
      const uint64_t t0 = t_read + s0;
      const uint32_t skew = (uint32_t)(t0 & 15u), tot = (s1 - s0) + skew;
      const bool staged = tot <= SAM_BIG_IMAGE;
      if (r < n) {
        SamOut<1> o;
        o.p = staged ? img + skew + (mine - s0) : A.text + t_read + mine;
        o.n = 0;
        sam_emit_record<1>(A, o, i, m0 + A.big_ord[m0 + r], r == 0, name, name_l, off, l, nullptr, nullptr);
      }
      __syncwarp();
      if (staged) {
        char *dst = A.text + (t0 - skew);
        for (uint32_t c = lane * 16u; c < tot; c += 512u) {
          if (c >= skew && c + 16u <= tot) *reinterpret_cast<uint4 *>(dst + c) = *reinterpret_cast<const uint4 *>(img + c);
          else
            for (uint32_t k = max(c, skew); k < min(c + 16u, tot); ++k) dst[k] = img[k];
        }
        __syncwarp();
      }
    }
  }
}
