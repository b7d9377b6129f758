/* index.c -- the succinct hash index on the host: build (`FEM index`), save, load.
 *
 * The arrays are the reference's (src/index.h:7-14): lut[h] = number of sampled q-grams with hash
 * < h, occ = their (record << 32 | position) words grouped by hash, ascending inside a group
 * (src/index.c:57-98).  The reference collects (hash, location) pairs and radix-sorts them twice;
 * here the same arrays come from one counting sort -- count, prefix-sum, scatter in ascending
 * (record, position) order, which leaves every group ascending with no second sort.
 * File layout (src/index.c:133-168), frozen: int k; int step; uint32 lut[4^k+1]; size_t n;
 * uint64 occ[n]; native endianness.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "femhost.h"
#include "host_internal.h"

const uint8_t femhost_base_code[256] = {
    ['A'] = 1, ['a'] = 1, ['C'] = 2, ['c'] = 2, ['G'] = 3, ['g'] = 3, ['T'] = 4, ['t'] = 4};
#define CODE(c) FEM_CODE(c)

/* calls fn(hash, record, position) for every sampled q-gram of every record
 * (positions 0, step, 2*step, ... while the q-gram lies inside the record; src/index.c:65) */
#define FOR_EACH_SAMPLE(ref, k, step, BODY)                                                        \
  for (uint32_t rec_ = 0; rec_ < (ref)->n_records; ++rec_) {                                       \
    const char *s_ = (ref)->bases + (ref)->base_off[rec_];                                         \
    const uint32_t len_ = femhost_record_length(ref, rec_);                                        \
    uint32_t h_ = 0;                                                                               \
    for (uint32_t i_ = 0; i_ < len_; ++i_) { /* rolling 2-bit hash, ambiguous base -> A */         \
      const uint8_t c_ = CODE(s_[i_]);                                                             \
      h_ = ((h_ << 2) | (c_ < 4 ? c_ : 0u)) & mask;                                                \
      if (i_ + 1 >= (uint32_t)(k) && (i_ + 1 - (uint32_t)(k)) % (uint32_t)(step) == 0) {           \
        const uint32_t pos_ = i_ + 1 - (uint32_t)(k);                                              \
        const uint32_t hash = h_;                                                                  \
        const uint64_t loc = ((uint64_t)rec_ << 32) | pos_;                                        \
        BODY                                                                                       \
      }                                                                                            \
    }                                                                                              \
  }

int femhost_index_build(const femhost_records *ref, int k, int step, femhost_index *out) {
  memset(out, 0, sizeof *out);
  if (k < 1 || k > 15 || step < 1) return femhost_set_error(-1, "bad index parameters k=%d step=%d", k, step);
  const uint32_t mask = (uint32_t)((1ull << (2 * k)) - 1);
  const size_t lut_n = ((size_t)1 << (2 * k)) + 1;
  uint32_t *lut = (uint32_t *)calloc(lut_n, sizeof(uint32_t));
  if (!lut) return femhost_set_error(-3, "out of memory (lookup table)");
  size_t total = 0;
  FOR_EACH_SAMPLE(ref, k, step, { (void)loc; ++lut[hash + 1]; ++total; })
  if (total > 0xFFFFFFFFull) { free(lut); return femhost_set_error(-1, "more than 2^32-1 sampled q-grams"); }
  for (size_t i = 1; i < lut_n; ++i) lut[i] += lut[i - 1];
  uint64_t *occ = (uint64_t *)malloc((total ? total : 1) * sizeof(uint64_t));
  uint32_t *fill = (uint32_t *)malloc((lut_n - 1) * sizeof(uint32_t));
  if (!occ || !fill) { free(lut); free(occ); free(fill); return femhost_set_error(-3, "out of memory (occurrence table)"); }
  memcpy(fill, lut, (lut_n - 1) * sizeof(uint32_t));
  FOR_EACH_SAMPLE(ref, k, step, { occ[fill[hash]++] = loc; })
  free(fill);
  out->k = k; out->step = step; out->lut = lut; out->lut_n = lut_n; out->occ = occ; out->n_occ = total;
  return 0;
}

int femhost_index_save(const femhost_index *idx, const char *path) {
  FILE *f = fopen(path, "wb");
  if (!f) return femhost_set_error(-1, "Cannot open index file %s for writing", path);
  const int k = idx->k, step = idx->step;
  const size_t n = idx->n_occ;
  int ok = fwrite(&k, sizeof(int), 1, f) == 1 && fwrite(&step, sizeof(int), 1, f) == 1 &&
           fwrite(idx->lut, sizeof(uint32_t), idx->lut_n, f) == idx->lut_n && fwrite(&n, sizeof(size_t), 1, f) == 1 &&
           fwrite(idx->occ, sizeof(uint64_t), n, f) == n;
  ok = fclose(f) == 0 && ok;
  return ok ? 0 : femhost_set_error(-1, "short write to index file %s", path);
}

int femhost_index_load(const char *path, femhost_index *out) {
  memset(out, 0, sizeof *out);
  FILE *f = path ? fopen(path, "rb") : NULL;
  if (!f) return femhost_set_error(-1, "Cannot open index file %s", path ? path : "(null)");
  int k = 0, step = 0;
  if (fread(&k, sizeof(int), 1, f) != 1 || fread(&step, sizeof(int), 1, f) != 1 || k < 1 || k > 15 || step < 1) {
    fclose(f);
    return femhost_set_error(-1, "bad index header in %s", path);
  }
  out->k = k; out->step = step;
  out->lut_n = ((size_t)1 << (2 * k)) + 1;
  out->lut = (uint32_t *)malloc(out->lut_n * sizeof(uint32_t));
  size_t n = 0;
  if (!out->lut || fread(out->lut, sizeof(uint32_t), out->lut_n, f) != out->lut_n || fread(&n, sizeof(size_t), 1, f) != 1) goto bad;
  out->occ = (uint64_t *)malloc((n ? n : 1) * sizeof(uint64_t));
  if (!out->occ || fread(out->occ, sizeof(uint64_t), n, f) != n) goto bad;
  out->n_occ = n;
  fclose(f);
  return 0;
bad:
  fclose(f);
  femhost_index_free(out);
  return femhost_set_error(-1, "index file %s is truncated", path);
}

void femhost_index_free(femhost_index *idx) {
  free(idx->lut);
  free(idx->occ);
  memset(idx, 0, sizeof *idx);
}
