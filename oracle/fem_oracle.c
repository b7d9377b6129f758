/* fem_oracle.c -- CPU restatement of the reference's `FEM map` hot path (plain C, scalar).
 *
 * TEST INFRASTRUCTURE ONLY -- see fem_oracle.h.  Parity: pinned against oracle/_ref/FEM_ref_*
 * (the reference compiled from /root/reference) and the golden SAM under tests/golden/.
 *
 * The code follows the reference's sequential algorithm step by step (so that it is evidently
 * the same algorithm) but is written from scratch around flat arrays; the product (fem_b200/)
 * uses a different, set-based formulation on the GPU and is checked against this file.
 * Citations are file:line relative to /root/reference.
 */
#include "fem_oracle.h"

#include <ctype.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

/* ------------------------------------------------------------------------------------------ */
/* small growable buffers                                                                      */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
  uint64_t *a;
  size_t n, m;
} u64vec;

static void u64_push(u64vec *v, uint64_t x) {
  if (v->n == v->m) {
    v->m = v->m ? v->m * 2 : 64;
    v->a = (uint64_t *)realloc(v->a, v->m * sizeof(uint64_t));
  }
  v->a[v->n++] = x;
}

static void buf_reserve(char **buf, size_t *cap, size_t need) {
  if (need > *cap) {
    size_t m = *cap ? *cap : 256;
    while (m < need) m *= 2;
    *buf = (char *)realloc(*buf, m);
    *cap = m;
  }
}
static void buf_putn(char **buf, size_t *len, size_t *cap, const char *s, size_t n) {
  buf_reserve(buf, cap, *len + n + 1);
  memcpy(*buf + *len, s, n);
  *len += n;
}
static void buf_putc(char **buf, size_t *len, size_t *cap, char c) { buf_putn(buf, len, cap, &c, 1); }
static void buf_putll(char **buf, size_t *len, size_t *cap, long long v) {
  char tmp[32];
  int n = snprintf(tmp, sizeof tmp, "%lld", v);
  buf_putn(buf, len, cap, tmp, (size_t)n);
}

/* ------------------------------------------------------------------------------------------ */
/* context                                                                                     */
/* ------------------------------------------------------------------------------------------ */
struct fo_ctx {
  uint32_t n;
  char **names;
  char **seqs;
  uint32_t *lens;
  int idx_k, idx_step;
  uint32_t *lut;
  size_t n_occ;
  uint64_t *occ;
  int k, step, e, a;
};

/* src/utils.h:72-81 : A/a 0, C/c 1, G/g 2, T/t 3, everything else 4 */
static inline uint8_t base_code(char c) {
  switch (c) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': return 3;
    default: return 4;
  }
}

/* ------------------------------------------------------------------------------------------ */
/* FASTA/FASTQ record reader with klib kseq semantics (src/kseq.h:185-226)                      */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
  unsigned char *d;
  size_t n, p;
  int last_char;
} rec_reader;

typedef struct {
  const char *name; size_t name_l;
  char *seq; size_t seq_l, seq_m;
  char *qual; size_t qual_l, qual_m;
  char *namebuf; size_t name_m;
} rec;

static int rr_open(rec_reader *r, const char *path) {
  gzFile f = gzopen(path, "r");
  if (!f) return -1;
  size_t m = 1 << 20;
  r->d = (unsigned char *)malloc(m);
  r->n = 0;
  for (;;) {
    if (r->n == m) { m *= 2; r->d = (unsigned char *)realloc(r->d, m); }
    int got = gzread(f, r->d + r->n, (unsigned)((m - r->n) > (1u << 30) ? (1u << 30) : (m - r->n)));
    if (got <= 0) break;
    r->n += (size_t)got;
  }
  gzclose(f);
  r->p = 0;
  r->last_char = 0;
  return 0;
}
static inline int rr_getc(rec_reader *r) { return r->p < r->n ? r->d[r->p++] : -1; }

/* append the rest of the current line (without '\n', trailing '\r' dropped when the string is
 * longer than one char -- src/kseq.h:146) ; returns -1 at EOF with nothing read */
static int rr_append_line(rec_reader *r, char **s, size_t *l, size_t *m) {
  if (r->p >= r->n) return -1;
  size_t b = r->p;
  while (r->p < r->n && r->d[r->p] != '\n') ++r->p;
  size_t len = r->p - b;
  if (*l + len + 2 > *m) { *m = (*l + len + 2) * 2; *s = (char *)realloc(*s, *m); }
  memcpy(*s + *l, r->d + b, len);
  *l += len;
  if (r->p < r->n) ++r->p; /* eat '\n' */
  if (*l > 1 && (*s)[*l - 1] == '\r') --*l;
  (*s)[*l] = 0;
  return 0;
}

/* returns seq length, -1 EOF, -2 truncated quality */
static long rr_next(rec_reader *r, rec *o) {
  int c;
  if (r->last_char == 0) {
    while ((c = rr_getc(r)) >= 0 && c != '>' && c != '@') {}
    if (c < 0) return -1;
    r->last_char = c;
  }
  o->seq_l = o->qual_l = 0;
  /* name: up to the first whitespace */
  if (r->p >= r->n) return -1;
  size_t b = r->p;
  while (r->p < r->n && !isspace(r->d[r->p])) ++r->p;
  o->name_l = r->p - b;
  if (o->name_l + 1 > o->name_m) { o->name_m = o->name_l + 64; o->namebuf = (char *)realloc(o->namebuf, o->name_m); }
  memcpy(o->namebuf, r->d + b, o->name_l);
  o->namebuf[o->name_l] = 0;
  o->name = o->namebuf;
  c = r->p < r->n ? r->d[r->p++] : -1;
  if (c != '\n' && c >= 0) { /* skip the comment */
    while (r->p < r->n && r->d[r->p] != '\n') ++r->p;
    if (r->p < r->n) ++r->p;
  }
  if (!o->seq) { o->seq_m = 256; o->seq = (char *)malloc(o->seq_m); }
  o->seq[0] = 0;
  while ((c = rr_getc(r)) >= 0 && c != '>' && c != '+' && c != '@') {
    if (c == '\n') continue;
    if (o->seq_l + 2 > o->seq_m) { o->seq_m *= 2; o->seq = (char *)realloc(o->seq, o->seq_m); }
    o->seq[o->seq_l++] = (char)c;
    o->seq[o->seq_l] = 0;
    rr_append_line(r, &o->seq, &o->seq_l, &o->seq_m);
  }
  if (c == '>' || c == '@') r->last_char = c;
  if (c != '+') { if (c < 0) r->last_char = 0; return (long)o->seq_l; }
  while ((c = rr_getc(r)) >= 0 && c != '\n') {}
  if (c == -1) return -2;
  if (!o->qual) { o->qual_m = 256; o->qual = (char *)malloc(o->qual_m); }
  while (rr_append_line(r, &o->qual, &o->qual_l, &o->qual_m) >= 0 && o->qual_l < o->seq_l) {}
  r->last_char = 0;
  if (o->seq_l != o->qual_l) return -2;
  return (long)o->seq_l;
}
static void rec_free(rec *o) { free(o->seq); free(o->qual); free(o->namebuf); }

/* ------------------------------------------------------------------------------------------ */
/* reference records                                                                           */
/* ------------------------------------------------------------------------------------------ */
fo_ctx *fo_create_reference(uint32_t n, const char *const *names, const char *const *seqs,
                            const uint32_t *lens) {
  fo_ctx *c = (fo_ctx *)calloc(1, sizeof *c);
  c->n = n;
  c->names = (char **)calloc(n ? n : 1, sizeof(char *));
  c->seqs = (char **)calloc(n ? n : 1, sizeof(char *));
  c->lens = (uint32_t *)calloc(n ? n : 1, sizeof(uint32_t));
  for (uint32_t i = 0; i < n; ++i) {
    c->names[i] = strdup(names[i]);
    c->seqs[i] = (char *)malloc((size_t)lens[i] + 1);
    memcpy(c->seqs[i], seqs[i], lens[i]);
    c->seqs[i][lens[i]] = 0;
    c->lens[i] = lens[i];
  }
  c->k = 12; c->step = 3; c->e = 2; c->a = 1; /* src/FEM_map.c:67-70 */
  return c;
}

/* src/sequence_batch.c:82-121: every record with a non-empty sequence, in file order */
fo_ctx *fo_open_reference(const char *fasta_path) {
  rec_reader r;
  if (rr_open(&r, fasta_path) != 0) return NULL;
  fo_ctx *c = (fo_ctx *)calloc(1, sizeof *c);
  size_t m = 0;
  rec o; memset(&o, 0, sizeof o);
  long l;
  while ((l = rr_next(&r, &o)) >= 0) {
    if (l == 0) continue;
    if (c->n == m) {
      m = m ? m * 2 : 16;
      c->names = (char **)realloc(c->names, m * sizeof(char *));
      c->seqs = (char **)realloc(c->seqs, m * sizeof(char *));
      c->lens = (uint32_t *)realloc(c->lens, m * sizeof(uint32_t));
    }
    c->names[c->n] = strdup(o.name);
    c->seqs[c->n] = (char *)malloc((size_t)l + 1);
    memcpy(c->seqs[c->n], o.seq, (size_t)l + 1);
    c->lens[c->n] = (uint32_t)l;
    ++c->n;
  }
  rec_free(&o);
  free(r.d);
  c->k = 12; c->step = 3; c->e = 2; c->a = 1;
  return c;
}

void fo_close(fo_ctx *c) {
  if (!c) return;
  for (uint32_t i = 0; i < c->n; ++i) { free(c->names[i]); free(c->seqs[i]); }
  free(c->names); free(c->seqs); free(c->lens); free(c->lut); free(c->occ); free(c);
}
uint32_t fo_num_records(const fo_ctx *c) { return c->n; }
uint32_t fo_record_length(const fo_ctx *c, uint32_t i) { return c->lens[i]; }
const char *fo_record_name(const fo_ctx *c, uint32_t i) { return c->names[i]; }
const char *fo_record_bases(const fo_ctx *c, uint32_t i) { return c->seqs[i]; }
int fo_index_k(const fo_ctx *c) { return c->idx_k; }
int fo_index_step(const fo_ctx *c) { return c->idx_step; }
size_t fo_index_num_occurrences(const fo_ctx *c) { return c->n_occ; }
const uint32_t *fo_index_lookup_table(const fo_ctx *c) { return c->lut; }
const uint64_t *fo_index_occurrences(const fo_ctx *c) { return c->occ; }

int fo_set_params(fo_ctx *c, int k, int step, int e, int a) {
  c->k = k > 0 ? k : c->idx_k;
  c->step = step > 0 ? step : c->idx_step;
  if (e < 0 || e > 7 || a < 0 || a > 2) return -1; /* src/FEM_map.c:30-41 */
  c->e = e; c->a = a;
  return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* hashing: src/utils.h:83-117                                                                  */
/* ------------------------------------------------------------------------------------------ */
/* 2 bits per base, first base most significant, ambiguous base (or past the end) hashed as A */
static uint32_t hash_one(const char *s, size_t len, size_t pos, int k) {
  uint32_t mask = (uint32_t)((1ull << (2 * k)) - 1), h = 0;
  for (int i = 0; i < k; ++i) {
    uint8_t b = pos + (size_t)i < len ? base_code(s[pos + i]) : 4;
    h = ((h << 2) | (b < 4 ? b : 0)) & mask;
  }
  return h;
}

int fo_hash_all_seeds(const char *text, int l, int k, uint32_t *hashes) {
  int n_s = l - k + 1, amb = 0;
  uint32_t mask = (uint32_t)((1ull << (2 * k)) - 1);
  uint32_t h = hash_one(text, (size_t)l, 0, k);
  hashes[0] = h;
  for (int i = 1; i < n_s; ++i) {
    uint8_t b = base_code(text[i + k - 1]);
    if (b < 4) h = ((h << 2) | b) & mask;
    else { h = (h << 2) & mask; ++amb; } /* only *incoming* ambiguous bases are counted (:113) */
    hashes[i] = h;
  }
  return amb;
}

/* src/sequence_batch.h:90-98: 3^code, code 4 -> 7 -> 'N'; output is upper case */
void fo_reverse_complement(const char *seq, int l, char *out) {
  static const char back[8] = {'A', 'C', 'G', 'T', 'N', 'N', 'N', 'N'};
  for (int i = 0; i < l; ++i) out[i] = back[3 ^ base_code(seq[l - 1 - i])];
  out[l] = 0;
}

/* ------------------------------------------------------------------------------------------ */
/* index: src/index.c:57-98                                                                     */
/* ------------------------------------------------------------------------------------------ */
static int cmp_u64(const void *a, const void *b) {
  uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
  return x < y ? -1 : x > y;
}

/* Result of construct_index: L = prefix sums of bucket sizes, O = locations grouped by hash,
 * ascending inside a bucket (:93).  Counting sort here; the reference's radix sorts give the same
 * arrays because a bucket's content is a set of distinct locations. */
int fo_build_index(fo_ctx *c, int k, int step) {
  if (k < 1 || k > 15 || step < 1) return -1;
  size_t lut_n = ((size_t)1 << (2 * k)) + 1;
  free(c->lut); free(c->occ);
  c->lut = (uint32_t *)calloc(lut_n, sizeof(uint32_t));
  size_t total = 0;
  for (uint32_t s = 0; s < c->n; ++s)
    for (uint32_t p = 0; p + (uint32_t)k - 1 < c->lens[s]; p += (uint32_t)step) {
      ++c->lut[hash_one(c->seqs[s], c->lens[s], p, k) + 1];
      ++total;
    }
  for (size_t i = 1; i < lut_n; ++i) c->lut[i] += c->lut[i - 1];
  c->occ = (uint64_t *)malloc((total ? total : 1) * sizeof(uint64_t));
  uint32_t *fill = (uint32_t *)malloc(lut_n * sizeof(uint32_t));
  memcpy(fill, c->lut, lut_n * sizeof(uint32_t));
  for (uint32_t s = 0; s < c->n; ++s)
    for (uint32_t p = 0; p + (uint32_t)k - 1 < c->lens[s]; p += (uint32_t)step)
      c->occ[fill[hash_one(c->seqs[s], c->lens[s], p, k)]++] = ((uint64_t)s << 32) | p;
  free(fill);
  /* records are visited in ascending (s, p), so every bucket is already ascending; keep the
   * sort as a restatement of :93 for clarity on tiny inputs only */
  if (total < (1u << 16))
    for (size_t h = 0; h + 1 < lut_n; ++h)
      if (c->lut[h + 1] - c->lut[h] > 1)
        qsort(c->occ + c->lut[h], c->lut[h + 1] - c->lut[h], sizeof(uint64_t), cmp_u64);
  c->n_occ = total;
  c->idx_k = k; c->idx_step = step;
  return 0;
}

/* file layout src/index.c:133-168: int k; int step; uint32 L[4^k+1]; size_t n; uint64 O[n] */
int fo_save_index(const fo_ctx *c, const char *path) {
  FILE *f = fopen(path, "wb");
  if (!f) return -1;
  size_t lut_n = ((size_t)1 << (2 * c->idx_k)) + 1;
  int ok = fwrite(&c->idx_k, sizeof(int), 1, f) == 1 && fwrite(&c->idx_step, sizeof(int), 1, f) == 1 &&
           fwrite(c->lut, sizeof(uint32_t), lut_n, f) == lut_n &&
           fwrite(&c->n_occ, sizeof(size_t), 1, f) == 1 &&
           fwrite(c->occ, sizeof(uint64_t), c->n_occ, f) == c->n_occ;
  fclose(f);
  return ok ? 0 : -1;
}

int fo_load_index(fo_ctx *c, const char *path) {
  FILE *f = fopen(path, "rb");
  if (!f) return -1;
  int k, step;
  if (fread(&k, sizeof(int), 1, f) != 1 || fread(&step, sizeof(int), 1, f) != 1 || k < 1 || k > 15) { fclose(f); return -1; }
  size_t lut_n = ((size_t)1 << (2 * k)) + 1, n;
  free(c->lut); free(c->occ);
  c->lut = (uint32_t *)malloc(lut_n * sizeof(uint32_t));
  if (fread(c->lut, sizeof(uint32_t), lut_n, f) != lut_n || fread(&n, sizeof(size_t), 1, f) != 1) { fclose(f); return -1; }
  c->occ = (uint64_t *)malloc((n ? n : 1) * sizeof(uint64_t));
  if (fread(c->occ, sizeof(uint64_t), n, f) != n) { fclose(f); return -1; }
  fclose(f);
  c->n_occ = n; c->idx_k = k; c->idx_step = step;
  return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* group seeding + candidate filter: src/filter.c                                               */
/* ------------------------------------------------------------------------------------------ */
typedef struct { uint32_t hash, start, freq; } seed_t; /* src/utils.h:119-124 */

/* src/filter.c:3-43.  seeds[0..n_g) of one group, s = group-index spacing, Q seeds wanted.
 * Returns the DP minimum, *n_sel = number of seeds the traceback produced (right-most first). */
static uint32_t select_seeds(const seed_t *seeds, int n_g, int s, int Q, uint32_t inf,
                             seed_t *sel, int *n_sel) {
  int R = Q + 1, C = n_g - Q * s + 2;
  uint32_t *M = (uint32_t *)calloc((size_t)R * C, sizeof(uint32_t));
  uint8_t *D = (uint8_t *)calloc((size_t)R * C, 1);
  for (int r = 1; r < R; ++r) { M[r * C] = inf; D[r * C] = 3; }
  for (int c = 0; c < C; ++c) { M[c] = 0; D[c] = 3; }
  for (int r = 1; r < R; ++r)
    for (int c = 1; c < C; ++c) {
      uint32_t with_seed = M[(r - 1) * C + c] + seeds[c + (r - 1) * s - 1].freq;
      if (with_seed < M[r * C + c - 1]) { M[r * C + c] = with_seed; D[r * C + c] = 2; }
      else { M[r * C + c] = M[r * C + c - 1]; D[r * C + c] = 1; }
    }
  int r = R - 1, c = C - 1, n = 0;
  while (D[r * C + c] != 3) {
    if (D[r * C + c] == 2) { sel[n++] = seeds[c + (r - 1) * s - 1]; --r; }
    else --c;
  }
  uint32_t best = M[(R - 1) * C + C - 1];
  free(M); free(D);
  *n_sel = n;
  return best;
}

/* src/filter.c:80-116: fold the Q occurrence runs into one sorted multiset of (occ - start) */
static void merge_seed_locations(const fo_ctx *ctx, const seed_t *sel, int Q, u64vec *acc, u64vec *tmp) {
  for (int si = 0; si < Q; ++si) {
    const uint64_t *run = ctx->occ + ctx->lut[sel[si].hash];
    size_t ai = 0, ri = 0, rn = sel[si].freq;
    int last = si == Q - 1;
    tmp->n = 0;
    while (ai < acc->n || (!last && ri < rn)) {
      if (ai < acc->n) {
        if (ri < rn) {
          if ((uint32_t)run[ri] < sel[si].start) { ++ri; continue; }
          uint64_t v = run[ri] - sel[si].start;
          if (v <= acc->a[ai]) { u64_push(tmp, v); ++ri; }
          else u64_push(tmp, acc->a[ai++]);
        } else u64_push(tmp, acc->a[ai++]);
      } else {
        if ((uint32_t)run[ri] >= sel[si].start) u64_push(tmp, run[ri] - sel[si].start);
        ++ri;
      }
    }
    u64vec sw = *acc; *acc = *tmp; *tmp = sw;
  }
}

long fo_generate_candidates(const fo_ctx *ctx, const char *text, int l, uint64_t **cands,
                            size_t *cap, uint32_t *n_before) {
  const int k = ctx->k, step = ctx->step, e = ctx->e, a = ctx->a, Q = e + 1 + a;
  *n_before = 0;
  int s = k / step + (k % step > 0);       /* :162-165 */
  int n_s = l - k + 1;                     /* :166 */
  if (n_s <= 0) return -1;                 /* reference: assert(:167) */
  if (Q > n_s / step) return 0;            /* :168-172 */
  uint32_t *hashes = (uint32_t *)malloc((size_t)n_s * sizeof(uint32_t));
  if (fo_hash_all_seeds(text, l, k, hashes) > e) { free(hashes); return 0; } /* :176-182 */
  /* the reference runs into UB when a group is too short for the DP table */
  for (int g = 0; g < step; ++g)
    if ((n_s - g) / step - Q * s + 2 < 2) { free(hashes); return -1; }

  u64vec acc = {0, 0, 0}, tmp = {0, 0, 0}, surv = {0, 0, 0}, cur = {0, 0, 0}, nxt = {0, 0, 0};
  seed_t *seeds = (seed_t *)malloc((size_t)n_s * sizeof(seed_t));
  seed_t sel[16];
  long result = 0;
  for (int g = 0; g < step; ++g) {          /* :190-213 */
    int n_g = (n_s - g) / step;
    for (int j = 0; j < n_g; ++j) {
      int p = g + j * step;
      seeds[j].hash = hashes[p];
      seeds[j].start = (uint32_t)p;
      seeds[j].freq = ctx->lut[hashes[p] + 1] - ctx->lut[hashes[p]];
    }
    int n_sel = 0;
    *n_before += select_seeds(seeds, n_g, s, Q, (uint32_t)ctx->n_occ, sel, &n_sel);
    if (n_sel != Q) { result = -1; break; }  /* reference: uninitialised seeds (UB) */
    /* :204 qsort by frequency only; glibc's qsort is a stable merge sort here */
    for (int i = 1; i < Q; ++i) {
      seed_t t = sel[i];
      int j = i;
      while (j > 0 && sel[j - 1].freq > t.freq) { sel[j] = sel[j - 1]; --j; }
      sel[j] = t;
    }
    acc.n = 0;
    merge_seed_locations(ctx, sel, Q, &acc, &tmp);
    /* :118-131 keep v[i] when `a` further elements lie within +e */
    surv.n = 0;
    for (size_t i = 0; i < acc.n; ++i) {
      size_t in_range = 1;
      while (i + in_range < acc.n && acc.a[i + in_range] <= acc.a[i] + (uint64_t)e) {
        ++in_range;
        if (in_range > (size_t)a) break;
      }
      if (in_range > (size_t)a) u64_push(&surv, acc.a[i]);
    }
    /* :45-78 merge with the running list, pushing only values > last pushed + e */
    nxt.n = 0;
    size_t i1 = 0, i2 = 0;
    while (i1 < cur.n || i2 < surv.n) {
      uint64_t v;
      if (i1 < cur.n && (i2 >= surv.n || cur.a[i1] < surv.a[i2])) v = cur.a[i1++];
      else v = surv.a[i2++];
      if (nxt.n == 0 || v > nxt.a[nxt.n - 1] + (uint64_t)e) u64_push(&nxt, v);
    }
    u64vec sw = cur; cur = nxt; nxt = sw;
  }
  if (result == 0) {
    /* :133-144 range filter; the stored candidate is the band's left edge */
    size_t n = 0;
    if (cur.n > *cap) { *cap = cur.n * 2; *cands = (uint64_t *)realloc(*cands, *cap * sizeof(uint64_t)); }
    for (size_t i = 0; i < cur.n; ++i) {
      uint32_t seq = (uint32_t)(cur.a[i] >> 32), pos = (uint32_t)cur.a[i];
      uint32_t len = ctx->lens[seq];
      if (pos >= (uint32_t)e && pos + (uint32_t)l + (uint32_t)e < len) (*cands)[n++] = cur.a[i] - (uint64_t)e;
    }
    result = (long)n;
  } else *n_before = 0;
  free(hashes); free(seeds); free(acc.a); free(tmp.a); free(surv.a); free(cur.a); free(nxt.a);
  return result;
}

/* ------------------------------------------------------------------------------------------ */
/* verification: src/align.c:4-51, 102-277                                                      */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
  uint32_t peq[5], vp, vn, mask;
  int errors;
} band_t;

static inline const char *window(const fo_ctx *ctx, uint64_t cand) {
  return ctx->seqs[cand >> 32] + (uint32_t)cand;
}

static void band_init(band_t *b, const char *pattern, int e, int width) {
  memset(b, 0, sizeof *b);
  b->mask = width == 16 ? 0xFFFFu : 0xFFFFFFFFu;
  for (int i = 0; i < 2 * e; ++i) b->peq[base_code(pattern[i])] |= 1u << i; /* :104-107 / :180-200 */
}

/* one column, src/align.c:118-133 (width 32) / :214-255 (width 16: every op wraps at 16 bits) */
static inline uint32_t band_step(band_t *b, const char *pattern, const char *text, int i, int e) {
  const uint32_t m = b->mask;
  b->peq[base_code(pattern[i + 2 * e])] |= 1u << (2 * e);
  uint32_t X = b->peq[base_code(text[i])] | b->vn;
  uint32_t D0 = ((((b->vp + (X & b->vp)) & m) ^ b->vp) | X) & m;
  uint32_t HN = b->vp & D0;
  uint32_t HP = (b->vn | (~(b->vp | D0) & m)) & m;
  X = D0 >> 1;
  b->vn = X & HP;
  b->vp = (HN | (~(X | HP) & m)) & m;
  b->errors += 1 - (int)(D0 & 1u);
  for (int c = 0; c < 5; ++c) b->peq[c] >>= 1;
  return D0;
}

/* tail walk, :135-146 / :257-275: first strict minimum wins */
static int band_finish(const band_t *b, int l, int e, int *end_offset) {
  int err = b->errors, best = err;
  *end_offset = l - 1;
  for (int i = 0; i < 2 * e; ++i) {
    err += (int)((b->vp >> i) & 1u) - (int)((b->vn >> i) & 1u);
    if (err < best) { best = err; *end_offset = l + i; }
  }
  return best;
}

int fo_banded_edit_distance(const fo_ctx *ctx, uint64_t cand, const char *text, int l, int width,
                            int *end_offset) {
  const int e = ctx->e;
  const char *pattern = window(ctx, cand);
  band_t b;
  band_init(&b, pattern, e, width);
  for (int i = 0; i < l; ++i) {
    band_step(&b, pattern, text, i, e);
    if (width == 32 && b.errors > 3 * e) return e + 1; /* :128-130 */
  }
  return band_finish(&b, l, e, end_offset);
}

uint32_t fo_verify_candidates(const fo_ctx *ctx, const char *text, int l, int direction,
                              const uint64_t *cands, uint32_t n, fo_mapping *out) {
  const int e = ctx->e;
  uint32_t n_out = 0, n_vec = n / 8, n_tail = n % 8;
  for (uint32_t v = 0; v < n_vec; ++v) { /* :16-32: eight 16-bit lanes in lock step */
    band_t b[8];
    const char *pat[8];
    int dist[8], endo[8], stopped = 0;
    for (int li = 0; li < 8; ++li) {
      pat[li] = window(ctx, cands[v * 8 + li]);
      band_init(&b[li], pat[li], e, 16);
      endo[li] = l - 1; /* :17-19 */
    }
    for (int i = 0; i < l && !stopped; ++i) {
      int over = 0;
      for (int li = 0; li < 8; ++li) {
        band_step(&b[li], pat[li], text, i, e);
        over += b[li].errors > 3 * e;
      }
      if (over == 8) { /* :247-252: leave only when ALL lanes are hopeless */
        for (int li = 0; li < 8; ++li) dist[li] = b[li].errors;
        stopped = 1;
      }
    }
    if (!stopped)
      for (int li = 0; li < 8; ++li) dist[li] = band_finish(&b[li], l, e, &endo[li]);
    for (int li = 0; li < 8; ++li)
      if (dist[li] <= e) {
        fo_mapping m; memset(&m, 0, sizeof m);
        m.direction = (uint8_t)direction; m.edit_distance = (uint8_t)dist[li];
        m.candidate_position = cands[v * 8 + li]; m.end_position_offset = endo[li];
        out[n_out++] = m;
      }
  }
  for (uint32_t t = 0; t < n_tail; ++t) { /* :33-49 scalar 32-bit */
    uint64_t cand = cands[n_vec * 8 + t];
    int endo = -l;
    int d = fo_banded_edit_distance(ctx, cand, text, l, 32, &endo);
    if (d <= e) {
      fo_mapping m; memset(&m, 0, sizeof m);
      m.direction = (uint8_t)direction; m.edit_distance = (uint8_t)d;
      m.candidate_position = cand; m.end_position_offset = endo;
      out[n_out++] = m;
    }
  }
  return n_out;
}

/* ------------------------------------------------------------------------------------------ */
/* mapping order: src/align.c:53-57, src/ksort.h:98-151                                         */
/* ------------------------------------------------------------------------------------------ */
static inline uint64_t sort_key(const fo_mapping *m) {
  return ((uint64_t)m->edit_distance << 60) | ((uint64_t)m->direction << 59) |
         (m->candidate_position + (uint64_t)(int64_t)m->end_position_offset);
}

static void insertion_sort(fo_mapping *b, fo_mapping *e) { /* ksort.h:104-113, stable */
  for (fo_mapping *i = b + 1; i < e; ++i)
    if (sort_key(i) < sort_key(i - 1)) {
      fo_mapping t = *i, *j;
      for (j = i; j > b && sort_key(&t) < sort_key(j - 1); --j) *j = *(j - 1);
      *j = t;
    }
}

/* ksort.h:114-146: in-place MSD "American flag" pass on one byte, then recurse per bucket */
static void flag_sort(fo_mapping *beg, fo_mapping *end, int shift) {
  fo_mapping *lo[256], *hi[256];
  size_t cnt[256] = {0};
  for (fo_mapping *i = beg; i != end; ++i) ++cnt[(sort_key(i) >> shift) & 255];
  fo_mapping *p = beg;
  for (int d = 0; d < 256; ++d) { lo[d] = p; p += cnt[d]; hi[d] = p; }
  for (int d = 0; d < 256;) {
    if (lo[d] != hi[d]) {
      int t = (int)((sort_key(lo[d]) >> shift) & 255);
      if (t != d) { /* follow the displacement cycle until an element for bucket d turns up */
        fo_mapping carry = *lo[d];
        do {
          fo_mapping sw = *lo[t];
          *lo[t]++ = carry;
          carry = sw;
          t = (int)((sort_key(&carry) >> shift) & 255);
        } while (t != d);
        *lo[d]++ = carry;
      } else ++lo[d];
    } else ++d;
  }
  if (shift) {
    int next = shift > 8 ? shift - 8 : 0;
    fo_mapping *b = beg;
    for (int d = 0; d < 256; ++d) {
      fo_mapping *e2 = hi[d];
      if (e2 - b > 64) flag_sort(b, e2, next);
      else if (e2 - b > 1) insertion_sort(b, e2);
      b = e2;
    }
  }
}

void fo_sort_mappings(fo_mapping *m, size_t n) {
  if (n <= 64) insertion_sort(m, m + n);
  else flag_sort(m, m + n, 56);
}

/* ------------------------------------------------------------------------------------------ */
/* alignment traceback + MD: src/align.c:279-544                                                */
/* ------------------------------------------------------------------------------------------ */
static void md_tag(const char *ref, const char *read, const uint32_t *cigar, int n_cigar, char *md) {
  int matches = 0, rp = 0, qp = 0;
  char *o = md;
  for (int ci = 0; ci < n_cigar; ++ci) {
    int op = (int)(cigar[ci] & 15), len = (int)(cigar[ci] >> 4);
    if (op == 0) {
      for (int j = 0; j < len; ++j, ++rp, ++qp) {
        if (ref[rp] == read[qp]) ++matches;   /* raw chars, :513 */
        else {
          if (matches) { o += sprintf(o, "%d", matches); matches = 0; }
          *o++ = ref[rp];
        }
      }
    } else if (op == 1) qp += len;
    else if (op == 2) {
      if (matches) { o += sprintf(o, "%d", matches); matches = 0; }
      *o++ = '^';
      for (int j = 0; j < len; ++j) *o++ = ref[rp++];
    }
  }
  if (matches) o += sprintf(o, "%d", matches);
  *o = 0;
}

int fo_generate_alignment(const fo_ctx *ctx, uint64_t cand, const char *text, int l,
                          int edit_distance, int end_offset, uint32_t *cigar, int *n_cigar,
                          char *md) {
  const int e = ctx->e;
  const char *pattern = window(ctx, cand);
  int start = end_offset - l + 1; /* :285 */
  int mism = 0;
  for (int i = 0; i < l; ++i) mism += text[i] != pattern[start + i]; /* :289-293 */
  if (mism == 0) { /* :294-300 */
    cigar[0] = (uint32_t)l << 4;
    *n_cigar = 1;
    md_tag(pattern + start, text, cigar, 1, md);
    return start;
  }
  /* :303-338 recompute the band (32-bit words) keeping D0 and HP of every column */
  uint32_t *D0s = (uint32_t *)malloc((size_t)l * sizeof(uint32_t));
  uint32_t *HPs = (uint32_t *)malloc((size_t)l * sizeof(uint32_t));
  {
    uint32_t peq[5] = {0, 0, 0, 0, 0}, VP = 0, VN = 0;
    for (int i = 0; i < 2 * e; ++i) peq[base_code(pattern[i])] |= 1u << i;
    for (int i = 0; i < l; ++i) {
      peq[base_code(pattern[i + 2 * e])] |= 1u << (2 * e);
      uint32_t X = peq[base_code(text[i])] | VN;
      uint32_t D0 = ((VP + (X & VP)) ^ VP) | X;
      uint32_t HN = VP & D0, HP = VN | ~(VP | D0);
      X = D0 >> 1;
      VN = X & HP;
      VP = HN | ~(X | HP);
      D0s[i] = D0; HPs[i] = HP;
      for (int c = 0; c < 5; ++c) peq[c] >>= 1;
    }
  }
  /* :340-440 walk back from the end cell until `edit_distance` errors are consumed.  Run-length
   * ops are collected from the read's 3' end; 'S' marks errors at the very end of the read. */
  char *ops = (char *)calloc((size_t)l + 2, 1);
  int *cnt = (int *)calloc((size_t)l + 2, sizeof(int));
  int bit = end_offset - l + 1, tp = l - 1, endp = end_offset, nerr = 0, nops = 0;
  char pre = 'S';
  int pre_n = 1;
#define D0BIT(t) ((D0s[t] >> bit) & 1u)
#define HPBIT(t) ((HPs[t] >> bit) & 1u)
  if (D0BIT(tp) && pattern[endp] == text[tp]) { --tp; --endp; pre = 'M'; }
  else if (!D0BIT(tp)) { --tp; --endp; ++nerr; pre = 'S'; }
  else if (D0BIT(tp) && HPBIT(tp)) { --tp; ++bit; ++nerr; pre = 'S'; ++start; }
  else { free(D0s); free(HPs); free(ops); free(cnt); *n_cigar = 0; md[0] = 0; return -1; } /* :367 assert */
  while (tp >= 0 && nerr != edit_distance) {
    if (D0BIT(tp) && pattern[endp] == text[tp]) { /* match */
      --tp; --endp;
      if (pre != 'M') { ops[nops] = pre; cnt[nops++] = pre_n; pre = 'M'; pre_n = 1; }
      else ++pre_n;
    } else if (!D0BIT(tp)) { /* mismatch */
      --tp; --endp; ++nerr;
      if (pre == 'S') ++pre_n;
      else if (pre != 'M') { ops[nops] = pre; cnt[nops++] = pre_n; pre = 'M'; pre_n = 1; }
      else ++pre_n;
    } else if (D0BIT(tp) && HPBIT(tp)) { /* insertion: read base only */
      --tp; ++bit; ++nerr;
      if (pre == 'S') ++pre_n;
      else if (pre != 'I') { ops[nops] = pre; cnt[nops++] = pre_n; pre = 'I'; pre_n = 1; }
      else ++pre_n;
      ++start;
    } else { /* deletion: reference base only */
      --bit; --endp; ++nerr;
      if (pre != 'D') { ops[nops] = pre; cnt[nops++] = pre_n; pre = 'D'; pre_n = 1; }
      else ++pre_n;
      --start;
    }
  }
#undef D0BIT
#undef HPBIT
  int last; /* index of the last op written, :445-459 */
  if (tp >= 0) {
    if (pre != 'M') { ops[nops] = pre; cnt[nops++] = pre_n; ops[nops] = 'M'; cnt[nops] = tp + 1; }
    else { ops[nops] = 'M'; cnt[nops] = pre_n + tp + 1; }
  } else { ops[nops] = pre; cnt[nops] = pre_n; }
  last = nops;
  int first = 0;
  if (ops[0] == 'S') { cnt[1] += cnt[0]; first = 1; } /* :466-469 fold the 3' errors */
  int nc = 0;
  for (int i = last; i >= first; --i) {
    uint32_t code = ops[i] == 'M' ? 0u : ops[i] == 'I' ? 1u : ops[i] == 'D' ? 2u : 15u;
    cigar[nc++] = ((uint32_t)cnt[i] << 4) | code;
  }
  *n_cigar = nc;
  md_tag(pattern + start, text, cigar, nc, md);
  free(D0s); free(HPs); free(ops); free(cnt);
  return start;
}

/* ------------------------------------------------------------------------------------------ */
/* whole reads                                                                                 */
/* ------------------------------------------------------------------------------------------ */
size_t fo_map_read(const fo_ctx *ctx, const char *seq, int l, fo_mapping **maps, size_t *cap,
                   fo_stats *st) {
  uint64_t *cands = NULL;
  size_t ccap = 0, n_maps = 0;
  char *rc = (char *)malloc((size_t)l + 1);
  fo_reverse_complement(seq, l, rc);
  int infeasible = 0;
  ++st->num_reads;
  for (int dir = 0; dir < 2; ++dir) { /* src/map.c:29-49 */
    const char *text = dir ? rc : seq;
    uint32_t before = 0;
    long n = fo_generate_candidates(ctx, text, l, &cands, &ccap, &before);
    if (n < 0) { infeasible = 1; n = 0; }
    st->num_candidates_before_filter += before;
    st->num_candidates += (uint64_t)n;
    if (n > 0) {
      if (n_maps + (size_t)n > *cap) {
        *cap = (n_maps + (size_t)n) * 2;
        *maps = (fo_mapping *)realloc(*maps, *cap * sizeof(fo_mapping));
      }
      uint32_t got = fo_verify_candidates(ctx, text, l, dir, cands, (uint32_t)n, *maps + n_maps);
      st->num_mappings += got;
      n_maps += got;
    }
  }
  st->num_infeasible_reads += (uint64_t)infeasible;
  if (n_maps) { ++st->num_mapped_reads; fo_sort_mappings(*maps, n_maps); }
  free(cands); free(rc);
  return n_maps;
}

/* htslib hts.c:73-94 */
static char nt16_char(char c) {
  static const char *str = "=ACMGRSVTWYHKDBN";
  static const char *letters = "ACMGRSVTWYHKDBN"; /* codes 1..15 */
  if (c == '=') return '=';
  if (c >= '0' && c <= '3') return str[1 << (c - '0')];
  char u = (char)toupper((unsigned char)c);
  const char *p = u ? strchr(letters, u) : NULL;
  return p ? *p : 'N';
}

void fo_format_header(const fo_ctx *ctx, char **buf, size_t *len, size_t *cap) {
  for (uint32_t i = 0; i < ctx->n; ++i) { /* src/output_queue.c:104-108 */
    buf_putn(buf, len, cap, "@SQ\tSN:", 7);
    buf_putn(buf, len, cap, ctx->names[i], strlen(ctx->names[i]));
    buf_putn(buf, len, cap, "\tLN:", 4);
    buf_putll(buf, len, cap, (int)ctx->lens[i]);
    buf_putc(buf, len, cap, '\n');
  }
}

void fo_format_sam(const fo_ctx *ctx, const char *name, const char *seq, const char *qual, int l,
                   const fo_mapping *maps, size_t n, char **buf, size_t *len, size_t *cap) {
  char *rc = (char *)malloc((size_t)l + 1);
  uint32_t *cigar = (uint32_t *)malloc(((size_t)l + 2) * sizeof(uint32_t));
  char *md = (char *)malloc(4 * (size_t)l + 64);
  fo_reverse_complement(seq, l, rc);
  for (size_t mi = 0; mi < n; ++mi) { /* src/align.c:70-89 */
    const fo_mapping *m = &maps[mi];
    int n_cigar = 0;
    int start = fo_generate_alignment(ctx, m->candidate_position, m->direction ? rc : seq, l,
                                      m->edit_distance, m->end_position_offset, cigar, &n_cigar, md);
    uint32_t pos = (uint32_t)((uint32_t)start + (uint32_t)m->candidate_position); /* :80 */
    int flag = (m->direction ? 16 : 0) | (mi ? 256 : 0);
    /* htslib sam.c:3004-3061 */
    buf_putn(buf, len, cap, name, strlen(name)); buf_putc(buf, len, cap, '\t');
    buf_putll(buf, len, cap, flag); buf_putc(buf, len, cap, '\t');
    const char *rn = ctx->names[m->candidate_position >> 32];
    buf_putn(buf, len, cap, rn, strlen(rn)); buf_putc(buf, len, cap, '\t');
    buf_putll(buf, len, cap, (long long)pos + 1); buf_putc(buf, len, cap, '\t');
    buf_putn(buf, len, cap, "255\t", 4);
    if (n_cigar == 0) buf_putc(buf, len, cap, '*');
    for (int ci = 0; ci < n_cigar; ++ci) {
      buf_putll(buf, len, cap, cigar[ci] >> 4);
      buf_putc(buf, len, cap, "MIDNSHP=XB??????"[cigar[ci] & 15]);
    }
    buf_putn(buf, len, cap, "\t*\t0\t0\t", 7);
    if (mi == 0 && l > 0) { /* primary carries SEQ/QUAL in the FASTQ's orientation (:79,:87) */
      for (int i = 0; i < l; ++i) buf_putc(buf, len, cap, nt16_char(seq[i]));
      buf_putc(buf, len, cap, '\t');
      if ((uint8_t)(qual[0] - 33) == 0xff) buf_putc(buf, len, cap, '*');
      else buf_putn(buf, len, cap, qual, (size_t)l);
    } else buf_putn(buf, len, cap, "*\t*", 3);
    buf_putn(buf, len, cap, "\tNM:i:", 6);
    buf_putll(buf, len, cap, m->edit_distance);
    buf_putn(buf, len, cap, "\tMD:Z:", 6);
    buf_putn(buf, len, cap, md, strlen(md));
    buf_putc(buf, len, cap, '\n');
  }
  free(rc); free(cigar); free(md);
}

int fo_map_file(const fo_ctx *ctx, const char *fastq_path, const char *sam_path, fo_stats *stats) {
  rec_reader r;
  if (rr_open(&r, fastq_path) != 0) return -1;
  FILE *out = fopen(sam_path, "wb");
  if (!out) { free(r.d); return -1; }
  char *buf = NULL;
  size_t len = 0, cap = 0, mcap = 0;
  fo_mapping *maps = NULL;
  memset(stats, 0, sizeof *stats);
  fo_format_header(ctx, &buf, &len, &cap);
  rec o; memset(&o, 0, sizeof o);
  long l;
  int rc = 0;
  while ((l = rr_next(&r, &o)) != -1) {
    if (l == 0) continue;                 /* src/sequence_batch.c:50-52 */
    if (l < 0) { rc = -2; break; }        /* :63-68 */
    size_t n = fo_map_read(ctx, o.seq, (int)l, &maps, &mcap, stats);
    if (n) fo_format_sam(ctx, o.name, o.seq, o.qual_l ? o.qual : o.seq, (int)l, maps, n, &buf, &len, &cap);
    if (len > (1u << 22)) { fwrite(buf, 1, len, out); len = 0; }
  }
  fwrite(buf, 1, len, out);
  fclose(out);
  rec_free(&o); free(r.d); free(buf); free(maps);
  return rc;
}
