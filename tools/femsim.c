/* femsim -- seeded synthetic data for the `FEM map` path.
 *
 *   femsim ref   <seed> <total_bp> <n_records> <repeat_permille> <out.fa>
 *   femsim reads <seed> <ref.fa> <n_reads> <read_len> <max_edits> <out.fq>
 *
 * ref:   upper-case ACGT, i.i.d. uniform, split into n_records records of (nearly) equal length
 *        named chr1..chrN, 60 columns.  repeat_permille > 0 overwrites that share of every record
 *        with tandem arrays (unit 2..200 bp, 10..500 copies, 1% divergence per copy) and copies of
 *        60 interspersed families (50 x 300 bp, 10 x 6 kbp; 5..15% substitutions per copy).
 * reads: single-end, fixed length, uniform start (>= max_edits+1 from both record ends), strand
 *        50/50, m ~ U{0..max_edits} edits per read (80% substitution, 10% insertion, 10% deletion),
 *        quality 'I', name r<idx>_<rec>_<pos>_<strand>_<m>.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static uint64_t rng_s[4];
static uint64_t splitmix(uint64_t *x) {
  uint64_t z = (*x += 0x9e3779b97f4a7c15ull);
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}
static void rng_seed(uint64_t seed) { for (int i = 0; i < 4; ++i) rng_s[i] = splitmix(&seed); }
static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
static inline uint64_t rng(void) { /* xoshiro256** */
  uint64_t r = rotl(rng_s[1] * 5, 7) * 9, t = rng_s[1] << 17;
  rng_s[2] ^= rng_s[0]; rng_s[3] ^= rng_s[1]; rng_s[1] ^= rng_s[2]; rng_s[0] ^= rng_s[3];
  rng_s[2] ^= t; rng_s[3] = rotl(rng_s[3], 45);
  return r;
}
static inline uint64_t rng_below(uint64_t n) { return (uint64_t)(((__uint128_t)rng() * n) >> 64); }

static const char ACGT[4] = {'A', 'C', 'G', 'T'};

static void mutate_copy(char *dst, const char *src, size_t n, unsigned permille) {
  for (size_t i = 0; i < n; ++i) {
    dst[i] = src[i];
    if (rng_below(1000) < permille) dst[i] = ACGT[rng_below(4)];
  }
}

static int cmd_ref(int argc, char **argv) {
  if (argc < 7) return 2;
  uint64_t seed = strtoull(argv[2], 0, 10), total = strtoull(argv[3], 0, 10);
  unsigned n_rec = (unsigned)atoi(argv[4]), rep = (unsigned)atoi(argv[5]);
  FILE *f = fopen(argv[6], "w");
  if (!f || n_rec == 0) return 1;
  rng_seed(seed);
  /* interspersed families are shared by all records */
  enum { NF = 60 };
  char *fam[NF]; size_t fam_len[NF];
  for (int i = 0; i < NF; ++i) {
    fam_len[i] = i < 50 ? 300 : 6000;
    fam[i] = (char *)malloc(fam_len[i]);
    for (size_t j = 0; j < fam_len[i]; ++j) fam[i][j] = ACGT[rng_below(4)];
  }
  uint64_t per = total / n_rec;
  char *buf = (char *)malloc(per + n_rec + 1);
  for (unsigned r = 0; r < n_rec; ++r) {
    uint64_t len = r + 1 == n_rec ? total - per * (n_rec - 1) : per;
    for (uint64_t i = 0; i < len; i += 32) {
      uint64_t w = rng();
      for (int j = 0; j < 32 && i + j < len; ++j, w >>= 2) buf[i + j] = ACGT[w & 3];
    }
    if (rep) {
      uint64_t budget = len * rep / 1000, used = 0;
      while (used < budget) {
        if (rng_below(2)) { /* tandem array */
          size_t unit = 2 + rng_below(199), copies = 10 + rng_below(491);
          size_t span = unit * copies;
          if (span + 1 >= len) continue;
          uint64_t p = rng_below(len - span);
          for (size_t c = 1; c < copies; ++c) mutate_copy(buf + p + c * unit, buf + p, unit, 10);
          used += span;
        } else {
          int fi = (int)rng_below(NF);
          if (fam_len[fi] + 1 >= len) continue;
          uint64_t p = rng_below(len - fam_len[fi]);
          mutate_copy(buf + p, fam[fi], fam_len[fi], 50 + (unsigned)rng_below(101));
          used += fam_len[fi];
        }
      }
    }
    fprintf(f, ">chr%u\n", r + 1);
    for (uint64_t i = 0; i < len; i += 60) {
      size_t n = len - i < 60 ? (size_t)(len - i) : 60;
      fwrite(buf + i, 1, n, f);
      fputc('\n', f);
    }
  }
  fclose(f);
  return 0;
}

static char comp(char c) { return c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : 'N'; }

static int cmd_reads(int argc, char **argv) {
  if (argc < 8) return 2;
  uint64_t seed = strtoull(argv[2], 0, 10);
  FILE *fa = fopen(argv[3], "r");
  uint64_t n_reads = strtoull(argv[4], 0, 10);
  int l = atoi(argv[5]), e = atoi(argv[6]);
  FILE *fq = fopen(argv[7], "w");
  if (!fa || !fq) return 1;
  /* load the FASTA (our own 60-column output) */
  size_t cap = 1 << 20, n = 0, n_rec = 0, rec_cap = 64;
  char *g = (char *)malloc(cap);
  uint64_t *rec_off = (uint64_t *)malloc(rec_cap * sizeof(uint64_t));
  char line[4096];
  while (fgets(line, sizeof line, fa)) {
    if (line[0] == '>') {
      if (n_rec + 2 > rec_cap) { rec_cap *= 2; rec_off = (uint64_t *)realloc(rec_off, rec_cap * sizeof(uint64_t)); }
      rec_off[n_rec++] = n;
      continue;
    }
    size_t ll = strlen(line);
    while (ll && (line[ll - 1] == '\n' || line[ll - 1] == '\r')) --ll;
    if (n + ll > cap) { while (n + ll > cap) cap *= 2; g = (char *)realloc(g, cap); }
    memcpy(g + n, line, ll);
    n += ll;
  }
  rec_off[n_rec] = n;
  fclose(fa);
  rng_seed(seed);
  char *w = (char *)malloc((size_t)l + 4 * (size_t)e + 16), *out = (char *)malloc((size_t)l + 1);
  setvbuf(fq, NULL, _IOFBF, 1 << 22);
  for (uint64_t ri = 0; ri < n_reads; ++ri) {
    size_t rec; uint64_t rlen, p;
    int span = l + e + 1;
    do { /* uniform over bases => records weighted by length */
      uint64_t gpos = rng_below(n);
      size_t lo = 0, hi = n_rec;
      while (hi - lo > 1) { size_t mid = (lo + hi) / 2; if (rec_off[mid] <= gpos) lo = mid; else hi = mid; }
      rec = lo; rlen = rec_off[rec + 1] - rec_off[rec]; p = gpos - rec_off[rec];
    } while (rlen < (uint64_t)(span + 2 * e + 2) || p < (uint64_t)e + 1 || p + span + e + 1 > rlen);
    int wl = span;
    memcpy(w, g + rec_off[rec] + p, (size_t)wl);
    int m = (int)rng_below((uint64_t)e + 1);
    for (int k = 0; k < m; ++k) {
      unsigned kind = (unsigned)rng_below(10);
      int at = (int)rng_below((uint64_t)l);
      if (kind < 8) { char c; do c = ACGT[rng_below(4)]; while (c == w[at]); w[at] = c; }
      else if (kind == 8) { memmove(w + at + 1, w + at, (size_t)(wl - at)); w[at] = ACGT[rng_below(4)]; ++wl; }
      else { memmove(w + at, w + at + 1, (size_t)(wl - at - 1)); --wl; }
    }
    int strand = (int)rng_below(2);
    if (strand) for (int i = 0; i < l; ++i) out[i] = comp(w[l - 1 - i]);
    else memcpy(out, w, (size_t)l);
    out[l] = 0;
    fprintf(fq, "@r%llu_%zu_%llu_%c_%d\n%s\n+\n", (unsigned long long)ri, rec, (unsigned long long)p, strand ? '-' : '+', m, out);
    for (int i = 0; i < l; ++i) fputc('I', fq);
    fputc('\n', fq);
  }
  fclose(fq);
  return 0;
}

int main(int argc, char **argv) {
  int rc = 2;
  if (argc >= 2 && !strcmp(argv[1], "ref")) rc = cmd_ref(argc, argv);
  else if (argc >= 2 && !strcmp(argv[1], "reads")) rc = cmd_reads(argc, argv);
  if (rc == 2) fprintf(stderr, "usage: femsim ref <seed> <total_bp> <n_records> <repeat_permille> <out.fa>\n"
                               "       femsim reads <seed> <ref.fa> <n_reads> <read_len> <max_edits> <out.fq>\n");
  return rc;
}
