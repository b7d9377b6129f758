/* sam.c -- from device mappings to SAM text on the host.
 *
 * Stands in for src/align.c:56-92 (process_mappings), :279-499 (generate_alignment), :501-544
 * (generate_MD_tag), :546-632 (generate_bam1_t) and the text htslib 1.10.2 prints for such a
 * record (extern/htslib/sam.c:3004-3061), plus the header of src/output_queue.c:93-116.  No
 * bam1_t is built: the line is written straight into a text buffer, byte for byte what
 * sam_write1 would print for the record the reference fills in.
 */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>

#include "femhost.h"
#include "host_internal.h"

double femhost_now(void) {
  struct timeval tv;
  gettimeofday(&tv, NULL);
  return (double)tv.tv_sec + 1e-6 * (double)tv.tv_usec;
}

/* ---- text buffer --------------------------------------------------------------------------------- */
int femhost_str_reserve(femhost_str *s, size_t extra) {
  if (s->l + extra + 1 <= s->m) return 0;
  size_t m = s->m ? s->m : 4096;
  while (m < s->l + extra + 1) m *= 2;
  char *p = (char *)realloc(s->s, m);
  if (!p) return femhost_set_error(-3, "out of memory (%zu bytes of SAM text)", m);
  s->s = p;
  s->m = m;
  return 0;
}
void femhost_str_free(femhost_str *s) {
  free(s->s);
  s->s = NULL;
  s->l = s->m = 0;
}
/* the callers reserve first; these never grow */
static inline void put_mem(femhost_str *s, const void *p, size_t n) { memcpy(s->s + s->l, p, n); s->l += n; }
static inline void put_char(femhost_str *s, char c) { s->s[s->l++] = c; }
static inline void put_uint(femhost_str *s, uint64_t v) {
  char tmp[24];
  int n = 0;
  do { tmp[n++] = (char)('0' + v % 10); v /= 10; } while (v);
  while (n) s->s[s->l++] = tmp[--n];
}
static inline char *fmt_uint(char *o, unsigned v) {
  char tmp[12];
  int n = 0;
  do { tmp[n++] = (char)('0' + v % 10); v /= 10; } while (v);
  while (n) *o++ = tmp[--n];
  return o;
}

int femhost_sam_header(const femhost_records *ref, femhost_str *out) {
  for (uint32_t i = 0; i < ref->n_records; ++i) {
    const char *name = femhost_record_name(ref, i);
    const size_t nl = strlen(name);
    if (femhost_str_reserve(out, nl + 48)) return -3;
    put_mem(out, "@SQ\tSN:", 7);
    put_mem(out, name, nl);
    put_mem(out, "\tLN:", 4);
    const int32_t len = (int32_t)femhost_record_length(ref, i); /* printed with %d, src/output_queue.c:107 */
    if (len < 0) { put_char(out, '-'); put_uint(out, (uint64_t)(-(int64_t)len)); }
    else put_uint(out, (uint64_t)len);
    put_char(out, '\n');
  }
  return 0;
}

/* ---- alignment: start position, CIGAR, MD ------------------------------------------------------- */
/* src/align.c:501-544.  ref = window + start.  Raw characters are compared. */
static void md_text(const char *ref, const char *read, const uint32_t *cigar, int n_cigar, char *md) {
  unsigned run = 0;
  char *o = md;
  for (int ci = 0; ci < n_cigar; ++ci) {
    const unsigned op = cigar[ci] & 15u, len = cigar[ci] >> 4;
    if (op == 0) { /* M */
      for (unsigned j = 0; j < len; ++j) {
        if (ref[j] == read[j]) { ++run; continue; }
        if (run) { o = fmt_uint(o, run); run = 0; }
        *o++ = ref[j];
      }
      ref += len;
      read += len;
    } else if (op == 1) { /* I */
      read += len;
    } else { /* D */
      if (run) { o = fmt_uint(o, run); run = 0; }
      *o++ = '^';
      memcpy(o, ref, len);
      o += len;
      ref += len;
    }
  }
  if (run) o = fmt_uint(o, run);
  *o = 0;
}

int femhost_alignment(const char *window, const char *text, int l, int e, int edit_distance, int end_offset,
                      uint32_t *cigar, int *n_cigar, char *md) {
  int start = end_offset - l + 1; /* src/align.c:285 */
  *n_cigar = 0;
  md[0] = 0;
  if (l <= 0 || start < 0) return femhost_set_error(FEMHOST_ALIGN_ERROR, "alignment: bad end offset %d for read length %d", end_offset, l);
  int diff = 0;
  for (int i = 0; i < l; ++i) diff += text[i] != window[start + i];
  if (diff == 0) { /* :294-300 */
    cigar[0] = (uint32_t)l << 4;
    *n_cigar = 1;
    md_text(window + start, text, cigar, 1, md);
    return start;
  }
  /* :303-338: the band again, 32-bit words, keeping D0 and HP of every column */
  uint32_t D0s[l], HPs[l];
  {
    uint32_t peq[5] = {0, 0, 0, 0, 0}, VP = 0, VN = 0;
    const uint32_t top = 1u << (2 * e);
    for (int i = 0; i < 2 * e; ++i) peq[FEM_CODE(window[i])] |= 1u << i;
    for (int i = 0; i < l; ++i) {
      peq[FEM_CODE(window[i + 2 * e])] |= top;
      uint32_t X = peq[FEM_CODE(text[i])] | VN;
      const uint32_t D0 = ((VP + (X & VP)) ^ VP) | X;
      const uint32_t HN = VP & D0, HP = VN | ~(VP | D0);
      X = D0 >> 1;
      VN = X & HP;
      VP = HN | ~(X | HP);
      D0s[i] = D0;
      HPs[i] = HP;
      peq[0] >>= 1; peq[1] >>= 1; peq[2] >>= 1; peq[3] >>= 1; peq[4] >>= 1;
    }
  }
  /* :340-440: walk back from the end cell until `edit_distance` errors are used up.  Runs are
   * collected from the read's 3' end; errors before the first match sit in a pseudo run 'S' that
   * is folded into the run after it (:466-469).  The walk starts in an empty 'S' run, which makes
   * the reference's special first step (:345-368) the general step. */
  char ops[l + 2];
  int cnt[l + 2];
  int n_ops = 0, bit = start, tp = l - 1, rp = end_offset, used = 0;
  char run = 'S';
  int run_n = 0;
#define CLOSE_RUN(next) do { ops[n_ops] = run; cnt[n_ops++] = run_n; run = (next); run_n = 1; } while (0)
  while (tp >= 0 && (used != edit_distance || tp == l - 1)) {
    const uint32_t d0 = (D0s[tp] >> bit) & 1u;
    if (d0 && window[rp] == text[tp]) { /* match */
      --tp; --rp;
      if (run != 'M') CLOSE_RUN('M'); else ++run_n;
    } else if (!d0) { /* mismatch */
      --tp; --rp; ++used;
      if (run == 'S' || run == 'M') ++run_n; else CLOSE_RUN('M');
    } else if ((HPs[tp] >> bit) & 1u) { /* insertion: a read base only */
      --tp; ++bit; ++used; ++start;
      if (run == 'S' || run == 'I') ++run_n; else CLOSE_RUN('I');
    } else { /* deletion: a reference base only */
      if (tp == l - 1 && n_ops == 0 && run == 'S' && run_n == 0)
        return femhost_set_error(FEMHOST_ALIGN_ERROR, "alignment: deletion at the read end (the reference asserts, src/align.c:367)");
      --bit; --rp; ++used; --start;
      if (run != 'D') CLOSE_RUN('D'); else ++run_n;
    }
  }
  /* :441-459: whatever is left of the read matches */
  if (tp >= 0) {
    if (run != 'M') { CLOSE_RUN('M'); run_n = tp + 1; }
    else run_n += tp + 1;
  }
  ops[n_ops] = run;
  cnt[n_ops] = run_n;
#undef CLOSE_RUN
  int first = 0;
  if (ops[0] == 'S') { /* :466-469 */
    if (n_ops == 0) return femhost_set_error(FEMHOST_ALIGN_ERROR, "alignment: read consists of errors only");
    cnt[1] += cnt[0];
    first = 1;
  }
  int nc = 0;
  for (int i = n_ops; i >= first; --i) cigar[nc++] = ((uint32_t)cnt[i] << 4) | (ops[i] == 'M' ? 0u : ops[i] == 'I' ? 1u : 2u);
  *n_cigar = nc;
  md_text(window + start, text, cigar, nc, md);
  return start;
}

/* ---- one read's records -------------------------------------------------------------------------- */
/* what sam_format1 prints for a base after generate_bam1_t stored seq_nt16_table[c]
 * (extern/htslib/hts.c:73-94): IUPAC letters (either case) as upper case, '=' as '=', digits
 * 0..3 as ACGT, anything else 'N' */
static char g_seq_out[256];
static pthread_once_t g_seq_out_once = PTHREAD_ONCE_INIT;
static void init_seq_out(void) {
  static const char iupac[] = "=ACMGRSVTWYHKDBN";
  char t[256];
  memset(t, 'N', sizeof t);
  for (int i = 0; iupac[i]; ++i) {
    t[(unsigned char)iupac[i]] = iupac[i];
    if (iupac[i] >= 'A') t[(unsigned char)(iupac[i] + 32)] = iupac[i];
  }
  t['0'] = 'A'; t['1'] = 'C'; t['2'] = 'G'; t['3'] = 'T';
  t[0] = 'N';
  memcpy(g_seq_out, t, sizeof t);
}

/* src/sequence_batch.h:90-98 */
static void make_rc(const char *seq, int l, char *rc) {
  static const char back[8] = {'A', 'C', 'G', 'T', 'N', 'N', 'N', 'N'};
  for (int i = 0; i < l; ++i) rc[i] = back[3 ^ FEM_CODE(seq[l - 1 - i])];
  rc[l] = 0;
}

int femhost_sam_read(const femhost_records *ref, const char *name, const char *seq, const char *qual, int l, int e,
                     femgpu_mapping *maps, femgpu_alignment *alns, size_t n, femhost_str *out) {
  if (n == 0) return 0; /* src/map.c:50: unmapped reads print nothing */
  pthread_once(&g_seq_out_once, init_seq_out); /* several SAM workers call this at once */
  femgpu_sort_mappings_aln(maps, alns, n); /* src/align.c:57; the device alignments follow their mappings */
  const size_t name_l = strlen(name);
  char rc[l + 1], md[4 * l + 64];
  uint32_t cigar[l + 2];
  int have_rc = 0;
  for (size_t mi = 0; mi < n; ++mi) {
    const femgpu_mapping *m = &maps[mi];
    const uint32_t rec = (uint32_t)(m->candidate_position >> 32), cpos = (uint32_t)m->candidate_position;
    if (rec >= ref->n_records) return femhost_set_error(-1, "mapping names record %u of %u", rec, ref->n_records);
    int n_cigar = 0, start;
    const char *window = ref->bases + ref->base_off[rec] + cpos;
    if (alns && alns[mi].n_cigar != FEMGPU_ALN_HOST) { /* start and CIGAR came from the device (k_trace) */
      const femgpu_alignment *al = &alns[mi];
      start = al->start;
      n_cigar = al->n_cigar;
      for (int ci = 0; ci < n_cigar; ++ci) cigar[ci] = ((uint32_t)(al->cigar[ci] >> 2) << 4) | (al->cigar[ci] & 3u);
      if (al->md_len <= FEMGPU_MD_MAX) { /* the MD text came along as well: no reference access on the host */
        memcpy(md, al->md, al->md_len);
        md[al->md_len] = 0;
      } else {
        if (m->direction && !have_rc) { make_rc(seq, l, rc); have_rc = 1; }
        md_text(window + start, m->direction ? rc : seq, cigar, n_cigar, md);
      }
    } else {
      if (m->direction && !have_rc) { make_rc(seq, l, rc); have_rc = 1; }
      start = femhost_alignment(window, m->direction ? rc : seq, l, e, m->edit_distance, m->end_position_offset, cigar,
                                &n_cigar, md);
      if (start == FEMHOST_ALIGN_ERROR) return -2;
    }
    const uint32_t pos = (uint32_t)start + cpos; /* src/align.c:80 */
    const char *rname = femhost_record_name(ref, rec);
    const size_t rname_l = strlen(rname), md_l = strlen(md);
    if (femhost_str_reserve(out, name_l + rname_l + md_l + (size_t)n_cigar * 12 + 2 * (size_t)l + 96)) return -3;
    put_mem(out, name, name_l);
    put_char(out, '\t');
    put_uint(out, (uint64_t)((m->direction ? 16 : 0) | (mi ? 256 : 0)));
    put_char(out, '\t');
    put_mem(out, rname, rname_l);
    put_char(out, '\t');
    put_uint(out, (uint64_t)pos + 1);
    put_mem(out, "\t255\t", 5);
    if (n_cigar == 0) put_char(out, '*');
    for (int ci = 0; ci < n_cigar; ++ci) {
      put_uint(out, cigar[ci] >> 4);
      put_char(out, "MID"[cigar[ci] & 3]);
    }
    put_mem(out, "\t*\t0\t0\t", 7);
    if (mi == 0) { /* the primary record carries SEQ/QUAL in the file's orientation (:79,:87) */
      char *o = out->s + out->l;
      for (int i = 0; i < l; ++i) o[i] = g_seq_out[(unsigned char)seq[i]];
      out->l += (size_t)l;
      put_char(out, '\t');
      if ((unsigned char)(qual[0] - 33) == 0xff) put_char(out, '*'); /* htslib sam.c:3047 */
      else put_mem(out, qual, (size_t)l);
    } else {
      put_mem(out, "*\t*", 3);
    }
    put_mem(out, "\tNM:i:", 6);
    put_uint(out, m->edit_distance);
    put_mem(out, "\tMD:Z:", 6);
    put_mem(out, md, md_l);
    put_char(out, '\n');
  }
  return 0;
}
