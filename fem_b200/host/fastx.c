/* fastx.c -- FASTA/FASTQ records into flattened batches (include/femhost.h).
 *
 * Record semantics are those of the reference's reader (klib kseq, src/kseq.h:185-226, driven by
 * src/sequence_batch.c:44-121): header char '>' or '@', name = text up to the first white space,
 * multi-line sequences, '+' line, quality lines until as long as the sequence, a trailing '\r' of
 * a line is dropped, zero-length records are skipped, a short quality string is an error.
 * The implementation is a block scanner (memchr over a 4 MiB window) writing straight into the
 * batch's flat arrays.
 */
#include <ctype.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

#include "femhost.h"
#include "host_internal.h"

static __thread char g_err[512];
const char *femhost_last_error(void) { return g_err; }
int femhost_set_error(int code, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return code;
}

/* ---- records ----------------------------------------------------------------------------------- */
void femhost_records_init(femhost_records *r, int pinned) {
  memset(r, 0, sizeof *r);
  r->pinned = pinned;
}

static void *grow(void *p, size_t old_bytes, size_t new_bytes, int pinned) {
  if (!pinned) return realloc(p, new_bytes);
  void *q = femgpu_alloc_pinned(new_bytes);
  if (!q) return NULL;
  if (p) { memcpy(q, p, old_bytes); femgpu_free_pinned(p); }
  return q;
}

int femhost_records_reserve(femhost_records *r, size_t n_records, size_t n_bases, size_t n_names) {
  if (n_records + 1 > r->cap_records) {
    size_t m = r->cap_records ? r->cap_records : 1024;
    while (m < n_records + 1) m *= 2;
    uint64_t *bo = (uint64_t *)grow(r->base_off, r->cap_records * 8, m * 8, r->pinned);
    uint64_t *no = (uint64_t *)grow(r->name_off, r->cap_records * 8, m * 8, r->pinned);
    uint8_t *hq = (uint8_t *)grow(r->has_qual, r->cap_records, m, r->pinned);
    if (bo) r->base_off = bo;
    if (no) r->name_off = no;
    if (hq) r->has_qual = hq;
    if (!bo || !no || !hq) return femhost_set_error(-3, "out of memory (%zu records)", m);
    if (r->cap_records == 0) { r->base_off[0] = 0; r->name_off[0] = 0; }
    r->cap_records = m;
  }
  if (n_bases > r->cap_bases) {
    size_t m = r->cap_bases ? r->cap_bases : (1u << 16);
    while (m < n_bases) m *= 2;
    char *b = (char *)grow(r->bases, r->cap_bases, m, r->pinned);
    char *q = (char *)grow(r->quals, r->cap_bases, m, r->pinned);
    if (b) r->bases = b;
    if (q) r->quals = q;
    if (!b || !q) return femhost_set_error(-3, "out of memory (%zu bases)", m);
    r->cap_bases = m;
  }
  if (n_names > r->cap_names) {
    size_t m = r->cap_names ? r->cap_names : (1u << 14);
    while (m < n_names) m *= 2;
    char *nm = (char *)grow(r->names, r->cap_names, m, r->pinned);
    if (!nm) return femhost_set_error(-3, "out of memory (%zu name bytes)", m);
    r->names = nm;
    r->cap_names = m;
  }
  return 0;
}

void femhost_records_free(femhost_records *r) {
  if (r->pinned) { /* names and qualities too: the device's SAM text reads them (femgpu_set_sam_inputs) */
    femgpu_free_pinned(r->bases); femgpu_free_pinned(r->base_off);
    femgpu_free_pinned(r->quals); femgpu_free_pinned(r->has_qual); femgpu_free_pinned(r->names); femgpu_free_pinned(r->name_off);
  } else {
    free(r->bases); free(r->base_off);
    free(r->quals); free(r->has_qual); free(r->names); free(r->name_off);
  }
  int pinned = r->pinned;
  memset(r, 0, sizeof *r);
  r->pinned = pinned;
}

/* ---- stream ------------------------------------------------------------------------------------- */
#define FX_BUF (4 << 20)
struct femhost_reader {
  gzFile f;
  unsigned char *buf;
  int begin, end, is_eof, err;
  int last_char;
};

femhost_reader *femhost_reader_open(const char *path) {
  gzFile f = gzopen(path, "r");
  if (!f) { femhost_set_error(-1, "Cannot find sequence file!"); return NULL; } /* src/sequence_batch.c:33 */
  gzbuffer(f, 1 << 20);
  femhost_reader *rd = (femhost_reader *)calloc(1, sizeof *rd);
  rd->f = f;
  rd->buf = (unsigned char *)malloc(FX_BUF);
  return rd;
}

void femhost_reader_close(femhost_reader *rd) {
  if (!rd) return;
  gzclose(rd->f);
  free(rd->buf);
  free(rd);
}

/* refill; returns 0 when no more data */
static inline int fx_fill(femhost_reader *rd) {
  if (rd->is_eof) return 0;
  rd->begin = 0;
  rd->end = gzread(rd->f, rd->buf, FX_BUF);
  if (rd->end <= 0) {
    if (rd->end < 0) rd->err = 1;
    rd->end = 0;
    rd->is_eof = 1;
    return 0;
  }
  return 1;
}
static inline int fx_getc(femhost_reader *rd) {
  if (rd->begin >= rd->end && !fx_fill(rd)) return rd->err ? -3 : -1;
  return rd->buf[rd->begin++];
}

/* Appends the rest of the current line to dst[*l ..] (dst = *pdst, grown through `reserve`);
 * drops a trailing '\r' when the whole string is longer than one char (src/kseq.h:146).
 * Returns -1 when nothing could be read because the stream is at its end. */
typedef int (*fx_reserve_fn)(void *ctx, size_t need);
static int fx_line(femhost_reader *rd, char **pdst, size_t base, size_t *l, fx_reserve_fn reserve, void *ctx) {
  int got = 0;
  for (;;) {
    if (rd->begin >= rd->end && !fx_fill(rd)) break;
    got = 1;
    const unsigned char *p = rd->buf + rd->begin;
    const int avail = rd->end - rd->begin;
    const unsigned char *nl = (const unsigned char *)memchr(p, '\n', (size_t)avail);
    const int n = nl ? (int)(nl - p) : avail;
    if (reserve(ctx, base + *l + (size_t)n + 1)) return -3;
    memcpy(*pdst + base + *l, p, (size_t)n);
    *l += (size_t)n;
    rd->begin += n + (nl ? 1 : 0);
    if (nl) break;
  }
  if (!got) return rd->err ? -3 : -1;
  if (*l > 1 && (*pdst)[base + *l - 1] == '\r') --*l;
  return 0;
}

/* name: text up to the first white-space char; *delim receives that char (0 at end of stream) */
static int fx_name(femhost_reader *rd, femhost_records *r, size_t base, size_t *l, int *delim) {
  int got = 0;
  *delim = 0;
  *l = 0;
  for (;;) {
    if (rd->begin >= rd->end && !fx_fill(rd)) break;
    got = 1;
    int i = rd->begin;
    while (i < rd->end && !isspace(rd->buf[i])) ++i;
    const int n = i - rd->begin;
    if (femhost_records_reserve(r, 0, 0, base + *l + (size_t)n + 1)) return -3;
    memcpy(r->names + base + *l, rd->buf + rd->begin, (size_t)n);
    *l += (size_t)n;
    rd->begin = i + 1;
    if (i < rd->end) { *delim = rd->buf[i]; break; }
  }
  if (!got) return rd->err ? -3 : -1;
  r->names[base + *l] = 0;
  return 0;
}

static int reserve_bases(void *ctx, size_t need) { return femhost_records_reserve((femhost_records *)ctx, 0, need, 0); }

/* One record appended as record number r->n_records (not yet counted).  Returns its sequence
 * length, -1 end of file, -2 truncated quality, -3 stream / memory error. */
/* Fast path for the common shape: a four-line FASTQ record that lies completely inside [p, end)
 * -- "@name[ comment]\n" "SEQ\n" "+...\n" "QUAL\n" with |QUAL| == |SEQ| > 0 and no '\r'.  For such a
 * record the general scanner below produces exactly these fields, so it is parsed with four
 * memchr calls and appended as record r->n_records (not counted yet).  Returns the first byte
 * after the record, or NULL when the bytes at p are not of that shape (nothing is consumed;
 * *oom is set when memory ran out instead). */
static const unsigned char *fastq4_record(const unsigned char *p, const unsigned char *end, femhost_records *r, int *oom) {
  if (p >= end || *p != '@') return NULL;
  const unsigned char *e1 = (const unsigned char *)memchr(p, '\n', (size_t)(end - p));
  if (!e1 || e1 + 1 >= end) return NULL;
  const unsigned char *s = e1 + 1;
  if (*s == '\n' || *s == '>' || *s == '@' || *s == '+') return NULL;
  const unsigned char *e2 = (const unsigned char *)memchr(s, '\n', (size_t)(end - s));
  if (!e2 || e2 + 1 >= end || e2[1] != '+' || e2[-1] == '\r') return NULL;
  const unsigned char *e3 = (const unsigned char *)memchr(e2 + 1, '\n', (size_t)(end - e2 - 1));
  if (!e3) return NULL;
  const size_t sl = (size_t)(e2 - s);
  const unsigned char *q = e3 + 1;
  if ((size_t)(end - q) < sl + 1 || q[sl] != '\n' || q[sl - 1] == '\r') return NULL;
  if (memchr(q, '\n', sl)) return NULL; /* a shorter first quality line: general path */
  const unsigned char *n0 = p + 1, *n1 = n0;
  while (n1 < e1 && !isspace(*n1)) ++n1;
  const size_t nl = (size_t)(n1 - n0);
  const uint32_t i = r->n_records;
  if (i == 0 && r->cap_records == 0 && femhost_records_reserve(r, 1, 16, 16)) { *oom = 1; return NULL; }
  const size_t nb = r->name_off[i], sb = r->base_off[i];
  if (femhost_records_reserve(r, (size_t)i + 2, sb + sl + 2, nb + nl + 2)) { *oom = 1; return NULL; }
  memcpy(r->names + nb, n0, nl);
  r->names[nb + nl] = 0;
  memcpy(r->bases + sb, s, sl);
  memcpy(r->quals + sb, q, sl);
  r->name_off[i + 1] = nb + nl + 1;
  r->base_off[i + 1] = sb + sl;
  r->has_qual[i] = 1;
  return q + sl + 1;
}

static long fx_record_fast(femhost_reader *rd, femhost_records *r) {
  int oom = 0;
  const unsigned char *next = fastq4_record(rd->buf + rd->begin, rd->buf + rd->end, r, &oom);
  if (!next) return oom ? -3 : -4;
  rd->begin = (int)(next - rd->buf);
  return (long)(r->base_off[r->n_records + 1] - r->base_off[r->n_records]);
}

/* Whole buffer of four-line records (the parallel reader of engine.c): replaces the content of
 * *out.  Returns 0 when [p, end) was consumed exactly, 1 when some record is not of the fast
 * shape (the caller falls back to the sequential reader), -3 out of memory. */
int femhost_parse_fastq4(const unsigned char *p, const unsigned char *end, femhost_records *out) {
  out->n_records = 0;
  out->n_bases = 0;
  if (femhost_records_reserve(out, 1, 16, 16)) return -3;
  int oom = 0;
  const unsigned char *first = p;
  while (p < end) {
    const unsigned char *next = fastq4_record(p, end, out, &oom);
    if (!next) return oom ? -3 : 1;
    ++out->n_records;
    if (p == first && next < end) {
      /* room for the whole buffer at once, from the first record's shape (reads of one run look alike): the arrays are
       * pinned, and growing them by doubling costs an allocation, a copy and a release of pinned memory each time */
      const size_t rec = (size_t)(next - first), est = (size_t)(end - first) / rec + 64;
      const size_t nb = (size_t)out->base_off[1] + 1, nn = (size_t)out->name_off[1] + 1;
      if (femhost_records_reserve(out, est, est * nb + est * nb / 32, est * nn + est * nn / 4)) return -3;
    }
    p = next;
  }
  out->n_bases = out->base_off[out->n_records];
  return 0;
}

/* First record start after the first line break of buf[0..n) (at 0 itself when at_file_start):
 * a line that begins with '@' whose second next line begins with '+'.  A quality line may begin
 * with '@', but then the second next line is a sequence line.  Returns the offset, n when the
 * buffer holds no further line (end of file), or -1 when no decision is possible. */
long femhost_fastq4_sync(const unsigned char *buf, size_t n, int at_file_start, int at_file_end) {
  const unsigned char *p = buf, *end = buf + n;
  if (!at_file_start) {
    p = (const unsigned char *)memchr(p, '\n', n);
    if (!p) return at_file_end ? (long)n : -1;
    ++p;
  }
  while (p < end) {
    const unsigned char *l1 = (const unsigned char *)memchr(p, '\n', (size_t)(end - p));
    if (!l1) return at_file_end ? (long)n : -1;
    if (*p == '@') {
      const unsigned char *l2 = l1 + 1 < end ? (const unsigned char *)memchr(l1 + 1, '\n', (size_t)(end - l1 - 1)) : NULL;
      if (!l2 || l2 + 1 >= end) return -1;
      if (l2[1] == '+') return (long)(p - buf);
    }
    p = l1 + 1;
  }
  return at_file_end ? (long)n : -1;
}

static long fx_record(femhost_reader *rd, femhost_records *r) {
  int c;
  if (rd->last_char == 0) {
    const long fast = fx_record_fast(rd, r);
    if (fast != -4) return fast;
  }
  if (rd->last_char == 0) { /* jump to the next header line */
    while ((c = fx_getc(rd)) >= 0 && c != '>' && c != '@') {}
    if (c < 0) return c;
    rd->last_char = c;
  }
  const uint32_t i = r->n_records;
  if (femhost_records_reserve(r, (size_t)i + 1, 0, 0)) return -3;
  const size_t nb = r->name_off[i], sb = r->base_off[i];
  size_t nl = 0, sl = 0, ql = 0;
  int rc = fx_name(rd, r, nb, &nl, &c);
  if (rc < 0) return rc;
  if (c != '\n') { /* comment: read and dropped */
    for (;;) {
      if (rd->begin >= rd->end && !fx_fill(rd)) break;
      const unsigned char *p = rd->buf + rd->begin;
      const unsigned char *e = (const unsigned char *)memchr(p, '\n', (size_t)(rd->end - rd->begin));
      if (e) { rd->begin += (int)(e - p) + 1; break; }
      rd->begin = rd->end;
    }
  }
  if (reserve_bases(r, sb + 2)) return -3;
  while ((c = fx_getc(rd)) >= 0 && c != '>' && c != '+' && c != '@') {
    if (c == '\n') continue;
    if (reserve_bases(r, sb + sl + 2)) return -3;
    r->bases[sb + sl++] = (char)c;
    if (fx_line(rd, &r->bases, sb, &sl, reserve_bases, r) == -3) return -3;
  }
  if (c == '>' || c == '@') rd->last_char = c;
  r->name_off[i + 1] = nb + nl + 1;
  r->base_off[i + 1] = sb + sl;
  r->has_qual[i] = 0;
  if (c != '+') return c == -3 ? -3 : (long)sl; /* FASTA */
  while ((c = fx_getc(rd)) >= 0 && c != '\n') {}
  if (c == -1) return -2;
  if (c == -3) return -3;
  /* quality lines land in r->quals at the same offsets; the reserve keeps both arrays in step */
  while ((rc = fx_line(rd, &r->quals, sb, &ql, reserve_bases, r)) >= 0 && ql < sl) {}
  if (rc == -3) return -3;
  rd->last_char = 0;
  if (sl != ql) return -2;
  r->has_qual[i] = ql != 0;
  return (long)sl;
}

int64_t femhost_reader_next(femhost_reader *rd, uint32_t max_records, femhost_records *out) {
  out->n_records = 0;
  out->n_bases = 0;
  if (femhost_records_reserve(out, 1, 16, 16)) return -3;
  while (out->n_records < max_records) {
    long l = fx_record(rd, out);
    if (l == 0) continue; /* src/sequence_batch.c:50-52 */
    if (l > 0) {
      ++out->n_records;
      out->n_bases += (uint64_t)l;
    } else {
      if (l != -1) /* src/sequence_batch.c:63-66 */
        return femhost_set_error((int)l, "Didn't reach the end of sequence file, which might be corrupted!");
      break;
    }
  }
  return out->n_records;
}

int femhost_load_all(const char *path, femhost_records *out) {
  femhost_reader *rd = femhost_reader_open(path);
  if (!rd) return -1;
  int64_t n = femhost_reader_next(rd, 0xFFFFFFF0u, out);
  femhost_reader_close(rd);
  return n < 0 ? (int)n : 0;
}

/* The parallel reader's decomposition, run serially (tests, and a reference for engine.c): the
 * file is cut at multiples of chunk_bytes, each range [sync(i*chunk), sync((i+1)*chunk)) is parsed
 * on its own with the four-line fast path and the ranges are concatenated.  Returns 0, 1 when the
 * file is not plain four-line FASTQ (callers use the sequential reader), negative on errors. */
int femhost_load_all_ranges(const char *path, uint64_t chunk_bytes, femhost_records *out) {
  FILE *f = fopen(path, "rb");
  if (!f) return femhost_set_error(-1, "Cannot find sequence file!");
  fseek(f, 0, SEEK_END);
  const uint64_t size = (uint64_t)ftell(f);
  fseek(f, 0, SEEK_SET);
  unsigned char *buf = (unsigned char *)malloc(size + 1);
  if (!buf || fread(buf, 1, size, f) != size) { fclose(f); free(buf); return femhost_set_error(-3, "cannot read %s", path); }
  fclose(f);
  out->n_records = 0;
  out->n_bases = 0;
  int rc = femhost_records_reserve(out, 1, 16, 16) ? -3 : 0;
  if (!rc && (size < 4 || buf[0] != '@')) rc = 1;
  if (chunk_bytes == 0) chunk_bytes = 1;
  femhost_records part;
  femhost_records_init(&part, 0);
  const size_t win = 1 << 18;
  for (uint64_t a = 0; !rc && a < size; a += chunk_bytes) {
    uint64_t lim[2];
    for (int j = 0; j < 2 && !rc; ++j) {
      const uint64_t at = a + (uint64_t)j * chunk_bytes;
      if (at == 0) lim[j] = 0;
      else if (at >= size) lim[j] = size;
      else {
        const size_t n = at + win > size ? (size_t)(size - at) : win;
        const long o = femhost_fastq4_sync(buf + at, n, 0, at + n == size);
        if (o < 0) rc = 1;
        else lim[j] = at + (uint64_t)o;
      }
    }
    if (rc) break;
    rc = femhost_parse_fastq4(buf + lim[0], buf + lim[1], &part);
    if (rc) break;
    const uint64_t nb = out->name_off[out->n_records], sb = out->base_off[out->n_records];
    if (femhost_records_reserve(out, (size_t)out->n_records + part.n_records + 1, sb + part.n_bases + 2,
                                nb + part.name_off[part.n_records] + 2)) { rc = -3; break; }
    memcpy(out->bases + sb, part.bases, part.n_bases);
    memcpy(out->quals + sb, part.quals, part.n_bases);
    memcpy(out->names + nb, part.names, part.name_off[part.n_records]);
    for (uint32_t i = 0; i < part.n_records; ++i) {
      out->base_off[out->n_records + i + 1] = sb + part.base_off[i + 1];
      out->name_off[out->n_records + i + 1] = nb + part.name_off[i + 1];
      out->has_qual[out->n_records + i] = 1;
    }
    out->n_records += part.n_records;
    out->n_bases += part.n_bases;
  }
  femhost_records_free(&part);
  free(buf);
  return rc;
}

/* One byte-range shard of a plain four-line FASTQ file: records that start in
 * [sync(size*i/n), sync((i+1)*size/n)), exactly the cut femhost_map_file makes for
 * femhost_map_options.shard_index / shard_count, so the shards concatenate to the file. */
int femhost_load_shard(const char *path, uint32_t shard_index, uint32_t shard_count, femhost_records *out) {
  if (shard_count == 0 || shard_index >= shard_count) return femhost_set_error(-1, "load: shard %u of %u", shard_index, shard_count);
  FILE *f = fopen(path, "rb");
  if (!f) return femhost_set_error(-1, "Cannot find sequence file!");
  fseeko(f, 0, SEEK_END);
  const uint64_t size = (uint64_t)ftello(f);
  out->n_records = 0;
  out->n_bases = 0;
  int rc = femhost_records_reserve(out, 1, 16, 16) ? -3 : 0;
  unsigned char head = 0;
  fseeko(f, 0, SEEK_SET);
  if (!rc && (size < 4 || fread(&head, 1, 1, f) != 1 || head != '@')) rc = 1;
  const size_t win = 1 << 18;
  unsigned char *w = (unsigned char *)malloc(win), *buf = NULL;
  uint64_t lim[2] = {0, 0};
  for (int j = 0; j < 2 && !rc; ++j) {
    const uint64_t at = (uint64_t)((__uint128_t)size * (shard_index + (uint32_t)j) / shard_count);
    if (at == 0) lim[j] = 0;
    else if (at >= size) lim[j] = size;
    else {
      const size_t n = at + win > size ? (size_t)(size - at) : win;
      if (fseeko(f, (off_t)at, SEEK_SET) != 0 || fread(w, 1, n, f) != n) { rc = femhost_set_error(-3, "cannot read %s", path); break; }
      const long o = femhost_fastq4_sync(w, n, 0, at + n == size);
      if (o < 0) rc = 1;
      else lim[j] = at + (uint64_t)o;
    }
  }
  if (!rc && lim[1] > lim[0]) {
    const size_t n = (size_t)(lim[1] - lim[0]);
    buf = (unsigned char *)malloc(n + 1);
    if (!buf || fseeko(f, (off_t)lim[0], SEEK_SET) != 0 || fread(buf, 1, n, f) != n) rc = femhost_set_error(-3, "cannot read %s", path);
    else rc = femhost_parse_fastq4(buf, buf + n, out);
  }
  free(buf);
  free(w);
  fclose(f);
  return rc;
}
