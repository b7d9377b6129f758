/* probe_tmpfs_write.c -- how fast can T threads put S bytes of text into ONE tmpfs file?  (the end-to-end bound of
 * `FEM map` on a B200: 3.9 GB of SAM text per 10 M cfg3 reads).  No GPU needed.
 * Forms: A one shared mapping, page faults from the copying threads (the engine up to round 2)
 *        B one mapping per chunk with MAP_POPULATE, made by the thread that copies the chunk
 *        C one mapping per chunk, no populate
 *        D pwrite per chunk
 *        E one shared mapping, madvise(MADV_POPULATE_WRITE) on the chunk by the copying thread, then the copy
 *        F as the engine: one mapping per BATCH (16 MB) made by a coordinating thread, chunks of 768 KB populated and
 *          copied by the workers, the batch unmapped by the worker that finishes it
 *        G as F, with the pages allocated by a concurrent fallocate thread (4 MB steps) instead of up front
 *        H as E (one mapping, populate per chunk) with the concurrent fallocate thread of G
 *        I as A (one mapping, page faults) with the concurrent fallocate thread
 *        J as E on the first half of the file (allocated up front) while a fallocate thread allocates the second half:
 *          do allocator and copiers disturb each other when they are far apart in the file?
 * The file is fallocate'd first in forms A-F (timed separately).
 * gcc -O2 -o /tmp/probe_tmpfs_write tools/probe_tmpfs_write.c -lpthread ; /tmp/probe_tmpfs_write [GB] [threads] [chunk MB] */
#define _GNU_SOURCE
#include <fcntl.h>
#include <pthread.h>
#include <sched.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <time.h>
#include <unistd.h>
#ifndef MADV_POPULATE_WRITE
#define MADV_POPULATE_WRITE 23
#endif
static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }
static size_t S, CH;
static int fd, form, T;
static char *src, *shared_map;
static volatile size_t next_chunk;
#define BATCH (16u << 20)
#define FCHUNK (768u << 10)
typedef struct { char *map; size_t off, len; volatile int left; } fbatch;
static fbatch *fb;
static size_t n_fb;
static volatile size_t fb_ready; /* batches mapped so far */
static void *coordinator(void *a) {
  (void)a;
  for (size_t i = 0; i < n_fb; ++i) {
    fb[i].map = mmap(NULL, fb[i].len, PROT_READ | PROT_WRITE, MAP_SHARED, fd, (off_t)fb[i].off);
    if (fb[i].map == MAP_FAILED) { perror("mmap"); exit(1); }
    __sync_synchronize();
    fb_ready = i + 1;
  }
  return NULL;
}
static size_t falloc_from;
static double falloc_seconds;
static void *falloc_thread(void *a) {
  (void)a;
  const double t0 = now();
  for (size_t o = falloc_from; o < S; o += 4u << 20) if (fallocate(fd, FALLOC_FL_KEEP_SIZE, (off_t)o, 4 << 20)) { perror("fallocate"); break; }
  falloc_seconds = now() - t0;
  return NULL;
}
static void *work_f(void *a) {
  (void)a;
  const size_t per = BATCH / FCHUNK + (BATCH % FCHUNK != 0);
  for (;;) {
    const size_t c = __sync_fetch_and_add(&next_chunk, 1), bi = c / per, ci = c % per;
    if (bi >= n_fb) break;
    while (fb_ready <= bi) sched_yield();
    fbatch *b = &fb[bi];
    const size_t o = ci * FCHUNK;
    if (o < b->len) {
      const size_t n = FCHUNK < b->len - o ? FCHUNK : b->len - o;
      madvise(b->map + o, n, MADV_POPULATE_WRITE);
      memcpy(b->map + o, src + b->off + o, n);
    }
    if (__sync_sub_and_fetch(&b->left, 1) == 0) munmap(b->map, b->len);
  }
  return NULL;
}
static void *work(void *a) {
  (void)a;
  for (;;) {
    const size_t c = __sync_fetch_and_add(&next_chunk, 1), off = c * CH;
    if (off >= (form == 'J' ? S / 2 : S)) break;
    const size_t n = CH < S - off ? CH : S - off;
    if (form == 'A' || form == 'I') memcpy(shared_map + off, src + off, n);
    else if (form == 'B' || form == 'C') {
      char *m = mmap(NULL, n, PROT_READ | PROT_WRITE, MAP_SHARED | (form == 'B' ? MAP_POPULATE : 0), fd, (off_t)off);
      if (m == MAP_FAILED) { perror("mmap"); exit(1); }
      memcpy(m, src + off, n);
      munmap(m, n);
    } else if (form == 'D') {
      size_t done = 0;
      while (done < n) { const ssize_t w = pwrite(fd, src + off + done, n - done, (off_t)(off + done)); if (w <= 0) { perror("pwrite"); exit(1); } done += (size_t)w; }
    } else if (form == 'E' || form == 'H' || form == 'J') {
      madvise(shared_map + off, n, MADV_POPULATE_WRITE);
      memcpy(shared_map + off, src + off, n);
    }
  }
  return NULL;
}
int main(int argc, char **argv) {
  S = (size_t)((argc > 1 ? atof(argv[1]) : 2.0) * (double)(1ull << 30));
  T = argc > 2 ? atoi(argv[2]) : 6;
  CH = (size_t)(argc > 3 ? atoi(argv[3]) : 4) << 20;
  src = malloc(S);
  memset(src, 'x', S);
  const char *forms = argc > 4 ? argv[4] : "ABCDEFGHI";
  for (const char *f = forms; *f; ++f) {
    form = *f;
    const char *path = "/dev/shm/probe_tmpfs_write.bin";
    unlink(path);
    fd = open(path, O_RDWR | O_CREAT | O_TRUNC, 0644);
    if (fd < 0 || ftruncate(fd, (off_t)S)) { perror("file"); return 1; }
    double t0 = now();
    const int concurrent = form == 'G' || form == 'H' || form == 'I' || form == 'J', batched = form == 'F' || form == 'G';
    if (!concurrent && fallocate(fd, 0, 0, (off_t)S)) perror("fallocate");
    falloc_from = 0;
    if (form == 'J') { if (fallocate(fd, 0, 0, (off_t)(S / 2))) perror("fallocate"); falloc_from = S / 2; }
    const double tf = now() - t0 + 1e-9;
    shared_map = NULL;
    if (form == 'A' || form == 'E' || form == 'H' || form == 'I' || form == 'J') shared_map = mmap(NULL, S, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    next_chunk = 0;
    pthread_t th[64], tc, tfa;
    if (batched) {
      n_fb = (S + BATCH - 1) / BATCH;
      fb = calloc(n_fb, sizeof *fb);
      const size_t per = BATCH / FCHUNK + (BATCH % FCHUNK != 0);
      for (size_t i = 0; i < n_fb; ++i) { fb[i].off = i * BATCH; fb[i].len = BATCH < S - fb[i].off ? BATCH : S - fb[i].off; fb[i].left = (int)per; }
      fb_ready = 0;
    }
    t0 = now();
    if (concurrent) pthread_create(&tfa, NULL, falloc_thread, NULL);
    if (batched) pthread_create(&tc, NULL, coordinator, NULL);
    for (int i = 0; i < T; ++i) pthread_create(&th[i], NULL, batched ? work_f : work, NULL);
    for (int i = 0; i < T; ++i) pthread_join(th[i], NULL);
    const double tcopy = now() - t0;
    if (batched) { pthread_join(tc, NULL); free(fb); }
    if (concurrent) pthread_join(tfa, NULL);
    const double tw = now() - t0;
    t0 = now();
    if (shared_map) munmap(shared_map, S);
    const double tu = now() - t0;
    printf("form %c, %d threads, %zu MB chunks: fallocate %.1f GB/s, copies %.3f s = %.1f GB/s, munmap %.3f s\n", form, T, CH >> 20, S / tf / 1e9, tw, S / tw / 1e9, tu);
    if (concurrent) printf("        copiers alone %.3f s for %.1f GB = %.1f GB/s; concurrent fallocate %.3f s for %.1f GB = %.1f GB/s\n", tcopy, (form == 'J' ? S / 2 : S) / 1e9,
                           (form == 'J' ? S / 2 : S) / tcopy / 1e9, falloc_seconds, (S - falloc_from) / 1e9, (S - falloc_from) / falloc_seconds / 1e9);
    close(fd);
    unlink(path);
  }
  return 0;
}
