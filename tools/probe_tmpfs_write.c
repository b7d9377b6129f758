/* probe_tmpfs_write.c -- how fast can T threads put S bytes of text into ONE tmpfs file?  (the end-to-end bound of
 * `FEM map` on a B200: 3.9 GB of SAM text per 10 M cfg3 reads).  No GPU needed.
 * Forms: A one shared mapping, page faults from the copying threads (the engine up to round 2)
 *        B one mapping per chunk with MAP_POPULATE, made by the thread that copies the chunk
 *        C one mapping per chunk, no populate
 *        D pwrite per chunk
 *        E one shared mapping, madvise(MADV_POPULATE_WRITE) on the chunk by the copying thread, then the copy
 * The file is fallocate'd first in every form (timed separately).
 * gcc -O2 -o /tmp/probe_tmpfs_write tools/probe_tmpfs_write.c -lpthread ; /tmp/probe_tmpfs_write [GB] [threads] [chunk MB] */
#define _GNU_SOURCE
#include <fcntl.h>
#include <pthread.h>
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
static void *work(void *a) {
  (void)a;
  for (;;) {
    const size_t c = __sync_fetch_and_add(&next_chunk, 1), off = c * CH;
    if (off >= S) break;
    const size_t n = CH < S - off ? CH : S - off;
    if (form == 'A') memcpy(shared_map + off, src + off, n);
    else if (form == 'B' || form == 'C') {
      char *m = mmap(NULL, n, PROT_READ | PROT_WRITE, MAP_SHARED | (form == 'B' ? MAP_POPULATE : 0), fd, (off_t)off);
      if (m == MAP_FAILED) { perror("mmap"); exit(1); }
      memcpy(m, src + off, n);
      munmap(m, n);
    } else if (form == 'D') {
      size_t done = 0;
      while (done < n) { const ssize_t w = pwrite(fd, src + off + done, n - done, (off_t)(off + done)); if (w <= 0) { perror("pwrite"); exit(1); } done += (size_t)w; }
    } else if (form == 'E') {
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
  const char *forms = argc > 4 ? argv[4] : "ABCDE";
  for (const char *f = forms; *f; ++f) {
    form = *f;
    const char *path = "/dev/shm/probe_tmpfs_write.bin";
    unlink(path);
    fd = open(path, O_RDWR | O_CREAT | O_TRUNC, 0644);
    if (fd < 0 || ftruncate(fd, (off_t)S)) { perror("file"); return 1; }
    double t0 = now();
    if (fallocate(fd, 0, 0, (off_t)S)) perror("fallocate");
    const double tf = now() - t0;
    shared_map = NULL;
    if (form == 'A' || form == 'E') shared_map = mmap(NULL, S, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    next_chunk = 0;
    pthread_t th[64];
    t0 = now();
    for (int i = 0; i < T; ++i) pthread_create(&th[i], NULL, work, NULL);
    for (int i = 0; i < T; ++i) pthread_join(th[i], NULL);
    const double tw = now() - t0;
    t0 = now();
    if (shared_map) munmap(shared_map, S);
    const double tu = now() - t0;
    printf("form %c, %d threads, %zu MB chunks: fallocate %.1f GB/s, copies %.3f s = %.1f GB/s, munmap %.3f s\n", form, T, CH >> 20, S / tf / 1e9, tw, S / tw / 1e9, tu);
    close(fd);
    unlink(path);
  }
  return 0;
}
