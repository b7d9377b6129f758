// How fast can N threads fill a new tmpfs file?  Variants of the copy into the page cache.
// usage: shm_write <MB> <threads> <mode> [chunk KB]
//   mode 0: memcpy into a shared mapping (one fault per page)     1: MADV_POPULATE_WRITE per chunk, then memcpy
//   mode 2: fallocate the whole file first (timed), then memcpy   3: pwrite per chunk
//   mode 4: one private file per thread (memcpy into its mapping): is the limit per file or global?
#define _GNU_SOURCE
#include <fcntl.h>
#include <pthread.h>
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
static size_t total, chunk; static int n_thr, mode, fd; static char *map, *src; static volatile size_t next_chunk;
static void *run(void *arg) {
  long id = (long)arg; char *m = map; int f = fd;
  if (mode == 4) { char p[64]; snprintf(p, sizeof p, "/dev/shm/shm_write_%ld", id); f = open(p, O_RDWR | O_CREAT | O_TRUNC, 0600);
    if (ftruncate(f, total)) return NULL; m = mmap(NULL, total, PROT_READ | PROT_WRITE, MAP_SHARED, f, 0); }
  for (;;) {
    size_t c = __sync_fetch_and_add(&next_chunk, 1), off = c * chunk;
    if (off >= total) break;
    size_t n = off + chunk > total ? total - off : chunk;
    if (mode == 3) { if (pwrite(f, src + (off % (64u << 20)), n, off) != (ssize_t)n) perror("pwrite"); continue; }
    if (mode == 1) madvise(m + off, n, MADV_POPULATE_WRITE);
    memcpy(m + off, src + (off % (64u << 20)), n);
  }
  if (mode == 4) { char p[64]; snprintf(p, sizeof p, "/dev/shm/shm_write_%ld", id); munmap(m, total); close(f); unlink(p); }
  return NULL;
}
int main(int argc, char **argv) {
  total = (size_t)atol(argv[1]) << 20; n_thr = atoi(argv[2]); mode = atoi(argv[3]); chunk = (argc > 4 ? atol(argv[4]) : 1024) << 10;
  src = malloc((64u << 20) + chunk); memset(src, 'y', (64u << 20) + chunk);
  unlink("/dev/shm/shm_write_f");
  fd = open("/dev/shm/shm_write_f", O_RDWR | O_CREAT | O_TRUNC, 0600);
  double t0 = now();
  if (ftruncate(fd, total)) return 1;
  if (mode == 2 && fallocate(fd, 0, 0, total)) perror("fallocate");
  double t1 = now();
  map = mmap(NULL, total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  pthread_t th[64];
  for (long i = 0; i < n_thr; ++i) pthread_create(&th[i], NULL, run, (void *)i);
  for (int i = 0; i < n_thr; ++i) pthread_join(th[i], NULL);
  double t2 = now();
  printf("mode %d threads %2d chunk %5zu KB: prepare %.1f ms, fill %.1f ms = %.2f GB/s\n", mode, n_thr, chunk >> 10, 1e3 * (t1 - t0), 1e3 * (t2 - t1), total / (t2 - t1) / 1e9);
  munmap(map, total); close(fd); unlink("/dev/shm/shm_write_f");
  return 0;
}
