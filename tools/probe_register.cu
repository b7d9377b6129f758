// probe_register.cu -- can the SAM text's device-to-host copy land in the output file's pages directly?
// (VERDICT r1 #6a).  Times, on a tmpfs file: fallocate, mapping + MADV_POPULATE_WRITE, cudaHostRegister of the
// mapped span in pieces (one and two threads), the D2H copy into it, cudaHostUnregister; beside it the copy into
// ordinary pinned memory followed by CPU copies into the mapping (what the engine does today).
// nvcc -O2 -o /tmp/probe_register tools/probe_register.cu -lpthread ; /tmp/probe_register [GB] [span MB]
#ifndef _GNU_SOURCE
#define _GNU_SOURCE
#endif
#include <cuda_runtime.h>
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
static double now() { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
struct reg_job { char *p; size_t n, span; int rc; double s; };
static void *reg_main(void *a) {
  reg_job *j = (reg_job *)a;
  const double t0 = now();
  for (size_t o = 0; o < j->n; o += j->span)
    if (cudaHostRegister(j->p + o, j->span < j->n - o ? j->span : j->n - o, cudaHostRegisterDefault) != cudaSuccess) { j->rc = 1; break; }
  j->s = now() - t0;
  return NULL;
}
struct cp_job { char *d; const char *s; size_t n; };
static void *cp_main(void *a) { cp_job *j = (cp_job *)a; memcpy(j->d, j->s, j->n); return NULL; }
int main(int argc, char **argv) {
  const size_t S = (size_t)(argc > 1 ? atof(argv[1]) : 2.0) * (1ull << 30), span = (size_t)(argc > 2 ? atoi(argv[2]) : 256) << 20;
  CK(cudaSetDevice(0));
  char *dev; CK(cudaMalloc(&dev, S)); CK(cudaMemset(dev, 0x41, S));
  char *pin; CK(cudaMallocHost(&pin, S));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float ms;
  CK(cudaEventRecord(e0)); CK(cudaMemcpyAsync(pin, dev, S, cudaMemcpyDeviceToHost)); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
  CK(cudaEventElapsedTime(&ms, e0, e1)); printf("D2H into cudaMallocHost memory: %.1f GB/s\n", S / ms / 1e6);
  for (int threads = 1; threads <= 4; threads *= 2) {
    const char *path = "/dev/shm/probe_register.bin";
    unlink(path);
    const int fd = open(path, O_RDWR | O_CREAT | O_TRUNC, 0644);
    if (fd < 0 || ftruncate(fd, (off_t)S) != 0) { printf("no tmpfs file\n"); return 1; }
    double t0 = now();
    const int fa = fallocate(fd, 0, 0, (off_t)S);
    printf("[%d thread(s)] fallocate rc %d: %.3f s = %.1f GB/s\n", threads, fa, now() - t0, S / (now() - t0) / 1e9);
    t0 = now();
    char *mp = (char *)mmap(NULL, S, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (mp == MAP_FAILED) { printf("mmap failed\n"); return 1; }
    const int populate = argc > 3 ? atoi(argv[3]) : 1;
    const int mv = populate ? madvise(mp, S, MADV_POPULATE_WRITE) : -1;
    printf("  mmap + MADV_POPULATE_WRITE (%s) rc %d: %.3f s = %.1f GB/s\n", populate ? "done" : "skipped", mv, now() - t0, S / (now() - t0) / 1e9);
    // registration, `threads` threads on disjoint parts
    pthread_t th[4]; reg_job rj[4];
    t0 = now();
    for (int i = 0; i < threads; ++i) {
      const size_t part = (S / threads) & ~(span - 1);
      rj[i].p = mp + part * i; rj[i].n = i == threads - 1 ? S - part * i : part; rj[i].span = span; rj[i].rc = 0;
      pthread_create(&th[i], NULL, reg_main, &rj[i]);
    }
    int bad = 0;
    for (int i = 0; i < threads; ++i) { pthread_join(th[i], NULL); bad |= rj[i].rc; }
    const double treg = now() - t0;
    printf("  cudaHostRegister in %zu MB spans, %d thread(s): rc %d, %.3f s = %.1f GB/s\n", span >> 20, threads, bad, treg, S / treg / 1e9);
    if (bad) { printf("  registration failed: %s\n", cudaGetErrorString(cudaGetLastError())); }
    else {
      CK(cudaEventRecord(e0));
      for (int i = 0; i < threads; ++i) // a copy may not span two registrations
        for (size_t o = 0; o < rj[i].n; o += span) {
          const size_t off = (size_t)(rj[i].p - mp) + o, n = span < rj[i].n - o ? span : rj[i].n - o;
          CK(cudaMemcpyAsync(mp + off, dev + off, n, cudaMemcpyDeviceToHost));
        }
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
      CK(cudaEventElapsedTime(&ms, e0, e1)); printf("  D2H into the registered file mapping (one copy per span): %.1f GB/s\n", S / ms / 1e6);
      t0 = now();
      for (int i = 0; i < threads; ++i)
        for (size_t o = 0; o < rj[i].n; o += span) cudaHostUnregister(rj[i].p + o);
      printf("  cudaHostUnregister: %.3f s = %.1f GB/s\n", now() - t0, S / (now() - t0) / 1e9);
      printf("  file byte 0 / last: %d %d\n", mp[0], mp[S - 1]);
    }
    // today's form: CPU copies from pinned memory into the (allocated, populated) mapping
    for (int ct = 4; ct <= 8; ct += 4) {
      pthread_t ch[8]; cp_job cj[8];
      t0 = now();
      for (int i = 0; i < ct; ++i) { cj[i].d = mp + S / ct * i; cj[i].s = pin + S / ct * i; cj[i].n = S / ct; pthread_create(&ch[i], NULL, cp_main, &cj[i]); }
      for (int i = 0; i < ct; ++i) pthread_join(ch[i], NULL);
      printf("  %d CPU threads, pinned -> populated mapping: %.1f GB/s\n", ct, S / (now() - t0) / 1e9);
    }
    t0 = now();
    munmap(mp, S); close(fd); unlink(path);
    printf("  munmap + unlink: %.3f s\n", now() - t0);
  }
  return 0;
}
