#!/bin/bash
gcc -O2 -pthread -o /tmp/shm_write tools/micro/shm_write.c 2>/dev/null
cat /sys/kernel/mm/transparent_hugepage/shmem_enabled; df -h /dev/shm | tail -1; grep -c processor /proc/cpuinfo; grep "model name" /proc/cpuinfo | head -1
for t in 1 2 4 6 8 16; do for m in 0 1 2 3; do /tmp/shm_write 276 $t $m; done; done
for t in 2 4 8; do /tmp/shm_write 276 $t 4; done
/tmp/shm_write 276 6 1 256; /tmp/shm_write 276 6 1 4096; /tmp/shm_write 276 6 3 256; /tmp/shm_write 276 6 3 4096; /tmp/shm_write 276 6 0 64
