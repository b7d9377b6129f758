#!/usr/bin/env python
"""bench.py -- reads/sec of the `FEM map` hot path on B200 (BASELINE.json's metric).

  python bench.py --gpus N --steps K --warmup W            # this repository's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU code

Workload = BASELINE.json configs[2] (cfg3), the configuration the metric and the north_star targets name:
synthetic 3.1 Gbp reference (24 records), 10 M simulated 150 bp single-end reads, e=7, index window 12 /
step 5.  It fits one GPU (5.0 GB occurrence table + 0.8 GB packed reference).  STRONG scaling: ONE read
file of 10 M reads whatever N is; under torchrun rank r maps byte-range shard r of N of that file
(femhost_map_options.shard_index / shard_count) against its replica of the index -- no collective on the
data path.  A "step" is one pass of the hot path over the whole read set.  --workload cfg1|cfg2|cfg4p|cfg5
selects the other BASELINE.json configs (parity-test cases; their lines go under "other_workloads").

  value  reads/s with the reads already resident in HBM (femgpu_rerun_resident: seed -> candidates ->
         verify -> compaction -> traceback -> SAM text), CUDA-event time summed over the K steps, max over ranks
  e2e    reads/s of the whole `FEM map` run through the C ABI (femhost_map_file): FASTQ parse from the page
         cache, pinned H2D, kernels, D2H, SAM text, file written to tmpfs -- the same window as the reference's
         own "Time:" line (src/FEM_map.c:172-219).  N > 1 measures both product forms: one process per GPU
         (rank r writes shard r; the shard files in rank order are the SAM file) and ONE process driving N GPUs
         with one reader pool and one ordered writer (`FEM map --gpus N`, run by rank 0 while the other ranks
         wait on a CPU barrier); e2e.value is the faster, the other stands beside it, and the outputs are
         compared by sha256 with the single-GPU file.
  roofline      k_cand (dominant) and k_seed: algorithmic bytes / CUDA-event time vs MEASURED_PEAKS.json
  cpu_baseline  oracle/_ref/FEM_ref_hdr (the reference compiled from /root/reference, plus the
                two-line index-header patch) with -t <host cores> on a bounded sample
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[0..4]; cfg4 as written is infeasible in the reference and is restated as cfg4p
    "cfg1": dict(name="cfg1: 4.6 Mbp random reference (4 records), 100k x 100 bp SE reads, e=3, window 12 / step 5",
                 ref_bp=4_600_000, ref_records=4, repeat_permille=0, reads=100_000, read_len=100, e=3, a=1, k=12, step=5, seed=1),
    "cfg2": dict(name="cfg2: 100 Mbp random reference (10 records), 1M x 100 bp SE reads, e=5, window 12 / step 4",
                 ref_bp=100_000_000, ref_records=10, repeat_permille=0, reads=1_000_000, read_len=100, e=5, a=1, k=12, step=4, seed=2),
    "cfg3": dict(name="cfg3: 3.1 Gbp random reference (24 records), 10M x 150 bp SE reads, e=7, window 12 / step 5",
                 ref_bp=3_100_000_000, ref_records=24, repeat_permille=0, reads=10_000_000, read_len=150, e=7, a=1, k=12, step=5, seed=3),
    "cfg4p": dict(name="cfg4': 3.1 Gbp repeat-rich reference (10% repeats, 24 records), 10M x 150 bp SE reads, e=7, window 12 / step 2",
                  ref_bp=3_100_000_000, ref_records=24, repeat_permille=100, reads=10_000_000, read_len=150, e=7, a=1, k=12, step=2, seed=4),
    "cfg5": dict(name="cfg5: 3.1 Gbp random reference (24 records), 50M x 250 bp SE reads, e=7, window 14 / step 10",
                 ref_bp=3_100_000_000, ref_records=24, repeat_permille=0, reads=50_000_000, read_len=250, e=7, a=1, k=14, step=10, seed=3),
}
DEFAULT_WORKLOAD = "cfg3"  # the configuration BASELINE.json's metric and the north_star targets are quoted on
WORKLOAD = dict(WORKLOADS[DEFAULT_WORKLOAD])
OTHER_WORKLOADS = (("cfg2", 1_000_000), ("cfg5", 1_000_000), ("cfg4p", 1_000_000), ("cfg1", 100_000))  # (name, reads) under "other_workloads"
FEMSIM = os.path.join(ROOT, "tools", "femsim")
REF_HDR = os.path.join(ROOT, "oracle", "_ref", "FEM_ref_hdr")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--reads", type=int, default=0, help="reads of the read file, shared by all GPUs (default: the workload's)")
    ap.add_argument("--others", default="all", help="comma-separated workloads measured in sub-runs of this script and reported under "
                    "other_workloads (N=1 and the default workload only; 'all' = cfg2,cfg5,cfg4p at 1M reads, cfg1 at 0.1M; 'none')")
    ap.add_argument("--no-single-process", action="store_true", help="N > 1: skip the one-process-N-GPUs end-to-end leg")
    ap.add_argument("--ref-bp", type=int, default=0)
    ap.add_argument("--data-dir", default=os.environ.get("FEM_BENCH_DIR", "/tmp/fem_bench"))
    ap.add_argument("--cpu-sample", type=int, default=200_000, help="reads of the workload the CPU baseline maps")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-threads", type=int, default=0, help="-t of the reference's CPU runs (default: all host cores; BASELINE.json configs[0] names -t 8)")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the end-to-end leg (default: --steps)")
    ap.add_argument("--e2e-threads", type=int, default=0, help="host threads of the end-to-end leg (default: cores/ranks - 3)")
    ap.add_argument("--e2e-batch", type=int, default=0, help="reads per GPU batch of the end-to-end leg (default: the engine's)")
    args = ap.parse_args()
    WORKLOAD.clear()
    WORKLOAD.update(WORKLOADS[args.workload])
    args.reads = args.reads or WORKLOAD["reads"]
    args.ref_bp = args.ref_bp or WORKLOAD["ref_bp"]
    return args


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


def ensure_data(args, rank, world, barrier):
    """Synthetic inputs (untimed): reference + index + ONE read file, generated by rank 0, shared by all ranks."""
    from fem_b200 import host
    w = WORKLOAD
    # cfg3 and cfg5 share one 3.1 Gbp reference (same seed); index and reads are per workload
    d = os.path.join(args.data_dir, "ref%d_rec%d_rep%d_seed%d" % (args.ref_bp, w["ref_records"], w["repeat_permille"], w["seed"]))
    fa, idx = os.path.join(d, "ref.fa"), os.path.join(d, "index_k%d_s%d" % (w["k"], w["step"]))
    fq = os.path.join(d, "reads_%d_l%d_e%d.fq" % (args.reads, w["read_len"], w["e"]))
    if rank == 0:
        os.makedirs(d, exist_ok=True)
        if not os.path.exists(fa + ".ok"):
            t = time.time()
            subprocess.check_call([FEMSIM, "ref", str(w["seed"]), str(args.ref_bp), str(w["ref_records"]), str(w["repeat_permille"]), fa])
            open(fa + ".ok", "w").close()
            log("reference generated in %.1fs" % (time.time() - t))
        if not os.path.exists(idx + ".ok"):
            t = time.time()
            subprocess.check_call([host.FEM_BINARY, "index", str(w["k"]), str(w["step"]), fa, idx, "--gpu"], stderr=subprocess.DEVNULL)
            open(idx + ".ok", "w").close()
            log("index built in %.1fs" % (time.time() - t))
        if not os.path.exists(fq + ".ok"):
            t = time.time()
            subprocess.check_call([FEMSIM, "reads", str(1000 + w["seed"]), fa, str(args.reads), str(w["read_len"]), str(w["e"]), fq])
            open(fq + ".ok", "w").close()
            log("%d reads simulated in %.1fs" % (args.reads, time.time() - t))
    barrier()
    return fa, idx, fq


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.proc, self.lines = gpu_index, None, []
        self.nvml, self.samples, self.halt = None, [], False

    def start(self):
        # NVML polled every 2 ms from a thread: the timed region of the resident leg is tens of milliseconds,
        # shorter than nvidia-smi's start-up; nvidia-smi -lms stays as the fall-back
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            idx = int(vis.split(",")[self.gpu]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else self.gpu
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _poll(self):
        n = self.nvml
        while not self.halt:
            try:
                sm = float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM))
                try:
                    why = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    why = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                try:
                    power = n.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                except Exception:
                    power = 0.0
                self.samples.append((sm, why, power))
            except Exception:
                pass
            time.sleep(0.002)

    def _stop_nvml(self):
        self.halt = True
        self.t.join(timeout=2)
        bits = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40, "sw_power_cap": 0x4}
        sm = sorted(x[0] for x in self.samples)
        reasons = sorted(k for k, b in bits.items() if any(x[1] & b for x in self.samples))
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(sm), "power_w_max": max((x[2] for x in self.samples), default=0.0), "source": "nvml polled with 2 ms pauses, from the start of the resident leg to the end of the end-to-end leg"}

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.nvml:
            return self._stop_nvml()
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.t.join(timeout=2)
        sm, mx, reasons, power = [], 0, set(), 0.0
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx = max(mx, float(f[2])); power = max(power, float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": power}


def head_fastq(src, dst, n_reads):
    """first n_reads records of a 4-line FASTQ (the bounded CPU sample)"""
    if os.path.exists(dst + ".ok"):
        return dst
    with open(src, "rb") as f, open(dst, "wb") as g:
        for _ in range(4 * n_reads):
            line = f.readline()
            if not line:
                break
            g.write(line)
    open(dst + ".ok", "w").close()
    return dst


def time_reference(fa, idx, fq, e, a, threads, out_sam):
    """One run of the reference's own CPU `FEM map`; returns (seconds of its "Time:" window, counters)."""
    p = subprocess.run([REF_HDR, "map", "-e", str(e), "-a", str(a), "-t", str(threads), "--ref", fa, "--index", idx,
                        "--read1", fq, "-o", out_sam], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    if p.returncode != 0:
        raise RuntimeError("reference FEM map failed: " + p.stderr[-500:])
    secs, reads = None, None
    for line in p.stderr.splitlines():
        if line.startswith("Time:"):
            secs = float(line.split()[1].rstrip("s"))
        elif line.startswith("The number of read:"):
            reads = int(line.split(":")[1])
    return secs, reads


def sam_groups(path):
    """SAM records grouped by QNAME in file order (reference -t N interleaves reads but keeps each read's
    records contiguous and ordered); header lines under the key None."""
    groups = {}
    with open(path, "rb") as f:
        for line in f:
            key = None if line.startswith(b"@") else line.split(b"\t", 1)[0]
            groups.setdefault(key, []).append(line)
    return groups


def cpu_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def config_dict(args, world):
    w = WORKLOAD
    return {"workload": w["name"] if args.reads == w["reads"] else w["name"] + " -- run with %d reads" % args.reads,
            "reads": args.reads, "reads_per_gpu": args.reads / world, "read_len": w["read_len"], "e": w["e"], "a": w["a"],
            "window": w["k"], "step": w["step"], "reference_bp": args.ref_bp,
            "sharding": "one read file; rank r maps byte-range shard r of %d; index replicated per GPU; no collective on the data path" % world,
            "l2": "inputs larger than L2 (index %.0f MB + reads, re-streamed every step)" % (4.0 * (4 ** w["k"] + 1) / 1e6 + 8.0 * args.ref_bp / w["step"] / 1e6)}


def scaling_kind(args):
    return "strong"  # the read set is fixed (BASELINE.json configs[2] and [4] are fixed read sets split over 1/2/4/8 GPUs)


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return 0
    w = WORKLOAD
    if not os.access(REF_HDR, os.X_OK):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/FEM_ref_hdr was not built (no /root/reference at build time)"}))
        return 0
    fa, idx, fq = ensure_data(args, 0, 1, lambda: None)
    sample = head_fastq(fq, fq + ".cpu%d" % args.cpu_sample, args.cpu_sample)
    cores = args.cpu_threads or cpu_cores()
    out = "/dev/shm/fem_bench_ref.sam" if os.path.isdir("/dev/shm") else os.path.join(args.data_dir, "ref.sam")
    times = []
    for i in range(args.warmup + args.steps):
        secs, n = time_reference(fa, idx, sample, w["e"], w["a"], cores, out)
        if i >= args.warmup:
            times.append(secs)
    try:
        os.remove(out)
    except OSError:
        pass
    total = sum(times)
    value = n * len(times) / total
    line = {"impl": "reference", "metric": "reads/sec (all locations)", "value": value, "unit": "reads/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True,
            "scaling": scaling_kind(args), "vs_baseline": None, "dtype": "u32/u64 integer", "data": "synthetic", "config": config_dict(args, world),
            "cpu_baseline": {"value": value, "unit": "reads/s", "cores": cores, "kind": "reference",
                             "sample": "first %d reads of the workload per step, FEM_ref_hdr map -t %d, its own Time: window "
                                       "(src/FEM_map.c:172-219: FASTQ read + map + SAM write; the reference and index load of every "
                                       "invocation is outside it)" % (n, cores)},
            "e2e": {"value": value, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def sha256_files(paths):
    import hashlib
    h = hashlib.sha256()
    for p in paths:
        with open(p, "rb") as f:
            while True:
                b = f.read(1 << 24)
                if not b:
                    break
                h.update(b)
    return h.hexdigest()


def run_ours(args):
    import ctypes as C
    rank, local_rank, world = dist_env()
    # stdout carries exactly one JSON line: whatever libraries print there (NCCL's version banner) goes to stderr
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    import numpy as np
    import torch
    torch.cuda.set_device(local_rank)
    dist, cpu_group = None, None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        cpu_group = dist.new_group(backend="gloo")  # ranks parked on the CPU (an NCCL barrier would spin on their GPUs)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def cpu_barrier():
        if dist is not None:
            dist.barrier(group=cpu_group)

    def reduce_ranks(x, op="max"):
        if dist is None:
            return x
        # the device by name: the single-process leg below leaves another device current on rank 0's thread
        t = torch.tensor([x], dtype=torch.float64, device=torch.device("cuda", local_rank))
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
        return float(t.item())

    from fem_b200 import host
    from fem_b200.api import Params, Result, SamInputs, Stats, _check, load_library
    w = WORKLOAD
    fa, idx, fq = ensure_data(args, rank, world, barrier)
    lib = load_library()
    t0 = time.time()
    hm = host.HostMapper(fa, idx, 1, [local_rank])
    ctx = hm.lib.femhost_mapper_context(hm.h, 0)
    reads = host.load_shard(fq, rank, world)  # the very records femhost_map_file maps for this rank's shard
    assert reads is not None, "the simulated read file is plain four-line FASTQ"
    n_mine = len(reads)
    n_reads = int(round(reduce_ranks(float(n_mine), "sum")))
    assert n_reads == args.reads, (n_reads, args.reads)
    log("rank %d: reference + index resident, shard of %d reads parsed in %.1fs" % (rank, n_mine, time.time() - t0))
    params, batch, res = Params(0, 0, w["e"], w["a"]), reads.as_batch(), Result()

    # ---- resident leg: `value` -----------------------------------------------------------------
    # names and qualities too: the SAM text is built on the device (k_sam), as in the end-to-end run
    sam_in = SamInputs(reads.c.names, C.cast(reads.c.name_off, C.c_void_p), reads.c.quals, C.cast(reads.c.has_qual, C.c_void_p))
    _check(lib.femgpu_set_sam_inputs(ctx, 0, C.byref(sam_in)))
    _check(lib.femgpu_map_batch(ctx, C.byref(params), C.byref(batch), C.byref(res)))  # stages the reads in HBM
    stats = res.stats.as_tuple()
    sam_text_bytes = int(res.sam_off[n_mine]) if res.sam_off else 0
    sam_host_reads = int(np.ctypeslib.as_array(res.sam_host, shape=(n_mine,)).sum()) if res.sam_off else None
    tot, st = C.c_float(0), (C.c_float * 6)()
    for _ in range(max(args.warmup, 3)):
        _check(lib.femgpu_rerun_resident(ctx, C.byref(tot), C.byref(st)))
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    wall0 = time.perf_counter()
    dev_ms, stage_ms, launches = 0.0, [0.0] * 6, 0
    ms5, nl = (C.c_float * 6)(), C.c_uint32(0)
    for _ in range(args.steps):
        _check(lib.femgpu_rerun_resident(ctx, C.byref(tot), C.byref(st)))
        dev_ms += tot.value
        stage_ms = [a + b for a, b in zip(stage_ms, st)]
        _check(lib.femgpu_last_timings(ctx, C.byref(ms5), C.byref(nl)))
        launches += nl.value
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - wall0)
    work = (C.c_uint64 * 4)()
    _check(lib.femgpu_last_work(ctx, C.byref(work)))
    dev_ms = reduce_ranks(dev_ms)
    wall_ms = reduce_ranks(wall_ms)
    launches = int(reduce_ranks(float(launches), "sum"))
    value = n_reads * args.steps / (dev_ms * 1e-3)

    # ---- end-to-end leg, form 1: one process per GPU, rank r maps shard r into its own part of the output ------
    cores = cpu_cores()
    threads = args.e2e_threads or max(2, cores // world - 3)  # -t; the engine adds its readers, GPU threads and copiers
    shm = "/dev/shm" if os.path.isdir("/dev/shm") else args.data_dir
    out = shm + "/fem_bench.sam.part%04d" % rank
    e2e_steps = args.e2e_steps or args.steps
    shard = (rank, world) if world > 1 else None

    def map_shard():
        return hm.map_file(fq, out, w["e"], w["a"], threads=threads, batch_reads=args.e2e_batch, no_header=rank > 0, shard=shard)

    map_shard()  # warm-up: pinned staging buffers, page cache
    # every step writes a new file (what a user's run does): removing the previous step's output is not timed
    e2e_stats, tm, e2e_s = None, None, 0.0
    for _ in range(e2e_steps):
        if os.path.exists(out):
            os.remove(out)
        barrier()
        e0 = time.perf_counter()
        e2e_stats, tm = map_shard()
        e2e_s += time.perf_counter() - e0
    barrier()
    e2e_s = reduce_ranks(e2e_s)
    sam_bytes = reduce_ranks(float(os.path.getsize(out)), "sum")
    assert e2e_stats[:5] == stats[:5], (e2e_stats, stats)
    h2d = reduce_ranks(float(tm["h2d_bytes"]), "sum")
    d2h = reduce_ranks(float(tm["d2h_bytes"]), "sum")
    forms = {"process_per_gpu": {"value": n_reads * e2e_steps / e2e_s, "unit": "reads/s", "steps": e2e_steps, "host_threads_per_process": threads,
                                 "phases_s_rank0": {k: tm[k] for k in ("wall", "parse", "gpu_wait", "format", "write")},
                                 "what": "femhost_map_file per rank on byte-range shard r of %d: FASTQ parse -> pinned H2D -> kernels -> D2H of the SAM text -> "
                                         "part file r on tmpfs (the parts in rank order are the SAM file); max over ranks" % world}}
    totals = [int(round(reduce_ranks(float(x), "sum"))) for x in stats]
    stats_all = tuple(totals)
    clocks = sampler.stop()  # sampled from the start of the resident leg to the end of this end-to-end leg

    # ---- end-to-end leg, form 2 (N > 1): ONE process, N GPUs, one input, one ordered output (`FEM map --gpus N`) ---
    identical = None
    if world > 1 and not args.no_single_process:
        parts = [shm + "/fem_bench.sam.part%04d" % r for r in range(world)]
        one = shm + "/fem_bench_single_process.sam"
        sp = None
        if rank == 0:
            try:
                sha_parts = sha256_files(parts)
                t1 = time.time()
                hmN = host.HostMapper(fa, idx, world, list(range(world)))  # index read once, peer copies to the other devices
                t_create = time.time() - t1
                thrN = args.e2e_threads or max(2, cores - 2 * world - 4)
                hmN.map_file(fq, one, w["e"], w["a"], threads=thrN, batch_reads=args.e2e_batch)  # warm-up
                sp_s, sp_tm, sp_stats = 0.0, None, None
                for _ in range(e2e_steps):
                    os.remove(one)
                    e0 = time.perf_counter()
                    sp_stats, sp_tm = hmN.map_file(fq, one, w["e"], w["a"], threads=thrN, batch_reads=args.e2e_batch)
                    sp_s += time.perf_counter() - e0
                sha_one = sha256_files([one])
                hmN.close()
                # the single-GPU file of the same reads, from this rank's own one-GPU mapper
                hm.map_file(fq, one, w["e"], w["a"], threads=thrN, batch_reads=args.e2e_batch)
                sha_n1 = sha256_files([one])
                os.remove(one)
                identical = sha_one == sha_n1 and sha_parts == sha_n1
                sp = {"value": n_reads * e2e_steps / sp_s, "unit": "reads/s", "steps": e2e_steps, "host_threads": thrN,
                      "phases_s": {k: sp_tm[k] for k in ("wall", "parse", "gpu_wait", "format", "write")},
                      "h2d_bytes_per_step": sp_tm["h2d_bytes"], "d2h_bytes_per_step": sp_tm["d2h_bytes"],
                      "mapper_create_s": t_create, "stats_equal": tuple(sp_stats[:5]) == stats_all[:5],
                      "identical_to_n1": sha_one == sha_n1, "process_per_gpu_parts_identical_to_n1": sha_parts == sha_n1, "sha256": sha_one,
                      "what": "femhost_map_file with n_gpus = %d in ONE process (rank 0; the other ranks wait on a CPU barrier): one reader pool, "
                              "2 threads per GPU, one ordered writer, one SAM file on tmpfs" % world}
            except Exception as ex:  # the leg is reported, never fatal for the line
                sp = {"value": None, "error": repr(ex)[:300]}
        cpu_barrier()
        torch.cuda.set_device(local_rank)
        if rank == 0:
            forms["single_process"] = sp
    if os.path.exists(out):
        os.remove(out)
    best = max((k for k in forms if forms[k] and forms[k].get("value")), key=lambda k: forms[k]["value"]) if rank == 0 else "process_per_gpu"
    e2e_value = forms[best]["value"]
    if best == "single_process":
        h2d, d2h = forms[best]["h2d_bytes_per_step"], forms[best]["d2h_bytes_per_step"]

    # ---- roofline of the dominant kernel -----------------------------------------------------------
    # per-stage times and work counters summed over ranks: bytes / (sum of ms / N) = the whole job's rate per GPU x N
    hbm_peak, peak_src = load_peaks()
    per = [reduce_ranks(m, "max") / args.steps for m in stage_ms]  # ms per step: seed, candidates, verify, compaction, traceback, SAM text
    n_s = w["read_len"] - w["k"] + 1
    cols, n_cand, c_before, n_maps = [int(round(reduce_ranks(float(x), "sum"))) for x in work]
    sam_text_total = reduce_ranks(float(sam_text_bytes), "sum")
    algo = {"k_seed": 16.0 * n_s * n_reads + w["read_len"] * n_reads,          # two LUT words per q-gram and strand + the read
            "k_cand": 8.0 * c_before,                                           # occurrence words of the selected seeds
            "k_verify": n_cand * ((w["read_len"] + 2 * w["e"] + 3) // 4) + 16.0 * n_maps,  # 2-bit window per candidate
            "k_compact": 16.0 * n_maps + 4.0 * n_cand,
            "k_trace": n_maps * (2.0 * ((w["read_len"] + 2 * w["e"] + 3) // 4) + w["read_len"] + 40.0)}  # window twice, read, 40-byte record
    # text written + what it is built from (read, qualities, name, 16-byte mapping, 88-byte alignment record)
    algo["k_sam"] = sam_text_total + 2.0 * w["read_len"] * stats_all[1] + 104.0 * n_maps
    names = ["k_seed", "k_cand", "k_verify", "k_compact", "k_trace", "k_sam"]
    # per GPU: every rank ran 1/N of the bytes in `per` milliseconds (max over ranks)
    kernels = {n: {"ms": per[i], "algorithmic_bytes": algo[n], "gbs_per_gpu": algo[n] / world / (per[i] * 1e-3) / 1e9 if per[i] > 0 else None}
               for i, n in enumerate(names)}
    cap = None
    try:  # DRAM bytes per launch, from the committed ncu --set full captures of this workload
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            cap = json.load(f).get(args.workload)
    except (OSError, ValueError):
        pass

    def roofline_of(i):
        ach = kernels[names[i]]["gbs_per_gpu"]
        r = {"kernel": names[i], "bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak if ach else None,
             "traffic": None, "traffic_source": None, "algorithmic_bytes": algo[names[i]] / world, "peak_source": peak_src,
             "share_of_step": per[i] / sum(per), "per": "GPU and launch (1/%d of the read set)" % world}
        if cap and names[i] in cap:
            r["traffic"] = cap[names[i]] / cap["reads"] * n_reads / world
            r["traffic_source"] = cap["source"] + (" (scaled from %d to %d reads)" % (cap["reads"], n_reads // world) if cap["reads"] != n_reads // world else "")
            # what the kernel really moves (a random 8-byte gather costs a 64-byte DRAM burst), over the same time
            r["traffic_gbs"] = r["traffic"] / (per[i] * 1e-3) / 1e9
            r["traffic_frac_of_peak"] = r["traffic_gbs"] / hbm_peak
        return r

    top = max(range(len(names)), key=lambda i: per[i])
    roofline = roofline_of(top)
    roofline_seed = roofline_of(0)
    # INT32 (ALU pipe) peak: the probe's three readings beside the nominal 64 lanes x SMs x the clock sampled during the probe
    probe_clock = ClockSampler(local_rank)
    probe_clock.start()
    giops = (C.c_double * 3)()
    _check(lib.femgpu_measure_int32_peak(ctx, C.byref(giops)))
    pc = probe_clock.stop()
    sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
    nominal = sms * 64 * (pc.get("sm_mhz") or 0) * 1e6 / 1e9
    lop3 = float(giops[0])
    peak_int = lop3 if nominal and 0.9 * nominal <= lop3 <= 1.02 * nominal else (nominal or lop3)
    int_ops = 24.0 * cols + 12.0 * w["e"] * n_cand  # 24 integer ops per executed column
    ach_int = int_ops / world / (per[2] * 1e-3) / 1e9 if per[2] > 0 else None
    ncu_alu = cap.get("k_verify_pipe_alu_pct") if cap else None
    verify_int32 = {"kernel": "k_verify", "bound": "int32", "achieved": ach_int, "peak": peak_int, "unit": "Gop/s",
                    "frac": ach_int / peak_int if ach_int and peak_int else None,
                    "peak_source": "probe LOP3 %.0f / IADD3 %.0f / SHF+LOP3 %.0f Gop/s (femgpu_measure_int32_peak, k_peak.cuh); nominal %d SMs x 64 lanes x %.0f MHz "
                                   "sampled during the probe = %.0f Gop/s; the LOP3 reading is used when it lies within 0.90..1.02 of nominal, else nominal"
                                   % (giops[0], giops[1], giops[2], sms, pc.get("sm_mhz") or 0, nominal),
                    "ncu_pipe_alu_pct": ncu_alu, "ncu_source": cap.get("k_verify_source") if cap else None,
                    "columns_executed": cols, "candidates": n_cand, "ops_counted": "24 per executed column + 12e per candidate (src/align.c:118-133, SURVEY.md 8(d))",
                    "note": "the plane form executes 20 SASS instructions per column (profiles/r2_sass_k_verify_column.txt): LOP3 fuses pairs of the 24 counted "
                            "operations, so the op-count fraction can pass the pipe's INSTRUCTION rate; ncu_pipe_alu_pct is the utilisation of the pipe itself"}

    line = {"metric": "reads/sec (all locations)", "value": value, "value_scope": "resident: reads already in HBM, every kernel from seeding to the SAM text",
            "unit": "reads/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": dev_ms / args.steps, "wall_ms_per_step": wall_ms / args.steps,
            "higher_is_better": True, "scaling": scaling_kind(args), "vs_baseline": None, "dtype": "u32/u64 integer", "data": "synthetic",
            "config": config_dict(args, world), "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "reads/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "sam_bytes_per_step": sam_bytes, "form": best, "forms": forms, "outputs_identical_to_single_gpu": identical},
            "gpu_launches": launches, "roofline": roofline, "roofline_seed": roofline_seed, "kernels": kernels, "roofline_verify_int32": verify_int32,
            "stats": dict(zip(("reads", "mapped_reads", "candidates_before_filter", "candidates", "mappings", "infeasible"), stats_all)),
            "device_sam": {"text_bytes": sam_text_total, "reads_left_to_host": sam_host_reads}}

    # ---- CPU baseline beside it (rank 0, N=1 only) ---------------------------------------------------
    if world == 1 and not args.no_cpu_baseline and os.access(REF_HDR, os.X_OK):
        sample = head_fastq(fq, fq + ".cpu%d" % args.cpu_sample, args.cpu_sample)
        ref_threads = args.cpu_threads or cores
        secs, n = time_reference(fa, idx, sample, w["e"], w["a"], ref_threads, out + ".ref")
        # untimed: the same sample through this repository's path, compared record for record
        hm.map_file(sample, out + ".ours", w["e"], w["a"], threads=threads)
        ref_g, our_g = sam_groups(out + ".ref"), sam_groups(out + ".ours")
        line["parity_vs_reference"] = {"reads": n, "sam_records": sum(len(v) for k, v in ref_g.items() if k is not None),
                                       "identical": ref_g == our_g,
                                       "how": "header + per-read record groups (content and order) of oracle/_ref/FEM_ref_hdr -t %d vs femhost_map_file on the same reads" % ref_threads}
        os.remove(out + ".ref")
        os.remove(out + ".ours")
        line["cpu_baseline"] = {"value": n / secs, "unit": "reads/s", "cores": ref_threads, "kind": "reference",
                                "sample": "first %d reads of the workload, oracle/_ref/FEM_ref_hdr map -t %d, its own Time: window" % (n, ref_threads)}
    elif world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = {"value": None, "unit": "reads/s", "cores": cores, "kind": "reference", "sample": "oracle/_ref/FEM_ref_hdr not built"}
    hm.close()
    reads.close()

    # ---- the other BASELINE.json configs, each in a sub-run of this script (N = 1 only) ------------------
    if world == 1 and args.others and args.others != "none" and args.workload == DEFAULT_WORKLOAD:
        wanted = dict(OTHER_WORKLOADS)
        pick = [x for x in (wanted if args.others == "all" else args.others.split(",")) if x in wanted and x != args.workload]
        line["other_workloads"] = {}
        for name in pick:
            cmd = [sys.executable, os.path.abspath(__file__), "--workload", name, "--reads", str(wanted[name]), "--steps", "3", "--warmup", "3",
                   "--cpu-sample", str(min(args.cpu_sample, wanted[name] // 5)), "--data-dir", args.data_dir, "--others", "none"]
            try:
                p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=900)
                sub = json.loads(p.stdout.strip().splitlines()[-1])
                line["other_workloads"][name] = {k: sub.get(k) for k in ("value", "ms_per_step", "config", "e2e", "roofline", "roofline_seed", "kernels",
                                                                           "roofline_verify_int32", "stats", "parity_vs_reference", "cpu_baseline")}
            except Exception as ex:
                line["other_workloads"][name] = {"error": repr(ex)[:300]}

    if rank == 0:
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
