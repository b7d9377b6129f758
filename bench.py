#!/usr/bin/env python
"""bench.py -- reads/sec of the `FEM map` hot path on B200 (BASELINE.json's metric).

  python bench.py --gpus N --steps K --warmup W            # this repository's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU code

Workload = BASELINE.json configs[1]: synthetic 100 Mbp reference (10 records), 1 M simulated
100 bp single-end reads, e=5, index window 12 / step 4 -- per GPU (weak scaling: every rank maps
its own 1 M reads against the replicated index; there is no collective on the data path).
A "step" is one pass of the hot path over the rank's reads.

  value  reads/s with the reads already resident in HBM (femgpu_rerun_resident: seed ->
         candidates -> verify -> compaction -> traceback), CUDA-event time summed over the K steps, max over ranks
  e2e    reads/s of the whole `FEM map` run through the C ABI (femhost_map_file): FASTQ parse from
         the page cache, pinned H2D, kernels, D2H, mapping order + traceback + SAM text, file written
         to tmpfs -- the same window as the reference's own "Time:" line (src/FEM_map.c:172-219)
  roofline      dominant kernel: algorithmic bytes / its CUDA-event time vs MEASURED_PEAKS.json
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
                 ref_bp=3_100_000_000, ref_records=24, repeat_permille=0, reads=50_000_000, read_len=250, e=7, a=1, k=14, step=10, seed=5),
}
WORKLOAD = dict(WORKLOADS["cfg2"])  # the metric's single-GPU configuration; --workload selects another
FEMSIM = os.path.join(ROOT, "tools", "femsim")
REF_HDR = os.path.join(ROOT, "oracle", "_ref", "FEM_ref_hdr")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--reads", type=int, default=0, help="reads per GPU (default: the workload's)")
    ap.add_argument("--ref-bp", type=int, default=0)
    ap.add_argument("--data-dir", default=os.environ.get("FEM_BENCH_DIR", "/tmp/fem_bench"))
    ap.add_argument("--cpu-sample", type=int, default=200_000, help="reads of the workload the CPU baseline maps")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    """Synthetic inputs (untimed): reference + index shared by all ranks, one reads file per rank."""
    from fem_b200 import host
    w = WORKLOAD
    d = os.path.join(args.data_dir, "ref%d_rep%d_k%d_s%d" % (args.ref_bp, w["repeat_permille"], w["k"], w["step"]))
    os.makedirs(d, exist_ok=True)
    fa, idx = os.path.join(d, "ref.fa"), os.path.join(d, "index")
    if rank == 0:
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
    barrier()
    fq = os.path.join(d, "reads_%d_l%d_e%d_r%d.fq" % (args.reads, w["read_len"], w["e"], rank))
    if not os.path.exists(fq + ".ok"):
        subprocess.check_call([FEMSIM, "reads", str(1000 + w["seed"] + rank), fa, str(args.reads), str(w["read_len"]),
                               str(w["e"]), fq])
        open(fq + ".ok", "w").close()
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
    return {"workload": w["name"], "reads_per_gpu": args.reads, "read_len": w["read_len"], "e": w["e"], "a": w["a"],
            "window": w["k"], "step": w["step"], "reference_bp": args.ref_bp, "sharding": "reads sharded per GPU, index replicated, no collective",
            "l2": "inputs larger than L2 (index %.0f MB + reads, re-streamed every step)" % (4.0 * (4 ** w["k"] + 1) / 1e6 + 8.0 * args.ref_bp / w["step"] / 1e6)}


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
    cores = cpu_cores()
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
            "scaling": "weak", "vs_baseline": None, "dtype": "u32/u64 integer", "data": "synthetic", "config": config_dict(args, world),
            "cpu_baseline": {"value": value, "unit": "reads/s", "cores": cores, "kind": "reference",
                             "sample": "first %d reads of the workload per step, FEM_ref_hdr map -t %d, its own Time: window" % (n, cores)},
            "e2e": {"value": value, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def run_ours(args):
    import ctypes as C
    rank, local_rank, world = dist_env()
    # stdout carries exactly one JSON line: whatever libraries print there (NCCL's version banner) goes to stderr
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    import torch
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    from fem_b200 import host
    from fem_b200.api import Params, Result, SamInputs, Stats, _check, load_library
    w = WORKLOAD
    fa, idx, fq = ensure_data(args, rank, world, barrier)
    lib = load_library()
    t0 = time.time()
    hm = host.HostMapper(fa, idx, 1, [local_rank])
    ctx = hm.lib.femhost_mapper_context(hm.h, 0)
    reads = host.load_all(fq)
    n_reads = len(reads)
    log("rank %d: reference + index resident, %d reads parsed in %.1fs" % (rank, n_reads, time.time() - t0))
    params, batch, res = Params(0, 0, w["e"], w["a"]), reads.as_batch(), Result()

    # ---- resident leg: `value` -----------------------------------------------------------------
    # names and qualities too: the SAM text is built on the device (k_sam), as in the end-to-end run
    sam_in = SamInputs(reads.c.names, C.cast(reads.c.name_off, C.c_void_p), reads.c.quals, C.cast(reads.c.has_qual, C.c_void_p))
    _check(lib.femgpu_set_sam_inputs(ctx, 0, C.byref(sam_in)))
    _check(lib.femgpu_map_batch(ctx, C.byref(params), C.byref(batch), C.byref(res)))  # stages the reads in HBM
    stats = res.stats.as_tuple()
    sam_text_bytes = int(res.sam_off[n_reads]) if res.sam_off else 0
    import numpy as np
    sam_host_reads = int(np.ctypeslib.as_array(res.sam_host, shape=(n_reads,)).sum()) if res.sam_off else None
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
    dev_ms = max_over_ranks(dev_ms)
    wall_ms = max_over_ranks(wall_ms)
    value = world * n_reads * args.steps / (dev_ms * 1e-3)

    # ---- end-to-end leg: the whole `FEM map` run through the C ABI, host buffers -----------------
    cores = cpu_cores()
    threads = args.e2e_threads or max(2, cores // world - 3)  # -t; the engine adds its readers, GPU threads and copiers
    out = ("/dev/shm" if os.path.isdir("/dev/shm") else args.data_dir) + "/fem_bench_r%d.sam" % rank
    e2e_steps = args.e2e_steps or args.steps
    hm.map_file(fq, out, w["e"], w["a"], threads=threads, batch_reads=args.e2e_batch)  # warm-up: pinned staging buffers, page cache
    # every step writes a new file (what a user's run does): removing the previous step's output is not timed
    e2e_stats, tm, e2e_s = None, None, 0.0
    for _ in range(e2e_steps):
        if os.path.exists(out):
            os.remove(out)
        barrier()
        e0 = time.perf_counter()
        e2e_stats, tm = hm.map_file(fq, out, w["e"], w["a"], threads=threads, batch_reads=args.e2e_batch)
        e2e_s += time.perf_counter() - e0
    barrier()
    clocks = sampler.stop()  # sampled from the start of the resident leg to the end of the end-to-end leg
    e2e_s = max_over_ranks(e2e_s)
    sam_bytes = os.path.getsize(out)
    os.remove(out)
    assert e2e_stats[:5] == stats[:5], (e2e_stats, stats)
    e2e_value = world * n_reads * e2e_steps / e2e_s

    # ---- roofline of the dominant kernel -----------------------------------------------------------
    hbm_peak, peak_src = load_peaks()
    per = [m / args.steps for m in stage_ms]  # ms per step: seed, candidates, verify, compaction, traceback, SAM text
    n_s = w["read_len"] - w["k"] + 1
    cols, n_cand, c_before, n_maps = [int(x) for x in work]
    algo = {"k_seed": 16.0 * n_s * n_reads + w["read_len"] * n_reads,          # two LUT words per q-gram and strand + the read
            "k_cand": 8.0 * c_before,                                           # occurrence words of the selected seeds
            "k_verify": n_cand * ((w["read_len"] + 2 * w["e"] + 3) // 4) + 16.0 * n_maps,  # 2-bit window per candidate
            "k_compact": 16.0 * n_maps + 4.0 * n_cand,
            "k_trace": n_maps * (2.0 * ((w["read_len"] + 2 * w["e"] + 3) // 4) + w["read_len"] + 40.0)}  # window twice, read, 40-byte record
    # text written + what it is built from (read, qualities, name, 16-byte mapping, 88-byte alignment record)
    algo["k_sam"] = sam_text_bytes + 2.0 * w["read_len"] * stats[1] + 104.0 * n_maps
    names = ["k_seed", "k_cand", "k_verify", "k_compact", "k_trace", "k_sam"]
    kernels = {n: {"ms": per[i], "algorithmic_bytes": algo[n], "gbs": algo[n] / (per[i] * 1e-3) / 1e9 if per[i] > 0 else None}
               for i, n in enumerate(names)}
    top = max(range(len(names)), key=lambda i: per[i])
    ach = kernels[names[top]]["gbs"]
    traffic, traffic_src = None, None
    try:  # DRAM bytes of that kernel per launch, from the committed ncu --set full capture of this workload
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            cap = json.load(f).get(args.workload)
        if cap and names[top] in cap:
            traffic = cap[names[top]] / cap["reads"] * n_reads
            traffic_src = cap["source"] + (" (scaled from %d to %d reads)" % (cap["reads"], n_reads) if cap["reads"] != n_reads else "")
    except (OSError, ValueError):
        pass
    roofline = {"kernel": names[top], "bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                "traffic": traffic, "traffic_source": traffic_src, "algorithmic_bytes": algo[names[top]],
                "peak_source": peak_src, "share_of_step": per[top] / sum(per)}
    if traffic:  # what the kernel really moves (random 8-byte gathers cost a 64-byte DRAM burst each), for the same time
        roofline["traffic_gbs"] = traffic / (per[top] * 1e-3) / 1e9
        roofline["traffic_frac_of_peak"] = roofline["traffic_gbs"] / hbm_peak
    giops = (C.c_double * 3)()
    _check(lib.femgpu_measure_int32_peak(ctx, C.byref(giops)))
    int_ops = 24.0 * cols + 12.0 * w["e"] * n_cand  # 24 integer ops per executed column
    verify_int32 = {"kernel": "k_verify", "bound": "int32", "achieved": int_ops / (per[2] * 1e-3) / 1e9 if per[2] > 0 else None,
                    "peak": giops[0], "unit": "Gop/s", "peak_source": "femgpu_measure_int32_peak (LOP3 %.0f, IADD3 %.0f, SHF+LOP3 %.0f Gop/s)" % tuple(giops),
                    "columns_executed": cols, "candidates": n_cand}
    if verify_int32["achieved"]:
        verify_int32["frac"] = verify_int32["achieved"] / giops[0]

    line = {"metric": "reads/sec (all locations)", "value": value, "unit": "reads/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": dev_ms / args.steps, "wall_ms_per_step": wall_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32/u64 integer", "data": "synthetic",
            "config": config_dict(args, world), "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "reads/s", "h2d_bytes_per_step": tm["h2d_bytes"], "d2h_bytes_per_step": tm["d2h_bytes"],
                    "steps": e2e_steps, "sam_bytes_per_step": sam_bytes, "host_threads": threads,
                    "phases_s": {k: tm[k] for k in ("wall", "parse", "gpu_wait", "format", "write")},
                    "what": "femhost_map_file: FASTQ parse -> pinned H2D -> kernels -> D2H -> mapping order, traceback, SAM text -> file on tmpfs"},
            "gpu_launches": launches, "roofline": roofline, "kernels": kernels, "roofline_verify_int32": verify_int32,
            "stats": dict(zip(("reads", "mapped_reads", "candidates_before_filter", "candidates", "mappings", "infeasible"), stats)),
            "device_sam": {"text_bytes": sam_text_bytes, "reads_left_to_host": sam_host_reads}}

    # ---- CPU baseline beside it (rank 0, N=1 only) ---------------------------------------------------
    if world == 1 and not args.no_cpu_baseline and os.access(REF_HDR, os.X_OK):
        sample = head_fastq(fq, fq + ".cpu%d" % args.cpu_sample, args.cpu_sample)
        secs, n = time_reference(fa, idx, sample, w["e"], w["a"], cores, out + ".ref")
        # untimed: the same sample through this repository's path, compared record for record
        hm.map_file(sample, out + ".ours", w["e"], w["a"], threads=threads)
        ref_g, our_g = sam_groups(out + ".ref"), sam_groups(out + ".ours")
        line["parity_vs_reference"] = {"reads": n, "sam_records": sum(len(v) for k, v in ref_g.items() if k is not None),
                                       "identical": ref_g == our_g,
                                       "how": "header + per-read record groups (content and order) of oracle/_ref/FEM_ref_hdr -t %d vs femhost_map_file on the same reads" % cores}
        os.remove(out + ".ref")
        os.remove(out + ".ours")
        line["cpu_baseline"] = {"value": n / secs, "unit": "reads/s", "cores": cores, "kind": "reference",
                                "sample": "first %d reads of the workload, oracle/_ref/FEM_ref_hdr map -t %d, its own Time: window" % (n, cores)}
    elif world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = {"value": None, "unit": "reads/s", "cores": cores, "kind": "reference", "sample": "oracle/_ref/FEM_ref_hdr not built"}
    hm.close()
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
