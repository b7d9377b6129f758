"""One process per GPU: `FEM map` with the reads sharded across ranks.

Reads are independent (src/map.c:27-56 keeps no cross-read state but additive counters), so rank r
maps a contiguous slice of the read file against its own replica of the index; there is NO
collective on the data path.  torch.distributed carries only the control plane: a barrier around
the timed region, the max-over-ranks time, the sum of the five counters, one flag.  Rank 0 prints
the @SQ header, the other ranks none, so the shard files concatenated in rank order are the
single-process SAM file byte for byte (= the reference's -t 1 order).

Slices: a plain four-line FASTQ file is cut by BYTE ranges (rank r takes the records that start in
[size*r/world, size*(r+1)/world), femhost_map_options.shard_index/shard_count): no counting pass and
no rank parses reads it does not map.  Any other input (FASTA, gzip, multi-line or CRLF records) makes
every rank fall back -- together, after a max-reduce of one flag -- to read-COUNT slices
shard_range(n, r, world), with n from the real parser (femhost_count_records).

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 -m fem_b200.shard \\
        -e 5 -t 8 --ref ref.fa --index idx --read1 reads.fq -o out.sam
"""
import argparse
import os
import shutil
import sys
import time


def shard_range(n_reads, rank, world):
    """(skip, count) of rank's contiguous slice: the first n % world ranks get one read more."""
    if world < 1 or not 0 <= rank < world or n_reads < 0:
        raise ValueError("bad shard request: n=%r rank=%r world=%r" % (n_reads, rank, world))
    base, extra = divmod(n_reads, world)
    skip = rank * base + min(rank, extra)
    return skip, base + (1 if rank < extra else 0)


def concatenate(parts, out_path):
    """rank-ordered shard files -> one SAM file"""
    with open(out_path, "wb") as out:
        for p in parts:
            with open(p, "rb") as f:
                shutil.copyfileobj(f, out, 1 << 24)


class Control:
    """the control plane: barrier, max and sum over ranks (gloo on CPU, nccl on GPUs); world 1 needs none"""

    def __init__(self, dist=None, device=None):
        self.dist, self.device = dist, device
        self.rank = dist.get_rank() if dist else 0
        self.world = dist.get_world_size() if dist else 1

    def barrier(self):
        if self.dist:
            self.dist.barrier()

    def _reduce(self, values, op):
        if not self.dist:
            return list(values)
        import torch
        t = torch.tensor(list(values), dtype=torch.float64 if op == "max" else torch.int64, device=self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return t.tolist()

    def max(self, x):
        return float(self._reduce([float(x)], "max")[0])

    def sum(self, counters):
        return tuple(int(v) for v in self._reduce([int(c) for c in counters], "sum"))


def map_sharded(ctl, map_shard, count_reads, out_path, part_dir=None, map_byte_shard=None):
    """Every rank maps its slice into a part file, then rank 0 concatenates.

    map_byte_shard(rank, world, part_path, no_header) -> counters, or None when byte-range slices do not apply
    to the input (optional; tried first);  map_shard(skip, count, part_path, no_header) -> counters maps the
    read-count slice shard_range(count_reads(), rank, world); count None = to the end of the file (a world of
    one needs no count).  Returns (summed counters, max-over-ranks seconds of the mapping phase); the summed read
    counter is checked against the parser's count on the fallback path."""
    part_dir = part_dir or os.path.dirname(os.path.abspath(out_path))
    part = os.path.join(part_dir, os.path.basename(out_path) + ".part%04d" % ctl.rank)
    ctl.barrier()
    t0 = time.perf_counter()
    counters = map_byte_shard(ctl.rank, ctl.world, part, ctl.rank != 0) if map_byte_shard and ctl.world > 1 else None
    if ctl.world == 1:
        total = tuple(map_shard(0, None, part, False))
    elif ctl.max(0.0 if counters is not None else 1.0) > 0.0:  # some rank could not: all ranks take read-count slices
        n_reads = count_reads() if callable(count_reads) else int(count_reads)
        skip, count = shard_range(n_reads, ctl.rank, ctl.world)
        counters = map_shard(skip, count, part, ctl.rank != 0)
        total = ctl.sum(counters)
        if total[0] != n_reads:
            raise RuntimeError("sharded run mapped %d reads, the parser counts %d" % (total[0], n_reads))
    else:
        total = ctl.sum(counters)
    seconds = ctl.max(time.perf_counter() - t0)
    ctl.barrier()  # every part file is complete
    if ctl.rank == 0:
        parts = [os.path.join(part_dir, os.path.basename(out_path) + ".part%04d" % r) for r in range(ctl.world)]
        concatenate(parts, out_path)
        for p in parts:
            os.remove(p)
    ctl.barrier()
    return total, seconds


def main(argv=None):
    ap = argparse.ArgumentParser(description="FEM map, one process per GPU (launch under torchrun)")
    ap.add_argument("-e", type=int, default=2)
    ap.add_argument("-a", type=int, default=1)
    ap.add_argument("-t", type=int, default=1)
    ap.add_argument("--ref", required=True)
    ap.add_argument("--index", required=True)
    ap.add_argument("--read1", required=True)
    ap.add_argument("-o", required=True)
    args = ap.parse_args(argv)
    rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
    import torch
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctl = Control(dist, torch.device("cuda", local) if dist else None)
    from fem_b200 import host
    hm = host.HostMapper(args.ref, args.index, 1, [local])  # raises without a device: there is no CPU fallback

    def map_byte_shard(r, n, part, no_header):
        try:
            st, _ = hm.map_file(args.read1, part, args.e, args.a, threads=args.t, no_header=no_header, shard=(r, n))
        except host.NotFastq4:
            return None
        return st[:5]

    def map_shard(skip, count, part, no_header):
        # count == 0: more ranks than reads (rank 0 still prints the header of an empty run); None: to the end
        st, _ = hm.map_file(args.read1, part, args.e, args.a, threads=args.t, skip_reads=skip, max_reads=count or 0,
                            no_header=no_header, no_reads=count == 0)
        return st[:5]

    total, seconds = map_sharded(ctl, map_shard, lambda: host.count_records(args.read1), args.o, map_byte_shard=map_byte_shard)
    hm.close()
    if rank == 0:
        for label, v in zip(("read", "mapped read", "candidate before additional q-gram filter", "candidate", "mapping"), total):
            print("The number of %s: %d" % (label, v), file=sys.stderr)
        print("Time: %fs (max over %d ranks)" % (seconds, world), file=sys.stderr)
    if dist:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
