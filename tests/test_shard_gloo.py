"""The N>1 host logic on CPU: world_size-2 (and 3) gloo process groups run fem_b200.shard.map_sharded
with the ORACLE standing in for the per-rank GPU mapper (tests may use the oracle; the product's
map_shard calls femhost_map_file on the rank's GPU).  Checked: contiguous shards, header only on rank
0, rank-ordered concatenation == the single-process SAM byte for byte, counters summed over ranks,
max-over-ranks time, ranks without reads."""
import os
import socket

import pytest

import oracle_lib as ol
import synth
from fem_b200 import shard


def test_shard_ranges_partition_the_reads():
    for n in (0, 1, 2, 3, 7, 100, 1_000_003):
        for world in (1, 2, 3, 8):
            pos = 0
            for r in range(world):
                skip, count = shard.shard_range(n, r, world)
                assert skip == pos and count in (n // world, n // world + 1)
                pos += count
            assert pos == n
    with pytest.raises(ValueError):
        shard.shard_range(10, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, workdir, n_reads, e, reads_name="reads.fq", byte_shards=False):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ctl = shard.Control(dist)
        oracle = ol.Oracle(fasta=os.path.join(workdir, "ref.fa"))
        oracle.load_index(os.path.join(workdir, "idx"))
        oracle.set_params(e, 1)
        from fem_b200 import host
        reads_path = os.path.join(workdir, reads_name)

        def map_records(recs, part, no_header):
            fq = part + ".fq"
            with open(fq, "wb") as f:  # records without qualities (FASTA input) stay FASTA
                f.write(b"".join(b">%s\n%s\n" % (n, s) if q is None else b"@%s\n%s\n+\n%s\n" % (n, s, q) for n, s, q in recs))
            st = oracle.map_file(fq, part + ".full")
            with open(part + ".full", "rb") as f, open(part, "wb") as g:
                g.write(b"".join(l for l in f if not (no_header and l.startswith(b"@"))))
            os.remove(fq)
            os.remove(part + ".full")
            return st.as_tuple()[:5]

        def map_shard(skip, count, part, no_header):  # read-count slice, cut with the real parser
            recs = host.load_all(reads_path).as_list()
            return map_records(recs[skip:] if count is None else recs[skip:skip + count], part, no_header)

        def map_byte_shard(r, n, part, no_header):  # the cut femhost_map_file makes for shard_index / shard_count
            recs = host.load_shard(reads_path, r, n)
            return None if recs is None else map_records(recs.as_list(), part, no_header)

        total, seconds = shard.map_sharded(ctl, map_shard, lambda: host.count_records(reads_path),
                                           os.path.join(workdir, "sharded.sam"), map_byte_shard=map_byte_shard if byte_shards else None)
        assert seconds > 0
        # every rank sees the same reduced values
        with open(os.path.join(workdir, "result.%d" % rank), "w") as f:
            f.write(repr((total, round(seconds, 9))))
        oracle.close()
    finally:
        dist.destroy_process_group()


def _write_multiline_fastq(path, reads, crlf=False):
    """the same records in shapes the byte-range cut refuses: sequence over two lines, or CRLF line ends"""
    nl = b"\r\n" if crlf else b"\n"
    with open(path, "wb") as f:
        for name, seq, qual in reads:
            name, seq, qual = (x if isinstance(x, bytes) else x.encode() for x in (name, seq, qual))
            if crlf:
                f.write(b"@" + name + nl + seq + nl + b"+" + nl + qual + nl)
            else:
                h = len(seq) // 2
                f.write(b"@" + name + nl + seq[:h] + nl + seq[h:] + nl + b"+" + nl + qual[:h] + nl + qual[h:] + nl)


@pytest.mark.parametrize("world,n_reads,shape", [(2, 301, "count"), (3, 2, "count"), (2, 0, "count"), (2, 301, "bytes"), (3, 257, "bytes"),
                                                 (3, 2, "bytes"), (2, 0, "bytes"), (2, 150, "multiline"), (3, 150, "fasta.gz")])
def test_sharded_map_equals_single_process(tmp_path, world, n_reads, shape):
    import torch.multiprocessing as mp
    ol.build()
    workdir = str(tmp_path)
    records = synth.random_reference(5, [120_000, 80_000], 0.1)
    reads = synth.simulate_reads(6, records, n_reads, 100, 3)
    synth.write_fasta(os.path.join(workdir, "ref.fa"), records)
    synth.write_fastq(os.path.join(workdir, "reads.fq"), reads)
    oracle = ol.Oracle(fasta=os.path.join(workdir, "ref.fa"))
    oracle.build_index(12, 3)
    oracle.save_index(os.path.join(workdir, "idx"))
    oracle.set_params(3, 1)
    whole = oracle.map_file(os.path.join(workdir, "reads.fq"), os.path.join(workdir, "whole.sam")).as_tuple()[:5]
    oracle.close()
    from fem_b200 import host
    reads_name = "reads.fq"
    if shape == "multiline":  # byte-range shards refuse it on every rank -> read-count shards from the real parser
        reads_name = "reads_ml.fq"
        _write_multiline_fastq(os.path.join(workdir, reads_name), reads)
    elif shape == "fasta.gz":  # two lines per record, gzip: what `newlines // 4` got wrong
        import gzip
        reads_name = "reads.fa.gz"
        with gzip.open(os.path.join(workdir, reads_name), "wb") as f:
            for name, seq, _ in reads:
                f.write(b">%s\n%s\n" % tuple(x if isinstance(x, bytes) else x.encode() for x in (name, seq)))
    if reads_name != "reads.fq":  # the single-process answer for this very file
        oracle = ol.Oracle(fasta=os.path.join(workdir, "ref.fa"))
        oracle.load_index(os.path.join(workdir, "idx"))
        oracle.set_params(3, 1)
        whole = oracle.map_file(os.path.join(workdir, reads_name), os.path.join(workdir, "whole.sam")).as_tuple()[:5]
        oracle.close()
    assert host.count_records(os.path.join(workdir, reads_name)) == n_reads
    mp.spawn(_rank_main, args=(world, _free_port(), workdir, n_reads, 3, reads_name, shape != "count"), nprocs=world, join=True)
    with open(os.path.join(workdir, "whole.sam"), "rb") as f, open(os.path.join(workdir, "sharded.sam"), "rb") as g:
        assert f.read() == g.read()
    results = [eval(open(os.path.join(workdir, "result.%d" % r)).read()) for r in range(world)]
    assert all(r == results[0] for r in results)
    assert tuple(results[0][0]) == tuple(whole)
    assert not [p for p in os.listdir(workdir) if ".part" in p]
