"""GPU tests of the whole drop-in path: the `FEM` command line (fem_b200/bin/FEM -> libfemhost.so ->
libfemgpu.so) must write the reference's SAM file byte for byte and print the same five counters.
Checked against the golden fixtures (reference binaries' own output) and, on larger seeded inputs,
against the oracle restatement (which tests/test_oracle_vs_ref.py pins to the reference)."""
import glob
import gzip
import json
import os
import re
import subprocess

import pytest

import oracle_lib as ol
import synth

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
TAGS = sorted(os.path.basename(p)[:-5] for p in glob.glob(os.path.join(GOLD, "*.json")))


def run_fem_map(fa, idx, fq, sam, e, a, threads=4, extra=()):
    from fem_b200.host import FEM_BINARY
    p = subprocess.run([FEM_BINARY, "map", "-e", str(e), "-a", str(a), "-t", str(threads), "--ref", fa, "--index", idx,
                        "--read1", fq, "-o", sam, *extra], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert p.returncode == 0, p.stderr[-2000:]
    vals = []
    for tag in ("The number of read:", "The number of mapped read:", "The number of candidate before additional q-gram filter:",
                "The number of candidate:", "The number of mapping:"):
        m = re.search(re.escape(tag) + r" (\d+)", p.stderr)
        assert m, p.stderr
        vals.append(int(m.group(1)))
    return tuple(vals)


@pytest.mark.parametrize("tag", TAGS)
def test_cli_reproduces_golden_sam(tag, tmp_path):
    from fem_b200.host import FEM_BINARY
    with open(os.path.join(GOLD, tag + ".json")) as f:
        meta = json.load(f)
    fa, fq = (os.path.join(GOLD, "%s.%s" % (tag, s)) for s in ("ref.fa.gz", "reads.fq.gz"))
    with gzip.open(os.path.join(GOLD, tag + ".sam.gz"), "rb") as f:
        want = f.read()
    idx, sam = str(tmp_path / "idx"), str(tmp_path / "out.sam")
    subprocess.run([FEM_BINARY, "index", str(meta["k"]), str(meta["step"]), fa, idx], check=True, stderr=subprocess.DEVNULL)
    extra = ("--asis",) if meta["binary"] == "asis" else ()
    counters = run_fem_map(fa, idx, fq, sam, meta["e"], meta["a"], extra=extra)
    c = meta["counters"]
    assert counters == (c["num_reads"], c["num_mapped_reads"], c["num_candidates_before_filter"], c["num_candidates"],
                        c["num_mappings"])
    with open(sam, "rb") as f:
        assert f.read() == want


BIG = [
    ("cli_cfg1", [600_000, 400_000, 300_000], 0.02, 30_000, 100, 3, 1, 12, 5),
    ("cli_cfg2", [900_000, 500_000], 0.02, 20_000, 100, 5, 1, 12, 4),
    ("cli_cfg3", [800_000, 400_000], 0.05, 10_000, 150, 7, 1, 12, 5),
    ("cli_rep", [300_000, 200_000], 0.40, 4_000, 150, 7, 1, 12, 2),
]


@pytest.fixture(scope="module", params=BIG, ids=[c[0] for c in BIG])
def big(request, tmp_path_factory):
    from fem_b200.host import FEM_BINARY
    tag, lens, rep, n_reads, l, e, a, k, step = request.param
    d = tmp_path_factory.mktemp(tag)
    records = synth.random_reference(sum(map(ord, tag)), lens, rep)
    reads = synth.simulate_reads(sum(map(ord, tag)) + 1, records, n_reads, l, e)
    fa, fq, idx = str(d / "ref.fa"), str(d / "reads.fq"), str(d / "idx")
    synth.write_fasta(fa, records)
    synth.write_fastq(fq, reads)
    subprocess.run([FEM_BINARY, "index", str(k), str(step), fa, idx], check=True, stderr=subprocess.DEVNULL)
    o = ol.Oracle(fasta=fa)
    o.load_index(idx)
    o.set_params(e, a)
    want_sam = str(d / "oracle.sam")
    st = o.map_file(fq, want_sam)
    o.close()
    with open(want_sam, "rb") as f:
        want = f.read()
    return dict(d=d, fa=fa, fq=fq, idx=idx, e=e, a=a, want=want, counters=st.as_tuple()[:5], n_reads=n_reads)


def test_cli_matches_oracle_on_larger_inputs(big):
    sam = str(big["d"] / "gpu.sam")
    # small batches: many batches in flight on both slots, several SAM threads
    counters = run_fem_map(big["fa"], big["idx"], big["fq"], sam, big["e"], big["a"], threads=8, extra=("--batch", "1500"))
    assert counters == big["counters"]
    with open(sam, "rb") as f:
        assert f.read() == big["want"]
    counters = run_fem_map(big["fa"], big["idx"], big["fq"], sam, big["e"], big["a"], threads=1)
    assert counters == big["counters"]
    with open(sam, "rb") as f:
        assert f.read() == big["want"]


def test_read_shards_concatenate_to_the_whole(big):
    """One process per GPU: contiguous read shards, outputs concatenated in shard order (8e)."""
    from fem_b200.host import HostMapper
    m = HostMapper(big["fa"], big["idx"], 1)
    n, parts, total, n_dev = big["n_reads"], [], [0] * 5, 0
    cuts = [0, n // 3, n // 3 + 1, n]
    for r in range(3):
        out = str(big["d"] / ("shard%d.sam" % r))
        st, _ = m.map_file(big["fq"], out, big["e"], big["a"], threads=4, batch_reads=997, skip_reads=cuts[r],
                           max_reads=cuts[r + 1] - cuts[r], no_header=r > 0)
        total = [x + y for x, y in zip(total, st[:5])]
        n_dev += _["device_sam_reads"]
        with open(out, "rb") as f:
            parts.append(f.read())
    m.close()
    assert n_dev > 0  # k_sam wrote records (the host only formats what it leaves)
    assert b"".join(parts) == big["want"]
    assert tuple(total) == big["counters"]


def test_byte_range_shards_concatenate_to_the_whole(big):
    """One process per GPU on a plain four-line FASTQ file: rank r maps the records that start in byte range r of n
    (femhost_map_options.shard_index / shard_count) -- no counting pass, nothing parsed twice.  The parts in rank order
    are the single-process file; input of another shape is refused with FEMHOST_ENOTFASTQ4, never mapped in part."""
    from fem_b200 import host
    m = host.HostMapper(big["fa"], big["idx"], 1)
    for n in (2, 3, 8):
        parts, total = [], [0] * 5
        for r in range(n):
            out = str(big["d"] / ("bshard%d.sam" % r))
            st, _ = m.map_file(big["fq"], out, big["e"], big["a"], threads=3, batch_reads=1201, no_header=r > 0, shard=(r, n))
            assert st[0] == len(host.load_shard(big["fq"], r, n))
            total = [x + y for x, y in zip(total, st[:5])]
            with open(out, "rb") as f:
                parts.append(f.read())
        assert b"".join(parts) == big["want"], n
        assert tuple(total) == big["counters"]
    zipped = str(big["d"] / "shard_reads.fq.gz")
    with open(big["fq"], "rb") as f, gzip.open(zipped, "wb", compresslevel=1) as g:
        g.write(f.read())
    with pytest.raises(host.NotFastq4):
        m.map_file(zipped, str(big["d"] / "never.sam"), big["e"], big["a"], threads=2, shard=(0, 2))
    # the read-count fall-back every rank then takes: slices cut with the real parser, explicit "no reads" for an empty slice
    n_reads = host.count_records(zipped)
    assert n_reads == big["n_reads"]
    parts = []
    for r, (skip, count) in enumerate([(0, n_reads // 2), (n_reads // 2, n_reads - n_reads // 2), (n_reads, 0)]):
        out = str(big["d"] / ("cshard%d.sam" % r))
        st, _ = m.map_file(zipped, out, big["e"], big["a"], threads=3, skip_reads=skip, max_reads=count, no_header=r > 0, no_reads=count == 0)
        assert st[0] == count
        with open(out, "rb") as f:
            parts.append(f.read())
    assert b"".join(parts) == big["want"] and parts[2] == b""
    m.close()


def test_stream_output_is_written_once(big):
    """with a FIFO (or /dev/stdout) as the output the parallel reader's restart would send the header and
    the first batches twice; a stream output therefore takes the sequential reader.  CRLF records and a trailing
    blank line are the inputs that trigger the restart on a regular file."""
    import threading
    from fem_b200.host import FEM_BINARY
    d = big["d"]
    crlf = str(d / "crlf.fq")
    with open(big["fq"], "rb") as f, open(crlf, "wb") as g:
        g.write(f.read().replace(b"\n", b"\r\n") + b"\r\n")
    to_file = str(d / "crlf_file.sam")
    run_fem_map(big["fa"], big["idx"], crlf, to_file, big["e"], big["a"], threads=4, extra=("--batch", "800"))
    with open(to_file, "rb") as f:
        want = f.read()
    assert want.count(b"\n") == big["want"].count(b"\n") and want.startswith(b"@SQ")
    for src in (crlf, big["fq"]):
        fifo = str(d / "out.fifo")
        if os.path.exists(fifo):
            os.remove(fifo)
        os.mkfifo(fifo)
        got = []
        t = threading.Thread(target=lambda: got.append(open(fifo, "rb").read()))
        t.start()
        p = subprocess.run([FEM_BINARY, "map", "-e", str(big["e"]), "-a", str(big["a"]), "-t", "4", "--ref", big["fa"], "--index", big["idx"],
                            "--read1", src, "-o", fifo, "--batch", "800"], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        t.join(timeout=120)
        assert p.returncode == 0, p.stderr[-1000:]
        assert got and got[0] == (want if src == crlf else big["want"])


def test_cli_gpus2_matches_single(big):
    """The product's multi-GPU form: ONE process, N GPUs, one input, one ordered output (`FEM map --gpus 2`; replaces
    src/map.c:54 + src/output_queue.c:46-58).  Byte-identical to the single-GPU file with the index cloned device to
    device (default) and loaded per device (FEM_CLONE_RESIDENT=0); femgpu_clone_resident across two devices."""
    from fem_b200.api import load_library
    if load_library().femgpu_device_count() < 2:
        pytest.skip("needs two GPUs")
    from fem_b200.host import FEM_BINARY, HostMapper
    for clone in ("1", "0"):
        sam = str(big["d"] / ("gpus2_%s.sam" % clone))
        p = subprocess.run([FEM_BINARY, "map", "-e", str(big["e"]), "-a", str(big["a"]), "-t", "6", "--ref", big["fa"], "--index", big["idx"],
                            "--read1", big["fq"], "-o", sam, "--gpus", "2", "--batch", "700"], stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                           env={**os.environ, "FEM_CLONE_RESIDENT": clone}, text=True)
        assert p.returncode == 0, p.stderr[-2000:]
        assert "onto 2 GPU(s)" in p.stderr
        with open(sam, "rb") as f:
            assert f.read() == big["want"], clone
    m = HostMapper(big["fa"], big["idx"], 2, [1, 0])  # device order is the caller's
    out = str(big["d"] / "gpus2_api.sam")
    st, tm = m.map_file(big["fq"], out, big["e"], big["a"], threads=4, batch_reads=500)
    m.close()
    assert st[:5] == big["counters"] and tm["n_batches"] >= big["n_reads"] // 500
    with open(out, "rb") as f:
        assert f.read() == big["want"]


def test_reader_paths_agree(big):
    """parallel byte-range reader (default on plain four-line FASTQ), forced sequential reader, and
    the automatic fall-back for multi-line / gzip input all give the same SAM file."""
    import gzip as gz
    d, sam = big["d"], str(big["d"] / "rp.sam")
    env = dict(os.environ)
    from fem_b200.host import FEM_BINARY

    def run(fq, extra_env=None):
        p = subprocess.run([FEM_BINARY, "map", "-e", str(big["e"]), "-a", str(big["a"]), "-t", "6", "--ref", big["fa"],
                            "--index", big["idx"], "--read1", fq, "-o", sam, "--batch", "700"],
                           stdout=subprocess.PIPE, stderr=subprocess.PIPE, env={**env, **(extra_env or {})})
        assert p.returncode == 0, p.stderr[-2000:]
        with open(sam, "rb") as f:
            return f.read()

    assert run(big["fq"], {"FEM_READER": "kseq"}) == big["want"]
    # the two ways the chunks reach the file: copies into a shared mapping / positional writes
    assert run(big["fq"], {"FEM_SAM_MMAP": "1"}) == big["want"]
    assert run(big["fq"], {"FEM_SAM_MMAP": "0"}) == big["want"]
    if os.path.isdir("/dev/shm"):  # tmpfs picks the mapped form by itself
        shm = "/dev/shm/fem_test_%d.sam" % os.getpid()
        try:
            run_fem_map(big["fa"], big["idx"], big["fq"], shm, big["e"], big["a"], threads=5, extra=("--batch", "900"))
            with open(shm, "rb") as f:
                assert f.read() == big["want"]
        finally:
            if os.path.exists(shm):
                os.remove(shm)
    assert run(big["fq"], {"FEM_DEVICE_TRACEBACK": "0"}) == big["want"]  # every alignment on the host
    assert run(big["fq"], {"FEM_DEVICE_SAM": "0"}) == big["want"]  # device alignments, text built on the host
    # sequences wrapped over two lines: not four-line FASTQ -> the run restarts with the kseq reader
    wrapped = str(d / "wrapped.fq")
    with open(big["fq"], "rb") as f, open(wrapped, "wb") as g:
        lines = f.read().split(b"\n")
        for i in range(0, len(lines) - 3, 4):
            s, q = lines[i + 1], lines[i + 3]
            h = len(s) // 2 if (i // 4) % 50 == 49 else len(s)
            g.write(lines[i] + b"\n" + s[:h] + b"\n" + (s[h:] + b"\n" if h < len(s) else b"") + b"+\n" + q + b"\n")
    assert run(wrapped) == big["want"]
    zipped = str(d / "reads.fq.gz")
    with open(big["fq"], "rb") as f, gz.open(zipped, "wb", compresslevel=1) as g:
        g.write(f.read())
    assert run(zipped) == big["want"]


def test_cli_ragged_low_complexity_and_empty_inputs(tmp_path):
    """the whole engine on a file that mixes read lengths from 12 to 1024 bases, N-only reads, duplicates and
    homopolymer / tandem-repeat reads with thousands of records (device SAM text, k_sam_big_*, host formatter
    and global-scratch candidates in one run), then on a file with one read and on an empty file"""
    import numpy as np
    from fem_b200.host import FEM_BINARY
    rng = np.random.default_rng(77)
    records = synth.random_reference(123, [150_000, 80_000, 2_000], 0.15)
    tail = bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), 20_000))
    records.append(("lowcomplexity", b"A" * 3000 + b"ACGT" * 900 + b"AC" * 1500 + tail))
    low = records[3][1]
    reads = synth.simulate_reads(11, records[:3], 1500, 100, 3, lengths=[12, 20, 35, 48, 64, 100, 101, 150, 251, 400, 1024])
    reads += [("polyA", b"A" * 100, b"I" * 100), ("allN", b"N" * 80, b"#" * 80), ("acgt", b"ACGT" * 25, b"I" * 100)]
    reads += [("t%d" % i, low[3000 + 4 * i:3100 + 4 * i], b"5" * 100) for i in range(60)]
    reads += [("dup%d" % i, reads[7][1], reads[7][2]) for i in range(50)]
    order = rng.permutation(len(reads))
    reads = [reads[i] for i in order]
    fa, fq, idx = str(tmp_path / "ref.fa"), str(tmp_path / "reads.fq"), str(tmp_path / "idx")
    synth.write_fasta(fa, records)
    synth.write_fastq(fq, reads)
    subprocess.run([FEM_BINARY, "index", "12", "3", fa, idx], check=True, stderr=subprocess.DEVNULL)
    o = ol.Oracle(fasta=fa)
    o.load_index(idx)
    o.set_params(3, 1)

    def both(fq_path, tag, **kw):
        want_sam, got_sam = str(tmp_path / (tag + ".want.sam")), str(tmp_path / (tag + ".sam"))
        st = o.map_file(fq_path, want_sam)
        counters = run_fem_map(fa, idx, fq_path, got_sam, 3, 1, **kw)
        assert counters == st.as_tuple()[:5]
        with open(want_sam, "rb") as f, open(got_sam, "rb") as g:
            assert g.read() == f.read(), tag
        return counters

    c = both(fq, "mixed", threads=6, extra=("--batch", "257"))
    assert c[1] > 500 and c[4] > 5000
    both(fq, "mixed_one_batch", threads=2)
    one = str(tmp_path / "one.fq")
    synth.write_fastq(one, reads[:1])
    both(one, "one")
    empty = str(tmp_path / "empty.fq")
    open(empty, "wb").close()
    want_sam, got_sam = str(tmp_path / "empty.want.sam"), str(tmp_path / "empty.sam")
    p = subprocess.run([FEM_BINARY, "map", "-e", "3", "-t", "2", "--ref", fa, "--index", idx, "--read1", empty, "-o", got_sam],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert p.returncode == 0, p.stderr[-1000:]
    with open(got_sam, "rb") as g:
        body = g.read()
    assert body.count(b"\n") == len(records) and all(line.startswith(b"@SQ") for line in body.splitlines())  # src/output_queue.c:93-116
    o.close()
