"""CPU tests against the committed golden fixtures (tests/golden/, produced by the reference
binaries -- see make_golden.py):
  * the oracle restatement reproduces every golden SAM byte for byte, plus the counters and the
    index file's sha256;
  * the product's host side (fem_b200/host/*.c through libfemhost.so) does the same for the parts
    that run on the host: FASTA/FASTQ parsing, `FEM index`, mapping order + traceback + SAM text
    (fed with the oracle's mappings in emission order -- no GPU here);
  * both shared libraries export every symbol their headers declare.
"""
import glob
import gzip
import hashlib
import json
import os
import re
import subprocess

import numpy as np
import pytest

import oracle_lib as ol

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = os.path.join(HERE, "golden")
TAGS = sorted(os.path.basename(p)[:-5] for p in glob.glob(os.path.join(GOLD, "*.json")))


def gold(tag):
    with open(os.path.join(GOLD, tag + ".json")) as f:
        meta = json.load(f)
    fa, fq = (os.path.join(GOLD, "%s.%s" % (tag, s)) for s in ("ref.fa.gz", "reads.fq.gz"))
    with gzip.open(os.path.join(GOLD, tag + ".sam.gz"), "rb") as f:
        sam = f.read()
    return meta, fa, fq, sam


def oracle_for(meta, fa):
    o = ol.Oracle(fasta=fa)
    o.build_index(meta["k"], meta["step"])
    if meta["binary"] == "asis":
        o.set_params(meta["e"], meta["a"], 12, 3)
    else:
        o.set_params(meta["e"], meta["a"])
    return o


def parse_fastq_gz(path):
    out = []
    with gzip.open(path, "rb") as f:
        lines = f.read().split(b"\n")
    for i in range(0, len(lines) - 3, 4):
        out.append((lines[i][1:].split()[0], lines[i + 1], lines[i + 3]))
    return out


def test_fixtures_present():
    assert len(TAGS) >= 6


@pytest.mark.parametrize("tag", TAGS)
def test_oracle_reproduces_golden(tag, tmp_path):
    meta, fa, fq, sam = gold(tag)
    o = oracle_for(meta, fa)
    idx = str(tmp_path / "idx")
    o.save_index(idx)
    with open(idx, "rb") as f:
        assert hashlib.sha256(f.read()).hexdigest() == meta["index_sha256"]
    out = str(tmp_path / "o.sam")
    st = o.map_file(fq, out)
    c = meta["counters"]
    assert st.as_tuple()[:5] == (c["num_reads"], c["num_mapped_reads"], c["num_candidates_before_filter"],
                                 c["num_candidates"], c["num_mappings"])
    with open(out, "rb") as f:
        assert f.read() == sam
    o.close()


@pytest.mark.parametrize("tag", TAGS)
def test_host_index_matches_golden(tag, tmp_path):
    from fem_b200 import host
    meta, fa, _, _ = gold(tag)
    ref = host.load_all(fa)
    ix = host.Index.build(ref, meta["k"], meta["step"])
    p = str(tmp_path / "idx")
    ix.save(p)
    with open(p, "rb") as f:
        assert hashlib.sha256(f.read()).hexdigest() == meta["index_sha256"]
    ix2 = host.Index.load(p)
    k, step, lut, occ = ix.arrays()
    k2, step2, lut2, occ2 = ix2.arrays()
    assert (k, step) == (k2, step2) and np.array_equal(lut, lut2) and np.array_equal(occ, occ2)
    # the command line writes the same file
    p3 = str(tmp_path / "idx3")
    subprocess.run([host.FEM_BINARY, "index", str(meta["k"]), str(meta["step"]), fa, p3], check=True,
                   stderr=subprocess.DEVNULL)
    with open(p3, "rb") as f:
        assert hashlib.sha256(f.read()).hexdigest() == meta["index_sha256"]


@pytest.mark.parametrize("tag", TAGS)
def test_host_reader_matches_python_parse(tag):
    from fem_b200 import host
    _, fa, fq, _ = gold(tag)
    expect = parse_fastq_gz(fq)
    got = [rec for batch in host.read_batches(fq, 64) for rec in batch]
    assert got == expect
    ref = host.load_all(fa)
    with gzip.open(fa, "rb") as f:
        chunks = f.read().split(b">")[1:]
    assert len(ref) == len(chunks)
    for i, ch in enumerate(chunks):
        head, _, body = ch.partition(b"\n")
        assert ref.name(i) == head.split()[0]
        assert ref.seq(i) == body.replace(b"\n", b"")


@pytest.mark.parametrize("tag", TAGS)
def test_host_sam_text_matches_golden(tag):
    """mapping order (src/align.c:57), traceback, CIGAR/MD and SAM text of the product's host code."""
    from fem_b200 import host
    from fem_b200.api import MAPPING_DTYPE
    meta, fa, fq, sam = gold(tag)
    o = oracle_for(meta, fa)
    ref = host.load_all(fa)
    text = [host.sam_header(ref)]
    most = 0
    for name, seq, qual in parse_fastq_gz(fq):
        emitted = []
        for d, strand in enumerate((seq, ol.revcomp(seq))):
            n, cands, _ = o.candidates(strand)
            if n > 0:
                emitted.extend(o.verify(strand, d, cands))
        maps = np.zeros(len(emitted), dtype=MAPPING_DTYPE)
        for j, (cand, end_off, direction, dist) in enumerate(emitted):
            maps[j] = (cand, end_off, direction, dist, (0, 0))
        most = max(most, len(emitted))
        text.append(host.sam_read(ref, name, seq, qual, meta["e"], maps))
    assert b"".join(text) == sam
    if tag == "hdr_e7_s2_rep":
        assert most > 64  # the in-place radix sort path of src/ksort.h:114-146 is exercised
    o.close()


def test_reader_edge_cases(tmp_path):
    """multi-line FASTA, comments, CRLF, blank lines, empty records, FASTQ after FASTA headers."""
    from fem_b200 import host
    p = str(tmp_path / "x.fa")
    with open(p, "wb") as f:
        f.write(b"junk before\n>r1 comment here\nACGT\nAC\n\nGT\n>empty\n>r2\tc\r\nAAAA\r\nCC\r\n@q1 x\nACGTN\n+\nIIIII\n@q2\nAC\nGT\n+q2\nII\nII\n")
    got = host.load_all(p).as_list()
    assert got == [(b"r1", b"ACGTACGT", None), (b"r2", b"AAAACC", None), (b"q1", b"ACGTN", b"IIIII"), (b"q2", b"ACGT", b"IIII")]
    bad = str(tmp_path / "bad.fq")
    with open(bad, "wb") as f:
        f.write(b"@q1\nACGT\n+\nII\n")
    with pytest.raises(Exception):
        host.load_all(bad)
    with pytest.raises(Exception):
        host.load_all(str(tmp_path / "missing.fa"))


def _declared(header):
    with open(os.path.join(ROOT, "include", header)) as f:
        src = f.read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = set(re.findall(r"\b(fem(?:gpu|host)_[a-z0-9_]+)\s*\(", src))
    inline = set(re.findall(r"static inline[^;{]*?\b(fem(?:gpu|host)_[a-z0-9_]+)\s*\(", src))
    return sorted(names - inline)


def test_libraries_export_every_declared_symbol():
    import ctypes
    from fem_b200.api import load_library
    from fem_b200.host import load_host_library
    dev, hst = load_library(), load_host_library()
    gpu_syms, host_syms = _declared("femgpu.h"), _declared("femhost.h")
    assert len(gpu_syms) >= 20 and len(host_syms) >= 18
    for s in gpu_syms:
        assert isinstance(getattr(dev, s), ctypes._CFuncPtr), s
    for s in host_syms:
        if s.startswith("femhost_"):
            assert isinstance(getattr(hst, s), ctypes._CFuncPtr), s
    assert dev.femgpu_abi_version() >= 1


def test_no_fallback_without_a_device():
    """Without a CUDA device the product must fail loudly, never compute on the host."""
    from fem_b200.api import FemGpuError, Mapper, load_library
    if load_library().femgpu_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(FemGpuError):
        Mapper(0)
    from fem_b200 import host
    _, fa, _, _ = gold(TAGS[0])
    with pytest.raises(FemGpuError):
        host.HostMapper(fa, fa, 1)


def test_parallel_reader_ranges_equal_sequential(tmp_path):
    """Byte-range parsing with record resynchronisation (engine.c's parallel reader) must give the
    sequential reader's records for every chunk size, also when quality lines begin with '@' or
    '+', and must decline anything that is not plain four-line FASTQ."""
    import random
    from fem_b200 import host
    rng = random.Random(5)
    p = str(tmp_path / "r.fq")
    with open(p, "wb") as f:
        for i in range(3000):
            l = rng.choice([1, 2, 36, 100, 151])
            seq = "".join(rng.choice("ACGTN") for _ in range(l))
            qual = "".join(rng.choice("@+I#>5") for _ in range(l))
            f.write(("@r%d some comment\n%s\n+%s\n%s\n" % (i, seq, "r%d" % i if i % 7 == 0 else "", qual)).encode())
    want = host.load_all(p).as_list()
    assert len(want) == 3000
    for chunk in (1, 7, 64, 1000, 4096, 100_000, 10_000_000):
        got = host.load_all_ranges(p, chunk)
        assert got is not None and got.as_list() == want, chunk
    for bad in (b"@a\nAC\nGT\n+\nIIII\n@b\nAC\n+\nII\n", b"@a\r\nAC\r\n+\r\nII\r\n" * 50, b">a\nACGT\n>b\nAC\n",
                b"x\n@a\nAC\n+\nII\n"):
        q = str(tmp_path / "bad.fq")
        with open(q, "wb") as f:
            f.write(bad)
        assert host.load_all_ranges(q, 16) is None
        assert host.load_shard(q, 1, 3) is None or bad.startswith(b"@a\nAC\nGT")  # the cut itself may land on clean lines
    # byte-range shards (femhost_map_options.shard_index / shard_count): a partition of the file in order
    for count in (1, 2, 3, 8, 64, 5000):
        got = []
        for i in range(count):
            part = host.load_shard(p, i, count)
            assert part is not None
            got += part.as_list()
        assert got == want, count


def test_count_records_uses_the_real_parser(tmp_path):
    """`newlines // 4` miscounts FASTA, gzip, multi-line records and a missing final newline."""
    import gzip
    from fem_b200 import host
    cases = {
        "plain.fq": (b"@a\nACGT\n+\nIIII\n@b\nAC\n+\nII\n", 2),
        "nofinal.fq": (b"@a\nACGT\n+\nIIII\n@b\nAC\n+\nII", 2),
        "multi.fq": (b"@a\nAC\nGT\n+\nII\nII\n@b\nAC\n+\nII\n", 2),
        "reads.fa": (b">a\nACGT\n>b\nAC\nGT\n>c\nA\n", 3),
        "empty_record.fa": (b">a\nACGT\n>b\n>c\nA\n", 2),  # src/sequence_batch.c:50-52: zero-length records are skipped
        "crlf.fq": (b"@a\r\nAC\r\n+\r\nII\r\n" * 5, 5),
        "none.fq": (b"", 0),
    }
    for name, (data, n) in cases.items():
        q = str(tmp_path / name)
        with open(q, "wb") as f:
            f.write(data)
        assert host.count_records(q) == n, name
        assert len(host.load_all(q)) == n, name
        with gzip.open(q + ".gz", "wb") as f:
            f.write(data)
        assert host.count_records(q + ".gz") == n, name + ".gz"


@pytest.mark.parametrize("n", [1, 2, 40, 64, 65, 300, 70_000])
def test_sort_with_alignments_follows_the_mapping_permutation(n):
    """femgpu_sort_mappings_aln = femgpu_sort_mappings on the records, alignments carried along;
    keys tie on purpose (src/ksort.h's in-place radix sort is not stable above 64 records)."""
    from fem_b200.api import ALIGNMENT_DTYPE, MAPPING_DTYPE, sort_mappings, sort_mappings_aln
    rng = np.random.default_rng(n)
    maps = np.zeros(n, dtype=MAPPING_DTYPE)
    maps["candidate_position"] = rng.integers(0, 50, size=n).astype(np.uint64) + (rng.integers(0, 2, size=n).astype(np.uint64) << np.uint64(32))
    maps["end_position_offset"] = rng.integers(99, 104, size=n)
    maps["direction"] = rng.integers(0, 2, size=n)
    maps["edit_distance"] = rng.integers(0, 4, size=n)
    alns = np.zeros(n, dtype=ALIGNMENT_DTYPE)
    alns["start"] = np.arange(n) % 30000
    alns["cigar"][:, 0] = np.arange(n) % 65521
    alns["cigar"][:, 1] = np.arange(n) // 65521
    ident = (alns["cigar"][:, 0].astype(np.int64) + 65521 * alns["cigar"][:, 1].astype(np.int64)).copy()
    want = sort_mappings(maps.copy())
    got_maps, got_alns = sort_mappings_aln(maps.copy(), alns.copy())
    assert np.array_equal(got_maps, want)
    src = got_alns["cigar"][:, 0].astype(np.int64) + 65521 * got_alns["cigar"][:, 1].astype(np.int64)
    assert sorted(src.tolist()) == ident.tolist()
    for f in ("candidate_position", "end_position_offset", "direction", "edit_distance"):
        assert np.array_equal(maps[f][src], got_maps[f])  # alignment i still belongs to mapping i
