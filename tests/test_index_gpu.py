"""GPU tests of index construction on the device (femgpu_build_index, `FEM index ... --gpu`): the
lookup table and occurrence table must equal the reference's own (src/index.c:57-98) entry for
entry -- checked against the oracle restatement (pinned to the reference binaries' index files by
tests/test_oracle_vs_ref.py), against the host builder, and against the golden fixtures' SAM through
an index that never left the device."""
import glob
import gzip
import json
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as ol
import synth

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")


def _poly(records_seed, lens, runs):
    """random records with long homopolymer / tandem runs: buckets far beyond the warp and block sort paths"""
    recs = synth.random_reference(records_seed, lens, 0.1)
    out = []
    rng = np.random.default_rng(records_seed)
    for (name, seq), (unit, copies) in zip(recs, runs):
        s = bytearray(seq)
        p = int(rng.integers(0, len(s) - len(unit) * copies))
        s[p:p + len(unit) * copies] = unit * copies
        out.append((name, bytes(s)))
    return out


# (tag, records, k, step)
def _cases():
    yield "k12_s5", synth.random_reference(11, [300_000, 200_000, 150_000], 0.0), 12, 5
    yield "k12_s4_rep", synth.random_reference(12, [400_000, 250_000], 0.3), 12, 4
    yield "k12_s1", synth.random_reference(13, [120_000, 80_000], 0.2), 12, 1
    yield "k14_s10", synth.random_reference(14, [600_000, 500_000], 0.0), 14, 10
    yield "k10_s3_N", synth.random_reference(15, [200_000, 100_000], 0.1, 0.01), 10, 3
    yield "k8_s2_dense", synth.random_reference(16, [500_000, 300_000], 0.0), 8, 2          # every bucket 3-10 entries
    yield "k6_s1_block", synth.random_reference(17, [700_000, 400_000], 0.0), 6, 1          # every bucket ~270: warp bitonic
    yield "k4_s1_huge", synth.random_reference(18, [900_000, 600_000], 0.0), 4, 1           # every bucket ~5900: block path
    yield "k3_s1_global", synth.random_reference(19, [900_000, 600_000], 0.0), 3, 1         # every bucket ~23000: global scratch
    yield "polyA", _poly(20, [300_000, 200_000], [(b"A", 40_000), (b"AC", 6_000)]), 12, 1   # one bucket of ~40k, two of ~6k
    yield "short_records", [("a", b"ACGTACGTAC"), ("b", b"ACG"), ("c", b"ACGTACGTACGTA"), ("d", b"A" * 12)], 12, 3
    yield "lower_case", [("a", synth.random_reference(21, [50_000])[0][1].lower())], 11, 2


CASES = list(_cases())


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_device_index_equals_oracle(case):
    from fem_b200.api import Mapper
    tag, records, k, step = case
    oracle = ol.Oracle(records=[(n, s.decode()) for n, s in records])
    oracle.build_index(k, step)
    ok, ostep, olut, oocc = oracle.index_arrays()
    m = Mapper(0)
    try:
        m.upload_reference(records)
        m.build_index(k, step)
        gk, gstep, lut, occ = m.download_index()
        assert (gk, gstep) == (ok, ostep) == (k, step)
        assert lut.shape == olut.shape and np.array_equal(lut, olut)
        assert occ.shape == oocc.shape and np.array_equal(occ, oocc)
        # building twice on one context (smaller, then the first again) leaves no stale state
        m.build_index(min(k, 9), step + 1)
        m.build_index(k, step)
        _, _, lut2, occ2 = m.download_index()
        assert np.array_equal(lut2, olut) and np.array_equal(occ2, oocc)
    finally:
        m.close()
        oracle.close()


def test_build_needs_reference_and_valid_parameters():
    from fem_b200.api import FemGpuError, Mapper
    m = Mapper(0)
    try:
        with pytest.raises(FemGpuError):
            m.build_index(12, 3)          # no reference yet
        m.upload_reference([("a", b"ACGT" * 100)])
        for k, step in ((0, 3), (16, 3), (12, 0)):
            with pytest.raises(FemGpuError):
                m.build_index(k, step)
        with pytest.raises(FemGpuError):
            m.download_index()            # failed builds leave no index behind
    finally:
        m.close()


def test_cli_index_gpu_writes_the_same_file(tmp_path):
    """`FEM index k step ref out --gpu` vs the host builder vs the oracle's save_index: identical bytes"""
    from fem_b200.host import FEM_BINARY
    records = synth.random_reference(31, [700_000, 300_000, 50], 0.15, 0.001)
    fa = str(tmp_path / "ref.fa")
    synth.write_fasta(fa, records)
    host_idx, gpu_idx, env_idx, orc_idx = (str(tmp_path / n) for n in ("host.idx", "gpu.idx", "env.idx", "oracle.idx"))
    subprocess.run([FEM_BINARY, "index", "12", "4", fa, host_idx], check=True, stderr=subprocess.DEVNULL)
    p = subprocess.run([FEM_BINARY, "index", "12", "4", fa, gpu_idx, "--gpu"], stderr=subprocess.PIPE, text=True)
    assert p.returncode == 0, p.stderr
    assert "Built index in" in p.stderr and "Saved index successfully" in p.stderr
    subprocess.run([FEM_BINARY, "index", "12", "4", fa, env_idx], check=True, stderr=subprocess.DEVNULL,
                   env=dict(os.environ, FEM_INDEX_GPU="0"))
    oracle = ol.Oracle(records=[(n, s.decode()) for n, s in records])
    oracle.build_index(12, 4)
    oracle.save_index(orc_idx)
    oracle.close()
    want = open(orc_idx, "rb").read()
    assert open(host_idx, "rb").read() == want
    assert open(gpu_idx, "rb").read() == want
    assert open(env_idx, "rb").read() == want


@pytest.mark.parametrize("tag", sorted(os.path.basename(p)[:-5] for p in glob.glob(os.path.join(GOLD, "hdr_*.json"))))
def test_golden_sam_through_a_device_built_index(tag, tmp_path):
    """reference -> HBM -> femgpu_build_index -> map: the golden SAM records without an index file"""
    from fem_b200.api import Mapper, sort_mappings
    with open(os.path.join(GOLD, tag + ".json")) as f:
        meta = json.load(f)
    records, name, seq = [], None, []
    with gzip.open(os.path.join(GOLD, tag + ".ref.fa.gz"), "rb") as f:
        for line in f.read().split(b"\n"):
            if line.startswith(b">"):
                if name is not None:
                    records.append((name, b"".join(seq)))
                name, seq = line[1:].split()[0].decode(), []
            elif line:
                seq.append(line)
    records.append((name, b"".join(seq)))
    with gzip.open(os.path.join(GOLD, tag + ".reads.fq.gz"), "rb") as f:
        lines = f.read().split(b"\n")
    reads = [(lines[i][1:].split()[0], lines[i + 1]) for i in range(0, len(lines) - 3, 4)]
    want = {}
    with gzip.open(os.path.join(GOLD, tag + ".sam.gz"), "rb") as f:
        for line in f:
            if not line.startswith(b"@"):
                fld = line.split(b"\t")
                nm = [x for x in fld[11:] if x.startswith(b"NM:i:")][0]
                want.setdefault(fld[0], []).append((int(fld[1]) & 16, fld[2].decode(), int(nm[5:])))
    m = Mapper(0)
    try:
        m.upload_reference(records)
        m.build_index(meta["k"], meta["step"])
        first, maps, stats = m.map_batch([r[1] for r in reads], meta["e"], meta["a"])
        c = meta["counters"]
        assert stats[:5] == (c["num_reads"], c["num_mapped_reads"], c["num_candidates_before_filter"], c["num_candidates"], c["num_mappings"])
        for i, (qname, _) in enumerate(reads):
            mine = maps[int(first[i]):int(first[i + 1])].copy()
            sort_mappings(mine)
            got = [(16 if int(x["direction"]) else 0, records[int(x["candidate_position"]) >> 32][0], int(x["edit_distance"])) for x in mine]
            assert got == want.get(qname, []), qname
    finally:
        m.close()
