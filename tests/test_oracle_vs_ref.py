"""Pins the CPU restatement (oracle/fem_oracle.c) to the reference itself.

Runs wherever oracle/_ref/FEM_ref_{asis,hdr} exist (they are compiled from /root/reference by
oracle/Makefile in the build container and travel to the GPU box as prebuilt files): the reference
binary builds the index and maps seeded synthetic reads with `-t 1`; the restatement must produce
the byte-identical SAM file and the same five stderr counters (src/FEM_map.c:214-218).
The cases follow SURVEY.md section 4: e 0..7, a 0..2, candidate counts around multiples of 8 (F6),
>64 mappings per read (F8), equal-frequency seeds (F9), reads at record edges, step != 3 (F1).
"""
import os

import pytest

import oracle_lib as ol
import synth

needs_ref = pytest.mark.skipif(not ol.have_reference_binaries(), reason="oracle/_ref reference binaries not built")

# (tag, record lengths, repeat_frac, n_reads, read length, e, a, k, step, binary)
CASES = [
    ("asis_e3_s3", [120_000, 60_000], 0.05, 1500, 100, 3, 1, 12, 3, "asis"),
    ("asis_e5_s3_rep", [80_000, 50_000], 0.40, 600, 100, 5, 1, 12, 3, "asis"),
    ("asis_e2_a0", [90_000], 0.30, 800, 100, 2, 0, 12, 3, "asis"),
    ("asis_e3_a2", [90_000, 30_000], 0.30, 800, 100, 3, 2, 12, 3, "asis"),
    ("asis_e0", [90_000], 0.20, 500, 100, 0, 1, 12, 3, "asis"),
    ("asis_e7_l250", [150_000, 50_000], 0.30, 300, 250, 7, 1, 12, 3, "asis"),
    ("hdr_cfg1_e3_s5", [200_000, 150_000, 100_000], 0.0, 2000, 100, 3, 1, 12, 5, "hdr"),
    ("hdr_cfg2_e5_s4", [200_000, 100_000], 0.0, 1000, 100, 5, 1, 12, 4, "hdr"),
    ("hdr_cfg3_e7_s5", [200_000, 100_000], 0.1, 600, 150, 7, 1, 12, 5, "hdr"),
    ("hdr_cfg4_e7_s2_rep", [100_000, 60_000], 0.45, 300, 150, 7, 1, 12, 2, "hdr"),
    ("hdr_cfg4_e5_s2_rep", [100_000, 60_000], 0.45, 400, 100, 5, 1, 12, 2, "hdr"),
    ("hdr_cfg5_k14_s10", [300_000, 200_000], 0.05, 400, 250, 7, 1, 14, 10, "hdr"),
    ("hdr_k10_s4_e1", [60_000, 30_000], 0.2, 500, 80, 1, 1, 10, 4, "hdr"),
    # step 5 index read by the unmodified binary: degenerate (nearly empty) output must match too (F1)
    ("asis_on_s5_index", [100_000], 0.0, 500, 100, 3, 1, 12, 5, "asis"),
]


def make_inputs(tmp_path, tag, lens, rep, n_reads, l, e):
    seed = sum(map(ord, tag))
    records = synth.random_reference(seed, lens, rep)
    reads = synth.simulate_reads(seed + 1, records, n_reads, l, e)
    # reads hugging both record edges exercise the range filter (src/filter.c:140)
    for ri, (name, seq) in enumerate(records):
        for off in range(0, 2 * e + 3):
            if len(seq) >= l + off:
                reads.append(("edgeL_%d_%d" % (ri, off), seq[off:off + l], b"I" * l))
                reads.append(("edgeR_%d_%d" % (ri, off), seq[len(seq) - l - off:len(seq) - off], b"I" * l))
    fa, fq = str(tmp_path / "ref.fa"), str(tmp_path / "reads.fq")
    synth.write_fasta(fa, records)
    synth.write_fastq(fq, reads)
    return fa, fq, len(reads)


@needs_ref
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_sam_and_counters_identical(case, tmp_path):
    tag, lens, rep, n_reads, l, e, a, k, step, which = case
    fa, fq, n_total = make_inputs(tmp_path, tag, lens, rep, n_reads, l, e)
    binary = ol.REF_ASIS if which == "asis" else ol.REF_HDR
    idx = str(tmp_path / "idx")
    ol.run_reference_index(binary, k, step, fa, idx)
    ref_sam = str(tmp_path / "ref.sam")
    ref_counters, _ = ol.run_reference(binary, fa, idx, fq, ref_sam, e, a, threads=1)

    o = ol.Oracle(fasta=fa)
    o.load_index(idx)
    if which == "asis":
        o.set_params(e, a, 12, 3)  # the unmodified binary never reads the header
    else:
        o.set_params(e, a)
    my_sam = str(tmp_path / "oracle.sam")
    st = o.map_file(fq, my_sam)
    assert st.as_tuple()[:5] == ref_counters
    assert st.num_infeasible_reads == 0
    assert ref_counters[0] == n_total
    with open(ref_sam, "rb") as f1, open(my_sam, "rb") as f2:
        a_, b_ = f1.read(), f2.read()
    if a_ != b_:
        la, lb = a_.split(b"\n"), b_.split(b"\n")
        for i, (x, y) in enumerate(zip(la, lb)):
            assert x == y, "first differing SAM line %d:\nref: %r\nour: %r" % (i, x[:300], y[:300])
        assert len(la) == len(lb)
    if which != "asis" or step == 3:
        assert ref_counters[1] > 0.5 * n_reads  # the case is not degenerate

    # the oracle's own index construction must reproduce the reference's index file byte for byte
    o2 = ol.Oracle(fasta=fa)
    o2.build_index(k, step)
    idx2 = str(tmp_path / "idx2")
    o2.save_index(idx2)
    with open(idx, "rb") as f1, open(idx2, "rb") as f2:
        assert f1.read() == f2.read()
    o.close()
    o2.close()


@needs_ref
def test_thread_count_changes_only_read_order(tmp_path):
    """-t 4 output equals -t 1 output as a multiset of per-read record groups."""
    fa, fq, _ = make_inputs(tmp_path, "threads", [150_000], 0.2, 3000, 100, 3)
    idx = str(tmp_path / "idx")
    ol.run_reference_index(ol.REF_ASIS, 12, 3, fa, idx)
    c1, _ = ol.run_reference(ol.REF_ASIS, fa, idx, fq, str(tmp_path / "t1.sam"), 3, 1, threads=1)
    c4, _ = ol.run_reference(ol.REF_ASIS, fa, idx, fq, str(tmp_path / "t4.sam"), 3, 1, threads=4)
    assert c1 == c4

    def groups(path):
        out, cur, name = [], [], None
        for line in open(path, "rb"):
            if line.startswith(b"@"):
                continue
            q = line.split(b"\t", 1)[0]
            if q != name and cur:
                out.append(b"".join(cur))
                cur = []
            name = q
            cur.append(line)
        if cur:
            out.append(b"".join(cur))
        return sorted(out)

    assert groups(str(tmp_path / "t1.sam")) == groups(str(tmp_path / "t4.sam"))
