"""INTEGRATION.md section B, executed: oracle/_ref/FEM_ref_b200 is the REFERENCE's own program (main, argument parsing,
input/output queues, process_mappings, htslib writer; compiled from /root/reference by oracle/Makefile) with one file
swapped -- oracle/stub/map_b200.c instead of src/map.c, i.e. one femgpu_map_batch call per batch where the reference
calls generate_group_seeding_candidates / verify_candidates per read and strand (src/map.c:27-49).  Its SAM file must
equal the patched reference binary's (-t 1: input order) byte for byte, and the five counters too."""
import os
import re
import subprocess

import pytest

import oracle_lib as ol
import synth

pytestmark = pytest.mark.gpu

STUB = os.path.join(os.path.dirname(ol.REF_HDR), "FEM_ref_b200")
COUNTERS = ("The number of read:", "The number of mapped read:", "The number of candidate before additional q-gram filter:",
            "The number of candidate:", "The number of mapping:")


def run(binary, fa, idx, fq, sam, e, a, threads):
    p = subprocess.run([binary, "map", "-e", str(e), "-a", str(a), "-t", str(threads), "--ref", fa, "--index", idx, "--read1", fq, "-o", sam],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert p.returncode == 0, p.stderr[-2000:]
    return tuple(int(re.search(re.escape(t) + r" (\d+)", p.stderr).group(1)) for t in COUNTERS)


@pytest.mark.parametrize("tag,lens,rep,n_reads,l,e,a,k,step", [
    ("stub_cfg3", [600_000, 300_000], 0.05, 12_000, 150, 7, 1, 12, 5),   # several input-queue batches of 10 000 reads
    ("stub_cfg2", [500_000], 0.0, 3_000, 100, 5, 1, 12, 4),
    ("stub_rep", [200_000, 100_000], 0.4, 1_500, 150, 7, 2, 12, 2),       # many mappings per read, a = 2
])
def test_reference_main_with_swapped_mapper_thread(tmp_path, tag, lens, rep, n_reads, l, e, a, k, step):
    if not (os.access(STUB, os.X_OK) and ol.have_reference_binaries()):
        pytest.skip("oracle/_ref/FEM_ref_b200 / FEM_ref_hdr not built (make -C oracle ref stub needs /root/reference)")
    records = synth.random_reference(sum(map(ord, tag)), lens, rep)
    reads = synth.simulate_reads(sum(map(ord, tag)) + 1, records, n_reads, l, e)
    fa, fq, idx = str(tmp_path / "ref.fa"), str(tmp_path / "reads.fq"), str(tmp_path / "idx")
    synth.write_fasta(fa, records)
    synth.write_fastq(fq, reads)
    subprocess.run([ol.REF_HDR, "index", str(k), str(step), fa, idx], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    want_sam, got_sam = str(tmp_path / "ref.sam"), str(tmp_path / "b200.sam")
    want = run(ol.REF_HDR, fa, idx, fq, want_sam, e, a, 1)
    got = run(STUB, fa, idx, fq, got_sam, e, a, 1)
    assert got == want
    with open(want_sam, "rb") as f, open(got_sam, "rb") as g:
        assert f.read() == g.read()
    assert want[1] > n_reads // 2
