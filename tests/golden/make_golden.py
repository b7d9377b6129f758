"""Generates the golden fixtures in this directory from the REFERENCE ITSELF.

Run in the build container (needs oracle/_ref/FEM_ref_{asis,hdr}, compiled from /root/reference by
`make -C oracle ref`):   python tests/golden/make_golden.py
For every case: seeded synthetic reference + reads (tests/synth.py) -> <case>.ref.fa.gz,
<case>.reads.fq.gz; the reference binary builds the index and maps with -t 1 -> <case>.sam.gz;
<case>.json holds the parameters, the five stderr counters and the sha256 of the index file.
"asis" = the unmodified reference binary, "hdr" = the same sources plus the two-line patch that
copies the index header's window/step into the arguments (oracle/Makefile).
"""
import gzip
import hashlib
import json
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as ol  # noqa: E402
import synth  # noqa: E402

# tag, record lengths, repeat_frac, n_reads, read length, e, a, k, step, binary
CASES = [
    ("asis_e3_s3", [40_000, 25_000], 0.15, 250, 100, 3, 1, 12, 3, "asis"),
    ("asis_e2_a0", [30_000], 0.30, 200, 100, 2, 0, 12, 3, "asis"),
    ("asis_e3_a2", [30_000], 0.30, 200, 100, 3, 2, 12, 3, "asis"),
    ("hdr_e5_s4", [50_000], 0.10, 200, 100, 5, 1, 12, 4, "hdr"),
    ("hdr_e7_s5_l150", [60_000], 0.10, 150, 150, 7, 1, 12, 5, "hdr"),
    ("hdr_e7_s2_rep", [40_000], 0.50, 100, 150, 7, 1, 12, 2, "hdr"),
]


def main():
    assert ol.have_reference_binaries(), "build oracle/_ref first: make -C oracle ref"
    for tag, lens, rep, n_reads, l, e, a, k, step, which in CASES:
        seed = sum(map(ord, tag))
        records = synth.random_reference(seed, lens, rep)
        reads = synth.simulate_reads(seed + 1, records, n_reads, l, e)
        for ri, (name, seq) in enumerate(records):  # reads hugging the record edges (range filter)
            for off in (0, 1, e, e + 1, 2 * e + 1):
                reads.append(("edgeL_%d_%d" % (ri, off), seq[off:off + l], b"I" * l))
                reads.append(("edgeR_%d_%d" % (ri, off), seq[len(seq) - l - off:len(seq) - off], b"I" * l))
        with tempfile.TemporaryDirectory() as tmp:
            fa, fq, idx, sam = (os.path.join(tmp, n) for n in ("ref.fa", "reads.fq", "idx", "out.sam"))
            synth.write_fasta(fa, records)
            synth.write_fastq(fq, reads)
            binary = ol.REF_ASIS if which == "asis" else ol.REF_HDR
            ol.run_reference_index(binary, k, step, fa, idx)
            counters, _ = ol.run_reference(binary, fa, idx, fq, sam, e, a, threads=1)
            for src, dst in ((fa, "ref.fa.gz"), (fq, "reads.fq.gz"), (sam, "sam.gz")):
                with open(src, "rb") as f, open(os.path.join(HERE, "%s.%s" % (tag, dst)), "wb") as raw:
                    with gzip.GzipFile(fileobj=raw, mode="wb", mtime=0, filename="") as g:
                        g.write(f.read())
            with open(idx, "rb") as f:
                digest = hashlib.sha256(f.read()).hexdigest()
        meta = dict(tag=tag, e=e, a=a, k=k, step=step, binary=which, read_length=l, n_reads=len(reads),
                    counters=dict(zip(("num_reads", "num_mapped_reads", "num_candidates_before_filter", "num_candidates",
                                       "num_mappings"), counters)), index_sha256=digest)
        with open(os.path.join(HERE, tag + ".json"), "w") as f:
            json.dump(meta, f, indent=1, sort_keys=True)
            f.write("\n")
        print(tag, counters)


if __name__ == "__main__":
    main()
