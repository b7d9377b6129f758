"""Seeded synthetic references and reads for the tests (numpy, no files needed)."""
import numpy as np

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
COMP = bytes.maketrans(b"ACGTacgtN", b"TGCAtgcaN")


def random_reference(seed, lengths, repeat_frac=0.0, n_frac=0.0):
    """list of (name, bytes).  repeat_frac: share of each record overwritten by copies of a few
    short families / tandem arrays; n_frac: share of bases replaced by 'N'."""
    rng = np.random.default_rng(seed)
    fams = [ACGT[rng.integers(0, 4, size=int(rng.integers(80, 400)))] for _ in range(6)]
    out = []
    for i, ln in enumerate(lengths):
        s = ACGT[rng.integers(0, 4, size=ln)].copy()
        budget = int(ln * repeat_frac)
        while budget > 0:
            if rng.random() < 0.5:
                f = fams[int(rng.integers(0, len(fams)))]
                if len(f) >= ln:
                    break
                p = int(rng.integers(0, ln - len(f)))
                c = f.copy()
                mut = rng.random(len(c)) < 0.04
                c[mut] = ACGT[rng.integers(0, 4, size=int(mut.sum()))]
                s[p:p + len(c)] = c
                budget -= len(c)
            else:
                unit = int(rng.integers(2, 40))
                copies = int(rng.integers(5, 60))
                if unit * copies >= ln:
                    break
                p = int(rng.integers(0, ln - unit * copies))
                s[p:p + unit * copies] = np.tile(s[p:p + unit], copies)
                budget -= unit * copies
        if n_frac > 0:
            s[rng.random(ln) < n_frac] = ord("N")
        out.append(("chr%d" % (i + 1), s.tobytes()))
    return out


def simulate_reads(seed, records, n_reads, length, max_edits, n_frac=0.0, lengths=None):
    """list of (name, seq bytes, qual bytes); substitutions 80% / insertions 10% / deletions 10%."""
    rng = np.random.default_rng(seed)
    weights = np.array([len(r[1]) for r in records], dtype=np.float64)
    weights /= weights.sum()
    out = []
    for i in range(n_reads):
        l = length if lengths is None else int(lengths[i % len(lengths)])
        while True:
            ri = int(rng.choice(len(records), p=weights))
            ref = records[ri][1]
            if len(ref) > l + 4 * max_edits + 8:
                break
        p = int(rng.integers(0, len(ref) - l - max_edits - 1))
        w = bytearray(ref[p:p + l + max_edits + 1])
        m = int(rng.integers(0, max_edits + 1))
        for _ in range(m):
            kind = int(rng.integers(0, 10))
            at = int(rng.integers(0, l))
            if kind < 8:
                w[at] = b"ACGT"[(b"ACGT".find(bytes([w[at]])) + 1 + int(rng.integers(0, 3))) % 4]
            elif kind == 8:
                w.insert(at, b"ACGT"[int(rng.integers(0, 4))])
            else:
                del w[at]
        seq = bytes(w[:l])
        if n_frac > 0:
            arr = np.frombuffer(seq, dtype=np.uint8).copy()
            arr[rng.random(l) < n_frac] = ord("N")
            seq = arr.tobytes()
        strand = int(rng.integers(0, 2))
        if strand:
            seq = seq.translate(COMP)[::-1]
        out.append(("r%d_%d_%d_%s_%d" % (i, ri, p, "-+"[1 - strand], m), seq, b"I" * l))
    return out


def write_fasta(path, records, width=60):
    with open(path, "wb") as f:
        for name, seq in records:
            f.write(b">" + name.encode() + b"\n")
            for i in range(0, len(seq), width):
                f.write(seq[i:i + width] + b"\n")


def write_fastq(path, reads):
    with open(path, "wb") as f:
        for name, seq, qual in reads:
            f.write(b"@" + name.encode() + b"\n" + seq + b"\n+\n" + qual + b"\n")
