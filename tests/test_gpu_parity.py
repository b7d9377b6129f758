"""GPU parity tests proper: the CUDA path, called through the C ABI (fem_b200.api -> libfemgpu.so),
against the CPU oracle (oracle/fem_oracle.c) on the same seeded inputs.  Bit-exact: this is integer
work, there is no tolerance."""
import numpy as np
import pytest

import oracle_lib as ol
import synth

pytestmark = pytest.mark.gpu

# (tag, record lengths, repeat_frac, ref N frac, n_reads, read length, e, a, k, step, read N frac)
CASES = [
    ("cfg1_e3_s5", [300_000, 200_000, 150_000], 0.0, 0.0, 1500, 100, 3, 1, 12, 5, 0.0),
    ("cfg2_e5_s4", [400_000, 350_000], 0.0, 0.0, 1200, 100, 5, 1, 12, 4, 0.0),
    ("cfg3_e7_s5", [500_000, 300_000], 0.0, 0.0, 800, 150, 7, 1, 12, 5, 0.0),
    ("cfg0_e3_s3", [300_000, 100_000], 0.05, 0.0, 1000, 100, 3, 1, 12, 3, 0.0),
    ("rep_e7_s2", [200_000, 150_000], 0.35, 0.0, 500, 150, 7, 1, 12, 2, 0.0),
    ("rep_e5_s2", [200_000, 150_000], 0.35, 0.0, 500, 100, 5, 1, 12, 2, 0.0),
    ("cfg5_k14_s10", [600_000, 500_000], 0.0, 0.0, 500, 250, 7, 1, 14, 10, 0.0),
    ("a0", [200_000, 100_000], 0.2, 0.0, 600, 100, 3, 0, 12, 3, 0.0),
    ("a2", [200_000, 100_000], 0.2, 0.0, 600, 100, 3, 2, 12, 3, 0.0),
    ("e0", [200_000, 100_000], 0.2, 0.0, 600, 100, 0, 1, 12, 3, 0.0),
    ("e1_k10", [100_000, 50_000], 0.1, 0.0, 600, 80, 1, 1, 10, 4, 0.0),
    ("withN", [200_000, 100_000], 0.1, 0.002, 800, 100, 4, 1, 12, 3, 0.004),
    # many short records: reads at record starts and ends (entries dropped for position < seed offset,
    # the 32-bit coordinate's pads, the range filter) on nearly every read
    ("edges_e3_s3", [1500] * 70 + [400, 130, 99, 40], 0.1, 0.0, 900, 100, 3, 1, 12, 3, 0.0),
    ("edges_e5_s2", [2200] * 50, 0.3, 0.0, 600, 120, 5, 1, 11, 2, 0.0),
    # dense indexes (window 8: 65 536 buckets, ~30 occurrences per seed as on the 3.1 Gbp configs): the occurrence runs
    # fill k_cand_reg's register slots, some overflow them, and chance pairs of near entries survive in most groups
    ("dense_q9", [2_500_000, 1_500_000], 0.0, 0.0, 300, 150, 7, 1, 8, 2, 0.0),
    ("dense_q7", [2_200_000, 1_800_000], 0.02, 0.0, 300, 120, 5, 1, 8, 2, 0.0),
    ("dense_q5", [1_200_000, 900_000], 0.05, 0.0, 300, 100, 3, 1, 8, 1, 0.0),
]


class Setup:
    def __init__(self, case):
        (self.tag, lens, rep, refn, n_reads, l, self.e, self.a, self.k, self.step, readn) = case
        seed = abs(hash(self.tag)) % (2 ** 31) if False else sum(map(ord, self.tag))
        self.records = synth.random_reference(seed, lens, rep, refn)
        self.reads = synth.simulate_reads(seed + 1, self.records, n_reads, l, self.e, readn)
        self.oracle = ol.Oracle(records=[(n, s.decode()) for n, s in self.records])
        self.oracle.build_index(self.k, self.step)
        self.oracle.set_params(self.e, self.a)
        from fem_b200.api import Mapper
        self.mapper = Mapper(0)
        k, step, lut, occ = self.oracle.index_arrays()
        self.mapper.upload_index(k, step, lut, occ)
        self.mapper.upload_reference(self.records)

    def close(self):
        self.mapper.close()
        self.oracle.close()


@pytest.fixture(scope="module", params=CASES, ids=[c[0] for c in CASES])
def setup(request):
    s = Setup(request.param)
    yield s
    s.close()


def oracle_candidates(s, seqs):
    first, cands, before, status = [0], [], [], []
    for seq in seqs:
        for text in (seq, ol.revcomp(seq)):
            n, c, b = s.oracle.candidates(text)
            status.append(2 if n < 0 else 0)
            cands.extend(c)
            before.append(b)
            first.append(len(cands))
    return np.array(first, np.uint64), np.array(cands, np.uint64), np.array(before, np.uint32), np.array(status, np.uint8)


def check_candidates(s, seqs):
    gf, gc, gb, gs = s.mapper.generate_group_seeding_candidates(seqs, s.e, s.a)
    of, oc, ob, os_ = oracle_candidates(s, seqs)
    # status 1 (reference returns 0 early) is indistinguishable from "ok, zero candidates" in the oracle
    assert np.array_equal(gs == 2, os_ == 2), "infeasible flags differ"
    assert np.array_equal(gb, ob), "candidates-before-filter counters differ"
    assert np.array_equal(gf, of), "candidate counts differ"
    assert np.array_equal(gc, oc), "candidate values differ"
    return len(oc)


def test_candidates_match_oracle(setup):
    seqs = [r[1] for r in setup.reads]
    n = check_candidates(setup, seqs)
    assert n > 0


def test_candidates_big_path(setup):
    """Tiny shared-memory capacities force read-strands through the global-scratch path."""
    seqs = [r[1] for r in setup.reads[:300]]
    setup.mapper.set_option("sort_capacity", 32)
    setup.mapper.set_option("acc_capacity", 2)
    try:
        for mode in (1, 2, 3, 4, 0):  # sort path for everything / hash path first / automatic (pair path when a >= 1)
            setup.mapper.set_option("cand_mode", mode)
            check_candidates(setup, seqs)
    finally:
        setup.mapper.set_option("cand_mode", 0)
        setup.mapper.set_option("sort_capacity", 1024)
        setup.mapper.set_option("acc_capacity", 256)


@pytest.mark.parametrize("slots", [32, 64, 4096])
def test_candidates_hash_path_table_sizes(setup, slots):
    """k_cand_hash with tables from far too small (most read-strands handed on to the sort path)
    to generous; and the sort path alone must agree with it."""
    seqs = [r[1] for r in setup.reads]
    setup.mapper.set_option("hash_slots", slots)
    setup.mapper.set_option("cand_mode", 2)
    try:
        check_candidates(setup, seqs)
    finally:
        setup.mapper.set_option("hash_slots", 0)
        setup.mapper.set_option("cand_mode", 0)


@pytest.mark.parametrize("bits,values", [(1024, 32), (1024, 2048), (4096, 128), (32768, 512), (0, 0)])
def test_candidates_pair_path_sizes(setup, bits, values):
    """k_cand_pair from a crowded bitmap (nearly every entry falsely flagged, most read-strands handed
    on to the sort path) and a tiny value buffer to generous ones; (0, 0) is the automatic plan."""
    seqs = [r[1] for r in setup.reads]
    setup.mapper.set_option("pair_bits", bits)
    setup.mapper.set_option("pair_values", values)
    setup.mapper.set_option("cand_mode", 6)  # k_cand_pair itself (3 would take its register form where the shape fits)
    try:
        check_candidates(setup, seqs)
    finally:
        setup.mapper.set_option("pair_bits", 0)
        setup.mapper.set_option("pair_values", 0)
        setup.mapper.set_option("cand_mode", 0)


@pytest.mark.parametrize("bits", [1024, 16384, 0])
def test_candidates_reg_path(setup, bits):
    """k_cand_reg (a group's entries in registers between the bitmap passes) forced whatever the index density
    (cand_mode 5; shapes Q = 5, 7, 9, the others take k_cand_pair), then excluded (cand_mode 6: k_cand_pair alone).
    A crowded bitmap floods the flagged list, so most read-strands go down the chain k_cand_reg -> k_cand_pair (list
    mode, length read on the device) -> k_cand."""
    seqs = [r[1] for r in setup.reads]
    setup.mapper.set_option("pair_bits", bits)
    try:
        for mode in (5, 6):
            setup.mapper.set_option("cand_mode", mode)
            check_candidates(setup, seqs)
    finally:
        setup.mapper.set_option("pair_bits", 0)
        setup.mapper.set_option("cand_mode", 0)


@pytest.mark.parametrize("shape", [0, 1, 4])
def test_candidates_small_path(setup, shape):
    """k_cand_small (all groups of a read-strand at once) forced on every case: dense cases hand most
    read-strands on to the sort path, sparse ones stay.  Shape 1: one seed per lane (at most 32 seeds,
    128 entries; wider cases take shape 4), 4: up to 128 seeds and 256 entries, 0: the automatic choice."""
    seqs = [r[1] for r in setup.reads]
    setup.mapper.set_option("cand_mode", 4)
    setup.mapper.set_option("small_shape", shape)
    try:
        check_candidates(setup, seqs)
    finally:
        setup.mapper.set_option("small_shape", 0)
        setup.mapper.set_option("cand_mode", 0)


def test_candidates_sort_path_only(setup):
    seqs = [r[1] for r in setup.reads]
    setup.mapper.set_option("cand_mode", 1)
    try:
        check_candidates(setup, seqs)
    finally:
        setup.mapper.set_option("cand_mode", 0)


def test_small_arena_retries(setup):
    seqs = [r[1] for r in setup.reads[:300]]
    setup.mapper.set_option("arena_candidates", 8)
    try:
        check_candidates(setup, seqs)
    finally:
        setup.mapper.set_option("arena_candidates", 0)


def test_mappings_match_oracle(setup):
    from fem_b200.api import sort_mappings
    seqs = [r[1] for r in setup.reads]
    first, maps, stats = setup.mapper.map_batch(seqs, setup.e, setup.a)
    ost = ol.Stats()
    n_checked = 0
    for i, seq in enumerate(seqs):
        mine = maps[int(first[i]):int(first[i + 1])].copy()
        # emission order (before the sort): forward hits in candidate order, then reverse hits
        if i < 200:
            exp = []
            for d, text in enumerate((seq, ol.revcomp(seq))):
                n, c, _ = setup.oracle.candidates(text)
                if n > 0:
                    exp.extend(setup.oracle.verify(text, d, c))
            got = [(int(m["candidate_position"]), int(m["end_position_offset"]), int(m["direction"]), int(m["edit_distance"])) for m in mine]
            assert got == exp, "read %d: emission order differs" % i
        exp_sorted = setup.oracle.map_read(seq, ost)
        sort_mappings(mine)
        got_sorted = [(int(m["candidate_position"]), int(m["end_position_offset"]), int(m["direction"]), int(m["edit_distance"])) for m in mine]
        assert got_sorted == exp_sorted, "read %d: sorted mappings differ" % i
        n_checked += len(exp_sorted)
    assert stats == ost.as_tuple()
    assert n_checked > 0


@pytest.mark.parametrize("mode", [0, 5, 6, 4, 2])
def test_repeated_runs_are_bit_identical(setup, mode):
    """compute-sanitizer is closed on this pool (profiles/r2v_sanitizer_closed_on_this_pool.txt), so a hazard between
    lanes that hand data through shared memory (bitmaps, flagged lists, survivor lists, SAM images) has to show up as
    what it causes: results that change from run to run.  Every candidate path maps the same batch six times; all
    outputs -- candidates, mappings in emission order, alignments, statistics -- must be bit-identical (and the first
    is checked against the oracle by the tests above)."""
    seqs = [r[1] for r in setup.reads]
    setup.mapper.set_option("cand_mode", mode)
    try:
        ref_c = setup.mapper.generate_group_seeding_candidates(seqs, setup.e, setup.a)
        ref_m = setup.mapper.map_batch(seqs, setup.e, setup.a)
        ref_a = setup.mapper.last_alignments.copy() if setup.mapper.last_alignments is not None else None
        for _ in range(5):
            c = setup.mapper.generate_group_seeding_candidates(seqs, setup.e, setup.a)
            assert all(np.array_equal(x, y) for x, y in zip(ref_c, c))
            m = setup.mapper.map_batch(seqs, setup.e, setup.a)
            assert np.array_equal(ref_m[0], m[0]) and np.array_equal(ref_m[1], m[1]) and ref_m[2] == m[2]
            if ref_a is not None:
                assert np.array_equal(ref_a, setup.mapper.last_alignments)
    finally:
        setup.mapper.set_option("cand_mode", 0)


def test_verify_forms_agree(setup):
    """k_verify's two forms of the Myers columns (src/align.c:102-277): bit planes (the default for candidates
    without ambiguous bases) and the general five-row form; same mappings, same statistics."""
    seqs = [r[1] for r in setup.reads]
    want = setup.mapper.map_batch(seqs, setup.e, setup.a)
    setup.mapper.set_option("verify_planes", 0)
    try:
        got = setup.mapper.map_batch(seqs, setup.e, setup.a)
    finally:
        setup.mapper.set_option("verify_planes", 1)
    assert np.array_equal(want[0], got[0]) and np.array_equal(want[1], got[1]) and want[2] == got[2]


def test_device_traceback_matches_oracle(setup):
    """k_trace: start, CIGAR and MD text of every mapping (src/align.c:279-544) against the oracle."""
    from fem_b200.api import ALN_HOST
    seqs = [r[1] for r in setup.reads[:400]]
    first, maps, _ = setup.mapper.map_batch(seqs, setup.e, setup.a)
    alns = setup.mapper.last_alignments
    assert alns is not None and len(alns) == len(maps)
    n_dev = n_md = 0
    for i, seq in enumerate(seqs):
        canonical = set(seq) <= set(b"ACGTN")
        for j in range(int(first[i]), int(first[i + 1])):
            m, al = maps[j], alns[j]
            text = ol.revcomp(seq) if m["direction"] else seq
            start, cigar, md = setup.oracle.alignment(int(m["candidate_position"]), text, int(m["edit_distance"]),
                                                      int(m["end_position_offset"]))
            if al["n_cigar"] == ALN_HOST:
                assert not canonical or setup.tag == "withN", "read %d left to the host without need" % i
                continue
            n_dev += 1
            got = [((int(c) >> 2) << 4) | (int(c) & 3) for c in al["cigar"][:al["n_cigar"]]]
            assert (int(al["start"]), got) == (start, cigar), "read %d mapping %d" % (i, j)
            if al["md_len"] != ALN_HOST:  # the MD text of src/align.c:501-544 as well
                n_md += 1
                assert al["md"][:al["md_len"]].decode() == md and len(md) == al["md_len"], "read %d mapping %d" % (i, j)
            else:
                assert len(md) > 47
    assert n_dev > 0 and n_md > 0
    setup.mapper.set_option("device_traceback", 0)
    try:
        setup.mapper.map_batch(seqs[:50], setup.e, setup.a)
        assert setup.mapper.last_alignments is None
    finally:
        setup.mapper.set_option("device_traceback", 1)


def test_device_sam_text_matches_oracle(setup, tmp_path):
    """k_sam: the records of every read the device formats (order, flags, SEQ/QUAL, NM, MD) are the
    oracle's SAM lines for that read; reads it leaves to the host own no bytes and are few."""
    reads = [(n, q, bytes(33 + (i * 7 + j * 3) % 61 for j in range(len(q)))) for i, (n, q, _) in enumerate(setup.reads[:700])]
    fq, want_sam = str(tmp_path / "r.fq"), str(tmp_path / "want.sam")
    synth.write_fastq(fq, reads)
    setup.oracle.map_file(fq, want_sam)
    want = {}
    with open(want_sam, "rb") as f:
        for line in f:
            if not line.startswith(b"@"):
                want.setdefault(line.split(b"\t", 1)[0], []).append(line)
    seqs, names = [r[1] for r in reads], [r[0].encode() for r in reads]
    quals = [r[2] if i % 7 else None for i, r in enumerate(reads)]  # some FASTA-like reads: QUAL '*'
    first, maps, _ = setup.mapper.map_batch_sam(seqs, names, quals, setup.e, setup.a)
    assert setup.mapper.last_sam is not None
    off, host, text = setup.mapper.last_sam
    assert len(text) == int(off[-1])
    n_dev = n_host = n_big = n_big_dev = 0
    for i, name in enumerate(names):
        got = text[int(off[i]):int(off[i + 1])]
        n = int(first[i + 1] - first[i])
        n_big += n > 64
        n_big_dev += n > 64 and not host[i]
        if host[i]:
            assert got == b"" and (n > 64 or not set(seqs[i]) <= set(b"ACGTN") or setup.tag == "withN"), "read %d" % i
            n_host += 1
            continue
        exp = b"".join(want.get(name, []))
        if quals[i] is None and n:  # the oracle's file carried qualities for every read
            lines = exp.split(b"\n")
            f0 = lines[0].split(b"\t")
            f0[10] = b"*"
            exp = b"\n".join([b"\t".join(f0)] + lines[1:])
        assert got == exp, "read %d: %r != %r" % (i, got[:300], exp[:300])
        n_dev += n > 0
    assert n_dev > 0 and n_dev >= n_host or setup.tag.startswith("rep")
    # reads with 65..2048 mappings go through k_sam_big_* unless two of their records share a sort key
    assert n_big == 0 or n_big_dev > 0, (n_big, n_big_dev)
    print("%s: %d reads formatted on the device (%d of %d with more than 64 records), %d on the host" % (setup.tag, n_dev, n_big_dev, n_big, n_host))
    # the inputs apply to one submit only
    setup.mapper.map_batch(seqs[:20], setup.e, setup.a)
    assert setup.mapper.last_sam is None


def test_clone_resident_maps_the_same(setup):
    """femgpu_clone_resident: a second context filled device to device (here on the same GPU; across GPUs in
    tools/gpu_n2.sh) gives the same mappings and the same device SAM text as the one that was uploaded to."""
    from fem_b200.api import Mapper
    seqs = [r[1] for r in setup.reads[:300]]
    names = [r[0].encode() for r in setup.reads[:300]]
    quals = [r[2] for r in setup.reads[:300]]
    f0, m0, st0 = setup.mapper.map_batch_sam(seqs, names, quals, setup.e, setup.a)
    sam0 = setup.mapper.last_sam
    other = Mapper(0)
    try:
        other.clone_resident_from(setup.mapper)
        f1, m1, st1 = other.map_batch_sam(seqs, names, quals, setup.e, setup.a)
        assert np.array_equal(f0, f1) and np.array_equal(m0, m1) and st0 == st1
        assert sam0 is not None and other.last_sam is not None
        assert np.array_equal(sam0[0], other.last_sam[0]) and sam0[2] == other.last_sam[2]
        k, step, lut, occ = other.download_index()
        k0, step0, lut0, occ0 = setup.mapper.download_index()
        assert (k, step) == (k0, step0) and np.array_equal(lut, lut0) and np.array_equal(occ, occ0)
    finally:
        other.close()


def test_double_buffered_submit_matches_sync(setup):
    seqs = [r[1] for r in setup.reads]
    half = len(seqs) // 2
    f_all, m_all, st_all = setup.mapper.map_batch(seqs, setup.e, setup.a)
    setup.mapper.submit(0, seqs[:half], setup.e, setup.a)
    setup.mapper.submit(1, seqs[half:], setup.e, setup.a)
    f0, m0, st0 = setup.mapper.wait(0)
    f1, m1, st1 = setup.mapper.wait(1)
    assert np.array_equal(np.concatenate([m0, m1]), m_all)
    assert np.array_equal(np.concatenate([f0, f1[1:] + f0[-1]]), f_all)
    assert tuple(a + b for a, b in zip(st0, st1)) == st_all


@pytest.mark.parametrize("mode", [1, 2, 3])
def test_seed_selection_forms_agree(setup, mode):
    """k_seed's DP forms (one lane per group / warp scan / several read-strands per warp) must select the same seeds."""
    seqs = [r[1] for r in setup.reads[:400]]
    setup.mapper.set_option("seed_dp_mode", mode)
    try:
        check_candidates(setup, seqs)
    finally:
        setup.mapper.set_option("seed_dp_mode", 0)
