"""GPU parity at the edges of the input space, through the C ABI against the CPU oracle: empty and
one-read batches, ragged read lengths from k up to the supported maximum (reads too short to seed are
the reference's crash cases: they must come back unmapped and counted), reads of only N, duplicated
reads, homopolymer / tandem-repeat reads with thousands of locations (global-scratch candidate path,
SAM records beyond the device formatter's limits), and a read longer than the maximum (left unmapped and counted,
not an error).  Bit-exact: integer work, no tolerance."""
import numpy as np
import pytest

import oracle_lib as ol
import synth

pytestmark = pytest.mark.gpu

K, STEP, E, A = 12, 3, 3, 1
LENGTHS = [12, 13, 20, 35, 47, 48, 64, 100, 101, 150, 200, 251, 400, 777, 1024]


class World:
    def __init__(self):
        rng = np.random.default_rng(4242)
        records = synth.random_reference(99, [150_000, 80_000, 2_000], 0.15)
        # a record of low complexity: homopolymer run, tandem repeat, then random sequence
        tail = bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), 20_000))
        records.append(("lowcomplexity", b"A" * 3000 + b"ACGT" * 900 + b"AC" * 1500 + tail))
        self.records = records
        self.oracle = ol.Oracle(records=[(n, s.decode()) for n, s in records])
        self.oracle.build_index(K, STEP)
        self.oracle.set_params(E, A)
        from fem_b200.api import Mapper
        self.mapper = Mapper(0)
        k, step, lut, occ = self.oracle.index_arrays()
        self.mapper.upload_index(k, step, lut, occ)
        self.mapper.upload_reference(records)

    def close(self):
        self.mapper.close()
        self.oracle.close()


@pytest.fixture(scope="module")
def world():
    w = World()
    yield w
    w.close()


def as_tuples(maps):
    return [(int(m["candidate_position"]), int(m["end_position_offset"]), int(m["direction"]), int(m["edit_distance"])) for m in maps]


def check_batch(w, reads, tmp_path, want_big=False):
    """mappings (reference order), statistics and the device's SAM text of `reads` against the oracle"""
    from fem_b200.api import sort_mappings
    seqs, names, quals = [r[1] for r in reads], [r[0].encode() for r in reads], [r[2] for r in reads]
    first, maps, stats = w.mapper.map_batch_sam(seqs, names, quals, E, A)
    ost = ol.Stats()
    for i, seq in enumerate(seqs):
        mine = maps[int(first[i]):int(first[i + 1])].copy()
        sort_mappings(mine)
        assert as_tuples(mine) == w.oracle.map_read(seq, ost), "read %d (%d bases)" % (i, len(seq))
    assert stats == ost.as_tuple()
    fq, want_sam = str(tmp_path / "r.fq"), str(tmp_path / "want.sam")
    synth.write_fastq(fq, reads)
    w.oracle.map_file(fq, want_sam)
    want = {}
    with open(want_sam, "rb") as f:
        for line in f:
            if not line.startswith(b"@"):
                want.setdefault(line.split(b"\t", 1)[0], []).append(line)
    if w.mapper.last_sam is None:
        assert len(seqs) == 0
        return stats, 0, 0
    off, host, text = w.mapper.last_sam
    n_dev = n_big_dev = 0
    for i, name in enumerate(names):
        got = text[int(off[i]):int(off[i + 1])]
        n = int(first[i + 1] - first[i])
        if host[i]:
            assert got == b"" and n > 64, "read %d left to the host with %d records" % (i, n)
            continue
        assert got == b"".join(want.get(name, [])), "read %d (%d bases, %d records)" % (i, len(seqs[i]), n)
        n_dev += n > 0
        n_big_dev += n > 64
    if want_big:
        assert n_big_dev > 0
    return stats, n_dev, n_big_dev


def test_empty_batch(world):
    first, maps, stats = world.mapper.map_batch([], E, A)
    assert list(first) == [0] and len(maps) == 0 and stats[:5] == (0, 0, 0, 0, 0)
    gf, gc, gb, gs = world.mapper.generate_group_seeding_candidates([], E, A)
    assert list(gf) == [0] and len(gc) == 0 and len(gb) == 0 and len(gs) == 0
    first, maps, stats = world.mapper.map_batch_sam([], [], [], E, A)
    assert list(first) == [0] and len(maps) == 0
    assert world.mapper.last_sam is None or world.mapper.last_sam[2] == b""
    # and the context still works
    reads = synth.simulate_reads(1, world.records[:2], 20, 100, E)
    assert world.mapper.map_batch([r[1] for r in reads], E, A)[2][1] > 0


def test_ragged_lengths_from_k_to_the_maximum(world, tmp_path):
    reads = synth.simulate_reads(5, world.records[:3], 450, 100, E, lengths=LENGTHS)
    assert {len(r[1]) for r in reads} == set(LENGTHS)
    stats, n_dev, _ = check_batch(world, reads, tmp_path)
    assert stats[1] > 200 and n_dev > 200
    assert stats[5] > 0  # the reads too short for e + 2 seeds per group are counted, not crashed on
    # candidates of both strands as well (status, before-filter counters, values)
    seqs = [r[1] for r in reads]
    gf, gc, gb, gs = world.mapper.generate_group_seeding_candidates(seqs, E, A)
    o_first, o_c, o_b, o_s = [0], [], [], []
    for seq in seqs:
        for text in (seq, ol.revcomp(seq)):
            n, c, b = world.oracle.candidates(text)
            o_s.append(2 if n < 0 else 0)
            o_c.extend(c)
            o_b.append(b)
            o_first.append(len(o_c))
    assert np.array_equal(gs == 2, np.array(o_s) == 2)
    assert np.array_equal(gb, np.array(o_b, np.uint32)) and np.array_equal(gf, np.array(o_first, np.uint64))
    assert np.array_equal(gc, np.array(o_c, np.uint64))


def test_one_read_and_duplicates(world, tmp_path):
    reads = synth.simulate_reads(6, world.records[:2], 3, 100, E)
    check_batch(world, reads[:1], tmp_path)
    dup = [("d%d" % i, reads[1][1], reads[1][2]) for i in range(333)]
    stats, n_dev, _ = check_batch(world, dup, tmp_path)
    assert stats[1] in (0, 333) and stats[4] % 333 == 0


def test_reads_of_n_and_zero_length(world, tmp_path):
    reads = [("allN", b"N" * 100, b"I" * 100), ("mostlyN", b"ACGTACGTACGTACGT" + b"N" * 84, b"I" * 100),
             ("fourN", b"N" * 4 + world.records[0][1][500:596], b"I" * 100), ("ok", world.records[0][1][1000:1100], b"I" * 100)]
    stats, _, _ = check_batch(world, reads, tmp_path)
    assert stats[0] == 4 and stats[1] >= 1
    # zero-length read inside a batch (the kseq reader never produces one, the C ABI accepts it): unmapped
    first, maps, st = world.mapper.map_batch([reads[3][1], b"", reads[3][1]], E, A)
    assert int(first[2] - first[1]) == 0 and int(first[1] - first[0]) == int(first[3] - first[2]) > 0


def test_low_complexity_reads(world, tmp_path):
    """thousands of locations per read: global-scratch candidate path, k_sam_big_* (65..2048 records) and the
    host's share beyond that"""
    low = world.records[3][1]
    reads = [("polyA", b"A" * 100, b"I" * 100), ("acgt", b"ACGT" * 25, b"I" * 100), ("ac", b"AC" * 50, b"I" * 100),
             ("edge", low[2950:3050], b"I" * 100), ("edge2", low[6560:6660], b"I" * 100)]
    reads += [("t%d" % i, low[3000 + 4 * i:3100 + 4 * i], b"5" * 100) for i in range(40)]
    reads += synth.simulate_reads(8, world.records[:2], 60, 100, E)
    stats, n_dev, n_big = check_batch(world, reads, tmp_path, want_big=True)
    assert stats[4] > 5000


def test_read_longer_than_the_maximum_is_left_unmapped(world):
    """one read beyond the seed kernel's shared-memory plan (1024 bases) used to abort the whole batch (and
    with it the `FEM map` run); it is now treated like the reads the reference cannot process -- left unmapped, counted
    in num_infeasible_reads -- while the other reads of the batch map as if it were not there."""
    from fem_b200.api import FemGpuError
    long_read = bytes(np.random.default_rng(3).choice(np.frombuffer(b"ACGT", dtype=np.uint8), 1025))
    reads = synth.simulate_reads(9, world.records[:2], 10, 100, E)
    seqs = [r[1] for r in reads]
    want_first, want_maps, want_stats = world.mapper.map_batch(seqs, E, A)
    first, maps, stats = world.mapper.map_batch(seqs[:4] + [long_read] + seqs[4:] + [long_read], E, A)
    assert stats[0] == 12 and stats[5] == want_stats[5] + 2 and stats[1:5] == want_stats[1:5]
    assert first[5] == first[4] and first[12] == first[11]  # no mappings for the long reads
    assert np.array_equal(maps, want_maps)
    assert world.mapper.map_batch(seqs, E, A)[2] == want_stats  # and the plan goes back to normal
    with pytest.raises(FemGpuError):
        world.mapper.map_batch([reads[0][1]], 8, A)  # src/FEM_map.c:30 "Wrong error threshold."
    with pytest.raises(FemGpuError):
        world.mapper.map_batch([reads[0][1]], E, 3)  # :38 "Wrong number of additional q-grams."
