"""femgpu_lookup_table_noalloc (include/femgpu.h): the entry point that allocates nothing -- its scratch is a static
device array inside libfemgpu.so.  It has no counterpart in the reference; it is checked against the existing entry
point that does the same arithmetic with cudaMalloc'ed scratch: the lookup table femgpu_build_index leaves on the
device (itself equal to the reference's, tests/test_index_gpu.py).  Integer work: the tables must be identical.

Collected last on purpose.  Passes alone in a fresh process and after the others:
nothing here depends on what earlier calls left in the static array, and two calls in a row of different sizes
show that the second does not see the first one's tile sums."""
import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu


def _existing_lookup_table(records, k, step):
    from fem_b200.api import Mapper
    m = Mapper(0)
    try:
        m.upload_reference(records)
        m.build_index(k, step)
        gk, gstep, lut, _ = m.download_index()
        assert (gk, gstep) == (k, step)
        return lut
    finally:
        m.close()


def _noalloc_lookup_table(counts):
    """counts (uint32, host) -> the table of the new entry point, device buffers from torch (the caller allocates)"""
    import torch
    from fem_b200 import api
    n = counts.shape[0]
    d_counts = torch.from_numpy(counts.view(np.int32).copy()).to("cuda:0")
    d_lut = torch.full((n + 1,), -1, dtype=torch.int32, device="cuda:0")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(0)
    api.lookup_table_noalloc(0, d_counts.data_ptr(), n, d_lut.data_ptr(), torch.cuda.current_stream(0).cuda_stream)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(0) == before
    return d_lut.cpu().numpy().view(np.uint32)


def test_two_calls_in_a_row_equal_the_existing_entry_point():
    # 4^12 buckets = the maximum (8192 tiles), then 4^9 + a ragged size: the second call uses 128 tiles of an array whose
    # 8192 entries hold the first call's sums
    big = _existing_lookup_table(synth.random_reference(31, [400_000, 300_000], 0.2), 12, 1)
    small = _existing_lookup_table(synth.random_reference(32, [150_000, 90_000], 0.0), 9, 2)
    assert big.shape[0] == (1 << 24) + 1 and small.shape[0] == (1 << 18) + 1
    for lut in (big, small, big, small):
        counts = np.diff(lut).astype(np.uint32)
        got = _noalloc_lookup_table(counts)
        assert got.shape == lut.shape and np.array_equal(got, lut)
    # a size that is no multiple of the tile, cut out of the same counts (prefix sums of a prefix)
    counts = np.diff(small).astype(np.uint32)[:100_003]
    got = _noalloc_lookup_table(counts)
    want = np.concatenate([[0], np.cumsum(counts, dtype=np.uint64)]).astype(np.uint32)
    assert np.array_equal(got, want) and np.array_equal(got, small[:100_004])
    one = _noalloc_lookup_table(np.array([7], dtype=np.uint32))
    assert one.tolist() == [0, 7]


def test_sizes_beyond_the_static_scratch_are_refused():
    import torch
    from fem_b200 import api
    d = torch.zeros(16, dtype=torch.int32, device="cuda:0")
    for n in (0, api.NOALLOC_MAX_BUCKETS + 1, 1 << 30):
        with pytest.raises(api.FemGpuError, match="buckets"):
            api.lookup_table_noalloc(0, d.data_ptr(), n, d.data_ptr())
    with pytest.raises(api.FemGpuError, match="null"):
        api.lookup_table_noalloc(0, 0, 16, d.data_ptr())
    with pytest.raises(api.FemGpuError, match="device"):
        api.lookup_table_noalloc(-1, d.data_ptr(), 15, d.data_ptr())
    torch.cuda.synchronize()
