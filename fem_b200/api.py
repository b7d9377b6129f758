"""ctypes binding of fem_b200/lib/libfemgpu.so (the C ABI declared in include/femgpu.h).

The names mirror the reference's call sites for this path:
  Mapper.generate_group_seeding_candidates  <->  src/filter.h:9   (both strands of a batch)
  Mapper.map_batch                          <->  src/map.c:27-49  (candidates + verify_candidates, src/align.h:13)
  sort_mappings                             <->  src/align.c:57   (radix_sort_mapping)

There is no CPU fallback: if the shared library or a CUDA device is missing, construction raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libfemgpu.so")


class FemGpuError(RuntimeError):
    pass


class Params(C.Structure):
    _fields_ = [("k", C.c_int32), ("step", C.c_int32), ("e", C.c_int32), ("a", C.c_int32)]


class MappingRec(C.Structure):
    _fields_ = [("candidate_position", C.c_uint64), ("end_position_offset", C.c_int32),
                ("direction", C.c_uint8), ("edit_distance", C.c_uint8), ("reserved", C.c_uint8 * 2)]


MAPPING_DTYPE = np.dtype([("candidate_position", "<u8"), ("end_position_offset", "<i4"),
                          ("direction", "u1"), ("edit_distance", "u1"), ("reserved", "u1", (2,))])
assert MAPPING_DTYPE.itemsize == 16 == C.sizeof(MappingRec)


class Batch(C.Structure):
    _fields_ = [("bases", C.c_void_p), ("offsets", C.c_void_p), ("n_reads", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("num_reads", "num_mapped_reads", "num_candidates_before_filter",
                                          "num_candidates", "num_mappings", "num_infeasible_reads")]

    def as_tuple(self):
        return tuple(int(getattr(self, n)) for n, _ in self._fields_)


class AlignmentRec(C.Structure):
    _fields_ = [("start", C.c_int16), ("n_cigar", C.c_uint8), ("md_len", C.c_uint8), ("cigar", C.c_uint16 * 18), ("md", C.c_char * 48)]


ALIGNMENT_DTYPE = np.dtype([("start", "<i2"), ("n_cigar", "u1"), ("md_len", "u1"), ("cigar", "<u2", (18,)), ("md", "S48")])
assert ALIGNMENT_DTYPE.itemsize == 88 == C.sizeof(AlignmentRec)
ALN_HOST = 0xFF


class Result(C.Structure):
    _fields_ = [("first", C.POINTER(C.c_uint64)), ("maps", C.POINTER(MappingRec)), ("stats", Stats),
                ("alns", C.POINTER(AlignmentRec)),
                ("sam_text", C.c_void_p), ("sam_off", C.POINTER(C.c_uint64)), ("sam_host", C.POINTER(C.c_uint8))]


class SamInputs(C.Structure):
    _fields_ = [("names", C.c_void_p), ("name_off", C.c_void_p), ("quals", C.c_void_p), ("has_qual", C.c_void_p),
                ("text_out", C.c_void_p), ("text_cap", C.c_uint64), ("off_out", C.c_void_p), ("host_out", C.c_void_p)]


class Candidates(C.Structure):
    _fields_ = [("first", C.POINTER(C.c_uint64)), ("cands", C.POINTER(C.c_uint64)),
                ("before", C.POINTER(C.c_uint32)), ("status", C.POINTER(C.c_uint8))]


EXPORTS = {
    "femgpu_abi_version": (C.c_int, []),
    "femgpu_device_count": (C.c_int, []),
    "femgpu_create": (C.c_void_p, [C.c_int]),
    "femgpu_destroy": (None, [C.c_void_p]),
    "femgpu_last_error": (C.c_char_p, []),
    "femgpu_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "femgpu_upload_index": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "femgpu_upload_index_file": (C.c_int, [C.c_void_p, C.c_char_p]),
    "femgpu_upload_reference": (C.c_int, [C.c_void_p, C.c_uint32, C.POINTER(C.c_char_p), C.POINTER(C.c_uint32)]),
    "femgpu_build_index": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "femgpu_index_sizes": (C.c_int, [C.c_void_p, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "femgpu_download_index": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "femgpu_save_index_file": (C.c_int, [C.c_void_p, C.c_char_p]),
    "femgpu_index_k": (C.c_int, [C.c_void_p]),
    "femgpu_index_step": (C.c_int, [C.c_void_p]),
    "femgpu_map_batch": (C.c_int, [C.c_void_p, C.POINTER(Params), C.POINTER(Batch), C.POINTER(Result)]),
    "femgpu_submit_batch": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(Params), C.POINTER(Batch)]),
    "femgpu_wait_batch": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(Result)]),
    "femgpu_set_reference_names": (C.c_int, [C.c_void_p, C.c_uint32, C.POINTER(C.c_char_p)]),
    "femgpu_clone_resident": (C.c_int, [C.c_void_p, C.c_void_p]),
    "femgpu_set_sam_inputs": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(SamInputs)]),
    "femgpu_candidates_batch": (C.c_int, [C.c_void_p, C.POINTER(Params), C.POINTER(Batch), C.POINTER(Candidates)]),
    "femgpu_sort_mappings": (None, [C.c_void_p, C.c_size_t]),
    "femgpu_sort_mappings_aln": (None, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "femgpu_last_timings": (C.c_int, [C.c_void_p, C.POINTER(C.c_float * 6), C.POINTER(C.c_uint32)]),
    "femgpu_alloc_pinned": (C.c_void_p, [C.c_size_t]),
    "femgpu_free_pinned": (None, [C.c_void_p]),
    "femgpu_measure_int32_peak": (C.c_int, [C.c_void_p, C.POINTER(C.c_double * 3)]),
    "femgpu_last_work": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64 * 4)]),
    "femgpu_rerun_resident": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float * 6)]),
    "femgpu_lookup_table_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
}

_lib = None


def load_library():
    """dlopen the in-tree shared library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FemGpuError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(make -C fem_b200/csrc); there is no CPU fallback" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in EXPORTS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def _check(rc):
    if rc != 0:
        raise FemGpuError("femgpu error %d: %s" % (rc, load_library().femgpu_last_error().decode(errors="replace")))


NOALLOC_MAX_BUCKETS = 1 << 24  # FEMGPU_NOALLOC_MAX_BUCKETS


def lookup_table_noalloc(device, counts_ptr, n_buckets, lut_ptr, stream=0):
    """femgpu_lookup_table_noalloc: lut[0] = 0, lut[h + 1] = counts[0] + ... + counts[h], from DEVICE pointers,
    queued on `stream` (a cudaStream_t as an integer), with its scratch in a static device array of the library
    -- no allocation of any kind.  No counterpart in the reference (see include/femgpu.h)."""
    _check(load_library().femgpu_lookup_table_noalloc(int(device), C.c_void_p(counts_ptr), int(n_buckets),
                                                      C.c_void_p(lut_ptr), C.c_void_p(stream)))


def sort_mappings_aln(maps: np.ndarray, alns: np.ndarray):
    """In-place reference order of one read's mappings with their device alignments carried along."""
    assert maps.dtype == MAPPING_DTYPE and alns.dtype == ALIGNMENT_DTYPE and maps.shape == alns.shape
    assert maps.flags.c_contiguous and alns.flags.c_contiguous
    load_library().femgpu_sort_mappings_aln(maps.ctypes.data, alns.ctypes.data, maps.shape[0])
    return maps, alns


def sort_mappings(maps: np.ndarray):
    """In-place reference order of one read's mappings (src/align.c:57)."""
    assert maps.dtype == MAPPING_DTYPE and maps.flags.c_contiguous
    load_library().femgpu_sort_mappings(maps.ctypes.data, maps.shape[0])
    return maps


class ReadBatch:
    """Flattened SequenceBatch (src/sequence_batch.h:15-24): bases back to back + offsets."""

    def __init__(self, reads):
        if isinstance(reads, ReadBatch):
            self.bases, self.offsets = reads.bases, reads.offsets
        else:
            lens = np.fromiter((len(r) for r in reads), dtype=np.uint64, count=len(reads))
            self.offsets = np.zeros(len(reads) + 1, dtype=np.uint64)
            np.cumsum(lens, out=self.offsets[1:])
            self.bases = np.frombuffer(b"".join(reads), dtype=np.uint8) if len(reads) else np.zeros(0, np.uint8)
        self.n_reads = self.offsets.shape[0] - 1

    @classmethod
    def from_arrays(cls, bases: np.ndarray, offsets: np.ndarray):
        b = cls.__new__(cls)
        b.bases = np.ascontiguousarray(bases, dtype=np.uint8)
        b.offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        b.n_reads = b.offsets.shape[0] - 1
        return b

    def c_struct(self):
        return Batch(self.bases.ctypes.data if self.bases.size else None, self.offsets.ctypes.data, self.n_reads)


class Mapper:
    """One GPU context: resident index + reference, batch mapping."""

    def __init__(self, device=0):
        self.lib = load_library()
        self.h = self.lib.femgpu_create(device)
        if not self.h:
            raise FemGpuError("femgpu_create(%d): %s" % (device, self.lib.femgpu_last_error().decode()))

    def close(self):
        if getattr(self, "h", None):
            self.lib.femgpu_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, key, value):
        _check(self.lib.femgpu_set_option(self.h, key.encode(), int(value)))

    # -- resident data ---------------------------------------------------------------------
    def load_index(self, path):
        """src/index.h:19 load_index"""
        _check(self.lib.femgpu_upload_index_file(self.h, os.fsencode(path)))

    def upload_index(self, k, step, lut: np.ndarray, occ: np.ndarray):
        lut = np.ascontiguousarray(lut, dtype=np.uint32)
        occ = np.ascontiguousarray(occ, dtype=np.uint64)
        _check(self.lib.femgpu_upload_index(self.h, k, step, lut.ctypes.data, lut.size,
                                            occ.ctypes.data if occ.size else None, occ.size))

    def build_index(self, k, step):
        """src/index.c:57-98 construct_index, on the device from the uploaded reference"""
        _check(self.lib.femgpu_build_index(self.h, k, step))

    def download_index(self):
        ln, on = C.c_size_t(0), C.c_size_t(0)
        _check(self.lib.femgpu_index_sizes(self.h, C.byref(ln), C.byref(on)))
        lut, occ = np.empty(ln.value, dtype=np.uint32), np.empty(on.value, dtype=np.uint64)
        _check(self.lib.femgpu_download_index(self.h, lut.ctypes.data, lut.size, occ.ctypes.data if occ.size else None, occ.size))
        return self.lib.femgpu_index_k(self.h), self.lib.femgpu_index_step(self.h), lut, occ

    def save_index(self, path):
        _check(self.lib.femgpu_save_index_file(self.h, os.fsencode(path)))

    def upload_reference(self, records):
        """records: list of (name, bytes)"""
        n = len(records)
        seqs = (C.c_char_p * n)(*[bytes(r[1]) for r in records])
        lens = (C.c_uint32 * n)(*[len(r[1]) for r in records])
        self._keep = (seqs, lens)
        _check(self.lib.femgpu_upload_reference(self.h, n, seqs, lens))
        names = (C.c_char_p * n)(*[r[0] if isinstance(r[0], bytes) else str(r[0]).encode() for r in records])
        _check(self.lib.femgpu_set_reference_names(self.h, n, names))

    def clone_resident_from(self, other):
        """index + reference + record names device to device from another Mapper (src/FEM_map.c:136-143 once per node)"""
        _check(self.lib.femgpu_clone_resident(self.h, other.h))

    # -- hot path -----------------------------------------------------------------------------
    @staticmethod
    def _params(e, a=1, k=0, step=0):
        return Params(k, step, e, a)

    def _collect(self, res, n_reads):
        first = np.ctypeslib.as_array(res.first, shape=(n_reads + 1,)).copy()
        n = int(first[-1])
        if n:
            maps = np.frombuffer(C.string_at(res.maps, n * 16), dtype=MAPPING_DTYPE).copy()
        else:
            maps = np.zeros(0, dtype=MAPPING_DTYPE)
        # device start/CIGAR per mapping (k_trace), parallel to maps; None when switched off
        self.last_alignments = (np.frombuffer(C.string_at(res.alns, n * 88), dtype=ALIGNMENT_DTYPE).copy()
                                if n and res.alns else None)
        # device SAM text (k_sam): per-read byte offsets, host flags, text; None without set_sam_inputs
        self.last_sam = None
        if res.sam_off:
            off = np.ctypeslib.as_array(res.sam_off, shape=(n_reads + 1,)).copy()
            host = np.ctypeslib.as_array(res.sam_host, shape=(max(n_reads, 1),))[:n_reads].copy()
            self.last_sam = (off, host, C.string_at(res.sam_text, int(off[-1])) if int(off[-1]) else b"")
        return first, maps, res.stats.as_tuple()

    def set_sam_inputs(self, slot, names, quals):
        """names: list of bytes; quals: list of bytes (same lengths as the reads) or None per read.
        Applies to the next submit/map_batch on the slot (map_batch uses slot 0)."""
        n = len(names)
        name_off = np.zeros(n + 1, dtype=np.uint64)
        np.cumsum(np.fromiter((len(x) + 1 for x in names), dtype=np.uint64, count=n), out=name_off[1:])
        nm = np.frombuffer(b"".join(x + b"\0" for x in names) or b"\0", dtype=np.uint8)
        has = np.fromiter((q is not None for q in quals), dtype=np.uint8, count=n)
        return nm, name_off, has

    def map_batch_sam(self, reads, names, quals, e, a=1, k=0, step=0):
        """map_batch with the SAM text formatted on the device: -> (first, maps, stats); text in last_sam."""
        b = ReadBatch(reads)
        nm, name_off, has = self.set_sam_inputs(0, names, quals)
        q = np.frombuffer(b"".join(qq if qq is not None else b" " * len(r) for qq, r in zip(quals, reads)) or b" ", dtype=np.uint8)
        si = SamInputs(nm.ctypes.data, name_off.ctypes.data, q.ctypes.data, has.ctypes.data)
        _check(self.lib.femgpu_set_sam_inputs(self.h, 0, C.byref(si)))
        return self.map_batch(b, e, a, k, step)

    def map_batch(self, reads, e, a=1, k=0, step=0):
        """-> (first[n+1], maps (structured array, emission order), stats tuple)"""
        b = ReadBatch(reads)
        p = self._params(e, a, k, step)
        cb = b.c_struct()
        res = Result()
        _check(self.lib.femgpu_map_batch(self.h, C.byref(p), C.byref(cb), C.byref(res)))
        return self._collect(res, b.n_reads)

    def submit(self, slot, reads, e, a=1, k=0, step=0):
        b = ReadBatch(reads)
        p = self._params(e, a, k, step)
        cb = b.c_struct()
        self._inflight = getattr(self, "_inflight", {})
        self._inflight[slot] = b
        _check(self.lib.femgpu_submit_batch(self.h, slot, C.byref(p), C.byref(cb)))

    def wait(self, slot):
        res = Result()
        _check(self.lib.femgpu_wait_batch(self.h, slot, C.byref(res)))
        b = self._inflight.pop(slot)
        return self._collect(res, b.n_reads)

    def generate_group_seeding_candidates(self, reads, e, a=1, k=0, step=0):
        """-> (first[2n+1], cands, before[2n], status[2n]); read-strand 2*i+s = strand s of read i."""
        b = ReadBatch(reads)
        p = self._params(e, a, k, step)
        cb = b.c_struct()
        out = Candidates()
        _check(self.lib.femgpu_candidates_batch(self.h, C.byref(p), C.byref(cb), C.byref(out)))
        n_rs = 2 * b.n_reads
        first = np.ctypeslib.as_array(out.first, shape=(n_rs + 1,)).copy()
        n = int(first[-1])
        cands = np.ctypeslib.as_array(out.cands, shape=(max(n, 1),))[:n].copy()
        before = np.ctypeslib.as_array(out.before, shape=(max(n_rs, 1),))[:n_rs].copy()
        status = np.ctypeslib.as_array(out.status, shape=(max(n_rs, 1),))[:n_rs].copy()
        return first, cands, before, status

    def last_timings(self):
        ms = (C.c_float * 6)()
        n = C.c_uint32(0)
        _check(self.lib.femgpu_last_timings(self.h, C.byref(ms), C.byref(n)))
        return list(ms), n.value

    def last_work(self):
        """(verify columns executed, candidates, selected-seed occurrences, mappings) of the last batch"""
        w = (C.c_uint64 * 4)()
        _check(self.lib.femgpu_last_work(self.h, C.byref(w)))
        return [int(x) for x in w]

    def int32_peak(self):
        """10^9 instr/s: (LOP3, IADD3, shift+logic)"""
        g = (C.c_double * 3)()
        _check(self.lib.femgpu_measure_int32_peak(self.h, C.byref(g)))
        return list(g)

    def rerun_resident(self):
        tot = C.c_float(0)
        st = (C.c_float * 6)()
        _check(self.lib.femgpu_rerun_resident(self.h, C.byref(tot), C.byref(st)))
        return tot.value, list(st)
