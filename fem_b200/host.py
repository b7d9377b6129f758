"""ctypes binding of fem_b200/lib/libfemhost.so (include/femhost.h): the host side of `FEM map`.

  Records / read_records      <->  SequenceBatch loaders, src/sequence_batch.c:44-121
  Index.build / save / load   <->  src/index.c:57-168
  sam_header / sam_read       <->  src/output_queue.c:93-116, src/align.c:56-92 (+ htslib text)
  HostMapper.map_file         <->  src/FEM_map.c:136-219 (the whole `FEM map` run)

The library links libfemgpu.so; loading it needs no GPU, mapping does (no CPU fallback).
"""
import ctypes as C
import os

import numpy as np

from .api import MAPPING_DTYPE, FemGpuError, Params, Stats, load_library

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libfemhost.so")
FEM_BINARY = os.path.join(_HERE, "bin", "FEM")


class RecordsStruct(C.Structure):
    _fields_ = [("n_records", C.c_uint32), ("n_bases", C.c_uint64), ("bases", C.c_void_p), ("quals", C.c_void_p),
                ("base_off", C.POINTER(C.c_uint64)), ("has_qual", C.POINTER(C.c_uint8)), ("names", C.c_void_p),
                ("name_off", C.POINTER(C.c_uint64)), ("cap_records", C.c_size_t), ("cap_bases", C.c_size_t),
                ("cap_names", C.c_size_t), ("pinned", C.c_int)]


class IndexStruct(C.Structure):
    _fields_ = [("k", C.c_int32), ("step", C.c_int32), ("lut", C.POINTER(C.c_uint32)), ("lut_n", C.c_size_t),
                ("occ", C.POINTER(C.c_uint64)), ("n_occ", C.c_size_t)]


class StrStruct(C.Structure):
    _fields_ = [("s", C.c_void_p), ("l", C.c_size_t), ("m", C.c_size_t)]


class MapOptions(C.Structure):
    _fields_ = [("params", Params), ("n_threads", C.c_int32), ("batch_reads", C.c_uint32), ("skip_reads", C.c_uint64),
                ("max_reads", C.c_uint64), ("no_header", C.c_int32), ("no_reads", C.c_int32), ("shard_index", C.c_uint32),
                ("shard_count", C.c_uint32)]


class MapTiming(C.Structure):
    _fields_ = [("wall", C.c_double), ("parse", C.c_double), ("gpu_wait", C.c_double), ("format", C.c_double),
                ("write", C.c_double), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64), ("sam_bytes", C.c_uint64),
                ("n_batches", C.c_uint32), ("n_launches", C.c_uint32), ("device_sam_reads", C.c_uint64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


EXPORTS = {
    "femhost_last_error": (C.c_char_p, []),
    "femhost_records_init": (None, [C.POINTER(RecordsStruct), C.c_int]),
    "femhost_records_free": (None, [C.POINTER(RecordsStruct)]),
    "femhost_reader_open": (C.c_void_p, [C.c_char_p]),
    "femhost_reader_close": (None, [C.c_void_p]),
    "femhost_reader_next": (C.c_int64, [C.c_void_p, C.c_uint32, C.POINTER(RecordsStruct)]),
    "femhost_load_all": (C.c_int, [C.c_char_p, C.POINTER(RecordsStruct)]),
    "femhost_load_all_ranges": (C.c_int, [C.c_char_p, C.c_uint64, C.POINTER(RecordsStruct)]),
    "femhost_load_shard": (C.c_int, [C.c_char_p, C.c_uint32, C.c_uint32, C.POINTER(RecordsStruct)]),
    "femhost_count_records": (C.c_int64, [C.c_char_p]),
    "femhost_index_build": (C.c_int, [C.POINTER(RecordsStruct), C.c_int, C.c_int, C.POINTER(IndexStruct)]),
    "femhost_index_build_device": (C.c_int, [C.POINTER(RecordsStruct), C.c_int, C.c_int, C.c_int, C.c_char_p, C.POINTER(C.c_double * 3)]),
    "femhost_index_save": (C.c_int, [C.POINTER(IndexStruct), C.c_char_p]),
    "femhost_index_load": (C.c_int, [C.c_char_p, C.POINTER(IndexStruct)]),
    "femhost_index_free": (None, [C.POINTER(IndexStruct)]),
    "femhost_str_free": (None, [C.POINTER(StrStruct)]),
    "femhost_sam_header": (C.c_int, [C.POINTER(RecordsStruct), C.POINTER(StrStruct)]),
    "femhost_alignment": (C.c_int, [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint32),
                                    C.POINTER(C.c_int), C.c_char_p]),
    "femhost_sam_read": (C.c_int, [C.POINTER(RecordsStruct), C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_int,
                                   C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(StrStruct)]),
    "femhost_mapper_create": (C.c_void_p, [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_int)]),
    "femhost_mapper_destroy": (None, [C.c_void_p]),
    "femhost_mapper_reference": (C.POINTER(RecordsStruct), [C.c_void_p]),
    "femhost_mapper_context": (C.c_void_p, [C.c_void_p, C.c_int]),
    "femhost_map_file": (C.c_int, [C.c_void_p, C.POINTER(MapOptions), C.c_char_p, C.c_char_p, C.POINTER(Stats),
                                   C.POINTER(MapTiming)]),
}

_lib = None


def load_host_library():
    global _lib
    if _lib is None:
        load_library()  # libfemgpu.so first (the host library links it); raises when missing
        if not os.path.exists(LIB_PATH):
            raise FemGpuError("%s is missing: build with `make -C fem_b200/host`" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in EXPORTS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def _err():
    return load_host_library().femhost_last_error().decode(errors="replace")


class Records:
    """Flattened sequence records owned by the C library."""

    def __init__(self):
        self.lib = load_host_library()
        self.c = RecordsStruct()
        self.lib.femhost_records_init(C.byref(self.c), 0)

    def __len__(self):
        return self.c.n_records

    def seq(self, i):
        lo, hi = self.c.base_off[i], self.c.base_off[i + 1]
        return C.string_at(self.c.bases + lo, hi - lo)

    def qual(self, i):
        if not self.c.has_qual[i]:
            return None
        lo, hi = self.c.base_off[i], self.c.base_off[i + 1]
        return C.string_at(self.c.quals + lo, hi - lo)

    def name(self, i):
        return C.string_at(self.c.names + self.c.name_off[i])

    def as_batch(self):
        """femgpu_batch view of these records (no copy; valid while the records live)."""
        from .api import Batch
        return Batch(self.c.bases, C.cast(self.c.base_off, C.c_void_p), self.c.n_records)

    def as_list(self):
        return [(self.name(i), self.seq(i), self.qual(i)) for i in range(len(self))]

    def close(self):
        if self.c is not None:
            self.lib.femhost_records_free(C.byref(self.c))
            self.c = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def load_all(path):
    r = Records()
    if r.lib.femhost_load_all(os.fsencode(path), C.byref(r.c)) != 0:
        raise FemGpuError(_err())
    return r


def load_all_ranges(path, chunk_bytes):
    """-> Records, or None when the file is not plain four-line FASTQ (parallel reader's fallback)."""
    r = Records()
    rc = r.lib.femhost_load_all_ranges(os.fsencode(path), chunk_bytes, C.byref(r.c))
    if rc == 1:
        r.close()
        return None
    if rc != 0:
        raise FemGpuError(_err())
    return r


ENOTFASTQ4 = -5  # femhost.h FEMHOST_ENOTFASTQ4


class NotFastq4(FemGpuError):
    """byte-range shards do not apply to this input: every rank falls back to read-count shards"""


def load_shard(path, shard_index, shard_count):
    """-> Records of byte-range shard shard_index of shard_count (the cut HostMapper.map_file(shard=...) makes),
    or None when the file is not plain four-line FASTQ."""
    r = Records()
    rc = r.lib.femhost_load_shard(os.fsencode(path), shard_index, shard_count, C.byref(r.c))
    if rc == 1:
        r.close()
        return None
    if rc != 0:
        raise FemGpuError(_err())
    return r


def count_records(path):
    """records the readers deliver for any accepted input (FASTA/FASTQ, multi-line, gzip)"""
    n = load_host_library().femhost_count_records(os.fsencode(path))
    if n < 0:
        raise FemGpuError(_err())
    return int(n)


def read_batches(path, batch):
    """Generator over lists of (name, seq, qual) in batches of `batch` records."""
    lib = load_host_library()
    rd = lib.femhost_reader_open(os.fsencode(path))
    if not rd:
        raise FemGpuError(_err())
    r = Records()
    try:
        while True:
            n = lib.femhost_reader_next(rd, batch, C.byref(r.c))
            if n < 0:
                raise FemGpuError(_err())
            if n == 0:
                break
            yield r.as_list()
    finally:
        lib.femhost_reader_close(rd)
        r.close()


class Index:
    def __init__(self):
        self.lib = load_host_library()
        self.c = IndexStruct()

    @classmethod
    def build(cls, records: Records, k, step):
        ix = cls()
        if ix.lib.femhost_index_build(C.byref(records.c), k, step, C.byref(ix.c)) != 0:
            raise FemGpuError(_err())
        return ix

    @staticmethod
    def build_on_device(records: Records, k, step, path, device=0):
        """`FEM index ... --gpu`: the same file from femgpu_build_index; returns (upload, build, save) seconds"""
        sec = (C.c_double * 3)()
        if load_host_library().femhost_index_build_device(C.byref(records.c), k, step, device, os.fsencode(path), C.byref(sec)) != 0:
            raise FemGpuError(_err())
        return tuple(sec)

    @classmethod
    def load(cls, path):
        ix = cls()
        if ix.lib.femhost_index_load(os.fsencode(path), C.byref(ix.c)) != 0:
            raise FemGpuError(_err())
        return ix

    def save(self, path):
        if self.lib.femhost_index_save(C.byref(self.c), os.fsencode(path)) != 0:
            raise FemGpuError(_err())

    def arrays(self):
        lut = np.ctypeslib.as_array(self.c.lut, shape=(self.c.lut_n,))
        occ = np.ctypeslib.as_array(self.c.occ, shape=(max(self.c.n_occ, 1),))[:self.c.n_occ]
        return self.c.k, self.c.step, lut, occ

    def close(self):
        if self.c is not None:
            self.lib.femhost_index_free(C.byref(self.c))
            self.c = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def sam_header(ref: Records) -> bytes:
    lib = load_host_library()
    s = StrStruct()
    if lib.femhost_sam_header(C.byref(ref.c), C.byref(s)) != 0:
        raise FemGpuError(_err())
    out = C.string_at(s.s, s.l) if s.l else b""
    lib.femhost_str_free(C.byref(s))
    return out


def sam_read(ref: Records, name: bytes, seq: bytes, qual: bytes, e: int, maps: np.ndarray, alns=None) -> bytes:
    """SAM lines of one read; `maps` (MAPPING_DTYPE, emission order) is sorted in place, `alns`
    (ALIGNMENT_DTYPE, the device's start/CIGAR per mapping, optional) follows the permutation."""
    lib = load_host_library()
    assert maps.dtype == MAPPING_DTYPE and maps.flags.c_contiguous
    s = StrStruct()
    rc = lib.femhost_sam_read(C.byref(ref.c), name, seq, qual, len(seq), e, maps.ctypes.data,
                              alns.ctypes.data if alns is not None else None, maps.shape[0], C.byref(s))
    out = C.string_at(s.s, s.l) if s.l else b""
    lib.femhost_str_free(C.byref(s))
    if rc != 0:
        raise FemGpuError(_err())
    return out


def alignment(window: bytes, text: bytes, e, dist, end_off):
    lib = load_host_library()
    l = len(text)
    cigar = (C.c_uint32 * (l + 2))()
    nc = C.c_int(0)
    md = C.create_string_buffer(4 * l + 64)
    start = lib.femhost_alignment(window, text, l, e, dist, end_off, cigar, C.byref(nc), md)
    return start, list(cigar[:nc.value]), md.value.decode()


class HostMapper:
    """Reference + index resident on the GPUs; map_file = one whole `FEM map` run."""

    def __init__(self, ref_path, index_path, n_gpus=1, gpu_ids=None):
        self.lib = load_host_library()
        ids = (C.c_int * n_gpus)(*gpu_ids) if gpu_ids else None
        self.h = self.lib.femhost_mapper_create(os.fsencode(ref_path), os.fsencode(index_path), n_gpus, ids)
        if not self.h:
            raise FemGpuError(_err())

    def map_file(self, reads_path, sam_path, e, a=1, threads=1, k=0, step=0, batch_reads=0, skip_reads=0, max_reads=0,
                 no_header=False, no_reads=False, shard=None):
        """shard=(index, count): byte-range shard of a plain four-line FASTQ file (raises NotFastq4 otherwise);
        skip_reads / max_reads / no_reads: read-count shard of any input."""
        si, sc = shard if shard else (0, 0)
        opt = MapOptions(Params(k, step, e, a), threads, batch_reads, skip_reads, max_reads, int(no_header), int(no_reads), si, sc)
        st, tm = Stats(), MapTiming()
        rc = self.lib.femhost_map_file(self.h, C.byref(opt), os.fsencode(reads_path),
                                       os.fsencode(sam_path) if sam_path else None, C.byref(st), C.byref(tm))
        if rc == ENOTFASTQ4:
            raise NotFastq4(_err())
        if rc != 0:
            raise FemGpuError(_err())
        return st.as_tuple(), tm.as_dict()

    def close(self):
        if getattr(self, "h", None):
            self.lib.femhost_mapper_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
