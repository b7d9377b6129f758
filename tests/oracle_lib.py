"""ctypes binding of oracle/_ref/liboracle.so (the CPU restatement, TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
"""
import ctypes as C
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB_PATH = os.path.join(ORACLE_DIR, "_ref", "liboracle.so")
REF_ASIS = os.path.join(ORACLE_DIR, "_ref", "FEM_ref_asis")
REF_HDR = os.path.join(ORACLE_DIR, "_ref", "FEM_ref_hdr")


class Mapping(C.Structure):
    _fields_ = [("candidate_position", C.c_uint64), ("end_position_offset", C.c_int32),
                ("direction", C.c_uint8), ("edit_distance", C.c_uint8), ("pad_", C.c_uint8 * 2)]


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("num_reads", "num_mapped_reads", "num_candidates_before_filter",
                                          "num_candidates", "num_mappings", "num_infeasible_reads")]

    def as_tuple(self):
        return tuple(getattr(self, n) for n, _ in self._fields_)


_lib = None


def build():
    """Compile the restatement (plain gcc, no reference tree needed)."""
    src = os.path.join(ORACLE_DIR, "fem_oracle.c")
    if (not os.path.exists(LIB_PATH)) or os.path.getmtime(LIB_PATH) < max(
            os.path.getmtime(src), os.path.getmtime(os.path.join(ORACLE_DIR, "fem_oracle.h"))):
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "port"])
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        vp, u32, u64, i32, sz = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int, C.c_size_t
        sig = {
            "fo_open_reference": (vp, [C.c_char_p]),
            "fo_create_reference": (vp, [u32, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p), C.POINTER(u32)]),
            "fo_close": (None, [vp]),
            "fo_num_records": (u32, [vp]),
            "fo_record_length": (u32, [vp, u32]),
            "fo_record_name": (C.c_char_p, [vp, u32]),
            "fo_record_bases": (vp, [vp, u32]),
            "fo_build_index": (i32, [vp, i32, i32]),
            "fo_load_index": (i32, [vp, C.c_char_p]),
            "fo_save_index": (i32, [vp, C.c_char_p]),
            "fo_index_k": (i32, [vp]),
            "fo_index_step": (i32, [vp]),
            "fo_index_num_occurrences": (sz, [vp]),
            "fo_index_lookup_table": (vp, [vp]),
            "fo_index_occurrences": (vp, [vp]),
            "fo_set_params": (i32, [vp, i32, i32, i32, i32]),
            "fo_reverse_complement": (None, [C.c_char_p, i32, C.c_char_p]),
            "fo_hash_all_seeds": (i32, [C.c_char_p, i32, i32, C.POINTER(u32)]),
            "fo_generate_candidates": (C.c_long, [vp, C.c_char_p, i32, C.POINTER(C.POINTER(u64)), C.POINTER(sz), C.POINTER(u32)]),
            "fo_verify_candidates": (u32, [vp, C.c_char_p, i32, i32, C.POINTER(u64), u32, C.POINTER(Mapping)]),
            "fo_banded_edit_distance": (i32, [vp, u64, C.c_char_p, i32, i32, C.POINTER(i32)]),
            "fo_sort_mappings": (None, [C.POINTER(Mapping), sz]),
            "fo_generate_alignment": (i32, [vp, u64, C.c_char_p, i32, i32, i32, C.POINTER(u32), C.POINTER(i32), C.c_char_p]),
            "fo_map_read": (sz, [vp, C.c_char_p, i32, C.POINTER(C.POINTER(Mapping)), C.POINTER(sz), C.POINTER(Stats)]),
            "fo_map_file": (i32, [vp, C.c_char_p, C.c_char_p, C.POINTER(Stats)]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class Oracle:
    """Reference records + index + parameters; thin object wrapper over the fo_* functions."""

    def __init__(self, fasta=None, records=None):
        L = lib()
        if fasta is not None:
            self.h = L.fo_open_reference(os.fsencode(fasta))
        else:
            n = len(records)
            names = (C.c_char_p * n)(*[r[0].encode() for r in records])
            seqs = (C.c_char_p * n)(*[r[1].encode() if isinstance(r[1], str) else bytes(r[1]) for r in records])
            lens = (C.c_uint32 * n)(*[len(r[1]) for r in records])
            self.h = L.fo_create_reference(n, names, seqs, lens)
        if not self.h:
            raise RuntimeError("oracle: cannot open reference")
        self.e = 2

    def close(self):
        if self.h:
            lib().fo_close(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- index ---------------------------------------------------------------------------
    def build_index(self, k, step):
        assert lib().fo_build_index(self.h, k, step) == 0

    def load_index(self, path):
        assert lib().fo_load_index(self.h, os.fsencode(path)) == 0, path

    def save_index(self, path):
        assert lib().fo_save_index(self.h, os.fsencode(path)) == 0

    def index_arrays(self):
        import numpy as np
        L = lib()
        k, n = L.fo_index_k(self.h), L.fo_index_num_occurrences(self.h)
        lut = np.ctypeslib.as_array(C.cast(L.fo_index_lookup_table(self.h), C.POINTER(C.c_uint32)), shape=((1 << (2 * k)) + 1,))
        occ = np.ctypeslib.as_array(C.cast(L.fo_index_occurrences(self.h), C.POINTER(C.c_uint64)), shape=(max(n, 1),))[:n]
        return k, L.fo_index_step(self.h), lut, occ

    def records(self):
        L = lib()
        out = []
        for i in range(L.fo_num_records(self.h)):
            ln = L.fo_record_length(self.h, i)
            out.append((L.fo_record_name(self.h, i).decode(), C.string_at(L.fo_record_bases(self.h, i), ln)))
        return out

    def set_params(self, e, a=1, k=0, step=0):
        assert lib().fo_set_params(self.h, k, step, e, a) == 0
        self.e = e

    # -- stages ----------------------------------------------------------------------------
    def candidates(self, text: bytes):
        L = lib()
        buf = C.POINTER(C.c_uint64)()
        cap = C.c_size_t(0)
        before = C.c_uint32(0)
        n = L.fo_generate_candidates(self.h, text, len(text), C.byref(buf), C.byref(cap), C.byref(before))
        out = [buf[i] for i in range(max(n, 0))]
        if buf:
            libc = C.CDLL(None)
            libc.free.argtypes = [C.c_void_p]
            libc.free(buf)
        return n, out, before.value

    def verify(self, text: bytes, direction, cands):
        L = lib()
        n = len(cands)
        arr = (C.c_uint64 * max(n, 1))(*cands)
        out = (Mapping * max(n, 1))()
        got = L.fo_verify_candidates(self.h, text, len(text), direction, arr, n, out)
        return [(m.candidate_position, m.end_position_offset, m.direction, m.edit_distance) for m in out[:got]]

    def map_read(self, seq: bytes, stats=None):
        """-> sorted mappings [(cand, end_off, dir, dist)] exactly as src/align.c:57 leaves them."""
        L = lib()
        maps = C.POINTER(Mapping)()
        cap = C.c_size_t(0)
        st = stats if stats is not None else Stats()
        n = L.fo_map_read(self.h, seq, len(seq), C.byref(maps), C.byref(cap), C.byref(st))
        out = [(maps[i].candidate_position, maps[i].end_position_offset, maps[i].direction, maps[i].edit_distance)
               for i in range(n)]
        if maps:
            libc = C.CDLL(None)
            libc.free.argtypes = [C.c_void_p]
            libc.free(maps)
        return out

    def alignment(self, cand, text: bytes, dist, end_off):
        L = lib()
        l = len(text)
        cigar = (C.c_uint32 * (l + 2))()
        nc = C.c_int(0)
        md = C.create_string_buffer(4 * l + 64)
        start = L.fo_generate_alignment(self.h, cand, text, l, dist, end_off, cigar, C.byref(nc), md)
        return start, list(cigar[:nc.value]), md.value.decode()

    def map_file(self, fastq, sam):
        st = Stats()
        rc = lib().fo_map_file(self.h, os.fsencode(fastq), os.fsencode(sam), C.byref(st))
        assert rc == 0, rc
        return st


def revcomp(seq: bytes) -> bytes:
    out = C.create_string_buffer(len(seq) + 1)
    lib().fo_reverse_complement(seq, len(seq), out)
    return out.raw[:len(seq)]


def have_reference_binaries():
    return os.access(REF_ASIS, os.X_OK) and os.access(REF_HDR, os.X_OK)


def run_reference(binary, ref_fa, index, reads_fq, out_sam, e, a=1, threads=1):
    """Run the compiled reference `FEM map`; returns the five stderr counters."""
    p = subprocess.run([binary, "map", "-e", str(e), "-a", str(a), "-t", str(threads), "--ref", ref_fa,
                        "--index", index, "--read1", reads_fq, "-o", out_sam],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    if p.returncode != 0:
        raise RuntimeError("reference failed rc=%d: %s" % (p.returncode, p.stderr[-400:]))
    vals = {}
    for line in p.stderr.splitlines():
        for key, tag in (("num_reads", "The number of read:"), ("num_mapped_reads", "The number of mapped read:"),
                         ("before", "The number of candidate before additional q-gram filter:"),
                         ("num_candidates", "The number of candidate:"), ("num_mappings", "The number of mapping:"),
                         ("time", "Time:")):
            if line.startswith(tag):
                vals[key] = line[len(tag):].strip().rstrip("s")
    return (int(vals["num_reads"]), int(vals["num_mapped_reads"]), int(vals["before"]),
            int(vals["num_candidates"]), int(vals["num_mappings"])), float(vals["time"])


def run_reference_index(binary, k, step, ref_fa, out):
    subprocess.run([binary, "index", str(k), str(step), ref_fa, out], check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
