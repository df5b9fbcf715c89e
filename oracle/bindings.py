"""ctypes bindings for the CPU oracle — TEST INFRASTRUCTURE ONLY.

Two libraries live behind this module:

* ``liboracle_tetrex.so``  — oracle/tetrex_oracle.c, our plain-C restatement (``orc_*``).
* ``_ref/libtetrex_ref.so`` — the UNMODIFIED reference (TetRex + lib/hibf + lib/re2) compiled from
  /root/reference by ``make -C oracle ref`` plus oracle/ref_harness.cpp (``ref_*``).  Present only
  where it was built (the dev container; the built .so travels to the GPU box).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_ORC_PATH = os.path.join(HERE, "liboracle_tetrex.so")
_REF_PATH = os.path.join(HERE, "_ref", "libtetrex_ref.so")

u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)
i32p = C.POINTER(C.c_int32)


def _p(a, t):
    return a.ctypes.data_as(t)


def build_oracle(force: bool = False) -> str:
    src = os.path.join(HERE, "tetrex_oracle.c")
    if force or not os.path.exists(_ORC_PATH) or os.path.getmtime(_ORC_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "oracle"], stdout=subprocess.DEVNULL)
    return _ORC_PATH


_orc = None
_ref = None


def orc():
    global _orc
    if _orc is None:
        lib = C.CDLL(build_oracle())
        lib.orc_decompose.restype = C.c_uint64
        lib.orc_decompose.argtypes = [C.c_int, C.c_int, C.c_int, u8p, C.c_uint64, u64p, C.c_uint64]
        lib.orc_update_kmer.restype = C.c_uint64
        lib.orc_update_kmer.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, u64p]
        lib.orc_peptide_maps.restype = None
        lib.orc_peptide_maps.argtypes = [C.c_int, u8p, u8p]
        lib.orc_compute_bitcount.restype = C.c_uint64
        lib.orc_compute_bitcount.argtypes = [C.c_uint64, C.c_float]
        lib.orc_hash_row.restype = C.c_uint64
        lib.orc_hash_row.argtypes = [C.c_uint64, C.c_int, C.c_uint64]
        lib.orc_emplace.restype = None
        lib.orc_emplace.argtypes = [u64p, C.c_uint64, C.c_uint64, C.c_int, u64p, u32p, C.c_uint64]
        lib.orc_bulk_contains.restype = None
        lib.orc_bulk_contains.argtypes = [u64p, C.c_uint64, C.c_uint64, C.c_int, u64p, C.c_uint64, u64p, C.c_int]
        lib.orc_bulk_contains_checksum.restype = C.c_uint64
        lib.orc_bulk_contains_checksum.argtypes = [u64p, C.c_uint64, C.c_uint64, C.c_int, u64p, C.c_uint64, C.c_int]
        lib.orc_hibf_membership.restype = None
        lib.orc_hibf_membership.argtypes = [C.c_uint64, C.c_uint64, u64p, u64p, u64p, C.POINTER(u64p), C.POINTER(u64p),
                                            C.POINTER(u64p), u64p, u64p, C.c_uint64, u64p, C.c_int]
        lib.orc_emplace_records.restype = None
        lib.orc_emplace_records.argtypes = [u64p, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                            u64p, u32p, C.c_uint64, C.c_int]
        _orc = lib
    return _orc


def have_ref() -> bool:
    return os.path.exists(_REF_PATH)


def ref():
    global _ref
    if _ref is None:
        lib = C.CDLL(_REF_PATH)
        lib.ref_last_error.restype = C.c_char_p
        lib.ref_decompose.restype = C.c_uint64
        lib.ref_decompose.argtypes = [C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_uint64, u64p, C.c_uint64]
        lib.ref_update_kmer.restype = C.c_uint64
        lib.ref_update_kmer.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, u64p]
        lib.ref_peptide_maps.restype = None
        lib.ref_peptide_maps.argtypes = [C.c_int, C.c_int, u8p, C.c_char_p]
        lib.ref_compute_bitcount.restype = C.c_uint64
        lib.ref_compute_bitcount.argtypes = [C.c_uint64, C.c_float]
        lib.ref_ibf_new.restype = C.c_void_p
        lib.ref_ibf_new.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32]
        lib.ref_ibf_free.argtypes = [C.c_void_p]
        lib.ref_ibf_params.argtypes = [C.c_void_p, u64p]
        lib.ref_ibf_words.restype = u64p
        lib.ref_ibf_words.argtypes = [C.c_void_p]
        lib.ref_ibf_emplace.argtypes = [C.c_void_p, u64p, u32p, C.c_uint64]
        lib.ref_ibf_bulk_contains.argtypes = [C.c_void_p, u64p, C.c_uint64, u64p, C.c_int]
        lib.ref_ibf_bulk_contains_checksum.restype = C.c_uint64
        lib.ref_ibf_bulk_contains_checksum.argtypes = [C.c_void_p, u64p, C.c_uint64, C.c_int]
        lib.ref_hibf_build.restype = C.c_void_p
        lib.ref_hibf_build.argtypes = [u64p, u64p, C.c_uint64, C.c_uint32, C.c_float, C.c_int]
        lib.ref_hibf_free.argtypes = [C.c_void_p]
        lib.ref_hibf_layout.restype = C.c_uint64
        lib.ref_hibf_layout.argtypes = [u64p, u64p, C.c_uint64, C.c_uint32, C.c_float, C.c_int, u64p, C.c_uint64]
        lib.ref_hibf_ibf_count.restype = C.c_uint64
        lib.ref_hibf_ibf_count.argtypes = [C.c_void_p]
        lib.ref_hibf_ibf_params.argtypes = [C.c_void_p, C.c_uint64, u64p]
        lib.ref_hibf_ibf_words.restype = u64p
        lib.ref_hibf_ibf_words.argtypes = [C.c_void_p, C.c_uint64]
        lib.ref_hibf_tables.restype = C.c_uint64
        lib.ref_hibf_tables.argtypes = [C.c_void_p, C.c_uint64, u64p, u64p, C.c_uint64]
        lib.ref_hibf_membership.argtypes = [C.c_void_p, u64p, C.c_uint64, u64p, C.c_int]
        lib.ref_index_build.restype = C.c_int
        lib.ref_index_build.argtypes = [C.POINTER(C.c_char_p), C.c_uint64, C.c_int, C.c_float, C.c_int, C.c_int,
                                        C.c_int, C.c_int, C.c_int, C.c_char_p]
        lib.ref_index_query.restype = C.c_int
        lib.ref_index_query.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, u64p, C.c_uint64, C.c_char_p]
        lib.ref_index_query_dibf.restype = C.c_int
        lib.ref_index_query_dibf.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.c_int, u64p, C.c_uint64,
                                             C.c_char_p]
        lib.ref_dindex_build.restype = C.c_int
        lib.ref_dindex_build.argtypes = [C.POINTER(C.c_char_p), C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, C.c_float,
                                         C.c_char_p, u64p]
        lib.ref_kgraph.restype = C.c_int
        lib.ref_kgraph.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, i32p, i32p, i32p, i32p, C.c_uint64,
                                   C.c_char_p, C.c_uint64]
        lib.ref_collect_flat.restype = C.c_int
        lib.ref_collect_flat.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int, u64p]
        lib.ref_collect_hibf.restype = C.c_int
        lib.ref_collect_hibf.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int, u64p]
        lib.ref_re2_find_all.restype = C.c_int64
        lib.ref_re2_find_all.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_uint64, u64p, u64p, C.c_uint64]
        lib.ref_hibf_from_tables.restype = C.c_void_p
        lib.ref_hibf_from_tables.argtypes = [C.c_uint64, C.c_uint64, u64p, u64p, u64p, C.POINTER(u64p), C.POINTER(u64p),
                                             C.POINTER(u64p), u64p]
        lib.ref_re2_scan_records.restype = C.c_int64
        lib.ref_re2_scan_records.argtypes = [C.c_char_p, C.c_int, C.c_void_p, u64p, C.c_uint64, C.c_int]
        lib.ref_re2_scan_records_hits.restype = C.c_int64
        lib.ref_re2_scan_records_hits.argtypes = [C.c_char_p, C.c_int, C.c_void_p, u64p, C.c_uint64, C.c_int, u32p, C.c_uint64]
        lib.ref_index_conjunction.restype = C.c_int
        lib.ref_index_conjunction.argtypes = [C.c_char_p, C.c_char_p, C.c_int, u64p, C.c_uint64, C.c_char_p]
        lib.ref_flat_set_script.restype = C.c_uint64
        lib.ref_flat_set_script.argtypes = [u64p, u8p, C.c_uint64, u64p]
        lib.ref_rh_pop_order.restype = C.c_uint64
        lib.ref_rh_pop_order.argtypes = [u64p, C.c_uint64, u32p]
        _ref = lib
    return _ref


# ----------------------------------------------------------------------------------------------- oracle (C port)
def decompose(molecule: int, k: int, reduction: int, seq: bytes) -> np.ndarray:
    buf = np.frombuffer(seq, dtype=np.uint8)
    cap = len(seq) + 2
    out = np.zeros(cap, dtype=np.uint64)
    n = orc().orc_decompose(molecule, k, reduction, _p(buf, u8p) if len(seq) else None, len(seq), _p(out, u64p), cap)
    return out[:n].copy()


def update_kmer(molecule: int, k: int, reduction: int, symbol: int, kmer: int):
    v = C.c_uint64(kmer)
    probe = orc().orc_update_kmer(molecule, k, reduction, symbol, C.byref(v))
    return probe, v.value


def peptide_maps(reduction: int):
    a = np.zeros(256, dtype=np.uint8)
    r = np.zeros(256, dtype=np.uint8)
    orc().orc_peptide_maps(reduction, _p(a, u8p), _p(r, u8p))
    return a, r


def compute_bitcount(n: int, fpr: float) -> int:
    return orc().orc_compute_bitcount(n, fpr)


class OracleIBF:
    """seqan::hibf::interleaved_bloom_filter restated: geometry + word array + emplace/bulk_contains."""

    def __init__(self, bins: int, bin_size: int, hash_funs: int):
        self.bins, self.bin_size, self.hash_funs = bins, bin_size, hash_funs
        self.bin_words = (bins + 63) // 64
        self.technical_bins = self.bin_words * 64
        self.hash_shift = 64 - int(bin_size).bit_length()
        self.words = np.zeros(self.bin_words * bin_size, dtype=np.uint64)

    def emplace(self, values: np.ndarray, bin_ids: np.ndarray):
        values = np.ascontiguousarray(values, dtype=np.uint64)
        bin_ids = np.ascontiguousarray(bin_ids, dtype=np.uint32)
        orc().orc_emplace(_p(self.words, u64p), self.bins, self.bin_size, self.hash_funs, _p(values, u64p),
                          _p(bin_ids, u32p), len(values))

    def emplace_records(self, molecule, k, reduction, text, rec_offsets, rec_bin, threads: int = 1):
        text = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
        offs = np.ascontiguousarray(rec_offsets, dtype=np.uint64)
        rb = np.ascontiguousarray(rec_bin, dtype=np.uint32)
        orc().orc_emplace_records(_p(self.words, u64p), self.bins, self.bin_size, self.hash_funs, molecule, k, reduction,
                                  text.ctypes.data, _p(offs, u64p), _p(rb, u32p), len(rb), threads)

    def bulk_contains(self, values: np.ndarray, threads: int = 1) -> np.ndarray:
        values = np.ascontiguousarray(values, dtype=np.uint64)
        out = np.zeros((len(values), self.bin_words), dtype=np.uint64)
        orc().orc_bulk_contains(_p(self.words, u64p), self.bins, self.bin_size, self.hash_funs, _p(values, u64p),
                                len(values), _p(out, u64p), threads)
        return out

    def bulk_contains_checksum(self, values: np.ndarray, threads: int = 1) -> int:
        values = np.ascontiguousarray(values, dtype=np.uint64)
        return orc().orc_bulk_contains_checksum(_p(self.words, u64p), self.bins, self.bin_size, self.hash_funs,
                                                _p(values, u64p), len(values), threads)


class HIBFTables:
    """Flattened HIBF: per-IBF geometry, words and the two per-technical-bin tables."""

    def __init__(self, n_user_bins, bins, bin_size, hash_funs, words, next_ibf, user_bin):
        self.n_user_bins = int(n_user_bins)
        self.bins = np.ascontiguousarray(bins, dtype=np.uint64)
        self.bin_size = np.ascontiguousarray(bin_size, dtype=np.uint64)
        self.hash_funs = np.ascontiguousarray(hash_funs, dtype=np.uint64)
        self.words = [np.ascontiguousarray(w, dtype=np.uint64) for w in words]
        self.next_ibf = [np.ascontiguousarray(w, dtype=np.uint64) for w in next_ibf]
        self.user_bin = [np.ascontiguousarray(w, dtype=np.uint64) for w in user_bin]
        self.table_len = np.array([len(t) for t in self.user_bin], dtype=np.uint64)
        self.n_ibfs = len(self.words)
        self.out_words = (self.n_user_bins + 63) // 64

    def ptr_arrays(self):
        mk = lambda arrs: (u64p * len(arrs))(*[_p(a, u64p) for a in arrs])
        return mk(self.words), mk(self.next_ibf), mk(self.user_bin)

    def membership(self, values: np.ndarray, threads: int = 1) -> np.ndarray:
        values = np.ascontiguousarray(values, dtype=np.uint64)
        out = np.zeros((len(values), self.out_words), dtype=np.uint64)
        w, nx, ub = self.ptr_arrays()
        orc().orc_hibf_membership(self.n_ibfs, self.n_user_bins, _p(self.bins, u64p), _p(self.bin_size, u64p),
                                  _p(self.hash_funs, u64p), w, nx, ub, _p(self.table_len, u64p), _p(values, u64p),
                                  len(values), _p(out, u64p), threads)
        return out

    def save(self, path):
        d = {"n_user_bins": np.uint64(self.n_user_bins), "bins": self.bins, "bin_size": self.bin_size,
             "hash_funs": self.hash_funs}
        for i in range(self.n_ibfs):
            d[f"words_{i}"] = self.words[i]
            d[f"next_{i}"] = self.next_ibf[i]
            d[f"user_{i}"] = self.user_bin[i]
        np.savez_compressed(path, **d)

    @staticmethod
    def load(path):
        z = np.load(path)
        n = len(z["bins"])
        return HIBFTables(int(z["n_user_bins"]), z["bins"], z["bin_size"], z["hash_funs"],
                          [z[f"words_{i}"] for i in range(n)], [z[f"next_{i}"] for i in range(n)],
                          [z[f"user_{i}"] for i in range(n)])


# ------------------------------------------------------------------------------ regex oracle (small cases only)
def find_all_python(pattern: str, posix_longest: bool, text: bytes):
    """Non-overlapping FindAndConsume sequence restated with Python's `re` (src/query.cpp:194-237).

    Python's backtracking engine IS leftmost-first for the syntax subset TetRex accepts, so the DNA
    flavour is re.search in a loop.  Leftmost-longest (RE2::POSIX) is restated by brute force: the
    leftmost start that has any match, then the longest end that fullmatches — O(L^2), small inputs only.
    After a match the remaining text is searched as a fresh string, like RE2::FindAndConsume does.
    """
    import re
    rx = re.compile(("(" + pattern + ")").encode("latin-1"))
    out = []
    p = 0
    L = len(text)
    while p <= L:
        rest = text[p:]
        m = rx.search(rest)
        if not m:
            break
        s = m.start()
        e = m.end()
        if posix_longest:
            # m.start() is the leftmost start under any semantics; find the longest end from there
            anch = pattern.endswith("$") and not pattern.endswith("\\$")
            core = pattern[:-1] if anch else pattern
            core = core[1:] if core.startswith("^") else core
            fx = re.compile(("(?:" + core + ")").encode("latin-1"))
            best = None
            for end in range(len(rest), s, -1):
                if anch and end != len(rest):
                    break
                if fx.fullmatch(rest, s, end):
                    best = end
                    break
            if best is None:
                break
            e = best
        if e == s:
            raise ValueError("pattern matches the empty string")
        out.append((p + s, p + e))
        p += e
    return out


# ------------------------------------------------------------------------------------------ reference wrappers
def ref_decompose(molecule, k, reduction, seq: bytes) -> np.ndarray:
    cap = len(seq) + 2
    out = np.zeros(cap, dtype=np.uint64)
    n = ref().ref_decompose(molecule, k, reduction, seq, len(seq), _p(out, u64p), cap)
    return out[:n].copy()


def ref_find_all(pattern: str, posix_longest: bool, text: bytes, cap: int = 1 << 16):
    s = np.zeros(cap, dtype=np.uint64)
    e = np.zeros(cap, dtype=np.uint64)
    n = ref().ref_re2_find_all(pattern.encode(), int(posix_longest), text, len(text), _p(s, u64p), _p(e, u64p), cap)
    if n < 0:
        raise ValueError(ref().ref_last_error().decode())
    return [(int(a), int(b)) for a, b in zip(s[:n], e[:n])]


class RefIBF:
    def __init__(self, bins, bin_size, hash_funs):
        self.h = ref().ref_ibf_new(bins, bin_size, hash_funs)
        if not self.h:
            raise ValueError(ref().ref_last_error().decode())
        p = np.zeros(7, dtype=np.uint64)
        ref().ref_ibf_params(self.h, _p(p, u64p))
        (self.bins, self.technical_bins, self.bin_size, self.hash_shift, self.bin_words, self.hash_funs,
         self.n_words) = [int(x) for x in p]

    def __del__(self):
        if getattr(self, "h", None):
            ref().ref_ibf_free(self.h)
            self.h = None

    def emplace(self, values, bin_ids):
        values = np.ascontiguousarray(values, dtype=np.uint64)
        bin_ids = np.ascontiguousarray(bin_ids, dtype=np.uint32)
        ref().ref_ibf_emplace(self.h, _p(values, u64p), _p(bin_ids, u32p), len(values))

    @property
    def words(self):
        return np.ctypeslib.as_array(ref().ref_ibf_words(self.h), shape=(self.n_words,)).copy()

    def bulk_contains(self, values, threads=1):
        values = np.ascontiguousarray(values, dtype=np.uint64)
        out = np.zeros((len(values), self.bin_words), dtype=np.uint64)
        ref().ref_ibf_bulk_contains(self.h, _p(values, u64p), len(values), _p(out, u64p), threads)
        return out

    def bulk_contains_checksum(self, values, threads=1):
        values = np.ascontiguousarray(values, dtype=np.uint64)
        return ref().ref_ibf_bulk_contains_checksum(self.h, _p(values, u64p), len(values), threads)

    def collect(self, molecule, k, reduction, regex: str, augment=False):
        out = np.zeros(self.bin_words, dtype=np.uint64)
        n = ref().ref_collect_flat(self.h, molecule, k, reduction, regex.encode(), int(augment), _p(out, u64p))
        if n < 0:
            raise ValueError(ref().ref_last_error().decode())
        return out


def ref_scan_records(pattern: str, posix_longest: bool, text: np.ndarray, rec_offsets: np.ndarray, threads: int) -> int:
    """RE2 FindAndConsume over in-memory records (the verify inner loop), OpenMP over records."""
    text = np.ascontiguousarray(text, dtype=np.uint8)
    offs = np.ascontiguousarray(rec_offsets, dtype=np.uint64)
    n = ref().ref_re2_scan_records(pattern.encode(), int(posix_longest), text.ctypes.data, _p(offs, u64p), len(offs) - 1, threads)
    if n < 0:
        raise ValueError(ref().ref_last_error().decode())
    return n


class RefHIBF:
    """seqan::hibf::hierarchical_interleaved_bloom_filter built by the reference from per-bin k-mer lists."""

    @classmethod
    def from_tables(cls, t: "HIBFTables"):
        self = cls.__new__(cls)
        self.user_bins = t.n_user_bins
        w, nx, ub = t.ptr_arrays()
        self.h = ref().ref_hibf_from_tables(t.n_ibfs, t.n_user_bins, _p(t.bins, u64p), _p(t.bin_size, u64p),
                                            _p(t.hash_funs, u64p), w, nx, ub, _p(t.table_len, u64p))
        if not self.h:
            raise ValueError(ref().ref_last_error().decode())
        return self

    def __init__(self, per_bin_values, hash_funs=3, fpr=0.05, threads=1):
        offs = np.zeros(len(per_bin_values) + 1, dtype=np.uint64)
        offs[1:] = np.cumsum([len(v) for v in per_bin_values])
        vals = np.ascontiguousarray(np.concatenate(per_bin_values), dtype=np.uint64)
        self.user_bins = len(per_bin_values)
        self.h = ref().ref_hibf_build(_p(vals, u64p), _p(offs, u64p), self.user_bins, hash_funs, fpr, threads)
        if not self.h:
            raise ValueError(ref().ref_last_error().decode())

    def __del__(self):
        if getattr(self, "h", None):
            ref().ref_hibf_free(self.h)
            self.h = None

    def tables(self) -> HIBFTables:
        n = ref().ref_hibf_ibf_count(self.h)
        bins, bs, hf, words, nxt, ub = [], [], [], [], [], []
        for i in range(n):
            p = np.zeros(7, dtype=np.uint64)
            ref().ref_hibf_ibf_params(self.h, i, _p(p, u64p))
            bins.append(p[0]); bs.append(p[2]); hf.append(p[5])
            words.append(np.ctypeslib.as_array(ref().ref_hibf_ibf_words(self.h, i), shape=(int(p[6]),)).copy())
            cap = int(p[1])
            a = np.zeros(cap, dtype=np.uint64)
            b = np.zeros(cap, dtype=np.uint64)
            ln = ref().ref_hibf_tables(self.h, i, _p(a, u64p), _p(b, u64p), cap)
            nxt.append(a[:ln].copy()); ub.append(b[:ln].copy())
        return HIBFTables(self.user_bins, bins, bs, hf, words, nxt, ub)

    def membership(self, values, threads=1):
        values = np.ascontiguousarray(values, dtype=np.uint64)
        w = (self.user_bins + 63) // 64
        out = np.zeros((len(values), w), dtype=np.uint64)
        ref().ref_hibf_membership(self.h, _p(values, u64p), len(values), _p(out, u64p), threads)
        return out

    def collect(self, molecule, k, reduction, regex: str, augment=False):
        w = (self.user_bins + 63) // 64
        out = np.zeros(w, dtype=np.uint64)
        n = ref().ref_collect_hibf(self.h, molecule, k, reduction, regex.encode(), int(augment), _p(out, u64p))
        if n < 0:
            raise ValueError(ref().ref_last_error().decode())
        return out


def ref_kgraph(regex: str, molecule: int, k: int, reduction: int = 0, augment: bool = False):
    """The k-graph the reference builds (construct_kgraph / construct_reduced_kgraph + determine_top_sort):
    (postfix, [(symbol, first, second, rank) per node id]) or None if the reference threw."""
    cap = 1 << 17
    sym = np.zeros(cap, dtype=np.int32)
    fi = np.zeros(cap, dtype=np.int32)
    se = np.zeros(cap, dtype=np.int32)
    rk = np.zeros(cap, dtype=np.int32)
    pf = C.create_string_buffer(1 << 16)
    i32p = C.POINTER(C.c_int32)
    n = ref().ref_kgraph(regex.encode(), molecule, k, reduction, int(augment), sym.ctypes.data_as(i32p), fi.ctypes.data_as(i32p),
                         se.ctypes.data_as(i32p), rk.ctypes.data_as(i32p), cap, pf, 1 << 16)
    if n < 0:
        return None
    return pf.value.decode(), [[int(sym[i]), int(fi[i]), int(se[i]), int(rk[i])] for i in range(n)]


def parse_layout_dump(w: np.ndarray, n_bins: int) -> dict:
    """The word dump of ref_hibf_layout / trxh_hibf_layout as a dict."""
    w = [int(x) for x in w]
    it = iter(w)
    out = {"tmax": next(it), "top_level_max_bin_id": next(it), "max_bins": [], "user_bins": []}
    for _ in range(next(it)):
        n = next(it)
        out["max_bins"].append(([next(it) for _ in range(n)], next(it)))
    for _ in range(next(it)):
        n = next(it)
        out["user_bins"].append(([next(it) for _ in range(n)], next(it), next(it), next(it)))
    out["kmer_counts"] = [next(it) for _ in range(n_bins)]
    return out


def ref_hibf_layout(values: np.ndarray, bin_offsets: np.ndarray, hash_funs: int, fpr: float, threads: int = 1) -> dict:
    """seqan::hibf's layout (sketches, k-mer count estimates, DP, rearrangement) for user bins given as k-mer lists."""
    values = np.ascontiguousarray(values, dtype=np.uint64)
    offs = np.ascontiguousarray(bin_offsets, dtype=np.uint64)
    n_bins = len(offs) - 1
    out = np.zeros(64 + 16 * n_bins + 4096, dtype=np.uint64)
    n = ref().ref_hibf_layout(_p(values, u64p), _p(offs, u64p), n_bins, hash_funs, fpr, threads, _p(out, u64p), len(out))
    if n == 0:
        raise RuntimeError(ref().ref_last_error().decode())
    if n > len(out):
        out = np.zeros(n, dtype=np.uint64)
        ref().ref_hibf_layout(_p(values, u64p), _p(offs, u64p), n_bins, hash_funs, fpr, threads, _p(out, u64p), len(out))
    return parse_layout_dump(out[:n], n_bins)


def ref_index_build(files, k, fpr, hash_count, molecule, flavor, reduction, out_path, threads=1):
    """drive_index's work with the reference's own classes: flavor 0 = flat IBF, 1 = HIBF. Writes out_path."""
    arr = (C.c_char_p * len(files))(*[f.encode() for f in files])
    rc = ref().ref_index_build(arr, len(files), k, fpr, hash_count, molecule, flavor, reduction, threads, out_path.encode())
    if rc != 0:
        raise RuntimeError(ref().ref_last_error().decode())


def ref_dindex_build(files, min_gap, max_gap, hash_count, fpr, out_path) -> np.ndarray:
    """drive_dindex's work with the reference's DGramIndex; writes out_path, returns the codes each bin buffered."""
    arr = (C.c_char_p * len(files))(*[f.encode() for f in files])
    per_bin = np.zeros(len(files), dtype=np.uint64)
    rc = ref().ref_dindex_build(arr, len(files), min_gap, max_gap, hash_count, fpr, out_path.encode(), _p(per_bin, u64p))
    if rc != 0:
        raise RuntimeError(ref().ref_last_error().decode())
    return per_bin


def ref_index_query(index_path, regex, verify=True, threads=1, stdout_path=None, augment=False, dibf_path=None):
    """process_query (+ iter_disk_search) of the reference on an index file: (hit words, captured stdout).
    augment = `-a`, dibf_path = `-g`."""
    import tempfile
    words = np.zeros(1 << 14, dtype=np.uint64)
    tmp = stdout_path or tempfile.mktemp(prefix="ref_stdout_")
    if dibf_path:
        n = ref().ref_index_query_dibf(index_path.encode(), regex.encode(), 1 if augment else 0, dibf_path.encode(),
                                       1 if verify else 0, threads, _p(words, u64p), len(words), tmp.encode())
    else:
        n = ref().ref_index_query(index_path.encode(), regex.encode(), 1 if augment else 0, 1 if verify else 0, threads,
                                  _p(words, u64p), len(words), tmp.encode())
    if n < 0:
        raise RuntimeError(ref().ref_last_error().decode())
    text = ""
    if verify and os.path.exists(tmp):
        text = open(tmp).read()
        if stdout_path is None:
            os.unlink(tmp)
    return words[:n].copy(), text


def ref_scan_records_hits(pattern: str, posix_longest: bool, text: np.ndarray, rec_offsets: np.ndarray, threads: int,
                          cap: int = 1 << 20) -> np.ndarray:
    """RE2 FindAndConsume over in-memory records: (n, 3) uint32 rows (record, start, end) in print order."""
    text = np.ascontiguousarray(text, dtype=np.uint8)
    offs = np.ascontiguousarray(rec_offsets, dtype=np.uint64)
    out = np.zeros((cap, 3), dtype=np.uint32)
    n = ref().ref_re2_scan_records_hits(pattern.encode(), int(posix_longest), text.ctypes.data, _p(offs, u64p), len(offs) - 1,
                                        threads, _p(out, u32p), cap)
    if n < 0:
        raise ValueError(ref().ref_last_error().decode())
    if n > cap:
        return ref_scan_records_hits(pattern, posix_longest, text, rec_offsets, threads, cap=int(n))
    return out[:n].copy()


def ref_flat_set_script(keys, ops) -> np.ndarray:
    """Iteration order of hibf's robin_hood::unordered_flat_set<uint64_t> after ops[i] = 0: insert(keys[i]) / 1: clear()."""
    k = np.ascontiguousarray(keys, dtype=np.uint64)
    o = np.ascontiguousarray(ops, dtype=np.uint8)
    out = np.zeros(max(len(k), 1), dtype=np.uint64)
    n = ref().ref_flat_set_script(_p(k, u64p), _p(o, u8p), len(k), _p(out, u64p))
    return out[:n].copy()


def ref_rh_pop_order(keys) -> np.ndarray:
    """Pop order of a robin_hood::unordered_map<uint64_t, ...> filled like OTFCollector::push fills a rank."""
    k = np.ascontiguousarray(keys, dtype=np.uint64)
    out = np.zeros(len(k), dtype=np.uint32)
    n = ref().ref_rh_pop_order(_p(k, u64p), len(k), _p(out, u32p))
    return out[:n].copy()


def ref_index_conjunction(index_path, regex_list, threads=1):
    """`tetrex query -c`: AND of the filters of the ':'-separated queries + RE2::Set verify: (hit words, stdout)."""
    import tempfile
    words = np.zeros(1 << 14, dtype=np.uint64)
    tmp = tempfile.mktemp(prefix="ref_stdout_")
    n = ref().ref_index_conjunction(index_path.encode(), regex_list.encode(), threads, _p(words, u64p), len(words), tmp.encode())
    if n < 0:
        raise RuntimeError(ref().ref_last_error().decode())
    text = ""
    if os.path.exists(tmp):
        text = open(tmp).read()
        os.unlink(tmp)
    return words[:n].copy(), text


def guarded(fn, *args):
    """Runs fn(*args) in a forked child and returns its (pickled) result, or the string "ABORTED" if the child died:
    the reference terminates or segfaults on some inputs (empty regex after trimming, UB in its reduced k-graph
    builder, short queries on a flat index), which must not take the test process with it."""
    import pickle
    r, w = os.pipe()
    pid = os.fork()
    if pid == 0:
        os.close(r)
        try:
            import faulthandler
            faulthandler.disable()  # an expected crash of the reference must not dump a traceback into the test log
            devnull = os.open(os.devnull, os.O_WRONLY)
            os.dup2(devnull, 2)
            os.write(w, pickle.dumps(fn(*args)))
        finally:
            os._exit(0)
    os.close(w)
    data = b""
    while True:
        chunk = os.read(r, 1 << 16)
        if not chunk:
            break
        data += chunk
    os.close(r)
    _, status = os.waitpid(pid, 0)
    return pickle.loads(data) if status == 0 and data else "ABORTED"
