"""Thin object layer over the C ABI — numpy in, numpy out.  Names follow the reference's domain
(IBF, HIBF, bins, k-mers, paths, records).  Every method is one C-ABI call; nothing is computed in
Python and nothing falls back to the CPU.

Host-buffer methods take numpy arrays; ``*_dev`` methods take raw device pointers (ints), e.g.
``tensor.data_ptr()`` of a torch CUDA tensor, and only enqueue work on the context's stream.
"""
from __future__ import annotations

import atexit
import ctypes as C
import sys
import time
import weakref

import numpy as np

from ._abi import TrxHibfDesc, TrxHit, check, lib, u64p

MOL_NA, MOL_AA = 0, 1
RED_BASE, RED_MURPHY, RED_LI = 0, 1, 2


def _np(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


def _ptr(a):
    return a.ctypes.data if a is not None else None


_live_contexts = weakref.WeakSet()


@atexit.register
def _close_contexts():
    # device handles go while the CUDA runtime is still loaded: module globals die in no particular order at interpreter
    # exit, and a trx_*_destroy after the runtime's own teardown is a crash
    for ctx in list(_live_contexts):
        try:
            ctx.close()
        except Exception:
            pass


class Context:
    """One CUDA device + one stream (trx_ctx_t).  stream = raw cudaStream_t or None."""

    def __init__(self, device: int = 0, stream: int | None = None):
        self._h = C.c_void_p()
        check(lib().trx_ctx_create(device, stream, C.byref(self._h)))
        self.device = device
        self._children = weakref.WeakSet()  # handles that must be destroyed before the context
        _live_contexts.add(self)

    def close(self):
        if self._h:
            for child in list(self._children):
                child.close()
            lib().trx_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        check(lib().trx_ctx_sync(self._h))

    @property
    def launches(self) -> int:
        return lib().trx_ctx_launches(self._h)

    def timer_start(self):
        check(lib().trx_timer_start(self._h))

    def timer_stop_ms(self) -> float:
        ms = C.c_float()
        check(lib().trx_timer_stop_ms(self._h, C.byref(ms)))
        return ms.value

    def decompose_records(self, molecule, k, reduction, residues: bytes | np.ndarray, rec_offsets):
        """decompose_record for every record; returns (values, emit_offsets)."""
        res = np.frombuffer(residues, dtype=np.uint8) if isinstance(residues, (bytes, bytearray)) else _np(residues, np.uint8)
        offs = _np(rec_offsets, np.uint64)
        n = len(offs) - 1
        cap = int(sum(lib().trx_emit_count(molecule, k, int(offs[i + 1] - offs[i])) for i in range(n)))
        out = np.zeros(max(cap, 1), dtype=np.uint64)
        out_offs = np.zeros(n + 1, dtype=np.uint64)
        check(lib().trx_decompose_records(self._h, molecule, k, reduction, _ptr(res), _ptr(offs), n, _ptr(out), cap,
                                          _ptr(out_offs)))
        return out[:cap], out_offs


class Comm:
    """NCCL communicator of the CUDA library for one context = one GPU (trx_comm_t)."""

    def __init__(self, ctx: Context, rank: int, world: int, unique_id: bytes):
        assert len(unique_id) == 128
        self.ctx, self.rank, self.world = ctx, rank, world
        self._h = C.c_void_p()
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        check(lib().trx_comm_create(ctx._h, rank, world, C.addressof(buf), C.byref(self._h)))
        ctx._children.add(self)

    @staticmethod
    def unique_id() -> bytes:
        buf = (C.c_uint8 * 128)()
        check(lib().trx_comm_unique_id(C.addressof(buf)))
        return bytes(buf)

    @classmethod
    def from_torch_dist(cls, ctx: Context):
        """One rank per GPU under torchrun: rank 0 draws the id, torch.distributed ships the 128 bytes."""
        import torch.distributed as dist
        rank, world = dist.get_rank(), dist.get_world_size()
        box = [cls.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        return cls(ctx, rank, world, box[0])

    @property
    def collectives(self) -> int:
        return lib().trx_comm_collectives(self._h)

    def any_alive_dev(self, d_alive: int, n: int):
        check(lib().trx_comm_any_alive_dev(self._h, d_alive, n))

    def allgather_dev(self, d_send: int, d_recv: int, bytes_per_rank: int):
        check(lib().trx_comm_allgather_dev(self._h, d_send, d_recv, bytes_per_rank))

    def bcast_dev(self, d_buf: int, nbytes: int, root: int = 0):
        check(lib().trx_comm_bcast_dev(self._h, d_buf, nbytes, root))

    def close(self):
        if self._h:
            lib().trx_comm_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def emit_count(molecule, k, length) -> int:
    return lib().trx_emit_count(molecule, k, length)


def peptide_maps(reduction):
    a = np.zeros(256, dtype=np.uint8)
    r = np.zeros(256, dtype=np.uint8)
    lib().trx_peptide_maps(reduction, a.ctypes.data_as(C.POINTER(C.c_uint8)), r.ctypes.data_as(C.POINTER(C.c_uint8)))
    return a, r


def compute_bitcount(n, fpr) -> int:
    return lib().trx_compute_bitcount(n, fpr)


def dgram_count(ctx, residues, rec_offsets, rec_bin, min_gap: int, max_gap: int, n_bins: int) -> np.ndarray:
    """dgram_buffer_[b].size() after DGramIndex::populate (include/dGramIndex.h:103-111,194-243), counted on the device."""
    res = np.frombuffer(residues, dtype=np.uint8) if isinstance(residues, (bytes, bytearray)) else _np(residues, np.uint8)
    offs = _np(rec_offsets, np.uint64)
    rb = _np(rec_bin, np.uint32)
    out = np.zeros(n_bins, dtype=np.uint64)
    check(lib().trx_dgram_count(ctx._h, _ptr(res), _ptr(offs), _ptr(rb), len(rb), min_gap, max_gap, n_bins, _ptr(out)))
    return out


def hll_sketch_records(ctx, molecule, k, reduction, residues, rec_offsets, rec_bin, n_bins: int, sketch_bits: int = 12) -> np.ndarray:
    """compute_sketches (lib/hibf/src/sketch/compute_sketches.cpp:27-43) on the device: (n_bins, 2**sketch_bits) uint8."""
    res = np.frombuffer(residues, dtype=np.uint8) if isinstance(residues, (bytes, bytearray)) else _np(residues, np.uint8)
    offs = _np(rec_offsets, np.uint64)
    rb = _np(rec_bin, np.uint32)
    out = np.zeros((n_bins, 1 << sketch_bits), dtype=np.uint8)
    check(lib().trx_hll_sketch_records(ctx._h, molecule, k, reduction, _ptr(res), _ptr(offs), _ptr(rb), len(rb), n_bins, sketch_bits,
                                       _ptr(out)))
    return out


class IBF:
    """Flat interleaved Bloom filter resident in HBM (seqan::hibf::interleaved_bloom_filter)."""

    def __init__(self, ctx: Context, bins: int, bin_size: int, hash_funs: int, shard_index: int = 0, shard_count: int = 1):
        self.ctx = ctx
        self._h = C.c_void_p()
        check(lib().trx_ibf_create(ctx._h, bins, bin_size, hash_funs, shard_index, shard_count, C.byref(self._h)))
        g = np.zeros(10, dtype=np.uint64)
        check(lib().trx_ibf_geometry(self._h, g.ctypes.data_as(u64p)))
        (self.bins, self.technical_bins, self.bin_size, self.hash_shift, self.bin_words, self.hash_funs, self.n_words,
         self.shard_word_begin, self.shard_words, self.shard_n_words) = [int(x) for x in g]
        ctx._children.add(self)

    def close(self):
        if self._h:
            lib().trx_ibf_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def device_words(self) -> int:
        return lib().trx_ibf_device_words(self._h)

    def clear(self):
        check(lib().trx_ibf_clear(self._h))

    def upload(self, words):
        w = _np(words, np.uint64)
        check(lib().trx_ibf_upload(self._h, _ptr(w), len(w)))

    def download(self) -> np.ndarray:
        w = np.zeros(self.n_words, dtype=np.uint64)
        check(lib().trx_ibf_download(self._h, _ptr(w), len(w)))
        return w

    def emplace_values(self, values, bin_ids):
        v = _np(values, np.uint64)
        b = _np(bin_ids, np.uint32)
        check(lib().trx_emplace_values(self._h, _ptr(v), _ptr(b), len(v)))

    def emplace_values_dev(self, d_values: int, d_bins: int, n: int):
        check(lib().trx_emplace_values_dev(self._h, d_values, d_bins, n))

    def emplace_records(self, molecule, k, reduction, residues, rec_offsets, rec_bin):
        res = np.frombuffer(residues, dtype=np.uint8) if isinstance(residues, (bytes, bytearray)) else _np(residues, np.uint8)
        offs = _np(rec_offsets, np.uint64)
        rb = _np(rec_bin, np.uint32)
        check(lib().trx_emplace_records(self._h, molecule, k, reduction, _ptr(res), _ptr(offs), _ptr(rb), len(rb)))

    def emplace_records_dev(self, molecule, k, reduction, d_residues: int, n_residues: int, d_rec_offsets: int,
                            d_rec_bin: int, n_records: int):
        check(lib().trx_emplace_records_dev(self._h, molecule, k, reduction, d_residues, n_residues, d_rec_offsets,
                                            d_rec_bin, n_records))

    def emplace_records_slab_dev(self, molecule, k, reduction, d_residues: int, n_residues: int, d_rec_offsets: int,
                                 d_rec_bin: int, n_records: int, first_bin: int, n_bins: int):
        check(lib().trx_emplace_records_slab_dev(self._h, molecule, k, reduction, d_residues, n_residues, d_rec_offsets,
                                                 d_rec_bin, n_records, first_bin, n_bins))

    def dgram_emplace(self, residues, rec_offsets, rec_bin, min_gap: int, max_gap: int):
        """init_ibf of the DGram index (include/dGramIndex.h:144-154): every gapped 3+3-mer code of every record."""
        res = np.frombuffer(residues, dtype=np.uint8) if isinstance(residues, (bytes, bytearray)) else _np(residues, np.uint8)
        offs = _np(rec_offsets, np.uint64)
        rb = _np(rec_bin, np.uint32)
        check(lib().trx_dgram_emplace(self._h, _ptr(res), _ptr(offs), _ptr(rb), len(rb), min_gap, max_gap))

    def bulk_contains(self, kmers, out: np.ndarray | None = None) -> np.ndarray:
        v = _np(kmers, np.uint64)
        if out is None:
            out = np.zeros((len(v), self.shard_words), dtype=np.uint64)
        check(lib().trx_bulk_contains(self._h, _ptr(v), len(v), _ptr(out)))
        return out

    def bulk_contains_noalloc(self, kmers, out: np.ndarray | None = None) -> np.ndarray:
        """bulk_contains through the library's fixed device scratch: no allocation on the device side at all."""
        v = _np(kmers, np.uint64)
        if out is None:
            out = np.zeros((len(v), self.shard_words), dtype=np.uint64)
        check(lib().trx_bulk_contains_noalloc(self._h, _ptr(v), len(v), _ptr(out)))
        return out

    @property
    def noalloc_max_kmers(self) -> int:
        return lib().trx_bulk_contains_noalloc_max_kmers(self._h)

    def bulk_contains_dev(self, d_kmers: int, n: int, d_out: int):
        check(lib().trx_bulk_contains_dev(self._h, d_kmers, n, d_out))


class Paths:
    """Pool of path bitvectors (CollectorsItem::path_) in HBM."""

    def __init__(self, ctx: Context, words_per_slot: int, capacity: int):
        self.ctx = ctx
        self.words_per_slot, self.capacity = words_per_slot, capacity
        self._h = C.c_void_p()
        check(lib().trx_paths_create(ctx._h, words_per_slot, capacity, C.byref(self._h)))
        ctx._children.add(self)

    def close(self):
        if self._h:
            lib().trx_paths_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def fill(self, first, n, bins):
        check(lib().trx_paths_fill(self._h, first, n, bins))

    def step(self, ibf: IBF, kmers, slots) -> np.ndarray:
        v = _np(kmers, np.uint64)
        s = _np(slots, np.uint32)
        alive = np.zeros(len(v), dtype=np.uint8)
        check(lib().trx_paths_step(ibf._h, self._h, _ptr(v), _ptr(s), len(v), _ptr(alive)))
        return alive

    def step_dev(self, ibf: IBF, d_kmers: int, d_slots: int, n: int, d_alive: int):
        check(lib().trx_paths_step_dev(ibf._h, self._h, d_kmers, d_slots, n, d_alive))

    def and_hits(self, d_hits: int, hit_index, slots) -> np.ndarray:
        hi = _np(hit_index, np.uint32)
        s = _np(slots, np.uint32)
        alive = np.zeros(len(s), dtype=np.uint8)
        check(lib().trx_paths_and_hits(self._h, d_hits, _ptr(hi), _ptr(s), len(s), _ptr(alive)))
        return alive

    def or_(self, dst, src):
        d = _np(dst, np.uint32)
        s = _np(src, np.uint32)
        check(lib().trx_paths_or(self._h, _ptr(d), _ptr(s), len(d)))

    def read(self, first, n) -> np.ndarray:
        out = np.zeros((n, self.words_per_slot), dtype=np.uint64)
        check(lib().trx_paths_read(self._h, first, n, _ptr(out)))
        return out


class HIBF:
    """Hierarchical IBF resident in HBM, built from flattened tables (oracle.bindings.HIBFTables-like)."""

    def __init__(self, ctx: Context, n_user_bins, bins, bin_size, hash_funs, words, next_ibf, user_bin):
        self.ctx = ctx
        n = len(words)
        self._keep = dict(bins=_np(bins, np.uint64), bin_size=_np(bin_size, np.uint64), hash_funs=_np(hash_funs, np.uint64),
                          words=[_np(w, np.uint64) for w in words], next_ibf=[_np(w, np.uint64) for w in next_ibf],
                          user_bin=[_np(w, np.uint64) for w in user_bin])
        k = self._keep
        table_len = np.array([len(t) for t in k["user_bin"]], dtype=np.uint64)
        mk = lambda arrs: (u64p * n)(*[a.ctypes.data_as(u64p) for a in arrs])
        desc = TrxHibfDesc(n, int(n_user_bins), k["bins"].ctypes.data_as(u64p), k["bin_size"].ctypes.data_as(u64p),
                           k["hash_funs"].ctypes.data_as(u64p), mk(k["words"]), mk(k["next_ibf"]), mk(k["user_bin"]),
                           table_len.ctypes.data_as(u64p))
        self._h = C.c_void_p()
        check(lib().trx_hibf_upload(ctx._h, C.byref(desc), C.byref(self._h)))
        self.n_user_bins = int(n_user_bins)
        self.out_words = int(lib().trx_hibf_user_bin_words(self._h))
        self._keep = None  # the device copy is complete; host arrays are not referenced any more
        ctx._children.add(self)

    @classmethod
    def from_tables(cls, ctx, t):
        return cls(ctx, t.n_user_bins, t.bins, t.bin_size, t.hash_funs, t.words, t.next_ibf, t.user_bin)

    def close(self):
        if self._h:
            lib().trx_hibf_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def membership(self, kmers, out: np.ndarray | None = None, stats: bool = False):
        v = _np(kmers, np.uint64)
        if out is None:
            out = np.zeros((len(v), self.out_words), dtype=np.uint64)
        st = np.zeros(2, dtype=np.uint64)
        check(lib().trx_hibf_membership(self._h, _ptr(v), len(v), _ptr(out), st.ctypes.data_as(u64p)))
        return (out, st) if stats else out

    def membership_ids(self, kmers, id_bytes: int = 2, cap: int | None = None, counts: np.ndarray | None = None,
                       ids: np.ndarray | None = None):
        """membership_for's own result form: (counts per k-mer, concatenated ascending user-bin ids).  counts / ids may be
        caller-owned (e.g. pinned) buffers; returns (counts, ids[:n_ids])."""
        v = _np(kmers, np.uint64)
        if counts is None:
            counts = np.zeros(len(v), dtype=np.uint32)
        dt = np.uint16 if id_bytes == 2 else np.uint32
        if ids is None:
            ids = np.zeros(cap if cap is not None else max(len(v) * 64, 1), dtype=dt)
        assert ids.dtype == dt and counts.dtype == np.uint32 and len(counts) >= len(v)
        n_ids = C.c_uint64()
        check(lib().trx_hibf_membership_ids(self._h, _ptr(v), len(v), _ptr(counts), _ptr(ids), id_bytes, len(ids), C.byref(n_ids), None))
        return counts[:len(v)], ids[:n_ids.value]

    def membership_dev(self, d_kmers: int, n: int, d_out: int):
        st = np.zeros(2, dtype=np.uint64)
        check(lib().trx_hibf_membership_dev(self._h, d_kmers, n, d_out, st.ctypes.data_as(u64p)))
        return st


class Regex:
    def __init__(self, pattern: str, posix_longest: bool):
        self._h = C.c_void_p()
        check(lib().trx_regex_compile(pattern.encode("latin-1"), int(posix_longest), C.byref(self._h)))
        self.pattern = pattern
        self.positions = lib().trx_regex_positions(self._h)

    def window_plan(self, byte_freq):
        """The class window the multi-pattern first stage would use (host only), or None."""
        f = _np(byte_freq, np.float64)
        assert len(f) == 256
        out = np.zeros(65, dtype=np.int32)
        cls = np.zeros(48, dtype=np.uint32)
        p = np.zeros(2, dtype=np.float64)
        check(lib().trx_regex_window_plan(self._h, _ptr(f), _ptr(out), _ptr(cls), _ptr(p)))
        if not out[0]:
            return None

        def members(words):
            return bytes(32 * w + b for w in range(8) for b in range(32) if (int(words[w]) >> b) & 1)

        sets = [members(cls[8 * k:8 * k + 8]) for k in range(int(out[1]))]
        confirm = [(members(out[7 + 10 * c:15 + 10 * c].view(np.uint32)), int(out[5 + 10 * c]), int(out[6 + 10 * c]))
                   for c in range(int(out[4]))]
        return {"len": int(out[1]), "off_lo": int(out[2]), "off_hi": int(out[3]), "classes": sets, "confirm": confirm,
                "pass": float(p[0]), "confirmed": float(p[1])}

    def filter_plan(self, byte_freq, reverse_strand=False, map256=None) -> dict:
        """The phase-1 prefilter a scan would use on a text with these relative byte frequencies (host only)."""
        f = _np(byte_freq, np.float64)
        assert len(f) == 256
        m = _np(map256, np.uint8) if map256 is not None else None
        out = np.zeros(136, dtype=np.int32)
        check(lib().trx_regex_filter_plan(self._h, _ptr(f), int(reverse_strand), _ptr(m) if m is not None else None, _ptr(out)))
        kind = int(out[0])
        if kind == 1:
            n = int(out[1] + out[2])
            return {"kind": "window", "start_delta": int(out[3]),
                    "literals": [(int(out[8 + j]), int(out[16 + j])) for j in range(n)]}  # (window byte, literal)
        if kind == 2:
            atoms = {}
            for i in range(int(out[2])):
                a = atoms.setdefault(int(out[56 + i]), [[], int(out[24 + i]), int(out[40 + i])])
                a[0].append(int(out[8 + i]))
            confirm = []
            for c in range(int(out[6])):
                words = out[74 + 10 * c:82 + 10 * c].view(np.uint32)
                byts = tuple(32 * w + b for w in range(8) for b in range(32) if (int(words[w]) >> b) & 1)
                confirm.append((byts, int(out[72 + 10 * c]), int(out[73 + 10 * c])))
            return {"kind": "atoms", "n_lit": int(out[1]), "start_lo": int(out[3]), "start_hi": int(out[4]),
                    "atoms": [(tuple(sorted(b)), lo, hi) for b, lo, hi in atoms.values()],  # (bytes of the class, lo, hi)
                    "confirm": confirm}  # second stage, same triples, offsets signed
        return {"kind": "none"}

    def close(self):
        if self._h:
            lib().trx_regex_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Text:
    """Residue store in HBM: records grouped by ascending bin."""

    def __init__(self, ctx: Context, residues, rec_offsets, rec_bin, n_bins: int):
        self.ctx = ctx
        res = np.frombuffer(residues, dtype=np.uint8) if isinstance(residues, (bytes, bytearray)) else _np(residues, np.uint8)
        self.rec_offsets = _np(rec_offsets, np.uint64)
        self.rec_bin = _np(rec_bin, np.uint32)
        self.n_bins = n_bins
        self._h = C.c_void_p()
        check(lib().trx_text_upload(ctx._h, _ptr(res), _ptr(self.rec_offsets), _ptr(self.rec_bin), len(self.rec_bin),
                                    n_bins, C.byref(self._h)))
        ctx._children.add(self)

    def close(self):
        if self._h:
            lib().trx_text_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    SCAN_SLOTS = 4  # TRX_SCAN_SLOTS

    def _candidates(self, candidate_bins):
        words = (self.n_bins + 63) // 64
        if candidate_bins is None:
            if getattr(self, "_all_bins", None) is None:
                cand = np.full(words, np.uint64(2 ** 64 - 1), dtype=np.uint64)
                if self.n_bins % 64:
                    cand[-1] = np.uint64((1 << (self.n_bins % 64)) - 1)
                self._all_bins = cand
            return self._all_bins
        cand = _np(candidate_bins, np.uint64)
        assert len(cand) == words
        return cand

    def scan_begin(self, regex: Regex, candidate_bins=None, reverse_strand=False, map256=None, cap=1 << 16) -> tuple:
        """First half of scan(): returns a ticket at once; up to SCAN_SLOTS scans may be in flight."""
        cand = self._candidates(candidate_bins)
        m = _np(map256, np.uint8) if map256 is not None else None
        ticket = C.c_int(-1)
        check(lib().trx_regex_scan_begin(self._h, regex._h, _ptr(cand), int(reverse_strand), _ptr(m) if m is not None else None,
                                         cap, C.byref(ticket)))
        return ticket.value, cap

    def scan_end(self, ticket: tuple) -> np.ndarray:
        """Second half: waits for that scan; (n,4) uint32 rows (bin, record, start, end) in reference print order."""
        tk, cap = ticket
        bufs = self.__dict__.setdefault("_slot_bufs", {})
        if tk not in bufs or len(bufs[tk]) < cap:
            bufs[tk] = (TrxHit * cap)()
        n = C.c_uint64()
        st = np.zeros(2, dtype=np.uint64)
        check(lib().trx_regex_scan_end(self._h, tk, C.addressof(bufs[tk]), C.byref(n), st.ctypes.data_as(u64p)))
        self.last_filter_literals = int(st[1])
        return np.frombuffer(bufs[tk], dtype=np.uint32, count=4 * n.value).reshape(-1, 4).copy()

    def scan_many(self, regexes, candidate_bins=None, reverse_strand=False, map256=None, cap=1 << 16, depth=3):
        """A batch of regexes over the same candidate bins, `depth` scans in flight; returns one hit array per regex."""
        assert 1 <= depth <= self.SCAN_SLOTS
        out, pending = [], []
        try:
            for rx in regexes:
                if len(pending) == depth:
                    out.append(self.scan_end(pending.pop(0)))
                pending.append(self.scan_begin(rx, candidate_bins, reverse_strand, map256, cap))
            while pending:
                out.append(self.scan_end(pending.pop(0)))
        finally:
            for tk in pending:  # an error above must not leave slots taken
                try:
                    self.scan_end(tk)
                except Exception:
                    pass
        return out

    def scan_multi(self, regexes, candidate_bins=None, cap=1 << 20, stats=False):
        """A batch of patterns in shared passes over the text (trx_regex_scan_multi).  candidate_bins: None (all bins),
        one bitmap for all patterns, or an (n_patterns, words) array.  Returns one (n,4) hit array per regex."""
        n = len(regexes)
        words = (self.n_bins + 63) // 64
        cand = self._candidates(None) if candidate_bins is None else _np(candidate_bins, np.uint64)
        stride = 0 if cand.ndim == 1 else words
        assert cand.shape[-1] == words and (cand.ndim == 1 or cand.shape[0] == n)
        handles = (C.c_void_p * max(n, 1))(*[rx._h for rx in regexes])
        # arrays of this call's own (pages are touched only where hits land): the per-pattern results are views into them
        # — the arrays of the previous call again if nobody holds a result of it any more (warm pages)
        prev = getattr(self, "_multi_bufs", None)
        if prev is not None and prev[0].shape[0] == cap and sys.getrefcount(prev[0]) == 2:
            hit_buf, pat_buf = prev
        else:
            hit_buf, pat_buf = np.empty((cap, 4), dtype=np.uint32), np.empty(cap, dtype=np.uint32)
            self._multi_bufs = (hit_buf, pat_buf)
        del prev
        total = C.c_uint64()
        per = np.zeros(max(n, 1), dtype=np.uint64)
        st = np.zeros(4, dtype=np.uint64)
        t0 = time.perf_counter()
        check(lib().trx_regex_scan_multi(self._h, handles, n, _ptr(cand), stride, _ptr(hit_buf), _ptr(pat_buf),
                                         cap, C.byref(total), _ptr(per), _ptr(st)))
        self.last_multi_call_seconds = time.perf_counter() - t0
        bounds = np.concatenate([[0], np.cumsum(per[:n])]).astype(np.int64).tolist()
        out = [hit_buf[bounds[i]:bounds[i + 1]] for i in range(n)]
        self.last_multi_stats = dict(zip(("residues", "shared", "single", "passes"), map(int, st)))
        return (out, self.last_multi_stats) if stats else out

    def scan(self, regex: Regex, candidate_bins=None, reverse_strand=False, map256=None, cap=1 << 20, stats=False):
        """Returns an (n,4) uint32 array of (bin, record, start, end) in reference print order."""
        words = (self.n_bins + 63) // 64
        if candidate_bins is None:
            if getattr(self, "_all_bins", None) is None:
                cand = np.full(words, np.uint64(2 ** 64 - 1), dtype=np.uint64)
                if self.n_bins % 64:
                    cand[-1] = np.uint64((1 << (self.n_bins % 64)) - 1)
                self._all_bins = cand
            cand = self._all_bins
        else:
            cand = _np(candidate_bins, np.uint64)
            assert len(cand) == words
        m = _np(map256, np.uint8) if map256 is not None else None
        if getattr(self, "_hit_cap", 0) != cap:
            self._hit_buf, self._hit_cap = (TrxHit * cap)(), cap
        hits = self._hit_buf
        n = C.c_uint64()
        st = np.zeros(2, dtype=np.uint64)
        check(lib().trx_regex_scan(self._h, regex._h, _ptr(cand), int(reverse_strand), _ptr(m) if m is not None else None,
                                   C.addressof(hits), cap, C.byref(n), st.ctypes.data_as(u64p)))
        arr = np.frombuffer(hits, dtype=np.uint32, count=4 * n.value).reshape(-1, 4).copy()
        self.last_filter_literals = int(st[1])
        return (arr, int(st[0])) if stats else arr
