"""Multi-GPU query driver: one process per GPU, each holding one shard of an index file.

The work is done by the C++ host layer (tetrex_b200/host, libtrxhost.so: DeviceIndex, process_query,
iter_disk_search) and, underneath, the CUDA kernels; this module only wires the three points at which shards must talk:

1. pruning — `path_.none()` (include/otf_collector.h:383) is a statement about ALL bins, so after every probe step the
   per-item alive bytes are OR-ed over the shards: `dist.any_alive` (all-reduce MAX of n bytes, NCCL on the GPU box);
2. the final candidate vector — each flat-IBF shard owns a block of 64-bin words: `dist.gather_words`;
3. output — every shard verifies (K4) and formats the lines of its own bins; shards own ascending bin ranges, so the
   concatenation in rank order is the reference's -t 1 print order: `gather_lines`.

A HIBF is small and replicated: every rank computes the whole candidate vector (no collective) and only the verify text
is sharded.  Nothing here computes on the CPU: without the built libraries the import of the handles fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import dist as tdist
from ._abi import LIB_PATH

HOST_LIB_PATH = os.path.join(os.path.dirname(LIB_PATH), "libtrxhost.so")
_ALIVE_FN = C.CFUNCTYPE(None, C.POINTER(C.c_uint8), C.c_uint64, C.c_void_p)
_host = None


class HostError(RuntimeError):
    pass


def host_lib():
    global _host
    if _host is None:
        if not os.path.exists(HOST_LIB_PATH):
            raise ImportError(f"{HOST_LIB_PATH} is missing: build it with `make -C tetrex_b200/host` "
                              "(python -c 'import __graft_entry__ as g; g.build()')")
        h = C.CDLL(HOST_LIB_PATH)
        vp, u64p = C.c_void_p, C.POINTER(C.c_uint64)
        h.trxh_last_error.restype = C.c_char_p
        h.trxh_open.argtypes = [vp, C.c_char_p, C.c_uint32, C.c_uint32, C.POINTER(vp)]
        h.trxh_close.argtypes = [vp]
        h.trxh_close.restype = None
        h.trxh_info.argtypes = [vp, u64p]
        h.trxh_process_query.argtypes = [vp, C.c_char_p, _ALIVE_FN, vp, u64p, u64p]
        h.trxh_verify.argtypes = [vp, u64p, C.c_char_p, C.POINTER(C.c_void_p), u64p, u64p]
        h.trxh_process_query_batch.argtypes = [vp, C.POINTER(C.c_char_p), C.c_uint64, u64p, C.POINTER(C.c_uint8), u64p]
        h.trxh_free.argtypes = [vp]
        h.trxh_free.restype = None
        h.trxh_attach_comm.argtypes = [vp, vp]
        h.trxh_preload_text.argtypes = [vp, C.c_int]
        h.trxh_set_augment.argtypes = [vp, C.c_int]
        h.trxh_attach_dgram.argtypes = [vp, C.c_char_p]
        h.trxh_dindex_build.argtypes = [vp, C.POINTER(C.c_char_p), C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_float,
                                        C.c_char_p, u64p]
        _host = h
    return _host


def _check(rc):
    if rc != 0:
        raise HostError(host_lib().trxh_last_error().decode(errors="replace"))


def dindex_build(ctx, files, min_gap: int = 3, max_gap: int = 21, hash_count: int = 3, fpr: float = 0.05,
                 out_path: str | None = None) -> np.ndarray:
    """`tetrex dindex` (src/dGramIndex.cpp:19-38) on ctx's device; returns the gapped codes each bin contributed."""
    arr = (C.c_char_p * len(files))(*[os.fsencode(f) for f in files])
    per_bin = np.zeros(len(files), dtype=np.uint64)
    _check(host_lib().trxh_dindex_build(ctx._h, arr, len(files), min_gap, max_gap, hash_count, fpr,
                                        os.fsencode(out_path) if out_path else None, per_bin.ctypes.data_as(C.POINTER(C.c_uint64))))
    return per_bin


class ShardedIndex:
    """Shard `shard_index` of `shard_count` of an index file, resident on the device of `ctx` (a tetrex_b200.Context).

    any_alive(alive: np.ndarray[uint8]) -> np.ndarray[uint8] must return the element-wise OR over all shards; the
    default is the process group's all-reduce (`tetrex_b200.dist.any_alive`)."""

    def __init__(self, ctx, path: str, shard_index: int | None = None, shard_count: int | None = None, any_alive=None,
                 comm=None, preload_text: bool = False):
        rank, world, _ = tdist.rank_world()
        self.shard_index = rank if shard_index is None else shard_index
        self.shard_count = world if shard_count is None else shard_count
        self._any_alive = any_alive or tdist.any_alive
        self._h = C.c_void_p()
        self._ctx = ctx  # keeps the context alive
        _check(host_lib().trxh_open(ctx._h, os.fsencode(path), self.shard_index, self.shard_count, C.byref(self._h)))
        info = np.zeros(8, dtype=np.uint64)
        _check(host_lib().trxh_info(self._h, info.ctypes.data_as(C.POINTER(C.c_uint64))))
        (self.bin_count, self.bin_words, hibf, dna, self.k, self.word_begin, self.words, self.reduction) = map(int, info)
        self.is_hibf, self.is_dna = bool(hibf), bool(dna)
        self.column_sharded = (not self.is_hibf) and self.shard_count > 1
        self.comm = comm  # tetrex_b200.api.Comm: pruning flags and candidate words travel over NCCL inside the C++ layer
        if comm is not None:
            _check(host_lib().trxh_attach_comm(self._h, comm._h))
        if preload_text:
            _check(host_lib().trxh_preload_text(self._h, 1))

        def _cb(ptr, n, _user):
            a = np.ctypeslib.as_array(ptr, shape=(int(n),))
            a[:] = self._any_alive(a)

        self._cb = _ALIVE_FN(_cb)  # must outlive every call

    def set_augment(self, on: bool = True):
        """`tetrex query -a`: high-complexity stretches of the k-graph become Gap nodes before the walk."""
        _check(host_lib().trxh_set_augment(self._h, 1 if on else 0))

    def attach_dgram(self, path: str | None):
        """`tetrex query -g FILE`: the DGram index the gapped items probe (None detaches)."""
        _check(host_lib().trxh_attach_dgram(self._h, os.fsencode(path) if path else None))

    def close(self):
        if self._h:
            host_lib().trxh_close(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def local_candidates(self, regex: str):
        """process_query on this shard: (bin_words uint64 words with only the shard's words filled in when the index is
        column-sharded, {probes, ranks, items})."""
        words = np.zeros(self.bin_words, dtype=np.uint64)
        stats = np.zeros(3, dtype=np.uint64)
        cb = self._cb if (self.column_sharded and self.comm is None) else C.cast(None, _ALIVE_FN)
        _check(host_lib().trxh_process_query(self._h, regex.encode(), cb, None, words.ctypes.data_as(C.POINTER(C.c_uint64)),
                                             stats.ctypes.data_as(C.POINTER(C.c_uint64))))
        return words, dict(zip(("probes", "ranks", "items"), map(int, stats)))

    def local_candidates_batch(self, regexes):
        """process_query for a list of patterns with one membership batch for all their k-mers (`tetrex query -f`):
        (n x bin_words words, n status bytes — 0 where the pattern was refused —, n x 3 stats)."""
        n = len(regexes)
        words = np.zeros((n, self.bin_words), dtype=np.uint64)
        status = np.zeros(n, dtype=np.uint8)
        stats = np.zeros((n, 3), dtype=np.uint64)
        arr = (C.c_char_p * max(n, 1))(*[r.encode() for r in regexes])
        _check(host_lib().trxh_process_query_batch(self._h, arr, n, words.ctypes.data_as(C.POINTER(C.c_uint64)),
                                                   status.ctypes.data_as(C.POINTER(C.c_uint8)),
                                                   stats.ctypes.data_as(C.POINTER(C.c_uint64))))
        return words, status, stats

    def verify_local(self, candidate_words, regex: str) -> tuple[str, int]:
        """iter_disk_search over the shard's own bins: (lines in print order, line count)."""
        cand = np.ascontiguousarray(candidate_words, dtype=np.uint64)
        assert len(cand) == self.bin_words
        text, length, lines = C.c_void_p(), C.c_uint64(), C.c_uint64()
        _check(host_lib().trxh_verify(self._h, cand.ctypes.data_as(C.POINTER(C.c_uint64)), regex.encode(), C.byref(text),
                                      C.byref(length), C.byref(lines)))
        try:
            out = C.string_at(text, length.value).decode(errors="replace")
        finally:
            host_lib().trxh_free(text)
        return out, int(lines.value)

    def query(self, regex: str, verify: bool = True, gather_words=None, gather_lines=None):
        """The whole query on all shards (collective: every rank calls it with the same regex).  Returns
        (candidate words of ALL bins, the output text on rank 0 / None elsewhere)."""
        words, _ = self.local_candidates(regex)
        if self.column_sharded and self.comm is None:
            local = words[self.word_begin:self.word_begin + self.words]
            words = (gather_words or tdist.gather_words)(local)
        if not verify:
            return words, None
        # an error on ONE rank (a moved FASTA file of its bins, …) must not leave the others waiting in the gather:
        # every rank enters the collective, the failure travels with it and is raised everywhere
        failure = None
        try:
            text, _ = self.verify_local(words, regex) if words.any() else ("", 0)
        except HostError as e:
            text, failure = "", f"shard {self.shard_index}: {e}"
        gathered = (gather_lines or gather_text)(text if failure is None else _FAILED + failure)
        if failure is not None:
            raise HostError(failure)
        if gathered is not None and _FAILED in gathered:
            raise HostError(gathered.split(_FAILED, 1)[1].split("\n", 1)[0])
        return words, gathered


_FAILED = "\x00trx-shard-failed\x00"


def gather_text(text: str):
    """Rank 0 gets the shards' outputs concatenated in rank order (= ascending bins); other ranks get None."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return text
    parts = [None] * dist.get_world_size() if dist.get_rank() == 0 else None
    dist.gather_object(text, parts, dst=0)
    return "".join(parts) if parts is not None else None
