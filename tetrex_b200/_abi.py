"""ctypes binding of include/tetrex_b200.h (libtetrex_cuda.so).  No fallback: if the library is
missing it is an error to use the package (build it with ``python __graft_entry__.py build``)."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libtetrex_cuda.so")

u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)
vp = C.c_void_p
vpp = C.POINTER(C.c_void_p)


class TrxHit(C.Structure):
    _fields_ = [("bin", C.c_uint32), ("record", C.c_uint32), ("start", C.c_uint32), ("end", C.c_uint32)]


class TrxHibfDesc(C.Structure):
    _fields_ = [("n_ibfs", C.c_uint64), ("n_user_bins", C.c_uint64), ("bins", u64p), ("bin_size", u64p),
                ("hash_funs", u64p), ("words", C.POINTER(u64p)), ("next_ibf", C.POINTER(u64p)),
                ("user_bin", C.POINTER(u64p)), ("table_len", u64p)]


# name → (restype, argtypes); exactly the symbols declared in include/tetrex_b200.h
PROTOTYPES = {
    "trx_last_error": (C.c_char_p, []),
    "trx_abi_version": (C.c_int, []),
    "trx_ctx_create": (C.c_int, [C.c_int, vp, vpp]),
    "trx_ctx_destroy": (None, [vp]),
    "trx_ctx_sync": (C.c_int, [vp]),
    "trx_ctx_stream": (vp, [vp]),
    "trx_ctx_device": (C.c_int, [vp]),
    "trx_ctx_launches": (C.c_uint64, [vp]),
    "trx_dev_alloc": (C.c_int, [vp, C.c_uint64, vpp]),
    "trx_dev_free": (C.c_int, [vp, vp]),
    "trx_dev_memset": (C.c_int, [vp, vp, C.c_int, C.c_uint64]),
    "trx_copy_h2d": (C.c_int, [vp, vp, vp, C.c_uint64]),
    "trx_copy_d2h": (C.c_int, [vp, vp, vp, C.c_uint64]),
    "trx_timer_start": (C.c_int, [vp]),
    "trx_timer_stop_ms": (C.c_int, [vp, C.POINTER(C.c_float)]),
    "trx_emit_count": (C.c_uint64, [C.c_int, C.c_uint32, C.c_uint64]),
    "trx_peptide_maps": (None, [C.c_int, u8p, u8p]),
    "trx_compute_bitcount": (C.c_uint64, [C.c_uint64, C.c_float]),
    "trx_decompose_records": (C.c_int, [vp, C.c_int, C.c_uint32, C.c_int, vp, vp, C.c_uint64, vp, C.c_uint64, vp]),
    "trx_ibf_create": (C.c_int, [vp, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, vpp]),
    "trx_ibf_destroy": (None, [vp]),
    "trx_ibf_geometry": (C.c_int, [vp, u64p]),
    "trx_ibf_device_words": (vp, [vp]),
    "trx_ibf_clear": (C.c_int, [vp]),
    "trx_ibf_upload": (C.c_int, [vp, vp, C.c_uint64]),
    "trx_ibf_download": (C.c_int, [vp, vp, C.c_uint64]),
    "trx_emplace_values": (C.c_int, [vp, vp, vp, C.c_uint64]),
    "trx_emplace_values_dev": (C.c_int, [vp, vp, vp, C.c_uint64]),
    "trx_emplace_records": (C.c_int, [vp, C.c_int, C.c_uint32, C.c_int, vp, vp, vp, C.c_uint64]),
    "trx_emplace_records_dev": (C.c_int, [vp, C.c_int, C.c_uint32, C.c_int, vp, C.c_uint64, vp, vp, C.c_uint64]),
    "trx_emplace_records_slab_dev": (C.c_int, [vp, C.c_int, C.c_uint32, C.c_int, vp, C.c_uint64, vp, vp, C.c_uint64, C.c_uint64,
                                               C.c_uint64]),
    "trx_dgram_count": (C.c_int, [vp, vp, vp, vp, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint64, vp]),
    "trx_dgram_count_dev": (C.c_int, [vp, vp, C.c_uint64, vp, vp, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint64, vp]),
    "trx_dgram_emplace": (C.c_int, [vp, vp, vp, vp, C.c_uint64, C.c_uint32, C.c_uint32]),
    "trx_dgram_emplace_dev": (C.c_int, [vp, vp, C.c_uint64, vp, vp, C.c_uint64, C.c_uint32, C.c_uint32]),
    "trx_hll_sketch_records": (C.c_int, [vp, C.c_int, C.c_uint32, C.c_int, vp, vp, vp, C.c_uint64, C.c_uint64, C.c_uint32, vp]),
    "trx_hll_sketch_records_dev": (C.c_int, [vp, C.c_int, C.c_uint32, C.c_int, vp, C.c_uint64, vp, vp, C.c_uint64, C.c_uint64,
                                             C.c_uint32, vp]),
    "trx_bulk_contains": (C.c_int, [vp, vp, C.c_uint64, vp]),
    "trx_bulk_contains_dev": (C.c_int, [vp, vp, C.c_uint64, vp]),
    "trx_bulk_contains_noalloc": (C.c_int, [vp, vp, C.c_uint64, vp]),
    "trx_bulk_contains_noalloc_max_kmers": (C.c_uint64, [vp]),
    "trx_paths_create": (C.c_int, [vp, C.c_uint64, C.c_uint64, vpp]),
    "trx_paths_destroy": (None, [vp]),
    "trx_paths_device_words": (vp, [vp]),
    "trx_paths_fill": (C.c_int, [vp, C.c_uint64, C.c_uint64, C.c_uint64]),
    "trx_paths_step": (C.c_int, [vp, vp, vp, vp, C.c_uint64, vp]),
    "trx_paths_step_dev": (C.c_int, [vp, vp, vp, vp, C.c_uint64, vp]),
    "trx_paths_step_sharded": (C.c_int, [vp, vp, vp, vp, vp, C.c_uint64, vp]),
    "trx_paths_and_hits": (C.c_int, [vp, vp, vp, vp, C.c_uint64, vp]),
    "trx_paths_or": (C.c_int, [vp, vp, vp, C.c_uint64]),
    "trx_paths_read": (C.c_int, [vp, C.c_uint64, C.c_uint64, vp]),
    "trx_hibf_upload": (C.c_int, [vp, C.POINTER(TrxHibfDesc), vpp]),
    "trx_hibf_destroy": (None, [vp]),
    "trx_hibf_membership": (C.c_int, [vp, vp, C.c_uint64, vp, u64p]),
    "trx_hibf_membership_dev": (C.c_int, [vp, vp, C.c_uint64, vp, u64p]),
    "trx_hibf_membership_ids": (C.c_int, [vp, vp, C.c_uint64, vp, vp, C.c_uint32, C.c_uint64, u64p, u64p]),
    "trx_hibf_user_bin_words": (C.c_uint64, [vp]),
    "trx_text_upload": (C.c_int, [vp, vp, vp, vp, C.c_uint64, C.c_uint64, vpp]),
    "trx_text_destroy": (None, [vp]),
    "trx_regex_compile": (C.c_int, [C.c_char_p, C.c_int, vpp]),
    "trx_regex_destroy": (None, [vp]),
    "trx_regex_positions": (C.c_int, [vp]),
    "trx_regex_dfa_info": (C.c_int, [vp, vp]),
    "trx_regex_dfa_starts": (C.c_uint32, [vp, vp, C.c_uint64, vp]),
    "trx_regex_dfa_match_end": (C.c_int64, [vp, vp, C.c_uint64, C.c_uint64]),
    "trx_regex_first_match_end": (C.c_int64, [vp, vp, C.c_uint64]),
    "trx_regex_scan": (C.c_int, [vp, vp, vp, C.c_int, vp, vp, C.c_uint64, u64p, u64p]),
    "trx_regex_scan_begin": (C.c_int, [vp, vp, vp, C.c_int, vp, C.c_uint64, C.POINTER(C.c_int)]),
    "trx_regex_scan_end": (C.c_int, [vp, C.c_int, vp, u64p, u64p]),
    "trx_regex_scan_multi": (C.c_int, [vp, vpp, C.c_uint32, vp, C.c_uint64, vp, vp, C.c_uint64, u64p, vp, vp]),
    "trx_regex_filter_plan": (C.c_int, [vp, vp, C.c_int, vp, vp]),
    "trx_regex_window_plan": (C.c_int, [vp, vp, vp, vp, vp]),
    "trx_comm_unique_id": (C.c_int, [vp]),
    "trx_comm_create": (C.c_int, [vp, C.c_int, C.c_int, vp, vpp]),
    "trx_comm_create_all": (C.c_int, [vpp, C.c_int, vpp]),
    "trx_comm_destroy": (None, [vp]),
    "trx_comm_rank": (C.c_int, [vp]),
    "trx_comm_world": (C.c_int, [vp]),
    "trx_comm_collectives": (C.c_uint64, [vp]),
    "trx_comm_version": (C.c_int, []),
    "trx_comm_any_alive_dev": (C.c_int, [vp, vp, C.c_uint64]),
    "trx_comm_allgather_dev": (C.c_int, [vp, vp, vp, C.c_uint64]),
    "trx_comm_bcast_dev": (C.c_int, [vp, vp, C.c_uint64, C.c_int]),
    "trx_comm_allgather": (C.c_int, [vp, vp, vp, C.c_uint64]),
}

_lib = None


class TetrexError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"[TRX {code}] {message}")
        self.code = code


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: tetrex_b200 has no CPU or PyTorch fallback; build the CUDA "
                              "library first (python -c 'import __graft_entry__ as g; g.build()')")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(handle, name)  # AttributeError = header and library disagree
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int):
    if rc != 0:
        raise TetrexError(rc, lib().trx_last_error().decode(errors="replace"))
