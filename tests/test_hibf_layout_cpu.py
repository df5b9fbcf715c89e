"""CPU: the host restatement of seqan::hibf's layout (host/hibf_layout.cpp: HyperLogLog sketches and estimates, size sort
and Jaccard clustering, the hierarchical / simple binning DPs) and of its k-mer set's iteration order (host/rh_order.hpp)
against golden vectors written by the compiled reference (tests/golden/hibf_layout.json, make_golden.py `hibf_layout`),
and live against oracle/_ref when it is present.  No device."""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

from tetrex_b200 import sharded, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "hibf_layout.json")))
u64p, u8p = C.POINTER(C.c_uint64), C.POINTER(C.c_uint8)


def host():
    h = sharded.host_lib()
    h.trxh_hibf_layout.restype = C.c_uint64
    h.trxh_hibf_layout.argtypes = [u64p, u64p, C.c_uint64, C.c_uint32, C.c_float, u64p, C.c_uint64]
    h.trxh_flat_set_script.restype = C.c_uint64
    h.trxh_flat_set_script.argtypes = [u64p, u8p, C.c_uint64, u64p]
    return h


def our_layout(values, offs, hash_funs, fpr):
    from oracle.bindings import parse_layout_dump
    n_bins = len(offs) - 1
    out = np.zeros(64 + 24 * n_bins + 4096, dtype=np.uint64)
    n = host().trxh_hibf_layout(values.ctypes.data_as(u64p), offs.ctypes.data_as(u64p), n_bins, hash_funs, fpr,
                                out.ctypes.data_as(u64p), len(out))
    assert 0 < n <= len(out), host().trxh_last_error()
    return parse_layout_dump(out[:n], n_bins)


def as_tuples(layout):
    return dict(tmax=layout["tmax"], top=layout["top_level_max_bin_id"], max_bins=[(list(a), b) for a, b in layout["max_bins"]],
                user_bins=[(list(a), b, c, d) for a, b, c, d in layout["user_bins"]], counts=list(layout["kmer_counts"]))


@pytest.mark.parametrize("case", range(len(GOLD["layouts"])))
def test_layout_equals_reference_golden(case):
    c = GOLD["layouts"][case]
    values, offs = synth.layout_collection(c["seed"], c["n_bins"], c["mean"], c["overlap"], c["skew"])
    got = our_layout(values, offs, c["hash_funs"], c["fpr"])
    assert as_tuples(got) == as_tuples(c["layout"])


def test_layouts_cover_the_interesting_shapes():
    shapes = [c["layout"] for c in GOLD["layouts"]]
    assert any(len(l["max_bins"]) == 0 for l in shapes)                      # everything fits the top level
    assert any(len(l["max_bins"]) >= 10 for l in shapes)                     # many merged bins
    assert any(any(u[2] > 1 for u in l["user_bins"]) for l in shapes)        # split user bins
    assert any(any(len(u[0]) >= 1 for u in l["user_bins"]) for l in shapes)  # user bins below the top level


@pytest.mark.parametrize("case", range(len(GOLD["set_scripts"])))
def test_kmer_set_iteration_order_equals_reference_golden(case):
    c = GOLD["set_scripts"][case]
    keys, ops = synth.set_script(c["seed"])
    assert len(keys) == c["operations"]
    out = np.zeros(len(keys), dtype=np.uint64)
    n = host().trxh_flat_set_script(keys.ctypes.data_as(u64p), ops.ctypes.data_as(u8p), len(keys), out.ctypes.data_as(u64p))
    assert n == c["n"]
    assert [str(int(x)) for x in out[:8][:n]] == c["head"]
    assert hashlib.sha256(out[:n].tobytes()).hexdigest() == c["sha256"]


def test_layout_live_against_compiled_reference():
    from oracle import bindings as ob
    if not ob.have_ref():
        pytest.skip("oracle/_ref not built")
    for seed, n_bins, mean, ov, sk, hf, fpr in [(41, 33, 900, 1, 1, 3, 0.05), (42, 200, 250, 1, 0, 2, 0.1), (43, 90, 400, 0, 1, 5, 0.02)]:
        values, offs = synth.layout_collection(seed, n_bins, mean, ov, sk)
        assert as_tuples(our_layout(values, offs, hf, fpr)) == as_tuples(ob.ref_hibf_layout(values, offs, hf, fpr, 1)), seed
