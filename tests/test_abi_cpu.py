"""CPU: the C-ABI library loads, exports every symbol include/tetrex_b200.h declares, its host-side
helpers agree with the oracle, and it refuses to run without a GPU (no CPU fallback)."""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest

import tetrex_b200 as t
from oracle import bindings as ob
from tetrex_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "tetrex_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(trx_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    names = declared_symbols()
    assert len(names) >= 45
    handle = C.CDLL(_abi.LIB_PATH)
    for n in names:
        assert hasattr(handle, n), f"{n} declared in include/tetrex_b200.h but not exported"
    # and the Python prototypes cover exactly the declared set
    assert sorted(_abi.PROTOTYPES) == names


def test_host_facade_exports_every_declared_symbol():
    """include/tetrex_host.h ↔ tetrex_b200/lib/libtrxhost.so (the C facade of the C++ host layer)."""
    from tetrex_b200 import sharded
    hdr = open(os.path.join(ROOT, "include", "tetrex_host.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(trxh_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 8
    handle = sharded.host_lib()
    for n in names:
        assert hasattr(handle, n), f"{n} declared in include/tetrex_host.h but not exported by libtrxhost.so"


def test_abi_version():
    assert t.lib().trx_abi_version() == 1


def test_host_helpers_match_oracle(golden_dir):
    g = json.load(open(os.path.join(golden_dir, "decompose.json")))
    for red in (0, 1, 2):
        a, r = t.peptide_maps(red)
        assert a.tolist() == g["maps"][str(red)]["aamap"]
        assert r.tolist() == g["maps"][str(red)]["redmap"]
    for c in g["bitcount"]:
        assert t.compute_bitcount(c["n"], c["fpr"]) == c["bits"]
    for mol, k, L in [(0, 3, 16), (0, 20, 19), (0, 20, 20), (1, 6, 5), (1, 6, 6), (1, 6, 350)]:
        assert t.emit_count(mol, k, L) == len(ob.decompose(mol, k, 0, b"A" * L))


def test_regex_compile_errors_and_positions():
    assert t.Regex("LMA(E|Q)GLYN", True).positions == 9
    assert t.Regex("G.{2,4}K", True).positions == 6
    assert t.Regex("A(C+|G+)T", False).positions == 4
    # up to 250 positions (one, two or four 64-bit words of automaton state)
    assert t.Regex("x" * 129, True).positions == 129
    assert t.Regex("M.{30,80}W.{20,60}K", True).positions == 143
    assert t.Regex("G.{249}", False).positions == 250
    for bad in ["A*", "(A|)", "", "A(", "A)", "[AB", "A{3,2}", "A**", "\\d+", "A{2", "(?i)A", "x" * 251, "A{251}", "(AC){126}", "G.{10,250}"]:
        with pytest.raises(t.TetrexError) as e:
            t.Regex(bad, True)
        assert e.value.code == -6, bad


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="a GPU is present")
def test_no_cpu_fallback_without_gpu():
    with pytest.raises(t.TetrexError) as e:
        t.Context(0)
    assert e.value.code == -3  # TRX_ENODEV
    assert "no CPU fallback" in str(e.value)
