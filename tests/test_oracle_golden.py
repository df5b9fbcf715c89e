"""CPU: the C restatement (oracle/tetrex_oracle.c) against the committed golden vectors, which were
generated from the unmodified reference (tests/golden/make_golden.py).  Pins the oracle."""
import json
import os

import numpy as np
import pytest

from oracle import bindings as ob


def test_decompose_golden(golden_dir):
    g = json.load(open(os.path.join(golden_dir, "decompose.json")))
    assert len(g["decompose"]) > 50
    for c in g["decompose"]:
        got = ob.decompose(c["molecule"], c["k"], c["reduction"], c["seq"].encode("latin-1"))
        assert [int(x) for x in got] == c["values"], (c["molecule"], c["k"], c["reduction"], len(c["seq"]))


def test_survey_known_answers():
    # SURVEY.md Appendix A.1 / A.2 (executed against the reference headers during the survey)
    def dna(s):
        v = 0
        for ch in s:
            v = (v << 2) | ((ch >> 1) & 3)
        return v

    def canon(s):
        comp = {65: 84, 67: 71, 71: 67, 84: 65}
        rc = bytes(comp[c] for c in reversed(s))
        return min(dna(s), dna(rc))

    got = ob.decompose(0, 3, 0, b"ACTGACT")
    want = [canon(x) for x in [b"ACT", b"CTA", b"TAC", b"ACT", b"CTG", b"TGA", b"GAC", b"ACT"]]
    assert [int(x) for x in got] == want
    assert [int(x) for x in ob.decompose(1, 3, 0, b"LMAEGx")] == [9536, 10243, 101, 3232]


def test_maps_and_bitcount_golden(golden_dir):
    g = json.load(open(os.path.join(golden_dir, "decompose.json")))
    for red in (0, 1, 2):
        a, r = ob.peptide_maps(red)
        assert a.tolist() == g["maps"][str(red)]["aamap"]
        assert r.tolist() == g["maps"][str(red)]["redmap"]
    for c in g["bitcount"]:
        assert ob.compute_bitcount(c["n"], c["fpr"]) == c["bits"]
    # 17 emits at fpr 0.05 → 106 bits
    assert ob.compute_bitcount(17, 0.05) == 106


def test_ibf_golden(golden_dir):
    z = np.load(os.path.join(golden_dir, "ibf.npz"))
    for i, (bins, bs, h) in enumerate(z["shapes"]):
        o = ob.OracleIBF(int(bins), int(bs), int(h))
        o.emplace(z[f"values_{i}"], z[f"bins_{i}"])
        assert np.array_equal(o.words, z[f"words_{i}"])
        assert np.array_equal(o.bulk_contains(z[f"query_{i}"]), z[f"hits_{i}"])
        assert np.array_equal(o.bulk_contains(z[f"query_{i}"], threads=2), z[f"hits_{i}"])


@pytest.mark.parametrize("name", ["hibf_small", "hibf_100", "hibf_300", "hibf_c2twin"])
def test_hibf_golden(golden_dir, name):
    tb = ob.HIBFTables.load(os.path.join(golden_dir, name + "_tables.npz"))
    q = np.load(os.path.join(golden_dir, name + "_query.npz"))
    assert np.array_equal(tb.membership(q["query"]), q["hits"])
    assert np.array_equal(tb.membership(q["query"], threads=2), q["hits"])


def test_regex_python_oracle_golden(golden_dir):
    cases = json.load(open(os.path.join(golden_dir, "regex.json")))
    assert len(cases) > 200
    for c in cases:
        got = ob.find_all_python(c["pattern"], c["posix"], c["text"].encode())
        assert got == [tuple(m) for m in c["matches"]], (c["pattern"], c["posix"], c["text"])


@pytest.mark.skipif(not ob.have_ref(), reason="oracle/_ref not built here")
def test_oracle_vs_compiled_reference_random():
    """Differential check against the compiled reference itself on fresh random inputs."""
    import random
    rng = random.Random(99)
    for k in range(1, 33):
        s = bytes(rng.choice(b"ACGTNacgt") for _ in range(k + rng.randrange(0, 80)))
        assert np.array_equal(ob.decompose(0, k, 0, s), ob.ref_decompose(0, k, 0, s)), k
    for k in range(1, 13):
        for red in (0, 1, 2):
            s = bytes(rng.choice(b"ACDEFGHIKLMNPQRSTVWYXBZJOU") for _ in range(k + rng.randrange(0, 80)))
            assert np.array_equal(ob.decompose(1, k, red, s), ob.ref_decompose(1, k, red, s)), (k, red)
    nrng = np.random.default_rng(3)
    o, r = ob.OracleIBF(200, 3001, 4), ob.RefIBF(200, 3001, 4)
    v = nrng.integers(0, 2 ** 64, 3000, dtype=np.uint64)
    b = nrng.integers(0, 200, 3000).astype(np.uint32)
    o.emplace(v, b)
    r.emplace(v, b)
    assert np.array_equal(o.words, r.words)
    assert np.array_equal(o.bulk_contains(v[:100]), r.bulk_contains(v[:100]))
