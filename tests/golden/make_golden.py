"""Generates tests/golden/*.npz / *.json from the UNMODIFIED reference compiled into
oracle/_ref/libtetrex_ref.so (make -C oracle ref).  Run in the dev container only (needs
/root/reference to have been compiled); the vectors are committed so that the GPU box and later
sessions can check the oracle and the CUDA path without the reference.

    python tests/golden/make_golden.py
"""
import json
import os
import random
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import bindings as ob  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
AA = b"ACDEFGHIKLMNPQRSTVWY"


def rand_seq(rng, alphabet, n):
    return bytes(rng.choice(alphabet) for _ in range(n))


def golden_decompose():
    rng = random.Random(11)
    cases = []
    for mol, k, red, alpha in [(0, 3, 0, b"ACGT"), (0, 8, 0, b"ACGTN"), (0, 16, 0, b"ACGT"), (0, 20, 0, b"ACGTacgtNRY"),
                               (0, 31, 0, b"ACGT"), (0, 32, 0, b"ACGT"), (1, 6, 0, AA + b"XBZJOU*"), (1, 3, 0, AA),
                               (1, 6, 1, AA + b"XBZ"), (1, 6, 2, AA + b"XBZ"), (1, 12, 0, AA)]:
        for L in [k, k + 1, 2 * k + 3, 97, 400]:
            s = rand_seq(rng, alpha, L)
            cases.append(dict(molecule=mol, k=k, reduction=red, seq=s.decode("latin-1"),
                              values=[int(x) for x in ob.ref_decompose(mol, k, red, s)]))
    cases.append(dict(molecule=0, k=3, reduction=0, seq="ACTGACT", values=[int(x) for x in ob.ref_decompose(0, 3, 0, b"ACTGACT")]))
    cases.append(dict(molecule=1, k=3, reduction=0, seq="LMAEGx", values=[int(x) for x in ob.ref_decompose(1, 3, 0, b"LMAEGx")]))
    maps = {}
    import ctypes as C
    for red in (0, 1, 2):
        a = np.zeros(256, np.uint8)
        r = C.create_string_buffer(256)
        ob.ref().ref_peptide_maps(6, red, a.ctypes.data_as(ob.u8p), r)
        maps[str(red)] = dict(aamap=a.tolist(), redmap=list(r.raw))
    bitcount = [dict(n=n, fpr=f, bits=int(ob.ref().ref_compute_bitcount(n, f)))
                for n in [1, 17, 1000, 192500, 3662109, 10 ** 9] for f in [0.05, 0.01, 0.3, 0.001]]
    json.dump(dict(decompose=cases, maps=maps, bitcount=bitcount), open(os.path.join(OUT, "decompose.json"), "w"))


def golden_ibf():
    rng = np.random.default_rng(2024)
    d = {}
    shapes = [(5, 106, 3), (64, 1000, 2), (65, 12345, 3), (1024, 9973, 3), (300, 7777, 5), (130, 4099, 1)]
    d["shapes"] = np.array(shapes, dtype=np.uint64)
    for i, (bins, bs, h) in enumerate(shapes):
        r = ob.RefIBF(bins, bs, h)
        v = rng.integers(0, 2 ** 63, 4000, dtype=np.uint64)
        b = rng.integers(0, bins, 4000).astype(np.uint32)
        r.emplace(v, b)
        q = np.concatenate([v[:300], rng.integers(0, 2 ** 63, 300, dtype=np.uint64)])
        d[f"values_{i}"], d[f"bins_{i}"], d[f"query_{i}"] = v, b, q
        d[f"words_{i}"] = r.words
        d[f"hits_{i}"] = r.bulk_contains(q)
    np.savez_compressed(os.path.join(OUT, "ibf.npz"), **d)


def golden_hibf():
    rng = np.random.default_rng(7)
    for name, n_ub, lo, hi in [("hibf_small", 5, 10, 40), ("hibf_100", 100, 50, 3000), ("hibf_300", 300, 20, 1500)]:
        per = [rng.integers(0, 2 ** 40, int(rng.integers(lo, hi)), dtype=np.uint64) for _ in range(n_ub)]
        rh = ob.RefHIBF(per, 3, 0.05, 1)
        tb = rh.tables()
        q = np.concatenate([p[:8] for p in per] + [rng.integers(0, 2 ** 40, 1500, dtype=np.uint64)])
        tb.save(os.path.join(OUT, name + "_tables.npz"))
        np.savez_compressed(os.path.join(OUT, name + "_query.npz"), query=q, hits=rh.membership(q))


REGEX_CORPUS_AA = ["LMA(E|Q)GLYN", "AC+D", "A(C|D)*E", "[ILV][ILV]K", "G.{2,4}K", "A{2,3}", "(AB|A)BC", "C.C", "^MK", "LL$",
                   "(A|AA)(A|AA)", "K[^P]T", "N[ST]", "W(A|C|D){1,3}W", "(GA)+", "R.?K", "P{3}", "Q+E?", "(L|LM)(MA|A)"]
REGEX_CORPUS_NA = ["A(C+|G+)T", "ACGT", "(A|AC)(CG|G)", "T+A", "G.{1,3}C", "(AC)*G", "CA?T", "^AC", "GT$", "(A|C)(A|C)G", "[AG]C[CT]",
                   "A{2,4}C", "(TA|T)A+", "N", "AC|ACG|ACGT"]


def golden_regex():
    rng = random.Random(5)
    cases = []
    for pats, alpha, posix in [(REGEX_CORPUS_AA, b"ACDEGKLMNPQRSTW", True), (REGEX_CORPUS_NA, b"ACGT", False),
                               (REGEX_CORPUS_NA, b"ACGT", True), (REGEX_CORPUS_AA, b"ABCDEKLMQ", False)]:
        for pat in pats:
            for L in [0, 1, 5, 40, 300]:
                text = rand_seq(rng, alpha, L)
                try:
                    m = ob.ref_find_all(pat, posix, text)
                except ValueError:
                    continue
                cases.append(dict(pattern=pat, posix=posix, text=text.decode(), matches=m))
    json.dump(cases, open(os.path.join(OUT, "regex.json"), "w"))


if __name__ == "__main__":
    assert ob.have_ref(), "build oracle/_ref first: make -C oracle ref"
    golden_decompose()
    golden_ibf()
    golden_hibf()
    golden_regex()
    print("golden vectors written to", OUT)
