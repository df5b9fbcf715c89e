"""CPU: the collector's rules (tetrex_b200/host/collector.cpp: k-graph walk in the reference's pop order, update_path, push /
absorb, split, match, `-a` Gap nodes without a DGram filter) against the UNMODIFIED reference's OTFCollector::collect
(include/otf_collector.h:341-393) — without a GPU.  trxh_process_query_with_rows walks the k-graph over membership rows
the caller brings; here they come from the C oracle's IBF (oracle/tetrex_oracle.c, itself pinned to seqan's), filled with
the k-mers of a small synthetic peptide collection.  tests/golden/collector_rows.json holds what the reference's collector
returned on a seqan IBF of the same content (tests/golden/make_golden.py collector_rows, from oracle/_ref); where
oracle/_ref is present the reference is also asked live."""
import ctypes as C
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

from oracle import bindings as ob  # noqa: E402

GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "collector_rows.json")))
ROWS_FN = C.CFUNCTYPE(None, C.POINTER(C.c_uint64), C.c_uint64, C.POINTER(C.c_uint64), C.c_void_p)


@pytest.fixture(scope="module")
def host():
    h = C.CDLL(os.path.join(ROOT, "tetrex_b200", "lib", "libtrxhost.so"))
    h.trxh_last_error.restype = C.c_char_p
    h.trxh_process_query_with_rows.argtypes = [C.c_int, C.c_uint32, C.c_int, C.c_uint64, C.c_char_p, C.c_int, ROWS_FN, C.c_void_p,
                                               C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    return h


def oracle_filter(k):
    import make_golden as mg
    lib = mg.collector_rows_library(seed=GOLD["seed"], bins=GOLD["bins"], k=k)
    per_bin = [np.concatenate([ob.decompose(1, k, 0, r) for r in recs]) for recs in lib]
    bin_size = ob.compute_bitcount(max(len(v) for v in per_bin), GOLD["fpr"])
    ibf = ob.OracleIBF(len(lib), bin_size, GOLD["hash_funs"])
    for b, v in enumerate(per_bin):
        ibf.emplace(v, np.full(len(v), b, dtype=np.uint32))
    return ibf, per_bin, bin_size


def run(host, ibf, k, regex, augment):
    words = ibf.bin_words
    asked = []

    def rows(kmers, n, out, _user):
        v = np.ctypeslib.as_array(kmers, shape=(n,)).copy()
        asked.append(len(v))
        np.ctypeslib.as_array(out, shape=(n * words,))[:] = ibf.bulk_contains(v).reshape(-1)

    out = np.zeros(words, dtype=np.uint64)
    stats = np.zeros(3, dtype=np.uint64)
    rc = host.trxh_process_query_with_rows(1, k, 0, ibf.bins, regex.encode(), int(augment), ROWS_FN(rows), None,
                                           out.ctypes.data_as(C.POINTER(C.c_uint64)), stats.ctypes.data_as(C.POINTER(C.c_uint64)))
    if rc != 0:
        return host.trxh_last_error().decode(), None, asked
    return out, stats, asked


@pytest.mark.parametrize("k", [4, 6])
def test_collector_over_supplied_rows_equals_reference(host, k):
    ibf, _, _ = oracle_filter(k)
    cases = [c for c in GOLD["cases"] if c["k"] == k]
    assert len(cases) >= 150
    mask = np.array([(1 << min(64, ibf.bins - 64 * w)) - 1 for w in range(ibf.bin_words)], dtype=np.uint64)
    wrong, pruned = [], 0
    for c in cases:
        got, stats, asked = run(host, ibf, k, c["regex"], c["augment"])
        assert not isinstance(got, str), (c["regex"], got)
        assert len(asked) <= 1, "one membership batch per query"
        want = np.array(c["words"], dtype=np.uint64) & mask  # run_collection ANDs the answer into the real bins
        if not np.array_equal(got, want):
            wrong.append((c["regex"], c["augment"], [hex(int(x)) for x in got], [hex(int(x)) for x in want]))
        if asked and int(stats[0]) < asked[0]:
            pruned += 1  # the pruned walk probed fewer k-mers than the dry run listed
    assert not wrong, f"{len(wrong)} of {len(cases)} differ from the reference's collector, first: {wrong[:2]}"
    assert pruned >= 5


@pytest.mark.skipif(not ob.have_ref(), reason="oracle/_ref not built on this machine")
def test_collector_over_supplied_rows_live_against_reference(host):
    """A fresh seeded batch through the compiled reference's collector on a seqan IBF with the same content."""
    import random
    k = 5
    ibf, per_bin, bin_size = oracle_filter(k)
    ref = ob.RefIBF(ibf.bins, bin_size, GOLD["hash_funs"])
    for b, v in enumerate(per_bin):
        ref.emplace(v, np.full(len(v), b, dtype=np.uint32))
    assert np.array_equal(ref.words, ibf.words)
    import make_golden as mg
    flat = [r for recs in mg.collector_rows_library(seed=GOLD["seed"], bins=GOLD["bins"], k=k) for r in recs]
    rng = random.Random(int.from_bytes(os.urandom(4), "little"))
    mask = np.array([(1 << min(64, ibf.bins - 64 * w)) - 1 for w in range(ibf.bin_words)], dtype=np.uint64)
    done = 0
    for _ in range(120):
        src = rng.choice(flat)
        at = rng.randint(0, len(src) - 10)
        core = src[at:at + rng.randint(k + 1, 10)].decode()
        i = rng.randrange(1, len(core) - 1)
        rx = rng.choice([core[:i] + "(" + core[i] + "|" + rng.choice("ACDEFGHIKLMNPQRSTVWY") + ")" + core[i + 1:],
                         core[:i] + "." + core[i + 1:], core[:i] + core[i] + "?" + core[i:],
                         core[:i] + core[i] + "{1,3}" + core[i + 1:], "(X" + core[:2] + "|" + core[1] + ")" + core[2:],
                         core[:i] + "(" + core[i:i + 2] + ")+" + core[i + 2:]])
        for augment in (False, True):
            want = ob.guarded(lambda: ref.collect(1, k, 0, rx, augment).tolist())
            if want == "ABORTED":
                continue
            got, _, _ = run(host, ibf, k, rx, augment)
            assert not isinstance(got, str), (rx, got)
            assert np.array_equal(got, np.array(want, dtype=np.uint64) & mask), (rx, augment)
            done += 1
    assert done >= 150
