"""GPU: the DGram index (`tetrex dindex`) and augmented queries (`tetrex query -a [-g FILE]`) against what the UNMODIFIED
reference produced (tests/golden/dgram.json, written by tests/golden/make_golden.py `dgram` from oracle/_ref):

 * the gapped 3+3-mer codes each bin contributes (DGramIndex::process_sequence, include/dGramIndex.h:194-243),
 * the index file, byte for byte (sizing by the largest bin, emplace, cereal layout; :128-154,305-309),
 * the candidate bins of process_query with OTFCollector::augment, with and without the DGram filter
   (include/otf_collector.h:216-245,290-312,395-493), and the lines iter_disk_search prints for them.
"""
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

import tetrex_b200 as t
from tetrex_b200 import sharded

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TETREX = os.path.join(ROOT, "tetrex_b200", "bin", "tetrex")
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "dgram.json")))


def read_records(path):
    recs, cur = [], None
    for line in open(os.path.join(ROOT, path), "rb"):
        if line.startswith(b">"):
            if cur is not None:
                recs.append(b"".join(cur))
            cur = []
        elif cur is not None:
            cur.append(line.rstrip(b"\n"))
    if cur is not None:
        recs.append(b"".join(cur))
    return recs


def library(files):
    residues, offs, rec_bin = [], [0], []
    for b, f in enumerate(files):
        for r in read_records(f):
            residues.append(r)
            offs.append(offs[-1] + len(r))
            rec_bin.append(b)
    return b"".join(residues), np.array(offs, dtype=np.uint64), np.array(rec_bin, dtype=np.uint32)


def codes_python(seq: bytes, lo: int, hi: int):
    """process_sequence restated (include/dGramIndex.h:194-243) — small inputs only."""
    alpha = {c: i for i, c in enumerate(b"ACDEFGHIKLMNPQRSTVWY")}
    out = []
    L = len(seq)
    if L < lo + 7:
        return out
    i = 2
    while i + lo + 3 < L:
        a = [alpha.get(seq[i - 2]), alpha.get(seq[i - 1]), alpha.get(seq[i])]
        if None not in a:
            for gap in range(lo, hi + 1):
                j = i + gap + 1
                if j + 2 >= L:
                    break
                b = [alpha.get(seq[j]), alpha.get(seq[j + 1]), alpha.get(seq[j + 2])]
                if None in b:
                    continue
                out.append(gap * 64000000 + a[0] * 3200000 + a[1] * 160000 + a[2] * 8000 + b[0] * 400 + b[1] * 20 + b[2])
        i += 1
    return out


@pytest.mark.parametrize("build", range(len(GOLD["builds"])))
def test_codes_per_bin_equal_reference(ctx, build):
    b = GOLD["builds"][build]
    residues, offs, rec_bin = library(b["files"])
    got = t.dgram_count(ctx, residues, offs, rec_bin, b["min"], b["max"], len(b["files"]))
    assert [int(x) for x in got] == b["per_bin"]


def test_emplaced_codes_are_exactly_the_reference_codes(ctx):
    """Every code of the restated process_sequence is in its bin; the filter equals one filled value by value (K1a)."""
    rng = np.random.default_rng(5)
    recs = []
    for n in [0, 5, 9, 10, 11, 12, 40, 300, 301, 77]:
        s = rng.choice(np.frombuffer(b"ACDEFGHIKLMNPQRSTVWYXBZacd*", dtype=np.uint8), n).tobytes()
        recs.append(s)
    rec_bin = np.array([i % 5 for i in range(len(recs))], dtype=np.uint32)
    offs = np.concatenate([[0], np.cumsum([len(r) for r in recs])]).astype(np.uint64)
    for lo, hi in [(3, 21), (0, 2), (1, 1), (5, 3), (250, 300)]:
        want_codes, want_bins = [], []
        for r, b in zip(recs, rec_bin):
            c = codes_python(r, lo, hi)
            want_codes += c
            want_bins += [int(b)] * len(c)
        counts = t.dgram_count(ctx, b"".join(recs), offs, rec_bin, lo, hi, 5)
        assert [int(x) for x in counts] == [want_bins.count(b) for b in range(5)], (lo, hi)
        f = t.IBF(ctx, 5, 4099, 3)
        f.dgram_emplace(b"".join(recs), offs, rec_bin, lo, hi)
        g = t.IBF(ctx, 5, 4099, 3)
        if want_codes:
            g.emplace_values(np.array(want_codes, dtype=np.uint64), np.array(want_bins, dtype=np.uint32))
        assert np.array_equal(f.download(), g.download()), (lo, hi)
        f.close()
        g.close()


def test_slab_route_sets_the_same_bits(ctx, monkeypatch):
    """Matrices far beyond L2 are filled slab by slab (64 bins at a time into a dense column, then folded in): same bits as
    the direct route, here forced on a small matrix of 200 bins with ragged and empty bins."""
    rng = np.random.default_rng(21)
    alpha = np.frombuffer(b"ACDEFGHIKLMNPQRSTVWYX", dtype=np.uint8)
    recs, rec_bin = [], []
    for b in range(200):
        for _ in range(int(rng.integers(0, 3))):
            recs.append(rng.choice(alpha, int(rng.integers(0, 400))).tobytes())
            rec_bin.append(b)
    offs = np.concatenate([[0], np.cumsum([len(r) for r in recs])]).astype(np.uint64)
    rec_bin = np.array(rec_bin, dtype=np.uint32)
    direct = t.IBF(ctx, 200, 30011, 3)
    direct.dgram_emplace(b"".join(recs), offs, rec_bin, 2, 9)
    monkeypatch.setenv("TRX_DGRAM_SLAB_MIN_MB", "0")
    slab = t.IBF(ctx, 200, 30011, 3)
    slab.dgram_emplace(b"".join(recs), offs, rec_bin, 2, 9)
    a, b = direct.download(), slab.download()
    assert a.any() and np.array_equal(a, b)
    direct.close()
    slab.close()


@pytest.mark.parametrize("build", range(len(GOLD["builds"])))
def test_dindex_file_is_byte_identical(ctx, build, tmp_path):
    b = GOLD["builds"][build]
    out = str(tmp_path / "own.dibf")
    cwd = os.getcwd()
    os.chdir(ROOT)  # the file stores the names as given: relative to the repository root, like the golden run
    try:
        per_bin = sharded.dindex_build(ctx, b["files"], b["min"], b["max"], 3, 0.05, out)
    finally:
        os.chdir(cwd)
    assert [int(x) for x in per_bin] == b["per_bin"]
    assert os.path.getsize(out) == b["bytes"]
    assert hashlib.sha256(open(out, "rb").read()).hexdigest() == b["sha256"]


def test_dindex_command_line(tmp_path):
    """`tetrex dindex -l 2 -u 3 OUT list.lst`: a .lst argument is a list of bins whose names are stored verbatim
    (src/dGramIndex.cpp:24-29) — the file equals the one the reference wrote for the same two bins."""
    tiny = GOLD["tiny"]
    lst = tmp_path / "bins.lst"
    lst.write_text("".join(f + "\n" for f in tiny["files"]))
    out = tmp_path / "tiny.dibf"
    r = subprocess.run([TETREX, "dindex", "-l", str(tiny["min"]), "-u", str(tiny["max"]), str(out), str(lst)], capture_output=True,
                       text=True, cwd=ROOT, timeout=300)
    assert r.returncode == 0, r.stderr
    assert open(out, "rb").read() == open(os.path.join(ROOT, "tests", "golden", "tiny.dibf"), "rb").read()


@pytest.mark.parametrize("build", range(len(GOLD["builds"])))
def test_augmented_queries_equal_reference(ctx, build, tmp_path):
    """`-a` alone (gap bounds 0..0, no filter: the reference's default-constructed DGramIndex) and `-a -g`: candidate bins
    of every golden query, and the printed lines where the reference got as far as printing."""
    b = GOLD["builds"][build]
    cases = [q for q in GOLD["queries"] if q["build"] == build and not q.get("reference_aborts")]
    assert cases
    dibf = str(tmp_path / "own.dibf")
    cwd = os.getcwd()
    os.chdir(ROOT)
    try:
        sharded.dindex_build(ctx, b["files"], b["min"], b["max"], 3, 0.05, dibf)
        ix = sharded.ShardedIndex(ctx, cases[0]["index"], 0, 1)
        ix.set_augment(True)
        n_lines = n_gapped = 0
        for with_dibf in (False, True):
            ix.attach_dgram(dibf if with_dibf else None)
            for q in [q for q in cases if q["dibf"] == with_dibf]:
                try:
                    words, stats = ix.local_candidates(q["regex"])
                except sharded.HostError as e:  # reference UB on reduced alphabets
                    assert "reduced alphabet" in str(e), (q["regex"], str(e))
                    continue
                got = [i for i in range(len(b["files"])) if int(words[i // 64]) >> (i % 64) & 1]
                assert got == q["bins"], (q["index"], q["regex"], with_dibf)
                if q["stdout"] is not None:
                    text, lines = ix.verify_local(words, q["regex"])
                    assert text == q["stdout"], (q["index"], q["regex"], with_dibf)
                    n_lines += lines
        ix.close()
    finally:
        os.chdir(cwd)
    assert n_lines > 0 or "murphy" in cases[0]["index"]


def test_query_command_line_with_augmentation(tmp_path):
    """`tetrex query -a -g FILE IDX REGEX` end to end through the binary, for the hand-written golden patterns."""
    b = GOLD["builds"][0]
    dibf = str(tmp_path / "own.dibf")
    r = subprocess.run([TETREX, "dindex", "-l", str(b["min"]), "-u", str(b["max"]), dibf] + b["files"], capture_output=True,
                       text=True, cwd=ROOT, timeout=300)
    assert r.returncode == 0, r.stderr
    done = 0
    for q in [q for q in GOLD["queries"] if q["build"] == 0 and q.get("stdout")][:24]:
        args = ["query", "-a"] + (["-g", dibf] if q["dibf"] else []) + [q["index"], q["regex"]]
        r = subprocess.run([TETREX] + args, capture_output=True, text=True, cwd=ROOT, timeout=300)
        assert r.returncode == 0, r.stderr
        assert r.stdout == q["stdout"], (q["regex"], q["dibf"], r.stderr)
        done += 1
    assert done >= 10
