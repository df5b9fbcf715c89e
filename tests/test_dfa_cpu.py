"""CPU: the determinised anchored matcher of K4's phase 2 (host-side subset construction in csrc/regex_scan.cu,
`match_longest_dfa` on the device) against RE2's POSIX leftmost-longest answers (tests/golden/regex.json, written by the
compiled reference's RE2) and against the position-automaton semantics on the config-3 pattern population.  No device:
trx_regex_dfa_match_end runs the same table walk on the host."""
import ctypes as C
import json
import os
import random

import numpy as np
import pytest

from tetrex_b200 import synth
from tetrex_b200._abi import lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def compile_posix(pattern):
    h = C.c_void_p()
    rc = lib().trx_regex_compile(pattern.encode("latin-1"), 1, C.byref(h))
    return h if rc == 0 else None


def dfa_end(h, text: bytes, start: int) -> int:
    buf = (C.c_uint8 * max(len(text), 1)).from_buffer_copy(text or b"\0")
    return lib().trx_regex_dfa_match_end(h, buf, len(text), start)


def test_dfa_reproduces_re2_posix_matches():
    cases = [c for c in json.load(open(os.path.join(ROOT, "tests", "golden", "regex.json"))) if c["posix"]]
    checked = with_dfa = 0
    by_pattern = {}
    for c in cases:
        by_pattern.setdefault(c["pattern"], []).append(c)
    for pattern, group in by_pattern.items():
        h = compile_posix(pattern)
        if h is None:
            continue
        info = (C.c_uint32 * 3)()
        if not lib().trx_regex_dfa_info(h, info):
            lib().trx_regex_destroy(h)
            continue
        with_dfa += 1
        assert 2 <= info[2] <= info[0] and info[1] >= 2
        anchored = pattern.startswith("^")  # phase 2 tries start 0 only; the matcher itself does not know about '^'
        for c in group:
            text = c["text"].encode("latin-1")
            pos = 0
            if anchored:
                want = c["matches"][0][1] if c["matches"] else -1
                assert dfa_end(h, text, 0) == want, (pattern, c["text"])
                checked += 1
                continue
            for s, e in c["matches"]:
                # FindAndConsume: no match starts between the previous end and s, the longest match at s ends at e
                for q in range(pos, s):
                    assert dfa_end(h, text, q) == -1, (pattern, c["text"], q)
                assert dfa_end(h, text, s) == e, (pattern, c["text"], s)
                pos = e
                checked += 1
            for q in range(pos, len(text)):
                assert dfa_end(h, text, q) == -1, (pattern, c["text"], q)
        lib().trx_regex_destroy(h)
    assert with_dfa >= 20 and checked >= 150, (with_dfa, checked)


def python_longest(pattern: str, text: str, start: int) -> int:
    """Longest match of a PROSITE-style pattern (classes, '.', {m}, {m,n}; no alternation) anchored at start, by
    set-of-positions simulation — independent of the library's automaton."""
    items, i = [], 0  # (set of chars or None for '.', lo, hi)
    while i < len(pattern):
        if pattern[i] == "[":
            j = pattern.index("]", i)
            body = pattern[i + 1:j]
            cls = ("^", set(body[1:])) if body.startswith("^") else ("", set(body))
            i = j + 1
        elif pattern[i] == ".":
            cls, i = None, i + 1
        else:
            cls, i = ("", {pattern[i]}), i + 1
        lo = hi = 1
        if i < len(pattern) and pattern[i] == "{":
            j = pattern.index("}", i)
            parts = pattern[i + 1:j].split(",")
            lo, hi = int(parts[0]), int(parts[-1])
            i = j + 1
        items.append((cls, lo, hi))

    def ok(cls, ch):
        if cls is None:
            return True
        neg, chars = cls
        return (ch not in chars) if neg else (ch in chars)
    ends = {start}
    for cls, lo, hi in items:
        nxt = set()
        for e in ends:
            p = e
            for n in range(1, hi + 1):
                if p >= len(text) or not ok(cls, text[p]):
                    break
                p += 1
                if n >= lo:
                    nxt.add(p)
        ends = nxt
        if not ends:
            return -1
    return max(ends)


def test_dfa_on_the_config3_pattern_population():
    """Every one of the 1000 PROSITE-style patterns either has a DFA that agrees with an independent simulation on texts
    built around instances of the pattern, or reports that it has none."""
    rng = random.Random(11)
    pats = synth.prosite_patterns(1000, 1000)
    aa = "ACDEFGHIKLMNPQRSTVWY"
    have = none = checked = 0
    sizes = []
    for p in pats:
        h = compile_posix(p)
        assert h is not None, p
        info = (C.c_uint32 * 3)()
        if not lib().trx_regex_dfa_info(h, info):
            none += 1
            lib().trx_regex_destroy(h)
            continue
        have += 1
        sizes.append(int(info[0]))
        if have % 4 == 0:  # a quarter of them through the simulation
            inst = synth.pattern_instance(p, rng).decode()
            text = "".join(rng.choice(aa) for _ in range(6)) + inst + "".join(rng.choice(aa) for _ in range(8)) + inst[:len(inst) // 2]
            raw = text.encode()
            for s in range(len(text)):
                assert dfa_end(h, raw, s) == python_longest(p, text, s), (p, text, s)
                checked += 1
        lib().trx_regex_destroy(h)
    assert have >= 900, (have, none)  # the subset automata of this population stay small
    assert checked > 5000
    assert max(sizes) <= 2048


def test_reverse_searcher_finds_exactly_the_match_starts():
    """Phase 1 for patterns without a selective necessary factor: the reverse searcher marks p iff a match starts at p —
    checked against the anchored matcher at every position, on the config-3 population and on hand-written patterns;
    unbounded or anchored patterns have no reverse searcher."""
    rng = random.Random(5)
    aa = "ACDEFGHIKLMNPQRSTVWY"
    pats = synth.prosite_patterns(300, 1000) + ["...[GHKM]V[FHKL].", "[ST].[RK]", "(LMAQ|GL.N|W{2})", "L.A.G", "A(C|DE)?F"]
    have = 0
    states = []
    for p in pats:
        h = compile_posix(p)
        inst = synth.pattern_instance(p, rng).decode() if "(" not in p else "LMAQGLANWWACFADEF"
        text = ("".join(rng.choice(aa) for _ in range(40)) + inst + "".join(rng.choice(aa) for _ in range(30)) + inst).encode()
        starts = (C.c_uint8 * len(text))()
        buf = (C.c_uint8 * len(text)).from_buffer_copy(text)
        m = lib().trx_regex_dfa_starts(h, buf, len(text), starts)
        if m == 0:
            lib().trx_regex_destroy(h)
            continue
        have += 1
        assert m >= len(inst) or "(" in p, (p, m)
        for s in range(len(text)):
            assert bool(starts[s]) == (dfa_end(h, text, s) >= 0), (p, text, s)
        lib().trx_regex_destroy(h)
    assert have >= 295, have
    for p in ["AQGL+YN", "^..[DE]", "[KR]{2,3}$", "A(CD)*E"]:
        h = compile_posix(p)
        buf = (C.c_uint8 * 8)()
        assert lib().trx_regex_dfa_starts(h, buf, 8, (C.c_uint8 * 8)()) == 0, p
        lib().trx_regex_destroy(h)
