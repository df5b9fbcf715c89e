"""Host: the class window of trx_regex_scan_multi's bit-parallel first stage (regex_window_multi) is a NECESSARY condition
of every match — checked against RE2's hits for BASELINE config 3's 1000 patterns on the /100 twin
(tests/golden/prosite_batch.npz, written from the unmodified reference's RE2) without a GPU: for every hit there is a
text position p in [start + off_lo, start + off_hi] where the window's classes hold at p .. p+len-1 and every confirm
check finds its class at p + d for a d of its range."""
import os

import numpy as np

import tetrex_b200 as t
from tetrex_b200 import synth


def holds(text, plan, start):
    n = len(text)
    for p in range(start + plan["off_lo"], start + plan["off_hi"] + 1):
        if p < 0 or p + plan["len"] > n:
            continue
        if not all(text[p + k] in plan["classes"][k] for k in range(plan["len"])):
            continue
        if all(any(0 <= p + d < n and text[p + d] in cls for d in range(lo, hi + 1)) for cls, lo, hi in plan["confirm"]):
            return True
    return False


def test_class_window_is_necessary_for_every_re2_hit(golden_dir):
    z = np.load(os.path.join(golden_dir, "prosite_batch.npz"))
    residues, offs, rec_bin, _ = synth.swissprot_like(5700, 1024)
    pats = synth.prosite_patterns(1000, 1000)
    synth.plant_patterns(residues, offs, pats)
    text = residues.tobytes()
    freq = np.bincount(residues, minlength=256).astype(np.float64)
    freq /= freq.sum()
    want = z["hits"]  # (pattern, record, start, end)
    bounds = np.searchsorted(want[:, 0], np.arange(len(pats) + 1))
    planned = with_checks = checked = 0
    for i, p in enumerate(pats):
        rx = t.Regex(p, True)
        plan = rx.window_plan(freq)
        rx.close()
        if plan is None:
            continue
        planned += 1
        with_checks += bool(plan["confirm"])
        assert 1 <= plan["len"] <= 6 and plan["off_hi"] - plan["off_lo"] <= 31 and plan["confirmed"] <= 1.2e-4
        hits = want[bounds[i]:bounds[i + 1]]
        for h in hits[:: max(1, len(hits) // 40)]:
            assert holds(text, plan, int(offs[int(h[1])]) + int(h[2])), (p, plan, h)
            checked += 1
    assert planned >= 950 and with_checks >= 10 and checked >= 2000, (planned, with_checks, checked)


def test_patterns_without_a_window():
    freq = np.full(256, 0.0)
    freq[[ord(c) for c in "ACDEFGHIKLMNPQRSTVWY"]] = 0.05
    for p, ok in [("...[GHKM]V[FHKL].", False), ("^MKV", False), ("WCWHW", True), ("A+", False), ("RT.{2,4}.C[DGHPR]L", True),
                  ("C.{2,4}C.{3}[LIVMFYWC].{8}H.{3,5}H", False)]:
        rx = t.Regex(p, True)
        assert (rx.window_plan(freq) is not None) == ok, p
        rx.close()


def test_plans_of_wide_patterns_are_necessary_for_every_re2_match(golden_dir):
    """Patterns of 130..250 symbol positions (tests/golden/regex_wide.json, RE2's own match lists): whatever prefilter the
    single-pattern path picks (trx_regex_filter_plan) and whatever class window the multi-pattern path picks
    (trx_regex_window_plan) must allow every match RE2 reports — offsets beyond 127 bytes must not wrap anywhere."""
    import json
    from test_filter_plan_cpu import AA_FREQ, flagged_starts, frequencies
    cases = json.load(open(os.path.join(golden_dir, "regex_wide.json")))
    checked = plans = windows = 0
    for c in cases:
        try:
            rx = t.Regex(c["pattern"], c["posix"])
        except t.TetrexError:
            continue
        text = np.frombuffer(c["text"].encode(), dtype=np.uint8)
        if len(text) == 0:
            rx.close()
            continue
        # planned for a text with protein / uniform DNA frequencies (a necessary condition is one on ANY text; the golden
        # texts use small alphabets, on whose own frequencies hardly any prefilter is selective enough to be chosen)
        freq = frequencies(AA_FREQ) if c["posix"] else frequencies(dict.fromkeys("ACGT", 1.0))
        plan = rx.filter_plan(freq)
        if plan["kind"] != "none":
            plans += 1
            ok = flagged_starts(plan, text)
            for a, _ in c["matches"]:
                assert ok[a], (c["pattern"], plan, a)
                checked += 1
        if c["posix"]:
            w = rx.window_plan(freq)
            if w is not None:
                windows += 1
                for a, _ in c["matches"]:
                    assert holds(c["text"].encode(), w, a), (c["pattern"], w, a)
                    checked += 1
        rx.close()
    assert plans >= 8 and windows >= 3 and checked >= 25, (plans, windows, checked)
