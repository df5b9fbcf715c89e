"""GPU parity tests for rows a9/a10: the verify scan against RE2 golden vectors (generated from the
compiled reference), the Python restatement, and — where oracle/_ref travelled — RE2 itself."""
import json
import os
import random

import numpy as np
import pytest

import tetrex_b200 as t
from oracle import bindings as ob

pytestmark = pytest.mark.gpu

COMP = bytes.maketrans(b"ACGTUMRWSYKVHDBNacgtumrwsykvhdbn", b"TGCAAKYWSRMBDHVNtgcaakywsrmbdhvn")


def build_text(ctx, seqs, rec_bin, n_bins):
    offs = np.zeros(len(seqs) + 1, dtype=np.uint64)
    offs[1:] = np.cumsum([len(s) for s in seqs])
    return t.Text(ctx, b"".join(seqs), offs, np.asarray(rec_bin, dtype=np.uint32), n_bins)


def expected_hits(seqs, rec_bin, pattern, posix, finder, cand=None, transform=None):
    out = []
    for r, s in enumerate(seqs):
        if cand is not None and rec_bin[r] not in cand:
            continue
        text = transform(s) if transform else s
        for a, b in finder(pattern, posix, text):
            out.append((rec_bin[r], r, a, b))
    return np.array(out, dtype=np.uint32).reshape(-1, 4)


def test_regex_golden_vectors(ctx, golden_dir):
    """Every (pattern, text) case of tests/golden/regex.json, grouped by pattern into one multi-record text."""
    cases = json.load(open(os.path.join(golden_dir, "regex.json")))
    groups = {}
    for c in cases:
        groups.setdefault((c["pattern"], c["posix"]), []).append(c)
    assert len(groups) >= 50
    for (pat, posix), cs in groups.items():
        seqs = [c["text"].encode() for c in cs]
        if sum(len(s) for s in seqs) == 0:
            continue
        rec_bin = list(range(len(seqs)))
        txt = build_text(ctx, seqs, rec_bin, len(seqs))
        got = txt.scan(t.Regex(pat, posix))
        want = [(rec_bin[r], r, a, b) for r, c in enumerate(cs) for a, b in c["matches"]]
        assert got.tolist() == [list(x) for x in want], (pat, posix)
        txt.close()


def test_regex_wide_patterns_equal_re2(ctx, golden_dir, monkeypatch):
    """Patterns of 129..250 symbol positions (`.{30,80}`-style PROSITE gaps; four 64-bit words of automaton state, the
    W = 4 instantiations of the scan kernels) against the reference's own RE2 (tests/golden/regex_wide.json, written by
    make_golden.py regex_wide): single scans with and without prefilter, and the same patterns through one
    multi-pattern call together with narrow ones.  What does not compile is over 250 positions or non-greedy."""
    cases = json.load(open(os.path.join(golden_dir, "regex_wide.json")))
    groups = {}
    for c in cases:
        groups.setdefault((c["pattern"], c["posix"]), []).append(c)
    done = 0
    for (pat, posix), cs in groups.items():
        try:
            rx = t.Regex(pat, posix)
        except t.TetrexError as e:
            assert "250 symbol positions" in str(e) or "non-greedy" in str(e), (pat, str(e))
            continue
        assert 128 < rx.positions <= 250, pat
        seqs = [c["text"].encode() for c in cs]
        rec_bin = list(range(len(seqs)))
        txt = build_text(ctx, seqs, rec_bin, len(seqs))
        want = [[rec_bin[r], r, a, b] for r, c in enumerate(cs) for a, b in c["matches"]]
        assert txt.scan(rx).tolist() == want, (pat, posix)
        monkeypatch.setenv("TRX_REGEX_FILTER", "off")
        assert txt.scan(rx).tolist() == want, (pat, posix, "no prefilter")
        monkeypatch.delenv("TRX_REGEX_FILTER")
        if posix:
            narrow = [t.Regex(p, True) for p in ("K", "[CH].{2}[LM]")]
            multi = txt.scan_multi([narrow[0], rx, narrow[1]], cap=1 << 18)
            assert multi[1].tolist() == want, (pat, "multi")
            for g, nrx in zip((multi[0], multi[2]), narrow):
                assert np.array_equal(g, txt.scan(nrx, cap=1 << 18)), pat
        txt.close()
        done += 1
    assert done >= 12


PATTERNS_AA = ["LMA(E|Q)GLYN", "[ILV][ILV]K", "G.{2,4}K", "N[ST][^P]", "(AB|A)BC", "K+R?E", "C.{0,3}C", "W(A|C|D){1,3}W",
               "^M", "K$", "(GA)+G", "A{3,5}", "[DE]{2}.[KR]{2}", "(A|AA)(A|AA)C"]
PATTERNS_NA = ["A(C+|G+)T", "ACGT", "(A|AC)(CG|G)", "T+A", "G.{1,3}C", "(AC)*G", "CA?T", "[AG]C[CT]", "A{2,4}C", "(TA|T)A+",
               "AC|ACG|ACGT", "^ACG", "TT$", "GATTACA", "(G|GG|GGG)T", "C(A|AT|ATG)G"]


@pytest.mark.parametrize("posix,patterns,alpha", [(True, PATTERNS_AA, b"ACDEGKLMNPQRSTWILVY"), (False, PATTERNS_NA, b"ACGT"),
                                                  (True, PATTERNS_NA, b"ACGT"), (False, PATTERNS_AA, b"ABCDEKLMQGWR")])
def test_regex_random_multi_bin(ctx, posix, patterns, alpha):
    """Ragged records (empty, 1 byte, chunk-straddling, multi-tile) in several bins, candidate subsets."""
    rng = random.Random(sum(alpha) + len(patterns))
    lengths = [0, 1, 2, 127, 128, 129, 255, 256, 257, 1000, 0, 5000, 40000, 33, 70000, 3, 0, 640, 128 * 256, 128 * 256 + 1, 90]
    seqs = [bytes(rng.choice(alpha) for _ in range(L)) for L in lengths]
    n_bins = 7
    rec_bin = sorted(rng.randrange(n_bins) for _ in seqs)
    txt = build_text(ctx, seqs, rec_bin, n_bins)
    finder = ob.ref_find_all if ob.have_ref() else ob.find_all_python
    for pat in patterns:
        rx = t.Regex(pat, posix)
        got, residues = txt.scan(rx, stats=True)
        want = expected_hits(seqs, rec_bin, pat, posix, finder)
        assert np.array_equal(got, want), (pat, posix, len(got), len(want))
        assert residues == sum(lengths)
        cand_bins = {1, 4, 6}
        cand = np.zeros(1, dtype=np.uint64)
        for b in cand_bins:
            cand[0] |= np.uint64(1) << np.uint64(b)
        got = txt.scan(rx, candidate_bins=cand)
        assert np.array_equal(got, expected_hits(seqs, rec_bin, pat, posix, finder, cand=cand_bins)), (pat, "subset")
    txt.close()


def test_regex_staging_variants_agree(ctx, monkeypatch):
    """The TMA-staged kernel and the direct-load variant (TRX_REGEX_STAGING=ldg) are the same algorithm."""
    rng = random.Random(3)
    seqs = [bytes(rng.choice(b"ACGT") for _ in range(L)) for L in [100000, 5, 300000, 77777]]
    txt = build_text(ctx, seqs, [0, 0, 1, 2], 3)
    rx = t.Regex("GATTA(C|G)A", False)
    a = txt.scan(rx)
    monkeypatch.setenv("TRX_REGEX_STAGING", "ldg")
    b = txt.scan(rx)
    assert len(a) > 0 and np.array_equal(a, b)
    txt.close()


def test_dfa_matcher_equals_position_automaton(ctx, monkeypatch):
    """Phase 2 resolves POSIX patterns with a determinised matcher (TRX_REGEX_DFA=off: the position automaton), and patterns
    without any necessary factor skip phase 1 when they have one: same hits either way and equal to RE2 / the Python
    restatement, including a pattern dense enough for the device-side hit formatting (> 1024 hits)."""
    rng = random.Random(77)
    aa = b"ACDEFGHIKLMNPQRSTVWY"
    seqs = [bytes(rng.choice(aa) for _ in range(L)) for L in [30000, 127, 128, 5, 0, 90000, 129, 4000, 1, 2, 3, 700]]
    rec_bin = [0, 0, 1, 1, 2, 3, 3, 4, 4, 4, 5, 5]
    txt = build_text(ctx, seqs, rec_bin, 6)
    finder = ob.ref_find_all if ob.have_ref() else ob.find_all_python
    cand_some = np.zeros(1, dtype=np.uint64)
    cand_some[0] = 0b101010
    for pat in ["...[GHKM]V[FHKL].", "[ST].[RK]", ".[ACDEL]D.{2,4}.[AEGQT]S", "L.A.G", "[LI]MAQ", "AQGL+YN", "^..[DE]", "[KR]{2,3}$",
                "C.{2,4}C.{3}[LIVMFYWC]", "(LMAQ|GL.N|W{2})"]:
        rx = t.Regex(pat, True)
        a = txt.scan(rx, cap=1 << 20)
        a_some = txt.scan(rx, candidate_bins=cand_some, cap=1 << 20)
        monkeypatch.setenv("TRX_REGEX_DFA", "off")
        b = txt.scan(rx, cap=1 << 20)
        monkeypatch.delenv("TRX_REGEX_DFA")
        monkeypatch.setenv("TRX_REGEX_STARTS", "off")  # no reverse searcher: every record of a candidate bin, every start
        c = txt.scan(rx, cap=1 << 20)
        monkeypatch.delenv("TRX_REGEX_STARTS")
        want = expected_hits(seqs, rec_bin, pat, True, finder)
        assert np.array_equal(a, want), pat
        assert np.array_equal(b, want), pat
        assert np.array_equal(c, want), pat
        assert np.array_equal(a_some, expected_hits(seqs, rec_bin, pat, True, finder, cand={1, 3, 5})), pat
    dense = txt.scan(t.Regex("[ST].[RK]", True), cap=1 << 20)
    assert len(dense) > 1024
    txt.close()


def test_factor_filter_equals_exact_automaton(ctx, monkeypatch):
    """Phase 1 has two variants (SIMD necessary-factor filter / exact automaton); both must lead to the same hits."""
    rng = random.Random(12)
    aa = b"ACDEFGHIKLMNPQRSTVWY"
    seqs = [bytes(rng.choice(aa) for _ in range(L)) for L in [3000, 127, 128, 5, 0, 60000, 129, 4000]]
    motif = b"LMAQGLYN"
    for r, pos in [(0, 5), (0, 120), (0, 2992), (1, 119), (2, 0), (2, 120), (5, 127), (5, 128 * 256 - 3), (5, 59992), (6, 121)]:
        b = bytearray(seqs[r])
        b[pos:pos + 8] = motif
        seqs[r] = bytes(b)
    rec_bin = [0, 0, 1, 1, 2, 3, 3, 4]
    txt = build_text(ctx, seqs, rec_bin, 5)
    finder = ob.ref_find_all if ob.have_ref() else ob.find_all_python
    dna = [bytes(rng.choice(b"ACGT") for _ in range(L)) for L in [50000, 130, 7]]
    dtxt = build_text(ctx, dna, [0, 1, 1], 2)
    revcomp = lambda s: s.translate(COMP)[::-1]
    for pat, posix, used in [("LMA(E|Q)GLYN", True, True), ("MA[EQ]GLY", True, True), ("L.A.G", True, True), ("[LI]MAQ", True, True),
                             ("(LMAQ|GLYN)", True, False), ("AQGL+YN", True, True), ("QGLYN$", True, True)]:
        rx = t.Regex(pat, posix)
        a = txt.scan(rx)
        assert (txt.last_filter_literals >= 2) == used, (pat, txt.last_filter_literals)
        monkeypatch.setenv("TRX_REGEX_FILTER", "off")
        b = txt.scan(rx)
        assert txt.last_filter_literals == 0
        monkeypatch.delenv("TRX_REGEX_FILTER")
        want = expected_hits(seqs, rec_bin, pat, posix, finder)
        assert np.array_equal(a, want) and np.array_equal(b, want), pat
        assert len(want) > 0, pat
    for pat in ["GATTACAG", "ACGTAC(G|T)", "TTGACA.{3}T"]:
        rx = t.Regex(pat, False)
        for rev in (False, True):
            a = dtxt.scan(rx, reverse_strand=rev)
            assert rev or dtxt.last_filter_literals >= 6, pat  # reverse: 'A' has two raw preimages (T, U) → fewer literals
            want = expected_hits(dna, [0, 1, 1], pat, False, finder, transform=revcomp if rev else None)
            assert np.array_equal(a, want), (pat, rev)
    txt.close()
    dtxt.close()


def test_atom_filter_equals_exact_automaton(ctx, monkeypatch):
    """Phase 1, gapped variant (literal masks + shifted mask algebra, regex_filter_atoms): PROSITE-style patterns with
    x(m,n) gaps, instances planted across chunk (128 B) and tile (32 KB) borders, at record ends and back to back."""
    from test_filter_plan_cpu import sample
    rng = random.Random(77)
    aa = "ACDEFGHIKLMNPQRSTVWY"
    weights = [8.25, 1.37, 5.45, 6.75, 3.86, 7.07, 2.27, 5.96, 5.84, 9.66, 2.42, 4.06, 4.70, 3.93, 5.53, 6.56, 5.34, 6.87, 1.08, 2.92]
    lengths = [70000, 127, 40, 0, 33000, 128 * 256, 128 * 256 + 5, 900, 66000]
    rec_bin = [0, 0, 1, 1, 2, 2, 3, 3, 4]
    finder = ob.ref_find_all if ob.have_ref() else ob.find_all_python
    for pat in ["C.{2,4}C.{3}[LIVMFYWC].{8}H.{3,5}H", "W.{2,4}C.{5,9}CW", "D.{0,3}EW", "M.{2}(C|W)H.{3,4}(WK|WR)C",
                "C.{12,16}W.{10,14}H.{2}M", "WM.{30,31}CH", "(C.{2,3}){2}WH", "C[ST]?W.{2,5}HM$",
                # weak literals + classes that only fit the confirm stage (evaluated at the surviving anchors)
                ".N.I.[GRS]", "R..[AV][MQW][DW]", "[LQTY]E.{2,4}V.{2,4}.F[IKMRY]", "H[IMPTV]G.[GHNPT]."]:
        body = pat.rstrip("$")
        seqs = [bytearray("".join(rng.choices(aa, weights=weights, k=L)).encode()) for L in lengths]
        planted = 0
        for r, L in enumerate(lengths):
            spots = [0, 60, 100, 120, 127, 128, 250, 32768 - 40, 32768 - 20, 32768 - 3, 32768, 65536 - 10, L - 60, L - 30]
            for pos in spots:
                inst = sample(body, aa, rng).encode()
                if pos < 0 or pos + len(inst) > L:
                    continue
                seqs[r][pos:pos + len(inst)] = inst
                planted += 1
            inst = sample(body, aa, rng).encode()  # one flush with the record end
            if len(inst) <= L:
                seqs[r][L - len(inst):] = inst
        seqs = [bytes(x) for x in seqs]
        txt = build_text(ctx, seqs, rec_bin, 5)
        rx = t.Regex(pat, True)
        monkeypatch.setenv("TRX_REGEX_FILTER", "atoms")
        a = txt.scan(rx)
        assert txt.last_filter_literals >= 2, (pat, txt.last_filter_literals)
        monkeypatch.setenv("TRX_REGEX_FILTER", "off")
        b = txt.scan(rx)
        assert txt.last_filter_literals == 0
        monkeypatch.delenv("TRX_REGEX_FILTER")
        c = txt.scan(rx)
        want = expected_hits(seqs, rec_bin, pat, True, finder)
        assert len(want) >= (5 if pat.endswith("$") else planted // 3), (pat, len(want), planted)
        assert np.array_equal(a, want), (pat, "atoms", len(a), len(want))
        assert np.array_equal(b, want) and np.array_equal(c, want), pat
        cand = np.array([0b10101], dtype=np.uint64)
        monkeypatch.setenv("TRX_REGEX_FILTER", "atoms")
        got = txt.scan(rx, candidate_bins=cand)
        monkeypatch.delenv("TRX_REGEX_FILTER")
        assert np.array_equal(got, expected_hits(seqs, rec_bin, pat, True, finder, cand={0, 2, 4})), (pat, "subset")
        txt.close()
    # DNA, both strands (leftmost-first semantics, complement folded into the byte table)
    revcomp = lambda s: s.translate(COMP)[::-1]
    for pat in ["TATA.{2,4}CGCG", "GG.{1,6}CCATGG", "CG.{3}(AT|TA)GC.{0,2}GG"]:
        dna = [bytearray(bytes(rng.choice(b"ACGT") for _ in range(L))) for L in [50000, 130, 7, 128 * 256 + 64]]
        for r, L in enumerate([50000, 130, 7, 128 * 256 + 64]):
            for pos in [0, 100, 125, 32768 - 8, 32768 - 1, L - 20]:
                for rev in (False, True):
                    inst = sample(pat, "ACGT", rng).encode()
                    inst = revcomp(inst) if rev else inst
                    q = pos + (40 if rev else 0)
                    if q >= 0 and q + len(inst) <= L:
                        dna[r][q:q + len(inst)] = inst
        dna = [bytes(x) for x in dna]
        dtxt = build_text(ctx, dna, [0, 1, 1, 1], 2)
        rx = t.Regex(pat, False)
        for rev in (False, True):
            monkeypatch.setenv("TRX_REGEX_FILTER", "atoms")
            a = dtxt.scan(rx, reverse_strand=rev)
            assert dtxt.last_filter_literals >= 1, (pat, rev)  # SIMD atoms; further positions may sit in the confirm stage
            monkeypatch.delenv("TRX_REGEX_FILTER")
            want = expected_hits(dna, [0, 1, 1, 1], pat, False, finder, transform=revcomp if rev else None)
            assert len(want) >= 4 and np.array_equal(a, want), (pat, rev, len(a), len(want))
        dtxt.close()


def test_pipelined_scans_equal_sequential(ctx):
    """trx_regex_scan_begin/_end: several scans in flight on one text give what one scan at a time gives, tickets can be
    ended in any order, and the fifth begin is refused."""
    rng = random.Random(21)
    aa = b"ACDEFGHIKLMNPQRSTVWY"
    seqs = [bytes(rng.choice(aa) for _ in range(L)) for L in [50000, 300, 0, 70000, 128 * 256 + 7, 5]]
    rec_bin = [0, 0, 1, 2, 2, 3]
    txt = build_text(ctx, seqs, rec_bin, 4)
    pats = ["LMA(E|Q)GLYN", "C.{2,4}C", "[ILV][ILV]K", "W.{2,4}C.{5,9}CW", "N[ST][^P]", "KK+", "MA[EQ]G", "D.{0,3}EW", "HH", "^M", "W$"]
    regs = [t.Regex(p, True) for p in pats]
    want = [txt.scan(rx) for rx in regs]
    assert sum(len(w) for w in want) > 100
    for depth in (1, 2, 4):
        got = txt.scan_many(regs, depth=depth)
        for p, a, b in zip(pats, got, want):
            assert np.array_equal(a, b), (p, depth)
    cand = np.array([0b0101], dtype=np.uint64)
    got = txt.scan_many(regs, candidate_bins=cand, depth=3)
    for p, a, rx in zip(pats, got, regs):
        assert np.array_equal(a, txt.scan(rx, candidate_bins=cand)), p
    # out-of-order end, slot exhaustion, bad tickets
    tickets = [txt.scan_begin(rx) for rx in regs[:4]]
    with pytest.raises(t.TetrexError) as e:
        txt.scan_begin(regs[4])
    assert e.value.code == -5  # TRX_ECAPACITY
    for i in (2, 0, 3, 1):
        assert np.array_equal(txt.scan_end(tickets[i]), want[i]), pats[i]
    with pytest.raises(t.TetrexError):
        txt.scan_end(tickets[0])  # already ended
    with pytest.raises(t.TetrexError):
        txt.scan_end((17, 16))
    # a scan that overflows its capacity reports it at _end and frees the slot
    tk = txt.scan_begin(regs[2], cap=3)
    with pytest.raises(t.TetrexError) as e:
        txt.scan_end(tk)
    assert e.value.code == -5
    assert np.array_equal(txt.scan(regs[2]), want[2])
    txt.close()


def test_reverse_strand(ctx):
    """reverse_verify_fasta_hit (src/query.cpp:167-191): the record is complemented through comp_tab and reversed."""
    rng = random.Random(9)
    seqs = [bytes(rng.choice(b"ACGTN") for _ in range(L)) for L in [16, 16, 500, 0, 1, 30000, 129]]
    rec_bin = [0, 1, 1, 2, 2, 3, 4]
    txt = build_text(ctx, seqs, rec_bin, 5)
    finder = ob.ref_find_all if ob.have_ref() else ob.find_all_python
    revcomp = lambda s: s.translate(COMP)[::-1]
    for pat in ["A(C+|G+)T", "ACG", "T+A", "(A|AC)(CG|G)", "G.{1,3}C"]:
        got = txt.scan(t.Regex(pat, False), reverse_strand=True)
        want = expected_hits(seqs, rec_bin, pat, False, finder, transform=revcomp)
        assert np.array_equal(got, want), pat
    txt.close()


def test_reduced_alphabet_map(ctx):
    """verify_reduced_fasta_hit (src/query.cpp:240-315): text bytes go through redmap_ before matching."""
    rng = random.Random(10)
    aa = b"ACDEFGHIKLMNPQRSTVWY"
    seqs = [bytes(rng.choice(aa) for _ in range(L)) for L in [50, 400, 3000, 129]]
    rec_bin = [0, 0, 1, 2]
    txt = build_text(ctx, seqs, rec_bin, 3)
    finder = ob.ref_find_all if ob.have_ref() else ob.find_all_python
    for red in (t.RED_MURPHY, t.RED_LI):
        _, redmap = t.peptide_maps(red)
        table = bytes(redmap.tolist())
        for pat in ["IIK", "B[KS]B", "A(B|F)G", "S{2,3}P"]:
            got = txt.scan(t.Regex(pat, True), map256=redmap)
            want = expected_hits(seqs, rec_bin, pat, True, finder, transform=lambda s: s.translate(table))
            assert np.array_equal(got, want), (red, pat)
    txt.close()


def test_many_hits_capacity(ctx):
    seqs = [b"AC" * 5000]
    txt = build_text(ctx, seqs, [0], 1)
    rx = t.Regex("AC", False)
    got = txt.scan(rx, cap=6000)
    assert len(got) == 5000 and got[:, 2].tolist() == list(range(0, 10000, 2))
    with pytest.raises(t.TetrexError) as e:
        txt.scan(rx, cap=100)
    assert e.value.code == -5
    txt.close()


def test_128_positions_two_word_state(ctx):
    rng = random.Random(4)
    motif = bytes(rng.choice(b"ACDEFGHIK") for _ in range(100))
    seqs = [bytes(rng.choice(b"ACDEFGHIK") for _ in range(3000)) + motif + b"KK" + motif[:50], motif * 3]
    txt = build_text(ctx, seqs, [0, 1], 2)
    pat = motif.decode()
    rx = t.Regex(pat, True)
    assert rx.positions == 100
    finder = ob.ref_find_all if ob.have_ref() else ob.find_all_python
    assert np.array_equal(txt.scan(rx), expected_hits(seqs, [0, 1], pat, True, finder))
    pat2 = motif[:70].decode() + ".{3,9}" + motif[79:90].decode()
    assert np.array_equal(txt.scan(t.Regex(pat2, True)), expected_hits(seqs, [0, 1], pat2, True, finder))
    txt.close()


def test_prosite_pattern_batch_equals_re2(ctx, golden_dir, monkeypatch):
    """BASELINE config 3's pattern population — synth.prosite_patterns(1000, 1000), 10 % planted — over the /100 twin of
    the Swissprot-shaped collection: EVERY hit of EVERY pattern equals what RE2 (POSIX longest, FindAndConsume, as
    include/query.h:146-149 + src/query.cpp:194-237 use it) returned for the same text (tests/golden/prosite_batch.npz,
    made by tests/golden/make_golden.py from oracle/_ref).  Run pipelined with the planner's own choice of phase 1, and
    again with each phase-1 kind forced, so window, atoms and the exact automaton all face the whole population."""
    from tetrex_b200 import synth
    z = np.load(os.path.join(golden_dir, "prosite_batch.npz"))
    residues, offs, rec_bin, _ = synth.swissprot_like(5700, 1024)
    pats = synth.prosite_patterns(1000, 1000)
    planted = synth.plant_patterns(residues, offs, pats)
    assert planted == z["planted"].tolist()
    crc = int(np.bitwise_xor.reduce(residues.astype(np.uint64) * np.arange(1, len(residues) + 1, dtype=np.uint64)))
    assert crc == int(z["text_crc"]), "the seeded twin is not the text the golden hits were made on"
    want_all = z["hits"]
    assert len(want_all) >= 10000
    bounds = np.searchsorted(want_all[:, 0], np.arange(len(pats) + 1))
    txt = t.Text(ctx, residues, offs, rec_bin, 1024)
    regexes = [t.Regex(p, True) for p in pats]

    def check(results, label):
        kinds_wrong = []
        for pi, got in enumerate(results):
            want = want_all[bounds[pi]:bounds[pi + 1], 1:]
            g = got[:, 1:]  # (record, start, end); bin = rec_bin[record]
            if not (np.array_equal(g, want) and np.array_equal(got[:, 0], rec_bin[got[:, 1]])):
                kinds_wrong.append((pi, pats[pi], len(g), len(want)))
        assert not kinds_wrong, f"{label}: {len(kinds_wrong)} patterns differ from RE2, first {kinds_wrong[:3]}"

    freq = np.bincount(residues, minlength=256).astype(np.float64)
    freq /= freq.sum()
    kinds = {"window": 0, "atoms": 0, "none": 0}
    for rx in regexes:
        kinds[rx.filter_plan(freq)["kind"]] += 1
    assert kinds["window"] > 100 and kinds["atoms"] > 50, kinds
    check(txt.scan_many(regexes, cap=1 << 16, depth=3), "planner's choice, 3 scans in flight")
    for forced in ("off", "atoms", "window"):
        monkeypatch.setenv("TRX_REGEX_FILTER", forced)
        sub = regexes if forced == "off" else regexes[::3]
        res = [txt.scan(rx, cap=1 << 16) for rx in sub]
        if forced != "off":
            full = [np.zeros((0, 4), dtype=np.uint32)] * len(regexes)
            for i, r in zip(range(0, len(regexes), 3), res):
                full[i] = r
            wrong = [i for i in range(0, len(regexes), 3)
                     if not np.array_equal(full[i][:, 1:], want_all[bounds[i]:bounds[i + 1], 1:])]
            assert not wrong, (forced, wrong[:5])
        else:
            check(res, "exact automaton for every pattern")
        monkeypatch.delenv("TRX_REGEX_FILTER")
    for rx in regexes:
        rx.close()
    txt.close()


def test_multi_pattern_scan_equals_re2_and_single_scans(ctx, golden_dir, monkeypatch):
    """trx_regex_scan_multi (one pass over the text per batch of <= 128 patterns: shared occurrence masks + per-pattern
    necessary-factor terms + one phase-2 launch) on BASELINE config 3's 1000 patterns over the /100 twin: every hit of
    every pattern equals RE2's (tests/golden/prosite_batch.npz); most patterns really take the shared pass; with
    per-pattern candidate bitmaps the answer equals trx_regex_scan's for each; TRX_REGEX_MULTI=off gives the same lists."""
    from tetrex_b200 import synth
    z = np.load(os.path.join(golden_dir, "prosite_batch.npz"))
    residues, offs, rec_bin, _ = synth.swissprot_like(5700, 1024)
    pats = synth.prosite_patterns(1000, 1000)
    synth.plant_patterns(residues, offs, pats)
    want_all = z["hits"]
    bounds = np.searchsorted(want_all[:, 0], np.arange(len(pats) + 1))
    txt = t.Text(ctx, residues, offs, rec_bin, 1024)
    regexes = [t.Regex(p, True) for p in pats]
    got, st = txt.scan_multi(regexes, cap=1 << 17, stats=True)
    assert st["shared"] >= 900 and st["shared"] + st["single"] == 1000 and st["passes"] >= 8, st
    wrong = [i for i in range(len(pats))
             if not (np.array_equal(got[i][:, 1:], want_all[bounds[i]:bounds[i + 1], 1:]) and
                     np.array_equal(got[i][:, 0], rec_bin[got[i][:, 1]]))]
    assert not wrong, f"{len(wrong)} patterns differ from RE2, first {[(i, pats[i]) for i in wrong[:3]]}"
    assert sum(len(g) for g in got) == len(want_all)
    # per-pattern candidate bins: pattern i may only be verified in its own bins
    rng = np.random.default_rng(3)
    sub = list(range(0, 1000, 7))
    cand = rng.integers(0, 2 ** 63, (len(sub), 16), dtype=np.uint64) | (rng.integers(0, 2 ** 63, (len(sub), 16), dtype=np.uint64) << np.uint64(1))
    multi = txt.scan_multi([regexes[i] for i in sub], candidate_bins=cand, cap=1 << 17)
    for k, i in enumerate(sub):
        assert np.array_equal(multi[k], txt.scan(regexes[i], candidate_bins=cand[k], cap=1 << 16)), pats[i]
    # the same batch with the shared pass switched off
    monkeypatch.setenv("TRX_REGEX_MULTI", "off")
    off, st_off = txt.scan_multi(regexes[:200], cap=1 << 17, stats=True)
    assert st_off["shared"] == 0 and st_off["single"] == 200
    assert all(np.array_equal(a, b) for a, b in zip(off, got[:200]))
    monkeypatch.delenv("TRX_REGEX_MULTI")
    # capacity: too small a buffer is reported, not truncated
    with pytest.raises(t.TetrexError) as e:
        txt.scan_multi(regexes, cap=100)
    assert e.value.code == -5
    # an empty batch and a batch of one
    assert txt.scan_multi([]) == []
    assert np.array_equal(txt.scan_multi([regexes[0]])[0], got[0])
    for rx in regexes:
        rx.close()
    txt.close()


def test_multi_pattern_scan_dna_and_edge_shapes(ctx):
    """Other alphabets and shapes through the shared pass: DNA literals, a pattern repeated in one batch, patterns that share
    no letters, matches straddling chunk / tile boundaries (records of 40 000 residues), anchored and unfilterable patterns
    (single path inside the same call)."""
    rng = random.Random(11)
    seqs, rec_bin = [], []
    for b in range(6):
        for r in range(3):
            n = rng.choice([50, 4000, 40000])
            s = bytearray(rng.choice(b"ACGT") for _ in range(n))
            for at in (0, 120, 125, n - 9, n // 2, 32760 if n > 33000 else 7):
                if 0 <= at <= n - 8:
                    s[at:at + 8] = rng.choice([b"GATTACAG", b"TTGACGTA", b"GGGCCCAT"])
            seqs.append(bytes(s))
            rec_bin.append(b)
    txt = build_text(ctx, seqs, rec_bin, 6)
    pats = ["GATTACAG", "TTGACGTA", "GAT{2}ACA", "GG(G|A)CCCAT", "GATTACAG", "TTGAC.TA", "^ACGT", "[AC]+GATTA", "T+", "GGGCCC(AT|TA)"]
    regexes = [t.Regex(p, False) for p in pats]
    got, st = txt.scan_multi(regexes, cap=1 << 20, stats=True)
    assert st["shared"] >= 5, st
    for p, rx, g in zip(pats, regexes, got):
        assert np.array_equal(g, txt.scan(rx, cap=1 << 20)), p
    cand = np.array([0b010110], dtype=np.uint64)
    for g, rx in zip(txt.scan_multi(regexes, candidate_bins=cand, cap=1 << 20), regexes):
        assert np.array_equal(g, txt.scan(rx, candidate_bins=cand, cap=1 << 20))
    for rx in regexes:
        rx.close()
    txt.close()


def test_multi_pattern_scan_overflowing_record_lists_and_result_views(ctx, monkeypatch):
    """A pattern whose class window survives in more records than a shared pass lists (32 768) reports nothing there and is
    scanned again on the single-pattern path inside the same call, its hits landing in the same device list as everybody
    else's; the list comes back sorted by (pattern, record, start).  40 000 short records, three patterns that occur in
    almost all of them, three that occur in a few, one that never occurs; with and without the class-window stage; and the
    arrays a call returns stay valid while the next call runs."""
    rng = random.Random(5)
    alpha = b"ACDEFGHIKLMNPQRSTVWY"
    seqs, rec_bin = [], []
    for r in range(40000):
        s = bytearray(rng.choice(alpha) for _ in range(rng.choice([24, 31, 40])))
        if r % 50:
            s[3:9] = b"WCWHWM"
        if r % 1000 == 7:
            s[12:18] = b"HHPGYK"
        if r % 3 == 0:
            s[18:23] = b"MMQWC"
        seqs.append(bytes(s))
        rec_bin.append(r // 400)
    txt = build_text(ctx, seqs, rec_bin, 100)
    pats = ["WCWHW", "HHPGYK", "CWHWM", "MMQWC", "HHP[GA]Y", "W.WHW", "YYYYWWWW", "[HK]HPGY"]
    regexes = [t.Regex(p, True) for p in pats]
    want = [txt.scan(rx, cap=1 << 20) for rx in regexes]
    assert len(want[0]) > 32768 and len(want[2]) > 32768 and len(want[6]) == 0 and 0 < len(want[1]) < 100
    got, st = txt.scan_multi(regexes, cap=1 << 20, stats=True)
    assert st["single"] >= 3 and st["shared"] + st["single"] == len(pats), st
    for p, g, w in zip(pats, got, want):
        assert np.array_equal(g, w), p
    again = txt.scan_multi(regexes[::-1], cap=1 << 20)  # another call while `got` is alive: `got` must not change
    for p, g, w, a in zip(pats, got, want, again[::-1]):
        assert np.array_equal(g, w) and np.array_equal(a, w), p
    monkeypatch.setenv("TRX_MULTI_WINDOW", "off")
    for p, g, w in zip(pats, txt.scan_multi(regexes, cap=1 << 20), want):
        assert np.array_equal(g, w), p
    monkeypatch.delenv("TRX_MULTI_WINDOW")
    total = sum(len(w) for w in want)
    with pytest.raises(t.TetrexError) as e:
        txt.scan_multi(regexes, cap=total - 1)
    assert e.value.code == -5
    assert sum(len(g) for g in txt.scan_multi(regexes, cap=total)) == total
    for rx in regexes:
        rx.close()
    txt.close()
