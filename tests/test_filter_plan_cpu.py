"""CPU: the phase-1 prefilter plans of K4 (fixed literal window / gapped atoms, trx_regex_filter_plan) are NECESSARY
conditions of every match.  The plan is derived on the host from the position automaton (dominator chain + walk
ranges, csrc/regex_scan.cu derive_filter / derive_atoms); here it is replayed in numpy over random text with planted
matches and every match start Python's `re` finds (all starts, overlapping ones included — a superset of what
FindAndConsume, lib/re2/re2/re2.cc:428-437, reports) must lie in a flagged range.  No GPU involved."""
import random
import re
import zlib

import numpy as np
import pytest

import tetrex_b200 as t

try:  # Python >= 3.11
    import re._parser as sre_parse
except ImportError:  # pragma: no cover
    import sre_parse

AA = "ACDEFGHIKLMNPQRSTVWY"
AA_FREQ = dict(zip(AA, [8.25, 1.37, 5.45, 6.75, 3.86, 7.07, 2.27, 5.96, 5.84, 9.66, 2.42, 4.06, 4.70, 3.93, 5.53, 6.56,
                        5.34, 6.87, 1.08, 2.92]))
COMP = bytes.maketrans(b"ACGTUMRWSYKVHDBNacgtumrwsykvhdbn", b"TGCAAKYWSRMBDHVNtgcaakywsrmbdhvn")


def sample(pattern: str, alphabet: str, rng: random.Random) -> str:
    """A random string the pattern matches (syntax subset TetRex's lexer accepts)."""
    def gen(items):
        out = []
        for op, arg in items:
            name = str(op)
            if name == "LITERAL":
                out.append(chr(arg))
            elif name == "NOT_LITERAL":
                out.append(rng.choice([c for c in alphabet if c != chr(arg)]))
            elif name == "ANY":
                out.append(rng.choice(alphabet))
            elif name == "IN":
                neg = any(str(o) == "NEGATE" for o, _ in arg)
                chars = [chr(a) for o, a in arg if str(o) == "LITERAL"]
                out.append(rng.choice([c for c in alphabet if c not in chars] if neg else chars))
            elif name in ("MAX_REPEAT", "MIN_REPEAT"):
                lo, hi, sub = arg
                hi = lo + 3 if hi > 64 else hi
                for _ in range(rng.randint(lo, hi)):
                    out.append(gen(sub))
            elif name == "SUBPATTERN":
                out.append(gen(arg[-1]))
            elif name == "BRANCH":
                out.append(gen(rng.choice(arg[1])))
            else:
                raise AssertionError(name)
        return "".join(out)
    return gen(sre_parse.parse(pattern))


def flagged_starts(plan: dict, text: np.ndarray) -> np.ndarray:
    """Boolean array: addresses at which the plan allows a match to start (exact replay, no chunk rounding)."""
    n = len(text)
    ok = np.zeros(n, dtype=bool)
    if plan["kind"] == "window":
        hit = np.ones(n, dtype=bool)
        for off, lit in plan["literals"]:
            m = np.zeros(n, dtype=bool)
            m[:n - off] = text[off:] == lit
            hit &= m
        for a in np.flatnonzero(hit):
            s = a + plan["start_delta"]
            if 0 <= s < n:
                ok[s] = True
        return ok
    assert plan["kind"] == "atoms"
    hit = np.ones(n, dtype=bool)
    lits = set()
    for byts, lo, hi in plan["atoms"]:
        assert 0 <= lo <= hi <= 31 and 1 <= len(byts) <= 3
        lits |= set(byts)
        m = np.zeros(n, dtype=bool)
        for d in range(lo, hi + 1):
            m[:n - d] |= np.isin(text[d:], byts)
        hit &= m
    assert len(lits) == plan["n_lit"] <= 4
    for byts, lo, hi in plan.get("confirm", []):  # second stage: evaluated at the surviving anchors, offsets signed
        assert -128 <= lo <= hi <= 127 and hi - lo <= 7 and len(byts) >= 1
        m = np.zeros(n, dtype=bool)
        member = np.isin(text, byts)
        for d in range(lo, hi + 1):
            if d >= 0:
                m[:n - d] |= member[d:]
            else:
                m[-d:] |= member[:n + d]
        hit &= m
    assert any(lo == 0 and hi == 0 for _, lo, hi in plan["atoms"]), "one atom is the anchor"
    for a in np.flatnonzero(hit):
        lo, hi = max(0, a + plan["start_lo"]), min(n - 1, a + plan["start_hi"])
        if lo <= hi:
            ok[lo:hi + 1] = True
    return ok


def frequencies(alphabet_freq: dict) -> np.ndarray:
    f = np.zeros(256)
    total = sum(alphabet_freq.values())
    for ch, v in alphabet_freq.items():
        f[ord(ch)] = v / total
    return f


PEPTIDE = [
    "C.{2,4}C.{3}[LIVMFYWC].{8}H.{3,5}H",   # PS00028 zinc finger C2H2
    "LMA(E|Q)GLYN", "MA[EQ]GLY", "L.A.G", "AQGL+YN", "QGLYN$", "^MKT.{1,3}W",
    "D.{0,3}EW", "C.{2}C.*H.{3}H", "W.{2,4}C.{5,9}CW", "(AB|CD)W.{1,2}M(K|R)H", "C(A|.{2,3})W.{0,2}M",
    "M.{2}(C|W)H.{3,4}(WK|WR)C", "HC+W.{3}M", "(C.{2,3}){2}WH", "C.{12,16}W.{10,14}H.{2}M", "W{2,3}.{4}C",
    "C[ST]?W.{2,5}HM", "WM.{30,31}CH", "C.{3}C.{3}C.{3}W", "N[^P][ST][^P]", "[RK].{2,3}[DE].{2,3}Y",
    "[DM]D[KR]K[HLMY][EFS][CI]", "[HI]{2,4}[CSV]{2,4}R[NP].E", "[EMVWY]F[FR]Y[CST]{2,4}.", "H[IMPTV]G.[GHNPT].",
    "[FHI][AGS].S.M", "R..[AV][MQW][DW]", "[CW][HM].{1,3}[CW]K", "GF[DMN].[CG].[INPSY]{2,4}.{2,4}[GQT][KR]",
    ".N.I.[GRS]", "...[GPST][NRSWY]F", ".[GRWY].I[DGNQR]E.", "[LQTY]E.{2,4}V.{2,4}.F[IKMRY]", "Y[GIMR]{2,4}.F..",
    "[AIQVY][DILRY][INRW][AEGKQ]{2,4}[ADFKM].M", "..[MV]{2,4}[EHLMW][FGLS][LV][DHPTY][PQTW]", "[ILV]{2}.K[DE]{1,3}[FYW].N",
]
DNA = ["GATTACAG", "ACGTAC(G|T)", "TTGACA.{3}T", "TATA.{2,4}CGCG", "GG.{1,6}CCATGG", "A(C+|G+)T", "CG.{3}(AT|TA)GC.{0,2}GG"]


@pytest.mark.parametrize("pattern", PEPTIDE)
def test_peptide_plan_is_necessary(pattern):
    rng = random.Random(zlib.crc32(pattern.encode()))
    freq = frequencies(AA_FREQ)
    rx = t.Regex(pattern, True)
    plan = rx.filter_plan(freq)
    body = pattern.strip("^$")
    # text: background residues with planted instances (also back to back and at both ends)
    parts = [sample(body, AA, rng), sample(body, AA, rng)]
    for _ in range(300):
        parts.append("".join(rng.choices(AA, weights=[AA_FREQ[c] for c in AA], k=rng.randint(0, 90))))
        parts.append(sample(body, AA, rng))
    text = "".join(parts)
    starts = [m.start() for m in re.finditer("(?=(" + body + "))", text)]
    assert len(starts) >= 300
    if plan["kind"] == "none":
        return
    ok = flagged_starts(plan, np.frombuffer(text.encode(), dtype=np.uint8))
    missed = [s for s in starts if not ok[s]]
    assert not missed, (pattern, plan, missed[:5], text[missed[0]:missed[0] + 60])
    assert ok.mean() < 0.2, (pattern, plan, ok.mean())  # and it still filters


@pytest.mark.parametrize("pattern", DNA)
@pytest.mark.parametrize("reverse", [False, True])
def test_dna_plan_is_necessary(pattern, reverse):
    rng = random.Random(zlib.crc32(pattern.encode()) + reverse)
    freq = frequencies({c: 1.0 for c in "ACGT"})
    rx = t.Regex(pattern, False)
    plan = rx.filter_plan(freq, reverse_strand=reverse)
    parts = []
    for _ in range(300):
        parts.append("".join(rng.choices("ACGT", k=rng.randint(0, 60))))
        parts.append(sample(pattern, "ACGT", rng))
    read = "".join(parts)  # what the automaton reads
    starts = [m.start() for m in re.finditer("(?=(" + pattern + "))", read)]
    assert len(starts) >= 300
    if plan["kind"] == "none":
        return
    if reverse:
        # the store holds the other strand: the scan reads it back to front through comp_tab (src/query.cpp:167-191)
        stored = read.encode().translate(COMP)[::-1]
        addr = [len(read) - 1 - s for s in starts]
    else:
        stored = read.encode()
        addr = starts
    ok = flagged_starts(plan, np.frombuffer(stored, dtype=np.uint8))
    missed = [a for a in addr if not ok[a]]
    assert not missed, (pattern, reverse, plan, missed[:5])


def test_plan_kinds():
    freq = frequencies(AA_FREQ)
    assert t.Regex("LMA(E|Q)GLYN", True).filter_plan(freq)["kind"] == "window"
    zf = t.Regex("C.{2,4}C.{3}[LIVMFYWC].{8}H.{3,5}H", True).filter_plan(freq)
    assert zf["kind"] == "atoms" and zf["n_lit"] == 2
    assert sorted(zf["atoms"]) == [((67,), 0, 0), ((67,), 3, 5), ((72,), 16, 18), ((72,), 20, 24)]
    assert (zf["start_lo"], zf["start_hi"]) == (0, 0)
    cl = t.Regex("[DM]D[KR]K[HLMY][EFS][CI]", True).filter_plan(freq)  # no literal window is selective enough; small classes are
    assert cl["kind"] == "atoms" and any(len(b) > 1 for b, _, _ in cl["atoms"])
    weak = t.Regex(".N.I.[GRS]", True).filter_plan(freq)  # two weak literals + a class that does not fit the byte budget
    assert weak["kind"] == "atoms" and weak["confirm"] and sorted(weak["confirm"][0][0]) == [ord(c) for c in "GRS"]
    assert t.Regex("(LMAQ|GLYN)", True).filter_plan(freq)["kind"] == "none"   # no necessary literal
    assert t.Regex("[ST].[RK]", True).filter_plan(freq)["kind"] == "none"
    # a reduced alphabet turns every literal into a class of raw bytes: no fixed literal window survives the byte map,
    # classes of up to three raw bytes can still be atoms (and wider ones confirm checks)
    _, redmap = t.peptide_maps(1)
    reduced = "".join(chr(redmap[ord(c)]) for c in "LMAQGLYN")
    assert t.Regex(reduced, True).filter_plan(freq, map256=redmap)["kind"] in ("none", "atoms")


@pytest.mark.parametrize("reduction", [1, 2])
@pytest.mark.parametrize("motif", ["LMAQGLYNCW", "CWHKDEPGCC", "HP.{2,3}WCM"])
def test_reduced_alphabet_plan_is_necessary(reduction, motif):
    """verify_reduced_fasta_hit (src/query.cpp:239-291): the RECORD is mapped through redmap_ and the reduced regex runs on
    that; the prefilter works on the raw text with the map folded into the byte classes."""
    rng = random.Random(zlib.crc32(motif.encode()) + reduction)
    _, redmap = t.peptide_maps(reduction)
    table = bytes(int(x) for x in redmap)
    pattern = "".join(chr(redmap[ord(c)]) if c.isalpha() else c for c in motif)  # reduce_query_alphabet on the literals
    freq = frequencies(AA_FREQ)
    plan = t.Regex(pattern, True).filter_plan(freq, map256=redmap)
    parts = []
    for _ in range(300):
        parts.append("".join(rng.choices(AA, weights=[AA_FREQ[c] for c in AA], k=rng.randint(0, 90))))
        inst = sample(pattern, "".join(sorted(set(chr(b) for b in table[65:91]))), rng)
        # a raw preimage of the reduced instance: any residue that maps to the wanted reduced letter
        parts.append("".join(rng.choice([a for a in AA if chr(redmap[ord(a)]) == ch] or [ch]) for ch in inst))
    raw = "".join(parts)
    mapped = raw.encode().translate(table).decode()
    starts = [m.start() for m in re.finditer("(?=(" + pattern + "))", mapped)]
    assert len(starts) >= 300
    if plan["kind"] == "none":
        return
    ok = flagged_starts(plan, np.frombuffer(raw.encode(), dtype=np.uint8))
    missed = [s for s in starts if not ok[s]]
    assert not missed, (pattern, plan, missed[:5])
