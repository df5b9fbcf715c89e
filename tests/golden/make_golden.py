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


REGEX_WIDE = [  # (pattern, posix, alphabet): 129..250 symbol positions after {m,n} expansion
    ("M.{30,80}W.{20,60}K", True, b"ACDEGKLMNPQRSTW"), ("C.{10,50}C.{10,50}H.{5,40}H", True, b"ACDEGHKLMN"),
    ("A{130}", True, b"AAAC"), ("(AC|GT){70}", True, b"ACGT"), ("[ACD]{100}E.{0,99}F", True, b"ACDEF"),
    ("L.{100,200}L", True, b"ACDEGKLMNPQRSTW"), ("G.{249}", True, b"ACDEG"), ("^M.{130}", True, b"ACDEGKLM"),
    (".{140}K$", True, b"ACDEGK"), ("W(AC|DE|KL){20,60}W", True, b"ACDEKLW"), ("ACGT.{100,150}TTGA", False, b"ACGT"),
    ("T(A|AC|ACG){30}T", False, b"ACGT"), ("(A|C)+G{130}", False, b"ACG"), ("GA.{0,200}?TC", False, b"ACGT"),
    ("C.{10,50}C.{10,50}H.{5,40}H", False, b"ACDEGHKLMN"), ("N[ACGT]{128}N", False, b"ACGTN"),
]


def golden_regex_wide():
    """RE2's FindAndConsume sequence (the reference's own RE2, oracle/_ref) for patterns of 129..250 positions."""
    rng = random.Random(17)
    cases = []
    for pat, posix, alpha in REGEX_WIDE:
        for L in [0, 90, 300, 2500]:
            text = bytearray(rand_seq(rng, alpha, L))
            if pat == "A{130}" and L >= 300:
                text[40:40 + 135] = b"A" * 135
                text[200:200 + 90] = b"A" * 90
            if pat == "(A|C)+G{130}" and L >= 300:
                text[100:100 + 140] = b"G" * 140
            if pat == "(AC|GT){70}" and L >= 300:
                text[30:30 + 150] = (b"ACGTGT" * 30)[:150]
            if pat == "[ACD]{100}E.{0,99}F" and L >= 300:
                for i in range(20, 125):
                    text[i] = rng.choice(b"ACD")
                text[125] = ord("E")
            if pat == "N[ACGT]{128}N" and L >= 300:
                for i in range(50, 260):
                    if text[i] == ord("N"):
                        text[i] = ord("A")
                text[60] = text[60 + 129] = ord("N")
            if pat == "W(AC|DE|KL){20,60}W" and L >= 300:
                text[10:10 + 62] = b"W" + b"ACDEKL" * 10 + b"W"
            if pat == "T(A|AC|ACG){30}T" and L >= 300:
                text[70:70 + 62] = b"T" + b"AC" * 30 + b"T"
            text = bytes(text)
            try:
                m = ob.ref_find_all(pat, posix, text)
            except ValueError:
                continue
            cases.append(dict(pattern=pat, posix=posix, text=text.decode(), matches=m))
    json.dump(cases, open(os.path.join(OUT, "regex_wide.json"), "w"))
    print(len(cases), "cases,", sum(len(c["matches"]) for c in cases), "matches;", sum(1 for c in cases if c["matches"]), "non-empty")


def collector_rows_library(seed=23, bins=96, k=4):
    """A small peptide collection for the CPU collector test: records that share planted motifs, so that walks prune, split,
    absorb and match.  Returns (per-bin list of records, alphabet)."""
    rng = random.Random(seed)
    alpha = b"ACDEFGHIKLMNPQRSTVWY"
    motifs = [b"LMAEGLYN", b"LMAQGLYN", b"CAHCWKKH", b"GATTWCAW", b"WWKDEFAC", b"MKVLAAGIV"]
    lib = []
    for b_ in range(bins):
        recs = []
        for _ in range(rng.randint(1, 3)):
            s_ = bytearray(rng.choice(alpha) for _ in range(rng.randint(20, 90)))
            for m in rng.sample(motifs, rng.randint(0, 2)):
                at = rng.randint(0, len(s_) - len(m))
                s_[at:at + len(m)] = m
            recs.append(bytes(s_))
        lib.append(recs)
    return lib


COLLECTOR_ROWS_REGEXES = [
    "LMA(E|Q)GLYN", "L.AQGLYN", "CAHCWKKH", "LMA[EQ]GL", "(XAC|C)AHCW", "W{2}KDEF", "G.TTWC?AW", "MKV(L|I)A+G", "A{2,3}GIV",
    "(LMA|CAH)(EGL|CWK)", "WKDE.{1,2}C", "M.V.A", "K+D", "LMAEGLYN|CAHCWKKH", "[LM][MA]AEG", "GAT{2}W", "(W|K)(K|D)(D|E)F",
    "CA?HC", "LM.{0,2}GLY", "QGLYN", "ZZZZZ", "AAAA", "LMA", "L(M|A)(A|E)(E|G)(G|L)(L|Y)", "(LM|L)(A|AE)GLYN",
]


def golden_collector_rows():
    """The reference's OTFCollector over a flat seqan IBF of collector_rows_library (oracle/_ref: RefIBF.collect) for k = 4
    and 6, plain and augmented: the CPU test replays the same queries through trxh_process_query_with_rows with membership
    rows from the C oracle's IBF."""
    cases = []
    for k in (4, 6):
        lib = collector_rows_library(k=k)
        bins = len(lib)
        per_bin = [np.concatenate([ob.ref_decompose(1, k, 0, r) for r in recs]) for recs in lib]
        bin_size = ob.compute_bitcount(max(len(v) for v in per_bin), 0.05)
        ibf = ob.RefIBF(bins, bin_size, 3)
        for b_, v in enumerate(per_bin):
            ibf.emplace(v, np.full(len(v), b_, dtype=np.uint32))
        rng = random.Random(k)
        regexes = list(COLLECTOR_ROWS_REGEXES)
        flat = [r for recs in lib for r in recs]
        for _ in range(60):  # grown around true substrings, like the GPU fuzz
            src = rng.choice(flat)
            at = rng.randint(0, len(src) - 9)
            core = src[at:at + rng.randint(k + 1, 9)].decode()
            i = rng.randrange(len(core))
            kind = rng.randrange(6)
            if kind == 0:
                rx = core[:i] + "(" + core[i] + "|" + rng.choice("ACDEFGHIKLMNPQRSTVWY") + ")" + core[i + 1:]
            elif kind == 1:
                rx = core[:i] + "." + core[i + 1:]
            elif kind == 2:
                rx = core[:i] + core[i] + "?" + core[i:]
            elif kind == 3:
                rx = core[:i] + "[" + core[i] + rng.choice("ACDEFGHIKLMNPQRSTVWY") + "]" + core[i + 1:]
            elif kind == 4:
                rx = core[:i] + core[i] + "{1,2}" + core[i + 1:]
            else:
                rx = "(X" + core[:2] + "|" + core[1] + ")" + core[2:]
            regexes.append(rx)
        for rx in regexes:
            for augment in (False, True):
                got = ob.guarded(lambda: ibf.collect(1, k, 0, rx, augment).tolist())
                if got == "ABORTED":
                    continue
                cases.append({"k": k, "regex": rx, "augment": augment, "words": [int(x) for x in got]})
    json.dump({"seed": 23, "bins": 96, "hash_funs": 3, "fpr": 0.05, "cases": cases},
              open(os.path.join(OUT, "collector_rows.json"), "w"))
    print(len(cases), "cases,", sum(1 for c in cases if any(c["words"])), "with candidates")


def golden_kgraph():
    """k-graphs of the reference (node symbols, two-slot arc map, ranks) for the host layer's construct_kgraph."""
    cases = []
    fixed = [("LMA(E|Q)GLYN", 1, 3, 0), ("LMA(E|Q)GLYN", 1, 6, 0), ("AC*GT", 1, 3, 0), ("A(CD)+EF", 1, 4, 0),
             ("AC{2,4}DE", 1, 3, 0), ("(AC){3}DEF", 1, 3, 0), ("AC?D.EF", 1, 3, 0), ("A[CD]E[^FG]HIK", 1, 3, 0),
             (".*ACDEF.{2}", 1, 4, 0), ("ACGT(A|C)T+G", 0, 4, 0), ("LMAEGLYN", 1, 3, 1), ("L.{0,2}MAC", 1, 3, 0),
             ("AC{0,1}DEF", 1, 3, 0), ("A(C|D)*EFG", 1, 3, 0), ("(A|C)(D|E)(F|G)HIK", 1, 4, 2), ("AB{3}", 1, 3, 0),
             ("A{2,3}(CD){1,2}EF", 1, 3, 0), ("^MK.*LL$", 1, 2, 0), ("C.{2,4}C.{3}[LIVMFYWC].{8}H.{3,5}H", 1, 3, 0),
             ("TATA[AT]A[AT](A|G)", 0, 5, 0), ("GG(ACT)+CC", 0, 6, 0), ("LM[AG]EGLYN", 1, 4, 1), ("W(KR|DE)LY?N", 1, 3, 2)]
    rng = random.Random(5)

    def gen(alpha, depth):
        r = rng.random()
        if depth == 0 or r < 0.45:
            return "".join(rng.choice(alpha) for _ in range(rng.randint(1, 4)))
        if r < 0.6:
            return "(" + gen(alpha, depth - 1) + "|" + gen(alpha, depth - 1) + ")"
        if r < 0.72:
            inner = "(" + gen(alpha, depth - 1) + ")" if rng.random() < 0.5 else rng.choice(alpha)
            return inner + rng.choice(["*", "+", "?", "{2}", "{1,3}", "{0,2}", "{3}", "{2,4}", "{0,1}"])
        if r < 0.82:
            return "[" + "".join(rng.sample(alpha, rng.randint(1, 3))) + "]"
        return gen(alpha, depth - 1) + gen(alpha, depth - 1)

    for _ in range(60):
        mol = rng.choice([0, 1])
        alpha = "ACGT" if mol == 0 else AA.decode()
        fixed.append(("".join(gen(alpha, 2) for _ in range(rng.randint(1, 3))), mol, rng.choice([2, 3, 4, 5, 6]), 0))
    for rx, mol, k, red in fixed:
        got = ob.ref_kgraph(rx, mol, k, red)
        assert got is not None, rx
        cases.append({"regex": rx, "molecule": mol, "k": k, "reduction": red, "postfix": got[0], "nodes": got[1]})
    with open(os.path.join(OUT, "kgraph.json"), "w") as f:
        json.dump(cases, f, separators=(",", ":"))
    print("kgraph.json:", len(cases), "cases")


CLI_QUERIES_AA = ["LMA(E|Q)GLYN", "WKC(DE)+FH", "MK[AG]PLW", "HHC.{2}CW", "AAAAAAA", "^MKTA", "C.{2,4}C.{3}[LIVMFYWC]", "G(LY|QP)NW?K",
                  "KR{2}T", "PW(AC|DE|FG)HI*K"]
CLI_QUERIES_NA = ["TATAAAAG", "GG(ACT)+CC", "ACGTAC(GG|TT)AC", "TTTTTTTTTT", "CAT[AC]GATTACA", "GAT{2,3}ACA"]


def write_bins(rng, directory, prefix, alphabet, n_bins, n_records, length, motifs):
    """Deterministic synthetic FASTA bins (committed under tests/data/): random residues with planted motifs."""
    os.makedirs(directory, exist_ok=True)
    files = []
    for b in range(n_bins):
        path = os.path.join(directory, f"{prefix}_{b}.fa")
        with open(path, "w") as f:
            for r in range(n_records):
                n = rng.randint(length // 2, length)
                seq = bytearray(rand_seq(rng, alphabet, n))
                for m in motifs:
                    if rng.random() < 0.08:
                        at = rng.randint(0, max(0, n - len(m)))
                        seq[at:at + len(m)] = m
                if b == 2 and r == 1:
                    seq = seq[:3]  # shorter than k: "RECORD TOO SHORT"
                f.write(f">{prefix}{b}_{r} synthetic record {r} of bin {b}\n")
                for i in range(0, len(seq), 60):
                    f.write(seq[i:i + 60].decode() + "\n")
        files.append(os.path.relpath(path, ROOT))
    return files


def guarded(fn, *args):
    """Runs fn in a forked child: the reference aborts the process on some inputs (documented in DESIGN.md)."""
    import pickle
    r, w = os.pipe()
    pid = os.fork()
    if pid == 0:
        os.close(r)
        try:
            os.write(w, pickle.dumps(fn(*args)))
        finally:
            os._exit(0)
    os.close(w)
    data = b""
    while True:
        chunk = os.read(r, 1 << 16)
        if not chunk:
            break
        data += chunk
    os.close(r)
    _, status = os.waitpid(pid, 0)
    return pickle.loads(data) if status == 0 and data else None


def golden_cli():
    """End-to-end fixtures for the host layer (tetrex index / query): the reference builds indexes over
    tests/data/* and answers a list of queries; index files, candidate bins and printed lines are committed."""
    rng = random.Random(2024)
    data = os.path.join(ROOT, "tests", "data")
    aa_motifs = [b"LMAEGLYN", b"LMAQGLYN", b"WKCDEDEFH", b"MKAPLW", b"HHCAKCW", b"AAAAAAAA", b"GLYNWK", b"GQPNK", b"KRRT",
                 b"PWDEHIIK", b"CAACLLLV", b"MKTA"]
    na_motifs = [b"TATAAAAG", b"GGACTACTCC", b"CTTTATA", b"ACGTACGGAC", b"GGAGTAGTCC", b"CATAGATTACA", b"TGTAATCGATG", b"GATTTACA"]
    aa_files = write_bins(rng, data, "pep", AA, 12, 6, 240, aa_motifs)
    na_files = write_bins(rng, data, "dna", b"ACGT", 9, 4, 400, na_motifs)
    big_files = write_bins(rng, data, "wide", AA, 70, 2, 120, aa_motifs)  # > 64 user bins: two-level HIBF
    cwd = os.getcwd()
    os.chdir(ROOT)  # indexes store the paths relative to the repo root
    out = []
    try:
        for name, files, mol, k, flavor, red, queries in [
                ("pep_flat_k4", aa_files, 1, 4, 0, 0, CLI_QUERIES_AA), ("pep_hibf_k4", aa_files, 1, 4, 1, 0, CLI_QUERIES_AA),
                ("pep_flat_k3_murphy", aa_files, 1, 3, 0, 1, ["LMAEGLYN", "MKAPLW", "WKCDEDEFH"]),
                ("dna_flat_k6", na_files, 0, 6, 0, 0, CLI_QUERIES_NA), ("dna_hibf_k8", na_files, 0, 8, 1, 0, CLI_QUERIES_NA),
                ("wide_hibf_k4", big_files, 1, 4, 1, 0, CLI_QUERIES_AA[:5])]:
            idx = os.path.join("tests", "golden", "cli", name + ".ibf")
            os.makedirs(os.path.dirname(idx), exist_ok=True)
            ob.ref_index_build(files, k, 0.05, 3, mol, flavor, red, idx)
            for q in queries:
                got = guarded(ob.ref_index_query, idx, q, True)
                if got is None:
                    print("reference aborts on", name, q)
                    out.append(dict(index=idx, files=files, molecule=mol, k=k, flavor=flavor, reduction=red, regex=q,
                                    reference_aborts=True))
                    continue
                words, text = got
                bins = [b for b in range(len(files)) if int(words[b // 64]) >> (b % 64) & 1]
                out.append(dict(index=idx, files=files, molecule=mol, k=k, flavor=flavor, reduction=red, regex=q, bins=bins,
                                stdout=text))
    finally:
        os.chdir(cwd)
    json.dump(out, open(os.path.join(OUT, "cli.json"), "w"), indent=0)
    print("cli.json:", len(out), "queries,", sum(len(c.get("stdout", "").splitlines()) for c in out), "lines")


def golden_c1():
    """BASELINE config 1 on the reference's own five files (data/dna_example_split/sequence{1..5}.fa, 27 bytes each,
    copied verbatim as test DATA to tests/data/dna_example_split/): the flat DNA IBF (k=3, h=3, FPR 0.05: 106 rows x 64
    technical bins = 848 bytes), the nucleotide HIBF, and the PEPTIDE HIBF the reference's CLI really builds for
    `tetrex index -n -k 3`, each queried with README.md:45-51's A(C+|G+)T."""
    import shutil
    src = "/root/reference/data/dna_example_split"
    dst = os.path.join(ROOT, "tests", "data", "dna_example_split")
    os.makedirs(dst, exist_ok=True)
    files = []
    for i in range(1, 6):
        shutil.copyfile(os.path.join(src, f"sequence{i}.fa"), os.path.join(dst, f"sequence{i}.fa"))
        files.append(os.path.relpath(os.path.join(dst, f"sequence{i}.fa"), ROOT))
    cwd = os.getcwd()
    os.chdir(ROOT)
    out = []
    try:
        for name, mol, flavor in [("c1_dna_flat_k3", 0, 0), ("c1_dna_hibf_k3", 0, 1), ("c1_pep_hibf_k3", 1, 1)]:
            idx = os.path.join("tests", "golden", "cli", name + ".ibf")
            ob.ref_index_build(files, 3, 0.05, 3, mol, flavor, 0, idx)
            for q in ["A(C+|G+)T", "ACT", "CCC(T|G)", "GA+TC"]:
                got = guarded(ob.ref_index_query, idx, q, True)
                if got is None:
                    out.append(dict(index=idx, files=files, molecule=mol, k=3, flavor=flavor, reduction=0, regex=q,
                                    reference_aborts=True))
                    continue
                words, text = got
                bins = [b for b in range(len(files)) if int(words[b // 64]) >> (b % 64) & 1]
                out.append(dict(index=idx, files=files, molecule=mol, k=3, flavor=flavor, reduction=0, regex=q, bins=bins,
                                stdout=text))
    finally:
        os.chdir(cwd)
    json.dump(out, open(os.path.join(OUT, "cli_c1.json"), "w"), indent=0)
    print("cli_c1.json:", len(out), "queries,", sum(len(c.get("stdout", "").splitlines()) for c in out), "lines")


def read_records(path):
    recs, cur = [], []
    for line in open(os.path.join(ROOT, path)):
        if line.startswith(">"):
            if cur:
                recs.append("".join(cur))
            cur = []
        else:
            cur.append(line.strip())
    if cur:
        recs.append("".join(cur))
    return recs


def fuzz_regexes(rng, records, alphabet, k, n):
    """Regexes grown around true substrings of the indexed records, so that paths survive the filter and the collector's
    split / absorb / prune rules all fire; one family is built to make items with different shift counts collide on the
    same sub-k-mer key (the symbol that packs to 0 — 'A' — in front of a shared suffix: `(XAC|C)DEF`)."""
    out = []
    zero = "A"
    while len(out) < n:
        rec = rng.choice(records)
        L = rng.randint(k + 1, k + 8)
        if len(rec) < L + 2:
            continue
        at = rng.randint(0, len(rec) - L)
        s = rec[at:at + L]
        fam = rng.random()
        if fam < 0.22:
            # colliding-prefix family
            m = rng.randint(1, 2)
            j = rng.randint(1, max(1, k - 2))
            r = "".join(rng.choice(alphabet) for _ in range(rng.randint(1, 2)))
            alt = [r + zero * j + s[:m], s[:m]]
            if rng.random() < 0.5:
                alt.reverse()
            if rng.random() < 0.3:
                alt.append("".join(rng.choice(alphabet) for _ in range(rng.randint(1, 3))) + s[:m])
            out.append("(" + "|".join(alt) + ")" + s[m:])
            continue
        if fam < 0.30:
            # two joins in a row: paths of three different lengths meet
            a, b = rng.randint(1, 2), rng.randint(1, 2)
            x = "".join(rng.choice(alphabet) for _ in range(rng.randint(1, 3)))
            out.append("(" + x + s[:a] + "|" + s[:a] + ")(" + zero + s[a:a + b] + "|" + s[a:a + b] + ")" + s[a + b:])
            continue
        parts, i = [], 0
        dots = 0
        while i < len(s):
            u = rng.random()
            c = s[i]
            if u < 0.52:
                parts.append(c)
            elif u < 0.64:
                extra = rng.sample(alphabet, rng.randint(1, 3))
                cls = sorted(set(extra + [c]))
                rng.shuffle(cls)
                parts.append("[" + "".join(cls) + "]")
            elif u < 0.76:
                w = rng.randint(1, min(3, len(s) - i))
                true = s[i:i + w]
                other = "".join(rng.choice(alphabet) for _ in range(rng.randint(1, 3)))
                if rng.random() < 0.3:
                    other = other[:1] + zero * rng.randint(1, 2) + true[-1:]
                alts = [true, other]
                rng.shuffle(alts)
                parts.append("(" + "|".join(alts) + ")")
                i += w - 1
            elif u < 0.81:
                parts.append(c + rng.choice(alphabet) + "?")
            elif u < 0.87:
                parts.append(c + rng.choice(["+", "*", "{1,2}", "{1,3}", "{1}"]))
                if parts[-1].endswith("*"):
                    parts[-1] = c + c + "*"
            elif u < 0.91 and i + 1 < len(s) and s[i + 1] == c:
                parts.append(c + rng.choice(["{2}", "{2,3}", "{1,2}", "+"]))
                i += 1
            elif u < 0.96 and dots < 2 and alphabet != "ACGT":
                parts.append(".")
                dots += 1
            elif dots < 1 and alphabet != "ACGT" and i + 2 < len(s):
                parts.append(".{1,2}")
                i += rng.randint(0, 1)
                dots += 2
            else:
                parts.append(c)
            i += 1
        out.append("".join(parts))
    return out


def golden_collector_fuzz():
    """>= 500 seeded regexes through the reference's process_query (preprocess, k-graph, OTFCollector::collect with its
    robin_hood pop order) on the reference-written index files of tests/golden/cli: the candidate-bin sets the host
    collector must reproduce bit for bit (tests/test_collector_vs_ref.py)."""
    rng = random.Random(20251105)
    cli = json.load(open(os.path.join(OUT, "cli.json"))) + json.load(open(os.path.join(OUT, "cli_c1.json")))
    seen, out = set(), []
    cwd = os.getcwd()
    os.chdir(ROOT)
    try:
        for c in cli:
            if c["index"] in seen:
                continue
            seen.add(c["index"])
            records = [r for f in c["files"] for r in read_records(f)]
            alphabet = "ACGT" if c["molecule"] == 0 else AA.decode()
            n = 30 if "c1_" in c["index"] else (40 if c["reduction"] else 90)
            aborted = 0
            for q in fuzz_regexes(rng, records, alphabet, c["k"], n):
                got = guarded(ob.ref_index_query, c["index"], q, False)
                if got is None:
                    aborted += 1
                    continue
                words = got[0]
                bins = [b for b in range(len(c["files"])) if int(words[b // 64]) >> (b % 64) & 1]
                out.append(dict(index=c["index"], regex=q, bins=bins))
            print(c["index"], n, "regexes,", aborted, "aborted in the reference")
    finally:
        os.chdir(cwd)
    with open(os.path.join(OUT, "collector_fuzz.json"), "w") as f:
        json.dump(out, f, indent=0)
    nonempty = sum(1 for c in out if c["bins"])
    print("collector_fuzz.json:", len(out), "cases,", nonempty, "with candidate bins")


def golden_prosite_batch():
    """BASELINE config 3's pattern population (synth.prosite_patterns(1000, 1000)) through RE2 exactly as
    iter_disk_search uses it (POSIX longest, FindAndConsume) over the /100 twin of the Swissprot-shaped collection
    (5700 records, 2.0 M residues) with 10 % of the patterns planted: every hit of every pattern."""
    from tetrex_b200 import synth
    residues, offs, rec_bin, _ = synth.swissprot_like(5700, 1024)
    pats = synth.prosite_patterns(1000, 1000)
    planted = synth.plant_patterns(residues, offs, pats)
    rows = []
    for pi, pat in enumerate(pats):
        h = ob.ref_scan_records_hits(pat, True, residues, offs, 8)
        rows.append(np.concatenate([np.full((len(h), 1), pi, dtype=np.uint32), h], axis=1))
    hits = np.concatenate(rows)
    np.savez_compressed(os.path.join(OUT, "prosite_batch.npz"), hits=hits, planted=np.array(planted, dtype=np.uint32),
                        text_crc=np.uint64(int(np.bitwise_xor.reduce(residues.astype(np.uint64) * np.arange(1, len(residues) + 1, dtype=np.uint64)))))
    per = np.bincount(hits[:, 0], minlength=len(pats))
    print("prosite_batch.npz:", len(hits), "hits,", int((per > 0).sum()), "patterns with hits,", len(planted), "planted")


def per_bin_kmers(residues, offs, rec_bin, n_bins, k=6):
    """What HIBFIndex::populate_index hands to seqan::hibf::config::input_fn (include/index_hibf.h:110-122): every
    peptide k-mer of every record of a bin, duplicates included, through the reference's decomposer."""
    per = [[] for _ in range(n_bins)]
    raw = residues.tobytes()
    for r in range(len(offs) - 1):
        per[int(rec_bin[r])].append(ob.ref_decompose(1, k, 0, raw[int(offs[r]):int(offs[r + 1])]))
    return [np.concatenate(p) if p else np.zeros(0, dtype=np.uint64) for p in per]


def golden_c2_twin_hibf():
    """The /100 twin of BASELINE config 2 (5700 Swissprot-shaped records in 1024 user bins, k = 6, h = 3, FPR 0.05) laid
    out and built by seqan::hibf itself (sketches, DP layout, split + merged bins) exactly as HIBFIndex::populate_index
    calls it; tables + the reference's membership answers for present and absent 6-mers."""
    from tetrex_b200 import synth
    residues, offs, rec_bin, _ = synth.swissprot_like(5700, 1024)
    per = per_bin_kmers(residues, offs, rec_bin, 1024)
    rh = ob.RefHIBF(per, 3, 0.05, 1)
    tb = rh.tables()
    rng = np.random.default_rng(1024)
    allk = np.concatenate(per)
    q = np.concatenate([rng.choice(allk, 6000), rng.integers(0, 1 << 30, 6000, dtype=np.uint64)])
    tb.save(os.path.join(OUT, "hibf_c2twin_tables.npz"))
    np.savez_compressed(os.path.join(OUT, "hibf_c2twin_query.npz"), query=q, hits=rh.membership(q, 8))
    split = sum(int((np.bincount(u[u < 2 ** 63].astype(np.int64)) > 1).sum()) for u in tb.user_bin)
    print("hibf_c2twin:", tb.n_ibfs, "IBFs,", sum(w.nbytes for w in tb.words) / 1e6, "MB,", split, "split user bins")


def golden_cli_ws():
    """kseq's raw-line semantics (include/kseq.h:198-202): sequence lines keep embedded / trailing blanks and tabs, only
    one trailing \\r goes, a line starting with '@' or '+' ends a FASTA record too.  Three small peptide bins written
    with such lines; the reference's record lengths, k-mers and printed start,end coordinates all depend on it."""
    data = os.path.join(ROOT, "tests", "data")
    bodies = [
        b">ws0_a first record\r\nMKTAYIAKQR \r\nQISFVKSHFS\t\r\nRQLEERLGLI\r\n>ws0_b\nLMAEGLYN  WKCDEDEFH\nGLYNWK\n",
        b">ws1_a\tcomment with tab\nHHCAKCW \n\n \nPWDEHIIK\n@ws1_b in the middle\nMKAPLWLMAQGLYN\nKRRT\n",
        b">ws2_a\nCAACLLLV\r\n\r\nGQPNK\t\tMKTA\n>ws2_b\nAAAAAAAA \nLMAEGLYN",
    ]
    files = []
    for b, body in enumerate(bodies):
        path = os.path.join(data, f"ws_{b}.fa")
        open(path, "wb").write(body)
        files.append(os.path.relpath(path, ROOT))
    cwd = os.getcwd()
    os.chdir(ROOT)
    out = []
    try:
        idx = os.path.join("tests", "golden", "cli", "ws_flat_k3.ibf")
        ob.ref_index_build(files, 3, 0.05, 3, 1, 0, 0, idx)
        for q in ["LMA(E|Q)GLYN", "GLYNWK", "MKTA", "KRRT", "W.PWDE", "LLLVGQ", "HFS.RQL", "PNK..MK", "AAAA"]:
            got = guarded(ob.ref_index_query, idx, q, True)
            if got is None:
                print("reference aborts on", q)
                continue
            words, text = got
            bins = [b for b in range(len(files)) if int(words[b // 64]) >> (b % 64) & 1]
            out.append(dict(index=idx, files=files, molecule=1, k=3, flavor=0, reduction=0, regex=q, bins=bins, stdout=text))
    finally:
        os.chdir(cwd)
    json.dump(out, open(os.path.join(OUT, "cli_ws.json"), "w"), indent=0)
    print("cli_ws.json:", len(out), "queries,", sum(len(c["stdout"].splitlines()) for c in out), "lines")
    for c in out:
        print(c["regex"], c["bins"], repr(c["stdout"]))


def write_twin_fasta(directory, residues, offs, rec_bin, n_bins):
    """The /100 twin as n_bins FASTA files (one record per line pair), names as bench.py writes them."""
    os.makedirs(directory, exist_ok=True)
    raw = residues.tobytes()
    files, r = [], 0
    for b in range(n_bins):
        path = os.path.join(directory, f"bin_{b:04d}.fa")
        with open(path, "wb") as f:
            while r < len(rec_bin) and rec_bin[r] == b:
                f.write(b">sp|S%07d|SYN_%d\n" % (r, b))
                f.write(raw[int(offs[r]):int(offs[r + 1])])
                f.write(b"\n")
                r += 1
        files.append(path)
    return files


def guarded_timed(fn, seconds, *args):
    """guarded() with a CPU-time limit in the child: the reference's enumeration explodes on wildcard-rich patterns."""
    import pickle
    import resource
    r, w = os.pipe()
    pid = os.fork()
    if pid == 0:
        os.close(r)
        try:
            resource.setrlimit(resource.RLIMIT_CPU, (seconds, seconds))
            resource.setrlimit(resource.RLIMIT_AS, (6 << 30, 6 << 30))
            os.dup2(os.open(os.devnull, os.O_WRONLY), 2)
            os.write(w, pickle.dumps(fn(*args)))
        finally:
            os._exit(0)
    os.close(w)
    data = b""
    while True:
        chunk = os.read(r, 1 << 16)
        if not chunk:
            break
        data += chunk
    os.close(r)
    _, status = os.waitpid(pid, 0)
    return pickle.loads(data) if status == 0 and data else None


def golden_prosite_filter():
    """BASELINE config 3's FILTER phase: the reference's process_query (k-graph + OTFCollector over a flat peptide IBF,
    k = 6) for the PROSITE-style patterns on the /100 twin, 1024 bins.  The index is built by the reference from FASTA files
    in a scratch directory; only the candidate-bin lists are committed — the GPU test rebuilds the (byte-identical) index
    with `tetrex index -i`.  Patterns whose enumeration does not finish in 3 CPU-seconds in the reference are skipped."""
    import tempfile
    from tetrex_b200 import synth
    residues, offs, rec_bin, _ = synth.swissprot_like(5700, 1024)
    pats = synth.prosite_patterns(1000, 1000)
    synth.plant_patterns(residues, offs, pats)
    d = tempfile.mkdtemp(prefix="trx_twin_")
    files = write_twin_fasta(d, residues, offs, rec_bin, 1024)
    idx = os.path.join(d, "twin_flat.ibf")
    ob.ref_index_build(files, 6, 0.05, 3, 1, 0, 0, idx)
    out, skipped = [], 0
    for pi, pat in enumerate(pats):
        got = guarded_timed(ob.ref_index_query, 3, idx, pat, False)
        if got is None:
            skipped += 1
            continue
        words = got[0]
        out.append(dict(pattern=pi, bins=[b for b in range(1024) if int(words[b // 64]) >> (b % 64) & 1]))
        if len(out) >= 400:
            break
    import shutil
    shutil.rmtree(d, ignore_errors=True)
    with open(os.path.join(OUT, "prosite_filter.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("prosite_filter.json:", len(out), "patterns,", skipped, "skipped (reference > 3 s or aborted),",
          sum(1 for c in out if c["bins"]), "with candidates, mean", sum(len(c["bins"]) for c in out) / len(out), "bins")


CONJ_CASES = [
    ("pep_hibf_k4", "WKC(DE)+FH:PWDEHI*K"), ("pep_hibf_k4", "PWDEHI*K:WKC(DE)+FH"), ("pep_hibf_k4", "WKCDEDEFH:MKAPLW:PWDEHIIK"),
    ("pep_hibf_k4", "PWDEHIIK:MKAPLW:WKCDEDEFH"), ("pep_hibf_k4", "MKAPLW:WKCDEDEFH:PWDEHIIK"), ("pep_hibf_k4", "HHC.{2}CW:KR{2}T"),
    ("pep_hibf_k4", "KRRT:HHCAKCW"), ("pep_hibf_k4", "LMA(E|Q)GLYN:WKCDEDEFH"), ("pep_hibf_k4", "LMAEGLYN:MKTA"),
    ("pep_hibf_k4", "KRRT"), ("pep_hibf_k4", "KRRT:KRRT"), ("pep_hibf_k4", "G(LY|QP)NW?K:CAACLLLV"),
    ("pep_flat_k4", "WKCDEDEFH:PWDEHIIK"), ("pep_flat_k4", "KRRT:PWDEHIIK:AAAA"),
    ("wide_hibf_k4", "AAAAAAAA:CAACLLLV"), ("wide_hibf_k4", "CAACLLLV:AAAAAAAA"), ("wide_hibf_k4", "MKTA:MKAPLW:WKCDEDEFH"),
    ("wide_hibf_k4", "WKCDEDEFH:MKAPLW:MKTA"), ("wide_hibf_k4", "LMAEGLYN:GLYNWK"), ("wide_hibf_k4", "GQPNK:LMAEGLYN:HHCAKCW"),
]


def golden_cli_conj():
    """`tetrex query -c "q1:q2:…"` (src/query.cpp:463-466, include/query.h:191-224,293-326): AND of the queries' filters,
    then RE2::Set per record of the candidate bins; the rule order inside a printed line comes out of RE2's DFA."""
    cwd = os.getcwd()
    os.chdir(ROOT)
    out = []
    try:
        cli = json.load(open(os.path.join(OUT, "cli.json")))
        for name, qs in CONJ_CASES:
            idx = os.path.join("tests", "golden", "cli", name + ".ibf")
            files = next(c["files"] for c in cli if c["index"] == idx)
            got = guarded(ob.ref_index_conjunction, idx, qs)
            if got is None:
                print("reference aborts on", name, qs)
                continue
            words, text = got
            out.append(dict(index=idx, regex=qs, bins=[b for b in range(len(files)) if int(words[b // 64]) >> (b % 64) & 1],
                            stdout=text))
    finally:
        os.chdir(cwd)
    json.dump(out, open(os.path.join(OUT, "cli_conj.json"), "w"), indent=0)
    print("cli_conj.json:", len(out), "cases,", sum(len(c["stdout"].splitlines()) for c in out), "lines")
    for c in out:
        print(c["regex"], c["bins"], repr(c["stdout"][:300]))


def gap_regexes(rng, records, n):
    """Patterns with a wildcard stretch between two literal anchors cut out of real records: what `-a` turns into a Gap
    node.  Left anchors of >= 3 residues and gaps inside the DGram bounds make gapped items; short anchors, short right
    sides, gaps outside the bounds, several wildcard stretches and mismatching right sides cover the other branches of
    gap_procedure / update_gapped (include/otf_collector.h:216-245,290-312)."""
    out = []
    while len(out) < n:
        rec = rng.choice(records)
        la, g, ra = rng.randint(1, 6), rng.randint(1, 9), rng.randint(0, 6)
        if len(rec) < la + g + ra + 12:
            continue
        at = rng.randint(0, len(rec) - (la + g + ra + 10))
        left, right = rec[at:at + la], rec[at + la + g:at + la + g + ra]
        u = rng.random()
        if u < 0.35:
            mid = ".{%d}" % g
        elif u < 0.5:
            mid = "." * g
        elif u < 0.75:
            lo = max(0, g - rng.randint(0, 2))
            mid = ".{%d,%d}" % (lo, g + rng.randint(0, 2))
        elif u < 0.85:
            a = rng.randint(1, g) if g > 1 else 1
            mid = ".{%d}" % a + (".{%d,%d}" % (g - a, g - a + 1) if g - a > 0 else "")
        else:
            mid = "[^P]{%d}" % g
        if right and rng.random() < 0.2:  # a right side that is not in the record
            j = rng.randrange(len(right))
            right = right[:j] + rng.choice(AA.decode().replace(right[j], "")) + right[j + 1:]
        tail = ""
        if rng.random() < 0.25:  # a second stretch further right
            g2 = rng.randint(1, 5)
            st = at + la + g + ra + g2
            tail = ".{%d}" % g2 + rec[st:st + rng.randint(1, 4)]
        out.append(left + mid + right + tail)
    return out


def golden_dgram():
    """`tetrex dindex` + `tetrex query -a [-g]` (include/dGramIndex.h, include/otf_collector.h:216-245,290-312,395-493):
    the reference builds DGram indexes over tests/data (per-bin code counts and the sha256 of the file it writes are kept —
    the files themselves are up to 1.3 MB), then answers gap patterns with augmentation, with and without the DGram
    filter.  A tiny two-bin DGram file is committed whole for the CPU-side reader/writer round trip."""
    import hashlib
    rng = random.Random(424242)
    cwd = os.getcwd()
    os.chdir(ROOT)
    out = dict(builds=[], queries=[])
    tmp = os.path.join("tests", "golden", "_tmp_dgram")
    os.makedirs(tmp, exist_ok=True)
    try:
        cli = json.load(open(os.path.join(OUT, "cli.json")))
        files_of = {c["index"]: c["files"] for c in cli}
        tiny = [os.path.join("tests", "data", "ws_%d.fa" % i) for i in range(2)]
        tiny = [f for f in tiny if os.path.exists(f)] or files_of["tests/golden/cli/pep_flat_k4.ibf"][:2]
        per = ob.ref_dindex_build(tiny, 2, 3, 3, 0.05, os.path.join("tests", "golden", "tiny.dibf"))
        out["tiny"] = dict(files=tiny, min=2, max=3, per_bin=[int(x) for x in per])
        hand = ["WKC.{3}EFH", "WKC...EFH", "WKCD.{3}FH", "PWD.{3,4}IK", "PWDE.{2,3}K", "LMA.GLYN", "LMA.{1}GLYN", "LMAE.{3}N",
                "CAAC.{2}LV", "CAA.{3}LV", "MKA.{2,4}W", "HHC.{3}W", "HHCA.{1,3}", "AAA.{2}AAA", "GQP.K", "WKCD.{4}H",
                "WKC.{5}EFH", "WKC.{3}EFW", "LMA(E|Q).{0,2}LYN", "WKC.{2}.{1,2}EFH", "GLY.{3}[ACDEFGHIKLMNPQRSTVWY]K",
                "KRRT.{3,8}[LIV]", "WKC.{3}EF.{2}[A-Z]", "WKCDEDE.{0}FH", "XKC.{3}EFH", "WKC.{3}XFH", "WKC.{3}BFH"]
        for name, lo, hi in [("pep_flat_k4", 3, 21), ("pep_hibf_k4", 1, 4), ("wide_hibf_k4", 3, 21), ("pep_flat_k3_murphy", 2, 6)]:
            idx = os.path.join("tests", "golden", "cli", name + ".ibf")
            files = files_of[idx]
            dpath = os.path.join(tmp, "%s_%d_%d.dibf" % (name, lo, hi))
            per = ob.ref_dindex_build(files, lo, hi, 3, 0.05, dpath)
            out["builds"].append(dict(name=name, files=files, min=lo, max=hi, per_bin=[int(x) for x in per],
                                      sha256=hashlib.sha256(open(dpath, "rb").read()).hexdigest(), bytes=os.path.getsize(dpath)))
            records = [r for f in files for r in read_records(f)]
            pats = hand + gap_regexes(rng, records, 60 if "wide" not in name else 30)
            for q in pats:
                for dibf in (None, dpath):
                    got = guarded_timed(ob.ref_index_query, 20, idx, q, True, 1, None, True, dibf)
                    if got is None:
                        # the reference dies in iter_disk_search when no k-mer was probed at all (every bit of the start
                        # vector, padding included, survives): keep at least the filter's answer
                        got = guarded_timed(ob.ref_index_query, 20, idx, q, False, 1, None, True, dibf)
                        if got is None:
                            out["queries"].append(dict(index=idx, build=len(out["builds"]) - 1, regex=q, dibf=bool(dibf),
                                                       reference_aborts=True))
                            continue
                        got = (got[0], None)
                    words, text = got
                    out["queries"].append(dict(index=idx, build=len(out["builds"]) - 1, regex=q, dibf=bool(dibf), stdout=text,
                                               bins=[b for b in range(len(files)) if int(words[b // 64]) >> (b % 64) & 1]))
    finally:
        import shutil
        shutil.rmtree(tmp, ignore_errors=True)
        os.chdir(cwd)
    json.dump(out, open(os.path.join(OUT, "dgram.json"), "w"), indent=0)
    qs = out["queries"]
    print("dgram.json:", len(out["builds"]), "builds,", len(qs), "queries,", sum(1 for c in qs if c.get("reference_aborts")),
          "aborts,", sum(1 for c in qs if c.get("bins")), "with candidate bins,", sum(1 for c in qs if c.get("stdout")), "with lines,",
          sum(1 for c in qs if "bins" in c and c["stdout"] is None), "filter only")
    # how often the DGram filter changes the answer, and how often augmentation does
    by = {}
    for c in qs:
        by.setdefault((c["index"], c["regex"]), {})[c["dibf"]] = c.get("bins")
    print("  DGram filter narrows the candidates of", sum(1 for v in by.values() if v.get(True) != v.get(False)), "of", len(by))


def golden_kgraph_aug():
    """k-graphs after OTFCollector::augment (include/otf_collector.h:395-493): catsites from detect_bad_graphs merged by rank,
    one Gap node per distinct path length (slot 2 of a Gap node = its length), Split/Ghost guards when there are several."""
    from tetrex_b200 import synth
    hand = ["AC.DEF", "LMA.{2,4}GLYN", "C.{2,4}C.{3}[LIVMFYWC].{8}H.{3,5}H", "A.?BCD", "A(.|CD)EFG", "AC.{1,2}.DE", "ACD.", "AC.+DE",
            "LM.{0,2}AC", "W.{3}.{2,3}K", "WKC.{3}EFH", "PWD.{3,4}IK", "[^P]{2}ACD", "AC[^DE]F.G", "(AC|.{2})DEF", "MK.{2}(L|.)W"]
    pats = hand + synth.prosite_patterns(50, 77)
    cases = []
    for k in (3, 4, 6):
        for rx in pats:
            got = guarded(ob.ref_kgraph, rx, 1, k, 0, True)
            if got is None:
                continue
            cases.append({"regex": rx, "k": k, "postfix": got[0], "nodes": got[1]})
    with open(os.path.join(OUT, "kgraph_aug.json"), "w") as f:
        json.dump(cases, f, separators=(",", ":"))
    print("kgraph_aug.json:", len(cases), "cases,", sum(1 for c in cases if any(n[0] == 259 for n in c["nodes"])), "with Gap nodes,",
          sum(1 for c in cases if sum(1 for n in c["nodes"] if n[0] == 259) > 1), "with several")


LAYOUT_CASES = [(1, 5, 300, 0, 0, 3, 0.05), (2, 12, 2000, 1, 0, 3, 0.05), (3, 64, 500, 1, 1, 2, 0.05), (4, 65, 500, 0, 1, 3, 0.01),
                (5, 70, 800, 1, 0, 3, 0.05), (6, 130, 300, 1, 1, 3, 0.05), (7, 300, 400, 1, 1, 4, 0.05), (8, 600, 150, 1, 0, 3, 0.05)]


def golden_hibf_layout():
    """seqan::hibf's layout (HyperLogLog estimates, size sort + clustering, DP, lower levels) for seeded collections
    (synth.layout_collection), one thread; and iteration orders of its k-mer set under seeded insert/clear scripts
    (synth.set_script)."""
    import hashlib
    from tetrex_b200 import synth
    out = {"layouts": [], "set_scripts": []}
    for seed, n_bins, mean, ov, sk, hf, fpr in LAYOUT_CASES:
        v, o = synth.layout_collection(seed, n_bins, mean, ov, sk)
        lay = ob.ref_hibf_layout(v, o, hf, fpr, 1)
        out["layouts"].append(dict(seed=seed, n_bins=n_bins, mean=mean, overlap=ov, skew=sk, hash_funs=hf, fpr=fpr, layout=lay))
    for seed in range(100, 116):
        keys, ops = synth.set_script(seed)
        order = ob.ref_flat_set_script(keys, ops)
        out["set_scripts"].append(dict(seed=seed, operations=len(keys), n=len(order), sha256=hashlib.sha256(order.tobytes()).hexdigest(),
                                       head=[str(int(x)) for x in order[:8]]))
    json.dump(out, open(os.path.join(OUT, "hibf_layout.json"), "w"), separators=(",", ":"))
    print("hibf_layout.json:", len(out["layouts"]), "layouts,", len(out["set_scripts"]), "set scripts")


def golden_rh_order():
    """Pop orders of the real robin_hood::unordered_map for key lists shaped like a rank's sub-k-mer keys."""
    rng = random.Random(99)
    cases = []
    for n in [1, 2, 3, 5, 7, 8, 9, 13, 20, 33, 64, 100, 257, 1000]:
        for bits in [6, 15, 25, 60]:
            keys = [rng.getrandbits(bits) for _ in range(n)]
            cases.append(dict(keys=[str(k) for k in keys], order=[int(x) for x in ob.ref_rh_pop_order(keys)]))
    json.dump(cases, open(os.path.join(OUT, "rh_order.json"), "w"))
    print("rh_order.json:", len(cases), "cases")


STEPS = {"decompose": golden_decompose, "ibf": golden_ibf, "hibf": golden_hibf, "regex": golden_regex, "regex_wide": golden_regex_wide, "collector_rows": golden_collector_rows,
         "kgraph": golden_kgraph, "cli": golden_cli, "c1": golden_c1, "collector_fuzz": golden_collector_fuzz, "rh_order": golden_rh_order,
         "prosite_batch": golden_prosite_batch, "c2_twin_hibf": golden_c2_twin_hibf, "cli_ws": golden_cli_ws, "prosite_filter": golden_prosite_filter, "cli_conj": golden_cli_conj, "dgram": golden_dgram, "kgraph_aug": golden_kgraph_aug, "hibf_layout": golden_hibf_layout}

if __name__ == "__main__":
    assert ob.have_ref(), "build oracle/_ref first: make -C oracle ref"
    for step in (sys.argv[1:] or list(STEPS)):
        STEPS[step]()
    print("golden vectors written to", OUT)
