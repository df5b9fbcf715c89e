"""GPU: `tetrex index` without -i builds the HIBF the reference builds (row a3): seqan::hibf's sketches, layout and
hierarchical build restated in host/hibf_layout.cpp + host/hibf_build.cpp, bits set by K1.  The reference numbers its IBFs
in an order shuffled with std::random_device, so files are compared through `tetrex hibf-canon`, a depth-first dump of
geometry, bit-matrix hashes and bin tables that does not depend on the numbering: equal dumps = the same answer (false
positives included) for every membership query.  Reference side: the golden index files of tests/golden/cli (written by
the compiled reference) and, for the 1024-bin twin of config 2, the compiled reference run live (oracle/_ref)."""
import json
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TETREX = os.path.join(ROOT, "tetrex_b200", "bin", "tetrex")
CASES = json.load(open(os.path.join(ROOT, "tests", "golden", "cli.json")))


def run(args, **kw):
    return subprocess.run([TETREX] + args, capture_output=True, text=True, cwd=ROOT, timeout=900, **kw)


def canon(path):
    out = run(["hibf-canon", path])
    assert out.returncode == 0, out.stderr
    return out.stdout


@pytest.mark.parametrize("name", ["pep_hibf_k4", "dna_hibf_k8", "wide_hibf_k4"])
def test_hibf_built_here_equals_reference_written_index(name, tmp_path):
    c = next(c for c in CASES if c["index"].endswith(name + ".ibf"))
    lst = tmp_path / "bins.lst"
    lst.write_text("\n".join(c["files"]) + "\n")
    args = ["index", "-k", str(c["k"])] + (["-n"] if c["molecule"] == 0 else [])
    out = run(args + [str(tmp_path / name), str(lst)])
    assert out.returncode == 0, out.stderr
    ours, ref = canon(str(tmp_path / (name + ".ibf"))), canon(c["index"])
    assert ours == ref
    assert ours.count("ibf bins") >= (2 if name.startswith("wide") else 1)


def test_config2_twin_hibf_equals_live_reference_build(tmp_path):
    """1024 user bins (the /100 twin of BASELINE config 2: 5700 Swissprot-shaped records, k = 6, h = 3, FPR 0.05): two
    levels, every user bin split — layout by HyperLogLog sketches, union estimates and clustering, sizes from distinct
    k-mer counts, split bins cut along the reference's hash-set iteration order."""
    from oracle import bindings as ob
    from tetrex_b200 import synth
    if not ob.have_ref():
        pytest.skip("oracle/_ref not built")
    residues, offs, rec_bin, _ = synth.swissprot_like(5700, 1024)
    raw = np.asarray(residues, dtype=np.uint8).tobytes()
    files, r = [], 0
    for b in range(1024):
        path = str(tmp_path / f"bin_{b:04d}.fa")
        with open(path, "wb") as f:
            while r < len(rec_bin) and rec_bin[r] == b:
                f.write(b">sp|S%07d|SYN_%d\n" % (r, b))
                f.write(raw[int(offs[r]):int(offs[r + 1])])
                f.write(b"\n")
                r += 1
        files.append(path)
    lst = tmp_path / "bins.lst"
    lst.write_text("\n".join(files) + "\n")
    import time
    t0 = time.perf_counter()
    out = run(["index", "-k", "6", str(tmp_path / "ours"), str(lst)])
    assert out.returncode == 0, out.stderr
    t1 = time.perf_counter()
    ob.ref_index_build(files, 6, 0.05, 3, 1, 1, 0, str(tmp_path / "ref.ibf"), threads=1)
    print(f"tetrex index (HIBF, 1024 bins): {t1 - t0:.1f} s; reference build, 1 thread: {time.perf_counter() - t1:.1f} s")
    ours, ref = canon(str(tmp_path / "ours.ibf")), canon(str(tmp_path / "ref.ibf"))
    assert ours.count("ibf bins") > 10 and " merged" in ours
    assert ours == ref
    # and the query path on top of it: same candidate bins as the reference's process_query on ITS file
    for q in ["LMA(E|Q)GLYN", "WKC(DE)+FH", "MK[AG]PLW"]:
        got = [int(x) for x in run(["bins", str(tmp_path / "ours.ibf"), q]).stdout.split()]
        words, _ = ob.guarded(ob.ref_index_query, str(tmp_path / "ref.ibf"), q, False)
        want = [b for b in range(1024) if int(words[b // 64]) >> (b % 64) & 1]
        assert got == want, q


def test_device_sketches_equal_hyperloglog_of_the_emitted_kmers(ctx):
    """trx_hll_sketch_records (K1's decomposition with a HyperLogLog sink) against hyperloglog::add
    (lib/hibf/src/sketch/hyperloglog.cpp:54-70) applied to the oracle's decompose_record output: registers bit-exact, DNA
    (wrap-around k-mers included) and peptides, ragged records, an empty bin."""
    import tetrex_b200 as t
    from oracle import bindings as ob
    rng = np.random.default_rng(8)

    def sketch(values, bits=12):
        reg = np.zeros(1 << bits, dtype=np.uint8)
        for v in values:
            p = int(v) * 0x9E3779B97F4A7C15
            h = (p & 0xFFFFFFFFFFFFFFFF) ^ (p >> 64)
            rest = ((h << bits) & 0xFFFFFFFFFFFFFFFF) | ((1 << bits) - 1)
            rank = 64 - rest.bit_length() + 1
            idx = h >> (64 - bits)
            reg[idx] = max(reg[idx], rank)
        return reg
    for mol, k, red, alpha in [(t.MOL_AA, 6, 0, b"ACDEFGHIKLMNPQRSTVWY"), (t.MOL_NA, 8, 0, b"ACGT"), (t.MOL_AA, 4, 1, b"ACDEFGHIKLMNPQRSTVWY")]:
        recs = [bytes(rng.choice(np.frombuffer(alpha, dtype=np.uint8), n)) for n in [300, 3, 0, 1500, 77, 2000, 9, 640]]
        rec_bin = np.array([0, 0, 1, 1, 3, 3, 3, 4], dtype=np.uint32)  # bin 2 stays empty
        offs = np.concatenate([[0], np.cumsum([len(r) for r in recs])]).astype(np.uint64)
        got = t.hll_sketch_records(ctx, mol, k, red, b"".join(recs), offs, rec_bin, 5)
        for b in range(5):
            vals = np.concatenate([ob.decompose(mol, k, red, r) for r, rb in zip(recs, rec_bin) if rb == b and len(r) >= k] or
                                  [np.zeros(0, dtype=np.uint64)])
            assert np.array_equal(got[b], sketch(vals)), (mol, k, b)
        assert not got[2].any() and got[0].any()
