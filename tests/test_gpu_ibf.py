"""GPU parity tests for rows a1–a8: decomposition, emplace, bulk_contains, path algebra, HIBF.
Everything goes through the C ABI (tetrex_b200.api is one ctypes call per method) and is compared
bit for bit with the oracle / golden vectors generated from the compiled reference."""
import json
import os
import random

import numpy as np
import pytest

import tetrex_b200 as t
from oracle import bindings as ob

pytestmark = pytest.mark.gpu

AA = b"ACDEFGHIKLMNPQRSTVWYXBZJOU"


def make_records(rng, alphabet, lengths):
    seqs = [bytes(rng.choice(alphabet) for _ in range(L)) for L in lengths]
    offs = np.zeros(len(seqs) + 1, dtype=np.uint64)
    offs[1:] = np.cumsum([len(s) for s in seqs])
    return seqs, b"".join(seqs), offs


# ------------------------------------------------------------------------------------------ a1 / a2
def test_decompose_golden(ctx, golden_dir):
    g = json.load(open(os.path.join(golden_dir, "decompose.json")))
    for c in g["decompose"]:
        s = c["seq"].encode("latin-1")
        vals, offs = ctx.decompose_records(c["molecule"], c["k"], c["reduction"], s, [0, len(s)])
        assert [int(x) for x in vals] == c["values"], (c["molecule"], c["k"], c["reduction"], len(s))


@pytest.mark.parametrize("mol,k,red,alpha", [(0, 3, 0, b"ACGT"), (0, 15, 0, b"ACGTN"), (0, 16, 0, b"ACGT"),
                                             (0, 20, 0, b"ACGTacgtN"), (0, 32, 0, b"ACGT"), (1, 6, 0, AA),
                                             (1, 1, 0, AA), (1, 12, 0, AA), (1, 6, 1, AA), (1, 5, 2, AA)])
def test_decompose_ragged_records(ctx, mol, k, red, alpha):
    rng = random.Random(k * 7 + mol)
    lengths = [0, 1, k - 1, k, k + 1, 2 * k, 0, 0, 33, 64, 127, 128, 129, 500, k, 0, 1000, 17]
    seqs, text, offs = make_records(rng, alpha, lengths)
    vals, eoffs = ctx.decompose_records(mol, k, red, text, offs)
    want = [ob.decompose(mol, k, red, s) for s in seqs]
    assert [int(x) for x in eoffs] == [0] + list(np.cumsum([len(w) for w in want]))
    assert np.array_equal(vals, np.concatenate(want))


# ------------------------------------------------------------------------------------------ a3 / a4
def test_emplace_values_golden(ctx, golden_dir):
    z = np.load(os.path.join(golden_dir, "ibf.npz"))
    for i, (bins, bs, h) in enumerate(z["shapes"]):
        f = t.IBF(ctx, int(bins), int(bs), int(h))
        f.emplace_values(z[f"values_{i}"], z[f"bins_{i}"])
        assert np.array_equal(f.download(), z[f"words_{i}"]), (bins, bs, h)
        assert np.array_equal(f.bulk_contains(z[f"query_{i}"]), z[f"hits_{i}"])
        f.clear()
        assert not f.download().any()
        f.close()


def test_ibf_geometry_and_errors(ctx):
    f = t.IBF(ctx, 5, 106, 3)
    assert (f.technical_bins, f.bin_words, f.hash_shift, f.n_words) == (64, 1, 57, 106)
    f.close()
    for args in [(0, 10, 3), (5, 0, 3), (5, 10, 0), (5, 10, 6)]:
        with pytest.raises(t.TetrexError) as e:
            t.IBF(ctx, *args)
        assert e.value.code == -1
    with pytest.raises(t.TetrexError):  # 3 shards do not divide 16 words
        t.IBF(ctx, 1024, 100, 3, 0, 3)
    f = t.IBF(ctx, 10, 100, 2)
    with pytest.raises(t.TetrexError) as e:
        f.emplace_values([1, 2], [3, 10])
    assert e.value.code == -7
    f.close()


@pytest.mark.parametrize("mol,k,red,alpha,bins", [(0, 3, 0, b"ACGT", 5), (0, 20, 0, b"ACGT", 70), (0, 16, 0, b"ACGTN", 64),
                                                  (0, 32, 0, b"ACGT", 9), (0, 11, 0, b"ACGT", 130), (1, 6, 0, AA, 200),
                                                  (1, 6, 1, AA, 33), (1, 4, 2, AA, 1024), (1, 12, 0, AA, 65)])
def test_emplace_records_equals_reference_build(ctx, mol, k, red, alpha, bins):
    """IBFIndex::populate_index: size by the largest bin, emplace every emitted k-mer (index_ibf.h:88-131)."""
    rng = random.Random(bins * 31 + k)
    lengths = [rng.choice([0, k - 1, k, k + 1, 40, 200, 333, 1000]) for _ in range(3 * bins)]
    seqs, text, offs = make_records(rng, alpha, lengths)
    rec_bin = np.sort(np.array([rng.randrange(bins) for _ in seqs], dtype=np.uint32))
    emits = [ob.decompose(mol, k, red, s) for s in seqs]
    per_bin = np.zeros(bins, dtype=np.int64)
    for e, b in zip(emits, rec_bin):
        per_bin[b] += len(e)
    assert sum(t.emit_count(mol, k, len(s)) for s in seqs) == per_bin.sum()
    bin_size = max(ob.compute_bitcount(int(per_bin.max()), 0.05), 1)
    want = ob.OracleIBF(bins, bin_size, 3)
    want.emplace(np.concatenate(emits), np.repeat(rec_bin, [len(e) for e in emits]))
    f = t.IBF(ctx, bins, bin_size, 3)
    f.emplace_records(mol, k, red, text, offs, rec_bin)
    got = f.download()
    assert np.array_equal(got, want.words)
    # idempotent: a second pass sets nothing new
    f.emplace_records(mol, k, red, text, offs, rec_bin)
    assert np.array_equal(f.download(), want.words)
    f.close()


def test_dna_example_split_c1(ctx):
    """BASELINE config 1: data/dna_example_split shape — 5 bins × one 16-nt record, k=3, h=3, FPR 0.05."""
    rng = random.Random(1)
    seqs, text, offs = make_records(rng, b"ACGT", [16] * 5)
    bin_size = t.compute_bitcount(17, 0.05)
    assert bin_size == 106
    f = t.IBF(ctx, 5, bin_size, 3)
    f.emplace_records(0, 3, 0, text, offs, np.arange(5, dtype=np.uint32))
    want = ob.OracleIBF(5, 106, 3)
    for b, s in enumerate(seqs):
        e = ob.decompose(0, 3, 0, s)
        assert len(e) == 17
        want.emplace(e, np.full(len(e), b, dtype=np.uint32))
    assert np.array_equal(f.download(), want.words)
    f.close()


@pytest.mark.parametrize("mol,k,alpha", [(0, 20, b"ACGT"), (1, 6, AA)])
def test_emplace_slab_route_equals_oracle(ctx, mol, k, alpha, monkeypatch):
    """K1's slab route (dense copy of the slab's word columns, L2-sliced passes, strided fold into the matrix): forced on
    a small matrix with a 1 MB slice size; records of bins inside the named range, a few OUTSIDE it (those must still land,
    straight in the matrix), a second slab into the same matrix, and the host-buffer entry point which derives the range
    itself — all bit-equal to the oracle's emplace."""
    import torch
    monkeypatch.setenv("TRX_EMPLACE_WINDOW_MB", "1")
    rng = random.Random(k)
    bins, bin_size = 2048, 60013  # 32 words x 60013 rows = 15 MB: more than ten 1 MB slices -> the whole-matrix route gives up
    f = t.IBF(ctx, bins, bin_size, 3)
    want = ob.OracleIBF(bins, bin_size, 3)
    for first, n_bins, strays in [(128, 64, [5, 2047]), (700, 30, []), (0, 64, [64])]:
        seqs, text, offs = make_records(rng, alpha, [rng.randint(k, 400) for _ in range(60)])
        rec_bin = np.array([first + rng.randrange(n_bins) for _ in seqs], dtype=np.uint32)
        for j, b in enumerate(strays):
            rec_bin[j] = b
        d_text = torch.frombuffer(bytearray(text), dtype=torch.uint8).cuda()
        d_offs = torch.from_numpy(offs.view(np.int64)).cuda()
        d_bin = torch.from_numpy(rec_bin.view(np.int32)).cuda()
        launches = ctx.launches
        f.emplace_records_slab_dev(mol, k, 0, d_text.data_ptr(), len(text), d_offs.data_ptr(), d_bin.data_ptr(), len(seqs),
                                   first, n_bins)
        ctx.sync()
        assert ctx.launches - launches == 2, "one pass into the 480 KB slab + the fold"
        want.emplace_records(mol, k, 0, text, offs, rec_bin)
        assert np.array_equal(f.download(), want.words), (first, n_bins)
    # host-buffer entry point: range taken from rec_bin
    seqs, text, offs = make_records(rng, alpha, [rng.randint(k, 300) for _ in range(40)])
    rec_bin = np.array([1000 + rng.randrange(20) for _ in seqs], dtype=np.uint32)
    f.emplace_records(mol, k, 0, text, offs, rec_bin)
    want.emplace_records(mol, k, 0, text, offs, rec_bin)
    assert np.array_equal(f.download(), want.words)
    # and with the route switched off the same bits
    monkeypatch.setenv("TRX_EMPLACE_SLAB", "off")
    g = t.IBF(ctx, bins, bin_size, 3)
    g.emplace_records(mol, k, 0, text, offs, rec_bin)
    only = ob.OracleIBF(bins, bin_size, 3)
    only.emplace_records(mol, k, 0, text, offs, rec_bin)
    assert np.array_equal(g.download(), only.words)
    g.close()
    f.close()


# ------------------------------------------------------------------------------------------------ a5
@pytest.mark.parametrize("bins,bs,h", [(5, 106, 3), (64, 5000, 1), (128, 5000, 2), (320, 3001, 3), (1024, 20011, 3),
                                       (1000, 999, 4), (8192, 4099, 3), (8192 + 64, 1031, 5), (2048, 1 << 16, 3)])
def test_bulk_contains_random(ctx, bins, bs, h):
    rng = np.random.default_rng(bins + bs)
    n = 20000
    v = rng.integers(0, 2 ** 64, n, dtype=np.uint64)
    b = rng.integers(0, bins, n).astype(np.uint32)
    o = ob.OracleIBF(bins, bs, h)
    o.emplace(v, b)
    f = t.IBF(ctx, bins, bs, h)
    f.upload(o.words)
    q = np.concatenate([v[:3000], rng.integers(0, 2 ** 64, 3000, dtype=np.uint64), np.array([0, 1, 2 ** 64 - 1], dtype=np.uint64)])
    got = f.bulk_contains(q)
    assert got.shape == (len(q), o.bin_words)
    assert np.array_equal(got, o.bulk_contains(q, threads=4))
    # no false negatives: the inserted (value, bin) pairs are all reported
    hit = (got[np.arange(3000), b[:3000] >> 6] >> (b[:3000].astype(np.uint64) & np.uint64(63))) & np.uint64(1)
    assert hit.all()
    assert f.bulk_contains(np.zeros(0, dtype=np.uint64)).shape == (0, o.bin_words)
    f.close()


@pytest.mark.parametrize("shards", [2, 4, 8])
def test_sharded_ibf_equals_whole(ctx, shards):
    """§8(e): 64-bin blocks of every row spread over `shards` handles; rows never straddle a shard."""
    bins, bs, h = 1024, 7001, 3
    rng = np.random.default_rng(shards)
    v = rng.integers(0, 2 ** 64, 30000, dtype=np.uint64)
    b = rng.integers(0, bins, 30000).astype(np.uint32)
    o = ob.OracleIBF(bins, bs, h)
    o.emplace(v, b)
    q = np.concatenate([v[:2000], rng.integers(0, 2 ** 64, 2000, dtype=np.uint64)])
    want = o.bulk_contains(q)
    merged = np.zeros(o.bin_words * bs, dtype=np.uint64).reshape(bs, o.bin_words)
    parts = []
    for s in range(shards):
        f = t.IBF(ctx, bins, bs, h, s, shards)
        assert f.shard_words == o.bin_words // shards and f.shard_word_begin == s * f.shard_words
        f.emplace_values(v, b)  # every shard sees all pairs and keeps its own bins
        parts.append(f.bulk_contains(q))
        w = f.download().reshape(bs, o.bin_words)
        merged[:, f.shard_word_begin:f.shard_word_begin + f.shard_words] = w[:, f.shard_word_begin:f.shard_word_begin + f.shard_words]
        f.close()
    assert np.array_equal(np.concatenate(parts, axis=1), want)
    assert np.array_equal(merged.reshape(-1), o.words)


# ------------------------------------------------------------------------------------------------ a6
@pytest.mark.parametrize("bins", [5, 64, 200, 1024, 8192])
def test_path_algebra(ctx, bins):
    bs, h = 2003, 3
    rng = np.random.default_rng(bins)
    v = rng.integers(0, 2 ** 64, 40 * bins, dtype=np.uint64)
    b = rng.integers(0, bins, len(v)).astype(np.uint32)
    o = ob.OracleIBF(bins, bs, h)
    o.emplace(v, b)
    f = t.IBF(ctx, bins, bs, h)
    f.upload(o.words)
    W = o.bin_words
    cap = 300
    p = t.Paths(ctx, W, cap)
    p.fill(0, 200, bins)   # bitvector(bins, true)
    p.fill(200, 100, 0)    # bitvector(bins)
    model = p.read(0, cap)
    full = np.zeros(W, dtype=np.uint64)
    for i in range(bins):
        full[i >> 6] |= np.uint64(1) << np.uint64(i & 63)
    assert (model[:200] == full).all() and not model[200:].any()
    # three rounds of probe+AND on distinct slots, then OR pushes with repeated destinations
    for rnd in range(3):
        slots = rng.permutation(200)[:150].astype(np.uint32)
        kmers = rng.choice(v, 150)
        alive = p.step(f, kmers, slots)
        hits = o.bulk_contains(kmers)
        model[slots] &= hits
        assert np.array_equal(alive, model[slots].any(axis=1).astype(np.uint8))
        assert np.array_equal(p.read(0, cap), model)
    dst = rng.integers(200, 300, 400).astype(np.uint32)
    src = rng.integers(0, 200, 400).astype(np.uint32)
    p.or_(dst, src)
    for d, s in zip(dst, src):
        model[d] |= model[s]
    assert np.array_equal(p.read(0, cap), model)
    with pytest.raises(t.TetrexError):
        p.or_([cap], [0])
    p.close()
    f.close()


# ------------------------------------------------------------------------------------------------ a8
@pytest.mark.parametrize("name", ["hibf_small", "hibf_100", "hibf_300", "hibf_c2twin"])
def test_hibf_membership_golden(ctx, golden_dir, name):
    """hibf_c2twin = the /100 twin of BASELINE config 2 (1024 user bins, peptide 6-mers) laid out by seqan::hibf itself:
    65 IBFs, 1007 of the 1024 user bins split over several technical bins."""
    tb = ob.HIBFTables.load(os.path.join(golden_dir, name + "_tables.npz"))
    q = np.load(os.path.join(golden_dir, name + "_query.npz"))
    h = t.HIBF.from_tables(ctx, tb)
    got, st = h.membership(q["query"], stats=True)
    assert np.array_equal(got, q["hits"])
    assert st[0] >= len(q["query"]) and st[1] >= 1
    # path AND against device-resident HIBF hit vectors (the HIBF route of update_path)
    import ctypes as C
    from tetrex_b200._abi import check, lib
    n = 64
    d_k, d_o = C.c_void_p(), C.c_void_p()
    check(lib().trx_dev_alloc(ctx._h, n * 8, C.byref(d_k)))
    check(lib().trx_dev_alloc(ctx._h, n * h.out_words * 8, C.byref(d_o)))
    kq = np.ascontiguousarray(q["query"][:n])
    check(lib().trx_copy_h2d(ctx._h, d_k, kq.ctypes.data, n * 8))
    h.membership_dev(d_k.value, n, d_o.value)
    p = t.Paths(ctx, h.out_words, n)
    p.fill(0, n, tb.n_user_bins)
    alive = p.and_hits(d_o.value, np.arange(n), np.arange(n))
    assert np.array_equal(p.read(0, n), q["hits"][:n])
    assert np.array_equal(alive, q["hits"][:n].any(axis=1).astype(np.uint8))
    check(lib().trx_dev_free(ctx._h, d_k))
    check(lib().trx_dev_free(ctx._h, d_o))
    p.close()
    h.close()


def test_hibf_large_batch_matches_oracle(ctx, golden_dir):
    tb = ob.HIBFTables.load(os.path.join(golden_dir, "hibf_300_tables.npz"))
    rng = np.random.default_rng(1)
    q = rng.integers(0, 2 ** 40, 300000, dtype=np.uint64)
    h = t.HIBF.from_tables(ctx, tb)
    assert np.array_equal(h.membership(q), tb.membership(q, threads=8))
    h.close()


@pytest.mark.parametrize("route", ["windows", "frontier"])
def test_hibf_seqan_layout_c2_twin_large_batch(ctx, golden_dir, route, monkeypatch):
    """K3's bench route (root + L2 windows + merge: batches > 2048 k-mers) and the frontier route on the seqan-built
    1024-user-bin layout, 400 000 k-mers (half present), against the C port — which the golden file pins to seqan's own
    membership_for on this very layout (test_oracle_golden / hibf_c2twin_query.npz)."""
    if route == "frontier":
        monkeypatch.setenv("TRX_HIBF_FUSED", "off")
    tb = ob.HIBFTables.load(os.path.join(golden_dir, "hibf_c2twin_tables.npz"))
    gq = np.load(os.path.join(golden_dir, "hibf_c2twin_query.npz"))
    assert np.array_equal(tb.membership(gq["query"], threads=8), gq["hits"])  # the port agrees with seqan here
    rng = np.random.default_rng(5)
    q = np.concatenate([rng.choice(gq["query"][:6000], 200000), rng.integers(0, 1 << 30, 200000, dtype=np.uint64)])
    rng.shuffle(q)
    h = t.HIBF.from_tables(ctx, tb)
    got, st = h.membership(q, stats=True)
    assert np.array_equal(got, tb.membership(q, threads=8))
    assert st[0] > len(q)  # merged bins were entered
    h.close()


# ---- shapes seqan's layout does not produce at the golden sizes: three levels, IBFs wider than 64 technical bins,
# per-IBF hash counts, user bins scattered over the whole result row, split bins — every kernel route of K3
MERGED = np.uint64(0xFFFFFFFFFFFFFFFF)
DELETED = np.uint64(0xFFFFFFFFFFFFFFFE)


def synth_hibf(seed, n_user, root_tb, child_tb, grand_tb, hash_funs, scatter, values_per_bin=150):
    """Random HIBF: root (root_tb technical bins) → children (child_tb) → grandchildren (grand_tb).  user bins are
    dealt to the three levels; `scatter` permutes user-bin ids so that neighbouring technical bins report into
    distant words of the result row.  Bits are set with the oracle's emplace."""
    rng = np.random.default_rng(seed)
    ids = rng.permutation(n_user) if scatter else np.arange(n_user)
    vals = {int(u): rng.integers(0, 2 ** 40, values_per_bin, dtype=np.uint64) for u in range(n_user)}
    it = iter(int(u) for u in ids)
    ibfs = []  # dicts: tb, h, table[(kind, x)], filled later

    def new_ibf(tb, h):
        ibfs.append(dict(tb=tb, h=h, table=[("deleted", 0)] * tb))
        return len(ibfs) - 1

    def take(k):
        out = []
        for _ in range(k):
            try:
                out.append(next(it))
            except StopIteration:
                break
        return out

    root = new_ibf(root_tb, hash_funs[0])
    n_children = min(6, root_tb // 4)
    for c in range(n_children):
        child = new_ibf(child_tb, hash_funs[1 % len(hash_funs)] if c % 2 else hash_funs[0])
        ibfs[root]["table"][2 * c] = ("merged", child)
        if c < 2:  # two children carry a third level
            g = new_ibf(grand_tb, hash_funs[2 % len(hash_funs)])
            ibfs[child]["table"][0] = ("merged", g)
            for tb_i, u in enumerate(take(grand_tb // 2)):
                ibfs[g]["table"][2 * tb_i] = ("user", u)
                ibfs[g]["table"][2 * tb_i + 1] = ("user", u)  # split bin: two technical bins, one user bin
        for tb_i, u in enumerate(take(child_tb // 2 - 1)):
            ibfs[child]["table"][child_tb - 1 - tb_i] = ("user", u)
    for tb_i, u in enumerate(take(root_tb // 4)):
        ibfs[root]["table"][root_tb - 1 - tb_i] = ("user", u)
    placed = {x for d in ibfs for kind, x in d["table"] if kind == "user"}

    def subtree_users(i):
        us = []
        for kind, x in ibfs[i]["table"]:
            if kind == "user":
                us.append(x)
            elif kind == "merged":
                us += subtree_users(x)
        return sorted(set(us))

    words, nxt, ub, bins, bsz, hf = [], [], [], [], [], []
    for i, d in enumerate(ibfs):
        per_tb = []
        for t_i, (kind, x) in enumerate(d["table"]):
            if kind == "user":
                v = vals[x]
                # a split bin stores alternating halves in its technical bins
                twins = [j for j, (k2, x2) in enumerate(d["table"]) if k2 == "user" and x2 == x]
                per_tb.append(v[twins.index(t_i)::len(twins)])
            elif kind == "merged":
                per_tb.append(np.concatenate([vals[u] for u in subtree_users(x)]) if subtree_users(x) else np.zeros(0, np.uint64))
            else:
                per_tb.append(np.zeros(0, np.uint64))
        size = max(64, int(max(len(v) for v in per_tb) * 5) | 1)
        f = ob.OracleIBF(d["tb"], size, d["h"])
        for t_i, v in enumerate(per_tb):
            if len(v):
                f.emplace(v, np.full(len(v), t_i, dtype=np.uint32))
        words.append(f.words)
        bins.append(d["tb"])
        bsz.append(size)
        hf.append(d["h"])
        nxt.append(np.array([x if kind == "merged" else i for kind, x in d["table"]], dtype=np.uint64))
        ub.append(np.array([MERGED if kind == "merged" else (np.uint64(x) if kind == "user" else DELETED) for kind, x in d["table"]],
                           dtype=np.uint64))
    tb = ob.HIBFTables(n_user, bins, bsz, hf, words, nxt, ub)
    present = np.concatenate([vals[u][:40] for u in sorted(placed)])
    query = np.concatenate([present, rng.integers(0, 2 ** 40, 3000, dtype=np.uint64)])
    rng.shuffle(query)
    return tb, query, vals, placed


@pytest.mark.parametrize("case", [
    dict(seed=1, n_user=400, root_tb=64, child_tb=64, grand_tb=64, hash_funs=[3], scatter=False),    # FAST shape, 3 levels
    dict(seed=2, n_user=900, root_tb=128, child_tb=192, grand_tb=64, hash_funs=[3], scatter=False),  # multi-word IBFs
    dict(seed=3, n_user=500, root_tb=64, child_tb=64, grand_tb=64, hash_funs=[2, 4, 3], scatter=False),  # per-IBF hash counts
    dict(seed=4, n_user=4000, root_tb=64, child_tb=64, grand_tb=64, hash_funs=[3], scatter=True),    # spans too wide: frontier route
    dict(seed=5, n_user=70000, root_tb=64, child_tb=128, grand_tb=64, hash_funs=[2], scatter=False),  # > 65535 user bins
])
def test_hibf_synthetic_shapes(ctx, case, monkeypatch):
    tb, query, vals, placed = synth_hibf(**case)
    want = tb.membership(query, threads=4)
    # no false negatives by construction: every inserted value reports its user bin
    some = sorted(placed)[:: max(1, len(placed) // 50)]
    probe = np.array([vals[u][7] for u in some], dtype=np.uint64)
    hits = tb.membership(probe)
    assert all((int(hits[j][u >> 6]) >> (u & 63)) & 1 for j, u in enumerate(some))
    if ob.have_ref() and case["n_user"] <= 4000:
        assert np.array_equal(ob.RefHIBF.from_tables(tb).membership(query[:500], 1), want[:500])  # the reference's own recursion
    for fused in ("on", "off", "auto"):
        if fused == "auto":  # default: batches of <= 2048 k-mers take the unsorted frontier route, larger ones the windowed one
            monkeypatch.delenv("TRX_HIBF_FUSED")
        else:
            monkeypatch.setenv("TRX_HIBF_FUSED", fused)
        h = t.HIBF.from_tables(ctx, tb)
        got, st = h.membership(query, stats=True)
        assert np.array_equal(got, want), (case, fused)
        assert st[1] == 3
        one = h.membership(query[:1])  # a single k-mer (collector batches are tiny)
        assert np.array_equal(one, want[:1])
        for m in (7, 2048, 2049):
            assert np.array_equal(h.membership(query[:m]), want[:m]), (case, fused, m)
        if fused == "auto":
            # membership_for's own result form: counts + ascending user-bin ids, compacted on the device
            bits = np.unpackbits(want.view(np.uint8), axis=1, bitorder="little")
            want_counts = bits.sum(axis=1).astype(np.uint32)
            want_ids = np.nonzero(bits)[1]
            for id_bytes in ((2, 4) if case["n_user"] <= 65536 else (4,)):
                counts, ids = h.membership_ids(query, id_bytes=id_bytes)
                assert np.array_equal(counts, want_counts) and np.array_equal(ids.astype(np.int64), want_ids), (case, id_bytes)
            if case["n_user"] > 65536:
                with pytest.raises(t.TetrexError):
                    h.membership_ids(query, id_bytes=2)
            with pytest.raises(t.TetrexError):  # too little room: an error, not a truncated answer
                h.membership_ids(query, id_bytes=4, cap=max(int(want_counts.sum()) - 1, 0))
            c0, i0 = h.membership_ids(query[:0], id_bytes=4)
            assert len(c0) == 0 and len(i0) == 0
        h.close()
