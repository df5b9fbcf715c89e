"""Seeded synthetic collections for bench.py and the full-size tests.

These are DATA generators plus the host-side planning of a Swissprot-shaped HIBF (which user bin goes
to which technical bin of which IBF, and how large each IBF is, using hibf's own sizing formulae:
lib/hibf/include/hibf/build/bin_size_in_bits.hpp:32-42, lib/hibf/src/layout/compute_fpr_correction.cpp:
18-37, lib/hibf/src/build/construct_ibf.cpp:34-45).  Every bit of every IBF is set on the GPU through
the C ABI (trx_emplace_records); nothing here hashes or probes.
"""
from __future__ import annotations

import math

import numpy as np

from . import api

# Swissprot amino-acid background frequencies (UniProtKB/Swiss-Prot release statistics), order below
AA_LETTERS = b"ACDEFGHIKLMNPQRSTVWY"
AA_FREQ = np.array([8.25, 1.38, 5.46, 6.71, 3.86, 7.07, 2.27, 5.91, 5.80, 9.65, 2.41, 4.06, 4.74, 3.93, 5.53, 6.65,
                    5.36, 6.85, 1.10, 2.92])


def swissprot_like(n_seqs: int = 570_000, n_bins: int = 1024, seed: int = 20250346, plant: bool = True):
    """Peptide records, log-normal lengths (mu 5.6, sigma 0.75, clamped to 20…5000), i.i.d. residues from the
    Swissprot background + 0.01 % X; record i goes to bin i % n_bins (fasta-splitter style round robin); the
    returned arrays are grouped by ascending bin.  With plant=True the README query's answer is planted:
    LMAQGLYN in 23 records of bin 346 and LMAEGLYN in 1 record of bin 811 (README.md:84-108)."""
    rng = np.random.default_rng(seed)
    lengths = np.clip(rng.lognormal(5.6, 0.75, n_seqs), 20, 5000).astype(np.int64)
    rec_bin = (np.arange(n_seqs) % n_bins).astype(np.uint32)
    order = np.argsort(rec_bin, kind="stable")
    lengths, rec_bin = lengths[order], rec_bin[order]
    offs = np.zeros(n_seqs + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(lengths)
    total = int(offs[-1])
    p = AA_FREQ / AA_FREQ.sum()
    cdf = np.cumsum(p)
    letters = np.frombuffer(AA_LETTERS, dtype=np.uint8)
    u = rng.random(total, dtype=np.float32)
    residues = letters[np.minimum(np.searchsorted(cdf, u), 19)]
    x_mask = rng.random(total, dtype=np.float32) < 1e-4
    residues[x_mask] = ord("X")
    planted = []
    if plant:
        for b, motif, count in [(346, b"LMAQGLYN", 23), (811, b"LMAEGLYN", 1)]:
            if b >= n_bins:
                continue
            recs = np.nonzero(rec_bin == b)[0][:count]
            for r in recs:
                pos = int(offs[r]) + 5
                residues[pos:pos + len(motif)] = np.frombuffer(motif, dtype=np.uint8)
                planted.append((b, int(r), 5, 5 + len(motif)))
    return residues, offs, rec_bin, planted


def prosite_patterns(n: int = 1000, seed: int = 1000, planted_fraction: float = 0.0):
    """PROSITE-style patterns as regexes (SURVEY.md §8(d) "C3 patterns", grammar of utils/tetrex_tools:130-153):
    6-14 elements, each a literal (60 %), a class of 2-5 letters (25 %), x (10 %), x(2,4) or class(2,4) (5 %)."""
    import random
    rng = random.Random(seed)
    aa = "ACDEFGHIKLMNPQRSTVWY"
    pats = []
    for _ in range(n):
        parts = []
        for _e in range(rng.randint(6, 14)):
            u = rng.random()
            if u < 0.60:
                parts.append(rng.choice(aa))
            elif u < 0.85:
                parts.append("[" + "".join(sorted(rng.sample(aa, rng.randint(2, 5)))) + "]")
            elif u < 0.95:
                parts.append(".")
            elif rng.random() < 0.5:
                parts.append(".{2,4}")
            else:
                parts.append("[" + "".join(sorted(rng.sample(aa, rng.randint(2, 5)))) + "]{2,4}")
        # the reference trims leading/trailing wildcards before building the k-graph; a verify regex may keep them
        pats.append("".join(parts))
    return pats


def pattern_instance(pattern: str, rng) -> bytes:
    """One string the PROSITE-style pattern (grammar of prosite_patterns) matches: a letter per literal, a member per
    class, a random residue per '.', the lower or upper count for a {m,n} repeat."""
    aa = "ACDEFGHIKLMNPQRSTVWY"
    out, i = [], 0
    while i < len(pattern):
        c = pattern[i]
        if c == "[":
            j = pattern.index("]", i)
            choices, i = pattern[i + 1:j], j + 1
        elif c == ".":
            choices, i = aa, i + 1
        else:
            choices, i = c, i + 1
        reps = 1
        if i < len(pattern) and pattern[i] == "{":
            j = pattern.index("}", i)
            lo, _, hi = pattern[i + 1:j].partition(",")
            reps = rng.randint(int(lo), int(hi or lo))
            i = j + 1
        out.extend(rng.choice(choices) for _ in range(reps))
    return "".join(out).encode()


def plant_patterns(residues, rec_offsets, patterns, seed: int = 1000, fraction: float = 0.10, max_records: int = 50):
    """SURVEY.md 8(d) "C3 patterns": 10 % of the patterns are planted into 1-50 random records each (in place, at a random
    offset inside the record).  Returns the indices of the planted patterns."""
    import random
    rng = random.Random(seed + 1)
    n_rec = len(rec_offsets) - 1
    planted = []
    for pi, pat in enumerate(patterns):
        if rng.random() >= fraction:
            continue
        planted.append(pi)
        for _ in range(rng.randint(1, max_records)):
            inst = np.frombuffer(pattern_instance(pat, rng), dtype=np.uint8)
            r = rng.randrange(n_rec)
            lo, hi = int(rec_offsets[r]), int(rec_offsets[r + 1])
            if hi - lo < len(inst):
                continue
            at = rng.randint(lo, hi - len(inst))
            residues[at:at + len(inst)] = inst
    return planted


def dna_like(n_bins: int, records_per_bin: int, record_len: int, seed: int = 8192):
    rng = np.random.default_rng(seed)
    n = n_bins * records_per_bin
    letters = np.frombuffer(b"ACGT", dtype=np.uint8)
    residues = letters[rng.integers(0, 4, n * record_len, dtype=np.uint8)]
    offs = (np.arange(n + 1, dtype=np.uint64) * np.uint64(record_len))
    rec_bin = np.repeat(np.arange(n_bins, dtype=np.uint32), records_per_bin)
    return residues, offs, rec_bin


def bin_size_in_bits(fpr: float, hash_funs: int, elements: int) -> int:
    return int(math.ceil(-(elements * hash_funs) / math.log(1 - math.exp(math.log(fpr) / hash_funs))))


def fpr_correction(fpr: float, hash_funs: int, split: int) -> float:
    if split <= 1:
        return 1.0
    numerator = math.log1p(-math.exp(math.log(fpr) / hash_funs))
    log_target = math.log1p(-math.exp(math.log1p(-fpr) / split))
    return numerator / math.log1p(-math.exp(log_target / hash_funs))


class HibfPlan:
    """Two-level layout for n_bins equally sized user bins with t_max = 64 (hibf's default for 1024 bins,
    lib/hibf/src/config.cpp:100-103): root IBF = 64 merged bins of n_bins/64 user bins each; one lower IBF per
    merged bin in which every user bin is split over 64/(n_bins/64) technical bins (record-wise)."""

    def __init__(self, n_bins: int, t_max: int = 64):
        assert n_bins <= t_max or (n_bins % t_max == 0 and t_max % (n_bins // t_max) == 0)
        self.n_bins, self.t_max = n_bins, t_max
        self.flat = n_bins <= t_max
        self.per_merged = 1 if self.flat else n_bins // t_max
        self.split = 1 if self.flat else t_max // self.per_merged


def build_swissprot_hibf(ibf_factory, residues, rec_offsets, rec_bin, n_bins: int, k: int = 6, hash_funs: int = 3,
                         fpr: float = 0.05, relaxed_fpr: float = 0.3):
    """Returns tables (dict of lists) of a Swissprot-shaped HIBF.  ibf_factory(bins, bin_size, hash_funs) must
    return an object with emplace_records(), download() and close(): ``lambda *a: api.IBF(ctx, *a)`` sets every
    bit on the GPU; bench.py's reference arm passes a CPU-oracle adapter instead so that arm never touches CUDA."""
    plan = HibfPlan(n_bins)
    assert not plan.flat
    lens = (rec_offsets[1:] - rec_offsets[:-1]).astype(np.int64)
    emits = np.maximum(lens - k + 1, 0)
    per_ub = np.bincount(rec_bin, weights=emits, minlength=n_bins).astype(np.int64)
    t_max, pm, split = plan.t_max, plan.per_merged, plan.split
    merged = np.add.reduceat(per_ub, np.arange(0, n_bins, pm))
    tables = dict(n_user_bins=n_bins, bins=[], bin_size=[], hash_funs=[], words=[], next_ibf=[], user_bin=[])
    MERGED = np.uint64(2 ** 64 - 1)

    # root: technical bin j = merged bin over user bins [j*pm, (j+1)*pm); sized for the fullest at the relaxed FPR
    root_size = bin_size_in_bits(relaxed_fpr, hash_funs, int(merged.max()))
    root = ibf_factory(t_max, root_size, hash_funs)
    root.emplace_records(api.MOL_AA, k, api.RED_BASE, residues, rec_offsets, (rec_bin // pm).astype(np.uint32))
    tables["bins"].append(t_max)
    tables["bin_size"].append(root_size)
    tables["hash_funs"].append(hash_funs)
    tables["words"].append(root.download())
    tables["next_ibf"].append(np.arange(1, t_max + 1, dtype=np.uint64))
    tables["user_bin"].append(np.full(t_max, MERGED, dtype=np.uint64))
    root.close()

    # lower IBFs: user bin u of merged bin j occupies technical bins [(u % pm)*split, +split); record r of that
    # user bin goes to sub-bin (r's rank inside the bin) % split
    first_rec = np.searchsorted(rec_bin, np.arange(n_bins + 1))
    rank_in_bin = np.arange(len(rec_bin)) - first_rec[rec_bin]
    tb_lower = ((rec_bin % pm) * split + rank_in_bin % split).astype(np.uint32)
    corr = fpr_correction(fpr, hash_funs, split)
    for j in range(t_max):
        r0, r1 = int(first_rec[j * pm]), int(first_rec[(j + 1) * pm])
        ub_emits = per_ub[j * pm:(j + 1) * pm]
        per_tb = int(math.ceil(ub_emits.max() / split))
        size = int(math.ceil(bin_size_in_bits(fpr, hash_funs, per_tb) * corr))
        f = ibf_factory(t_max, size, hash_funs)
        base = rec_offsets[r0]
        f.emplace_records(api.MOL_AA, k, api.RED_BASE, residues[int(base):int(rec_offsets[r1])],
                          rec_offsets[r0:r1 + 1] - base, tb_lower[r0:r1])
        tables["bins"].append(t_max)
        tables["bin_size"].append(size)
        tables["hash_funs"].append(hash_funs)
        tables["words"].append(f.download())
        tables["next_ibf"].append(np.full(t_max, j + 1, dtype=np.uint64))  # unused for user bins (hibf keeps the own index)
        tables["user_bin"].append((j * pm + np.arange(t_max) // split).astype(np.uint64))
        f.close()
    return tables


def layout_collection(seed: int, n_bins: int, mean: int, overlap: int, skew: int):
    """Seeded user bins as k-mer lists (values, offsets): sizes uniform or heavy-tailed, neighbours optionally sharing half
    their k-mers — input of the HIBF layout tests (tests/golden/hibf_layout.json)."""
    rng = np.random.default_rng(seed)
    pool = rng.integers(0, 2 ** 40, size=int(mean * n_bins * 1.5) + 10, dtype=np.uint64)
    vals, offs = [], [0]
    for b in range(n_bins):
        sz = max(1, int(mean * (rng.pareto(2) + 0.2 if skew else rng.uniform(0.5, 1.5))))
        if overlap and b % 3 and vals:
            base = vals[-1][:sz // 2]
            v = np.concatenate([base, rng.choice(pool, sz - len(base))])
        else:
            v = rng.integers(0, 2 ** 40, size=sz, dtype=np.uint64)
        vals.append(v.astype(np.uint64))
        offs.append(offs[-1] + len(v))
    return np.concatenate(vals), np.array(offs, dtype=np.uint64)


def set_script(seed: int):
    """Seeded script for a hash-set emulation test: (keys, ops) with ops[i] = 0: insert keys[i], 1: clear; a few segments of
    3 .. 120 000 keys, some with duplicates, most followed by a clear."""
    rng = np.random.default_rng(seed)
    keys, ops = [], []
    for _ in range(int(rng.integers(1, 5))):
        n = int(rng.choice([3, 50, 700, 9000, 120000]))
        k = rng.integers(0, 2 ** int(rng.choice([30, 40, 63])), size=n, dtype=np.uint64)
        if rng.random() < 0.5:
            k = np.concatenate([k, k[:n // 3]])
        keys.append(k)
        ops.append(np.zeros(len(k), dtype=np.uint8))
        if rng.random() < 0.7:
            keys.append(np.zeros(1, dtype=np.uint64))
            ops.append(np.ones(1, dtype=np.uint8))
    return np.concatenate(keys), np.concatenate(ops)
