"""bench_configs.py — the BASELINE.json configs other than the headline, each measured AT the config:

  C1  data/dna_example_split (5 DNA bins, k=3)            whole queries through the host path, ms/query
  C3  peptide flat IBF, 1024 bins, 1000 PROSITE patterns  K2 probes/s, filter + verify of the pattern batch
  C4  DNA flat IBF, 8192 bins, k=20, ~30 Gbp              K1 build streamed in slabs, K2 probes with canonical 20-mers,
                                                          column-sharded query step with its collectives
  C5  peptide flat IBF, 8192 bins, Murphy / Li alphabets  K1 k-mers/s
  C2' the headline's collection on the layout seqan::hibf itself computes (built by oracle/_ref on the box)

Every block carries `roofline` (algorithmic bytes of §8(d) / CUDA-event time / measured peak), a parity statement checked
at that size, and `cpu_baseline` (the reference's classes from oracle/_ref where they travelled, else the C port) on a
bounded sample.  Imported by bench.py only; the oracle is used as checker and CPU baseline, never on the measured path.
"""
from __future__ import annotations

import os
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))


class DevArray:
    """A raw device pointer dressed up for torch.as_tensor (zero copy) through __cuda_array_interface__."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 3}


def events(torch):
    return torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def roofline(achieved_gbs, peak, note, **more):
    d = {"bound": "hbm", "achieved": achieved_gbs, "peak": peak, "unit": "GB/s", "frac": achieved_gbs / peak, "traffic": None,
         "note": note}
    d.update(more)
    return d


def revcomp_canonical(v: np.ndarray, k: int) -> np.ndarray:
    """min(v, reverse complement) for 2-bit packed k-mers, k < 32, with TetRex's code (A0 C1 T2 G3, complement = code^2;
    include/nucleotide_decomposer.h:44-79) — how query k-mers are generated, not a checker."""
    x = v ^ np.uint64(0xAAAAAAAAAAAAAAAA)
    x = ((x >> np.uint64(2)) & np.uint64(0x3333333333333333)) | ((x & np.uint64(0x3333333333333333)) << np.uint64(2))
    x = ((x >> np.uint64(4)) & np.uint64(0x0F0F0F0F0F0F0F0F)) | ((x & np.uint64(0x0F0F0F0F0F0F0F0F)) << np.uint64(4))
    x = x.byteswap()
    x = x >> np.uint64(64 - 2 * k)
    return np.minimum(v, x)


# ------------------------------------------------------------------------------------------------------------ C4
C4_BINS, C4_K, C4_H, C4_RECORDS_PER_BIN, C4_RECORD_LEN = 8192, 20, 3, 4, 915_527
C4_SLAB_BINS = 64  # one 64-bin column word of the matrix = one slab of FASTA text (234 MB)


def c4_slab_text(torch, slab: int, n_bins: int = C4_SLAB_BINS):
    """Seeded ACGT text of bins [slab*64, slab*64 + n_bins), generated on the device (stand-in for FASTA parsing; the
    bytes depend on the slab number only, so any rank and the CPU checker can regenerate them)."""
    g = torch.Generator(device="cuda")
    g.manual_seed(8192_000 + slab)
    n = n_bins * C4_RECORDS_PER_BIN * C4_RECORD_LEN
    letters = torch.tensor(list(b"ACGT"), dtype=torch.uint8, device="cuda")
    return letters[torch.randint(0, 4, (n,), dtype=torch.int64, device="cuda", generator=g)]


def measure_c4(t, ctx, torch, rank, world, barrier, max_over_ranks, sum_over_ranks, peak, comm=None, bins=C4_BINS):
    """Builds THE C4 matrix (8192 bins x 22.83 M rows = 23.4 GB over all ranks) with K1 from text streamed slab by slab —
    include/index_ibf.h:101-131 is the loop this replaces; nothing of the collection is ever held whole, on the host or on
    the device.  Column shards: rank r holds words [r*128/S, (r+1)*128/S) of every row and reads only the text of its own
    bins (K1 needs no collective)."""
    out = {}
    emits_per_bin = C4_RECORDS_PER_BIN * (C4_RECORD_LEN + 1)  # L + 1 emits per record (the wrap-around quirk, F5)
    bin_size = t.compute_bitcount(emits_per_bin, 0.05)
    f = t.IBF(ctx, bins, bin_size, C4_H, shard_index=rank, shard_count=world)
    n_slabs = bins // C4_SLAB_BINS
    my_slabs = range(rank * n_slabs // world, (rank + 1) * n_slabs // world)
    recs = C4_SLAB_BINS * C4_RECORDS_PER_BIN
    d_offs = (torch.arange(recs + 1, dtype=torch.int64, device="cuda") * C4_RECORD_LEN)
    k1_ms, gen_s = 0.0, 0.0
    keep = {}
    barrier()
    wall0 = time.perf_counter()
    for slab in my_slabs:
        tg = time.perf_counter()
        text = c4_slab_text(torch, slab)
        d_bin = (torch.arange(recs, dtype=torch.int32, device="cuda") // C4_RECORDS_PER_BIN + slab * C4_SLAB_BINS).to(torch.int32)
        torch.cuda.synchronize()
        gen_s += time.perf_counter() - tg
        e0, e1 = events(torch)
        e0.record()
        f.emplace_records_slab_dev(t.MOL_NA, C4_K, 0, text.data_ptr(), text.numel(), d_offs.data_ptr(), d_bin.data_ptr(), recs,
                                   slab * C4_SLAB_BINS, C4_SLAB_BINS)
        e1.record()
        torch.cuda.synchronize()
        k1_ms += e0.elapsed_time(e1)
        if slab == my_slabs[0]:
            keep["text"] = text[:8 * C4_RECORDS_PER_BIN * C4_RECORD_LEN].cpu().numpy()  # first 8 bins of the rank's first slab
        del text
    barrier()
    wall = max_over_ranks(time.perf_counter() - wall0)
    k1_ms = max_over_ranks(k1_ms)
    n_kmers = float(bins * emits_per_bin)
    out["k1_build"] = {
        "kmers": n_kmers, "kmers_per_s": n_kmers / (k1_ms / 1e3), "kernel_ms": k1_ms,
        "wall_s_incl_text_generation": wall, "text_generation_s_per_rank": gen_s,
        "matrix_gb": bins / 8 * bin_size / 1e9, "matrix_gb_per_gpu": f.shard_n_words * 8 / 1e9, "bin_size": bin_size,
        "streaming": f"{n_slabs} slabs of {C4_SLAB_BINS} bins ({recs * C4_RECORD_LEN / 1e6:.0f} MB of text each), one "
                     "trx_emplace_records_slab_dev call per slab (bits collected in a dense 183 MB copy of the slab's word "
                     "column in L2-sized row slices, then folded into the interleaved matrix), no slab outlives its call",
        "roofline": roofline(n_kmers * (1 + 64 * C4_H) / (k1_ms / 1e3) / 1e9 / world, peak,
                             "per GPU; algorithmic = 1 B text + h x 64 B (32 B sector fill + 32 B write-back per atomicOr into "
                             "a matrix >> L2) = 193 B per emitted k-mer (SURVEY 8(d))", bytes_per_kmer=1 + 64 * C4_H,
                             kernel="emplace_dna_kernel")}
    return out, f, keep, bin_size


def c4_probe_and_check(t, ctx, torch, f, keep, bin_size, rank, world, barrier, max_over_ranks, peak, comm=None):
    """K2 on the matrix K1 just built: 2^24 uniform random canonical 20-mers + 2^20 20-mers of the collection, the SAME
    batch on every rank (= broadcast), each rank answering for its 64-bin column block; then parity at full size."""
    from oracle import bindings as ob
    res = {}
    rng = np.random.default_rng(8192)
    n_rand, n_present = 1 << 24, 1 << 20
    first_bins_text = keep["text"]  # 8 bins of this rank's first slab
    rec_offs = np.arange(8 * C4_RECORDS_PER_BIN + 1, dtype=np.uint64) * np.uint64(C4_RECORD_LEN)
    vals, emit_offs = ctx.decompose_records(t.MOL_NA, C4_K, 0, first_bins_text, rec_offs)
    pick = rng.integers(0, len(vals), n_present)
    present = vals[pick]
    present_bin = (np.searchsorted(emit_offs, pick, side="right") - 1) // C4_RECORDS_PER_BIN  # global bin on rank 0 (slab 0)
    rand = revcomp_canonical(rng.integers(0, 1 << (2 * C4_K), n_rand, dtype=np.uint64), C4_K)
    q = np.concatenate([present, rand])
    if world > 1:  # every rank probes rank 0's batch
        import torch.distributed as dist
        tq = torch.from_numpy(q.view(np.int64)).cuda()
        dist.broadcast(tq, src=0)
        tb = torch.from_numpy(present_bin.astype(np.int64)).cuda()
        dist.broadcast(tb, src=0)
        q = tq.cpu().numpy().view(np.uint64)
        present_bin = tb.cpu().numpy()
        del tq, tb
    d_q = torch.from_numpy(q.view(np.int64)).cuda()
    n = len(q)
    step = 1 << 21
    W = f.shard_words
    d_o = torch.empty(step * W, dtype=torch.int64, device="cuda")
    for _ in range(2):
        f.bulk_contains_dev(d_q.data_ptr(), step, d_o.data_ptr())
    barrier()
    e0, e1 = events(torch)
    e0.record()
    for s in range(0, n - step + 1, step):
        f.bulk_contains_dev(d_q.data_ptr() + 8 * s, step, d_o.data_ptr())
    e1.record()
    barrier()
    probed = (n // step) * step
    ms = max_over_ranks(e0.elapsed_time(e1))
    read_b, write_b = C4_H * W * 8, W * 8
    res["k2_probe"] = {
        "kmers": probed, "kmers_per_s": probed / (ms / 1e3), "ms": ms, "shards": world,
        "what": "real canonical 20-mers (2^20 of the collection + 2^24 uniform random) against all 8192 bins of the matrix K1 "
                "built; every rank probes the same batch against its column block, hit vectors stay in HBM",
        "roofline": roofline(probed * (read_b + write_b) / (ms / 1e3) / 1e9, peak,
                             f"per GPU: h x {W * 8} B row runs at random rows of a {f.shard_n_words * 8 / 1e9:.1f} GB shard + "
                             f"{W * 8} B hit vector written per k-mer", bytes_per_kmer_per_gpu=read_b + write_b,
                             kernel="probe_kernel")}
    del d_o
    # ---- parity at full size
    # (1) no false negatives: a k-mer of bin b has bit b set (on the rank that owns b)
    owner_lo, owner_hi = f.shard_word_begin * 64, (f.shard_word_begin + W) * 64
    chk = 1 << 16
    hv = f.bulk_contains(q[:chk])
    gb = present_bin[:chk]
    mine = (gb >= owner_lo) & (gb < owner_hi)
    local = (gb[mine] - owner_lo).astype(np.int64)
    got_bits = (hv[mine][np.arange(len(local)), local // 64] >> (local % 64).astype(np.uint64)) & np.uint64(1)
    res["present_kmers_hit_their_bin"] = bool(got_bits.all())
    # (2) K2 == AND of the rows the oracle's hash_and_fit names, read back from the device matrix (present + random)
    mat = torch.as_tensor(DevArray(f.device_words, (bin_size, W), "<i8"), device="cuda")
    sample = np.concatenate([np.arange(256), n_present + np.arange(256)])
    rows = np.array([[ob.orc().orc_hash_row(int(q[i]), h, bin_size) for h in range(C4_H)] for i in sample], dtype=np.int64)
    got_rows = mat[torch.from_numpy(rows.reshape(-1)).cuda()].cpu().numpy().view(np.uint64).reshape(len(sample), C4_H, W)
    res["k2_equals_oracle_rows"] = bool(np.array_equal(f.bulk_contains(q[sample]), np.bitwise_and.reduce(got_rows, axis=1)))
    # (3) K1 bit-exact at full size: this rank's 8 kept bins through the C oracle (decompose + hash_and_fit + set bit)
    # against the same 8 bit columns of the device matrix, all 22.83 M rows
    threads = os.cpu_count() or 1
    if rank == 0:
        o = ob.OracleIBF(64, bin_size, C4_H)
        rec_bin8 = np.repeat(np.arange(8, dtype=np.uint32), C4_RECORDS_PER_BIN)
        t0 = time.perf_counter()
        o.emplace_records(0, C4_K, 0, first_bins_text, rec_offs, rec_bin8, threads)
        cpu_s = time.perf_counter() - t0
        col = mat[:, 0].cpu().numpy().view(np.uint64)
        res["k1_bit_exact_8_bins_all_rows"] = bool(np.array_equal(col & np.uint64(0xFF), o.words & np.uint64(0xFF)))
        res["k1_checked_bits_set"] = int(np.bitwise_count(o.words & np.uint64(0xFF)).sum())
        n_k = 8 * C4_RECORDS_PER_BIN * (C4_RECORD_LEN + 1)
        res["cpu_baseline_k1"] = {"value": n_k / cpu_s, "unit": "k-mers/s", "cores": threads, "kind": "port",
                                  "sample": f"8 bins ({n_k} k-mers) into a 64-bin x {bin_size}-row matrix (183 MB): "
                                            "decompose_record + emplace of the C oracle, OpenMP over records"}
        del o
        # CPU probe figure on the /8 shard BASELINE.md promises: 1024 bins x 22.83 M rows = 2.9 GB, reference bulk_contains
        try:
            n_cpu = 1 << 20
            if ob.have_ref():
                r, kind = ob.RefIBF(1024, bin_size, C4_H), "reference"
            else:
                r, kind = ob.OracleIBF(1024, bin_size, C4_H), "port"
            r.bulk_contains_checksum(q[:1 << 14], threads)
            t0 = time.perf_counter()
            r.bulk_contains_checksum(q[:n_cpu], threads)
            dt = time.perf_counter() - t0
            t0 = time.perf_counter()
            r.bulk_contains_checksum(q[:n_cpu >> 3], 1)
            dt1 = time.perf_counter() - t0
            res["cpu_baseline_k2"] = {"value": n_cpu / dt, "unit": "k-mers/s against 1024 bins", "cores": threads, "kind": kind,
                                      "single_thread_value": (n_cpu >> 3) / dt1,
                                      "sample": f"{n_cpu} k-mers of the same batch, bulk_contains on the /8 shard shape "
                                                f"(1024 bins x {bin_size} rows = 2.9 GB, zero matrix: the access pattern "
                                                "depends on the hashes only), one agent per thread"}
            del r
        except MemoryError as e:
            res["cpu_baseline_k2"] = {"error": str(e)}
    del mat
    return res, d_q


def c4_sharded_query_step(t, ctx, torch, f, d_q, rank, world, barrier, max_over_ranks, peak, comm, n=1 << 20, steps=8):
    """The column-sharded QUERY path of SURVEY.md 8(e) on the C4 matrix, collectives included: the same batch of path
    items on every rank; per step one fused probe + AND into the rank's HBM-resident path slots (K2 MODE 1, 128/S words
    per slot) and one all-reduce(MAX) of the alive bytes over NCCL — `path_.none()` is global (include/otf_collector.h:383);
    after the last step the candidate vectors are all-gathered: ONE vector per query (the match accumulator, `path_matrix`
    of otf_collector.h:343,363), a query being 256 items of the batch here (config 3's patterns average 22 k probes per
    query, so this is generous).  Whole-job items/s, device time, max over ranks."""
    W = f.shard_words
    paths = t.Paths(ctx, W, n)
    d_slots = torch.arange(n, dtype=torch.int32, device="cuda")
    d_alive = torch.zeros(n, dtype=torch.uint8, device="cuda")
    items_per_query = 256
    n_queries = n // items_per_query
    d_all = torch.empty(n_queries * W * world, dtype=torch.int64, device="cuda") if world > 1 else None
    words_ptr = t.lib().trx_paths_device_words(paths._h)
    use_comm = comm is not None and world > 1

    def one_round():
        paths.fill(0, n, min(W * 64, f.bins - f.shard_word_begin * 64))
        for s in range(steps):
            paths.step_dev(f, d_q.data_ptr() + 8 * n * s, d_slots.data_ptr(), n, d_alive.data_ptr())
            if use_comm:
                comm.any_alive_dev(d_alive.data_ptr(), n)
        if use_comm:
            comm.allgather_dev(words_ptr, d_all.data_ptr(), n_queries * W * 8)
    one_round()
    barrier()
    c0 = comm.collectives if comm is not None else 0
    e0, e1 = events(torch)
    e0.record()
    one_round()
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    items = n * steps
    read_b = C4_H * W * 8 + 2 * W * 8  # rows + path slot read-modify-write
    out = {"items_per_s": items / (ms / 1e3), "ms": ms, "items_per_step": n, "steps": steps, "shards": world,
           "collectives_in_region": (comm.collectives - c0) if comm is not None else 0,
           "bytes_over_nvlink_per_rank": (steps * n + n_queries * W * 8 * (world - 1)) if world > 1 else 0,
           "queries": n_queries, "items_per_query": items_per_query,
           "alive_after_last_step": float(d_alive.float().mean().item()),
           "what": "fused probe + path AND (trx_paths_step_dev) + ncclAllReduce(MAX, 1 B/item) per step, ncclAllGather of one "
                   "candidate vector per query at the end; k-mer batch identical on all ranks",
           "roofline": roofline(items * read_b / (ms / 1e3) / 1e9, peak,
                                f"per GPU: (h + 2) x {W * 8} B per item and step (rows + slot read + slot write); the "
                                "collectives ride on the same stream and are inside the timed region",
                                bytes_per_item_per_gpu=read_b, kernel="probe_kernel<MODE 1> + ncclAllReduce")}
    paths.close()
    return out


# ------------------------------------------------------------------------------------------------------------ C5
def measure_c5(t, ctx, torch, residues, offs, peak, k=6, h=3, fpr=0.05, bins=8192):
    """Index-construction throughput (BASELINE config 5): the Swissprot-shaped collection re-split round robin into 8192
    bins, reduced-alphabet peptide mode (-r murphy, -r li): K1 = decompose + hash + set bit fused, straight from residues
    (include/index_ibf.h:101-131, include/peptide_decomposer.h:59-151)."""
    from oracle import bindings as ob
    n_rec = len(offs) - 1
    rec_bin = (np.arange(n_rec) % bins).astype(np.uint32)
    lens = (offs[1:] - offs[:-1]).astype(np.int64)
    emits = np.maximum(lens - k + 1, 0)
    per_bin = np.bincount(rec_bin, weights=emits, minlength=bins)
    bin_size = t.compute_bitcount(int(per_bin.max()), fpr)
    n_emits = int(emits.sum())
    d_res = torch.from_numpy(residues).cuda()
    d_offs = torch.from_numpy(offs.view(np.int64)).cuda()
    d_bin = torch.from_numpy(rec_bin.view(np.int32)).cuda()
    out = {"bins": bins, "bin_size": bin_size, "kmers": n_emits, "matrix_mb": bins / 8 * bin_size / 1e6}
    threads = os.cpu_count() or 1
    sub = 40_000  # records of the CPU sample
    for name, red in (("murphy", t.RED_MURPHY), ("li", t.RED_LI)):
        f = t.IBF(ctx, bins, bin_size, h)
        f.emplace_records_dev(t.MOL_AA, k, red, d_res.data_ptr(), len(residues), d_offs.data_ptr(), d_bin.data_ptr(), n_rec)
        torch.cuda.synchronize()
        reps = 3
        e0, e1 = events(torch)
        e0.record()
        for _ in range(reps):
            f.emplace_records_dev(t.MOL_AA, k, red, d_res.data_ptr(), len(residues), d_offs.data_ptr(), d_bin.data_ptr(), n_rec)
        e1.record()
        torch.cuda.synchronize()
        sec = e0.elapsed_time(e1) / reps / 1e3
        blk = {"kmers_per_s": n_emits / sec, "ms": sec * 1e3,
               "roofline": roofline(n_emits * (1 + 64 * h) / sec / 1e9, peak,
                                    "algorithmic = 1 B residue + h x 64 B per emitted k-mer (SURVEY 8(d): 193 B, the figure for "
                                    f"a matrix >> L2); this {bins / 8 * bin_size / 1e6:.0f} MB matrix is filled in L2-sized row "
                                    "slices, so most read-modify-writes never reach DRAM",
                                    bytes_per_kmer=1 + 64 * h, kernel="emplace_records_kernel<AA>")}
        # parity at the config's shape: the first `sub` records through the C oracle vs a device build of the same records
        f.clear()
        f.emplace_records_dev(t.MOL_AA, k, red, d_res.data_ptr(), int(offs[sub]), d_offs.data_ptr(), d_bin.data_ptr(), sub)
        got = f.download()
        o = ob.OracleIBF(bins, bin_size, h)
        t0 = time.perf_counter()
        o.emplace_records(1, k, red, residues[:int(offs[sub])], offs[:sub + 1], rec_bin[:sub], threads)
        cpu_s = time.perf_counter() - t0
        blk["matrix_bit_exact_vs_oracle_on_sample"] = bool(np.array_equal(got, o.words))
        blk["cpu_baseline"] = {"value": float(emits[:sub].sum()) / cpu_s, "unit": "k-mers/s", "cores": threads, "kind": "port",
                               "sample": f"first {sub} records ({int(emits[:sub].sum())} k-mers): decompose_record + emplace "
                                         "of the C oracle, OpenMP over records, same 8192-bin matrix shape"}
        out[name] = blk
        f.close()
        del o
    if ob.have_ref():  # the reference's own classes, single thread (its populate_index is serial over bins for a flat IBF)
        n1 = 2000
        r = ob.RefIBF(bins, bin_size, h)
        raw = residues.tobytes()
        t0 = time.perf_counter()
        cnt = 0
        for i in range(n1):
            v = ob.ref_decompose(1, k, 1, raw[int(offs[i]):int(offs[i + 1])])
            r.emplace(v, np.full(len(v), rec_bin[i], dtype=np.uint32))
            cnt += len(v)
        dt = time.perf_counter() - t0
        out["cpu_baseline_reference_1_thread"] = {"value": cnt / dt, "unit": "k-mers/s", "cores": 1, "kind": "reference",
                                                  "sample": f"first {n1} records, PeptideDecomposer::decompose_record (murphy) "
                                                            "+ interleaved_bloom_filter::emplace"}
    return out


# ------------------------------------------------------------------------------------------------------------ C3
def measure_c3_patterns(ix, patterns, scan_all_bins, budget_s=40.0):
    """BASELINE config 3 end to end through the reference-shaped host path (libtrxhost: process_query = regex front end,
    k-graph, collector with K2 + path algebra on the device; iter_disk_search = K4 on the candidate bins + formatting):
    as many of the 1000 PROSITE-style patterns as fit the time budget.  scan_all_bins(pattern) -> number of hits when
    every bin is verified (no filter).  The k-graph unrolls `*`/`+` to at most k-1 repetitions and the reference's filter is
    lossy for some quantified classes, so "filtered lines <= all-bin lines" is the invariant and the
    patterns where the filter loses lines are counted, not hidden."""
    from tetrex_b200 import sharded
    done, refused, lines, cand_bins, probes, fewer, more = 0, 0, 0, 0, 0, [], []
    t_filter = t_verify = 0.0
    t_start = time.perf_counter()
    for pat in patterns:
        if time.perf_counter() - t_start > budget_s:
            break
        t0 = time.perf_counter()
        try:
            words, st = ix.local_candidates(pat)
        except sharded.HostError:
            # without -a the enumeration of wildcard-rich patterns explodes; the collector refuses
            # beyond its live-path budget instead of running out of memory
            refused += 1
            continue
        t1 = time.perf_counter()
        _, n_lines = ix.verify_local(words, pat) if words.any() else ("", 0)
        t2 = time.perf_counter()
        t_filter += t1 - t0
        t_verify += t2 - t1
        done += 1
        lines += n_lines
        cand_bins += int(np.bitwise_count(words).sum())
        probes += st["probes"]
        want = scan_all_bins(pat)
        if n_lines < want:
            fewer.append((pat, n_lines, want))
        elif n_lines > want:
            more.append((pat, n_lines, want))
    total = t_filter + t_verify
    return {"patterns_run": done, "patterns_refused_too_many_live_paths": refused, "queries_per_s": done / max(total, 1e-9),
            "filter_us_per_query": t_filter / max(done, 1) * 1e6, "verify_us_per_query": t_verify / max(done, 1) * 1e6,
            "mean_candidate_bins": cand_bins / max(done, 1), "mean_probes_per_query": probes / max(done, 1), "lines": lines,
            "never_more_lines_than_scan_of_all_bins": not more, "patterns_where_the_filter_loses_lines": len(fewer),
            "examples_filter_loses_lines": fewer[:3],
            "what": "tetrex query semantics per pattern: process_query (filter, K2 on the device) then iter_disk_search (K4) "
                    "on the candidate bins only, one pattern at a time"}


def measure_c3_augmented(t, ctx, ix, files, workdir, patterns, scan_all_bins, budget_s=30.0):
    """Config 3's filter phase with `-a -g`: DGramIndex built by this repo on the device over the same 1024 files, then the
    pattern batch again with augmentation.  Reports the build (codes/s through count + emplace, file written) and the
    batch; the comparison with the plain leg is in refused patterns and filter time per query."""
    import os
    from tetrex_b200 import sharded
    path = os.path.join(workdir, "c3.dibf")
    t0 = time.perf_counter()
    per_bin = sharded.dindex_build(ctx, files, 3, 21, 3, 0.05, path)
    build_s = time.perf_counter() - t0
    out = {"dgram_index": {"gaps": [3, 21], "codes": int(per_bin.sum()), "largest_bin_codes": int(per_bin.max()),
                           "file_gb": os.path.getsize(path) / 1e9,
                           "build_s_incl_fasta_parse_download_and_file_write": round(build_s, 2)}}
    ix.set_augment(True)
    res_no = measure_c3_patterns(ix, patterns, scan_all_bins, budget_s=budget_s / 2)
    t0 = time.perf_counter()
    ix.attach_dgram(path)
    out["dgram_index"]["load_s"] = round(time.perf_counter() - t0, 2)
    res_g = measure_c3_patterns(ix, patterns, scan_all_bins, budget_s=budget_s / 2)
    # `tetrex query -f`: the filter phase of the whole motif file with ONE membership batch (process_query_batch)
    ix.local_candidates_batch(patterns[:16])
    t0 = time.perf_counter()
    words, status, st = ix.local_candidates_batch(patterns)
    dt = time.perf_counter() - t0
    same = True
    for i in range(0, len(patterns), max(1, len(patterns) // 40)):
        try:
            one, _ = ix.local_candidates(patterns[i])
            same = same and bool(status[i]) and bool((one == words[i]).all())
        except sharded.HostError:
            same = same and not status[i]
    out["motif_file_filter_batch"] = {
        "patterns": len(patterns), "answered": int(status.sum()), "patterns_per_s": len(patterns) / dt,
        "us_per_pattern": dt / len(patterns) * 1e6, "kmers_probed_by_the_pruned_walks": int(st[:, 0].sum()),
        "equals_one_pattern_at_a_time_on_a_sample": same,
        "what": "process_query for all patterns of a motif file (-a -g): dry runs of the k-graph walks, one membership batch "
                "for the union of their k-mers, the walks on host words"}
    ix.attach_dgram(None)
    ix.set_augment(False)
    for r in (res_no, res_g):
        r["what"] = "the same with tetrex query -a: OTFCollector::augment before the walk"
    res_g["what"] += ", DGram index attached (-g)"
    out["augment_only"] = res_no
    out["augment_with_dgram"] = res_g
    os.unlink(path)
    return out


def cpu_baseline_flat_probe(bins, bin_size, h, kmers):
    """The reference's bulk_contains on a flat IBF of the config's shape, all cores and one core."""
    from oracle import bindings as ob
    threads = os.cpu_count() or 1
    if ob.have_ref():
        r, kind = ob.RefIBF(bins, bin_size, h), "reference"
    else:
        r, kind = ob.OracleIBF(bins, bin_size, h), "port"
    n = min(len(kmers), 1 << 21)
    r.bulk_contains_checksum(kmers[:1 << 14], threads)
    t0 = time.perf_counter()
    r.bulk_contains_checksum(kmers[:n], threads)
    dt = time.perf_counter() - t0
    t0 = time.perf_counter()
    r.bulk_contains_checksum(kmers[:n >> 3], 1)
    dt1 = time.perf_counter() - t0
    return {"value": n / dt, "unit": "probes/s", "cores": threads, "kind": kind, "single_thread_value": (n >> 3) / dt1,
            "sample": f"{n} random 30-bit k-mers, bulk_contains on {bins} bins x {bin_size} rows (zero matrix, access pattern "
                      "depends on the hashes only), one containment agent per thread"}


# ------------------------------------------------------------------------------------------------------------ C1
def measure_c1(t, ctx, reps=300):
    """BASELINE config 1 on the reference's own five files (tests/data/dna_example_split, the reference-written index
    files of tests/golden/cli): whole queries A(C+|G+)T through the host path, resident and as a one-shot CLI process,
    beside the reference's own load + process_query + iter_disk_search on one host thread."""
    import subprocess
    from oracle import bindings as ob
    from tetrex_b200 import sharded
    out = {}
    q = "A(C+|G+)T"
    cwd = os.getcwd()
    os.chdir(ROOT)  # the index names its FASTA bins relative to the repo root
    try:
        for name in ("c1_dna_flat_k3", "c1_dna_hibf_k3"):
            idx = os.path.join("tests", "golden", "cli", name + ".ibf")
            ix = sharded.ShardedIndex(ctx, idx, 0, 1)
            words, _ = ix.local_candidates(q)
            text, n_lines = ix.verify_local(words, q)
            t0 = time.perf_counter()
            for _ in range(reps):
                w, _ = ix.local_candidates(q)
                ix.verify_local(w, q)
            dt = (time.perf_counter() - t0) / reps
            blk = {"ms_per_query_resident": dt * 1e3, "queries_per_s_resident": 1 / dt, "lines": n_lines,
                   "candidate_bins": int(np.bitwise_count(words).sum())}
            ix.close()
            tetrex = os.path.join(ROOT, "tetrex_b200", "bin", "tetrex")
            t0 = time.perf_counter()
            p = subprocess.run([tetrex, "query", idx, q], capture_output=True, text=True)
            blk["ms_one_shot_cli_process"] = (time.perf_counter() - t0) * 1e3
            blk["cli_output_equals_resident"] = p.stdout == text
            if ob.have_ref():
                ob.ref_index_query(idx, q, True, 1)
                n_ref = 50
                t0 = time.perf_counter()
                for _ in range(n_ref):
                    _, ref_text = ob.ref_index_query(idx, q, True, 1)
                dt_ref = (time.perf_counter() - t0) / n_ref
                blk["output_equals_reference"] = ref_text == text
                blk["cpu_baseline"] = {"value": dt_ref * 1e3, "unit": "ms/query", "cores": 1, "kind": "reference",
                                       "sample": f"{n_ref} x (load_ibf + process_query + iter_disk_search), -t 1, in process"}
            out[name] = blk
    finally:
        os.chdir(cwd)
    out["query"] = q
    return out


# ------------------------------------------------------------------------------------------- C2 on seqan's layout
def measure_c2_seqan_layout(t, ctx, torch, residues, offs, rec_bin, n_bins, q_host, peak, k=6, h=3, fpr=0.05, steps=10):
    """The headline's collection laid out and built by seqan::hibf itself — HIBFIndex::populate_index's call
    (include/index_hibf.h:110-129: input_fn, number_of_user_bins, hash functions, maximum_fpr, threads; everything else hibf
    defaults: sketches, DP layout, split + merged bins, FPR correction) executed by oracle/_ref on the box's host cores —
    uploaded and probed with the headline's k-mer batch.  The reference here produces the INPUT (an index file's content)
    and the expected answers; the measured path is K3."""
    from oracle import bindings as ob
    if not ob.have_ref():
        return {"unavailable": "oracle/_ref did not travel to this machine"}
    threads = os.cpu_count() or 1
    aamap, _ = t.peptide_maps(t.RED_BASE)
    t0 = time.perf_counter()
    n = len(residues)
    codes = aamap[residues].astype(np.uint64)
    v = np.zeros(n - k + 1, dtype=np.uint64)
    for j in range(k):
        v = (v << np.uint64(5)) | codes[j:n - k + 1 + j]
    del codes
    starts = np.arange(n - k + 1, dtype=np.uint64)
    rec = np.searchsorted(offs, starts, side="right") - 1
    inside = starts + np.uint64(k) <= offs[rec + 1]
    vals, bins_of = v[inside], rec_bin[rec[inside]]
    del v, starts, rec, inside
    bo = np.searchsorted(bins_of, np.arange(n_bins + 1))  # records are grouped by ascending bin
    per = [vals[bo[i]:bo[i + 1]] for i in range(n_bins)]
    prep_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    rh = ob.RefHIBF(per, h, fpr, threads)
    build_s = time.perf_counter() - t0
    tb = rh.tables()
    del per, vals
    hibf = t.HIBF.from_tables(ctx, tb)
    W = hibf.out_words
    batch = len(q_host)
    d_q = torch.from_numpy(q_host.view(np.int64)).cuda()
    d_out = torch.empty(batch * W, dtype=torch.int64, device="cuda")
    for _ in range(3):
        st = hibf.membership_dev(d_q.data_ptr(), batch, d_out.data_ptr())
    torch.cuda.synchronize()
    e0, e1 = events(torch)
    e0.record()
    for _ in range(steps):
        st = hibf.membership_dev(d_q.data_ptr(), batch, d_out.data_ptr())
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    n_chk = 1 << 15
    got = d_out[:n_chk * W].cpu().numpy().view(np.uint64).reshape(n_chk, W)
    t0 = time.perf_counter()
    want = rh.membership(q_host[:n_chk], threads)
    cpu_s = time.perf_counter() - t0
    split = sum(int((np.bincount(u[u < np.uint64(2 ** 63)].astype(np.int64)) > 1).sum()) for u in tb.user_bin)
    algo = 32.0 * h * float(st[0])
    out = {"value": batch / (ms / 1e3), "unit": "probes/s", "ms_per_step": ms, "kmers_per_step": batch,
           "layout": f"{tb.n_ibfs} IBFs, {sum(w.nbytes for w in tb.words) / 1e6:.0f} MB, {split} of {n_bins} user bins split over "
                     f"several technical bins, technical bins per IBF {sorted(set(int(b) for b in tb.bins))}",
           "built_by": f"seqan::hibf (oracle/_ref) on {threads} host threads in {build_s:.1f} s (+ {prep_s:.1f} s k-mer lists)",
           "visited_ibfs_per_kmer": float(st[0]) / batch,
           "answers_equal_reference_membership_for": bool(np.array_equal(got, want)), "answers_checked": n_chk,
           "roofline": roofline(algo / (ms / 1e3) / 1e9, peak, "32 B sector per row x h x visited IBFs (SURVEY 8(d)), "
                                "all K3 kernels of the step", algorithmic_bytes_per_step=algo,
                                kernel="hibf_root_kernel + hibf_window_kernel + hibf_merge_tile_kernel"),
           "cpu_baseline": {"value": n_chk / cpu_s, "unit": "probes/s", "cores": threads, "kind": "reference",
                            "sample": f"first {n_chk} k-mers of the batch, membership_for on the same index"}}
    hibf.close()
    return out
