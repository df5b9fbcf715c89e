#!/usr/bin/env python
"""bench.py — headline measurement of the TetRex hot path on B200 (contract: see the task statement).

Workload (BASELINE.json configs[1], "C2"): synthetic Swissprot-shaped peptide HIBF — 1024 user bins,
~570 k sequences / ~200 M residues, k = 6, 3 hash functions, FPR 0.05 — probed with a batch of 6-mers.
One step = one batch of k-mers through HIBF membership (K3) = the probe part of the query path; the
regex-verify part (K4) and the flat-IBF probe (K2) are measured beside it and reported under "extra".

  metric  IBF k-mer probes/s, whole job (all ranks), inputs resident in HBM
  e2e     the same through the host-buffer C-ABI call (pinned host k-mers in, pinned host bitvectors out)
  roofline.achieved = 32 B x hash functions x visited IBFs per step / step time

`--impl reference` times the reference's own CPU implementation (seqan::hibf membership_agent compiled
unmodified into oracle/_ref, else the C oracle port) on the host cores; it never touches CUDA.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "IBF k-mer probes/sec (HIBF membership, peptide 6-mers, 1024 user bins)"
UNIT = "probes/s"
K, HASH_FUNS, FPR, N_BINS = 6, 3, 0.05, 1024
QUERY = "LMA(E|Q)GLYN"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--seqs", type=int, default=570_000, help="sequences in the synthetic collection (570000 = C2)")
    ap.add_argument("--batch", type=int, default=1 << 22, help="k-mers per rank per step")
    ap.add_argument("--no-extra", action="store_true", help="skip the K2/K4 side measurements")
    ap.add_argument("--c4-bins", type=int, default=8192, help="bins of the C4 build (8192 = the config; smaller for smoke runs)")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._stop_evt = index, [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.05)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def make_queries(tables_kmers: np.ndarray, n: int, seed: int) -> np.ndarray:
    """Half k-mers that occur in the collection, half uniform random 30-bit values (mostly absent)."""
    rng = np.random.default_rng(seed)
    present = rng.choice(tables_kmers, n // 2)
    absent = rng.integers(0, 1 << (5 * K), n - n // 2, dtype=np.uint64)
    q = np.concatenate([present, absent])
    rng.shuffle(q)
    return np.ascontiguousarray(q, dtype=np.uint64)


class OracleIbfAdapter:
    """CPU builder used ONLY by the reference arm / cpu_baseline (oracle = checker side of the repo)."""

    def __init__(self, bins, bin_size, hash_funs, threads):
        from oracle import bindings as ob
        self.o = ob.OracleIBF(bins, bin_size, hash_funs)
        self.threads = threads

    def emplace_records(self, mol, k, red, residues, offs, rec_bin):
        self.o.emplace_records(mol, k, red, residues, offs, rec_bin, self.threads)

    def download(self):
        return self.o.words

    def close(self):
        pass


def collection_kmers(residues, offs, aamap, limit=3_000_000):
    """Packed 6-mers (peptide_decomposer packing) of the first `limit` residues, windows inside records only."""
    n = int(min(limit, len(residues)))
    codes = aamap[residues[:n]].astype(np.uint64)
    v = np.zeros(n - K + 1, dtype=np.uint64)
    for j in range(K):
        v = (v << np.uint64(5)) | codes[j:n - K + 1 + j]
    starts = np.arange(n - K + 1, dtype=np.uint64)
    rec = np.searchsorted(offs, starts, side="right") - 1
    inside = starts + np.uint64(K) <= offs[rec + 1]
    return v[inside]


def run_ours(args):
    import torch

    import tetrex_b200 as t
    from tetrex_b200 import synth

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    # torch's default stream has handle 0, which the C ABI reads as "make your own": give both sides one explicit
    # stream so that torch's CUDA events bracket exactly the stream the kernels are launched on
    tstream = torch.cuda.Stream()
    torch.cuda.set_stream(tstream)
    ctx = t.Context(local, tstream.cuda_stream)

    # ---------------- setup (untimed): collection, GPU-built HIBF replica, query batch
    t0 = time.time()
    residues, offs, rec_bin, planted = synth.swissprot_like(args.seqs, N_BINS)
    tables = synth.build_swissprot_hibf(lambda *a: t.IBF(ctx, *a), residues, offs, rec_bin, N_BINS, K, HASH_FUNS, FPR)
    hibf = t.HIBF(ctx, tables["n_user_bins"], tables["bins"], tables["bin_size"], tables["hash_funs"], tables["words"],
                  tables["next_ibf"], tables["user_bin"])
    aamap, _ = t.peptide_maps(t.RED_BASE)
    pool = collection_kmers(residues, offs, aamap)
    q_host = make_queries(pool, args.batch, 1234 + rank)
    setup_s = time.time() - t0
    W = hibf.out_words
    d_q = torch.from_numpy(q_host.view(np.int64)).cuda()
    d_out = torch.empty(args.batch * W, dtype=torch.int64, device="cuda")
    pin_q = torch.from_numpy(q_host.view(np.int64)).pin_memory()
    pin_out = torch.empty(args.batch * W, dtype=torch.int64).pin_memory()

    from tetrex_b200.dist import max_over_ranks, sum_over_ranks  # all-reduce over NCCL when world > 1

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident timing: K steps of K3 on the batch
    visited = 0
    for _ in range(args.warmup):
        st = hibf.membership_dev(d_q.data_ptr(), args.batch, d_out.data_ptr())
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = ctx.launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        st = hibf.membership_dev(d_q.data_ptr(), args.batch, d_out.data_ptr())
        visited += int(st[0])
    ev1.record()
    barrier()
    ms = max_over_ranks(ev0.elapsed_time(ev1))
    clocks = sampler.stop()
    launches = ctx.launches - launches0
    total_probes = sum_over_ranks(float(args.batch * args.steps))
    value = total_probes / (ms / 1e3)
    visited_per_step = visited / args.steps
    checksum_dev = int(d_out.sum().item())

    # ---------------- end to end: host buffers through the C-ABI call (H2D + kernels + D2H inside the region)
    q_np = pin_q.numpy().view(np.uint64)
    out_np = pin_out.numpy().view(np.uint64).reshape(args.batch, W)
    for _ in range(2):
        hibf.membership(q_np, out=out_np)
    barrier()
    e_steps = max(3, args.steps // 2)
    tw = time.perf_counter()
    for _ in range(e_steps):
        hibf.membership(q_np, out=out_np)
    torch.cuda.synchronize()
    e_ms = max_over_ranks((time.perf_counter() - tw) * 1e3)
    barrier()
    e2e_value = sum_over_ranks(float(args.batch * e_steps)) / (e_ms / 1e3)
    assert int(out_np.view(np.int64).sum()) == checksum_dev, "host-buffer path and device path disagree"
    # what bounds that number: the 128-byte row per k-mer going back over PCIe.  A plain pinned device→host copy of the
    # same result buffer, timed the same way, is the ceiling of any host-buffer API that returns whole bitvectors.
    pin_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    tw = time.perf_counter()
    for _ in range(3):
        pin_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    d2h_gbs = 3 * args.batch * W * 8 / (time.perf_counter() - tw) / 1e9
    e2e_pcie = {"bound": "pcie d2h", "achieved": args.batch * W * 8 / (e_ms / e_steps / 1e3) / 1e9, "peak": d2h_gbs, "unit": "GB/s",
                "frac": args.batch * W * 8 / (e_ms / e_steps / 1e3) / 1e9 / d2h_gbs,
                "note": "result bytes of one step / its wall time, against a plain pinned cudaMemcpy D2H of the same buffer "
                        "measured in this run (per rank)"}

    # ---------------- end to end in membership_for's OWN result form: seqan returns the ids of the user bins that hold the
    # value (hierarchical_interleaved_bloom_filter.hpp:321-355); TetRex turns them into a bitvector afterwards
    # (include/index_hibf.h:132-147).  trx_hibf_membership_ids compacts the rows on the device, so what crosses PCIe is a
    # 4-byte count and ~17 two-byte ids per k-mer instead of a 128-byte row.
    pin_counts = torch.empty(args.batch, dtype=torch.int32).pin_memory()
    pin_ids = torch.empty(args.batch * 48, dtype=torch.int16).pin_memory()
    cnt_np = pin_counts.numpy().view(np.uint32)
    ids_np = pin_ids.numpy().view(np.uint16)
    for _ in range(2):
        got_counts, got_ids = hibf.membership_ids(q_np, id_bytes=2, counts=cnt_np, ids=ids_np)
    barrier()
    tw = time.perf_counter()
    for _ in range(e_steps):
        got_counts, got_ids = hibf.membership_ids(q_np, id_bytes=2, counts=cnt_np, ids=ids_np)
    torch.cuda.synchronize()
    i_ms = max_over_ranks((time.perf_counter() - tw) * 1e3)
    barrier()
    ids_value = sum_over_ranks(float(args.batch * e_steps)) / (i_ms / 1e3)
    n_ids = int(len(got_ids))
    # same answer as the rows: every k-mer's count, and the ids of the first 8192 k-mers bit for bit
    assert np.array_equal(got_counts, np.bitwise_count(out_np).sum(axis=1).astype(np.uint32)), "ids form: counts differ from the rows"
    head = np.zeros((8192, W * 64), dtype=np.uint8)
    ends = np.cumsum(got_counts[:8192].astype(np.int64))
    head[np.repeat(np.arange(8192), got_counts[:8192]), got_ids[:ends[-1]].astype(np.int64)] = 1
    assert np.array_equal(np.packbits(head, axis=1, bitorder="little").view(np.uint64), out_np[:8192]), "ids form: ids differ from the rows"
    e2e_ids = {"value": ids_value, "unit": UNIT, "h2d_bytes_per_step": args.batch * 8,
               "d2h_bytes_per_step": args.batch * 4 + n_ids * 2, "ids_per_kmer": n_ids / args.batch, "steps": e_steps}

    # ---------------- end to end the way the collector uses the device (OTFCollector::update_path, include/otf_collector.h:
    # 247-279): k-mers and their path slots go in from pinned host memory, the hit vectors are ANDed into HBM-resident
    # path bitvectors, ONE alive byte per k-mer comes back — the per-k-mer bitvectors never cross PCIe
    collector_e2e = None
    if not args.no_extra:
        paths = t.Paths(ctx, W, args.batch)
        paths.fill(0, args.batch, N_BINS)
        ident = torch.arange(args.batch, dtype=torch.int32).pin_memory().numpy().view(np.uint32)  # pinned like the k-mers
        alive = torch.zeros(args.batch, dtype=torch.uint8).pin_memory().numpy()
        from tetrex_b200._abi import check, u64p  # noqa: F401
        import ctypes as C

        def collector_step():
            d_q.copy_(pin_q, non_blocking=True)  # 8 B per k-mer over PCIe, on the context's stream
            hibf.membership_dev(d_q.data_ptr(), args.batch, d_out.data_ptr())
            check(t.lib().trx_paths_and_hits(paths._h, d_out.data_ptr(), ident.ctypes.data_as(C.c_void_p),
                                             ident.ctypes.data_as(C.c_void_p), args.batch, alive.ctypes.data_as(C.c_void_p)))
        for _ in range(2):
            collector_step()
        barrier()
        tw = time.perf_counter()
        for _ in range(e_steps):
            collector_step()
        torch.cuda.synchronize()
        c_ms = max_over_ranks((time.perf_counter() - tw) * 1e3)
        barrier()
        collector_e2e = {"value": sum_over_ranks(float(args.batch * e_steps)) / (c_ms / 1e3), "unit": UNIT,
                         "h2d_bytes_per_step": args.batch * 16, "d2h_bytes_per_step": args.batch,
                         "what": "k-mers + slot ids in (pinned host), membership on the device, AND into HBM-resident path "
                                 "slots (trx_paths_and_hits), one alive byte per k-mer out",
                         "alive_fraction_after_steps": float(alive.mean())}
        paths.close()

    # ---------------- BASELINE config 4 (8192-bin DNA IBF, k = 20, 23.4 GB), bins sharded over the ranks in 64-bin column
    # blocks, the SAME k-mer batch on every rank (= broadcast): K2 bulk_contains of every k-mer against all 8192 bins
    # and the real thing: the matrix BUILT by K1 from streamed text, probed with canonical 20-mers, checked against the
    # oracle at full size, and the column-sharded collector step with its NCCL collectives (bench_configs.py)
    flat_c4, c4 = None, None
    comm = None
    if not args.no_extra and 128 % world == 0:
        import bench_configs as bc
        peak_now, _ = peaks()
        if world > 1:
            comm = t.Comm.from_torch_dist(ctx)
        flat_c4 = measure_c4_sharded_probe(t, ctx, torch, rank, world, barrier, max_over_ranks)
        try:
            c4, f4, keep4, bs4 = bc.measure_c4(t, ctx, torch, rank, world, barrier, max_over_ranks, sum_over_ranks, peak_now, comm,
                                               bins=args.c4_bins)
            more, d_q4 = bc.c4_probe_and_check(t, ctx, torch, f4, keep4, bs4, rank, world, barrier, max_over_ranks, peak_now, comm)
            c4.update(more)
            c4["sharded_query_step"] = bc.c4_sharded_query_step(t, ctx, torch, f4, d_q4, rank, world, barrier, max_over_ranks,
                                                                peak_now, comm)
            del d_q4
            f4.close()
            torch.cuda.empty_cache()
        except Exception as e:  # a side measurement never costs the headline line; on N > 1 every rank fails alike
            import traceback
            c4 = {"error": f"{type(e).__name__}: {e}"[:400], "trace": traceback.format_exc()[-600:]}

    line = None
    if rank == 0:
        peak, peak_src = peaks()
        algo_bytes = 32.0 * HASH_FUNS * visited_per_step
        traffic = None  # DRAM bytes per step of the K3 kernels, from the committed ncu --set full capture
        tpath = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r2_traffic.json")
        if os.path.exists(tpath):
            traffic = json.load(open(tpath))["k3_dram_bytes_per_kmer"] * args.batch
        step_s = ms / 1e3 / args.steps
        achieved = algo_bytes / step_s / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": "C2: synthetic Swissprot-shaped peptide HIBF, 1024 user bins, k=6, h=3, FPR 0.05",
                       "sequences": args.seqs, "residues": int(offs[-1]), "kmers_per_rank_per_step": args.batch,
                       "query_mix": "50% k-mers of the collection, 50% uniform random 30-bit values",
                       "hibf": f"{len(tables['words'])} IBFs x 64 technical bins, 2 levels, "
                               f"{sum(w.nbytes for w in tables['words']) / 1e6:.0f} MB, replicated per GPU",
                       "parallelism": f"replica per GPU, k-mer batches sharded x{world}, no data-path collective",
                       "l2": "working set (HIBF + 512 MiB of result bitvectors per step) exceeds the 126 MB L2; no flush",
                       "visited_ibfs_per_kmer": visited_per_step / args.batch, "setup_s": round(setup_s, 1)},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "kernel": "hibf_window_kernel (+ hibf_root_kernel, hibf_merge_tile_kernel: one step)",
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_step": algo_bytes,
                         "note": "algorithmic = 32-byte sector per 8-byte row read x h x visited IBFs (SURVEY 8(d)); achieved = that / "
                                 "the step's device time (all K3 kernels); traffic = measured DRAM bytes per step (ncu), lower than "
                                 "algorithmic because the child windows are served from L2"},
            "e2e": dict(e2e_ids, what="trx_hibf_membership_ids with pinned host buffers: k-mers in, per k-mer a count and the ids of "
                                      "its user bins out (membership_for's own result form), checked against the bitvector rows"),
            "gpu_launches": int(launches), "clocks": clocks,
        }
        line["e2e"]["bitvector_form"] = {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": args.batch * 8,
                                         "d2h_bytes_per_step": args.batch * W * 8, "steps": e_steps, "roofline": e2e_pcie,
                                         "what": "trx_hibf_membership: one row of ceil(user bins / 64) words per k-mer "
                                                 "(HIBFIndex::query's form) — bound by PCIe"}
        if collector_e2e is not None:
            line["e2e"]["collector_step"] = collector_e2e
        if flat_c4 is not None:
            flat_c4["roofline"]["peak"] = peak
            flat_c4["roofline"]["frac"] = flat_c4["roofline"]["achieved"] / peak
            line["flat_c4_sharded"] = flat_c4
        line["configs"] = {"C2": "the headline above (value / e2e / roofline / cpu_baseline)"}
        if c4 is not None:
            line["configs"]["C4"] = c4
        if world == 1:
            line["cpu_baseline"] = cpu_baseline(tables, q_host)
            if not args.no_extra:
                line["extra"] = extra_measurements(t, ctx, torch, residues, offs, rec_bin, planted, hibf, peak)
                import bench_configs as bc
                for key, fn in (("C1", lambda: bc.measure_c1(t, ctx)),
                                ("C5", lambda: bc.measure_c5(t, ctx, torch, residues, offs, peak)),
                                ("C2_seqan_layout", lambda: bc.measure_c2_seqan_layout(t, ctx, torch, residues, offs, rec_bin,
                                                                                       N_BINS, q_host, peak))):
                    try:
                        line["configs"][key] = fn()
                    except Exception as e:
                        import traceback
                        line["configs"][key] = {"error": f"{type(e).__name__}: {e}"[:400], "trace": traceback.format_exc()[-600:]}
                line["configs"]["C3"] = line["extra"].pop("C3", None)
    if comm is not None:
        comm.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        emit(line)


def host_path_queries(t, ctx, residues, offs, rec_bin, reps=100, on_index=None):
    """Writes the synthetic collection as 1024 FASTA files, builds a HIBF and a flat index with tetrex_b200/bin/tetrex
    (K1), loads them through libtrxhost.so and times whole queries: regex -> k-graph -> collector (K2/K3 + path algebra on
    the device, one host decision per rank of the graph) -> candidate bins -> K4 verify -> formatted lines."""
    import shutil
    import subprocess
    import tempfile
    from tetrex_b200 import sharded
    tetrex = os.path.join(ROOT, "tetrex_b200", "bin", "tetrex")
    d = tempfile.mkdtemp(prefix="trx_bench_")
    res = {}
    try:
        raw = np.asarray(residues, dtype=np.uint8).tobytes()
        o = [int(x) for x in offs]
        files, r = [], 0
        for b in range(N_BINS):
            path = os.path.join(d, f"bin_{b:04d}.fa")
            parts = []
            while r < len(rec_bin) and rec_bin[r] == b:
                parts.append(b">sp|S%07d|SYN_%d\n" % (r, b))
                parts.append(raw[o[r]:o[r + 1]])
                parts.append(b"\n")
                r += 1
            with open(path, "wb") as f:
                f.write(b"".join(parts))
            files.append(path)
        lst = os.path.join(d, "bins.lst")
        with open(lst, "w") as f:
            f.write("\n".join(files) + "\n")
        for name, flags in (("hibf", []), ("flat", ["-i"])):
            t0 = time.perf_counter()
            p = subprocess.run([tetrex, "index"] + flags + ["-k", str(K), os.path.join(d, name), lst], capture_output=True,
                               text=True, timeout=600)
            if p.returncode != 0:
                raise RuntimeError(p.stderr[-300:])
            build_s = time.perf_counter() - t0
            ix = sharded.ShardedIndex(ctx, os.path.join(d, name + ".ibf"), 0, 1, preload_text=True)
            words, text = ix.query(QUERY)
            lines = text.count("\n")
            t0 = time.perf_counter()
            for _ in range(reps):
                w, st = ix.local_candidates(QUERY)
            t1 = time.perf_counter()
            for _ in range(reps):
                ix.verify_local(w, QUERY)
            t2 = time.perf_counter()
            res[name] = {"queries_per_s": reps / (t2 - t0), "process_query_us": (t1 - t0) / reps * 1e6,
                         "iter_disk_search_us": (t2 - t1) / reps * 1e6, "candidate_bins": int(np.bitwise_count(words).sum()),
                         "lines": lines, "collector": st, "index_build_s_incl_process_start_and_fasta_parse": round(build_s, 2)}
            if on_index is not None:
                on_index(name, ix, files, d)
            ix.close()
        res["query"] = QUERY
        return res
    finally:
        shutil.rmtree(d, ignore_errors=True)


def measure_c4_sharded_probe(t, ctx, torch, rank, world, barrier, max_over_ranks, n=1 << 21, steps=5):
    """K2 at the C4 shape (8192 bins x 22.83 M rows = 23.4 GB, h = 3, 3072 B read per k-mer): rank r holds the
    column block r of `world` (128/world words of every row), every rank probes the same n k-mers, so one k-mer is
    answered for all 8192 bins in the time the slowest rank needs for its shard.  The matrix content does not influence
    the access pattern (rows are chosen by the hash of the k-mer), so it is filled with a byte pattern instead of 30 Gbp
    of emplaces; the hit vectors of the shards are written to HBM and stay there (the collector consumes them on the
    device).  Returns whole-job k-mers/s with device timing, max over ranks."""
    bins, hash_funs = 8192, 3
    bin_size = t.compute_bitcount(4 * (915_527 + 1), 0.05)  # 4 records of 915 527 nt per bin, L + 1 emits per record (F5)
    f = t.IBF(ctx, bins, bin_size, hash_funs, shard_index=rank, shard_count=world)
    from tetrex_b200._abi import check
    shard_words = f.shard_words
    check(t.lib().trx_dev_memset(ctx._h, f.device_words, 0x5A, f.shard_n_words * 8))
    g = torch.Generator(device="cuda")
    g.manual_seed(8192)
    q = torch.randint(0, 1 << 40, (n,), dtype=torch.int64, device="cuda", generator=g)
    o = torch.empty(n * shard_words, dtype=torch.int64, device="cuda")
    for _ in range(3):
        f.bulk_contains_dev(q.data_ptr(), n, o.data_ptr())
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        f.bulk_contains_dev(q.data_ptr(), n, o.data_ptr())
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1)) / steps
    read_b, write_b = hash_funs * shard_words * 8, shard_words * 8
    out = {"workload": "C4 shape: flat IBF 8192 bins x %d rows (%.1f GB), h=3, uniform random 40-bit k-mers, matrix filled "
                       "with a byte pattern" % (bin_size, bins / 8 * bin_size / 1e9),
           "shards": world, "shard_gb_per_gpu": f.shard_n_words * 8 / 1e9, "kmers_per_step": n, "ms_per_step": ms,
           "value": n / (ms / 1e3), "unit": "k-mers probed against all 8192 bins per second (whole job)",
           "roofline": {"bound": "hbm", "achieved": n * (read_b + write_b) / (ms / 1e3) / 1e9, "unit": "GB/s per GPU",
                        "note": "per k-mer and GPU: h x shard_words x 8 B of row reads (random %d-byte runs in a matrix far "
                                "larger than L2) + shard_words x 8 B of hit vector written" % (shard_words * 8),
                        "bytes_per_kmer_per_gpu": read_b + write_b}}
    del o, q
    f.close()
    return out


def cpu_baseline(tables, q_host, budget_s: float = 12.0):
    """The same HIBF probed on the host cores: seqan::hibf's own membership_agent (oracle/_ref) if it travelled,
    else the C port.  Bounded sample of the step's k-mer batch."""
    from oracle import bindings as ob
    threads = os.cpu_count() or 1
    tb = ob.HIBFTables(tables["n_user_bins"], tables["bins"], tables["bin_size"], tables["hash_funs"], tables["words"],
                       tables["next_ibf"], tables["user_bin"])
    if ob.have_ref():
        kind, probe = "reference", ob.RefHIBF.from_tables(tb).membership
    else:
        kind, probe = "port", tb.membership
    n = 1 << 16
    t0 = time.perf_counter()
    probe(q_host[:n], threads)
    dt = time.perf_counter() - t0
    n = int(min(len(q_host), max(n, n * budget_s / max(dt, 1e-6) / 2)))
    t0 = time.perf_counter()
    probe(q_host[:n], threads)
    dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": UNIT, "cores": threads, "kind": kind,
            "sample": f"first {n} k-mers of the step's batch, one membership agent per thread, OpenMP static"}


def extra_measurements(t, ctx, torch, residues, offs, rec_bin, planted, hibf, peak):
    """Side measurements with CUDA events: K4 verify scan (streaming roofline), K2 flat-IBF probe (C3 shape),
    K1 emplace, and whole regex queries/s for the config's query."""
    from oracle import bindings as ob
    out = {}
    stream = torch.cuda.current_stream()

    def timed(fn, reps):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps / 1e3

    # ---- K4: scan every residue of every bin for the config's query (all bins as candidates)
    txt = t.Text(ctx, residues, offs, rec_bin, N_BINS)
    rx = t.Regex(QUERY, True)
    hits = txt.scan(rx)
    want = sorted((b, r, s, e) for b, r, s, e in planted)
    out["verify_hits_match_planted"] = sorted(map(tuple, hits.tolist())) == want
    sec = timed(lambda: txt.scan(rx), 5)
    n_res = int(offs[-1])
    out["verify_residues_per_s"] = n_res / sec
    out["verify_roofline"] = {"bound": "hbm", "achieved": n_res / sec / 1e9, "peak": peak, "unit": "GB/s",
                              "frac": n_res / sec / 1e9 / peak, "note": "1 B per residue, all 1024 bins scanned, "
                              "includes phase 2, flag clears and the hit read-back"}
    if ob.have_ref():
        n_rec = min(len(offs) - 1, 60000)
        t0 = time.perf_counter()
        ob.ref_scan_records(QUERY, True, residues, offs[:n_rec + 1], os.cpu_count() or 1)
        dt = time.perf_counter() - t0
        out["verify_cpu_residues_per_s"] = {"value": int(offs[n_rec]) / dt, "cores": os.cpu_count() or 1,
                                            "kind": "reference (RE2 FindAndConsume, records in memory, no gz/kseq)",
                                            "sample": f"first {n_rec} records"}

    # ---- K4 on a PROSITE-style pattern (C2H2 zinc finger, PS00028): no fixed-offset literal window exists; phase 1 is
    # the gapped-atom prefilter (literal masks + shifted mask algebra), TRX_REGEX_FILTER=off = the exact automaton kernel
    prosite = "C.{2,4}C.{3}[LIVMFYWC].{8}H.{3,5}H"
    rx2 = t.Regex(prosite, True)
    h2 = txt.scan(rx2)
    sec2 = timed(lambda: txt.scan(rx2), 5)
    out["verify_prosite"] = {"pattern": prosite, "hits": int(len(h2)), "filter_atoms": int(txt.last_filter_literals),
                             "residues_per_s": n_res / sec2, "frac_of_hbm_peak": n_res / sec2 / 1e9 / peak}
    os.environ["TRX_REGEX_FILTER"] = "off"
    h2x = txt.scan(rx2)
    sec2x = timed(lambda: txt.scan(rx2), 2)
    os.environ.pop("TRX_REGEX_FILTER")
    out["verify_prosite"]["exact_automaton_residues_per_s"] = n_res / sec2x
    out["verify_prosite"]["hits_equal_exact_automaton"] = bool(np.array_equal(h2, h2x))
    out["verify_prosite"]["exact_automaton_note"] = ("TRX_REGEX_FILTER=off; this pattern's reverse searcher is too large for "
                                                     "the starts kernel: position automaton in phase 1, DFA in phase 2")
    # the config's own query without a prefilter: small reverse searcher → regex_scan_starts_dfa (exact match starts, one
    # table lookup per residue) + exact-start phase 2
    os.environ["TRX_REGEX_FILTER"] = "off"
    h1x = txt.scan(rx)
    sec1x = timed(lambda: txt.scan(rx), 3)
    os.environ.pop("TRX_REGEX_FILTER")
    out["verify_no_prefilter"] = {"pattern": QUERY, "residues_per_s": n_res / sec1x, "us_per_scan": sec1x * 1e6,
                                  "hits_equal_prefiltered_scan": bool(np.array_equal(h1x, hits)),
                                  "route": "reverse-searcher DFA phase 1 + exact-start phase 2"}
    os.environ["TRX_REGEX_FILTER"] = "off"
    os.environ["TRX_REGEX_DFA"] = "off"
    h2n = txt.scan(rx2)
    sec2n = timed(lambda: txt.scan(rx2), 2)
    os.environ.pop("TRX_REGEX_FILTER")
    os.environ.pop("TRX_REGEX_DFA")
    out["verify_prosite"]["position_automaton_residues_per_s"] = n_res / sec2n
    out["verify_prosite"]["hits_equal_position_automaton"] = bool(np.array_equal(h2, h2n))
    # the densest pattern of the config-3 batch: no necessary factor at all, 513 k hits
    rx3 = t.Regex("...[GHKM]V[FHKL].", True)
    h3 = txt.scan(rx3, cap=1 << 21)
    sec3 = timed(lambda: txt.scan(rx3, cap=1 << 21), 3)
    out["verify_dense"] = {"pattern": "...[GHKM]V[FHKL].", "hits": int(len(h3)), "residues_per_s": n_res / sec3,
                           "us_per_scan": sec3 * 1e6, "note": "whole call incl. sorting and returning the hits to the host"}
    txt.close()

    # ---- K4 at a size where the per-call latencies (one H2D, one memset, two launches, one D2H, one sync ~ 50 us) no longer
    # dominate: the same collection 8 times over (1.6 G residues, 8192 bins) — the streaming rate of the kernels
    reps8 = 8
    big_res = np.tile(np.asarray(residues, dtype=np.uint8), reps8)
    n_rec = len(offs) - 1
    big_offs = np.concatenate([np.asarray(offs[:-1], dtype=np.uint64) + np.uint64(i * n_res) for i in range(reps8)] +
                              [np.array([reps8 * n_res], dtype=np.uint64)])
    big_bin = np.concatenate([np.asarray(rec_bin, dtype=np.uint32) + np.uint32(i * N_BINS) for i in range(reps8)])
    big = t.Text(ctx, big_res, big_offs, big_bin, N_BINS * reps8)
    del big_res
    large = {}
    for name, rxx in (("window", rx), ("atoms", rx2)):
        hb = big.scan(rxx)
        secb = timed(lambda: big.scan(rxx), 3)
        large[name] = {"pattern": rxx.pattern, "hits": int(len(hb)), "residues_per_s": reps8 * n_res / secb,
                       "frac_of_hbm_peak": reps8 * n_res / secb / 1e9 / peak, "us_per_scan": secb * 1e6}
    large["residues"] = reps8 * n_res
    large["records"] = reps8 * n_rec
    out["verify_large"] = large
    big.close()
    txt = t.Text(ctx, residues, offs, rec_bin, N_BINS)

    # ---- K4 over BASELINE config 2's pattern population: 1000 PROSITE-style patterns, every bin scanned for each
    from tetrex_b200 import synth
    pats = synth.prosite_patterns(1000, 1000)
    kinds = {"window": 0, "atoms": 0, "exact": 0}
    freq = np.bincount(residues[:1 << 24], minlength=256).astype(np.float64)
    freq /= freq.sum()
    compiled = []
    for pat in pats:
        rxp = t.Regex(pat, True)
        kind = rxp.filter_plan(freq)["kind"]
        kinds["exact" if kind == "none" else kind] += 1
        compiled.append((rxp, kind))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    total_hits, overflow, per_kind_s = 0, 0, {"window": 0.0, "atoms": 0.0, "none": 0.0}
    for rxp, kind in compiled:
        t1 = time.perf_counter()
        try:
            total_hits += len(txt.scan(rxp))
        except t.TetrexError:
            overflow += 1
        per_kind_s[kind] += time.perf_counter() - t1
    dt = time.perf_counter() - t0
    # the same batch with three scans in flight (trx_regex_scan_begin/_end): kernels back to back, host latency hidden
    only_rx = [rxp for rxp, _ in compiled]
    txt.scan_many(only_rx[:8], cap=1 << 20)
    torch.cuda.synchronize()
    dt_piped = float("inf")
    for _ in range(2):  # best of two: a 0.1 s host-driven loop is easily disturbed
        t0 = time.perf_counter()
        piped = txt.scan_many(only_rx, cap=1 << 20, depth=3)
        dt_piped = min(dt_piped, time.perf_counter() - t0)
    out["verify_pattern_batch_pipelined"] = {"depth": 3, "scans_per_s": len(pats) / dt_piped,
                                             "residues_per_s": len(pats) * n_res / dt_piped,
                                             "frac_of_hbm_peak": len(pats) * n_res / dt_piped / 1e9 / peak,
                                             "hits": int(sum(len(h) for h in piped))}
    out["verify_pattern_batch"] = {
        "patterns": len(pats), "grammar": "SURVEY 8(d) C3: 6-14 elements, 60 % literal / 25 % class / 10 % x / 5 % (2,4) gaps",
        "phase1_kernel": kinds, "scans_per_s": len(pats) / dt, "residues_per_s": len(pats) * n_res / dt,
        "frac_of_hbm_peak": len(pats) * n_res / dt / 1e9 / peak, "hits": int(total_hits), "over_capacity": overflow,
        "mean_us_per_scan": {("exact" if k == "none" else k): (per_kind_s[k] / max(kinds["exact" if k == "none" else k], 1)) * 1e6
                             for k in per_kind_s}}
    pattern_hits = {pat: int(len(h)) for pat, h in zip(pats, piped)}
    # N4: the same batch in shared passes over the text (trx_regex_scan_multi: <= 128 patterns per pass)
    txt.scan_multi(only_rx[:128], cap=1 << 21)
    torch.cuda.synchronize()
    dt_multi = float("inf")
    for _ in range(3):
        t0 = time.perf_counter()
        multi, mst = txt.scan_multi(only_rx, cap=1 << 21, stats=True)
        dt_multi = min(dt_multi, time.perf_counter() - t0)
    out["verify_pattern_batch_multi"] = {
        "scans_per_s": len(pats) / dt_multi, "ms_total": dt_multi * 1e3, "hits": int(sum(len(h) for h in multi)),
        "hit_lists_equal_pattern_at_a_time": bool(all(np.array_equal(a, b) for a, b in zip(multi, piped))),
        "patterns_on_shared_passes": mst["shared"], "patterns_one_at_a_time": mst["single"], "shared_passes": mst["passes"],
        "text_bytes_read_per_pattern": mst["residues"] / max(mst["shared"], 1),
        "speedup_vs_pipelined": dt_piped / dt_multi,
        "c_call_ms": txt.last_multi_call_seconds * 1e3,
        "note": "one TMA pass over the 204 MB text per batch of <= 128 patterns instead of one per pattern: first stage "
                "regex_window_multi (class windows, bit-parallel over the patterns), hits sorted on the device, one copy out"}
    for rxp, _ in compiled:
        rxp.close()

    # ---- whole query: k-mers of the query -> K3 -> AND/OR on device -> K4 on the candidate bins
    aamap, _ = t.peptide_maps(t.RED_BASE)

    def pack(sq):
        v = 0
        for ch in sq:
            v = (v << 5) | int(aamap[ch])
        return v
    paths = [b"LMAEGLYN", b"LMAQGLYN"]
    kmers = np.array([pack(pth[i:i + K]) for pth in paths for i in range(len(pth) - K + 1)], dtype=np.uint64)

    def query_once():
        hv = hibf.membership(kmers)
        per_path = hv.reshape(len(paths), -1, hv.shape[1])
        cand = np.bitwise_or.reduce(np.bitwise_and.reduce(per_path, axis=1), axis=0)
        return cand, txt.scan(rx, candidate_bins=cand)
    cand, qhits = query_once()
    out["query_candidate_bins"] = int(np.bitwise_count(cand).sum())
    out["query_hits_match_planted"] = sorted(map(tuple, qhits.tolist())) == want
    t0 = time.perf_counter()
    reps = 50
    for _ in range(reps):
        query_once()
    out["regex_queries_per_s"] = reps / (time.perf_counter() - t0)
    out["regex_query"] = QUERY + " (filter on GPU via K3 + path AND/OR of its 6 k-mers on host words, verify via K4)"
    txt.close()

    # ---- the same query through the reference-shaped host path (C++: process_query = k-graph + collector with the path
    # algebra on the device, iter_disk_search = K4 + line formatting) on index files built by this repo's `tetrex index`
    try:
        import bench_configs as bc
        c3 = {}

        def on_index(name, ix, files, workdir):
            if name == "flat":
                c3["pattern_batch_filter_and_verify"] = bc.measure_c3_patterns(ix, pats, lambda p: pattern_hits[p], budget_s=30.0)
                # the same batch the way SURVEY.md Appendix B says config 3 is meant to run: `-a` (high-fan-out stretches
                # become Gap nodes) with a DGram index built on the device (`tetrex dindex`, gaps 3..21) as `-g`
                c3["pattern_batch_augmented"] = bc.measure_c3_augmented(t, ctx, ix, files, workdir, pats,
                                                                         lambda p: pattern_hits[p])
        out["host_path_queries"] = host_path_queries(t, ctx, residues, offs, rec_bin, on_index=on_index)
        out["C3"] = c3
    except Exception as e:  # never lose the bench line over a side measurement
        out["host_path_queries"] = {"error": f"{type(e).__name__}: {e}"[:300]}

    # ---- K2 / K1 on the flat C3 shape: 1024 bins, bin_size from the reference's sizing rule
    lens = (offs[1:] - offs[:-1]).astype(np.int64)
    per_bin = np.bincount(rec_bin, weights=np.maximum(lens - K + 1, 0), minlength=N_BINS)
    bin_size = t.compute_bitcount(int(per_bin.max()), FPR)
    flat = t.IBF(ctx, N_BINS, bin_size, HASH_FUNS)
    d_res = torch.from_numpy(residues).cuda()
    d_offs = torch.from_numpy(offs.view(np.int64)).cuda()
    d_bin = torch.from_numpy(rec_bin.view(np.int32)).cuda()
    n_emits = int(per_bin.sum())
    sec = timed(lambda: flat.emplace_records_dev(t.MOL_AA, K, 0, d_res.data_ptr(), len(residues), d_offs.data_ptr(),
                                                 d_bin.data_ptr(), len(rec_bin)), 3)
    out["emplace_kmers_per_s"] = n_emits / sec
    out["emplace_note"] = f"flat IBF 1024 bins x {bin_size} bits ({flat.n_words * 8 / 1e6:.0f} MB), {n_emits} k-mers, h=3"
    n = 1 << 22
    q = torch.randint(0, 1 << 30, (n,), dtype=torch.int64, device="cuda")
    o = torch.empty(n * flat.bin_words, dtype=torch.int64, device="cuda")
    sec = timed(lambda: flat.bulk_contains_dev(q.data_ptr(), n, o.data_ptr()), 10)
    algo = n * HASH_FUNS * flat.bin_words * 8
    out["flat_probes_per_s"] = n / sec
    out["flat_roofline"] = {"bound": "hbm", "achieved": algo / sec / 1e9, "peak": peak, "unit": "GB/s",
                            "frac": algo / sec / 1e9 / peak, "note": "h x bin_words x 8 B read per k-mer (384 B); the "
                            "154 MB matrix is mostly L2-resident, result write (128 B/k-mer) not counted"}
    c3 = out.setdefault("C3", {})
    c3["workload"] = (f"peptide flat IBF, 1024 bins x {bin_size} rows ({flat.n_words * 8 / 1e6:.0f} MB), k=6, h=3; "
                      "1000 PROSITE-style patterns")
    c3["k2_probes_per_s"] = n / sec
    c3["roofline"] = dict(out["flat_roofline"], traffic=None, kernel="probe_kernel<3,2,8,0>")
    c3["k1_kmers_per_s"] = out["emplace_kmers_per_s"]
    c3["verify_batch_all_bins"] = {"scans_per_s_multi_pattern_pass": out["verify_pattern_batch_multi"]["scans_per_s"],
                                   "multi_equals_pattern_at_a_time": out["verify_pattern_batch_multi"]["hit_lists_equal_pattern_at_a_time"],
                                   "scans_per_s_pipelined": out["verify_pattern_batch_pipelined"]["scans_per_s"],
                                   "scans_per_s_one_at_a_time": out["verify_pattern_batch"]["scans_per_s"],
                                   "hits": out["verify_pattern_batch"]["hits"]}
    try:
        import bench_configs as bc
        c3["cpu_baseline"] = bc.cpu_baseline_flat_probe(N_BINS, bin_size, HASH_FUNS, q.cpu().numpy().view(np.uint64))
    except Exception as e:
        c3["cpu_baseline"] = {"error": f"{type(e).__name__}: {e}"[:200]}
    flat.close()
    del q, o
    # ---- DGram index build on the device (`tetrex dindex`, include/dGramIndex.h:103-154,194-243): gaps 3..21 over the
    # same residues — count per bin, size like compute_bitcount, emplace every gapped 3+3-mer code
    try:
        import ctypes as C
        from tetrex_b200._abi import check
        d_cnt = torch.zeros(N_BINS, dtype=torch.int64, device="cuda")
        lo, hi = 3, 21
        cnt_s = timed(lambda: check(t.lib().trx_dgram_count_dev(ctx._h, d_res.data_ptr(), len(residues), d_offs.data_ptr(),
                                                                 d_bin.data_ptr(), len(rec_bin), lo, hi, N_BINS, d_cnt.data_ptr())), 3)
        codes = int(d_cnt.sum().item())
        dsize = t.compute_bitcount(int(d_cnt.max().item()), FPR)
        dflat = t.IBF(ctx, N_BINS, dsize, HASH_FUNS)
        emp_s = timed(lambda: check(t.lib().trx_dgram_emplace_dev(dflat._h, d_res.data_ptr(), len(residues), d_offs.data_ptr(),
                                                                   d_bin.data_ptr(), len(rec_bin), lo, hi)), 2)
        algo = codes * (HASH_FUNS * 64) + len(residues)
        c3["dgram_build"] = {"gaps": [lo, hi], "codes": codes, "matrix_gb": dflat.n_words * 8 / 1e9,
                             "count_ms": cnt_s * 1e3, "emplace_ms": emp_s * 1e3, "codes_per_s": codes / emp_s,
                             "roofline": {"bound": "hbm", "achieved": algo / emp_s / 1e9, "peak": peak, "unit": "GB/s",
                                          "frac": algo / emp_s / 1e9 / peak, "traffic": None, "kernel": "dgram_kernel<false>",
                                          "note": "h x 64 B per code (32 B sector fill + write-back per atomicOr into a "
                                                  "matrix >> L2, the K1 figure of SURVEY 8(d)) + 1 B per residue"}}
        dflat.close()
    except Exception as e:
        c3["dgram_build"] = {"error": f"{type(e).__name__}: {e}"[:200]}
    return out


def run_reference(args):
    """The reference's own CPU path for the same metric/config; rank 0 only, no CUDA anywhere."""
    if int(os.environ.get("RANK", 0)) != 0:
        return
    from oracle import bindings as ob
    from tetrex_b200 import synth
    threads = os.cpu_count() or 1
    residues, offs, rec_bin, _ = synth.swissprot_like(args.seqs, N_BINS)
    tables = synth.build_swissprot_hibf(lambda *a: OracleIbfAdapter(*a, threads), residues, offs, rec_bin, N_BINS, K,
                                        HASH_FUNS, FPR)
    tb = ob.HIBFTables(tables["n_user_bins"], tables["bins"], tables["bin_size"], tables["hash_funs"], tables["words"],
                       tables["next_ibf"], tables["user_bin"])
    aamap, _ = ob.peptide_maps(0)
    q = make_queries(collection_kmers(residues, offs, aamap), args.batch, 1234)
    if ob.have_ref():
        kind, probe = "reference", ob.RefHIBF.from_tables(tb).membership
    else:
        kind, probe = "port", tb.membership
    # bounded sample per step: sized from a probe run so that steps+warmup stay within ~2 minutes
    n0 = 1 << 15
    t0 = time.perf_counter()
    probe(q[:n0], threads)
    rate = n0 / (time.perf_counter() - t0)
    per_step = int(min(len(q), max(n0, rate * 100.0 / (args.steps + args.warmup))))
    for _ in range(args.warmup):
        probe(q[:per_step], threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        probe(q[:per_step], threads)
    dt = time.perf_counter() - t0
    value = per_step * args.steps / dt
    sample = f"{per_step} k-mers per step (prefix of the same seeded batch), one membership agent per thread"
    emit({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": "C2: synthetic Swissprot-shaped peptide HIBF, 1024 user bins, k=6, h=3, FPR 0.05",
                   "sequences": args.seqs, "residues": int(offs[-1]), "kmers_per_step": per_step,
                   "index": "same layout as the GPU arm, bits set by the CPU oracle (no CUDA in this arm)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


_REAL_STDOUT = None


def claim_stdout():
    """stdout must carry exactly one JSON line: libraries that write to fd 1 (NCCL prints its version banner
    there) are sent to stderr for the whole run; emit() writes to the saved descriptor."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(line: dict):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


if __name__ == "__main__":
    a = parse_args()
    claim_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
