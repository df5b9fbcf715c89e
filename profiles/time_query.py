"""Per-query latency of the reference-shaped host path (libtrxhost.so: process_query = k-graph + collector on the device,
iter_disk_search = K4 + line formatting) on the 256-bin test collection: python profiles/time_query.py [flat|hibf]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import tetrex_b200 as t
from tetrex_b200 import sharded
from sharded_collection import build
coll = build("/tmp/tq_coll")
ctx = t.Context(0)
for kind in (sys.argv[1:] or ["flat", "hibf"]):
    ix = sharded.ShardedIndex(ctx, coll[kind], 0, 1)
    for regex in ["LMA(E|Q)GLYN", "L.AQGLYN", "CAHCWKKH"]:
        w, st = ix.local_candidates(regex)
        ix.verify_local(w, regex)
        n = 200
        t0 = time.perf_counter()
        for _ in range(n):
            w, st = ix.local_candidates(regex)
        t1 = time.perf_counter()
        for _ in range(n):
            ix.verify_local(w, regex)
        t2 = time.perf_counter()
        print(f"{kind:5s} {regex:14s} process_query {1e6 * (t1 - t0) / n:8.1f} us ({st['ranks']} ranks, {st['probes']} probes, "
              f"{st['items']} items)   iter_disk_search {1e6 * (t2 - t1) / n:8.1f} us   -> {n / (t2 - t0):7.0f} queries/s")
    ix.close()
