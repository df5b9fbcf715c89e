"""The patterns of the config-3 batch that trx_regex_scan_multi sends down the single-pattern path: plan, hits, time per scan.
Run under `ncu --metrics gpu__time_duration.sum` for the per-kernel split (phase 1 / phase 2 / hit sort)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tetrex_b200 as t
from tetrex_b200 import synth
ctx = t.Context(0)
residues, offs, rec_bin, planted = synth.swissprot_like(570_000, 1024)
txt = t.Text(ctx, residues, offs, rec_bin, 1024)
pats = synth.prosite_patterns(1000, 1000)
regs = [t.Regex(p, True) for p in pats]
freq = np.bincount(np.asarray(residues, dtype=np.uint8), minlength=256).astype(np.float64)
freq /= freq.sum()
single = []
for i, r in enumerate(regs):
    h, st = txt.scan_multi([r], cap=1 << 21, stats=True)
    if st["single"]:
        single.append(i)
quick = "--once" in sys.argv
for i in single:
    plan = regs[i].filter_plan(freq)
    plan = plan["kind"]
    hits = txt.scan(regs[i], cap=1 << 21)
    ctx.sync()
    if quick:
        print(i, pats[i], len(hits))
        continue
    t0 = time.perf_counter()
    for _ in range(5):
        txt.scan(regs[i], cap=1 << 21)
    ctx.sync()
    print(f"{i:4d} {pats[i]:40s} hits {len(hits):7d}  {(time.perf_counter() - t0) / 5 * 1e6:8.0f} us/scan  plan {plan}")
