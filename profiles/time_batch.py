"""The 1000-pattern PROSITE-style batch (config 3) over all bins of the C2 text: phase-1 census and scans/s, sequential and
pipelined.  TRX_FILTER_MAX_PASS=x python profiles/time_batch.py  (A/B of the prefilter admission threshold)"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tetrex_b200 as t
from tetrex_b200 import synth
ctx = t.Context(0)
residues, offs, rec_bin, planted = synth.swissprot_like(570_000, 1024)
txt = t.Text(ctx, residues, offs, rec_bin, 1024)
freq = np.bincount(residues[:1 << 24], minlength=256).astype(np.float64); freq /= freq.sum()
pats = synth.prosite_patterns(1000, 1000)
regs = [t.Regex(p, True) for p in pats]
kinds = [r.filter_plan(freq)["kind"] for r in regs]
txt.scan_many(regs[:8], cap=1 << 20)
per = {"window": 0.0, "atoms": 0.0, "none": 0.0}
t0 = time.perf_counter()
hits = 0
for r, k in zip(regs, kinds):
    t1 = time.perf_counter(); hits += len(txt.scan(r)); per[k] += time.perf_counter() - t1
dt = time.perf_counter() - t0
t0 = time.perf_counter(); piped = txt.scan_many(regs, cap=1 << 20, depth=int(os.environ.get("DEPTH", 3))); dtp = time.perf_counter() - t0
cnt = {k: kinds.count(k) for k in per}
print(f"max_pass={os.environ.get('TRX_FILTER_MAX_PASS', 'default')}: census {cnt}  sequential {len(pats)/dt:7.0f} scans/s  pipelined {len(pats)/dtp:7.0f} scans/s  "
      f"mean us: " + ", ".join(f"{k} {per[k]/max(cnt[k],1)*1e6:.0f}" for k in per) + f"  hits {hits} / {sum(len(h) for h in piped)}")
# N4: shared passes over the text (trx_regex_scan_multi)
t0 = time.perf_counter(); txt.scan_multi(regs, cap=1 << 21); dt_first = time.perf_counter() - t0  # derives the multi plans
t0 = time.perf_counter(); multi, st = txt.scan_multi(regs, cap=1 << 21, stats=True); dtm = time.perf_counter() - t0
print(f"multi first call (plans derived) {dt_first*1e3:.1f} ms")
same = all(np.array_equal(a, b) for a, b in zip(multi, piped))
print(f"multi-pattern passes: {len(pats)/dtm:7.0f} scans/s ({dtm*1e3:.1f} ms for the batch, {st}), hit lists equal pipelined: {same}, "
      f"launches so far {ctx.launches}")
