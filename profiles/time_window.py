"""trx_regex_scan_multi on the 1000-pattern batch (config 3 patterns, C2 text): class-window first stage
(regex_window_multi) against the term-list first stage (TRX_MULTI_WINDOW=off) — same hits, time per batch."""
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
def timed(fn, reps=3):
    fn(); ctx.sync()
    best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); ctx.sync()
        best = min(best, time.perf_counter() - t0)
    return best
res = {}
for mode in ("on", "off"):
    os.environ["TRX_MULTI_WINDOW"] = mode
    os.environ["TRX_SCAN_TRACE"] = "1" if mode == "on" else ""
    if not os.environ["TRX_SCAN_TRACE"]:
        del os.environ["TRX_SCAN_TRACE"]
    got, st = txt.scan_multi(regs, cap=1 << 21, stats=True)
    res[mode] = got
    dt = timed(lambda: txt.scan_multi(regs, cap=1 << 21))
    d1 = timed(lambda: txt.scan_multi(regs[:128], cap=1 << 21))
    timed(lambda: txt.scan_multi(regs, cap=1 << 21), 1)
    print(f"window {mode}: 1000 patterns {dt*1e3:.1f} ms = {1000/dt:.0f} scans/s; first 128: {d1*1e3:.2f} ms; stats {st}; C call {txt.last_multi_call_seconds*1e3:.2f} ms", flush=True)
same = all(np.array_equal(a, b) for a, b in zip(res["on"], res["off"]))
print("same hits:", same, "total", sum(len(g) for g in res["on"]))
assert same
for r in regs:
    r.close()
txt.close()
ctx.close()
print("closed", flush=True)
