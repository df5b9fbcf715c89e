"""Where the time of trx_regex_scan_multi goes on the 1000-pattern batch (config 3 patterns, C2 text)."""
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
txt.scan_multi(regs[:128], cap=1 << 21)
single, shared = [], []
for i, r in enumerate(regs):
    h, st = txt.scan_multi([r], cap=1 << 21, stats=True)
    (single if st["single"] else shared).append(i)
    if st["single"]:
        print("single:", i, pats[i], len(h[0]), "hits")
def timed(fn, reps=3):
    fn(); ctx.sync()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    ctx.sync()
    return (time.perf_counter() - t0) / reps
a = timed(lambda: txt.scan_multi([regs[i] for i in shared], cap=1 << 21))
b = timed(lambda: txt.scan_many([regs[i] for i in single], cap=1 << 21, depth=3), 1)
c = timed(lambda: txt.scan_multi([regs[i] for i in shared[:128]], cap=1 << 21))
print(f"{len(shared)} shared-pass patterns: {a*1e3:.1f} ms = {a/len(shared)*1e6:.1f} us/pattern; one batch of 128: {c*1e3:.2f} ms; "
      f"{len(single)} single-path patterns pipelined: {b*1e3:.1f} ms")
