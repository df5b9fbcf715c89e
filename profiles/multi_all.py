"""All 1000 config-3 patterns through trx_regex_scan_multi on the C2 text, twice (for ncu launch lists); SINGLES=1 also
times the patterns the shared passes leave to the single-pattern path."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tetrex_b200 as t
from tetrex_b200 import synth
ctx = t.Context(0)
residues, offs, rec_bin, planted = synth.swissprot_like(570_000, 1024)
txt = t.Text(ctx, residues, offs, rec_bin, 1024)
pats = synth.prosite_patterns(1000, 1000)
regs = [t.Regex(p, True) for p in pats]
for _ in range(2):
    h, st = txt.scan_multi(regs, cap=1 << 21, stats=True)
print(st, sum(len(x) for x in h), "hits")
if os.environ.get("SINGLES"):
    for i, r in enumerate(regs):
        _, s1 = txt.scan_multi([r], cap=1 << 21, stats=True)
        if s1["single"]:
            txt.scan(r, cap=1 << 21)
            t0 = time.perf_counter(); n = len(txt.scan(r, cap=1 << 21)); dt = time.perf_counter() - t0
            print(f"single {i:4d} {dt*1e3:7.2f} ms {n:8d} hits  {pats[i]}")
txt.close(); ctx.close()
