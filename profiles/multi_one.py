"""One batch of 128 config-3 patterns through trx_regex_scan_multi on the C2 text (for ncu launch lists / captures)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tetrex_b200 as t
from tetrex_b200 import synth
ctx = t.Context(0)
residues, offs, rec_bin, planted = synth.swissprot_like(570_000, 1024)
txt = t.Text(ctx, residues, offs, rec_bin, 1024)
first = int(os.environ.get("FIRST", 0))
regs = [t.Regex(p, True) for p in synth.prosite_patterns(1000, 1000)[first:first + 128]]
for _ in range(2):
    h, st = txt.scan_multi(regs, cap=1 << 21, stats=True)
print(st, sum(len(x) for x in h), "hits")
