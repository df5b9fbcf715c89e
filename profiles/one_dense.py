"""One scan of the densest config-3 pattern (513 k hits, no necessary factor): reverse-searcher phase 1 + exact-start phase 2."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tetrex_b200 as t
from tetrex_b200 import synth
ctx = t.Context(0)
residues, offs, rec_bin, planted = synth.swissprot_like(570_000, 1024)
txt = t.Text(ctx, residues, offs, rec_bin, 1024)
rx = t.Regex(sys.argv[1] if len(sys.argv) > 1 else "...[GHKM]V[FHKL].", True)
for _ in range(3):
    h = txt.scan(rx, cap=1 << 21)
print(len(h), "hits")
