"""One traced trx_regex_scan per pattern (TRX_SCAN_TRACE=1 → host-clock breakdown on stderr) + per-kernel device times."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["TRX_SCAN_TRACE"] = "1"
import tetrex_b200 as t
from tetrex_b200 import synth
ctx = t.Context(0)
residues, offs, rec_bin, planted = synth.swissprot_like(570_000, 1024)
txt = t.Text(ctx, residues, offs, rec_bin, 1024)
for pat in ["LMA(E|Q)GLYN", "C.{2,4}C.{3}[LIVMFYWC].{8}H.{3,5}H"]:
    rx = t.Regex(pat, True)
    for _ in range(6):
        txt.scan(rx)
    sys.stderr.flush()
