"""Breaks one trx_regex_scan call into its parts with CUDA events / host clocks (diagnostic)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tetrex_b200 as t
from tetrex_b200 import synth
ctx = t.Context(0)
residues, offs, rec_bin, planted = synth.swissprot_like(570_000, 1024)
txt = t.Text(ctx, residues, offs, rec_bin, 1024)
pats = sys.argv[1:] or ["LMA(E|Q)GLYN", "G.{2,4}K[ILV]", "C.{2,4}C.{3}[LIVMFYWC].{8}H.{3,5}H", "W.{2,4}C.{5,9}CW",
                        "C.{2}C.{12}H.{3,5}H.{2}M", "D.{0,3}EW", "M.{2}(C|W)H.{3,4}(WK|WR)C"]
for pat, posix in [(p, True) for p in pats]:
    rx = t.Regex(pat, posix)
    for env in ("", "off"):
        if env:
            os.environ["TRX_REGEX_FILTER"] = env
        else:
            os.environ.pop("TRX_REGEX_FILTER", None)
        hits = txt.scan(rx)
        t0 = time.perf_counter()
        for _ in range(20):
            hits = txt.scan(rx)
        dt = (time.perf_counter() - t0) / 20
        print(f"{pat:40s} filter={'on ' if not env else 'off'} literals={txt.last_filter_literals} hits={len(hits):7d} "
              f"{dt*1e6:8.1f} us/scan  {len(residues)/dt/1e9:8.1f} GB/s  ({len(residues)/dt/1e9/6553.3*100:.1f}% of HBM peak)")
os.environ.pop("TRX_REGEX_FILTER", None)
