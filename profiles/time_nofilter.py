"""Patterns scanned WITHOUT a prefilter (TRX_REGEX_FILTER=off): the no-prefilter routes against the position automaton."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tetrex_b200 as t
from tetrex_b200 import synth
ctx = t.Context(0)
residues, offs, rec_bin, planted = synth.swissprot_like(570_000, 1024)
txt = t.Text(ctx, residues, offs, rec_bin, 1024)
n_res = int(offs[-1])
def timed(rx, reps=5, **env):
    for k, v in env.items(): os.environ[k] = v
    h = txt.scan(rx, cap=1 << 22); ctx.sync()
    t0 = time.perf_counter()
    for _ in range(reps): txt.scan(rx, cap=1 << 22)
    ctx.sync(); dt = (time.perf_counter() - t0) / reps
    for k in env: os.environ.pop(k)
    return len(h), dt
for pat in ["C.{2,4}C.{3}[LIVMFYWC].{8}H.{3,5}H", "LMA(E|Q)GLYN", "...[GHKM]V[FHKL].", ".[ACDEL]D.{2,4}.[AEGQT]S", "[ST].[RK]W", "W.{3}W"]:
    rx = t.Regex(pat, True)
    n0, d0 = timed(rx)
    n1, d1 = timed(rx, TRX_REGEX_FILTER="off")
    n2, d2 = timed(rx, TRX_REGEX_FILTER="off", TRX_REGEX_DFA="off")
    n3, d3 = timed(rx, TRX_REGEX_FILTER="off", TRX_REGEX_STARTS="off")
    assert n0 == n1 == n2 == n3
    print(f"{pat:40s} hits {n0:7d}  default {d0*1e6:7.0f} us | no prefilter {d1*1e6:7.0f} us = {n_res/d1/1e12:.2f} T res/s | "
          f"position automaton {d2*1e6:7.0f} us | every start {d3*1e6:7.0f} us")
