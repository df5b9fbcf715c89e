"""Runs each hot kernel of the path once or twice at full C2/C3 size in a fixed order, for ncu:
setup (65 emplace launches, skipped with -s 65), then [K1 emplace, K3 root+level, K4 phase1+phase2, K2 probe] x reps.
Usage on the GPU box (see profiles/README.md):  python profiles/prof_driver.py [reps]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tetrex_b200 as t  # noqa: E402
from tetrex_b200 import synth  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
K, H, FPR, NB = 6, 3, 0.05, 1024
s = torch.cuda.Stream()
torch.cuda.set_stream(s)
ctx = t.Context(0, s.cuda_stream)
residues, offs, rec_bin, planted = synth.swissprot_like(570_000, NB)
tables = synth.build_swissprot_hibf(lambda *a: t.IBF(ctx, *a), residues, offs, rec_bin, NB, K, H, FPR)
hibf = t.HIBF(ctx, tables["n_user_bins"], tables["bins"], tables["bin_size"], tables["hash_funs"], tables["words"],
              tables["next_ibf"], tables["user_bin"])
lens = (offs[1:] - offs[:-1]).astype(np.int64)
per_bin = np.bincount(rec_bin, weights=np.maximum(lens - K + 1, 0), minlength=NB)
flat = t.IBF(ctx, NB, t.compute_bitcount(int(per_bin.max()), FPR), H)
d_res = torch.from_numpy(residues).cuda()
d_offs = torch.from_numpy(offs.view(np.int64)).cuda()
d_bin = torch.from_numpy(rec_bin.view(np.int32)).cuda()
n = 1 << 20
rng = np.random.default_rng(0)
q = torch.from_numpy(rng.integers(0, 1 << 30, n, dtype=np.int64)).cuda()
o = torch.empty(n * 16, dtype=torch.int64, device="cuda")
txt = t.Text(ctx, residues, offs, rec_bin, NB)
rx = t.Regex("LMA(E|Q)GLYN", True)
torch.cuda.synchronize()
for _ in range(reps):
    flat.emplace_records_dev(t.MOL_AA, K, 0, d_res.data_ptr(), len(residues), d_offs.data_ptr(), d_bin.data_ptr(), len(rec_bin))
    hibf.membership_dev(q.data_ptr(), n, o.data_ptr())
    hits = txt.scan(rx)
    flat.bulk_contains_dev(q.data_ptr(), n, o.data_ptr())
    torch.cuda.synchronize()
print("prof_driver ok", len(hits), ctx.launches)
