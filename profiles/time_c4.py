"""K1 at the C4 shape without the rest of the bench: python profiles/time_c4.py [bins]  (env knobs: TRX_EMPLACE_WINDOW_MB,
TRX_EMPLACE_SLAB).  Builds `bins` bins (default 1024 = 16 slabs, 2.9 GB matrix) and prints k-mers/s + the parity checks."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench_configs as bc  # noqa: E402
import tetrex_b200 as t  # noqa: E402

bins = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
s = torch.cuda.Stream()
torch.cuda.set_stream(s)
ctx = t.Context(0, s.cuda_stream)
ident = lambda x: x
out, f, keep, bs = bc.measure_c4(t, ctx, torch, 0, 1, torch.cuda.synchronize, ident, ident, 6550.7, None, bins=bins)
k1 = out["k1_build"]
print(json.dumps({"bins": bins, "window_mb": os.environ.get("TRX_EMPLACE_WINDOW_MB", "64"), "slab": os.environ.get("TRX_EMPLACE_SLAB", "on"),
                  "kmers_per_s": k1["kmers_per_s"], "ms_per_slab": k1["kernel_ms"] / (bins // 64), "frac": k1["roofline"]["frac"]}))
f.close()
ctx.close()
