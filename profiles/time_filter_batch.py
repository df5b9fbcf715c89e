"""Filter phase of a motif file (1000 config-3 patterns, -a) on the /100 twin's flat index: process_query one pattern at a
time against process_query_batch (one membership batch per 256 patterns, dry runs and walks on the host threads)."""
import os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np
import make_golden as mg
import tetrex_b200 as t
from tetrex_b200 import sharded, synth
tmp = tempfile.mkdtemp()
residues, offs, rec_bin, _ = synth.swissprot_like(5700, 1024)
pats = synth.prosite_patterns(1000, 1000)
synth.plant_patterns(residues, offs, pats)
files = mg.write_twin_fasta(os.path.join(tmp, "twin"), residues, offs, rec_bin, 1024)
open(os.path.join(tmp, "bins.lst"), "w").write("\n".join(files) + "\n")
subprocess.run([os.path.join(ROOT, "tetrex_b200", "bin", "tetrex"), "index", "-i", "-k", "6", os.path.join(tmp, "flat"),
                os.path.join(tmp, "bins.lst")], check=True, capture_output=True)
ctx = t.Context(0)
ix = sharded.ShardedIndex(ctx, os.path.join(tmp, "flat.ibf"), 0, 1)
for augment in ((True,) if os.environ.get("AUGMENT_ONLY") else (False, True)):
    ix.set_augment(augment)
    ix.local_candidates_batch(pats[:32])
    t0 = time.perf_counter(); words, status, stats = ix.local_candidates_batch(pats); dt_b = time.perf_counter() - t0
    t0 = time.perf_counter(); same = True; refused = 0
    for i, p in enumerate(pats):
        try:
            w, st = ix.local_candidates(p)
            same = same and bool(status[i]) and np.array_equal(w, words[i])
        except sharded.HostError:
            refused += 1
            same = same and not status[i]
    dt_1 = time.perf_counter() - t0
    print(f"augment={augment}: batch {dt_b*1e3:.1f} ms ({1000/dt_b:.0f} patterns/s), one at a time {dt_1*1e3:.1f} ms "
          f"({1000/dt_1:.0f}/s), refused {refused}, same answers {same}, probes {int(stats[:,0].sum())}")
ix.close(); ctx.close()
