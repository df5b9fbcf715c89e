"""Synthetic 256-bin peptide collection + flat and HIBF index files for the multi-GPU tests:
python tests/sharded_collection.py DIR  →  DIR/flat.ibf, DIR/hibf.ibf (built with the repo's own `tetrex index`)."""
import os
import random
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TETREX = os.path.join(ROOT, "tetrex_b200", "bin", "tetrex")
AA = "ACDEFGHIKLMNPQRSTVWY"
MOTIFS = {3: "LMAQGLYN", 70: "LMAEGLYN", 130: "LMAQGLYN", 200: "LMAEGLYN", 255: "LMAQGLYN", 64: "CAHCWKKH"}
QUERIES = ["LMA(E|Q)GLYN", "CAHCWKKH", "LMAQGL", "MA(E|Q)GL(Y|F)N", "WWWWWWW", "L.AQGLYN"]


def build(d: str) -> dict:
    """4 words of 64 bins; motif instances planted in a few bins of every 64-bin column block."""
    os.makedirs(d, exist_ok=True)
    rng = random.Random(5)
    files = []
    for b in range(256):
        path = os.path.join(d, f"bin_{b:04d}.fa")
        with open(path, "w") as f:
            for r in range(3):
                seq = "".join(rng.choice(AA) for _ in range(rng.randint(60, 400)))
                if b in MOTIFS and r == 1:
                    seq = seq[:20] + MOTIFS[b] + seq[20:]
                f.write(f">b{b}_r{r} synthetic\n{seq}\n")
        files.append(path)
    lst = os.path.join(d, "bins.lst")
    with open(lst, "w") as f:
        f.write("\n".join(files) + "\n")
    for flags, name in ((["-i"], "flat"), ([], "hibf")):
        out = subprocess.run([TETREX, "index"] + flags + ["-k", "5", os.path.join(d, name), lst], capture_output=True,
                             text=True, cwd=ROOT, timeout=600)
        assert out.returncode == 0, out.stderr
    return {"flat": os.path.join(d, "flat.ibf"), "hibf": os.path.join(d, "hibf.ibf"), "files": files}


if __name__ == "__main__":
    print(build(sys.argv[1]))
