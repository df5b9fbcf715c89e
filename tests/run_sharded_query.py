"""torchrun driver for the multi-GPU query path (one process per GPU, NCCL):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tests/run_sharded_query.py INDEX.ibf REGEX [REGEX …]

Rank 0 prints, per regex, the candidate bins and the output lines; with TRX_EXPECT_CLI=1 it also runs the single-GPU
`tetrex query` and fails if the two outputs differ."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist

import tetrex_b200 as t
from tetrex_b200 import sharded


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = t.Context(local)
    index = sharded.ShardedIndex(ctx, sys.argv[1])
    ok = True
    for regex in sys.argv[2:]:
        words, text = index.query(regex)
        if rank == 0:
            bins = [b for b in range(index.bin_count) if (int(words[b // 64]) >> (b % 64)) & 1]
            print(f"== {regex}: {len(bins)} candidate bins {bins[:16]}{' …' if len(bins) > 16 else ''}; {text.count(chr(10))} lines")
            sys.stdout.write(text)
            if os.environ.get("TRX_EXPECT_CLI"):
                want = subprocess.run([os.path.join(ROOT, "tetrex_b200", "bin", "tetrex"), "query", sys.argv[1], regex],
                                      capture_output=True, text=True)
                same = want.stdout == text
                print(f"   equals single-GPU tetrex query: {same}")
                ok &= same
    index.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
