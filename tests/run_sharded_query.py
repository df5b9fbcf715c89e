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
    # default: the C++ layer reduces its pruning flags and gathers its candidate words over NCCL itself (trx_comm_*, on the
    # device); TRX_SHARD_CALLBACK=1 = the older route through a Python callback and torch.distributed
    comm = None if os.environ.get("TRX_SHARD_CALLBACK") else t.Comm.from_torch_dist(ctx)
    index = sharded.ShardedIndex(ctx, sys.argv[1], comm=comm)
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
    # throughput of the filter phase over the shards, collectives included (every rank runs the same walk)
    import time
    reps = int(os.environ.get("TRX_REPS", "200"))
    for regex in sys.argv[2:3]:
        for _ in range(10):
            index.local_candidates(regex)
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            _, st = index.local_candidates(regex)
        dt = time.perf_counter() - t0
        if rank == 0:
            how = "trx_comm (NCCL inside libtrxhost)" if comm is not None else "python callback + torch.distributed"
            coll = f", {comm.collectives} NCCL collectives so far" if comm is not None else ""
            print(f"== process_query x{reps} over {world} shards via {how}: {reps / dt:.0f} queries/s, "
                  f"{dt / reps * 1e6:.0f} us/query, {st}{coll}")
    index.close()
    if comm is not None:
        comm.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
