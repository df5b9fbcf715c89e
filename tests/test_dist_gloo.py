"""§8(e) host-side multi-GPU logic on CPU: world_size 2 over gloo.  The column-block partition of a flat IBF, the
global `alive` decision of the collector, the final word gather and the hit gather are exercised with the oracle
standing in for the per-rank device (tests may use oracle/; the product never does)."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from tetrex_b200 import dist as td


def test_unit_slice_partitions_everything():
    for n in [0, 1, 7, 64, 1000003]:
        for world in [1, 2, 3, 8]:
            parts = [td.unit_slice(n, r, world) for r in range(world)]
            assert parts[0].start == 0 and parts[-1].stop == n
            assert all(a.stop == b.start for a, b in zip(parts, parts[1:]))
            sizes = [p.stop - p.start for p in parts]
            assert max(sizes) - min(sizes) <= 1


def test_column_shard_rule():
    assert td.column_shard(128, 3, 8) == (48, 16)
    assert td.column_shard(16, 1, 2) == (8, 8)
    with pytest.raises(ValueError):
        td.column_shard(16, 0, 3)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from oracle import bindings as ob
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        bins, bin_size, h = 256, 4099, 3
        rng = np.random.default_rng(7)  # same stream on every rank: the host NFA walk is replicated
        values = rng.integers(0, 1 << 30, size=20000, dtype=np.uint64)
        vbins = rng.integers(0, bins, size=values.size, dtype=np.uint32)
        whole = ob.OracleIBF(bins, bin_size, h)
        whole.emplace(values, vbins)
        bin_words = whole.bin_words
        first, count = td.column_shard(bin_words, rank, world)
        # items = paths of 3 k-mers each; half of them made of inserted values of one bin (stay alive somewhere)
        n_items = 64
        kmers = rng.integers(0, 1 << 30, size=(n_items, 3), dtype=np.uint64)
        for i in range(0, n_items, 2):
            b = i % bins
            mine = values[vbins == b][:3]
            kmers[i, :mine.size] = mine
        full = np.full((n_items, bin_words), ~np.uint64(0), dtype=np.uint64)
        for j in range(3):
            full &= whole.bulk_contains(kmers[:, j])
        local = full[:, first:first + count].copy()      # what this rank's device holds
        alive_local = (local != 0).any(axis=1).astype(np.uint8)
        alive = td.any_alive(alive_local)
        assert np.array_equal(alive, (full != 0).any(axis=1).astype(np.uint8))
        assert alive[::2].all()
        # match accumulator: OR over live items, gathered to the whole vector on every rank
        acc_local = np.bitwise_or.reduce(local[alive.astype(bool)], axis=0)
        got = td.gather_words(acc_local)
        want = np.bitwise_or.reduce(full[alive.astype(bool)], axis=0)
        assert np.array_equal(got, want)
        # hits of this rank's bins → rank 0, sorted like the reference prints them
        my_bins = np.arange(first * 64, (first + count) * 64, dtype=np.uint32)
        hr = np.random.default_rng(100 + rank)
        hits = np.stack([hr.choice(my_bins, 50), hr.integers(0, 9, 50), hr.integers(0, 500, 50), hr.integers(500, 900, 50)],
                        axis=1).astype(np.uint32)
        merged = td.gather_hits(hits)
        if rank == 0:
            assert merged.shape == (50 * world, 4)
            keys = [tuple(r[:3]) for r in merged.tolist()]
            assert keys == sorted(keys)
        else:
            assert merged is None
        # output lines: every rank formats the lines of its own (ascending) bins, rank 0 gets them in rank = bin order
        from tetrex_b200 import sharded
        text = sharded.gather_text("".join(f"bin{b}\t>r\tLMA\t0,3\n" for b in my_bins[:3]))
        if rank == 0:
            want_lines = [f"bin{b}\t>r\tLMA\t0,3" for r in range(world) for b in
                          np.arange(td.column_shard(bin_words, r, world)[0] * 64, td.column_shard(bin_words, r, world)[0] * 64 + 3)]
            assert text.splitlines() == want_lines
        else:
            assert text is None
        assert td.max_over_ranks(float(rank + 1)) == float(world)
        assert td.sum_over_ranks(1.5) == 1.5 * world
        sl = td.unit_slice(values.size, rank, world)
        assert td.sum_over_ranks(float(sl.stop - sl.start)) == float(values.size)
        td.barrier()
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))
