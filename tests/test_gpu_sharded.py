"""GPU: the multi-GPU query path of SURVEY.md §8(e) on ONE device — S shards of the same index file live side by side
(one thread and one CUDA context handle per shard), the collector's pruning flags are OR-ed over the shards at every probe
step, each shard verifies its own bins.  Candidate vector and printed lines must equal the unsharded `tetrex query`."""
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

import tetrex_b200 as t
from tetrex_b200 import sharded

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from sharded_collection import QUERIES, TETREX  # noqa: E402


def run(args):
    return subprocess.run([TETREX] + args, capture_output=True, text=True, cwd=ROOT, timeout=600)


class ThreadAllReduce:
    """all-reduce(MAX) over the shard threads of one process — what NCCL does between the processes of a torchrun job"""

    def __init__(self, n):
        self.n, self.barrier, self.lock, self.acc, self.calls = n, threading.Barrier(n), threading.Lock(), None, 0

    def __call__(self, a):
        with self.lock:
            self.acc = a.copy() if self.acc is None else np.maximum(self.acc, a)
        self.barrier.wait()
        out = self.acc.copy()
        if self.barrier.wait() == 0:
            self.acc = None
            self.calls += 1
        self.barrier.wait()
        return out


@pytest.fixture(scope="module")
def collection(tmp_path_factory):
    from sharded_collection import build
    return build(str(tmp_path_factory.mktemp("sharded")))


# ("hibf", 8): more shards than 64-bin words (256 bins = 4 words) — four of the eight shards own no bins at all and must
# answer "no lines" instead of failing to upload an empty text (a 256-bin replicated HIBF on an 8-GPU box)
@pytest.mark.parametrize("kind,shards", [("flat", 2), ("flat", 4), ("hibf", 2), ("hibf", 3), ("hibf", 8)])
def test_sharded_query_equals_unsharded(ctx, collection, kind, shards):
    path = collection[kind]
    reduce_ = ThreadAllReduce(shards)
    # one context (stream + staging buffers) per shard thread — the C ABI's "one handle per host thread" rule
    ctxs = [t.Context(0) for _ in range(shards)]
    handles = [sharded.ShardedIndex(ctxs[s], path, s, shards, any_alive=reduce_) for s in range(shards)]
    assert handles[0].bin_count == 256 and handles[0].bin_words == 4
    if kind == "flat":
        assert [(h.word_begin, h.words) for h in handles] == [(s * 4 // shards, 4 // shards) for s in range(shards)]
    try:
        for regex in QUERIES:
            want = run(["query", path, regex])
            assert want.returncode == 0, want.stderr
            want_bins = [int(x) for x in run(["bins", path, regex]).stdout.split()]
            results = [None] * shards
            errors = []

            def work(s):
                try:
                    results[s] = handles[s].local_candidates(regex)
                except Exception as e:  # a dead thread would leave the others waiting at the barrier
                    errors.append(e)
                    reduce_.barrier.abort()

            threads = [threading.Thread(target=work, args=(s,)) for s in range(shards)]
            [th.start() for th in threads]
            [th.join() for th in threads]
            assert not errors, errors
            words = np.zeros(4, dtype=np.uint64)
            for s, (w, stats) in enumerate(results):
                if kind == "flat":
                    lo, n = handles[s].word_begin, handles[s].words
                    assert not w[:lo].any() and not w[lo + n:].any(), "a column shard reports its own words only"
                    assert stats == results[0][1], "every shard walks the same items (same pruning decisions)"
                else:
                    assert np.array_equal(w, results[0][0]), "a replicated HIBF gives every shard the whole vector"
                words |= w
            got_bins = [b for b in range(256) if (int(words[b // 64]) >> (b % 64)) & 1]
            assert got_bins == want_bins, (kind, shards, regex)
            text = "".join(h.verify_local(words, regex)[0] for h in handles)
            assert text == want.stdout, (kind, shards, regex)
        if kind == "flat":
            assert reduce_.calls > 0  # the pruning flags really crossed shards
    finally:
        for h in handles:
            h.close()
        for c in ctxs:
            c.close()


def test_column_shard_without_reduction_is_refused(ctx, collection):
    h = sharded.ShardedIndex(ctx, collection["flat"], 0, 2, any_alive=lambda a: a)
    try:
        h.column_sharded = False  # pretend the driver forgot the hook: the C++ layer must refuse, not answer for half the bins
        with pytest.raises(sharded.HostError, match="any_alive"):
            h.local_candidates("LMAQGLYN")
    finally:
        h.close()


def test_shard_count_must_divide_columns(ctx, collection):
    with pytest.raises(sharded.HostError, match="divide"):
        sharded.ShardedIndex(ctx, collection["flat"], 0, 3)


@pytest.mark.parametrize("kind", ["flat", "hibf"])
def test_cli_gpus_flag_equals_single_gpu(collection, kind):
    """`tetrex query --gpus N` (one thread and one shard per device, NCCL inside the library) prints what the single-GPU
    run prints.  Needs at least two devices: skipped on the one-GPU boxes of the default test run, exercised by
    `gpurun --gpus 2` (log under profiles/)."""
    import torch
    n = min(torch.cuda.device_count(), 4)
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    path = collection[kind]
    for q in QUERIES:
        one = run(["query", path, q])
        many = run(["query", "--gpus", str(n), path, q])
        assert one.returncode == 0 and many.returncode == 0, many.stderr
        assert many.stdout == one.stdout, (kind, q, many.stderr)
    env_run = subprocess.run([TETREX, "query", path, QUERIES[0]], capture_output=True, text=True, cwd=ROOT, timeout=600,
                             env=dict(os.environ, TETREX_GPUS=str(n)))
    assert env_run.returncode == 0 and env_run.stdout == run(["query", path, QUERIES[0]]).stdout
