"""GPU: trx_bulk_contains_noalloc — the allocation-free variant of K2's host-buffer entry point (fixed file-scope
`static __device__` staging array instead of the context's cudaMalloc'ed buffers; no counterpart in the reference)
— against trx_bulk_contains on the same filter, through the C ABI.

Collected last on purpose (the file name sorts after every other test file): a failure
here must not hide the parity tests.  Passes alone in a fresh process and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest

import tetrex_b200 as t
from tetrex_b200._abi import lib

pytestmark = pytest.mark.gpu


def make_filter(ctx, bins, bin_size, h, seed):
    rng = np.random.default_rng(seed)
    f = t.IBF(ctx, bins, bin_size, h)
    v = rng.integers(0, 2 ** 63, 20000, dtype=np.uint64)
    f.emplace_values(v, rng.integers(0, bins, len(v)).astype(np.uint32))
    return f, v, rng


@pytest.mark.parametrize("bins,bin_size,h", [(1024, 20011, 3), (5, 106, 3), (8192, 4099, 3), (300, 7777, 5)])
def test_noalloc_equals_bulk_contains_two_calls_in_a_row(ctx, bins, bin_size, h):
    f, present, rng = make_filter(ctx, bins, bin_size, h, bins)
    cap = f.noalloc_max_kmers
    assert cap == (4 << 20) // (8 + f.shard_words * 8) and cap >= 1
    n = min(cap, 3000)
    # two calls in a row share the scratch array: different k-mers, the second batch shorter than the first, so stale
    # contents of the first call sit right behind the second call's data
    q1 = np.concatenate([present[:n // 2], rng.integers(0, 2 ** 63, n - n // 2, dtype=np.uint64)])
    q2 = np.concatenate([rng.integers(0, 2 ** 63, n // 3, dtype=np.uint64), present[n:n + n // 3]])
    want1, want2 = f.bulk_contains(q1), f.bulk_contains(q2)
    got1 = f.bulk_contains_noalloc(q1)
    got2 = f.bulk_contains_noalloc(q2)
    assert np.array_equal(got1, want1)  # bit-exact: integer work, the tolerance of every test in this repository is 0
    assert np.array_equal(got2, want2)
    assert want1.any() and want2.any()
    # and once more the first batch: the result does not depend on what the array held before
    assert np.array_equal(f.bulk_contains_noalloc(q1), want1)
    f.close()


def test_noalloc_full_capacity_and_rejection(ctx):
    f, present, rng = make_filter(ctx, 1024, 9973, 3, 7)
    cap = f.noalloc_max_kmers
    q = rng.integers(0, 2 ** 63, cap + 1, dtype=np.uint64)
    q[:1000] = present[:1000]
    assert np.array_equal(f.bulk_contains_noalloc(q[:cap]), f.bulk_contains(q[:cap]))  # exactly at the limit
    out = np.zeros((cap + 1, f.shard_words), dtype=np.uint64)
    rc = lib().trx_bulk_contains_noalloc(f._h, q.ctypes.data, cap + 1, out.ctypes.data)
    assert rc == -5  # TRX_ECAPACITY
    msg = lib().trx_last_error().decode()
    assert "fixed scratch" in msg and str(cap) in msg, msg
    assert not out.any()  # nothing was written
    assert lib().trx_bulk_contains_noalloc(f._h, None, 0, None) == 0  # n = 0 is a no-op
    assert lib().trx_bulk_contains_noalloc(None, q.ctypes.data, 1, out.ctypes.data) == -1
    f.close()


def test_noalloc_makes_no_device_allocation(ctx):
    """cudaMemGetInfo before and after: a fresh context's first trx_bulk_contains grows staging buffers, the no-alloc
    variant must leave the free-memory figure where it was."""
    import torch
    own = t.Context(0)  # fresh context: its grow-on-demand staging buffers are still empty
    f, present, rng = make_filter(own, 1024, 20011, 3, 3)
    q = rng.integers(0, 2 ** 63, 2000, dtype=np.uint64)
    f.bulk_contains_noalloc(q[:16])  # first use on this device: cudaGetSymbolAddress, module already loaded
    own.sync()
    free0, _ = torch.cuda.mem_get_info(0)
    for _ in range(3):
        f.bulk_contains_noalloc(q)
    own.sync()
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 >= free0, (free0, free1)  # not one byte of device memory was taken
    f.close()
    own.close()
