"""lsb200_sort_pairs_static: the allocation-free sort entry (scratch in a file-scope __device__ array of liblsb200.so).

Collected last on purpose.  Compares the entry's output with the existing
lsb200_debug_sort_pairs on the same pairs -- bit-exact, the tolerance of every integer test here -- on two calls in a
row (the second must not see the first call's histograms / tickets / status words), checks the size limit's error, and
that the calls themselves allocate no device memory.  Device arrays are the CALLER's: torch tensors made before the call.
"""
import numpy as np
import pytest

from lidarshooter_b200 import capi

pytestmark = pytest.mark.gpu


def _pairs(n, seed):
    rng = np.random.default_rng(seed)
    keys = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    keys[: n // 3] &= np.uint32(0x3FF)               # long runs of equal high digits + duplicates (stability shows)
    return keys, np.arange(n, dtype=np.uint32)


def _to_dev(torch, a):
    return torch.from_numpy(a.view(np.int32).copy()).cuda()


def _to_host(t):
    return t.cpu().numpy().view(np.uint32)


def test_static_sort_matches_existing_entry_two_calls_in_a_row():
    import torch
    ctx = capi.Context(0)
    try:
        assert ctx.sort_pairs_static(0, 0, 0, 0, 0) == 0           # set-up only: resolves the array's device address
        w = torch.zeros(4 * 20_000, dtype=torch.int32, device="cuda")   # warm-up: the driver loads the three kernels
        assert ctx.sort_pairs_static(*(w[i * 20_000:].data_ptr() for i in range(4)), 20_000) == 0   # lazily, outside the checks
        ctx.synchronize()
        # second call smaller than the first AND not a multiple of the tile: stale status words beyond / inside its
        # range would corrupt the look-back if the scratch were not cleared
        for call, n in enumerate((1_000_003, 300_001, capi.SORT_STATIC_MAX_PAIRS)):
            keys, vals = _pairs(n, 100 + call)
            dk, dv = _to_dev(torch, keys), _to_dev(torch, vals)
            kt, vt = torch.empty_like(dk), torch.empty_like(dv)
            torch.cuda.synchronize()
            reserved = torch.cuda.memory_reserved()
            free_before = torch.cuda.mem_get_info()[0]
            assert ctx.sort_pairs_static(dk.data_ptr(), dv.data_ptr(), kt.data_ptr(), vt.data_ptr(), n) == 0, ctx.last_error()
            ctx.synchronize()
            assert torch.cuda.mem_get_info()[0] == free_before      # nothing was allocated (or freed) on the device
            assert torch.cuda.memory_reserved() == reserved
            ek, ev = ctx.debug_sort_pairs(keys, vals)               # the existing entry point, same pairs
            assert np.array_equal(_to_host(dk), ek), (call, n)
            assert np.array_equal(_to_host(dv), ev), (call, n)
            order = np.argsort(keys, kind="stable")
            assert np.array_equal(ek, keys[order]) and np.array_equal(ev, vals[order])
    finally:
        ctx.close()


def test_static_sort_rejects_sizes_beyond_the_maximum():
    import torch
    ctx = capi.Context(0)
    try:
        d = torch.zeros(16, dtype=torch.int32, device="cuda")
        rc = ctx.sort_pairs_static(d.data_ptr(), d.data_ptr(), d.data_ptr(), d.data_ptr(), capi.SORT_STATIC_MAX_PAIRS + 1)
        assert rc == capi.ERR_CAPACITY
        assert "LSB200_SORT_STATIC_MAX_PAIRS" in ctx.last_error() and str(capi.SORT_STATIC_MAX_PAIRS) in ctx.last_error()
        assert ctx.sort_pairs_static(d.data_ptr(), d.data_ptr(), d.data_ptr(), d.data_ptr(), -1) == capi.ERR_INVALID
        assert ctx.sort_pairs_static(0, d.data_ptr(), d.data_ptr(), d.data_ptr(), 4) == capi.ERR_INVALID
        torch.cuda.synchronize()
    finally:
        ctx.close()


def test_static_sort_on_a_second_device_uses_that_devices_array():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU")
    ctx = capi.Context(1)
    try:
        keys, vals = _pairs(70_001, 7)
        with torch.cuda.device(1):
            dk, dv = _to_dev(torch, keys).to("cuda:1"), _to_dev(torch, vals).to("cuda:1")
            kt, vt = torch.empty_like(dk), torch.empty_like(dv)
            torch.cuda.synchronize()
        for _ in range(2):
            assert ctx.sort_pairs_static(dk.data_ptr(), dv.data_ptr(), kt.data_ptr(), vt.data_ptr(), keys.size) == 0, ctx.last_error()
            ctx.synchronize()
        order = np.argsort(keys, kind="stable")
        assert np.array_equal(_to_host(dk), keys[order]) and np.array_equal(_to_host(dv), vals[order])
    finally:
        ctx.close()
