"""ngp_compact_alive_static: the allocation-free variant of ngp_compact_alive (scratch in a file-scope __device__
array of the library, include/ngp_b200.h).  Checked through the C ABI against the existing entry point, bit for
bit (integer work), on calls in a row that share the array.  Collected last on purpose (file name):
a failure here must not hide the parity tests.
"""
import ctypes as C

import pytest
import torch

from nerfsafetyvalidation_b200 import _cabi

pytestmark = pytest.mark.gpu

MAX_N = 1 << 22     # NGP_COMPACT_STATIC_MAX_N


def _alive(n, seed, keep=0.63):
    gen = torch.Generator().manual_seed(seed)
    x = torch.arange(n, dtype=torch.int32)
    x[torch.rand(n, generator=gen) >= keep] = -1
    return x


def _existing(lib, xd):
    n = xd.numel()
    out = torch.full((max(n, 1),), -7, dtype=torch.int32, device=xd.device)
    count = torch.full((1,), 0xFFFF, dtype=torch.int32, device=xd.device)
    scratch = torch.empty(int(lib.ngp_compact_scratch_bytes(max(n, 1))), dtype=torch.uint8, device=xd.device)
    _cabi.check(lib.ngp_compact_alive(_cabi.ptr(xd), n, _cabi.ptr(out), _cabi.ptr(count), _cabi.ptr(scratch),
                                      _cabi.stream()), "compact_alive")
    return out, count


def _static(lib, xd, out, count):
    _cabi.check(lib.ngp_compact_alive_static(_cabi.ptr(xd), xd.numel(), _cabi.ptr(out), _cabi.ptr(count),
                                             _cabi.stream()), "compact_alive_static")


def test_static_scratch_matches_existing_entry_point_on_calls_in_a_row(cuda):
    lib = _cabi.load()
    # large, then small (stale ticket/status words of the larger call must not leak), then large again, the maximum,
    # and sizes around the 2048-element tile
    sizes = [2_000_000, 2049, 640_000, 1, 2048, MAX_N, 33, 100_003]
    for i, n in enumerate(sizes):
        xd = _alive(n, seed=10 + i).to(cuda)
        ref_out, ref_count = _existing(lib, xd)
        out = torch.full((n,), -7, dtype=torch.int32, device=cuda)
        count = torch.full((1,), 0xFFFF, dtype=torch.int32, device=cuda)
        _static(lib, xd, out, count)
        k = int(count.item())
        assert k == int(ref_count.item()) == int((xd >= 0).sum().item()), f"n={n}: survivor count"
        assert torch.equal(out[:k], ref_out[:k]), f"n={n}: survivors differ from ngp_compact_alive"
        assert torch.equal(out[:k], xd[xd >= 0]), f"n={n}: not the stable compaction"
        assert bool((out[k:] == -7).all()), f"n={n}: wrote past the survivor count"
        # the same input again, straight away: the array still holds the call before it
        out2 = torch.full((n,), -7, dtype=torch.int32, device=cuda)
        count2 = torch.zeros(1, dtype=torch.int32, device=cuda)
        _static(lib, xd, out2, count2)
        assert int(count2.item()) == k and torch.equal(out2, out), f"n={n}: second call in a row differs"


def test_static_scratch_allocates_nothing_and_counts_launches(cuda):
    lib = _cabi.load()
    xd = _alive(640_000, seed=5).to(cuda)
    out = torch.empty_like(xd)
    count = torch.zeros(1, dtype=torch.int32, device=cuda)
    _static(lib, xd, out, count)                       # resolves the symbol address for this device
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(cuda)
    launches0 = _cabi.launch_count()
    for _ in range(8):
        _static(lib, xd, out, count)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(cuda)
    assert free1 >= free0, f"device memory shrank by {free0 - free1} bytes across allocation-free calls"
    assert _cabi.launch_count() == launches0 + 8
    assert torch.equal(out[: int(count.item())], xd[xd >= 0])


def test_static_scratch_edges_and_refusals(cuda):
    lib = _cabi.load()
    count = torch.full((1,), 99, dtype=torch.int32, device=cuda)
    # empty input: count cleared, nothing launched
    empty = torch.empty(0, dtype=torch.int32, device=cuda)
    _cabi.check(lib.ngp_compact_alive_static(_cabi.ptr(empty), 0, _cabi.ptr(empty), _cabi.ptr(count), _cabi.stream()))
    assert int(count.item()) == 0
    # nothing survives
    none = torch.full((5000,), -1, dtype=torch.int32, device=cuda)
    out = torch.full((5000,), -7, dtype=torch.int32, device=cuda)
    _static(lib, none, out, count)
    assert int(count.item()) == 0 and bool((out == -7).all())
    # beyond the built-in scratch: refused before anything is touched, with the limit in the message
    big = torch.zeros(16, dtype=torch.int32, device=cuda)
    count.fill_(99)
    rc = lib.ngp_compact_alive_static(_cabi.ptr(big), MAX_N + 1, _cabi.ptr(big), _cabi.ptr(count), _cabi.stream())
    assert rc != 0
    msg = _cabi.last_error()
    assert "exceeds" in msg and str(MAX_N) in msg, msg
    assert int(count.item()) == 99
    # host pointers are refused like everywhere else (no CPU fallback)
    host = torch.zeros(64, dtype=torch.int32)
    rc = lib.ngp_compact_alive_static(C.c_void_p(host.data_ptr()), 64, _cabi.ptr(out), _cabi.ptr(count), _cabi.stream())
    assert rc != 0 and "device pointer" in _cabi.last_error()
    # and the library still works afterwards
    xd = _alive(4097, seed=1).to(cuda)
    out = torch.empty_like(xd)
    _static(lib, xd, out, count)
    assert torch.equal(out[: int(count.item())], xd[xd >= 0])
