"""dpt_channel_sum_static: the per-channel sum for hosts that may not allocate device memory.

Some embedders of libdpt_b200.so cannot give a kernel scratch memory: their device memory was handed out at
start-up and later cudaMalloc calls are forbidden (fixed budgets shared with other tenants), or they run as a
plug-in without access to the host framework's allocator, or they cannot afford the device-wide
synchronisation an allocation implies.  For them the library carries a built-in scratch array (limit 512
channels, 592 KiB per GPU).  This test calls that entry point through the C ABI exactly as such a host would
(raw pointers, no scratch argument) and checks it against the ordinary entry point, twice in a row on
different data so that whatever the first call left in the shared array is shown not to leak into the second.

The file is named test_zz_* so that pytest collects it last: a failure here cannot hide the other tests in a
run that stops at the first failure.  It makes no assumption about what ran before it in the process.
"""
import ctypes as C

import pytest
import torch

from distributedpytorch_b200.ops import native as N

pytestmark = pytest.mark.gpu

MAX_C = 512  # DPT_CHANNEL_SUM_STATIC_MAX_C in include/dpt_b200.h


def _bind():
    L = N.lib()
    L.dpt_channel_sum_static.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_float, C.c_void_p]
    L.dpt_channel_sum_static.restype = C.c_int
    return L


def _random_bf16(shape, seed, device="cuda"):
    g = torch.Generator(device=device).manual_seed(seed)
    return (torch.rand(*shape, device=device, generator=g) * 2 - 1).to(torch.bfloat16)


def _both(L, x, scale):
    c = x.shape[-1]
    pixels = x.numel() // c
    stream = N.stream_ptr()
    got = torch.full((c,), float("nan"), device=x.device)
    want = torch.full((c,), float("nan"), device=x.device)
    before = torch.cuda.memory_allocated()
    rc = L.dpt_channel_sum_static(x.data_ptr(), got.data_ptr(), pixels, c, scale, stream)
    assert rc == 0, L.dpt_last_error().decode()
    assert torch.cuda.memory_allocated() == before  # nothing was taken from torch's allocator either
    N.channel_sum(x, want, scale=scale)  # existing entry point, caller-provided scratch
    torch.cuda.synchronize()
    return got, want


def test_static_scratch_matches_regular_entry_point_twice_in_a_row():
    L = _bind()
    # first call: wide tensor, fills most of the shared array
    x1 = _random_bf16((3, 40, 60, MAX_C), seed=101)
    got1, want1 = _both(L, x1, 0.5)
    assert torch.isfinite(got1).all()
    torch.testing.assert_close(got1, want1, atol=5e-2, rtol=1e-3)
    torch.testing.assert_close(got1, 0.5 * x1.float().sum(dim=(0, 1, 2)), atol=5e-2, rtol=1e-3)
    # second call right after it: narrower tensor, different data, stale partial sums still in the array
    x2 = _random_bf16((2, 24, 40, 64), seed=202)
    got2, want2 = _both(L, x2, 1.0)
    assert torch.isfinite(got2).all()
    torch.testing.assert_close(got2, want2, atol=5e-2, rtol=1e-3)
    torch.testing.assert_close(got2, x2.float().sum(dim=(0, 1, 2)), atol=5e-2, rtol=1e-3)


def test_static_scratch_rejects_oversize_with_a_clear_error():
    L = _bind()
    x = _random_bf16((1, 8, 8, MAX_C + 64), seed=303)
    out = torch.zeros(MAX_C + 64, device="cuda")
    rc = L.dpt_channel_sum_static(x.data_ptr(), out.data_ptr(), 64, MAX_C + 64, 1.0, N.stream_ptr())
    assert rc != 0
    msg = L.dpt_last_error().decode()
    assert "exceeds" in msg and str(MAX_C) in msg, msg
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0  # nothing was launched


def test_static_scratch_channel_count_that_is_not_a_multiple_of_16():
    """Even channel counts that are not multiples of 16 take the kernel's scalar path; same contract, same shared array."""
    L = _bind()
    for c, seed in ((38, 404), (6, 405)):
        x = _random_bf16((2, 20, 28, c), seed=seed)
        got, want = _both(L, x, 2.0)
        assert torch.isfinite(got).all()
        torch.testing.assert_close(got, want, atol=5e-2, rtol=1e-3)
        torch.testing.assert_close(got, 2.0 * x.float().sum(dim=(0, 1, 2)), atol=5e-2, rtol=1e-3)


@pytest.mark.multigpu
def test_static_scratch_on_a_second_device_and_foreign_stream():
    """Every GPU has its own copy of the array and its address is looked up per device; a stream that belongs to
    another device than the current one is refused instead of being given the wrong GPU's array."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    L = _bind()
    with torch.cuda.device(1):
        x = _random_bf16((2, 24, 40, 128), seed=506, device="cuda:1")
        got, want = _both(L, x, 1.0)
        torch.testing.assert_close(got, want, atol=5e-2, rtol=1e-3)
        torch.testing.assert_close(got, x.float().sum(dim=(0, 1, 2)), atol=5e-2, rtol=1e-3)
        foreign = torch.cuda.Stream(device=0)
        out = torch.zeros(128, device="cuda:1")
        rc = L.dpt_channel_sum_static(x.data_ptr(), out.data_ptr(), x.numel() // 128, 128, 1.0, foreign.cuda_stream)
        assert rc != 0 and "belongs to device" in L.dpt_last_error().decode()
    # and device 0 still works afterwards with its own array
    x0 = _random_bf16((1, 16, 16, 64), seed=507)
    got0, want0 = _both(L, x0, 1.0)
    torch.testing.assert_close(got0, want0, atol=5e-2, rtol=1e-3)
