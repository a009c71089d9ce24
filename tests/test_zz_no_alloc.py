"""b200ddp_channel_sum_static — the entry point for callers that cannot provide scratch memory.

Who needs it: code that runs where no device allocator is within reach (a plugin invoked from inside another
framework's step, with all GPU memory already owned by that framework's pool) or where cudaMalloc must not be
called (it synchronises the device and can fail under a memory cap). Such a caller still passes its input and
output buffers, but the per-CTA partial sums and tickets live in a fixed 258 KB area inside the library.

The test goes through the C ABI (ctypes), runs the new entry point twice in a row on different data — the area
is shared and keeps the previous call's contents, so the second call proves that what must start at zero is
re-zeroed — and checks both results against the regular entry point (caller-provided scratch), with the
tolerance the other reduction tests use. It also checks that a channel count beyond the built-in limit is
refused with a clear message instead of overrunning the area.

This file is named so that pytest collects it last: if it ever fails under `-x`, every other test has already run.
"""
import pytest
import torch

from b200ddp import _lib

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)


def _regular(L, dy, st):
    n, c, h, w = dy.shape
    out = torch.empty(c, device=DEV)
    scratch = torch.zeros(int(L.b200ddp_channel_sum_scratch_bytes(n, c)), dtype=torch.uint8, device=DEV)
    _lib.check(L.b200ddp_channel_sum(dy.data_ptr(), out.data_ptr(), n, c, h * w, scratch.data_ptr(), scratch.numel(),
                                     0, st), "channel_sum")
    return out


def test_static_scratch_variant_matches_regular_entry_point_on_two_calls_in_a_row():
    L = _lib.load()
    st = torch.cuda.current_stream(DEV).cuda_stream
    torch.manual_seed(2024)
    # first call also performs the one-time address lookup for this device (no stream capture is active here)
    shapes = [(9, 384, 13, 13), (128, 64, 55, 55)]      # second call: fewer channels, more splits, other sizes
    for shape in shapes:
        dy = torch.randn(*shape, device=DEV)
        want = _regular(L, dy, st)
        got = torch.empty(shape[1], device=DEV)
        _lib.check(L.b200ddp_channel_sum_static(dy.data_ptr(), got.data_ptr(), shape[0], shape[1],
                                                shape[2] * shape[3], 0, st), "channel_sum_static")
        torch.cuda.synchronize()
        assert torch.allclose(got, want, rtol=1e-5, atol=1e-3)
        assert torch.allclose(got.double(), dy.double().sum((0, 2, 3)), rtol=1e-5, atol=1e-3)


def test_static_scratch_variant_refuses_sizes_beyond_its_limit():
    L = _lib.load()
    st = torch.cuda.current_stream(DEV).cuda_stream
    dy = torch.randn(2, 513, 4, 4, device=DEV)
    out = torch.empty(513, device=DEV)
    rc = L.b200ddp_channel_sum_static(dy.data_ptr(), out.data_ptr(), 2, 513, 16, 0, st)
    assert rc == 1
    msg = _lib.last_error()
    assert "513" in msg and "512" in msg and "scratch" in msg
