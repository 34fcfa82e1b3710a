"""GPU: vsp_fr_fft_noalloc_dev (static device scratch, no allocation) against vsp_fr_fft_dev through the C ABI.

The entry point has no counterpart in the reference (include/vsp_b200.h says why it exists); what pins it is the
existing entry point, which tests/test_gpu_ntt.py holds bit-exact against the oracle.  Integer work: the comparison is
bit-exact, like every other NTT test here.  Named test_zz_* so that pytest collects it last:
a failure here must not hide the parity tests.  Works alone in a fresh process (it makes its own first
call on the device, which does the per-device set-up) and after the other tests (which leave the pool, the twiddle
tables of the allocating path and other contexts behind)."""
import ctypes as C

import numpy as np
import pytest

from oracle import fixtures as fx

pytestmark = pytest.mark.gpu

MAX_LOG = 16  # VSP_NOALLOC_FFT_MAX_M = 2^16 (include/vsp_b200.h)


def _both(vsp, ctx, torch, m, a, inverse):
    """the same input through the allocating and the non-allocating entry points; returns the two results"""
    lib = ctx._lib
    host = torch.from_numpy(a.view(np.int64))
    d_ref, d_new = host.cuda(), host.cuda()
    ctx.check(lib.vsp_fr_fft_dev(ctx._h, m, int(inverse), C.c_void_p(d_ref.data_ptr())))
    ctx.check(lib.vsp_fr_fft_noalloc_dev(ctx._h, m, int(inverse), C.c_void_p(d_new.data_ptr())))
    ctx.synchronize()
    return d_ref.cpu().numpy().view(np.uint64), d_new.cpu().numpy().view(np.uint64)


@pytest.mark.parametrize("log_n", [1, 2, 5, 9, 10, 11, 13, 15, 16])
def test_noalloc_fft_equals_fft_dev_twice_in_a_row(vsp, ctx, log_n):
    """two calls in a row per direction, different data each time: the second call finds the first one's contents in
    the shared scratch and must not be affected by them"""
    import torch
    m = 1 << log_n
    for call in range(2):
        a = fx.random_fr(m, seed=9000 + 10 * log_n + call)
        for inverse in (False, True):
            want, got = _both(vsp, ctx, torch, m, a, inverse)
            assert np.array_equal(got, want), (log_n, call, inverse)
    # and back: inverse(forward(a)) == a through the non-allocating path alone (evaluation_domain.cpp:103-121)
    d = torch.from_numpy(a.view(np.int64)).cuda()
    dom = vsp.make_evaluation_domain(ctx, m)
    dom.fft_noalloc_dev(d.data_ptr(), False)
    dom.fft_noalloc_dev(d.data_ptr(), True)
    ctx.synchronize()
    assert np.array_equal(d.cpu().numpy().view(np.uint64), a)


def test_noalloc_fft_sizes_interleaved(vsp, ctx):
    """a large transform, a small one, the large one again: stale scratch from another size changes nothing"""
    import torch
    for log_n in (16, 11, 16, 12, 14):
        m = 1 << log_n
        a = fx.random_fr(m, seed=777 + log_n)
        want, got = _both(vsp, ctx, torch, m, a, False)
        assert np.array_equal(got, want), log_n


def test_noalloc_fft_leaves_device_memory_alone(vsp, ctx):
    """free device memory is the same before and after a run of calls (first call made beforehand: it loads the kernels
    and does the per-device set-up), and the launch counter moves by the two passes of each transform"""
    import torch
    m = 1 << MAX_LOG
    d = torch.from_numpy(fx.random_fr(m, seed=5).view(np.int64)).cuda()
    lib = ctx._lib
    ctx.check(lib.vsp_fr_fft_noalloc_dev(ctx._h, m, 0, C.c_void_p(d.data_ptr())))
    ctx.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    launches0 = ctx.launch_count
    for _ in range(8):
        ctx.check(lib.vsp_fr_fft_noalloc_dev(ctx._h, m, 0, C.c_void_p(d.data_ptr())))
        ctx.check(lib.vsp_fr_fft_noalloc_dev(ctx._h, m, 1, C.c_void_p(d.data_ptr())))
    ctx.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
    assert ctx.launch_count - launches0 == 8 * 2 * 2


def test_noalloc_fft_rejects_what_it_cannot_hold(vsp, ctx):
    import torch
    lib_mod = __import__("importlib").import_module("vote-saver-protocol_b200._lib")
    E_INVALID, E_DOMAIN = lib_mod.VSP_E_INVALID, lib_mod.VSP_E_DOMAIN
    lib = ctx._lib
    d = torch.zeros(((1 << MAX_LOG) * 2 + 8, 4), dtype=torch.int64, device="cuda")
    p = C.c_void_p(d.data_ptr())
    # beyond the static scratch: a clear error, nothing launched
    launches0 = ctx.launch_count
    rc = lib.vsp_fr_fft_noalloc_dev(ctx._h, 1 << (MAX_LOG + 1), 0, p)
    assert rc == E_INVALID
    msg = (lib.vsp_last_error(ctx._h) or b"").decode()
    assert "exceeds the static scratch" in msg and "2^16" in msg
    # not a basic radix-2 size (step domain sizes go through vsp_fr_fft_dev), degenerate sizes, null pointers
    for m in (0, 1, 3, 24, (1 << 15) + (1 << 14)):
        assert lib.vsp_fr_fft_noalloc_dev(ctx._h, m, 0, p) == E_DOMAIN, m
    assert lib.vsp_fr_fft_noalloc_dev(ctx._h, 16, 0, None) == E_INVALID
    assert lib.vsp_fr_fft_noalloc_dev(None, 16, 0, p) == E_INVALID
    assert ctx.launch_count == launches0
    ctx.synchronize()
    assert not d.any().item()
