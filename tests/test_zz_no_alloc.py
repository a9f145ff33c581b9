"""GPU: the allocation-free entry point ``zt_instnorm_nhwc_f32_noalloc`` (include/zerotig.h) against the existing
``zt_instnorm_nhwc_f32``, called through the C-ABI (ctypes, raw device pointers).

The variant keeps its fp64 statistics in a file-scope ``static __device__`` array that every call shares and that
holds the previous call's sums, so the test makes TWO calls in a row with different inputs and shapes and checks
each against the caller-scratch entry point (same kernels, same order of operations: the outputs must be identical,
and both must be within the 5e-6 of tests/test_gpu_ops.py::test_instance_norm of the fp64 torch reference).  Sizes
beyond the static maximum must be rejected with a clear error.  The file is named so that pytest collects it last.
"""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

TOL = 5e-6      # tests/test_gpu_ops.py::test_instance_norm


def _ptr(t):
    return C.c_void_p(t.data_ptr() if t is not None else 0)


def _nhwc(x):
    return x.permute(0, 2, 3, 1).contiguous().cuda()


def _both(L, x, res, relu_inner, relu_outer):
    """(existing entry point, allocation-free entry point, fp64 torch reference), NCHW on the host."""
    B, Cn, H, W = x.shape
    xin, rin = _nhwc(x), (_nhwc(res) if res is not None else None)
    out_a, out_b = torch.full_like(xin, float("nan")), torch.full_like(xin, float("nan"))
    stats = torch.empty(B * Cn * 2, dtype=torch.float64, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = L.zt_instnorm_nhwc_f32(_ptr(xin), Cn, _ptr(out_a), Cn, _ptr(rin), Cn, _ptr(stats), B, H * W, Cn, 1e-5,
                                int(relu_inner), int(relu_outer), st)
    assert rc == 0, L.zt_last_error()
    rc = L.zt_instnorm_nhwc_f32_noalloc(_ptr(xin), Cn, _ptr(out_b), Cn, _ptr(rin), Cn, B, H * W, Cn, 1e-5,
                                        int(relu_inner), int(relu_outer), st)
    assert rc == 0, L.zt_last_error()
    torch.cuda.synchronize()
    ref = F.instance_norm(x.double())
    if relu_inner:
        ref = F.relu(ref)
    if res is not None:
        ref = ref + res.double()
    if relu_outer:
        ref = F.relu(ref)
    back = lambda o: o.view(B, H, W, Cn).permute(0, 3, 1, 2).cpu()
    return back(out_a), back(out_b), ref


def test_instnorm_noalloc_matches_existing_entry_point_twice():
    from zero_tig_b200 import _lib
    L = _lib.lib()
    g = torch.Generator().manual_seed(77)
    # call 1: two images x 128 channels (the fnet layer-3 shape class), residual + both ReLUs, large mean
    x1 = torch.randn(2, 128, 23, 31, generator=g) * 3 + 5
    r1 = torch.randn(2, 128, 23, 31, generator=g)
    a1, b1, ref1 = _both(L, x1, r1, True, True)
    # call 2, straight after: other values, fewer channels, no residual -> stale sums of call 1 would show
    x2 = torch.randn(2, 64, 40, 36, generator=g) * 0.5 - 2
    a2, b2, ref2 = _both(L, x2, None, True, False)
    for a, b, ref in ((a1, b1, ref1), (a2, b2, ref2)):
        assert torch.isfinite(b).all()
        assert (a.double() - ref).abs().max().item() < TOL
        assert (b.double() - ref).abs().max().item() < TOL
        assert (a - b).abs().max().item() < TOL
    # the maximum itself (B*C == ZT_INSTNORM_NOALLOC_MAX_BC) is accepted
    x3 = torch.randn(2, 512, 9, 11, generator=g)
    a3, b3, ref3 = _both(L, x3, None, False, False)
    assert (b3.double() - ref3).abs().max().item() < TOL and (a3 - b3).abs().max().item() < TOL


def test_instnorm_noalloc_rejects_sizes_beyond_the_static_scratch():
    from zero_tig_b200 import _lib
    L = _lib.lib()
    Cn, B, HW = 516, 2, 64                      # B*C = 1032 > 1024
    x = torch.randn(B * HW, Cn, device="cuda")
    out = torch.zeros_like(x)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = L.zt_instnorm_nhwc_f32_noalloc(_ptr(x), Cn, _ptr(out), Cn, _ptr(None), 0, B, HW, Cn, 1e-5, 0, 0, st)
    assert rc != 0
    msg = L.zt_last_error().decode()
    assert "ZT_INSTNORM_NOALLOC_MAX_BC" in msg and "1032" in msg, msg
    torch.cuda.synchronize()
    assert out.abs().max().item() == 0.0        # nothing was launched
