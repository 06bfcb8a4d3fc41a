"""fb_round_single_noalloc: the round kernel for hosts that must not allocate device memory.

Some embedders cannot call cudaMalloc when a round is aggregated — they gave all free HBM to their
own pool at start-up, or cannot accept the device-wide stall an allocation implies, or run as a plug-in
whose host forbids allocator calls after initialisation.  For them the library carries the clipping
scratch inside its own image (a fixed device array) and exposes one extra C entry point.  This test
drives that entry point through the C ABI (ctypes on the .so, no Python wrapper in between) and checks
it against the existing fb_round_launch — whose scratch the caller supplies — on two calls in a row,
because the built-in array keeps its contents between calls and the second call must not be disturbed
by what the first one left.  The file name sorts last so that a `-x` run has already executed every
other test when it gets here.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _lib():
    from fedml_b200.native.build import build
    path = build()
    L = C.CDLL(path)
    L.fb_last_error.restype = C.c_char_p
    L.fb_round_single_noalloc.restype = C.c_int
    L.fb_round_single_noalloc.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_float, C.c_float,
                                          C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
    L.fb_round_single_noalloc_max_clients.restype = C.c_int
    return L


def _existing_entry_point(L, stack, weights, inv_total, clip, g, P, P_train, call_no, scratch):
    """fb_round_launch on one rank with caller-provided scratch (the path that already had tests)."""
    from fedml_b200.ops.native import FbRoundParams
    p = FbRoundParams()
    p.clients, p.ld, p.C, p.weights = stack.data_ptr(), stack.stride(0), stack.shape[0], weights.data_ptr()
    p.P, p.P_train, p.inv_total, p.inv_total_buf = P, P_train, inv_total, inv_total
    p.global_model, p.partial = g.data_ptr(), scratch["partial"].data_ptr()
    p.rank, p.G, p.epoch = 0, 1, call_no
    p.clip_bound, p.grad_scale, p.bias_corr1, p.bias_corr2 = clip, 1.0, 1.0, 1.0
    p.sqnorm_part, p.clip_scale_out = scratch["sq"].data_ptr(), scratch["cs"].data_ptr()
    grid = L.fb_round_grid(C.c_int64(P), C.c_int32(1), C.c_int32(1))
    p.grid_bar, p.grid_bar_target = scratch["bar"].data_ptr(), grid * call_no
    rc = L.fb_round_launch(C.byref(p), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, L.fb_last_error()


def test_noalloc_round_matches_existing_entry_point_twice():
    L = _lib()
    P, P_train, Cn = 4 * 6007, 4 * 5500, 11
    torch.manual_seed(0)
    g_new = torch.randn(P, device=DEV)
    g_old = g_new.clone()
    grid = L.fb_round_grid(C.c_int64(P), C.c_int32(1), C.c_int32(1))
    scratch = {"partial": torch.zeros(P, device=DEV), "sq": torch.zeros(grid * Cn, device=DEV),
               "cs": torch.zeros(Cn, device=DEV), "bar": torch.zeros(4, dtype=torch.int32, device=DEV)}
    cs_new = torch.zeros(Cn, device=DEV)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for call_no in (1, 2):
        stack = g_new[None, :] + (0.5 if call_no == 1 else 0.02) * torch.randn(Cn, P, device=DEV)
        weights = torch.randint(5, 60, (Cn,), device=DEV).float()
        inv_total = 1.0 / float(weights.sum())
        g64, x64, w64 = g_new.double().clone(), stack.double(), weights.double()
        rc = L.fb_round_single_noalloc(stack.data_ptr(), stack.stride(0), Cn, weights.data_ptr(), inv_total, 3.0,
                                       g_new.data_ptr(), P, P_train, cs_new.data_ptr(), stream)
        assert rc == 0, L.fb_last_error()
        _existing_entry_point(L, stack, weights, inv_total, 3.0, g_old, P, P_train, call_no, scratch)
        torch.cuda.synchronize()
        torch.testing.assert_close(g_new, g_old, rtol=1e-5, atol=2e-5)
        torch.testing.assert_close(cs_new, scratch["cs"], rtol=1e-5, atol=1e-6)
        # first call clips (noise 0.5 >> bound), second does not: both regimes are exercised
        assert (float(cs_new.max()) < 1.0) == (call_no == 1)
        # and an independent fp64 restatement of the formula in the header
        d = x64[:, :P_train] - g64[:P_train]
        s_c = 1.0 / torch.clamp(d.norm(dim=1) / 3.0, min=1.0)
        ref = (w64[:, None] * x64).sum(0) * inv_total
        ref[:P_train] = (w64[:, None] * (g64[:P_train] + s_c[:, None] * d)).sum(0) * inv_total
        torch.testing.assert_close(g_new.double(), ref, rtol=1e-5, atol=2e-5)
    assert torch.isfinite(g_new).all()


def test_noalloc_round_rejects_oversize_clearly():
    L = _lib()
    limit = L.fb_round_single_noalloc_max_clients()
    P = 4 * 64
    stack = torch.zeros(limit + 1, P, device=DEV)
    w = torch.ones(limit + 1, device=DEV)
    g = torch.zeros(P, device=DEV)
    rc = L.fb_round_single_noalloc(stack.data_ptr(), P, limit + 1, w.data_ptr(), 1.0, 1.0, g.data_ptr(), P, P, None,
                                   C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = L.fb_last_error().decode()
    assert str(limit) in msg and "fb_round_launch" in msg
