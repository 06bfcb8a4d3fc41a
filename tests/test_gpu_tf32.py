"""TF32 path (fp32 operands in HBM, tcgen05 kind::tf32, fp32 accumulate and output): the precision the reference trains
at (stock ``nn.Conv2d`` under cuDNN's TF32 default, python/fedml/simulation/single_process/fedavg/
my_model_trainer_classification.py:18-91).  Oracles are plain PyTorch in fp32/fp64 with TF32 disabled; the operand
truncation of the tensor core (10 mantissa bits kept) bounds the error at ~2^-10 relative per product."""
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


@pytest.fixture(scope="module", autouse=True)
def _lib(native_lib):
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    return native_lib


def _trunc(t):
    """What the tensor core sees of an fp32 operand: sign, exponent, top 10 mantissa bits."""
    return (t.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)


@pytest.mark.parametrize("mode", ["nt", "nn", "tn"])
@pytest.mark.parametrize("G,M,N,K", [(3, 200, 96, 136), (2, 64, 768, 320), (5, 136, 24, 72), (2, 640, 256, 256)])
def test_mm_tf32_modes(mode, G, M, N, K):
    from fedml_b200.ops import transformer as tf
    torch.manual_seed(0)
    r = lambda *s: torch.randn(*s, device=DEV)   # noqa: E731
    if mode == "nt":
        a, b = r(G, M, K), r(G, N, K)
        ref = _trunc(a).double() @ _trunc(b).double().transpose(1, 2)
    elif mode == "nn":
        a, b = r(G, M, K), r(G, K, N)
        ref = _trunc(a).double() @ _trunc(b).double()
    else:
        a, b = r(G, K, M), r(G, K, N)
        ref = _trunc(a).double().transpose(1, 2) @ _trunc(b).double()
    assert tf.mm_supported(mode, a, b)
    out = tf.mm(mode, a, b)
    assert out.dtype == torch.float32
    # against the truncated-operand product only fp32 accumulation order differs
    torch.testing.assert_close(out.double(), ref, rtol=1e-4, atol=1e-3)
    # and against the exact fp32 product the TF32 operand rounding shows: ~1e-3 relative of the row scale
    exact = {"nt": lambda: a.double() @ b.double().transpose(1, 2), "nn": lambda: a.double() @ b.double(),
             "tn": lambda: a.double().transpose(1, 2) @ b.double()}[mode]()
    assert float((out.double() - exact).abs().max()) < 4e-3 * float(exact.abs().max())


def test_mm_tf32_strided_weights_bias_relu_and_reduce_add():
    """Operands as strided views (1x1-conv weights read in place from rows of a parameter arena), bias + ReLU epilogue,
    split-K slices reduce-adding into an fp32 output."""
    from fedml_b200 import ops
    from fedml_b200.ops import transformer as tf
    torch.manual_seed(1)
    G, M, N, K = 4, 320, 64, 96
    arena = torch.randn(G, 3 * N * K + 8, device=DEV)
    w = arena[:, 4:4 + N * K].view(G, N, K)            # 16-byte aligned, group stride = arena row
    a = torch.randn(G, M, K, device=DEV)
    bias = torch.randn(G, N, device=DEV)
    out = tf.mm("nt", a, w, bias=bias, relu=True)
    ref = torch.relu(_trunc(a).double() @ _trunc(w).double().transpose(1, 2) + bias.double()[:, None, :])
    torch.testing.assert_close(out.double(), ref, rtol=1e-4, atol=1e-3)
    # TN with split-K (weight gradient shape): [G, Kpix, cout]^T x [G, Kpix, cin]
    Kp = 2560
    dy, x = torch.randn(G, Kp, 64, device=DEV), torch.randn(G, Kp, 32, device=DEV)
    assert ops.grouped_gemm_tn_supported(dy, x)
    garena = torch.zeros(G, 64 * 32 + 16, device=DEV)
    gout = garena[:, 8:8 + 64 * 32].view(G, 64, 32)
    ops.grouped_gemm_tn(dy, x, out=gout, out_is_zeroed=True)
    ref = _trunc(dy).double().transpose(1, 2) @ _trunc(x).double()
    torch.testing.assert_close(gout.double(), ref, rtol=1e-4, atol=2e-3)
    assert float(garena[:, :8].abs().sum()) == 0 and float(garena[:, 8 + 64 * 32:].abs().sum()) == 0


@pytest.mark.parametrize("B,H,cin,cout,stride", [(10, 32, 16, 16, 1), (4, 16, 32, 32, 1), (10, 32, 32, 32, 2), (6, 8, 64, 64, 1),
                                                  (4, 8, 128, 128, 1), (8, 16, 64, 128, 2), (10, 32, 8, 16, 1)])
def test_implicit_conv3x3_tf32_forward_and_dgrad(B, H, cin, cout, stride):
    import torch.nn.functional as F
    from fedml_b200 import ops
    torch.manual_seed(0)
    C = 3
    x = torch.randn(C, B, H, H, cin, device=DEV)
    w = torch.randn(C, cout, cin, 3, 3, device=DEV) * 0.1
    wm = w.permute(0, 1, 3, 4, 2).reshape(C, cout, 9 * cin).contiguous()
    assert ops.conv3x3_supported(x, cout, stride)
    y = ops.conv3x3(x, wm, stride)
    assert y.dtype == torch.float32
    for c in range(C):
        ref = F.conv2d(_trunc(x[c]).double().permute(0, 3, 1, 2), _trunc(w[c]).double(), stride=stride, padding=1).permute(0, 2, 3, 1)
        torch.testing.assert_close(y[c].double(), ref, rtol=1e-4, atol=1e-4 * float(ref.abs().max()))
    if stride == 1 and ops.conv3x3_supported(torch.empty(C, B, H, H, cout, device=DEV), cin, 1):
        dy = torch.randn(C, B, H, H, cout, device=DEV)
        wt = wm.transpose(1, 2).contiguous()
        dx = ops.conv3x3_dgrad(dy, wt)
        for c in range(C):
            ref = F.conv_transpose2d(_trunc(dy[c]).double().permute(0, 3, 1, 2), _trunc(w[c]).double(), stride=1,
                                     padding=1).permute(0, 2, 3, 1)
            torch.testing.assert_close(dx[c].double(), ref, rtol=1e-4, atol=1e-4 * float(ref.abs().max()))


@pytest.mark.parametrize("C,B,H,cin,cout,stride", [(40, 4, 16, 32, 32, 1), (3, 10, 32, 32, 32, 2),
                                                    (2, 6, 8, 64, 64, 1), (2, 4, 8, 128, 128, 1), (3, 8, 16, 64, 128, 2),
                                                    (2, 16, 4, 256, 64, 1)])
def test_implicit_conv3x3_tf32_wgrad(C, B, H, cin, cout, stride):
    import torch.nn.functional as F
    from fedml_b200 import ops
    torch.manual_seed(0)
    x = torch.randn(C, B, H, H, cin, device=DEV)
    Ho = (H - 1) // stride + 1
    dy = torch.randn(C, B, Ho, Ho, cout, device=DEV)
    assert ops.conv3x3_wgrad_supported(x, dy, stride)
    arena = torch.zeros(C, cout * 9 * cin + 64, device=DEV)
    out = arena[:, 32:32 + cout * 9 * cin].view(C, cout, 9 * cin)
    dw = ops.conv3x3_wgrad(x, dy, stride, out=out, out_is_zeroed=True)
    for c in range(min(C, 3)):
        w = torch.zeros(cout, cin, 3, 3, device=DEV, dtype=torch.float64, requires_grad=True)
        y = F.conv2d(_trunc(x[c]).double().permute(0, 3, 1, 2), w, stride=stride, padding=1)
        y.backward(_trunc(dy[c]).double().permute(0, 3, 1, 2))
        ref = w.grad.permute(0, 2, 3, 1).reshape(cout, 9 * cin)
        torch.testing.assert_close(dw[c].double(), ref, rtol=1e-4, atol=1e-4 * float(ref.abs().max()))
    assert float(arena[:, :32].abs().sum()) == 0 and float(arena[:, 32 + cout * 9 * cin:].abs().sum()) == 0
