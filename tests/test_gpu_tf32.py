"""TF32 path (fp32 operands in HBM, tcgen05 kind::tf32, fp32 accumulate and output): the precision the reference trains
at (stock ``nn.Conv2d`` under cuDNN's TF32 default, python/fedml/simulation/single_process/fedavg/
my_model_trainer_classification.py:18-91).  Oracles are plain PyTorch in fp32/fp64 with TF32 disabled; the operand
truncation of the tensor core (10 mantissa bits kept) bounds the error at ~2^-10 relative per product."""
import os

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
                                                    (2, 16, 4, 256, 64, 1),
                                                    # 16 input channels: one zero-padded 32-column tile per tap, 16 columns stored
                                                    (13, 10, 32, 16, 16, 1), (3, 4, 16, 16, 32, 2), (100, 2, 8, 16, 64, 1)])
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


@pytest.mark.parametrize("ch,M,relu,with_res", [(16, 640, True, False), (64, 10240, True, True), (256, 333, False, False),
                                                (512, 64, True, True), (8, 5000, True, False)])
def test_bn_stacked_fp32_activations(ch, M, relu, with_res):
    """BatchNorm(+residual, +ReLU) forward / backward on fp32 activations vs the plain PyTorch oracle."""
    from fedml_b200 import ops
    from fedml_b200.models.batched import TorchBackend
    C, P = 5, 4096
    torch.manual_seed(ch + M)
    x = torch.randn(C, 1, 1, M, ch, device=DEV) * 2 + 0.5
    res = torch.randn(C, 1, 1, M, ch, device=DEV) if with_res else None
    arena = torch.randn(C, P, device=DEV)
    gamma, beta = arena[:, :ch], arena[:, ch:2 * ch]
    rm0, rv0 = torch.randn(C, ch, device=DEV), torch.rand(C, ch, device=DEV) + 0.5
    active = torch.tensor([1.0, 0.0, 1.0, 1.0, 0.0], device=DEV)
    be = TorchBackend(torch.float32)
    rm_a, rv_a, nbt_a = rm0.clone(), rv0.clone(), torch.zeros(C, device=DEV)
    rm_b, rv_b, nbt_b = rm0.clone(), rv0.clone(), torch.zeros(C, device=DEV)
    y_ref, mean_ref, rstd_ref = be.bn_fwd(x, res, gamma, beta, rm_a, rv_a, nbt_a, active, 0.1, 1e-5, relu, True)
    y, mean, rstd = ops.bn_fwd_stacked(x, res, gamma, beta, rm_b, rv_b, nbt_b, active, 0.1, 1e-5, relu, True)
    assert y.dtype == torch.float32
    torch.testing.assert_close(mean, mean_ref, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(rstd, rstd_ref, rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(y, y_ref, rtol=1e-3, atol=1e-3)
    torch.testing.assert_close(rm_b, rm_a, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(rv_b, rv_a, rtol=1e-3, atol=1e-4)
    assert torch.equal(nbt_b, active)
    dy = torch.randn_like(x)
    dx_ref, dres_ref, dg_ref, db_ref = be.bn_bwd(dy, x, y_ref, gamma, mean_ref, rstd_ref, relu, with_res)
    dx, dres, dg, db = ops.bn_bwd_stacked(dy, x, y, gamma, mean, rstd, relu, with_res)
    scale = float(dx_ref.abs().max())
    torch.testing.assert_close(db, db_ref, rtol=1e-3, atol=1e-2)
    torch.testing.assert_close(dg, dg_ref, rtol=1e-3, atol=2e-2)
    torch.testing.assert_close(dx, dx_ref, rtol=2e-3, atol=2e-3 * scale)
    if with_res:
        torch.testing.assert_close(dres, dres_ref, rtol=1e-5, atol=1e-6)
    elif relu:            # mask recomputed from x instead of reading y: identical in fp32 up to values within round-off of 0
        dx2, _, dg2, db2 = ops.bn_bwd_stacked(dy, x, None, gamma, mean, rstd, relu, False, beta=beta)
        assert float(((dx2 - dx).abs() > 2e-3 * scale).float().mean()) < 1e-3
    two = (torch.randn_like(x), torch.randn_like(x))      # gradient arriving over two paths is summed on load
    dx3, _, _, _ = ops.bn_bwd_stacked(two[0], x, y, gamma, mean, rstd, relu, False, dy2=two[1])
    dx3_ref, _, _, _ = be.bn_bwd(two, x, y_ref, gamma, mean_ref, rstd_ref, relu, False)
    torch.testing.assert_close(dx3, dx3_ref, rtol=2e-3, atol=2e-3 * float(dx3_ref.abs().max()))
    y_e_ref, _, _ = be.bn_fwd(x, res, gamma, beta, rm_a, rv_a, nbt_a, active, 0.1, 1e-5, relu, False)
    y_e, _, _ = ops.bn_fwd_stacked(x, res, gamma, beta, rm_b, rv_b, nbt_b, active, 0.1, 1e-5, relu, False)
    torch.testing.assert_close(y_e, y_e_ref, rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("H,ch,stride", [(32, 16, 1), (16, 32, 2), (8, 64, 1), (7, 8, 2)])
def test_im2col_col2im_fp32(H, ch, stride):
    from fedml_b200 import ops
    from fedml_b200.models.batched import TorchBackend
    C, B = 3, 2
    torch.manual_seed(H)
    x = torch.randn(C, B, H, H, ch, device=DEV)
    be = TorchBackend(torch.float32)
    cols = ops.im2col3x3(x, stride)
    assert cols.dtype == torch.float32 and torch.equal(cols, be.im2col3x3(x, stride))
    d = torch.randn_like(cols)
    torch.testing.assert_close(ops.col2im3x3(d, B, H, H, ch, stride), be.col2im3x3(d, B, H, H, ch, stride), rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("block", ["bottleneck", "basic"])
def test_native_tf32_resnet_step_matches_fp32_oracle(block):
    """Whole client-batched training step with fp32 activations / TF32 tensor-core products vs the plain PyTorch fp32
    back-end (TF32 off): every gradient of every client within TF32 round-off of the oracle, no library fallback for
    convolutions, BatchNorm or im2col."""
    from fedml_b200.models.batched import NativeBackend, TorchBackend, make_batched
    from fedml_b200.models.cv import BasicBlock, Bottleneck, ResNet
    from fedml_b200.parallel.arena import ArenaLayout
    torch.manual_seed(0)
    C, B = 4, 8
    blk = Bottleneck if block == "bottleneck" else BasicBlock
    models = [ResNet(blk, [1, 1, 1], num_classes=10) for _ in range(C)]
    layout = ArenaLayout(models[0])
    stack = torch.stack([layout.flatten(m.state_dict()) for m in models]).to(DEV)
    x = torch.randint(0, 256, (C, B, 32, 32, 3), dtype=torch.uint8, device=DEV)
    y = torch.randint(0, 10, (C, B), device=DEV)
    active = torch.ones(C, device=DEV)
    outs = {}
    for name, be in (("torch", TorchBackend(torch.float32)), ("native", NativeBackend(torch.float32))):
        st, gst = stack.clone(), torch.zeros_like(stack)
        W, G = layout.stacked_views(st), layout.stacked_views(gst)
        bm = make_batched(models[0], be)
        loss = bm.train_step(x, y, W, G, active)
        outs[name] = (loss.clone(), gst.clone(), st.clone(), dict(getattr(be, "fallbacks", {})))
    bad = {k: v for k, v in outs["native"][3].items() if k in ("bn_fwd", "bn_bwd", "im2col", "col2im", "gemm_nt")}
    assert not bad, outs["native"][3]
    torch.testing.assert_close(outs["native"][0], outs["torch"][0], rtol=2e-3, atol=2e-3)
    gn, gt = outs["native"][1][:, :layout.P_train], outs["torch"][1][:, :layout.P_train]
    cos = torch.nn.functional.cosine_similarity(gn, gt, dim=1)
    rel = (gn - gt).norm(dim=1) / gt.norm(dim=1)
    print("tf32 whole-step: cos", cos.tolist(), "rel", rel.tolist())
    # TF32 operands (10 mantissa bits) through ~10 BatchNorm-normalised layers at batch 8: measured cos 0.9995-0.9999,
    # relative error 1-3 % of the gradient norm (the bf16 path of the same step: cos > 0.98, 20 %)
    assert float(cos.min()) > 0.999, cos
    assert float(rel.max()) < 5e-2, rel
    # running statistics (buffers) agree
    torch.testing.assert_close(outs["native"][2][:, layout.P_train:], outs["torch"][2][:, layout.P_train:], rtol=2e-3, atol=2e-3)


def test_tf32_fused_sgd_restage_matches_separate_optimizer_step():
    """fp32-compute train_step ending in fb_sgd_restage_f32 (in-place 1x1 / fc weights, fp32 staged 3x3 operands) ==
    train_step + separate stacked SGD, over three steps."""
    from fedml_b200 import ops
    from fedml_b200.models.batched import NativeBackend, make_batched
    from fedml_b200.models.cv import Bottleneck, ResNet
    from fedml_b200.parallel.arena import ArenaLayout
    torch.manual_seed(0)
    C, B = 3, 8
    model = ResNet(Bottleneck, [1, 1, 1], num_classes=10)
    layout = ArenaLayout(model)
    base = torch.stack([layout.flatten(model.state_dict()) for _ in range(C)]).to(DEV)
    base += 0.01 * torch.randn_like(base)
    x = torch.randint(0, 256, (C, B, 32, 32, 3), dtype=torch.uint8, device=DEV)
    y = torch.randint(0, 10, (C, B), device=DEV)
    active = torch.ones(C, device=DEV)
    lr = torch.tensor([0.05, 0.0, 0.02], device=DEV)          # client 1 frozen
    results = []
    for fused in (False, True):
        st, gst = base.clone(), torch.zeros_like(base)
        W, G = layout.stacked_views(st), layout.stacked_views(gst)
        ex = make_batched(model, NativeBackend(torch.float32))
        if fused:
            ex.set_fused_sgd(lr)
            ex.restage(W)
        snaps = []
        for _ in range(3):
            ex.train_step(x, y, W, G, active)
            if fused:
                assert ex.fused_applied
            else:
                ops.sgd_step_stacked(st, gst, layout.P_train, lr)
            snaps.append(st.clone())
        results.append(snaps)
    a, b = results
    assert torch.equal(a[2][1, :layout.P_train], base[1, :layout.P_train]) and torch.equal(b[2][1, :layout.P_train], base[1, :layout.P_train])
    # step 1 starts from identical weights: only the order of the split-K reduce-adds differs
    torch.testing.assert_close(b[0], a[0], rtol=1e-5, atol=2e-6)
    # later steps: a last-bit difference in an fp32 operand can move its TF32 value by 2^-10, so the two runs drift apart
    # at TF32 resolution, not at fp32 resolution
    torch.testing.assert_close(b[2], a[2], rtol=1e-2, atol=2e-3)
    assert not torch.equal(a[2][0], base[0])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("G,M,N,K", [(5, 10240, 16, 64), (3, 200, 96, 136), (100, 640, 64, 256), (2, 640, 512, 256),
                                     (7, 1000, 24, 72), (300, 130, 32, 32)])
def test_gemm_epilogue_column_statistics(dtype, G, M, N, K):
    """Column sums / sums of squares of the GEMM output accumulated by the epilogue (BatchNorm batch statistics of a
    1x1 convolution's output) vs torch on the stored output."""
    from fedml_b200 import ops
    torch.manual_seed(G + M + N)
    a = torch.randn(G, M, K, device=DEV).to(dtype)
    b = (torch.randn(G, N, K, device=DEV) * 0.2).to(dtype)
    stats = torch.zeros(G, 2, N, device=DEV)
    d = ops.grouped_gemm(a, b, dtype, stats=stats)
    d0 = ops.grouped_gemm(a, b, dtype)
    assert torch.equal(d, d0)                                      # the tile walk differs, the result must not
    df = d.double()
    torch.testing.assert_close(stats[:, 0].double(), df.sum(dim=1), rtol=1e-4, atol=1e-3 * M ** 0.5)
    torch.testing.assert_close(stats[:, 1].double(), (df * df).sum(dim=1), rtol=1e-4, atol=1e-3 * M ** 0.5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("C,B,H,cin,cout,stride", [(4, 10, 32, 16, 16, 1), (6, 4, 16, 32, 64, 2), (3, 6, 8, 64, 64, 1)])
def test_conv3x3_epilogue_statistics_feed_batchnorm(dtype, C, B, H, cin, cout, stride):
    """conv3x3(..., stats=) then BatchNorm from those statistics == conv3x3 then BatchNorm with its own statistics pass."""
    from fedml_b200 import ops
    torch.manual_seed(1)
    x = torch.randn(C, B, H, H, cin, device=DEV).to(dtype)
    wm = (torch.randn(C, cout, 9 * cin, device=DEV) * 0.1).to(dtype)
    stats = torch.zeros(C, 2, cout, device=DEV)
    y = ops.conv3x3(x, wm, stride, stats=stats)
    assert torch.equal(y, ops.conv3x3(x, wm, stride))
    yd = y.double().reshape(C, -1, cout)
    torch.testing.assert_close(stats[:, 0].double(), yd.sum(dim=1), rtol=1e-4, atol=1e-2)
    torch.testing.assert_close(stats[:, 1].double(), (yd * yd).sum(dim=1), rtol=1e-4, atol=1e-2)
    arena = torch.randn(C, 4 * cout, device=DEV)
    gamma, beta = arena[:, :cout], arena[:, cout:2 * cout]
    active = torch.ones(C, device=DEV)
    outs = []
    for st in (None, stats):
        rm, rv, nbt = torch.zeros(C, cout, device=DEV), torch.ones(C, cout, device=DEV), torch.zeros(C, device=DEV)
        o = ops.bn_fwd_stacked(y, None, gamma, beta, rm, rv, nbt, active, 0.1, 1e-5, True, True, stats=st)
        outs.append((o, rm, rv, nbt))
    (ya, ma, ra), (yb, mb, rb) = outs[0][0], outs[1][0]
    torch.testing.assert_close(mb, ma, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(rb, ra, rtol=1e-3, atol=1e-5)
    torch.testing.assert_close(yb.float(), ya.float(), rtol=1e-2 if dtype == torch.bfloat16 else 1e-3, atol=2e-2 if dtype == torch.bfloat16 else 1e-3)
    for i in (1, 2, 3):
        torch.testing.assert_close(outs[1][i], outs[0][i], rtol=1e-3, atol=1e-5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("C,B,H,ch,ncls", [(5, 10, 8, 256, 10), (3, 7, 4, 512, 100), (2, 16, 2, 64, 3)])
def test_fused_classifier_head_matches_oracle(dtype, C, B, H, ch, ncls):
    """pool -> fc -> cross-entropy -> fc gradients -> d(h) in one launch vs the composition of PyTorch ops (fp32)."""
    from fedml_b200.models.batched import NativeBackend, TorchBackend
    torch.manual_seed(C * 100 + ncls)
    h = torch.randn(C, B, H, H, ch, device=DEV).to(dtype)
    arena = torch.randn(C, ncls * ch + ncls + 8, device=DEV) * 0.1
    garena = torch.zeros(2, C, ncls * ch + ncls + 8, device=DEV)
    fc_w, fc_b = arena[:, :ncls * ch].view(C, ncls, ch), arena[:, ncls * ch:ncls * ch + ncls]
    y = torch.randint(0, ncls, (C, B), device=DEV)
    y[0, -2:] = -1                                            # padding rows
    scale = 1.0 / (y >= 0).sum(dim=1).clamp_min(1).float()
    res = []
    for i, be in enumerate((TorchBackend(torch.float32), NativeBackend(dtype))):
        g_w, g_b = garena[i][:, :ncls * ch].view(C, ncls, ch), garena[i][:, ncls * ch:ncls * ch + ncls]
        loss, dh = be.head_fwd_bwd(h if i else h.float(), fc_w, fc_b, y, scale, g_w, g_b)
        res.append((loss, dh, g_w, g_b))
        assert not getattr(be, "fallbacks", {})
    (l0, d0, w0, b0), (l1, d1, w1, b1) = res
    torch.testing.assert_close(l1, l0, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(w1, w0, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(b1, b0, rtol=1e-4, atol=1e-6)
    tol = dict(rtol=1e-4, atol=1e-7) if dtype == torch.float32 else dict(rtol=1e-2, atol=1e-5)
    torch.testing.assert_close(d1.float(), d0, **tol)
    assert float(garena[:, :, ncls * ch + ncls:].abs().sum()) == 0


@pytest.mark.parametrize("model,dataset,dtype", [("resnet56", "cifar10", "tf32"), ("resnet56", "cifar10", "bf16"), ("lr", "mnist", "fp32")])
def test_simulator_native_evaluation_matches_eager(model, dataset, dtype):
    """NCCLSimulator.evaluate on the native kernels (batched eval forward + fused fb_eval_counts) == the eager PyTorch
    evaluation of the same global model."""
    import fedml_b200 as fedml
    from fedml_b200.arguments import Arguments
    from fedml_b200.simulation import SimulatorNCCL
    a = fedml.init(Arguments.from_dict(dict(
        training_type="simulation", backend="NCCL", dataset=dataset, model=model, client_num_in_total=8, client_num_per_round=4,
        comm_round=1, epochs=1, batch_size=10, learning_rate=0.03, partition_method="homo", frequency_of_the_test=0,
        synthetic_train_size=8 * 40, log_level="ERROR", using_gpu=True, compute_dtype=dtype, random_seed=0, mnist_client_num=8)))
    ds, ncls = fedml.data.load(a)
    torch.manual_seed(0)
    sim = SimulatorNCCL(a, torch.device(DEV), ds, fedml.models.create(a, ncls)).simulator
    sim.train_one_round(0)
    from fedml_b200.ops import native
    before = native.LAUNCHES[0]
    nat = sim.evaluate(0)
    assert nat.get("eval_path") == "native" and native.LAUNCHES[0] > before
    a.native_eval = False
    eag = sim.evaluate(0)
    assert eag.get("eval_path") is None
    tol = 3e-2 if dtype == "bf16" else 5e-3
    for k in ("train_acc", "test_acc"):
        assert abs(nat[k] - eag[k]) <= tol + 1e-9, (k, nat[k], eag[k])
    for k in ("train_loss", "test_loss"):
        assert abs(nat[k] - eag[k]) <= tol * max(1.0, abs(eag[k])), (k, nat[k], eag[k])
    assert 0.0 <= nat["test_precision"] <= 1.0 and 0.0 <= nat["test_recall"] <= 1.0


def test_augment_kernel_matches_oracle_and_runs_inside_the_training_graph():
    """csrc/augment.cu == the CPU oracle byte for byte (crop / flip / Cutout from one Philox block per sample), with a
    host nonce and with a device nonce; the simulator with ``augment: true`` trains with CUDA graphs on."""
    from fedml_b200 import ops
    from fedml_b200.data.augment import augment_cifar_reference
    g = torch.Generator().manual_seed(0)
    x = torch.randint(0, 256, (300, 32, 32, 3), dtype=torch.uint8, generator=g)
    ref = augment_cifar_reference(x, seed=(5 << 33) | 77, nonce=9, first_index=1000)
    got = ops.augment_cifar(x.to(DEV), (5 << 33) | 77, 9, first_index=1000)
    assert torch.equal(got.cpu(), ref)
    nonce = torch.tensor([9], dtype=torch.int64, device=DEV)
    assert torch.equal(ops.augment_cifar(x.to(DEV), (5 << 33) | 77, nonce, first_index=1000).cpu(), ref)
    nonce.fill_(10)
    assert not torch.equal(ops.augment_cifar(x.to(DEV), (5 << 33) | 77, nonce, first_index=1000).cpu(), ref)
    x1 = torch.randint(0, 256, (7, 16, 24, 1), dtype=torch.uint8, generator=g)          # other shapes / channel counts
    assert torch.equal(ops.augment_cifar(x1.to(DEV), 3, 1, pad=2, cutout=6, mean_u8=(99,)).cpu(),
                       augment_cifar_reference(x1, 3, 1, pad=2, cutout=6, mean_u8=(99,)))
    import fedml_b200 as fedml
    from fedml_b200.arguments import Arguments
    from fedml_b200.simulation import SimulatorNCCL
    a = fedml.init(Arguments.from_dict(dict(
        training_type="simulation", backend="NCCL", dataset="cifar10", model="resnet56", client_num_in_total=8, client_num_per_round=4,
        comm_round=2, epochs=1, batch_size=10, learning_rate=0.03, partition_method="homo", frequency_of_the_test=0,
        synthetic_train_size=8 * 40, log_level="ERROR", using_gpu=True, compute_dtype="tf32", random_seed=0, augment=True)))
    ds, ncls = fedml.data.load(a)
    sim = SimulatorNCCL(a, torch.device(DEV), ds, fedml.models.create(a, ncls)).simulator
    before = sim.global_flat.clone()
    for r in range(2):
        sim.train_one_round(r)
    torch.cuda.synchronize()
    assert sim.graph is not None and sim._static.get("aug_nonce") is not None and int(sim._static["aug_nonce"]) > 0
    assert torch.isfinite(sim.global_flat).all() and not torch.equal(before, sim.global_flat)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("ch,HW,cpg,relu,with_res", [(64, 1024, 2, True, False), (128, 64, 2, True, True), (512, 16, 2, False, False),
                                                      (32, 256, 8, True, True)])
def test_groupnorm_stacked_kernels(dtype, ch, HW, cpg, relu, with_res):
    """GroupNorm(+residual, +ReLU) forward / backward on the BatchNorm kernels (rows = client x sample, group pooling in
    the apply prologue) vs the plain PyTorch formulation."""
    from fedml_b200 import ops
    from fedml_b200.models.batched import TorchBackend
    C, B = 3, 6
    torch.manual_seed(ch + HW)
    x = (torch.randn(C, B, 1, HW, ch, device=DEV) * 2 + 0.5).to(dtype)
    res = torch.randn(C, B, 1, HW, ch, device=DEV).to(dtype) if with_res else None
    arena = torch.randn(C, 4096, device=DEV)
    gamma, beta = arena[:, :ch], arena[:, ch:2 * ch]
    be = TorchBackend(torch.float32)
    y_ref, mean_ref, rstd_ref = be.gn_fwd(x.float(), None if res is None else res.float(), gamma, beta, cpg, 1e-5, relu)
    y, mean, rstd = ops.gn_fwd_stacked(x, res, gamma, beta, cpg, 1e-5, relu)
    tol = dict(rtol=1e-3, atol=1e-3) if dtype == torch.float32 else dict(rtol=2e-2, atol=3e-2)
    torch.testing.assert_close(mean, mean_ref, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(rstd, rstd_ref, rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(y.float(), y_ref, **tol)
    dy = torch.randn_like(x)
    dx_ref, dres_ref, dg_ref, db_ref = be.gn_bwd(dy.float(), x.float(), y_ref, gamma, mean_ref, rstd_ref, cpg, relu, with_res)
    y_mask = y_ref.to(dtype)                                    # same ReLU mask in both arms
    dx, dres, dg, db = ops.gn_bwd_stacked(dy, x, y_mask, gamma, mean, rstd, cpg, relu, with_res)
    scale = float(dx_ref.abs().max())
    gt = dict(rtol=2e-3, atol=2e-2) if dtype == torch.float32 else dict(rtol=3e-2, atol=0.5)
    torch.testing.assert_close(db, db_ref, **gt)
    torch.testing.assert_close(dg, dg_ref, **gt)
    torch.testing.assert_close(dx.float(), dx_ref, rtol=3e-2, atol=(3e-3 if dtype == torch.float32 else 4e-2) * scale)
    if with_res:
        torch.testing.assert_close(dres.float(), dres_ref, rtol=1e-2, atol=1e-2)


def test_native_groupnorm_resnet18_step_matches_fp32_oracle():
    """resnet18_gn-style network (BasicBlock, GroupNorm with two channels per group) on the native tf32 path: whole
    client-batched training step vs the plain PyTorch back-end, no library fallback."""
    from fedml_b200.models.batched import NativeBackend, TorchBackend, make_batched
    from fedml_b200.models.cv import BasicBlock, ResNet
    from fedml_b200.parallel.arena import ArenaLayout
    torch.manual_seed(0)
    C, B = 4, 8
    model = ResNet(BasicBlock, [1, 1, 1], num_classes=10, stem="cifar", group_norm=2)
    layout = ArenaLayout(model)
    stack = torch.stack([layout.flatten(model.state_dict()) for _ in range(C)]).to(DEV)
    stack += 0.01 * torch.randn_like(stack)
    x = torch.randint(0, 256, (C, B, 32, 32, 3), dtype=torch.uint8, device=DEV)
    y = torch.randint(0, 10, (C, B), device=DEV)
    outs = {}
    for name, be in (("torch", TorchBackend(torch.float32)), ("native", NativeBackend(torch.float32))):
        st, gst = stack.clone(), torch.zeros_like(stack)
        bm = make_batched(model, be)
        loss = bm.train_step(x, y, layout.stacked_views(st), layout.stacked_views(gst), torch.ones(C, device=DEV))
        outs[name] = (loss.clone(), gst.clone(), dict(getattr(be, "fallbacks", {})))
    assert not outs["native"][2], outs["native"][2]
    torch.testing.assert_close(outs["native"][0], outs["torch"][0], rtol=2e-3, atol=2e-3)
    gn, gt = outs["native"][1][:, :layout.P_train], outs["torch"][1][:, :layout.P_train]
    cos = torch.nn.functional.cosine_similarity(gn, gt, dim=1)
    assert float(cos.min()) > 0.999, cos


@pytest.mark.gpu
@pytest.mark.parametrize("resident", ["0", "1", "persist", "persist64"])
def test_bn_single_launch_kernels_small_ragged_shapes(resident):
    """Fused (per-client barrier) and slab-resident (shared-memory slices, bulk TMA in and out) BatchNorm on shapes whose
    rows do not divide by the slice count, fp32 and bf16, with / without residual, second gradient, residual gradient and
    recomputed ReLU mask, against the fp32 PyTorch restatement in tools/bench_bn.py.  The switches are read once per process,
    hence the subprocess."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, FB_BN_FUSED="1", FB_BN_RESIDENT="1" if resident.startswith("persist") else resident)
    if resident.startswith("persist"):          # persistent pipelined kernels; 64 KB budget = several items per CTA
        env["FB_BN_PERSIST"] = "1"
        if resident == "persist64":
            env["FB_BN_PERSIST_KB"] = "64"
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "bench_bn.py"), "--small"], env=env, capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0 and "BN_SMALL_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
