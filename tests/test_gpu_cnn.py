"""FEMNIST CNN_DropOut on the native back-end (reference model/cv/cnn.py:74-142): max-pool kernels and the whole
client-batched step against the fp32 PyTorch restatement (``TorchBackend``)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("H,W,ch,off,out_hw", [(26, 26, 64, (1, 1), (12, 12)), (8, 12, 8, (0, 0), None), (9, 7, 16, (1, 0), (4, 3))])
def test_maxpool2x2_kernels_match_oracle(dtype, H, W, ch, off, out_hw):
    from fedml_b200.models.batched import NativeBackend, TorchBackend
    torch.manual_seed(H * 31 + ch)
    nb, tb = NativeBackend(dtype), TorchBackend(torch.float32)
    x = torch.randn(3, 2, H, W, ch, device=DEV).to(dtype)
    x[0, 0, :4, :4] = 0.5                                   # ties: the first element of a window takes the gradient
    x[1] = x[1].relu()                                      # ReLU output: windows that are all zero
    y = nb.maxpool2x2_fwd(x, off, out_hw)
    yr = tb.maxpool2x2_fwd(x.float(), off, out_hw)
    torch.testing.assert_close(y.float(), yr, rtol=0, atol=0)
    dy = torch.randn_like(y)
    for relu_mask in (False, True):
        dx = nb.maxpool2x2_bwd(dy, x, off, relu_mask)
        dxr = tb.maxpool2x2_bwd(dy.float(), x.float(), off, relu_mask)
        torch.testing.assert_close(dx.float(), dxr, rtol=0, atol=0)
    assert not nb.fallbacks, nb.fallbacks


@pytest.mark.parametrize("dtype,tol", [(torch.bfloat16, 0.05), (torch.float32, 5e-3)])
def test_native_cnn_dropout_step_matches_fp32_oracle(dtype, tol):
    from fedml_b200.models.batched import NativeBackend, TorchBackend, make_batched
    from fedml_b200.models.cv import CNN_DropOut
    from fedml_b200.parallel.arena import ArenaLayout
    torch.manual_seed(1)
    C, B = 4, 10
    model = CNN_DropOut(only_digits=False)
    model.dropout_1.p = model.dropout_2.p = 0.0
    layout = ArenaLayout(model)
    stack = torch.stack([layout.flatten(CNN_DropOut(only_digits=False).state_dict()) for _ in range(C)]).to(DEV)
    x = torch.randn(C, B, 784, device=DEV)
    y = torch.randint(0, 62, (C, B), device=DEV)
    outs = {}
    for name, be in (("native", NativeBackend(dtype)), ("torch", TorchBackend(torch.float32))):
        g = torch.zeros_like(stack)
        bm = make_batched(model, be)
        loss = bm.train_step(x, y, layout.stacked_views(stack), layout.stacked_views(g), torch.ones(C, device=DEV))
        outs[name] = (loss.float(), g, bm.logits(x, layout.stacked_views(stack)).float())
        if name == "native":
            assert not be.fallbacks, be.fallbacks           # every primitive of the step ran on this repo's kernels
    torch.testing.assert_close(outs["native"][0], outs["torch"][0], rtol=tol, atol=tol)
    torch.testing.assert_close(outs["native"][2], outs["torch"][2], rtol=tol, atol=tol)
    gn, gt = outs["native"][1][:, :layout.P_train], outs["torch"][1][:, :layout.P_train]
    cos = torch.nn.functional.cosine_similarity(gn, gt, dim=1)
    # tf32: a fc-1 unit whose pre-activation is within rounding of zero flips its ReLU mask and with it a whole 9216-wide row
    # of the weight gradient (seen: 1 of 1280 units, cosine 0.998); everything else agrees to 1e-4
    assert float(cos.min()) > 0.995 and float(cos.median()) > (0.995 if dtype == torch.bfloat16 else 0.9995), cos


def test_native_cnn_dropout_runs_with_dropout_in_a_cuda_graph():
    """bf16 step with dropout on, captured and replayed: losses differ between replays (new masks) and stay finite."""
    from fedml_b200.models.batched import NativeBackend, make_batched
    from fedml_b200.models.cv import CNN_DropOut
    from fedml_b200.parallel.arena import ArenaLayout
    torch.manual_seed(2)
    C, B = 3, 8
    model = CNN_DropOut(only_digits=True)
    layout = ArenaLayout(model)
    stack = torch.stack([layout.flatten(model.state_dict()) for _ in range(C)]).to(DEV)
    g = torch.zeros_like(stack)
    be = NativeBackend(torch.bfloat16)
    bm = make_batched(model, be)
    x = torch.randn(C, B, 784, device=DEV)
    y = torch.randint(0, 10, (C, B), device=DEV)
    W, G, act = layout.stacked_views(stack), layout.stacked_views(g), torch.ones(C, device=DEV)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(2):
            bm.train_step(x, y, W, G, act)
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        loss = bm.train_step(x, y, W, G, act)
    graph.replay(); torch.cuda.synchronize(); l1 = loss.clone()
    graph.replay(); torch.cuda.synchronize(); l2 = loss.clone()
    assert torch.isfinite(l1).all() and torch.isfinite(l2).all() and not torch.equal(l1, l2)
    assert not be.fallbacks, be.fallbacks


def test_fp32_elementwise_kernels_match_bf16_forms_and_oracle():
    """fb_act_fwd/bwd_f32, fb_dropout_f32, fb_colsum_f32: same masks and maths as the bf16 kernels, fp32 storage."""
    from fedml_b200.ops import transformer as T
    torch.manual_seed(0)
    x = torch.randn(3, 40, 64, device=DEV)
    dy = torch.randn_like(x)
    for kind, ref in (("relu", torch.relu), ("gelu", torch.nn.functional.gelu)):
        torch.testing.assert_close(T.act_fwd(x, kind), ref(x), rtol=1e-5, atol=1e-6)
        xr = x.clone().requires_grad_(True)
        ref(xr).backward(dy)
        torch.testing.assert_close(T.act_bwd(dy, x, kind), xr.grad, rtol=1e-4, atol=1e-5)
    step = torch.tensor([7], dtype=torch.int32, device=DEV)
    d32 = T.dropout(x, 0.3, 11, 5, step)
    d16 = T.dropout(x.bfloat16(), 0.3, 11, 5, step)
    assert torch.equal(d32 == 0, d16 == 0)                                # one mask for both storage types
    kept = d32 != 0
    torch.testing.assert_close(d32[kept], (x / 0.7)[kept], rtol=1e-6, atol=1e-6)
    assert 0.2 < float((~kept).float().mean()) < 0.4
    out = torch.zeros(3, 64, device=DEV)
    torch.testing.assert_close(T.colsum(x, out), x.sum(1), rtol=1e-5, atol=1e-5)
