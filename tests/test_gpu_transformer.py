"""K12: transformer kernels and the NT / NN / TN grouped GEMM with two-level groups vs PyTorch fp32 oracles."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


@pytest.fixture(scope="module", autouse=True)
def _lib(native_lib):
    return native_lib


def _bf(*shape, scale=1.0):
    return (torch.randn(*shape, device=DEV) * scale).to(torch.bfloat16)


@pytest.mark.parametrize("mode", ["nt", "nn", "tn"])
@pytest.mark.parametrize("G,M,N,K", [(3, 200, 96, 136), (2, 64, 768, 320), (5, 136, 24, 72)])
def test_mm_modes_one_level(mode, G, M, N, K):
    from fedml_b200.ops import transformer as tf
    torch.manual_seed(0)
    if mode == "nt":
        a, b = _bf(G, M, K), _bf(G, N, K)
        ref = a.float() @ b.float().transpose(1, 2)
    elif mode == "nn":
        a, b = _bf(G, M, K), _bf(G, K, N)
        ref = a.float() @ b.float()
    else:
        a, b = _bf(G, K, M), _bf(G, K, N)
        ref = a.float().transpose(1, 2) @ b.float()
    out = tf.mm(mode, a, b, out_dtype=torch.float32)
    torch.testing.assert_close(out, ref, rtol=1e-3, atol=5e-2)
    out16 = tf.mm(mode, a, b)
    torch.testing.assert_close(out16.float(), ref, rtol=2e-2, atol=2e-1)


def test_mm_bias_relu_and_unaligned_extents():
    """Inner extents that are not multiples of 8 (classifier with 20 labels) on 16-byte aligned strides."""
    from fedml_b200.ops import transformer as tf
    torch.manual_seed(1)
    C, B, D, ncls = 4, 8, 64, 20
    x = _bf(C, B, D)
    w = _bf(C, ncls, D, scale=0.2)
    bias = torch.randn(C, ncls, device=DEV)
    logits = tf.mm("nt", x, w, out_dtype=torch.float32, bias=bias)
    torch.testing.assert_close(logits, x.float() @ w.float().transpose(1, 2) + bias[:, None, :], rtol=1e-3, atol=2e-2)
    y = tf.mm("nt", x, w, bias=bias, relu=True)
    torch.testing.assert_close(y.float(), torch.relu(x.float() @ w.float().transpose(1, 2) + bias[:, None, :]), rtol=2e-2,
                               atol=5e-2)
    dl = torch.zeros(C, B, 24, device=DEV, dtype=torch.bfloat16)[:, :, :ncls]
    dl.copy_(torch.randn(C, B, ncls, device=DEV))
    gw = torch.zeros(C, ncls * D + 8, device=DEV)
    out = gw[:, :ncls * D].view(C, ncls, D)
    tf.mm("tn", dl, x, out=out)
    torch.testing.assert_close(out, dl.float().transpose(1, 2) @ x.float(), rtol=1e-3, atol=2e-2)
    assert float(gw[:, ncls * D:].abs().sum()) == 0
    dx = tf.mm("nn", dl, w)
    torch.testing.assert_close(dx.float(), dl.float() @ w.float(), rtol=2e-2, atol=5e-2)


def test_attention_products_in_place_heads():
    """Q K^T, P V and the four gradient products read / write heads in place inside [CB, S, 3D] buffers."""
    from fedml_b200.ops import transformer as tf
    torch.manual_seed(2)
    CB, S, H, dh = 6, 64, 4, 32
    D = H * dh
    qkv = _bf(CB, S, 3 * D, scale=0.5)
    heads = lambda t, j: t.view(CB, S, t.shape[-1] // D, H, dh)[:, :, j].permute(0, 2, 1, 3)   # noqa: E731
    q, k, v = (heads(qkv, j) for j in range(3))
    scores = tf.mm("nt", q, k, out_dtype=torch.float32)
    ref_s = q.float() @ k.float().transpose(-1, -2)
    torch.testing.assert_close(scores, ref_s, rtol=1e-3, atol=2e-2)
    P = torch.softmax(ref_s / math.sqrt(dh), dim=-1).to(torch.bfloat16)
    ctx = torch.full((CB, S, D), 7.0, device=DEV, dtype=torch.bfloat16)
    tf.mm("nn", P, v, out=heads(ctx, 0))
    ref_ctx = (P.float() @ v.float()).permute(0, 2, 1, 3).reshape(CB, S, D)
    torch.testing.assert_close(ctx.float(), ref_ctx, rtol=2e-2, atol=2e-2)
    dctx = _bf(CB, S, D)
    dO = heads(dctx, 0)
    dqkv = torch.zeros_like(qkv)
    dq, dk, dv = (heads(dqkv, j) for j in range(3))
    tf.mm("tn", P, dO, out=dv)
    dP = tf.mm("nt", dO, v, out_dtype=torch.float32)
    torch.testing.assert_close(dP, dO.float() @ v.float().transpose(-1, -2), rtol=1e-3, atol=5e-2)
    dS = _bf(CB, H, S, S, scale=0.1)
    tf.mm("nn", dS, k, out=dq)
    tf.mm("tn", dS, q, out=dk)
    torch.testing.assert_close(dv.float(), P.float().transpose(-1, -2) @ dO.float(), rtol=2e-2, atol=5e-2)
    torch.testing.assert_close(dq.float(), dS.float() @ k.float(), rtol=2e-2, atol=5e-2)
    torch.testing.assert_close(dk.float(), dS.float().transpose(-1, -2) @ q.float(), rtol=2e-2, atol=5e-2)


def test_embed_and_layernorm_kernels():
    from fedml_b200.models.batched import NativeBackend, TorchBackend
    torch.manual_seed(3)
    C, B, S, D, V = 3, 4, 16, 96, 211
    ld = V * D + 32 * D + 2 * D + 12
    ld = (ld + 3) // 4 * 4
    arena = torch.randn(C, ld, device=DEV) * 0.5
    garena = torch.zeros_like(arena)
    word, pos = arena[:, :V * D].view(C, V, D), arena[:, V * D:V * D + 32 * D].view(C, 32, D)
    gamma, beta = arena[:, V * D + 32 * D:][:, :D], arena[:, V * D + 33 * D:][:, :D]
    gword, gpos = garena[:, :V * D].view(C, V, D), garena[:, V * D:V * D + 32 * D].view(C, 32, D)
    dgam, dbet = garena[:, V * D + 32 * D:][:, :D], garena[:, V * D + 33 * D:][:, :D]
    ids = torch.randint(0, V, (C, B * S), device=DEV)
    nb, tb = NativeBackend(torch.bfloat16), TorchBackend(torch.float32)
    e = nb.embed_fwd(ids, word, pos, S)
    e_ref = tb.embed_fwd(ids, word, pos, S)
    torch.testing.assert_close(e.float(), e_ref, rtol=1e-2, atol=1e-2)
    other = _bf(C, B * S, D)
    y, z, mean, rstd = nb.ln_fwd(e, other, gamma, beta, 1e-12)
    y_ref, z_ref, mean_ref, rstd_ref = tb.ln_fwd(e.float(), other.float(), gamma, beta, 1e-12)
    torch.testing.assert_close(z.float(), z_ref, rtol=1e-2, atol=1e-2)
    torch.testing.assert_close(mean, mean_ref, rtol=1e-2, atol=1e-2)
    torch.testing.assert_close(y.float(), y_ref, rtol=2e-2, atol=3e-2)
    dy, dy2 = _bf(C, B * S, D), _bf(C, B * S, D)
    dz = nb.ln_bwd(dy, dy2, z, mean, rstd, gamma, dgam, dbet)
    g_ref, b_ref = torch.zeros(C, D, device=DEV), torch.zeros(C, D, device=DEV)
    dz_ref = tb.ln_bwd(dy.float(), dy2.float(), z.float(), mean, rstd, gamma, g_ref, b_ref)
    torch.testing.assert_close(dz.float(), dz_ref, rtol=2e-2, atol=3e-2)
    torch.testing.assert_close(dgam, g_ref, rtol=1e-3, atol=1e-2)
    torch.testing.assert_close(dbet, b_ref, rtol=1e-3, atol=1e-2)
    nb.embed_bwd(ids, dz, gword, gpos, S)
    gw_ref, gp_ref = torch.zeros(C, V, D, device=DEV), torch.zeros(C, 32, D, device=DEV)
    tb.embed_bwd(ids, dz.float(), gw_ref, gp_ref, S)
    torch.testing.assert_close(gword, gw_ref, rtol=1e-4, atol=1e-3)
    torch.testing.assert_close(gpos, gp_ref, rtol=1e-4, atol=1e-3)
    # fused sparse SGD: the same scatter applied to the parameter rows with word_scale = -lr (0 freezes a client)
    lr = torch.tensor([0.1, 0.0, 0.3], device=DEV)
    w_before = word.clone()
    nb.embed_bwd(ids, dz, word, gpos, S, word_scale=-lr)
    torch.testing.assert_close(word, w_before - lr[:, None, None] * gw_ref, rtol=1e-4, atol=1e-4)
    assert torch.equal(word[1], w_before[1])
    assert not nb.fallbacks, nb.fallbacks


def test_softmax_act_dropout_colsum_kernels():
    from fedml_b200.models.batched import NativeBackend, TorchBackend
    from fedml_b200.ops import transformer as tf
    torch.manual_seed(4)
    CB, H, S = 5, 3, 40
    nb, tb = NativeBackend(torch.bfloat16), TorchBackend(torch.float32)
    scores = torch.randn(CB, H, S, S, device=DEV) * 3
    mask = (torch.rand(CB, S, device=DEV) > 0.3).float()
    mask[:, 0] = 1
    P, Pd = nb.softmax_fwd(scores, mask, H, 0.25)
    P_ref, _ = tb.softmax_fwd(scores, mask, H, 0.25)
    assert Pd is P
    torch.testing.assert_close(P.float(), P_ref, rtol=1e-2, atol=4e-3)
    dP = torch.randn(CB, H, S, S, device=DEV)
    dS = nb.softmax_bwd(dP, P, 0.25)
    torch.testing.assert_close(dS.float(), tb.softmax_bwd(dP, P.float(), 0.25), rtol=2e-2, atol=2e-3)
    # fused attention dropout == separate dropout kernel with the same key; backward regenerates the same mask
    step = torch.tensor([5], dtype=torch.int32, device=DEV)
    P2, Pd2 = nb.softmax_fwd(scores, mask, H, 0.25, 0.2, 1234, 17, step)
    assert torch.equal(P2, P)
    Pd_sep = tf.dropout(P, 0.2, 1234, 17, step)
    assert torch.equal(Pd2, Pd_sep)
    keep = Pd2 != 0
    frac = float(keep[P != 0].float().mean())
    assert 0.75 < frac < 0.85, frac
    dS2 = nb.softmax_bwd(dP, P, 0.25, 0.2, 1234, 17, step)
    g = dP * keep / 0.8
    ref2 = 0.25 * P.float() * (g - (g * P.float()).sum(-1, keepdim=True))
    torch.testing.assert_close(dS2.float(), ref2, rtol=2e-2, atol=2e-3)
    step.add_(1)
    assert not torch.equal(tf.dropout(P, 0.2, 1234, 17, step), Pd_sep)     # a new step draws a new mask
    # activations
    x = _bf(7, 33, 64, scale=2.0)
    dy = _bf(7, 33, 64)
    for kind in ("gelu", "relu"):
        torch.testing.assert_close(nb.act_fwd(x, kind).float(), tb.act_fwd(x.float(), kind), rtol=1e-2, atol=1e-2)
        torch.testing.assert_close(nb.act_bwd(dy, x, kind).float(), tb.act_bwd(dy.float(), x.float(), kind), rtol=1e-2, atol=1e-2)
    # bias gradients, also on a strided column block
    wide = _bf(4, 300, 3 * 64)
    out = torch.zeros(4, 200, device=DEV)
    nb.colsum(wide[:, :, 64:128], out[:, 8:72])
    torch.testing.assert_close(out[:, 8:72], wide[:, :, 64:128].float().sum(1), rtol=1e-3, atol=5e-2)
    assert float(out[:, :8].abs().sum()) == 0 and float(out[:, 72:].abs().sum()) == 0
    assert not nb.fallbacks, nb.fallbacks


@pytest.mark.parametrize("dropout", [0.0, 0.1])
def test_native_batched_distilbert_step(dropout):
    """Whole client-batched DistilBERT step on the native back-end vs the fp32 PyTorch back-end (dropout off), and a
    smoke run with dropout on (finite loss, no library fall-backs)."""
    from fedml_b200.models.batched import NativeBackend, TorchBackend, make_batched
    from fedml_b200.models.nlp import DistilBertForSequenceClassification
    from fedml_b200.parallel.arena import ArenaLayout
    torch.manual_seed(0)
    C, B, S = 3, 4, 32
    mk = lambda: DistilBertForSequenceClassification(num_labels=20, seq_classif_dropout=dropout, vocab_size=512, dim=128,   # noqa: E731
                                                     n_layers=2, n_heads=4, hidden_dim=256, max_pos=64, dropout=dropout,
                                                     attention_dropout=dropout)
    models = [mk() for _ in range(C)]
    layout = ArenaLayout(models[0])
    stack = torch.stack([layout.flatten(m.state_dict()) for m in models]).to(DEV)
    stack += 0.02 * torch.randn_like(stack)
    ids = torch.randint(1, 512, (C, B, S), device=DEV)
    ids[:, :, 24:] = 0
    y = torch.randint(0, 20, (C, B), device=DEV)
    y[2, 3] = -1
    scale = 1.0 / (y >= 0).sum(dim=1).float()
    active = torch.ones(C, device=DEV)
    outs = {}
    for name, be in (("torch", TorchBackend(torch.float32)), ("native", NativeBackend(torch.bfloat16))):
        st, gst = stack.clone(), torch.zeros_like(stack)
        W, G = layout.stacked_views(st), layout.stacked_views(gst)
        bm = make_batched(models[0], be)
        loss = bm.train_step(ids, y, W, G, active, row_scale=scale)
        outs[name] = (loss.clone(), gst.clone(), getattr(be, "fallbacks", {}), bm.logits(ids, W).float())
    assert not outs["native"][2], outs["native"][2]
    assert torch.isfinite(outs["native"][0]).all() and torch.isfinite(outs["native"][1]).all()
    if dropout == 0.0:
        torch.testing.assert_close(outs["native"][0], outs["torch"][0], rtol=3e-2, atol=3e-2)
        gn, gt = outs["native"][1], outs["torch"][1]
        cos = torch.nn.functional.cosine_similarity(gn, gt, dim=1)
        assert float(cos.min()) > 0.98, cos
        rel = (gn - gt).norm(dim=1) / gt.norm(dim=1)
        assert float(rel.max()) < 0.2, rel
    torch.testing.assert_close(outs["native"][3], outs["torch"][3], rtol=5e-2, atol=5e-2)


@pytest.mark.parametrize("mode,G,M,N,K", [("nt", 3, 1024, 768, 320), ("nt", 2, 640, 384, 200), ("nn", 2, 1000, 256, 136),
                                          ("tn", 2, 512, 3072, 328), ("nt", 1, 2048, 130 * 8, 512)])
def test_mm_cta_pair_tiles(mode, G, M, N, K):
    """Shapes large enough for the cta_group::2 path (256 x bn tiles shared by two CTAs of a cluster), incl. ragged edges."""
    from fedml_b200.ops import transformer as tf
    torch.manual_seed(0)
    if mode == "nt":
        a, b = _bf(G, M, K), _bf(G, N, K)
        ref = a.float() @ b.float().transpose(1, 2)
    elif mode == "nn":
        a, b = _bf(G, M, K), _bf(G, K, N)
        ref = a.float() @ b.float()
    else:
        a, b = _bf(G, K, M), _bf(G, K, N)
        ref = a.float().transpose(1, 2) @ b.float()
    bias = torch.randn(G, N, device=DEV)
    out = tf.mm(mode, a, b, out_dtype=torch.float32, bias=bias)
    torch.testing.assert_close(out, ref + bias[:, None, :], rtol=1e-3, atol=5e-2)
    out16 = tf.mm(mode, a, b, relu=True)
    torch.testing.assert_close(out16.float(), torch.relu(ref), rtol=2e-2, atol=2e-1)


def test_distilbert_fused_sgd_restage_matches_separate_step():
    """DistilBERT executor: optimizer step + bf16 re-staging in one launch (plus the sparse embedding update) == the
    separate fused-SGD kernel over the dense gradient arena, over two steps."""
    from fedml_b200 import ops
    from fedml_b200.models.batched import NativeBackend, make_batched
    from fedml_b200.models.nlp import DistilBertForSequenceClassification
    from fedml_b200.parallel.arena import ArenaLayout
    torch.manual_seed(0)
    C, B, S = 3, 4, 32
    model = DistilBertForSequenceClassification(num_labels=20, seq_classif_dropout=0.0, vocab_size=512, dim=128, n_layers=2,
                                                n_heads=4, hidden_dim=256, max_pos=64, dropout=0.0, attention_dropout=0.0)
    layout = ArenaLayout(model)
    base = torch.stack([layout.flatten(model.state_dict()) for _ in range(C)]).to(DEV)
    base += 0.02 * torch.randn_like(base)
    ids = torch.randint(1, 512, (C, B, S), device=DEV)
    y = torch.randint(0, 20, (C, B), device=DEV)
    active = torch.ones(C, device=DEV)
    lr = torch.tensor([0.05, 0.0, 0.02], device=DEV)
    results = []
    for fused in (False, True):
        st, gst = base.clone(), torch.zeros_like(base)
        W, G = layout.stacked_views(st), layout.stacked_views(gst)
        ex = make_batched(model, NativeBackend(torch.bfloat16))
        if fused:
            ex.set_sparse_sgd(lr)
            ex.set_fused_sgd(lr)
            ex.restage(W)
        for _ in range(2):
            ex.train_step(ids, y, W, G, active)
            if fused:
                assert ex.fused_applied
            else:
                ops.sgd_step_stacked(st, gst, layout.P_train, lr)
        results.append(st.clone())
    a, b = results
    assert torch.equal(b[1], base[1])
    torch.testing.assert_close(b, a, rtol=1e-3, atol=2e-5)
    assert not torch.equal(a[0], base[0])
