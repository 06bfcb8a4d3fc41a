"""Numerics of the hand-written sm_100a kernels against plain PyTorch fp32/fp64 oracles."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


@pytest.fixture(scope="module", autouse=True)
def _lib(native_lib):
    return native_lib


def _stack(C, P, ld=None, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    ld = ld or P
    buf = torch.randn(C, ld, generator=g).to(DEV)
    return buf[:, :P] if ld != P else buf, buf


def test_weighted_accumulate_fp32_and_bf16():
    from fedml_b200 import ops
    C, P = 13, 4 * 1237
    x, _ = _stack(C, P, ld=P + 64)
    w = torch.rand(C, device=DEV) * 50 + 1
    out = torch.full((P,), 3.0, device=DEV)
    ops.weighted_accumulate(x, w, out, scale=0.25, beta=0.5)
    ref = 0.5 * 3.0 + 0.25 * (w[:, None].double() * x.double()).sum(0)
    torch.testing.assert_close(out.double(), ref, rtol=1e-5, atol=1e-4)
    xb = x.to(torch.bfloat16).contiguous()
    out2 = torch.zeros(P, device=DEV)
    ops.weighted_accumulate(xb, w, out2)
    ref2 = (w[:, None].double() * xb.double()).sum(0)
    torch.testing.assert_close(out2.double(), ref2, rtol=1e-5, atol=1e-3)


def test_broadcast_rows():
    from fedml_b200 import ops
    P, C = 4 * 999, 7
    src = torch.randn(P, device=DEV)
    rows = torch.zeros(C, P + 8, device=DEV)
    ops.broadcast_rows(src, rows[:, :P])
    assert torch.equal(rows[:, :P], src.expand(C, P))
    assert float(rows[:, P:].abs().sum()) == 0.0


@pytest.mark.parametrize("name,kw", [("sgd", {}), ("sgd", {"momentum": 0.9}), ("adam", {}), ("yogi", {}), ("adagrad", {})])
def test_server_opt_matches_oracle(name, kw):
    from fedml_b200 import ops
    from fedml_b200.core.aggregation import ServerOptimizer
    n = 4 * 511
    torch.manual_seed(1)
    w0 = torch.randn(n)
    oracle = ServerOptimizer(name, lr=0.1, momentum=kw.get("momentum", 0.0), beta1=0.9, beta2=0.99, eps=1e-3)
    w_ref = {"p": w0.clone()}
    w = w0.clone().to(DEV)
    m = torch.zeros(n, device=DEV); v = torch.zeros(n, device=DEV)
    for step in range(1, 4):
        avg = w_ref["p"] - 0.05 * torch.randn(n)
        w_ref = oracle.step(w_ref, {"p": avg}, ["p"])
        ops.server_opt(w, avg.to(DEV), m, v, name, lr=0.1, momentum=kw.get("momentum", 0.0), beta1=0.9, beta2=0.99,
                       eps=1e-3, step=step)
        torch.testing.assert_close(w.cpu(), w_ref["p"], rtol=2e-5, atol=2e-6)


def test_sqnorm_diff_and_philox():
    from fedml_b200 import ops
    C, P = 5, 4 * 3001
    x, _ = _stack(C, P)
    ref = torch.randn(P, device=DEV)
    sq = ops.sqnorm_diff(x, ref)
    torch.testing.assert_close(sq.double(), ((x - ref).double() ** 2).sum(1), rtol=1e-5, atol=1e-3)
    z = ops.philox_normal(1 << 20, seed=1234, stream_id=3, substream=7, device=DEV)
    assert abs(float(z.mean())) < 5e-3 and abs(float(z.std()) - 1.0) < 5e-3
    z2 = ops.philox_normal(1 << 20, seed=1234, stream_id=3, substream=7, device=DEV)
    assert torch.equal(z, z2)
    z3 = ops.philox_normal(1 << 20, seed=1234, stream_id=3, substream=8, device=DEV)
    assert abs(float((z * z3).mean())) < 5e-3


def _oracle_round(clients, weights, g, P_train, clip, sigma, seed, round_idx, opt_cfg, m, v, step, inv_total):
    """fp64 restatement of the fused round for all clients of all ranks concatenated."""
    from fedml_b200 import ops
    X = torch.cat([c for c in clients if c is not None], 0).double()
    w = torch.cat([c for c in weights if c is not None], 0).double()
    gd = g.double()
    contrib = X.clone()
    if clip > 0 or sigma > 0:
        diff = X[:, :P_train] - gd[:P_train]
        if clip > 0:
            nrm = diff.norm(dim=1)
            diff = diff / torch.clamp(nrm / clip, min=1.0)[:, None]
        contrib[:, :P_train] = gd[:P_train] + diff
        if sigma > 0:
            for c in range(X.shape[0]):
                z = ops.philox_normal(g.numel(), seed, round_idx, c, g.device)[:P_train].double()
                contrib[c, :P_train] += sigma * z
    avg = (w[:, None] * contrib).sum(0) * inv_total
    return avg


@pytest.mark.parametrize("G,two_shot", [(1, False), (2, False), (2, True), (4, False), (4, True)])
@pytest.mark.parametrize("dp", [False, True])
def test_round_kernel_emulated_matches_oracle(G, two_shot, dp):
    from fedml_b200.parallel.round_engine import DPConfig, ServerOptConfig, emulated_round
    P, P_train = 4 * 5003, 4 * 4500
    torch.manual_seed(G * 10 + int(two_shot))
    g = torch.randn(P, device=DEV)
    counts = [5, 0, 3, 4][:G] if G > 1 else [6]
    clients = [(g + 0.1 * torch.randn(c, P, device=DEV)) if c else None for c in counts]
    weights = [(torch.randint(10, 100, (c,), device=DEV).float()) if c else None for c in counts]
    total = sum(float(w.sum()) for w in weights if w is not None)
    dpc = DPConfig(clip=2.0, sigma=0.01, seed=99) if dp else DPConfig()
    outs, _ = emulated_round(clients, weights, g, P_train, 1.0 / total, two_shot, dp=dpc, round_idx=3)
    ref = _oracle_round(clients, weights, g, P_train, dpc.clip, dpc.sigma, dpc.seed, 3, None, None, None, 1, 1.0 / total)
    for r in range(G):
        torch.testing.assert_close(outs[r].double(), ref, rtol=1e-5, atol=2e-5)
    for r in range(1, G):   # every rank holds the bit-identical model
        assert torch.equal(outs[0], outs[r])


def test_round_kernel_g_invariance_and_fedadam():
    """Same clients split over 1, 2 and 4 emulated ranks give the same FedAdam result (tolerance:
    partial sums are grouped differently), and two consecutive rounds keep moments consistent."""
    from fedml_b200.core.aggregation import ServerOptimizer
    from fedml_b200.parallel.round_engine import ServerOptConfig, emulated_round
    P, P_train, K = 4 * 2048, 4 * 2000, 8
    torch.manual_seed(7)
    g0 = torch.randn(P, device=DEV)
    opt = ServerOptConfig("adam", lr=0.05, beta1=0.9, beta2=0.99, eps=1e-3)
    results = []
    for G in (1, 2, 4):
        g = g0.clone()
        state = None
        for rnd in range(2):
            torch.manual_seed(100 + rnd)
            X = g0 + 0.05 * torch.randn(K, P, device=DEV)
            w = torch.arange(1, K + 1, device=DEV).float()
            per = K // G
            clients = [X[r * per:(r + 1) * per] for r in range(G)]
            weights = [w[r * per:(r + 1) * per].contiguous() for r in range(G)]
            outs, state = emulated_round(clients, weights, g, P_train, 1.0 / float(w.sum()), G > 1 and rnd == 1, opt=opt,
                                         opt_step=rnd + 1, epoch=rnd + 1, state=state)
        results.append(outs[0].clone())
    # Adam divides by sqrt(v): a last-bit difference in the fp32 partial-sum grouping is amplified
    # where g ~ 0, hence the looser bound than for plain averaging
    torch.testing.assert_close(results[0], results[1], rtol=1e-3, atol=5e-4)
    torch.testing.assert_close(results[0], results[2], rtol=1e-3, atol=5e-4)
    # oracle for the G=1 path
    oracle = ServerOptimizer("adam", lr=0.05, beta1=0.9, beta2=0.99, eps=1e-3)
    wref = {"p": g0[:P_train].cpu().clone(), "b": g0[P_train:].cpu().clone()}
    for rnd in range(2):
        torch.manual_seed(100 + rnd)
        X = (g0 + 0.05 * torch.randn(K, P, device=DEV)).cpu()
        w = torch.arange(1, K + 1).float()
        avg = (w[:, None] * X).sum(0) / w.sum()
        wref = oracle.step(wref, {"p": avg[:P_train], "b": avg[P_train:]}, ["p"])
    torch.testing.assert_close(results[0][:P_train].cpu(), wref["p"], rtol=1e-3, atol=5e-4)
    torch.testing.assert_close(results[0][P_train:].cpu(), wref["b"], rtol=1e-5, atol=1e-5)


def test_round_engine_single_rank_api():
    from fedml_b200.models import LogisticRegression
    from fedml_b200.parallel.arena import ArenaLayout
    from fedml_b200.parallel.round_engine import RoundEngine
    model = LogisticRegression(784, 10)
    layout = ArenaLayout(model)
    eng = RoundEngine(layout, DEV)
    flat = layout.flatten(model.state_dict(), device=DEV)
    eng.set_global(flat)
    C = 9
    stack = flat[None, :].repeat(C, 1) + 0.01 * torch.randn(C, layout.P, device=DEV)
    w = torch.rand(C, device=DEV) + 0.5
    out = eng.aggregate(stack, w, 1.0 / float(w.sum()))
    ref = (w[:, None].double() * stack.double()).sum(0) / w.double().sum()
    torch.testing.assert_close(out.double(), ref, rtol=1e-5, atol=1e-6)
    sd = layout.unflatten(out)
    assert list(sd.keys()) == list(model.state_dict().keys())


@pytest.mark.parametrize("K", [1, 2, 7, 10, 64, 257])
def test_median_and_trimmed_mean(K):
    from fedml_b200 import ops
    P = 1000 + K
    torch.manual_seed(K)
    x = torch.randn(K, P, device=DEV)
    x[:, :50] = torch.randint(0, 3, (K, 50), device=DEV).float()   # heavy ties
    x[:, 50:80:3] = 5.0     # all clients equal on some coordinates, neighbours differ (lanes of one warp take both cases)
    med = ops.coordinate_median(x)
    assert torch.equal(med, torch.median(x, dim=0).values)
    for trim in {0, K // 10, (K - 1) // 2}:
        if 2 * trim >= K:
            continue
        tm = ops.trimmed_mean(x, trim)
        ref = torch.sort(x.double(), dim=0).values[trim:K - trim].mean(0)
        torch.testing.assert_close(tm.double(), ref, rtol=1e-5, atol=1e-5)


def test_krum_scores_vs_oracle():
    from fedml_b200 import ops
    from fedml_b200.core.aggregation import krum_scores
    K, P, f = 20, 512, 3
    torch.manual_seed(0)
    x = torch.randn(K, P, device=DEV)
    x[:f] += 5.0
    gram = (x.double() @ x.double().t()).float()
    s = ops.krum_scores(gram, f)
    ref = krum_scores(x.cpu(), f)
    torch.testing.assert_close(s.cpu().double(), ref, rtol=2e-3, atol=1e-1)
    assert int(torch.argmin(s)) == int(torch.argmin(ref))


def test_finite_field_ops():
    from fedml_b200 import ops
    p = 2 ** 31 - 1
    rng = np.random.default_rng(0)
    x = torch.randn(10007, device=DEV) * 3
    q = ops.ff_quantize(x, 2 ** 16, p)
    qn = np.round(x.cpu().numpy().astype(np.float64) * 2 ** 16).astype(np.int64) % p
    assert np.array_equal(q.cpu().numpy().astype(np.int64) & 0xFFFFFFFF, qn)
    back = ops.ff_dequantize(q, 2 ** 16, p)
    torch.testing.assert_close(back, x, rtol=0, atol=1.0 / 2 ** 16)
    U = rng.integers(0, p, size=(6, 9), dtype=np.int64)
    X = rng.integers(0, p, size=(9, 1001), dtype=np.int64)
    out = ops.ff_matmul(torch.from_numpy(U.astype(np.int32)).to(DEV), torch.from_numpy(X.astype(np.int32)).to(DEV), p)
    ref = np.zeros((6, 1001), dtype=object)
    for i in range(6):
        for j in range(9):
            ref[i] = (ref[i] + int(U[i, j]) * X[j].astype(object)) % p
    assert np.array_equal(out.cpu().numpy().astype(np.int64), ref.astype(np.int64))
    # pairwise masks cancel
    a = ops.ff_mask(q, [11, 22], [1, 1], p)
    b = ops.ff_mask(q, [11, 22], [-1, -1], p)
    s = ops.ff_add_(a.clone(), b, p)
    two_q = (2 * qn) % p
    assert np.array_equal(s.cpu().numpy().astype(np.int64) & 0xFFFFFFFF, two_q)
    assert not torch.equal(a, q)


def test_sgd_step_stacked_variants():
    from fedml_b200 import ops
    C, P, n = 6, 4 * 300, 4 * 250
    torch.manual_seed(3)
    w = torch.randn(C, P, device=DEV); g = torch.randn(C, P, device=DEV)
    lr = torch.tensor([0.1, 0.2, 0.0, 0.05, 0.1, 0.3], device=DEV)
    anchor = torch.randn(P, device=DEV)
    w0 = w.clone()
    ops.sgd_step_stacked(w, g, n, lr, weight_decay=0.01, mu=0.1, anchor=anchor)
    d = g[:, :n] + 0.01 * w0[:, :n] + 0.1 * (w0[:, :n] - anchor[:n])
    ref = w0.clone(); ref[:, :n] -= lr[:, None] * d
    torch.testing.assert_close(w, ref, rtol=1e-5, atol=1e-6)
    assert torch.equal(w[2], w0[2])
    # momentum + FedNova cum_grad, two steps
    w = w0.clone(); buf = torch.zeros_like(w); cum = torch.zeros_like(w)
    ref_w = w0.clone(); ref_buf = None; ref_cum = torch.zeros_like(w)
    for step in range(2):
        g = torch.randn(C, P, device=DEV)
        ops.sgd_step_stacked(w, g, n, lr, momentum=0.9, mom_buf=buf, first_step=(step == 0), cum_grad=cum)
        d = g[:, :n]
        ref_buf = d.clone() if ref_buf is None else 0.9 * ref_buf + d
        act = (lr != 0)[:, None]
        ref_cum[:, :n] += torch.where(act, lr[:, None] * ref_buf, torch.zeros_like(ref_buf))
        ref_w[:, :n] -= lr[:, None] * ref_buf
    torch.testing.assert_close(w, ref_w, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(cum, ref_cum, rtol=1e-5, atol=1e-6)
    # bf16 gradients
    w = w0.clone(); gb = torch.randn(C, P, device=DEV).to(torch.bfloat16)
    ops.sgd_step_stacked(w, gb, n, lr)
    ref = w0.clone(); ref[:, :n] -= lr[:, None] * gb[:, :n].float()
    torch.testing.assert_close(w, ref, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("ncls,pre_sigmoid", [(10, False), (10, True), (62, False), (1000, False)])
def test_softmax_xent_and_eval(ncls, pre_sigmoid):
    from fedml_b200 import ops
    R = 77
    torch.manual_seed(ncls)
    logits = torch.randn(R, ncls, device=DEV) * 3
    labels = torch.randint(0, ncls, (R,), device=DEV)
    loss, grad = ops.softmax_xent(logits, labels, pre_sigmoid=pre_sigmoid)
    lg = logits.clone().requires_grad_(True)
    z = torch.sigmoid(lg) if pre_sigmoid else lg
    ref_rows = torch.nn.functional.cross_entropy(z, labels, reduction="none")
    ref_rows.mean().backward()
    torch.testing.assert_close(loss, ref_rows.detach(), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(grad, lg.grad, rtol=1e-3, atol=1e-6)
    counts = torch.zeros(2 + 3 * ncls, dtype=torch.int32, device=DEV)
    loss_sum = torch.zeros(1, device=DEV)
    ops.eval_counts(logits, labels, counts, loss_sum, pre_sigmoid=pre_sigmoid)
    pred = z.detach().argmax(1)
    assert int(counts[0]) == int((pred == labels).sum()) and int(counts[1]) == R
    torch.testing.assert_close(loss_sum[0], ref_rows.detach().sum(), rtol=1e-4, atol=1e-3)
    tp = torch.bincount(labels[pred == labels], minlength=ncls)
    fp = torch.bincount(pred[pred != labels], minlength=ncls)
    fn = torch.bincount(labels[pred != labels], minlength=ncls)
    assert torch.equal(counts[2:2 + ncls].long(), tp)
    assert torch.equal(counts[2 + ncls:2 + 2 * ncls].long(), fp)
    assert torch.equal(counts[2 + 2 * ncls:].long(), fn)


@pytest.mark.parametrize("ch,M,relu,with_res", [(16, 640, True, False), (64, 10240, True, True), (256, 333, False, False),
                                                (512, 64, True, True)])
def test_bn_stacked_fwd_bwd_vs_torch_backend(ch, M, relu, with_res):
    from fedml_b200 import ops
    from fedml_b200.models.batched import TorchBackend
    C, P = 5, 4096
    torch.manual_seed(ch + M)
    B, H, W = 1, 1, M
    x = (torch.randn(C, B, H, W, ch, device=DEV) * 2 + 0.5).to(torch.bfloat16)
    res = torch.randn(C, B, H, W, ch, device=DEV).to(torch.bfloat16) if with_res else None
    arena = torch.randn(C, P, device=DEV)
    gamma, beta = arena[:, :ch], arena[:, ch:2 * ch]
    rm0, rv0 = torch.randn(C, ch, device=DEV), torch.rand(C, ch, device=DEV) + 0.5
    active = torch.tensor([1.0, 0.0, 1.0, 1.0, 0.0], device=DEV)
    be = TorchBackend(torch.bfloat16)
    rm_a, rv_a, nbt_a = rm0.clone(), rv0.clone(), torch.zeros(C, device=DEV)
    rm_b, rv_b, nbt_b = rm0.clone(), rv0.clone(), torch.zeros(C, device=DEV)
    y_ref, mean_ref, rstd_ref = be.bn_fwd(x, res, gamma, beta, rm_a, rv_a, nbt_a, active, 0.1, 1e-5, relu, True)
    y, mean, rstd = ops.bn_fwd_stacked(x, res, gamma, beta, rm_b, rv_b, nbt_b, active, 0.1, 1e-5, relu, True)
    torch.testing.assert_close(mean, mean_ref, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(rstd, rstd_ref, rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(y.float(), y_ref.float(), rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(rm_b, rm_a, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(rv_b, rv_a, rtol=1e-3, atol=1e-4)
    assert torch.equal(nbt_b, active)
    dy = torch.randn_like(x)
    dx_ref, dres_ref, dg_ref, db_ref = be.bn_bwd(dy, x, y, gamma, mean, rstd, relu, with_res)
    dx, dres, dg, db = ops.bn_bwd_stacked(dy, x, y, gamma, mean, rstd, relu, with_res)
    torch.testing.assert_close(db, db_ref, rtol=1e-3, atol=1e-2)
    torch.testing.assert_close(dg, dg_ref, rtol=1e-3, atol=2e-2)
    torch.testing.assert_close(dx.float(), dx_ref.float(), rtol=3e-2, atol=3e-2)
    if with_res:
        torch.testing.assert_close(dres.float(), dres_ref.float(), rtol=1e-2, atol=1e-2)
    elif relu:   # mask recomputed from x instead of reading y (bf16 rounding of y can flip values that round to 0)
        dx2, _, dg2, db2 = ops.bn_bwd_stacked(dy, x, None, gamma, mean, rstd, relu, False, beta=beta)
        flips = ((y == 0) != ((x.float().reshape(C, -1, ch) - mean[:, None]) * (rstd * gamma)[:, None] + beta[:, None] <= 0)
                 .reshape(y.shape)).float().mean()
        assert float(flips) < 1e-2
        torch.testing.assert_close(db2, db_ref, rtol=5e-2, atol=0.5)
        torch.testing.assert_close(dx2.float(), dx_ref.float(), rtol=5e-2, atol=0.2)
    # eval mode uses the running statistics
    y_e_ref, _, _ = be.bn_fwd(x, res, gamma, beta, rm_a, rv_a, nbt_a, active, 0.1, 1e-5, relu, False)
    y_e, _, _ = ops.bn_fwd_stacked(x, res, gamma, beta, rm_b, rv_b, nbt_b, active, 0.1, 1e-5, relu, False)
    torch.testing.assert_close(y_e.float(), y_e_ref.float(), rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("H,ch,stride", [(32, 16, 1), (16, 32, 2), (8, 64, 1), (7, 8, 2)])
def test_im2col_col2im_vs_torch_backend(H, ch, stride):
    from fedml_b200 import ops
    from fedml_b200.models.batched import TorchBackend
    C, B = 3, 2
    torch.manual_seed(H)
    x = torch.randn(C, B, H, H, ch, device=DEV).to(torch.bfloat16)
    be = TorchBackend(torch.bfloat16)
    cols_ref = be.im2col3x3(x, stride)
    cols = ops.im2col3x3(x, stride)
    assert torch.equal(cols, cols_ref)
    d = torch.randn_like(cols)
    dx_ref = be.col2im3x3(d.float(), B, H, H, ch, stride)
    dx = ops.col2im3x3(d, B, H, H, ch, stride)
    torch.testing.assert_close(dx.float(), dx_ref, rtol=1e-2, atol=2e-2)


@pytest.mark.parametrize("G,M,N,K", [(1, 128, 16, 64), (3, 200, 16, 16), (2, 1000, 64, 144), (2, 640, 256, 64),
                                     (1, 130, 10, 256), (2, 256, 288, 32), (4, 2560, 32, 288), (1, 512, 512, 1024)])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float32])
def test_grouped_gemm_nt_tcgen05(G, M, N, K, out_dtype):
    from fedml_b200 import ops
    torch.manual_seed(M + N + K)
    a = torch.randn(G, M, K, device=DEV).to(torch.bfloat16)
    b = torch.randn(G, N, K, device=DEV).to(torch.bfloat16)
    d = ops.grouped_gemm(a, b, out_dtype)
    ref = torch.bmm(a.float(), b.float().transpose(1, 2))
    tol = dict(rtol=2e-2, atol=2e-1) if out_dtype == torch.bfloat16 else dict(rtol=1e-3, atol=1e-2)
    torch.testing.assert_close(d.float(), ref, **tol)


def test_grouped_gemm_nt_bias_relu_and_strided_groups():
    from fedml_b200 import ops
    G, M, N, K = 3, 300, 48, 80
    big = torch.randn(G, M + 10, K + 8, device=DEV).to(torch.bfloat16)
    a = big[:, :M, :K]                      # strided rows (lda = K + 8) and group stride
    b = torch.randn(G, N, K, device=DEV).to(torch.bfloat16)
    bias = torch.randn(G, N, device=DEV)
    d = ops.grouped_gemm(a, b, torch.float32, relu=True, bias=bias)
    ref = torch.relu(torch.bmm(a.float(), b.float().transpose(1, 2)) + bias[:, None, :])
    torch.testing.assert_close(d, ref, rtol=1e-3, atol=1e-2)


@pytest.mark.parametrize("G,Kred,M,N", [(1, 64, 64, 64), (2, 640, 16, 144), (3, 1000, 64, 16), (2, 2560, 128, 576),
                                        (1, 100, 256, 64), (2, 128, 200, 72)])
def test_grouped_gemm_tn_tcgen05(G, Kred, M, N):
    from fedml_b200 import ops
    torch.manual_seed(Kred + M + N)
    a = torch.randn(G, Kred, M, device=DEV).to(torch.bfloat16)
    b = torch.randn(G, Kred, N, device=DEV).to(torch.bfloat16)
    d = ops.grouped_gemm_tn(a, b)
    ref = torch.bmm(a.float().transpose(1, 2), b.float())
    torch.testing.assert_close(d, ref, rtol=1e-3, atol=2e-2)


@pytest.mark.parametrize("K,n", [(20, 5000), (64, 100000), (130, 33000)])
def test_gram_split_k_matches_fp64(K, n):
    from fedml_b200 import ops
    torch.manual_seed(K)
    x = torch.randn(K, n, device=DEV) * 0.01
    ld = (n + 63) // 64 * 64
    rows = ops.ptr_table([x.data_ptr() + k * n * 4 for k in range(K)], DEV)
    hi, lo = ops.krum_pack(rows, K, None, 0, n, ld, split=True)
    g = ops.gram_bf16(hi, lo)
    ref = (x.double() @ x.double().t())
    torch.testing.assert_close(g.double(), ref, rtol=1e-4, atol=1e-6)
    g_hi = ops.gram_bf16(hi, None)
    torch.testing.assert_close(g_hi.double(), ref, rtol=3e-2, atol=2e-3)   # bf16-only operands: ~1e-4 rounding noise per entry


@pytest.mark.parametrize("G", [1, 2, 4])
def test_dp_allreduce_step_emulated(G):
    """Cheetah kernel: G emulated ranks in one cooperative launch == mean-gradient SGD with momentum on every replica."""
    import ctypes as C
    from fedml_b200 import ops
    from fedml_b200.ops.native import FbDpStepParams, check, lib, stream_ptr
    n = 4 * 3001
    torch.manual_seed(G)
    w0 = torch.randn(n, device=DEV)
    params = [w0.clone() for _ in range(G)]
    grads = [torch.randn(n, device=DEV) for _ in range(G)]
    moms = [torch.zeros(n, device=DEV) for _ in range(G)]
    flags = [torch.zeros(ops.round_flag_bytes(G) // 4, dtype=torch.int32, device=DEV) for _ in range(G)]
    tab = lambda ts: torch.tensor([t.data_ptr() for t in ts], dtype=torch.int64, device=DEV)
    tg, tp, tf = tab(grads), tab(params), tab(flags)
    ref_w, ref_m = w0.clone(), torch.zeros(n, device=DEV)
    for step in range(1, 3):
        for g in grads:
            g.copy_(torch.randn(n, device=DEV))
        ps = []
        for r in range(G):
            p = FbDpStepParams()
            p.peer_grad, p.peer_param, p.peer_flags, p.mom = tg.data_ptr(), tp.data_ptr(), tf.data_ptr(), moms[r].data_ptr()
            p.n, p.rank, p.G, p.epoch = n, r, G, step
            p.lr, p.momentum, p.weight_decay, p.inv_world, p.nesterov = 0.1, 0.9, 0.01, 1.0 / G, 0
            ps.append(p)
        arr = (FbDpStepParams * G)(*ps)
        dev_arr = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(DEV)
        check(lib().fb_dp_allreduce_step_emulated(C.c_void_p(dev_arr.data_ptr()), C.c_int32(G), C.c_int64(n),
                                                  C.c_void_p(stream_ptr(torch.device(DEV)))), "dp emulated")
        torch.cuda.synchronize()
        gmean = torch.stack(grads).mean(0) + 0.01 * ref_w
        ref_m = 0.9 * ref_m + gmean
        ref_w = ref_w - 0.1 * ref_m
        for r in range(G):
            torch.testing.assert_close(params[r], ref_w, rtol=1e-5, atol=1e-5)
        for r in range(1, G):
            assert torch.equal(params[0], params[r])


def test_native_batched_resnet_step_matches_torch_backend():
    """Whole client-batched training step on the native back-end (tcgen05 GEMMs, stacked BN, im2col, staged weights,
    gradients written straight into the arena) vs the fp32 PyTorch back-end."""
    from fedml_b200.models.batched import NativeBackend, TorchBackend, make_batched
    from fedml_b200.models.cv import Bottleneck, ResNet
    from fedml_b200.parallel.arena import ArenaLayout
    torch.manual_seed(0)
    C, B = 4, 8
    models = [ResNet(Bottleneck, [1, 1, 1], num_classes=10) for _ in range(C)]
    layout = ArenaLayout(models[0])
    stack = torch.stack([layout.flatten(m.state_dict()) for m in models]).to(DEV)
    x = torch.randint(0, 256, (C, B, 32, 32, 3), dtype=torch.uint8, device=DEV)
    y = torch.randint(0, 10, (C, B), device=DEV)
    active = torch.ones(C, device=DEV)
    outs = {}
    for name, be in (("torch", TorchBackend(torch.float32)), ("native", NativeBackend(torch.bfloat16))):
        st, gst = stack.clone(), torch.zeros_like(stack)
        W, G = layout.stacked_views(st), layout.stacked_views(gst)
        bm = make_batched(models[0], be)
        loss = bm.train_step(x, y, W, G, active)
        outs[name] = (loss.clone(), gst.clone(), st.clone(), getattr(be, "fallbacks", {}))
    assert not {k: v for k, v in outs["native"][3].items() if k in ("bn_fwd", "bn_bwd", "im2col", "col2im", "gemm_nt")}, outs["native"][3]
    torch.testing.assert_close(outs["native"][0], outs["torch"][0], rtol=5e-2, atol=5e-2)
    gn, gt = outs["native"][1][:, :layout.P_train], outs["torch"][1][:, :layout.P_train]
    cos = torch.nn.functional.cosine_similarity(gn, gt, dim=1)
    assert float(cos.min()) > 0.98, cos                      # bf16 activations: direction of the full gradient preserved
    rel = (gn - gt).norm(dim=1) / gt.norm(dim=1)
    assert float(rel.max()) < 0.2, rel
    torch.testing.assert_close(outs["native"][2][:, layout.P_train:], outs["torch"][2][:, layout.P_train:], rtol=5e-2, atol=5e-2)


def test_grouped_gemm_tn_split_k_reduce_add_in_place():
    """Split-K slices accumulate with TMA reduce-stores into a pre-zeroed strided output (rows of a gradient arena)."""
    from fedml_b200 import ops
    G, Kred, M, N, P = 3, 2560, 32, 64, 4096
    torch.manual_seed(0)
    a = torch.randn(G, Kred, M, device=DEV).to(torch.bfloat16)
    b = torch.randn(G, Kred, N, device=DEV).to(torch.bfloat16)
    arena = torch.zeros(G, P, device=DEV)
    out = arena[:, 128:128 + M * N].view(G, M, N)
    r = ops.grouped_gemm_tn(a, b, out=out, out_is_zeroed=True)
    assert r.data_ptr() == out.data_ptr()
    ref = torch.bmm(a.float().transpose(1, 2), b.float())
    torch.testing.assert_close(out, ref, rtol=1e-3, atol=5e-2)
    assert float(arena[:, :128].abs().sum()) == 0 and float(arena[:, 128 + M * N:].abs().sum()) == 0


@pytest.mark.parametrize("B,ch,H,W", [(4, 3, 64, 32), (2, 3, 512, 512), (3, 1, 8, 128)])
def test_amp_normalize_fft_kernel_matches_numpy_oracle(B, ch, H, W):
    """K14: shared-memory radix-2 FFT amplitude normalisation vs the reference's NumPy complex128 maths."""
    from fedml_b200 import ops
    from fedml_b200.simulation.sp.hs_fft import amplitude_normalize, amplitude_normalize_numpy
    torch.manual_seed(0)
    x = torch.rand(B, ch, H, W, device=DEV)
    run = torch.rand(ch, H, W, device=DEV) * 50
    launches = ops.native.LAUNCHES[0] if hasattr(ops, "native") else None
    y0, a0, b0 = amplitude_normalize(x)                       # first batch: running amplitude = batch mean
    r0, ra0, rb0 = amplitude_normalize_numpy(x.cpu().numpy())
    y1, a1, b1 = amplitude_normalize(x, run, momentum=0.1)
    r1, ra1, rb1 = amplitude_normalize_numpy(x.cpu().numpy(), run.cpu().numpy().astype(np.float64), 0.1)
    scale = float(np.abs(rb0).max())
    for got, ref in ((b0, rb0), (a0, ra0), (a1, ra1), (b1, rb1)):
        np.testing.assert_allclose(got.cpu().numpy(), ref, rtol=2e-4, atol=2e-5 * scale)
    np.testing.assert_allclose(y0.cpu().numpy(), r0, rtol=1e-3, atol=2e-4)
    np.testing.assert_allclose(y1.cpu().numpy(), r1, rtol=1e-3, atol=2e-4 * max(1.0, float(np.abs(r1).max())))
    from fedml_b200.ops import native
    assert launches is None or native.LAUNCHES[0] > launches      # the hand-written kernel ran, not torch.fft


@pytest.mark.parametrize("B,H,cin,cout,stride", [(10, 32, 16, 16, 1), (4, 16, 32, 32, 1), (10, 32, 32, 32, 2), (6, 8, 64, 64, 1),
                                                  (4, 8, 128, 128, 1), (8, 16, 64, 128, 2), (2, 4, 256, 64, 1)])
def test_implicit_conv3x3_forward_and_dgrad(B, H, cin, cout, stride):
    """tcgen05 implicit-GEMM 3x3 convolution (TMA-gathered taps, no im2col) vs F.conv2d, forward and input gradient."""
    import torch.nn.functional as F
    from fedml_b200 import ops
    torch.manual_seed(0)
    C = 3
    x = torch.randn(C, B, H, H, cin, device=DEV).to(torch.bfloat16)
    w = (torch.randn(C, cout, cin, 3, 3, device=DEV) * 0.1).to(torch.bfloat16)
    wm = w.permute(0, 1, 3, 4, 2).reshape(C, cout, 9 * cin).contiguous()
    assert ops.conv3x3_supported(x, cout, stride)
    y = ops.conv3x3(x, wm, stride)
    for c in range(C):
        ref = F.conv2d(x[c].float().permute(0, 3, 1, 2), w[c].float(), stride=stride, padding=1).permute(0, 2, 3, 1)
        torch.testing.assert_close(y[c].float(), ref, rtol=2e-2, atol=2e-2 * float(ref.abs().max()))
    if stride == 1:
        dy = torch.randn(C, B, H, H, cout, device=DEV).to(torch.bfloat16)
        wt = wm.transpose(1, 2).contiguous()                        # [C, 9*cin, cout]
        dx = ops.conv3x3_dgrad(dy, wt)
        for c in range(C):
            ref = F.conv_transpose2d(dy[c].float().permute(0, 3, 1, 2), w[c].float(), stride=1, padding=1).permute(0, 2, 3, 1)
            torch.testing.assert_close(dx[c].float(), ref, rtol=2e-2, atol=2e-2 * float(ref.abs().max()))


@pytest.mark.parametrize("C,B,H,cin,cout,stride", [(3, 10, 32, 16, 16, 1), (40, 4, 16, 32, 32, 1), (3, 10, 32, 32, 32, 2),
                                                    (2, 6, 8, 64, 64, 1), (2, 4, 8, 128, 128, 1), (3, 8, 16, 64, 128, 2),
                                                    (2, 16, 4, 256, 64, 1)])
def test_implicit_conv3x3_wgrad(C, B, H, cin, cout, stride):
    """Weight gradient without im2col: dy MN-major x TMA-gathered input windows, split-K reduce-add, vs autograd."""
    import torch.nn.functional as F
    from fedml_b200 import ops
    torch.manual_seed(0)
    x = torch.randn(C, B, H, H, cin, device=DEV).to(torch.bfloat16)
    Ho = (H - 1) // stride + 1
    dy = torch.randn(C, B, Ho, Ho, cout, device=DEV).to(torch.bfloat16)
    assert ops.conv3x3_wgrad_supported(x, dy, stride)
    arena = torch.zeros(C, cout * 9 * cin + 64, device=DEV)
    out = arena[:, 32:32 + cout * 9 * cin].view(C, cout, 9 * cin)
    dw = ops.conv3x3_wgrad(x, dy, stride, out=out, out_is_zeroed=True)
    for c in range(min(C, 3)):
        w = torch.zeros(cout, cin, 3, 3, device=DEV, requires_grad=True)
        y = F.conv2d(x[c].float().permute(0, 3, 1, 2), w, stride=stride, padding=1)
        y.backward(dy[c].float().permute(0, 3, 1, 2))
        ref = w.grad.permute(0, 2, 3, 1).reshape(cout, 9 * cin)
        torch.testing.assert_close(dw[c], ref, rtol=2e-2, atol=2e-2 * float(ref.abs().max()))
    assert float(arena[:, :32].abs().sum()) == 0 and float(arena[:, 32 + cout * 9 * cin:].abs().sum()) == 0


@pytest.mark.parametrize("block", ["bottleneck", "basic"])
def test_fused_sgd_restage_matches_separate_optimizer_step(block):
    """train_step ending in fb_sgd_restage (optimizer + bf16 re-staging in one launch, k x k gradients read from the
    GEMM-layout scratch) == train_step + un-staging + fused SGD kernel + staging, over three steps."""
    from fedml_b200 import ops
    from fedml_b200.models.batched import NativeBackend, make_batched
    from fedml_b200.models.cv import BasicBlock, Bottleneck, ResNet
    from fedml_b200.parallel.arena import ArenaLayout
    torch.manual_seed(0)
    C, B = 3, 8
    blk = Bottleneck if block == "bottleneck" else BasicBlock
    model = ResNet(blk, [1, 1, 1], num_classes=10)
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
        ex = make_batched(model, NativeBackend(torch.bfloat16))
        if fused:
            ex.set_fused_sgd(lr)
            ex.restage(W)
        for _ in range(3):
            ex.train_step(x, y, W, G, active)
            if fused:
                assert ex.fused_applied
            else:
                ops.sgd_step_stacked(st, gst, layout.P_train, lr)
        results.append(st.clone())
    a, b = results
    assert torch.equal(a[1, :layout.P_train], base[1, :layout.P_train]) and torch.equal(b[1, :layout.P_train], base[1, :layout.P_train])
    torch.testing.assert_close(b, a, rtol=1e-4, atol=1e-5)
    assert not torch.equal(a[0], base[0])
