"""BatchNorm kernels: numerics against a plain fp32 PyTorch restatement, and device time per call.

Run once per setting (the switches are read once per process):
    FB_BN_RESIDENT=0 python tools/bench_bn.py            # fused / two-pass kernels
    FB_BN_RESIDENT=1 python tools/bench_bn.py            # slab-resident kernels (default)
    FB_BN_FUSED=1 python tools/bench_bn.py --small       # small ragged shapes through the single-launch kernels
Prints one JSON line per case: max relative errors and microseconds (CUDA events, L2 flushed between timed calls)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def reference(x, res, gamma, beta, eps, relu, dy, dy2, use_y):
    xf = x.float()
    C, ch = x.shape[0], x.shape[-1]
    flat = xf.reshape(C, -1, ch)
    mean = flat.mean(1)
    var = flat.var(1, unbiased=False)
    rstd = (var + eps).rsqrt()
    sh = (C,) + (1,) * (x.dim() - 2) + (ch,)
    xhat = (xf - mean.view(sh)) * rstd.view(sh)
    pre = xhat * gamma.view(sh) + beta.view(sh) + (res.float() if res is not None else 0)
    y = pre.relu() if relu else pre
    d = dy.float() + (dy2.float() if dy2 is not None else 0)
    if relu:
        d = d * (pre > 0)
    dflat, xh = d.reshape(C, -1, ch), xhat.reshape(C, -1, ch)
    dbeta, dgamma = dflat.sum(1), (dflat * xh).sum(1)
    M = dflat.shape[1]
    dx = (gamma * rstd).view(sh) * (d - dbeta.view(sh) / M - xhat * dgamma.view(sh) / M)
    return y, mean, rstd, dx, d, dgamma, dbeta


def rel(a, b):
    return float((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-6))


def time_us(fn, flush, iters=10):
    fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.add_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters * 1e3


def run_case(ops, dtype, C, B, H, W, ch, with_res, relu, with_dy2, want_dres, keep_y, flush, timed=True):
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(C * 131 + ch + H)
    mk = lambda: torch.randn(C, B, H, W, ch, device=dev, generator=g).to(dtype)      # noqa: E731
    x, dy = mk(), mk()
    res = mk() if with_res else None
    dy2 = mk() if with_dy2 else None
    arena = torch.randn(C, 2 * ch + 8, device=dev, generator=g)
    gamma, beta = arena[:, :ch], arena[:, ch:2 * ch]
    rm, rv = torch.zeros(C, ch, device=dev), torch.ones(C, ch, device=dev)
    nbt, act = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    y, mean, rstd = ops.bn_fwd_stacked(x, res, gamma, beta, rm, rv, nbt, act, 0.1, 1e-5, relu, True)
    use_y = keep_y or with_res
    dx, dres, dgm, dbt = ops.bn_bwd_stacked(dy, x, y if use_y else None, gamma, mean, rstd, relu, want_dres, beta=beta, dy2=dy2)
    ry, rmean, rrstd, rdx, rd, rdg, rdb = reference(x, res, gamma, beta, 1e-5, relu, dy, dy2, use_y)
    out = {"dtype": str(dtype).split(".")[-1], "shape": [C, B, H, W, ch], "res": with_res, "relu": relu, "dy2": with_dy2,
           "dres": want_dres, "y_mask": use_y,
           "err": {"y": rel(y, ry), "mean": rel(mean, rmean), "rstd": rel(rstd, rrstd), "dx": rel(dx, rdx), "dgamma": rel(dgm, rdg),
                   "dbeta": rel(dbt, rdb), "dres": rel(dres, rd) if want_dres else None,
                   "running_mean": rel(rm, 0.1 * rmean)}}
    if timed:
        out["fwd_us"] = round(time_us(lambda: ops.bn_fwd_stacked(x, res, gamma, beta, rm, rv, nbt, act, 0.1, 1e-5, relu, True), flush), 1)
        out["bwd_us"] = round(time_us(lambda: ops.bn_bwd_stacked(dy, x, y if use_y else None, gamma, mean, rstd, relu, want_dres,
                                                                  beta=beta, dy2=dy2), flush), 1)
        nb = x.numel() * x.element_size()
        out["fwd_min_bytes_mb"] = round(nb * (2 + with_res) / 1e6, 1)
        out["bwd_min_bytes_mb"] = round(nb * (3 + with_dy2 + want_dres + (use_y and relu)) / 1e6, 1)
        out["fwd_tbs"] = round(out["fwd_min_bytes_mb"] / out["fwd_us"], 3)
        out["bwd_tbs"] = round(out["bwd_min_bytes_mb"] / out["bwd_us"], 3)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--small", action="store_true", help="small ragged shapes only (numerics)")
    ap.add_argument("--clients", type=int, default=100)
    args = ap.parse_args()
    from fedml_b200.native.build import build
    build()
    from fedml_b200 import ops
    flush = torch.zeros(64 << 20, device="cuda:0")        # 256 MB > L2
    env = {k: os.environ.get(k) for k in ("FB_BN_RESIDENT", "FB_BN_PERSIST", "FB_BN_PERSIST_KB", "FB_BN_FUSED", "FB_BN_RES_KB_FWD", "FB_BN_RES_KB_BWD")}
    print(json.dumps({"env": env}))
    worst = 0.0
    if args.small:
        cases = []
        for dtype in (torch.float32, torch.bfloat16):
            cases += [(dtype, 3, 5, 7, 9, 16, True, True, True, True, True), (dtype, 2, 3, 13, 11, 64, False, True, False, False, False),
                      (dtype, 4, 10, 32, 32, 16, False, False, True, True, False), (dtype, 1, 10, 32, 32, 64, True, True, False, True, True),
                      (dtype, 5, 2, 8, 8, 256, False, True, False, False, True), (dtype, 2, 10, 16, 16, 128, True, True, True, False, True),
                      # more work items than SMs: the persistent kernels pipeline several items per CTA
                      (dtype, 40, 2, 8, 8, 256, True, True, True, True, True), (dtype, 37, 3, 9, 8, 64, False, True, False, False, False)]
        for cs in cases:
            r = run_case(ops, *cs, flush=flush, timed=False)
            print(json.dumps(r))
            tol = 2e-2 if cs[0] == torch.bfloat16 else 2e-4
            bad = [k for k, v in r["err"].items() if v is not None and not v < tol]
            assert not bad, (bad, r)
        print("BN_SMALL_OK")
        return
    C = args.clients
    for dtype in (torch.float32, torch.bfloat16):
        # the bottleneck ResNet-56 layers of the headline step: 64 ch @ 32x32 (block outputs, residual), 16 ch @ 32x32 (inner),
        # 128 @ 16x16, 256 @ 8x8
        for (H, ch, with_res, dy2, dres) in ((32, 64, True, False, True), (32, 64, False, True, False), (32, 16, False, False, False),
                                             (16, 128, True, False, True), (8, 256, True, False, True)):
            r = run_case(ops, dtype, C, 10, H, H, ch, with_res, True, dy2, dres, False, flush)
            print(json.dumps(r))
            worst = max(worst, max(v for v in r["err"].values() if v is not None))
    print("BN_BENCH_DONE worst_err", worst)


if __name__ == "__main__":
    main()
