"""Per-shape timing of the grouped GEMMs of one client-batched DistilBERT step (C clients) vs cuBLAS bmm."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from fedml_b200.native.build import build  # noqa: E402

build()
from fedml_b200.ops import transformer as tf  # noqa: E402

DEV = "cuda:0"
C, B, S, D, F, H = int(os.environ.get("K", "8")), 8, 128, 768, 3072, 12
dh, R, CB = D // H, B * S, C * B


def bf(*s):
    return torch.randn(*s, device=DEV, dtype=torch.bfloat16)


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3   # us


qkv = bf(CB, S, 3 * D)
heads = lambda t, j: t.view(CB, S, t.shape[-1] // D, H, dh)[:, :, j].permute(0, 2, 1, 3)   # noqa: E731
q, k, v = (heads(qkv, j) for j in range(3))
P = bf(CB, H, S, S)
ctx = bf(C, R, D)
x, xF = bf(C, R, D), bf(C, R, F)
Wqkv, Wo, W1, W2 = bf(C, 3 * D, D), bf(C, D, D), bf(C, F, D), bf(C, D, F)
gD, gF = torch.zeros(C, D, D, device=DEV), torch.zeros(C, F, D, device=DEV)
cases = [
    ("qkv   nt", "nt", x, Wqkv, None, None, 2 * C * R * 3 * D * D),
    ("score nt", "nt", q, k, None, torch.float32, 2 * CB * H * S * S * dh),
    ("PV    nn", "nn", P, v, heads(ctx, 0), None, 2 * CB * H * S * S * dh),
    ("out   nt", "nt", x, Wo, None, None, 2 * C * R * D * D),
    ("ffn1  nt", "nt", x, W1, None, None, 2 * C * R * D * F),
    ("ffn2  nt", "nt", xF, W2, None, None, 2 * C * R * D * F),
    ("dX f2 nn", "nn", x, W2, None, None, 2 * C * R * D * F),
    ("dX f1 nn", "nn", xF, W1, None, None, 2 * C * R * D * F),
    ("dW f1 tn", "tn", xF, x, gF, None, 2 * C * R * D * F),
    ("dW o  tn", "tn", x, x, gD, None, 2 * C * R * D * D),
    ("dV    tn", "tn", P, heads(ctx, 0), heads(torch.empty_like(qkv), 2), None, 2 * CB * H * S * S * dh),
    ("dQ    nn", "nn", P, k, heads(torch.empty_like(qkv), 0), None, 2 * CB * H * S * S * dh),
]
tot = 0.0
for name, mode, a, b, out, odt, flops in cases:
    us = timeit(lambda: tf.mm(mode, a, b, out=out, out_dtype=odt or torch.bfloat16))
    af, bfm = a.float(), b.float()
    if a.dim() == 3:
        a3, b3 = a, b
        ref = {"nt": lambda: torch.bmm(a3, b3.transpose(1, 2)), "nn": lambda: torch.bmm(a3, b3),
               "tn": lambda: torch.bmm(a3.transpose(1, 2), b3)}[mode]
        us_ref = timeit(ref)
    else:
        ref = {"nt": lambda: a @ b.transpose(-1, -2), "nn": lambda: a @ b, "tn": lambda: a.transpose(-1, -2) @ b}[mode]
        us_ref = timeit(ref)
    tot += us
    print(f"{name}: {us:8.1f} us  {flops / us / 1e6:7.1f} TFLOP/s   cuBLAS {us_ref:8.1f} us {flops / us_ref / 1e6:7.1f} TFLOP/s")
print(f"sum {tot:.0f} us")
