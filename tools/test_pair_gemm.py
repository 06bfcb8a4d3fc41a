"""Correctness + timing of the CTA-pair (cta_group::2) grouped GEMM against torch (run with FB_GEMM_PAIR=1 / 0)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedml_b200.ops import transformer as tf
DEV = "cuda:0"
torch.manual_seed(0)
bf = lambda *s: (torch.randn(*s, device=DEV) * 0.5).to(torch.bfloat16)
def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / n * 1e3
cases = [("nt", 8, 1024, 2304, 768), ("nt", 8, 1024, 768, 3072), ("nn", 8, 1024, 3072, 768), ("tn", 8, 768, 3072, 1024),
         ("nt", 3, 640, 384, 200), ("nn", 2, 520, 128, 136), ("tn", 2, 512, 256, 320), ("nt", 1, 4096, 4096, 4096)]
for mode, G, M, N, K in cases:
    if mode == "nt": a, b = bf(G, M, K), bf(G, N, K); ref = a.float() @ b.float().transpose(1, 2)
    elif mode == "nn": a, b = bf(G, M, K), bf(G, K, N); ref = a.float() @ b.float()
    else: a, b = bf(G, K, M), bf(G, K, N); ref = a.float().transpose(1, 2) @ b.float()
    out = tf.mm(mode, a, b, out_dtype=torch.float32)
    torch.cuda.synchronize()
    err = float((out - ref).abs().max()) / float(ref.abs().max())
    us = timeit(lambda: tf.mm(mode, a, b))
    print(f"{mode} G={G} M={M} N={N} K={K}: rel err {err:.2e}  {us:8.1f} us  {2*G*M*N*K/us/1e6:7.1f} TFLOP/s", flush=True)
