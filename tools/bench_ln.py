"""Micro-benchmark of the fused add + LayerNorm forward / backward kernels at DistilBERT shapes against torch.nn.functional.layer_norm."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedml_b200.ops import transformer as tf
DEV = "cuda:0"
C, R, D = int(os.environ.get("K", "8")), 1024, 768
bf = lambda *s: torch.randn(*s, device=DEV, dtype=torch.bfloat16)
dy, dy2, z = bf(C, R, D), bf(C, R, D), bf(C, R, D)
mean, rstd = torch.randn(C, R, device=DEV), torch.rand(C, R, device=DEV) + 0.5
gamma = torch.randn(C, D, device=DEV); dg = torch.zeros(C, D, device=DEV); db = torch.zeros(C, D, device=DEV)
def t(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / n * 1e3
us = t(lambda: tf.ln_bwd(dy, dy2, z, mean, rstd, gamma, dg, db))
mb = 4 * C * R * D * 2 / 1e6
print(f"ln_bwd rpw={os.environ.get('FB_LN_ROWS_PER_WARP')} cps={os.environ.get('FB_LN_CTAS_PER_SM')}: {us:.1f} us  {mb/us*1e-3*1e3:.0f} GB/s")
us = t(lambda: tf.add_ln_fwd(dy, dy2, gamma, gamma, 1e-12))
print(f"ln_fwd: {us:.1f} us {4*C*R*D*2/1e6/us*1e3:.0f} GB/s")
out = torch.zeros(C, D, device=DEV)
us = t(lambda: tf.colsum(dy, out))
print(f"colsum: {us:.1f} us {C*R*D*2/1e6/us*1e3:.0f} GB/s")
