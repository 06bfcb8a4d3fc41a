"""Short single-GPU program exercising the K12 / K14 kernels at DistilBERT bench shapes (target for `ncu --set full`)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedml_b200.native.build import build
build()
from fedml_b200 import ops
from fedml_b200.ops import transformer as tf
DEV = "cuda:0"
C, B, S, D, F, H = 8, 8, 128, 768, 3072, 12
dh, R, CB = D // H, B * S, C * B
bf = lambda *s: torch.randn(*s, device=DEV, dtype=torch.bfloat16)
torch.manual_seed(0)
x, xF = bf(C, R, D), bf(C, R, F)
W1, W2 = bf(C, F, D), bf(C, D, F)
qkv = bf(CB, S, 3 * D)
heads = lambda t, j: t.view(CB, S, t.shape[-1] // D, H, dh)[:, :, j].permute(0, 2, 1, 3)
q, k, v = (heads(qkv, j) for j in range(3))
gF = torch.zeros(C, F, D, device=DEV)
gamma, beta = torch.randn(C, D, device=DEV), torch.randn(C, D, device=DEV)
dg, db = torch.zeros(C, D, device=DEV), torch.zeros(C, D, device=DEV)
mask = torch.ones(CB, S, device=DEV)
ctx = bf(C, R, D)
img = torch.rand(8, 3, 512, 512, device=DEV)
for it in range(2):
    u = tf.mm("nt", x, W1)                                   # FFN-1 forward
    dx = tf.mm("nn", x, W2)                                  # FFN-2 data gradient (B operand MN-major)
    tf.mm("tn", xF, x, out=gF)                               # FFN-1 weight gradient into the arena
    sc = tf.mm("nt", q, k, out_dtype=torch.float32)          # Q K^T, two-level groups, heads in place
    P, Pd = tf.attn_softmax_fwd(sc, mask, H, 0.125, 0.1, 1, 17, None)
    tf.mm("nn", Pd, v, out=heads(ctx, 0))                    # P V, merged heads written in place
    dS = tf.attn_softmax_bwd(sc, P, 0.125, 0.1, 1, 17, None)
    y, z, mean, rstd = tf.add_ln_fwd(x, ctx, gamma, beta, 1e-12)
    dz = tf.ln_bwd(x, ctx, z, mean, rstd, gamma, dg, db)
    a = tf.act_fwd(xF, "gelu"); da = tf.act_bwd(xF, xF, "gelu")
    tf.colsum(xF, torch.zeros(C, F, device=DEV))
    yn, amp, bat = ops.amp_normalize(img)
torch.cuda.synchronize()
print("ok")
