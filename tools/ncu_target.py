"""Short single-GPU program exercising the hot kernels at bench shapes (target for `ncu --set full`)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedml_b200.native.build import build
build()
from fedml_b200 import ops
from fedml_b200.models import resnet56
from fedml_b200.parallel.arena import ArenaLayout
from fedml_b200.parallel.round_engine import RoundEngine
dev = torch.device("cuda:0")
C = 100
torch.manual_seed(0)
# grouped GEMM: layer-1 1x1 conv forward (M=B*H*W=10240, N=16, K=64) and a 3x3 weight gradient
a = torch.randn(C, 10240, 64, device=dev).to(torch.bfloat16); b = torch.randn(C, 16, 64, device=dev).to(torch.bfloat16)
dy = torch.randn(C, 10240, 16, device=dev).to(torch.bfloat16); cols = torch.randn(C, 10240, 144, device=dev).to(torch.bfloat16)
x = torch.randn(C, 10, 32, 32, 64, device=dev).to(torch.bfloat16)
arena = torch.randn(C, 4096, device=dev); gamma, beta = arena[:, :64], arena[:, 64:128]
rm, rv, nbt, act = torch.zeros(C, 64, device=dev), torch.ones(C, 64, device=dev), torch.zeros(C, device=dev), torch.ones(C, device=dev)
layout = ArenaLayout(resnet56(10)); eng = RoundEngine(layout, dev, max_clients=C)
stack = torch.randn(C, layout.P, device=dev); w = torch.rand(C, device=dev) + 1
for it in range(3):
    d = ops.grouped_gemm(a, b)
    dw = ops.grouped_gemm_tn(dy, cols)
    y, mean, rstd = ops.bn_fwd_stacked(x, None, gamma, beta, rm, rv, nbt, act, 0.1, 1e-5, True, True)
    dx, _, dg, db = ops.bn_bwd_stacked(x, x, None, gamma, mean, rstd, True, False, beta=beta)
    cols3 = ops.im2col3x3(x[..., :16].contiguous(), 1)
    eng.aggregate(stack, w, 1.0 / float(w.sum()))
    med = ops.coordinate_median(stack[:64, :4 * 65536])
torch.cuda.synchronize()
print("ok")
