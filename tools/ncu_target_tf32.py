"""Short single-GPU program with the hot kernels of the tf32 headline step at bench shapes (target for `ncu --set full`)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedml_b200.native.build import build
build()
from fedml_b200 import ops
from fedml_b200.ops import transformer as tf
dev = torch.device("cuda:0")
C = 100
torch.manual_seed(0)
a = torch.randn(C, 10240, 64, device=dev); b = torch.randn(C, 16, 64, device=dev)            # layer-1 1x1 conv forward (NT)
dy = torch.randn(C, 10240, 16, device=dev); wfull = torch.randn(C, 16, 64, device=dev)          # its data gradient (NN)
cols = torch.randn(C, 10240, 144, device=dev)                                                  # 3x3 weight gradient via im2col (TN)
x = torch.randn(C, 10, 32, 32, 64, device=dev)
x16 = torch.randn(C, 10, 32, 32, 16, device=dev); wm16 = torch.randn(C, 16, 144, device=dev)
arena = torch.randn(C, 4096, device=dev); gamma, beta = arena[:, :64], arena[:, 64:128]
rm, rv, nbt, act = torch.zeros(C, 64, device=dev), torch.ones(C, 64, device=dev), torch.zeros(C, device=dev), torch.ones(C, device=dev)
for it in range(3):
    d = ops.grouped_gemm(a, b)
    dx = tf.mm("nn", dy, wfull)
    dw = ops.grouped_gemm_tn(dy, cols)
    yc = ops.conv3x3(x16, wm16, 1)
    y, mean, rstd = ops.bn_fwd_stacked(x, None, gamma, beta, rm, rv, nbt, act, 0.1, 1e-5, True, True)
    dxb, _, dg, db = ops.bn_bwd_stacked(x, x, None, gamma, mean, rstd, True, False, beta=beta)
torch.cuda.synchronize()
print("ok")
