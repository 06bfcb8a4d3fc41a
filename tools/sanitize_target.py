"""Small single-GPU program for compute-sanitizer (memcheck / racecheck / synccheck / initcheck): every kernel family once,
on small shapes (the tools slow kernels down 10-100x), including the emulated multi-rank round with DP clipping (flag
protocol + intra-rank grid barrier) and the bounded-spin time-out path."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from fedml_b200.native.build import build  # noqa: E402

build()
from fedml_b200 import ops  # noqa: E402
from fedml_b200.ops import transformer as tf  # noqa: E402
from fedml_b200.parallel.round_engine import DPConfig, RoundEngine, ServerOptConfig, emulated_round  # noqa: E402

dev = "cuda:0"
torch.manual_seed(0)
# round kernel: 3 emulated ranks, DP clip + noise, one-shot and two-shot, FedAdam
P, Pt = 4 * 1031, 4 * 900
g = torch.randn(P, device=dev)
for two_shot in (False, True):
    clients = [g + 0.1 * torch.randn(c, P, device=dev) if c else None for c in (3, 0, 2)]
    weights = [torch.rand(c, device=dev) + 1 if c else None for c in (3, 0, 2)]
    tot = sum(float(w.sum()) for w in weights if w is not None)
    emulated_round(clients, weights, g, Pt, 1.0 / tot, two_shot, opt=ServerOptConfig("adam", lr=0.01), dp=DPConfig(1.0, 0.01, 7))
RoundEngine.check_spin_errors()
# GEMM (bf16 + tf32, NT / NN / TN), implicit conv, BN, im2col, head, staging, robust, secagg, augment
for dt in (torch.bfloat16, torch.float32):
    a, b = torch.randn(2, 200, 72, device=dev).to(dt), torch.randn(2, 40, 72, device=dev).to(dt)
    tf.mm("nt", a, b); tf.mm("nn", a, torch.randn(2, 72, 40, device=dev).to(dt)); tf.mm("tn", torch.randn(2, 72, 200, device=dev).to(dt), torch.randn(2, 72, 40, device=dev).to(dt))
    x = torch.randn(2, 4, 8, 8, 32, device=dev).to(dt)
    wm = torch.randn(2, 32, 9 * 32, device=dev).to(dt)
    y = ops.conv3x3(x, wm, 1)
    ops.conv3x3_dgrad(y, wm.view(2, 32, 9, 32).permute(0, 2, 3, 1).reshape(2, 9 * 32, 32).contiguous())
    ops.conv3x3_wgrad(x, y, 1)
    arena = torch.randn(2, 256, device=dev)
    rm, rv, nbt, act = torch.zeros(2, 32, device=dev), torch.ones(2, 32, device=dev), torch.zeros(2, device=dev), torch.ones(2, device=dev)
    yb, mean, rstd = ops.bn_fwd_stacked(y, x, arena[:, :32], arena[:, 32:64], rm, rv, nbt, act, 0.1, 1e-5, True, True)
    ops.bn_bwd_stacked(torch.randn_like(y), y, yb, arena[:, :32], mean, rstd, True, True)
    cols = ops.im2col3x3(x, 2); ops.col2im3x3(ops.im2col3x3(x, 1), 4, 8, 8, 32, 1)
    garena = torch.zeros(2, 10 * 32 + 16, device=dev)
    ops.head_fwd_bwd(yb, arena[:, :0].new_zeros(2, 10, 32) + 0.1, arena[:, :10].contiguous(), torch.randint(0, 10, (2, 4), device=dev), None,
                     garena[:, :320].view(2, 10, 32), garena[:, 320:330])
st = torch.randn(200, 4096, device=dev)
ops.coordinate_median(st[:9]); ops.trimmed_mean(st[:33], 3); ops.coordinate_median(st)
q = ops.ff_quantize(torch.randn(1000, device=dev), 2 ** 16, 2 ** 31 - 1)
ops.ff_mask(q, [11, 22], [1, -1], 2 ** 31 - 1, nonce=3)
ops.augment_cifar(torch.randint(0, 256, (8, 32, 32, 3), dtype=torch.uint8, device=dev), 1, 2)
torch.cuda.synchronize()
print("sanitize target ok")
