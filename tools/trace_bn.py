"""Per-CTA timeline of the slab-resident BatchNorm forward kernel (csrc/bn_resident.cuh, fb_bn_set_trace): where a CTA's time
goes - TMA load, statistics, client barrier, partial sums + coefficients, apply, store drain - and how CTAs overlap."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from fedml_b200.native.build import build
build()
from fedml_b200 import ops
from fedml_b200.ops.native import lib

dev = torch.device("cuda:0")
Cn, B, H, ch = int(os.environ.get("C", 100)), 10, int(os.environ.get("H", 32)), int(os.environ.get("CH", 64))
dtype = torch.float32 if os.environ.get("DTYPE", "fp32") == "fp32" else torch.bfloat16
with_res = bool(int(os.environ.get("RES", "0")))
x = torch.randn(Cn, B, H, H, ch, device=dev).to(dtype)
res = torch.randn_like(x) if with_res else None
arena = torch.randn(Cn, 2 * ch + 8, device=dev)
gamma, beta = arena[:, :ch], arena[:, ch:2 * ch]
rm, rv = torch.zeros(Cn, ch, device=dev), torch.ones(Cn, ch, device=dev)
nbt, act = torch.zeros(Cn, device=dev), torch.ones(Cn, device=dev)
run = lambda: ops.bn_fwd_stacked(x, res, gamma, beta, rm, rv, nbt, act, 0.1, 1e-5, True, True)      # noqa: E731
for _ in range(3):
    run()
torch.cuda.synchronize()
trace = torch.zeros(Cn * 128 * 8, dtype=torch.int64, device=dev)
assert lib().fb_bn_set_trace(C.c_void_p(trace.data_ptr())) == 0
flush = torch.zeros(64 << 20, device=dev); flush.add_(1.0)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); run(); e1.record(); torch.cuda.synchronize()
assert lib().fb_bn_set_trace(C.c_void_p(0)) == 0
t = trace.cpu().numpy().reshape(-1, 8)
t = t[t[:, 0] > 0]
S = len(t) // Cn
t0 = t[:, 0].min()
names = ["wait for TMA load", "statistics + reduce", "client barrier", "partial sums + coefficients", "apply (smem)", "store drain"]
print(f"{len(t)} CTAs, S={S}, kernel {e0.elapsed_time(e1) * 1e3:.1f} us (events), span {(t[:, 6].max() - t0) / 1e3:.1f} us, "
      f"bytes in+out {x.numel() * x.element_size() * (2 + with_res) / 1e6:.0f} MB")
for i, n in enumerate(names):
    d = (t[:, i + 1] - t[:, i]) / 1e3
    print(f"  {n:30s} mean {d.mean():6.2f} us  p50 {np.median(d):6.2f}  p90 {np.percentile(d, 90):6.2f}  max {d.max():6.2f}")
life = (t[:, 6] - t[:, 0]) / 1e3
print(f"  CTA lifetime mean {life.mean():.2f} us; start times: first wave {np.sum(t[:, 0] - t0 < 2000)} CTAs within 2 us; "
      f"distinct SMs {len(np.unique(t[:, 7]))}")
# concurrency: CTAs alive over time, and CTAs in a data phase (loading or storing)
grid = np.linspace(t0, t[:, 6].max(), 60)
alive = [(np.sum((t[:, 0] <= g) & (t[:, 6] > g))) for g in grid]
moving = [(np.sum(((t[:, 0] <= g) & (t[:, 1] > g)) | ((t[:, 5] <= g) & (t[:, 6] > g)))) for g in grid]
print("  alive  :", " ".join(f"{a:3d}" for a in alive[::3]))
print("  moving :", " ".join(f"{a:3d}" for a in moving[::3]))
