"""Robust aggregation at the BASELINE size on one GPU: K rows x P coordinates (ResNet-18: 11.2 M), median / trimmed mean."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedml_b200.native.build import build
build()
from fedml_b200 import ops
K, P = int(os.environ.get("K", "512")), int(os.environ.get("P", str(11173960 // 4 * 4)))
x = torch.randn(K, P, device="cuda:0")
out = torch.empty(P, device="cuda:0")
for mode in ("median", "trimmed_mean"):
    for path in ("runs", "bitwise") if K > 128 else ("plain",):
        if path == "bitwise":
            fn = lambda: ops.robust_select(rows, K, 0, P, mode, K // 10 if mode != "median" else 0, outs, 1)   # noqa: E731
            rows = ops.ptr_table([x.data_ptr() + k * P * 4 for k in range(K)], x.device)
            outs = ops.ptr_table([out.data_ptr()], x.device)
        else:
            fn = (lambda: ops.coordinate_median(x, out, inplace=True)) if mode == "median" else (lambda: ops.trimmed_mean(x, K // 10, out, inplace=True))
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(f"K={K} P={P} {mode:13s} {path:8s} {ms:8.2f} ms  {K * P * 4 / ms / 1e6:7.1f} GB/s of input", flush=True)
