"""Micro-benchmark: tcgen05 grouped GEMM vs cuBLAS bmm on the client-batched ResNet-56 shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedml_b200.native.build import build
build()
from fedml_b200 import ops
dev = "cuda:0"
C = int(os.environ.get("C", "100"))
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timeit(fn, n=5):
    for _ in range(2): fn()
    ts = []
    for _ in range(n):
        flush.zero_(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts) * 1e3

print(f"{'kind':5} {'M':>6} {'N':>4} {'K':>5} | {'tc us':>8} {'GB/s':>7} | {'bmm us':>8} {'GB/s':>7}")
for (M, N, K) in [(10240, 16, 64), (10240, 16, 144), (10240, 64, 16), (10240, 144, 16), (2560, 32, 128), (2560, 32, 288),
                  (2560, 128, 32), (640, 64, 256), (640, 64, 576), (640, 256, 64), (10240, 16, 72)]:
    a = torch.randn(C, M, K, device=dev).to(torch.bfloat16); b = torch.randn(C, N, K, device=dev).to(torch.bfloat16)
    by = C * (M * K + N * K + M * N) * 2
    t1 = timeit(lambda: ops.grouped_gemm(a, b)); t2 = timeit(lambda: torch.bmm(a, b.transpose(1, 2)))
    print(f"{'NT':5} {M:6d} {N:4d} {K:5d} | {t1:8.1f} {by/t1/1e3:7.0f} | {t2:8.1f} {by/t2/1e3:7.0f}")
for (Kr, M, N) in [(10240, 16, 64), (10240, 16, 144), (10240, 64, 16), (2560, 32, 288), (2560, 128, 32), (640, 64, 576), (640, 256, 64)]:
    a = torch.randn(C, Kr, M, device=dev).to(torch.bfloat16); b = torch.randn(C, Kr, N, device=dev).to(torch.bfloat16)
    by = C * (Kr * M + Kr * N) * 2 + C * M * N * 4
    t1 = timeit(lambda: ops.grouped_gemm_tn(a, b)); t2 = timeit(lambda: torch.bmm(a.transpose(1, 2), b))
    print(f"{'TN':5} {Kr:6d} {M:4d} {N:5d} | {t1:8.1f} {by/t1/1e3:7.0f} | {t2:8.1f} {by/t2/1e3:7.0f}")
