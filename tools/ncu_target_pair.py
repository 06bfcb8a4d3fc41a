"""ncu target: the CTA-pair grouped GEMM on a large square problem and on the FFN shapes of the DistilBERT step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedml_b200.ops import transformer as tf
DEV = "cuda:0"
bf = lambda *s: (torch.randn(*s, device=DEV) * 0.5).to(torch.bfloat16)
a, b = bf(1, 4096, 4096), bf(1, 4096, 4096)
x, W1, xF = bf(8, 1024, 768), bf(8, 3072, 768), bf(8, 1024, 3072)
for _ in range(3):
    tf.mm("nt", a, b)
    tf.mm("nt", x, W1)
    tf.mm("nn", xF, W1)
    tf.mm("tn", xF, x, out_dtype=torch.float32)
torch.cuda.synchronize()
print("ok")
