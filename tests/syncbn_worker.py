"""gloo worker for the SyncBN test: each rank gets half of a batch; output, input gradient, parameter gradients and
running statistics must equal plain BatchNorm on the whole batch."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import torch.nn as nn  # noqa: E402

from fedml_b200.models.sync_batchnorm import convert_model  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
torch.manual_seed(0)
ref = nn.Sequential(nn.Conv2d(3, 6, 3, padding=1), nn.BatchNorm2d(6), nn.ReLU(), nn.Conv2d(6, 4, 1), nn.BatchNorm2d(4)).double()
import copy
net = convert_model(copy.deepcopy(ref))
x = torch.randn(8, 3, 5, 5, dtype=torch.float64)
t = torch.randn(8, 4, 5, 5, dtype=torch.float64)
# reference: whole batch on one process
xr = x.clone().requires_grad_(True)
((ref(xr) - t) ** 2).sum().backward()
# replicas: half a batch each; parameter gradients are summed over ranks (what DDP's all-reduce would do, unscaled)
lo, hi = rank * 4, rank * 4 + 4
xs = x[lo:hi].clone().requires_grad_(True)
((net(xs) - t[lo:hi]) ** 2).sum().backward()
ok = torch.allclose(xs.grad, xr.grad[lo:hi], rtol=1e-9, atol=1e-10)
for (n1, p1), (n2, p2) in zip(net.named_parameters(), ref.named_parameters()):
    g = p1.grad.clone()
    dist.all_reduce(g)
    ok = ok and torch.allclose(g, p2.grad, rtol=1e-8, atol=1e-9)
for (n1, b1), (n2, b2) in zip(net.named_buffers(), ref.named_buffers()):
    ok = ok and torch.allclose(b1.double(), b2.double(), rtol=1e-9, atol=1e-10)
assert list(net.state_dict().keys()) == list(ref.state_dict().keys())
net.eval(); ref.eval()
ok = ok and torch.allclose(net(x[lo:hi]), ref(x[lo:hi]), rtol=1e-9, atol=1e-10)
print(f"SYNCBN_OK {rank}" if ok else f"SYNCBN_MISMATCH {rank}", flush=True)
dist.destroy_process_group()
