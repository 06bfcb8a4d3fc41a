"""Cheetah (fused reduce-scatter + SGD + all-gather over peer memory, bucketed and overlapped with backward) vs
DistributedDataParallel + torch.optim.SGD, same eager model, same batch.  torchrun --nproc-per-node N tools/bench_cheetah.py
Env: MODEL=resnet18|distilbert, B=per-GPU batch, STEPS, BUCKETS.  Device-timed, max over ranks."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import torch.nn as nn  # noqa: E402

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
dev = torch.device(f"cuda:{local}")
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
from fedml_b200.native.build import build  # noqa: E402
if rank == 0:
    build()
dist.barrier()
from fedml_b200.parallel.cheetah import CheetahEngine  # noqa: E402

MODEL, B, STEPS = os.environ.get("MODEL", "resnet18"), int(os.environ.get("B", "64")), int(os.environ.get("STEPS", "20"))
BUCKETS = int(os.environ.get("BUCKETS", "8"))


def make():
    torch.manual_seed(0)
    if MODEL == "distilbert":
        from fedml_b200.models.nlp import DistilBertForSequenceClassification
        m = DistilBertForSequenceClassification(num_labels=4)
        x = torch.randint(1, 30000, (B, 128), device=dev)
    else:
        from fedml_b200.models import resnet18
        m = resnet18(num_classes=10)
        x = torch.randn(B, 3, 32, 32, device=dev)
    y = torch.randint(0, 4 if MODEL == "distilbert" else 10, (B,), device=dev)
    return m, x, y


def timed(fn):
    for _ in range(5):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(STEPS):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / STEPS], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


res = {"model": MODEL, "n_gpus": world, "per_gpu_batch": B}
crit = nn.CrossEntropyLoss()
# no communication at all: forward + backward + local SGD (lower bound for any data-parallel step)
m, x, y = make(); m = m.to(dev); opt = torch.optim.SGD(m.parameters(), lr=0.05, momentum=0.9)
def local_step():
    opt.zero_grad(set_to_none=False); crit(m(x), y).backward(); opt.step()
res["local_ms"] = timed(local_step)
res["params_m"] = sum(p.numel() for p in m.parameters()) / 1e6
del m, opt
# DDP + torch.optim.SGD
m, x, y = make(); ddp = nn.parallel.DistributedDataParallel(m.to(dev), device_ids=[local]); opt = torch.optim.SGD(ddp.parameters(), lr=0.05, momentum=0.9)
def ddp_step():
    opt.zero_grad(set_to_none=False); crit(ddp(x), y).backward(); opt.step()
res["ddp_ms"] = timed(ddp_step)
del ddp, opt, m
for name, kw in (("cheetah_ms", dict(n_buckets=1, overlap=False)), ("cheetah_overlap_ms", dict(n_buckets=BUCKETS, overlap=True))):
    m, x, y = make()
    eng = CheetahEngine(m, dev, lr=0.05, momentum=0.9, **kw)
    def ch_step():
        eng.zero_grad(); crit(eng.model(x), y).backward(); eng.step()
    res[name] = timed(ch_step)
    res[name.replace("_ms", "_buckets")] = len(eng.buckets)
    del eng, m
for k in ("ddp_ms", "cheetah_ms", "cheetah_overlap_ms"):
    res[k.replace("_ms", "_exposed_comm_ms")] = round(res[k] - res["local_ms"], 3)
if rank == 0:
    print(json.dumps(res), flush=True)
dist.destroy_process_group()
