"""BASELINE.json configs #3 and #4 end to end on one GPU (or under torchrun): ResNet-18, CIFAR-10-shaped data.

  CONFIG=fedopt_dp   FedOpt (FedAdam server optimizer) + DP clip / noise
  CONFIG=median | trimmed_mean | krum   Byzantine-robust FedAvg with 10 % malicious clients
Env: K = clients per round (default 64), ROUNDS (3).  Prints one JSON line (device-timed, max over ranks).
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import fedml_b200 as fedml  # noqa: E402
from fedml_b200.arguments import Arguments  # noqa: E402
from fedml_b200.native.build import build  # noqa: E402
from fedml_b200.ops import native  # noqa: E402

RANK, WORLD, LOCAL = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
if RANK == 0:
    build()
CONFIG = os.environ.get("CONFIG", "fedopt_dp")
K, ROUNDS = int(os.environ.get("K", "64")), int(os.environ.get("ROUNDS", "3"))
extra = {"fedopt_dp": dict(federated_optimizer="FedOpt", server_optimizer="adam", server_lr=0.01, dp_clip=1.0, dp_sigma=0.001),
         "median": dict(federated_optimizer="FedAvg", robust_agg="median", byzantine_fraction=0.1),
         "trimmed_mean": dict(federated_optimizer="FedAvg", robust_agg="trimmed_mean", byzantine_fraction=0.1),
         "krum": dict(federated_optimizer="FedAvg", robust_agg="krum", byzantine_fraction=0.1, krum_m=max(1, K // 2))}[CONFIG]
total = max(2 * K, 64)
args = Arguments.from_dict(dict(
    training_type="simulation", backend="NCCL", dataset="cifar10", model="resnet18", client_num_in_total=total,
    client_num_per_round=K, comm_round=ROUNDS + 2, epochs=1, batch_size=10, learning_rate=0.03, partition_method="homo",
    using_gpu=True, frequency_of_the_test=0, log_level="WARNING", synthetic_train_size=total * 50, **extra))
args = fedml.init(args)
dev = torch.device(f"cuda:{LOCAL}")
torch.cuda.set_device(dev)
if WORLD > 1:
    dist.init_process_group("nccl", device_id=dev)
    dist.barrier()
ds, ncls = fedml.data.load(args)
args._federated_dataset.pin()
model = fedml.models.create(args, ncls)
from fedml_b200.simulation import SimulatorNCCL  # noqa: E402

sim = SimulatorNCCL(args, dev, ds, model).simulator
for r in range(2):
    sim.train_one_round(r)
torch.cuda.synchronize()
if WORLD > 1:
    dist.barrier()
native.LAUNCHES[0] = 0
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for r in range(2, 2 + ROUNDS):
    sim.train_one_round(r)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / ROUNDS
if WORLD > 1:
    t = torch.tensor([ms], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
be = getattr(sim.batched, "be", None) if sim.batched is not None else None
if RANK == 0:
    print(json.dumps({"config": CONFIG, "model": "resnet18 (11.2M params)", "clients_per_round": K, "n_gpus": WORLD,
                      "ms_per_round": round(ms, 2), "client_updates_per_s": round(K * 1e3 / ms, 1),
                      "phase_ms": {k: round(v / (ROUNDS + 2), 2) for k, v in sim.timer.collect().items()},
                      "my_launches_per_round": native.LAUNCHES[0] // ROUNDS, "library_fallbacks": getattr(be, "fallbacks", None),
                      "mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 1)}))
if WORLD > 1:
    dist.destroy_process_group()
