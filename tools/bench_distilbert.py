"""FedNLP config (BASELINE.json #5): DistilBERT-base text classification, C clients trained together on one GPU.

  python tools/bench_distilbert.py                  # product: client-batched native executor (bf16)
  MODE=eager python tools/bench_distilbert.py       # reference behaviour: one client after another, PyTorch eager

Env: K = clients per round over all GPUs (default 8; launch with torchrun for several GPUs), B = batch size (8), S = sequence length (128),
N = samples per client (32), ROUNDS = timed rounds (3).  Prints one JSON line.
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import fedml_b200 as fedml  # noqa: E402
from fedml_b200.arguments import Arguments  # noqa: E402
from fedml_b200.native.build import build  # noqa: E402
from fedml_b200.ops import native  # noqa: E402

import torch.distributed as dist  # noqa: E402

RANK, WORLD, LOCAL = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
if RANK == 0:
    build()
K, B, S = int(os.environ.get("K", "8")), int(os.environ.get("B", "8")), int(os.environ.get("S", "128"))
N, ROUNDS = int(os.environ.get("N", "32")), int(os.environ.get("ROUNDS", "3"))
MODE = os.environ.get("MODE", "batched")
total = max(4 * K, 32)
args = Arguments.from_dict(dict(
    training_type="simulation", backend="NCCL", federated_optimizer="FedAvg", dataset="20news", model="distilbert",
    client_num_in_total=total, client_num_per_round=K, comm_round=ROUNDS + 2, epochs=1, batch_size=B, learning_rate=0.01,
    partition_method="homo", using_gpu=True, frequency_of_the_test=0, log_level="WARNING", max_seq_length=S,
    synthetic_train_size=total * N, batched_clients=(MODE == "batched"), shuffle_clients=False))
args = fedml.init(args)
dev = torch.device(f"cuda:{LOCAL}")
torch.cuda.set_device(dev)
if WORLD > 1:
    dist.init_process_group("nccl", device_id=dev)
    dist.barrier()
ds, ncls = fedml.data.load(args)
if getattr(args, "_federated_dataset", None) is not None:
    args._federated_dataset.pin()
model = fedml.models.create(args, ncls)
n_params = sum(p.numel() for p in model.parameters())
n_emb = sum(p.numel() for n, p in model.named_parameters() if "embeddings" in n and "LayerNorm" not in n)
from fedml_b200.simulation import SimulatorNCCL  # noqa: E402

sim = SimulatorNCCL(args, dev, ds, model).simulator
for r in range(2):
    sim.train_one_round(r)
torch.cuda.synchronize()
if WORLD > 1:
    dist.barrier()
    torch.cuda.synchronize()
native.LAUNCHES[0] = 0
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
for r in range(2, 2 + ROUNDS):
    sim.train_one_round(r)
e1.record()
torch.cuda.synchronize()
wall = (time.perf_counter() - t0) / ROUNDS * 1e3
ms = e0.elapsed_time(e1) / ROUNDS
if WORLD > 1:                                   # max over ranks, timed on the device
    t = torch.tensor([ms, wall], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, wall = float(t[0]), float(t[1])
steps = N // B
tokens = K * N * S
# 6 * (non-embedding params) FLOPs per token + attention (12 * L * S * D per token, fwd + bwd)
L, D = 6, 768
flops = tokens * (6 * (n_params - n_emb) + 12 * L * S * D)
be = getattr(sim.batched, "be", None) if sim.batched is not None else None
if RANK == 0:
    print(json.dumps({
        "config": f"DistilBERT-base ({n_params / 1e6:.1f}M params), {K} clients/round over {WORLD} GPU(s) x {N} samples, batch {B}, seq {S}, {steps} local steps",
        "n_gpus": WORLD, "phase_ms": {k: round(v / (ROUNDS + 2), 2) for k, v in sim.timer.collect().items()},
        "mode": MODE, "ms_per_round": round(ms, 2), "wall_ms_per_round": round(wall, 2), "ms_per_step": round(ms / steps, 2),
        "client_updates_per_s": round(K * 1e3 / ms, 2), "tokens_per_s": round(tokens * 1e3 / ms),
        "model_tflops": round(flops / ms / 1e9, 1), "my_launches_per_round": native.LAUNCHES[0] // ROUNDS,
        "graph_nodes_per_step": getattr(sim, "_graph_launches", None),
        "library_fallbacks": getattr(be, "fallbacks", None), "mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 1)}))
if WORLD > 1:
    dist.destroy_process_group()
