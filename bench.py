#!/usr/bin/env python
"""Headline benchmark: FL rounds/sec and client-updates/sec of the multi-GPU ("NCCL") simulator.

Config (BASELINE.json #2): FedAvg, ResNet-56 (reference architecture, 591,322 params), CIFAR-10-shaped
synthetic data, 1000 clients / 100 per round, batch 10, 1 local epoch (5 SGD steps per client), SGD lr 0.03,
random-init weights.  A "step" is one FL round: sample clients, stage their data host->device, local
training of all sampled clients, fused aggregation + redistribution of the global model.

  python bench.py --gpus N --steps K --warmup W            product arm (fedml_b200)
  python bench.py --impl reference --gpus N ...            the unmodified reference (baseline/reference_arm.py)

Timing: per-round CUDA events on the launching stream (device time, max over ranks) for `value`; the
`e2e` figure is wall-clock around the same public call (`NCCLSimulator.train_one_round`) including host
staging, the pinned H2D copy of that round's inputs and a D2H read of the round's mean loss.  L2 is
flushed between rounds (512 MiB memset) and the per-round working set (>= 240 MB of client models plus
activations) is larger than L2 anyway.  Clocks / throttle reasons are sampled with nvidia-smi during the
timed region.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = dict(model="resnet56", dataset="cifar10", client_num_in_total=1000, client_num_per_round=100, epochs=1,
           batch_size=10, learning_rate=0.03, samples_per_client=50, partition_method="homo")
METRIC = "client_updates_per_sec"


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); smax = float(r[2])
            except Exception:
                continue
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            for n, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        busy = [v for v in sm if smax and v > 0.5 * smax] or sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


def run_product(a):
    import torch
    import torch.distributed as dist
    import fedml_b200 as fedml
    from fedml_b200.arguments import Arguments
    from fedml_b200.native.build import build
    from fedml_b200.ops import native
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if rank == 0:
        build()
    args = Arguments.from_dict(dict(
        training_type="simulation", backend="NCCL", federated_optimizer="FedAvg", dataset=CFG["dataset"],
        model=CFG["model"], client_num_in_total=CFG["client_num_in_total"],
        client_num_per_round=CFG["client_num_per_round"], comm_round=a.warmup + a.steps, epochs=CFG["epochs"],
        batch_size=CFG["batch_size"], learning_rate=CFG["learning_rate"], partition_method=CFG["partition_method"],
        using_gpu=True, frequency_of_the_test=0, log_level="WARNING", compute_dtype=a.dtype,
        synthetic_train_size=CFG["client_num_in_total"] * CFG["samples_per_client"], random_seed=0))
    args = fedml.init(args)
    device = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(device)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=device)
        dist.barrier()
    dataset, class_num = fedml.data.load(args)
    args._federated_dataset.pin()
    model = fedml.models.create(args, class_num)
    from fedml_b200.simulation import SimulatorNCCL
    sim = SimulatorNCCL(args, device, dataset, model).simulator
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=device)

    def barrier():
        torch.cuda.synchronize(device)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    def one_round(r):
        sim.train_one_round(r)
        return float(sim._static["loss"].mean()) if sim._static is not None else 0.0     # D2H read of the round's loss

    for r in range(a.warmup):
        one_round(r)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    native.LAUNCHES[0] = 0
    h2d0 = sim.h2d_bytes
    dev_ms, wall_ms = [], []
    for i in range(a.steps):
        flush.zero_()                               # L2 flush between timed iterations (not timed)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        loss = one_round(a.warmup + i)
        e1.record()
        torch.cuda.synchronize(device)
        wall = (time.perf_counter() - t0) * 1e3
        t = torch.tensor([e0.elapsed_time(e1), wall], device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)    # max over ranks
        dev_ms.append(float(t[0])); wall_ms.append(float(t[1]))
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = native.LAUNCHES[0]
    if rank == 0:
        K = CFG["client_num_per_round"]
        ms = sum(dev_ms) / len(dev_ms)
        e2e_ms = sum(wall_ms) / len(wall_ms)
        phases = {k: round(v / (a.warmup + a.steps), 3) for k, v in sim.timer.collect().items()}
        fallbacks = getattr(getattr(sim.batched, "be", None), "fallbacks", {}) if sim.batched is not None else {}
        out = {"metric": METRIC, "value": K * 1e3 / ms, "unit": "client-updates/s", "n_gpus": world, "steps": a.steps,
               "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
               "dtype": a.dtype, "data": "synthetic (CIFAR-10-shaped uint8 images, random-init weights)",
               "rounds_per_sec": 1e3 / ms,
               "config": {"model": "resnet56", "algorithm": "FedAvg", "clients_total": CFG["client_num_in_total"],
                          "clients_per_round": K, "global_batch": K * CFG["batch_size"], "batch_size": CFG["batch_size"],
                          "local_epochs": 1, "local_steps": CFG["samples_per_client"] // CFG["batch_size"],
                          "seq_len": None, "parallelism": f"client-parallel x{world} (clients sharded over GPUs)",
                          "l2": "flushed between rounds (512 MiB memset) and working set > L2",
                          "timing": "CUDA events per round, max over ranks"},
               "clocks": clocks,
               "e2e": {"value": K * 1e3 / e2e_ms, "unit": "client-updates/s", "ms_per_step": e2e_ms,
                       "h2d_bytes_per_step": (sim.h2d_bytes - h2d0) // max(a.steps, 1), "d2h_bytes_per_step": 4,
                       "api": "fedml_b200.simulation.SimulatorNCCL(...).simulator.train_one_round"},
               "gpu_launches": launches, "phase_ms_per_round": phases,
               "library_fallbacks_per_run": fallbacks, "impl": "fedml_b200"}
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_reference(a):
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    try:
        import reference_arm
        cfg = dict(CFG)
        sampler = ClockSampler(0)
        if rank == 0:
            sampler.start()
        res = reference_arm.run(cfg, a.steps, a.warmup, world)
        clocks = sampler.stop() if rank == 0 else None
    except Exception as exc:  # noqa: BLE001
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": f"{type(exc).__name__}: {exc}"[:300]}), flush=True)
        return
    if rank != 0 or res is None:
        return
    K = CFG["client_num_per_round"]
    out = {"impl": "reference", "metric": METRIC, "value": res["client_updates_per_sec"], "unit": "client-updates/s",
           "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": res["ms_per_step"],
           "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "fp32 (cuDNN TF32 conv default)",
           "data": "synthetic (CIFAR-10-shaped, random-init weights)", "rounds_per_sec": res["rounds_per_sec"],
           "config": {"model": "resnet56", "algorithm": "FedAvg", "clients_total": CFG["client_num_in_total"],
                      "clients_per_round": K, "batch_size": CFG["batch_size"], "local_epochs": 1,
                      "parallelism": "reference SP simulator: sequential clients on one GPU",
                      "timing": "wall clock with cuda synchronize at round boundaries"},
           "clocks": clocks, "gpus_used": 1,
           "note": ("reference snapshot has no working multi-GPU simulator (SimulatorNCCL stub, MPI needs mpi4py): "
                    "single-GPU SP simulator result") if world > 1 else "reference SP simulator via its public API"}
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="fedml_b200", choices=["fedml_b200", "reference"])
    ap.add_argument("--dtype", default="bf16")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_product(a)


if __name__ == "__main__":
    main()
