#!/usr/bin/env python
"""Benchmarks of the multi-GPU ("NCCL") simulator: FL rounds/sec and client-updates/sec (BASELINE.json metric).

Default config = BASELINE.json #2, the headline: FedAvg, ResNet-56 (reference architecture, 591,322 params),
CIFAR-10-shaped synthetic data, 1000 clients / 100 per round, batch 10, 1 local epoch (5 SGD steps per client), SGD lr
0.03, random-init weights.  A "step" is one FL round: sample clients, stage their data host->device, local training of all
sampled clients, fused aggregation + redistribution of the global model.

  python bench.py --gpus N --steps K --warmup W            product arm (fedml_b200), default --dtype tf32
  python bench.py --impl reference --gpus N ...            the unmodified reference (baseline/reference_arm.py)
  python bench.py --config NAME                            other BASELINE.json configs (see CONFIGS below)

Precision: the reference trains in fp32 with cuDNN's TF32 convolutions (stock nn.Conv2d, no autocast).  The product
arm's default is the same: fp32 activations and weights in HBM, products on tcgen05 kind::tf32 with fp32 accumulation
(`--dtype tf32`).  The bf16 path (`--dtype bf16`) is reported as the extra key "bf16" of the same JSON line.

Timing: per-round CUDA events on the launching stream (device time, max over ranks) for `value`; the `e2e` figure is
wall-clock around the same public call (`NCCLSimulator.train_one_round`) including host staging, the pinned H2D copy of
that round's inputs and a D2H read of the round's mean loss.  L2 is flushed between rounds (512 MiB memset) and the
per-round working set is larger than L2 anyway.  Clocks / throttle reasons are sampled with nvidia-smi during the timed
region.
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

METRIC = "client_updates_per_sec"

# name -> simulator arguments + what the reference arm can do with it
CONFIGS = {
    # BASELINE.json #2 (headline)
    "resnet56_fedavg": dict(
        model="resnet56", dataset="cifar10", client_num_in_total=1000, client_num_per_round=100, epochs=1, batch_size=10,
        learning_rate=0.03, samples_per_client=50, partition_method="homo", federated_optimizer="FedAvg", extra={},
        reference="sp_fedavg", desc="FedAvg ResNet-56, CIFAR-10-shaped, 1000 clients / 100 per round"),
    # BASELINE.json #3
    "resnet18_fedopt_dp": dict(
        model="resnet18", dataset="cifar10", client_num_in_total=1024, client_num_per_round=512, epochs=1, batch_size=10,
        learning_rate=0.03, samples_per_client=50, partition_method="homo", federated_optimizer="FedOpt",
        extra=dict(server_optimizer="adam", server_lr=0.01, dp_clip=1.0, dp_sigma=0.001),
        reference=None, reference_note="reference lacks ResNet-18(BN), DP clip+noise on FedOpt and a multi-GPU simulator "
                                       "(SURVEY 0.2); its FedOpt exists only for the SP/MPI simulators",
        desc="FedOpt (FedAdam) + DP clip/noise, ResNet-18, 512 clients per round"),
    # BASELINE.json #4
    "resnet18_robust_median": dict(
        model="resnet18", dataset="cifar10", client_num_in_total=1024, client_num_per_round=512, epochs=1, batch_size=10,
        learning_rate=0.03, samples_per_client=50, partition_method="homo", federated_optimizer="FedAvg",
        extra=dict(robust_agg="median", byzantine_fraction=0.1), reference=None,
        reference_note="reference's coordinate median is a CPU torch.median over stacked state_dicts in the MPI simulator "
                       "only (needs mpi4py); no ResNet-18(BN)",
        desc="Byzantine-robust FedAvg (coordinate median), ResNet-18, 10 % malicious, 512 clients per round"),
    "resnet18_robust_trimmed": dict(
        model="resnet18", dataset="cifar10", client_num_in_total=1024, client_num_per_round=512, epochs=1, batch_size=10,
        learning_rate=0.03, samples_per_client=50, partition_method="homo", federated_optimizer="FedAvg",
        extra=dict(robust_agg="trimmed_mean", byzantine_fraction=0.1), reference=None,
        reference_note="trimmed mean is absent from the reference (SURVEY 0.2)",
        desc="Byzantine-robust FedAvg (trimmed mean), ResNet-18, 10 % malicious, 512 clients per round"),
    "resnet18_robust_krum": dict(
        model="resnet18", dataset="cifar10", client_num_in_total=1024, client_num_per_round=512, epochs=1, batch_size=10,
        learning_rate=0.03, samples_per_client=50, partition_method="homo", federated_optimizer="FedAvg",
        extra=dict(robust_agg="krum", byzantine_fraction=0.1, krum_m=256), reference=None,
        reference_note="Krum is absent from the reference (SURVEY 0.2)",
        desc="Byzantine-robust FedAvg (multi-Krum), ResNet-18, 10 % malicious, 512 clients per round"),
    # BASELINE.json #5 (bf16 by the config's own statement)
    "distilbert_fednlp": dict(
        model="distilbert", dataset="20news", client_num_in_total=1024, client_num_per_round=256, epochs=1, batch_size=8,
        learning_rate=0.01, samples_per_client=32, partition_method="homo", federated_optimizer="FedAvg",
        extra=dict(max_seq_length=128, shuffle_clients=False), reference=None, force_dtype="bf16",
        reference_note="DistilBERT / FedNLP is not in the reference snapshot (SURVEY 0.2)",
        desc="FedNLP DistilBERT-base text classification, 256 clients per round, bf16, seq 128"),
    # BASELINE.json #1 (CPU, single process)
    "sp_mnist_lr": dict(
        model="lr", dataset="mnist", client_num_in_total=1000, client_num_per_round=10, epochs=1, batch_size=10,
        learning_rate=0.03, samples_per_client=60, partition_method="hetero", federated_optimizer="FedAvg", extra={},
        reference="sp_fedavg_cpu", cpu=True,
        desc="single-process FedAvg logistic regression, MNIST-shaped synthetic data, 1000 clients / 10 per round (CPU)"),
}


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE,
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


def _sim_args(cfg, a, dtype, rounds):
    from fedml_b200.arguments import Arguments
    d = dict(training_type="simulation", backend="NCCL", federated_optimizer=cfg["federated_optimizer"],
             dataset=cfg["dataset"], model=cfg["model"], client_num_in_total=cfg["client_num_in_total"],
             client_num_per_round=cfg["client_num_per_round"], comm_round=rounds, epochs=cfg["epochs"],
             batch_size=cfg["batch_size"], learning_rate=cfg["learning_rate"], partition_method=cfg["partition_method"],
             using_gpu=True, frequency_of_the_test=0, log_level="WARNING", compute_dtype=dtype,
             synthetic_train_size=cfg["client_num_in_total"] * cfg["samples_per_client"], random_seed=0)
    d.update(cfg.get("extra", {}))
    return Arguments.from_dict(d)


def _measure(cfg, a, dtype, device, rank, world, local_rank, with_clocks=True):
    """Build a simulator in ``dtype`` and time ``a.steps`` rounds after ``a.warmup`` warm-up rounds."""
    import torch
    import torch.distributed as dist
    import fedml_b200 as fedml
    from fedml_b200.ops import native
    args = fedml.init(_sim_args(cfg, a, dtype, a.warmup + a.steps))
    dataset, class_num = fedml.data.load(args)
    if getattr(args, "_federated_dataset", None) is not None:
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
        return float(sim._static["loss"].mean()) if getattr(sim, "_static", None) is not None else float(sim.global_flat[0])

    for r in range(a.warmup):
        one_round(r)
    sampler = ClockSampler(local_rank)
    if rank == 0 and with_clocks:
        sampler.start()
    barrier()
    sim.timer.reset()                                # phase breakdown over the timed rounds only
    native.LAUNCHES[0] = 0
    h2d0 = sim.h2d_bytes
    dev_ms, wall_ms = [], []
    for i in range(a.steps):
        flush.zero_()                               # L2 flush between timed iterations (not timed)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        one_round(a.warmup + i)                     # includes the D2H read of the round's mean loss
        e1.record()
        torch.cuda.synchronize(device)
        wall = (time.perf_counter() - t0) * 1e3
        t = torch.tensor([e0.elapsed_time(e1), wall], device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)    # max over ranks
        dev_ms.append(float(t[0])); wall_ms.append(float(t[1]))
    barrier()
    if sim.engine is not None:
        sim.engine.check_spin_errors()              # a cross-rank wait that gave up would invalidate the numbers
    clocks = sampler.stop() if (rank == 0 and with_clocks) else None
    launches = native.LAUNCHES[0]
    phases = {k: round(v / a.steps, 3) for k, v in sim.timer.collect().items()}
    fallbacks = dict(getattr(getattr(sim.batched, "be", None), "fallbacks", {}) or {}) if sim.batched is not None else {}
    # every timed round stages exactly one round of inputs (its own or, prefetched during it, the next one's)
    h2d = (sim.h2d_bytes - h2d0) // max(a.steps, 1)
    res = dict(ms=sum(dev_ms) / len(dev_ms), e2e_ms=sum(wall_ms) / len(wall_ms), clocks=clocks, launches=launches,
               phases=phases, fallbacks=fallbacks, h2d=h2d,
               mem_gb=round(torch.cuda.max_memory_allocated(device) / 2 ** 30, 1))
    del sim, flush
    torch.cuda.empty_cache()
    return res


def run_sp_cpu(a, cfg):
    """BASELINE.json #1: the single-process simulator on the CPU (no GPU involved), timed with the wall clock."""
    import fedml_b200 as fedml
    from fedml_b200.arguments import Arguments
    args = Arguments.from_dict(dict(
        training_type="simulation", backend="sp", federated_optimizer="FedAvg", dataset=cfg["dataset"], model=cfg["model"],
        client_num_in_total=cfg["client_num_in_total"], client_num_per_round=cfg["client_num_per_round"],
        comm_round=a.warmup + a.steps, epochs=1, batch_size=cfg["batch_size"], learning_rate=cfg["learning_rate"],
        partition_method=cfg["partition_method"], using_gpu=False, frequency_of_the_test=10 ** 9, log_level="WARNING",
        synthetic_train_size=cfg["client_num_in_total"] * cfg["samples_per_client"], random_seed=0))
    args = fedml.init(args)
    import torch
    dataset, class_num = fedml.data.load(args)
    model = fedml.models.create(args, class_num)
    from fedml_b200.simulation.sp.fedavg_api import FedAvgAPI
    api = FedAvgAPI(args, torch.device("cpu"), dataset, model)
    api._local_test_on_all_clients = lambda *x, **k: {}      # the metric counts training rounds only (both arms)
    stamps = []
    orig = api._client_sampling if hasattr(api, "_client_sampling") else None
    t_all = time.perf_counter()
    if orig is not None:
        def stamped(*x, **k):
            stamps.append(time.perf_counter())
            return orig(*x, **k)
        api._client_sampling = stamped
    api.train()
    stamps.append(time.perf_counter())
    if len(stamps) >= a.warmup + a.steps + 1:
        per = [b - c for c, b in zip(stamps[:-1], stamps[1:])][a.warmup:]
        ms = 1e3 * sum(per) / len(per)
    else:
        ms = 1e3 * (time.perf_counter() - t_all) / (a.warmup + a.steps)
    K = cfg["client_num_per_round"]
    # the same rounds on the client-batched engine (NCCL simulator, CPU tensors): all sampled clients train as one stacked model
    batched = None
    try:
        import copy
        from fedml_b200.simulation import SimulatorNCCL
        b_args = copy.copy(args)
        b_args.backend, b_args.frequency_of_the_test = "NCCL", 0
        sim = SimulatorNCCL(b_args, torch.device("cpu"), dataset, fedml.models.create(args, class_num)).simulator
        per_b = []
        for r in range(a.warmup + a.steps):
            t0 = time.perf_counter()
            sim.train_one_round(r)
            per_b.append(time.perf_counter() - t0)
        b_ms = 1e3 * sum(per_b[a.warmup:]) / a.steps
        batched = {"ms_per_step": b_ms, "value": K * 1e3 / b_ms, "engine": "client-batched (simulation.nccl on CPU tensors), same rounds"}
    except Exception as exc:  # noqa: BLE001 - the headline of this config is the SP simulator above
        batched = {"error": repr(exc)}
    print(json.dumps({"metric": METRIC, "value": K * 1e3 / ms, "unit": "client-updates/s", "n_gpus": 0, "steps": a.steps,
                      "client_batched": batched,
                      "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "none",
                      "vs_baseline": None, "dtype": "fp32", "data": "synthetic (MNIST-shaped)", "rounds_per_sec": 1e3 / ms,
                      "config": {"name": a.config, "desc": cfg["desc"], "model": "lr", "clients_per_round": K,
                                 "parallelism": "single process, CPU"}, "gpu_launches": 0, "impl": "fedml_b200"}), flush=True)


def run_product(a):
    cfg = CONFIGS[a.config]
    if cfg.get("cpu"):
        return run_sp_cpu(a, cfg)
    import torch
    import torch.distributed as dist
    from fedml_b200.native.build import build
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if rank == 0:
        build()
    device = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(device)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=device)
        dist.barrier()
    dtype = cfg.get("force_dtype") or a.dtype
    res = _measure(cfg, a, dtype, device, rank, world, local_rank)
    extra_bf16 = None
    if dtype != "bf16" and not a.no_bf16:
        try:
            r2 = _measure(cfg, a, "bf16", device, rank, world, local_rank, with_clocks=False)
            extra_bf16 = {"value": cfg["client_num_per_round"] * 1e3 / r2["ms"], "ms_per_step": r2["ms"],
                          "e2e_value": cfg["client_num_per_round"] * 1e3 / r2["e2e_ms"],
                          "note": "same config with bf16 activations / kind::f16 products (lower precision than the "
                                  "reference arm; not the headline)"}
        except Exception as exc:  # noqa: BLE001
            extra_bf16 = {"error": f"{type(exc).__name__}: {exc}"[:200]}
    if rank == 0:
        K = cfg["client_num_per_round"]
        ms, e2e_ms = res["ms"], res["e2e_ms"]
        out = {"metric": METRIC, "value": K * 1e3 / ms, "unit": "client-updates/s", "n_gpus": world, "steps": a.steps,
               "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
               "dtype": dtype, "data": "synthetic (shape of the named dataset, random-init weights)",
               "rounds_per_sec": 1e3 / ms,
               "config": {"name": a.config, "desc": cfg["desc"], "model": cfg["model"],
                          "algorithm": cfg["federated_optimizer"], "clients_total": cfg["client_num_in_total"],
                          "clients_per_round": K, "global_batch": K * cfg["batch_size"], "batch_size": cfg["batch_size"],
                          "local_epochs": 1, "local_steps": cfg["samples_per_client"] // cfg["batch_size"],
                          "seq_len": cfg.get("extra", {}).get("max_seq_length"),
                          "parallelism": f"client-parallel x{world} (clients sharded over GPUs)",
                          "precision": {"tf32": "fp32 activations/weights in HBM, tcgen05 kind::tf32 products, fp32 accumulate "
                                                "(= the reference arm: fp32 with cuDNN TF32 convolutions)",
                                        "fp32": "as tf32", "bf16": "bf16 activations, kind::f16 products, fp32 master weights"
                                        }.get(dtype, dtype),
                          "l2": "flushed between rounds (512 MiB memset) and working set > L2",
                          "timing": "CUDA events per round, max over ranks"},
               "clocks": res["clocks"],
               "e2e": {"value": K * 1e3 / e2e_ms, "unit": "client-updates/s", "ms_per_step": e2e_ms,
                       "h2d_bytes_per_step": res["h2d"], "d2h_bytes_per_step": 4,
                       "api": "fedml_b200.simulation.SimulatorNCCL(...).simulator.train_one_round"},
               "gpu_launches": res["launches"], "phase_ms_per_round": res["phases"],
               "library_fallbacks_per_run": res["fallbacks"], "mem_gb": res["mem_gb"], "bf16": extra_bf16,
               "impl": "fedml_b200"}
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_reference(a):
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    cfg = dict(CONFIGS[a.config])
    if cfg.get("reference") is None:
        if rank == 0:
            print(json.dumps({"impl": "reference", "config": a.config, "unavailable": cfg.get("reference_note", "no reference path")}),
                  flush=True)
        return
    try:
        import reference_arm
        sampler = ClockSampler(0)
        if rank == 0:
            sampler.start()
        cfg["cpu"] = bool(cfg.get("cpu"))
        res = reference_arm.run(cfg, a.steps, a.warmup, world)
        clocks = sampler.stop() if rank == 0 else None
    except Exception as exc:  # noqa: BLE001
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": f"{type(exc).__name__}: {exc}"[:300]}), flush=True)
        return
    if rank != 0 or res is None:
        return
    K = cfg["client_num_per_round"]
    out = {"impl": "reference", "metric": METRIC, "value": res["client_updates_per_sec"], "unit": "client-updates/s",
           "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": res["ms_per_step"],
           "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
           "dtype": "fp32" if cfg.get("cpu") else "fp32 (cuDNN TF32 conv default)",
           "data": "synthetic (shape of the named dataset, random-init weights)", "rounds_per_sec": res["rounds_per_sec"],
           "config": {"name": a.config, "desc": cfg["desc"], "model": cfg["model"], "algorithm": "FedAvg",
                      "clients_total": cfg["client_num_in_total"], "clients_per_round": K, "batch_size": cfg["batch_size"],
                      "local_epochs": 1,
                      "parallelism": "reference SP simulator: sequential clients on one " + ("CPU process" if cfg.get("cpu") else "GPU"),
                      "timing": "wall clock with cuda synchronize at round boundaries"},
           "clocks": clocks, "gpus_used": 0 if cfg.get("cpu") else 1,
           "note": ("reference snapshot has no working multi-GPU simulator (SimulatorNCCL stub, MPI needs mpi4py): "
                    "single-GPU SP simulator result") if world > 1 else "reference SP simulator via its public API"}
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="fedml_b200", choices=["fedml_b200", "reference"])
    ap.add_argument("--dtype", default="tf32", choices=["tf32", "fp32", "bf16"])
    ap.add_argument("--config", default="resnet56_fedavg", choices=sorted(CONFIGS))
    ap.add_argument("--no-bf16", action="store_true", help="skip the extra bf16 measurement")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_product(a)


if __name__ == "__main__":
    main()
