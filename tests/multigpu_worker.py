"""torchrun worker for the real multi-GPU checks (one process per GPU, NCCL bootstrap only)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device(f"cuda:{local}")
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    from fedml_b200.models import LogisticRegression
    from fedml_b200.parallel.arena import ArenaLayout
    from fedml_b200.parallel.round_engine import DPConfig, RoundEngine, ServerOptConfig
    out = {"world": world}
    # ---- fused round: one-shot and two-shot against a gloo-free oracle (all ranks generate all clients)
    for two_shot in (False, True):
        for opt in (ServerOptConfig(), ServerOptConfig("adam", lr=0.05)):
            layout = type("L", (), {"P": 4 * 50021, "P_train": 4 * 48000})()
            eng = RoundEngine(layout, dev, opt=opt, two_shot=two_shot, max_clients=8)
            torch.manual_seed(0)
            g0 = torch.randn(layout.P, device=dev)
            eng.set_global(g0)
            K = 4 * world
            ref = g0.double().clone()
            m = torch.zeros(layout.P_train, dtype=torch.float64, device=dev); v = torch.zeros_like(m)
            for rnd in range(3):
                torch.manual_seed(100 + rnd)
                X = g0 + 0.05 * torch.randn(K, layout.P, device=dev)
                w = torch.arange(1, K + 1, device=dev).float()
                mine = slice(rank * 4, rank * 4 + 4)
                outg = eng.aggregate(X[mine].contiguous(), w[mine].contiguous(), 1.0 / float(w.sum()), rank * 4, rnd)
                avg = (w[:, None].double() * X.double()).sum(0) / w.double().sum()
                if opt.name == "adam":
                    g = ref[:layout.P_train] - avg[:layout.P_train]
                    m = 0.9 * m + 0.1 * g; v = 0.99 * v + 0.01 * g * g
                    t = rnd + 1
                    ref[:layout.P_train] -= (0.05 / (1 - 0.9 ** t)) * m / (v.sqrt() / (1 - 0.99 ** t) ** 0.5 + 1e-3)
                    ref[layout.P_train:] = avg[layout.P_train:]
                else:
                    ref = avg
                torch.cuda.synchronize()
                err = float((outg.double() - ref).abs().max())
                gathered = [torch.zeros_like(outg) for _ in range(world)]
                dist.all_gather(gathered, outg.clone())
                same = all(torch.equal(gathered[0], t_) for t_ in gathered)
                out[f"round_{'two' if two_shot else 'one'}shot_{opt.name}_r{rnd}"] = {"err": err, "identical": same}
            out.setdefault("provider", eng.region.provider)
            out.setdefault("multicast", bool(eng.region.multicast_ptr))
    # ---- Cheetah fused DP step vs all_reduce oracle
    from fedml_b200.parallel.cheetah import CheetahEngine
    torch.manual_seed(1)
    model = LogisticRegression(784, 10)
    eng = CheetahEngine(model, dev, lr=0.1, momentum=0.9)
    ref_w = eng.flat[:eng.n].clone(); ref_m = torch.zeros_like(ref_w)
    for step in range(3):
        torch.manual_seed(1000 * step + rank)
        x = torch.rand(32, 784, device=dev); y = torch.randint(0, 10, (32,), device=dev)
        eng.zero_grad()
        torch.nn.functional.cross_entropy(eng.model(x), y).backward()
        g = eng.grad[:eng.n].clone()
        dist.all_reduce(g); g /= world
        ref_m = 0.9 * ref_m + g; ref_w = ref_w - 0.1 * ref_m
        eng.step()
        torch.cuda.synchronize()
        out[f"cheetah_step{step}"] = float((eng.flat[:eng.n] - ref_w).abs().max())
    # ---- Cheetah with bucketed overlap (hooks launch each bucket's fused step on a side stream during backward)
    from fedml_b200.models import resnet56
    torch.manual_seed(2)
    net = resnet56(class_num=10)
    eng2 = CheetahEngine(net, dev, lr=0.05, momentum=0.9, weight_decay=1e-4, n_buckets=6, overlap=True)
    ref_w2 = eng2.flat[:eng2.n].clone(); ref_m2 = torch.zeros_like(ref_w2)
    out["cheetah_overlap_buckets"] = len(eng2.buckets)
    for step in range(3):
        torch.manual_seed(2000 * step + rank)
        x = torch.randn(16, 3, 32, 32, device=dev); y = torch.randint(0, 10, (16,), device=dev)
        eng2.zero_grad()
        eng2.model.train()
        torch.nn.functional.cross_entropy(eng2.model(x), y).backward()
        early = sum(eng2._launched)                      # buckets already in flight when backward returns
        torch.cuda.synchronize()
        g = eng2.grad[:eng2.n].clone()
        eng2.step()
        dist.all_reduce(g); g /= world
        d = g + 1e-4 * ref_w2
        ref_m2 = 0.9 * ref_m2 + d; ref_w2 = ref_w2 - 0.05 * ref_m2
        torch.cuda.synchronize()
        out[f"cheetah_overlap_step{step}"] = float((eng2.flat[:eng2.n] - ref_w2).abs().max())
        out[f"cheetah_overlap_early{step}"] = int(early)
    # ---- NCCL simulator end-to-end on real ranks: robust aggregators over peer rows + G-invariance of the global model
    import fedml_b200 as fedml
    from fedml_b200.arguments import Arguments
    from fedml_b200.simulation import SimulatorNCCL
    for tag, extra in (("fedavg", {}), ("median", {"robust_agg": "median", "byzantine_fraction": 0.25}),
                       ("trimmed", {"robust_agg": "trimmed_mean", "byzantine_fraction": 0.25, "trim_ratio": 0.25}),
                       ("krum", {"robust_agg": "krum", "byzantine_fraction": 0.25, "krum_m": 2}),
                       ("fedopt_dp", {"federated_optimizer": "FedOpt", "server_optimizer": "adam", "server_lr": 0.01, "dp_clip": 1.0})):
        base = dict(training_type="simulation", backend="NCCL", dataset="mnist", model="lr", client_num_per_round=8, comm_round=2,
                    batch_size=10, learning_rate=0.03, frequency_of_the_test=0, synthetic_train_size=2000, mnist_client_num=16,
                    log_level="WARNING", using_gpu=True, compute_dtype="fp32", use_cuda_graphs=False)
        base.update(extra)
        a = fedml.init(Arguments.from_dict(base))
        ds, ncls = fedml.data.load(a)
        torch.manual_seed(0)
        model = fedml.models.create(a, ncls)
        sim = SimulatorNCCL(a, dev, ds, model).simulator
        for r in range(2):
            sim.train_one_round(r)
        torch.cuda.synchronize()
        flat = sim.global_flat.clone()
        # oracle: the same rounds with the CPU/gloo-free single-process path of the simulator (world-independent maths)
        gathered = [torch.zeros_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        out[f"sim_{tag}"] = {"identical": all(torch.equal(gathered[0], g_) for g_ in gathered), "finite": bool(torch.isfinite(flat).all()),
                             "norm": float(flat.norm())}
    # ---- ResNets on the tf32 native path (CUDA graphs on) over the ranks: the global model after ONE round (two local steps)
    # is written to a file; the test compares it element-wise with one-process runs on the native path and on the plain
    # PyTorch fp32 back-end (FB_SIM_BACKEND=torch).  One round only: at random init with batch 10 these nets amplify a
    # last-bit difference by orders of magnitude per step (profiles/precision_sensitivity.md), so longer trajectories of
    # two correct implementations are not comparable element-wise.
    dump = os.environ.get("FB_DUMP_DIR")
    if dump:
        be = os.environ.get("FB_SIM_BACKEND", "native")
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
        for mname in ("resnet18", "resnet56"):
            base = dict(training_type="simulation", backend="NCCL", dataset="cifar10", model=mname, client_num_in_total=16,
                        client_num_per_round=8, comm_round=2, epochs=1, batch_size=10, learning_rate=0.03, partition_method="homo",
                        frequency_of_the_test=0, synthetic_train_size=16 * 20, log_level="WARNING", using_gpu=True,
                        compute_dtype="tf32", batched_backend=be, use_cuda_graphs=(be == "native"), random_seed=0)
            a = fedml.init(Arguments.from_dict(base))
            ds, ncls = fedml.data.load(a)
            torch.manual_seed(0)
            model = fedml.models.create(a, ncls)
            sim = SimulatorNCCL(a, dev, ds, model).simulator
            init = sim.global_flat.clone()
            sim.train_one_round(0)
            torch.cuda.synchronize()
            if sim.engine is not None:
                sim.engine.check_spin_errors()
            flat = sim.global_flat.clone()
            gathered = [torch.zeros_like(flat) for _ in range(world)]
            dist.all_gather(gathered, flat)
            out[f"sim_{mname}"] = {"identical": all(torch.equal(gathered[0], g_) for g_ in gathered),
                                   "finite": bool(torch.isfinite(flat).all()), "norm": float(flat.norm())}
            if rank == 0:
                torch.save({"init": init.cpu(), "flat": flat.cpu(), "P_train": sim.P_train},
                           os.path.join(dump, f"{mname}_{be}_{world}.pt"))
            del sim
            torch.cuda.empty_cache()
        # checkpoint with several ranks is a collective (two-shot FedAdam: the sharded moments are gathered), then resume
        base = dict(training_type="simulation", backend="NCCL", dataset="mnist", model="lr", client_num_per_round=8, comm_round=4,
                    batch_size=10, learning_rate=0.03, frequency_of_the_test=0, synthetic_train_size=2000, mnist_client_num=16,
                    log_level="WARNING", using_gpu=True, compute_dtype="fp32", use_cuda_graphs=False, federated_optimizer="FedOpt",
                    server_optimizer="adam", server_lr=0.01, two_shot=True)
        a = fedml.init(Arguments.from_dict(base))
        ds, ncls = fedml.data.load(a)
        torch.manual_seed(0)
        model = fedml.models.create(a, ncls)
        sim = SimulatorNCCL(a, dev, ds, model).simulator
        for r in range(2):
            sim.train_one_round(r)
        ck = os.path.join(dump, f"ckpt_lr_{be}_{world}")
        sim.save(ck)
        dist.barrier()
        for r in range(2, 4):
            sim.train_one_round(r)
        torch.cuda.synchronize()
        straight = sim.global_flat.clone()
        torch.manual_seed(0)
        sim2 = SimulatorNCCL(a, dev, ds, fedml.models.create(a, ncls)).simulator
        assert sim2.resume(ck) == 2
        for r in range(2, 4):
            sim2.train_one_round(r)
        torch.cuda.synchronize()
        out["resume_fedadam_two_shot"] = {"max_diff": float((sim2.global_flat - straight).abs().max()),
                                          "two_shot": bool(sim.engine.two_shot) if sim.engine is not None else None}
    if rank == 0:
        print("MULTIGPU_RESULT " + json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
