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
    if rank == 0:
        print("MULTIGPU_RESULT " + json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
