"""Aggregation-path benchmark for the BASELINE.json configs (run under torchrun, one rank per GPU).

For each (model size P, clients per round K, algorithm) it times, on the device (CUDA events, max over ranks):
  fused   — ONE launch of the round kernel (weighting [+ DP clip/noise] + cross-GPU reduce over peer memory +
            server optimizer + redistribution), one-shot or two-shot;
  nccl    — the "baseline, not the product": local weighted sum with torch ops, dist.all_reduce (NCCL), torch
            server-optimizer step (what a straightforward NCCL simulator does);
and reports the roofline fraction of the fused path: roofline time = max(local HBM bytes / 6428 GB/s,
NVLink bytes in+out per GPU / link peak per direction).  The link peak is MEASURED on the box first (``LINK`` line): every
rank copies 256 MiB out of the next rank's peer-mapped buffer with a plain device copy (ordinary loads over NVLink, all ranks
at once); 770 GB/s is only the fallback when there is a single rank.
"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

HBM, NVL = 6428e9, 770e9


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    dev = torch.device(f"cuda:{local}"); torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from fedml_b200.parallel.round_engine import DPConfig, RoundEngine, ServerOptConfig
    cfgs = [("resnet56_fedavg", 600320, 100, "none", False), ("resnet18_fedadam_dp", 11191264, 512, "adam", True),
            ("resnet18_fedadam_dp_aggnoise", 11191264, 512, "adam", "agg"), ("distilbert_fedavg", 66968852 // 4 * 4, 256, "none", False)]
    out = []
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    global NVL
    if world > 1:
        from fedml_b200.parallel.symmetric import SymmetricRegion, _RawCudaBuffer
        nb = 256 << 20
        reg = SymmetricRegion(nb + 4096, dev)
        reg.carve("buf", nb)
        reg.tensor("buf", torch.uint8).fill_(rank + 1)
        reg.barrier()
        peer = torch.as_tensor(_RawCudaBuffer(reg.peer_ptr("buf", (rank + 1) % world), nb), device=dev)
        dst = torch.empty(nb, dtype=torch.uint8, device=dev)
        for _ in range(2):
            dst.copy_(peer)
        ts = []
        for _ in range(5):
            torch.cuda.synchronize(); dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); dst.copy_(peer); e1.record(); torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX); ts.append(float(t))
        assert int(dst[12345]) == (rank + 1) % world + 1
        NVL = nb / (sorted(ts)[2] * 1e-3)
        if rank == 0:
            print("LINK " + json.dumps({"G": world, "peer_copy_GBps_per_dir": round(NVL / 1e9, 1), "bytes": nb,
                                        "how": "dst.copy_(peer-mapped buffer of rank+1), all ranks at once, median of 5, max over ranks"}), flush=True)
        reg.barrier()
        del reg, peer, dst
    for name, P, K, opt, dp in cfgs:
        C = max(1, K // world)
        layout = type("L", (), {"P": P, "P_train": P})()
        for two_shot in ((False, True) if world > 1 else (False,)):
            eng = RoundEngine(layout, dev, opt=ServerOptConfig(opt, lr=0.01) if opt != "none" else None,
                              dp=DPConfig(1.0, 0.001, 7, aggregate_noise=(dp == "agg")) if dp else None, two_shot=two_shot,
                              max_clients=C)
            stack = torch.randn(C, P, device=dev) * 0.01
            w = torch.rand(C, device=dev) + 1.0
            eng.set_global(torch.zeros(P))
            def fused():
                eng.aggregate(stack, w, 1.0 / (float(w.sum()) * world), weight_sq_sum=float((w * w).sum()) * world)
            # NCCL(+torch) baseline
            g = torch.zeros(P, device=dev); m = torch.zeros(P, device=dev); v = torch.zeros(P, device=dev)
            def nccl():
                part = (w[:, None] * stack).sum(0) if not dp else None
                if dp:
                    diff = stack - g
                    nrm = diff.norm(dim=1)
                    diff = diff / torch.clamp(nrm / 1.0, min=1.0)[:, None]
                    diff.add_(torch.randn_like(diff), alpha=0.001)
                    part = (w[:, None] * (g + diff)).sum(0)
                part /= float(w.sum()) * world
                if world > 1:
                    dist.all_reduce(part)
                if opt == "adam":
                    gr = g - part
                    m.mul_(0.9).add_(gr, alpha=0.1); v.mul_(0.99).addcmul_(gr, gr, value=0.01)
                    g.addcdiv_(m, v.sqrt().add_(1e-3), value=-0.01)
                else:
                    g.copy_(part)
            res = {}
            for label, fn in (("fused", fused), ("nccl", nccl)):
                if label == "nccl" and two_shot:
                    continue
                for _ in range(3):
                    fn()
                ts = []
                for _ in range(5):
                    flush.zero_(); torch.cuda.synchronize()
                    if world > 1:
                        dist.barrier()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(); fn(); e1.record(); torch.cuda.synchronize()
                    t = torch.tensor([e0.elapsed_time(e1)], device=dev)
                    if world > 1:
                        dist.all_reduce(t, op=dist.ReduceOp.MAX)
                    ts.append(float(t))
                res[label] = sorted(ts)[len(ts) // 2]
            passes = 2 if dp else 1
            hbm_bytes = passes * C * P * 4 + 2 * P * 4
            nvl_bytes = P * 4 * (world - 1) / world if two_shot else P * 4 * (world - 1)   # per direction per GPU
            roof_ms = max(hbm_bytes / HBM, (nvl_bytes / NVL) if world > 1 else 0) * 1e3
            rec = {"config": name, "G": world, "P": P, "clients_per_gpu": C, "mode": "two_shot" if two_shot else "one_shot",
                   "multicast": bool(eng._mc_partial) and two_shot, "fused_ms": round(res["fused"], 4), "roofline_ms": round(roof_ms, 4), "roofline_frac": round(roof_ms / res["fused"], 3),
                   "bound": "HBM" if hbm_bytes / HBM >= (nvl_bytes / NVL if world > 1 else 0) else "NVLink"}
            if "nccl" in res:
                rec["nccl_torch_ms"] = round(res["nccl"], 4); rec["speedup_vs_nccl"] = round(res["nccl"] / res["fused"], 2)
            out.append(rec)
            del eng, stack
            torch.cuda.empty_cache()
    # pure cross-GPU part (C=1 per rank: reduce + redistribute a model-sized vector): NVLink bandwidth sweep
    if world > 1:
        for name, P in (("resnet18_vec", 11191264), ("distilbert_vec_fp32", 66968852 // 4 * 4)):
            layout = type("L", (), {"P": P, "P_train": P})()
            for two_shot in (False, True):
                eng = RoundEngine(layout, dev, two_shot=two_shot, max_clients=1)
                x = torch.randn(1, P, device=dev); w1 = torch.ones(1, device=dev)
                for _ in range(3):
                    eng.aggregate(x, w1, 1.0 / world)
                ts = []
                for _ in range(5):
                    flush.zero_(); torch.cuda.synchronize(); dist.barrier()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(); eng.aggregate(x, w1, 1.0 / world); e1.record(); torch.cuda.synchronize()
                    t = torch.tensor([e0.elapsed_time(e1)], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX); ts.append(float(t))
                ms = sorted(ts)[2]
                buf = torch.randn(P, device=dev)
                for _ in range(3):
                    dist.all_reduce(buf)
                tn = []
                for _ in range(5):
                    torch.cuda.synchronize(); dist.barrier()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(); dist.all_reduce(buf); e1.record(); torch.cuda.synchronize()
                    t = torch.tensor([e0.elapsed_time(e1)], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX); tn.append(float(t))
                nvl_bytes = P * 4 * (world - 1) / world if two_shot else P * 4 * (world - 1)
                out.append({"config": name, "G": world, "mode": "two_shot" if two_shot else "one_shot",
                            "multicast": bool(eng._mc_partial) and two_shot, "fused_ms": round(ms, 4),
                            "nccl_allreduce_ms": round(sorted(tn)[2], 4), "nvlink_GBps_per_dir": round(nvl_bytes / ms / 1e6, 1),
                            "frac_of_measured_link": round(nvl_bytes / NVL * 1e3 / ms, 3)})
                del eng
                torch.cuda.empty_cache()
    if rank == 0:
        for r in out:
            print("AGG " + json.dumps(r), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
