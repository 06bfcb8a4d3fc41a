"""Kernel-time vs wall-time of one ResNet-56 round at small client counts (graphs on): how launch-bound is it?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import fedml_b200 as fedml
from fedml_b200.arguments import Arguments
from fedml_b200.native.build import build
build()
K = int(os.environ.get("K", "13"))
args = Arguments.from_dict(dict(training_type="simulation", backend="NCCL", federated_optimizer="FedAvg", dataset="cifar10", model="resnet56",
    client_num_in_total=1000, client_num_per_round=K, comm_round=12, epochs=1, batch_size=10, learning_rate=0.03, partition_method="homo",
    using_gpu=True, frequency_of_the_test=0, log_level="WARNING", synthetic_train_size=50000,
    compute_dtype=os.environ.get("DTYPE", "tf32")))
args = fedml.init(args); dev = torch.device("cuda:0")
ds, ncls = fedml.data.load(args); args._federated_dataset.pin(); model = fedml.models.create(args, ncls)
from fedml_b200.simulation import SimulatorNCCL
sim = SimulatorNCCL(args, dev, ds, model).simulator
for r in range(3): sim.train_one_round(r)
torch.cuda.synchronize()
if os.environ.get("TIME_ONLY"):          # plain CUDA-event timing of 8 rounds, no profiler
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(3, 11): sim.train_one_round(r)
    e1.record(); torch.cuda.synchronize()
    print(f"TIME K={K} dtype={args.compute_dtype} FB_BN_FUSED={os.environ.get('FB_BN_FUSED')} FB_BN_RESIDENT={os.environ.get('FB_BN_RESIDENT')}: "
          f"{e0.elapsed_time(e1) / 8:.3f} ms/round")
    sys.exit(0)
from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    sim.train_one_round(3); torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
tot = sum(e.device_time if hasattr(e, "device_time") else e.cuda_time for e in ev)
t0 = min(e.time_range.start for e in ev); t1 = max(e.time_range.end for e in ev)
print(f"K={K}: {len(ev)} device activities, kernel time sum {tot/1e3:.2f} ms, span {(t1-t0)/1e3:.2f} ms")
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=60))
