"""Kernel split of the Krum aggregation (ResNet-18, 64 clients) - torch.profiler around sim._aggregate_robust."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import fedml_b200 as fedml
from fedml_b200.arguments import Arguments
K = 64
args = Arguments.from_dict(dict(training_type="simulation", backend="NCCL", federated_optimizer="FedAvg", dataset="cifar10", model="resnet18",
    client_num_in_total=128, client_num_per_round=K, comm_round=4, epochs=1, batch_size=10, learning_rate=0.03, partition_method="homo",
    using_gpu=True, frequency_of_the_test=0, log_level="WARNING", synthetic_train_size=128 * 20, robust_agg="krum", byzantine_fraction=0.1, krum_m=32))
args = fedml.init(args); dev = torch.device("cuda:0")
ds, ncls = fedml.data.load(args); model = fedml.models.create(args, ncls)
from fedml_b200.simulation import SimulatorNCCL
sim = SimulatorNCCL(args, dev, ds, model).simulator
sim.train_one_round(0); torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    sim.train_one_round(1); torch.cuda.synchronize()
rows = [e for e in prof.key_averages() if any(t in e.key for t in ("krum", "gram", "weighted", "round_kernel", "robust", "reduce", "gemm_kernel<0", "accumulate", "elementwise", "sort", "topk"))]
for e in sorted(rows, key=lambda e: -e.device_time_total)[:14]:
    print(f"{e.key[:70]:70s} {e.device_time_total/1e3:8.2f} ms  x{e.count}")
