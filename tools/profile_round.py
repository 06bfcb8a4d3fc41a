"""Kernel-time breakdown of one training round (torch.profiler, no graphs) — development aid."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import fedml_b200 as fedml
from fedml_b200.arguments import Arguments
from fedml_b200.native.build import build
build()
K = int(os.environ.get("K", "100"))
args = Arguments.from_dict(dict(training_type="simulation", backend="NCCL", federated_optimizer="FedAvg", dataset="cifar10",
    model=os.environ.get("MODEL", "resnet56"), client_num_in_total=1000, client_num_per_round=K, comm_round=3, epochs=1, batch_size=10,
    learning_rate=0.03, partition_method="homo", using_gpu=True, frequency_of_the_test=0, log_level="WARNING",
    synthetic_train_size=50000, use_cuda_graphs=False))
args = fedml.init(args)
dev = torch.device("cuda:0")
ds, ncls = fedml.data.load(args); model = fedml.models.create(args, ncls)
from fedml_b200.simulation import SimulatorNCCL
sim = SimulatorNCCL(args, dev, ds, model).simulator
sim.train_one_round(0); torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    sim.train_one_round(1); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=70))
