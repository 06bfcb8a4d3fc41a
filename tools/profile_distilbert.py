"""Per-kernel time breakdown of one client-batched DistilBERT round (torch.profiler / CUPTI, graphs off)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import fedml_b200 as fedml  # noqa: E402
from fedml_b200.arguments import Arguments  # noqa: E402
from fedml_b200.native.build import build  # noqa: E402

build()
K, B, S, N = int(os.environ.get("K", "8")), int(os.environ.get("B", "8")), int(os.environ.get("S", "128")), int(os.environ.get("N", "32"))
total = max(4 * K, 32)
args = Arguments.from_dict(dict(
    training_type="simulation", backend="NCCL", federated_optimizer="FedAvg", dataset="20news", model="distilbert",
    client_num_in_total=total, client_num_per_round=K, comm_round=4, epochs=1, batch_size=B, learning_rate=0.01,
    partition_method="homo", using_gpu=True, frequency_of_the_test=0, log_level="WARNING", max_seq_length=S,
    synthetic_train_size=total * N, use_cuda_graphs=False, shuffle_clients=False))
args = fedml.init(args)
dev = torch.device("cuda:0")
ds, ncls = fedml.data.load(args)
model = fedml.models.create(args, ncls)
from fedml_b200.simulation import SimulatorNCCL  # noqa: E402

sim = SimulatorNCCL(args, dev, ds, model).simulator
sim.train_one_round(0)
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile  # noqa: E402

with profile(activities=[ProfilerActivity.CUDA]) as prof:
    sim.train_one_round(1)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=int(os.environ.get("ROWS", "40")), max_name_column_width=70))
