"""Per-kernel time split of one client-batched ResNet-18 round (torch.profiler)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import fedml_b200 as fedml
from fedml_b200.arguments import Arguments
from fedml_b200.native.build import build
build()
K = int(os.environ.get("K", "64"))
args = Arguments.from_dict(dict(training_type="simulation", backend="NCCL", federated_optimizer="FedAvg", dataset="cifar10", model="resnet18",
    client_num_in_total=2 * K, client_num_per_round=K, comm_round=6, epochs=1, batch_size=10, learning_rate=0.03, partition_method="homo",
    using_gpu=True, frequency_of_the_test=0, log_level="WARNING", synthetic_train_size=2 * K * 50))
args = fedml.init(args); dev = torch.device("cuda:0")
ds, ncls = fedml.data.load(args); args._federated_dataset.pin(); model = fedml.models.create(args, ncls)
from fedml_b200.simulation import SimulatorNCCL
sim = SimulatorNCCL(args, dev, ds, model).simulator
for r in range(2): sim.train_one_round(r)
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    sim.train_one_round(2); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=16, max_name_column_width=60))
