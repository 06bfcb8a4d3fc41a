"""ncu target: two client-batched ResNet-18 rounds without CUDA graphs (kernel filter picks what to profile)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import fedml_b200 as fedml
from fedml_b200.arguments import Arguments
K = int(os.environ.get("K", "32"))
args = Arguments.from_dict(dict(training_type="simulation", backend="NCCL", federated_optimizer="FedAvg", dataset="cifar10", model="resnet18",
    client_num_in_total=2 * K, client_num_per_round=K, comm_round=3, epochs=1, batch_size=10, learning_rate=0.03, partition_method="homo",
    using_gpu=True, frequency_of_the_test=0, log_level="WARNING", synthetic_train_size=2 * K * 20, use_cuda_graphs=False))
args = fedml.init(args); dev = torch.device("cuda:0")
ds, ncls = fedml.data.load(args); model = fedml.models.create(args, ncls)
from fedml_b200.simulation import SimulatorNCCL
sim = SimulatorNCCL(args, dev, ds, model).simulator
sim.train_one_round(0); torch.cuda.synchronize(); print("ok")
