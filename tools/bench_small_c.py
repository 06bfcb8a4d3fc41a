"""Launch-bound regime probe: one GPU training C clients per round (what each of 8 GPUs sees at K=100)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import fedml_b200 as fedml
from fedml_b200.arguments import Arguments
from fedml_b200.native.build import build
from fedml_b200.ops import native
build()
K = int(os.environ.get("K", "13"))
args = Arguments.from_dict(dict(training_type="simulation", backend="NCCL", federated_optimizer="FedAvg", dataset="cifar10", model="resnet56",
    client_num_in_total=1000, client_num_per_round=K, comm_round=12, epochs=1, batch_size=10, learning_rate=0.03, partition_method="homo",
    using_gpu=True, frequency_of_the_test=0, log_level="WARNING", synthetic_train_size=50000))
args = fedml.init(args); dev = torch.device("cuda:0")
ds, ncls = fedml.data.load(args); args._federated_dataset.pin(); model = fedml.models.create(args, ncls)
from fedml_b200.simulation import SimulatorNCCL
sim = SimulatorNCCL(args, dev, ds, model).simulator
for r in range(3): sim.train_one_round(r)
torch.cuda.synchronize(); native.LAUNCHES[0] = 0
t0 = time.perf_counter()
for r in range(3, 9): sim.train_one_round(r)
torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 6 * 1e3
print(f"K={K}: {dt:.2f} ms/round, {dt/5:.2f} ms/step, my launches/step={native.LAUNCHES[0]/30:.0f}, graph nodes/step=", end="")
try:
    import re
    sim.graph.debug_dump  # noqa
except Exception:
    pass
print(getattr(sim, "_graph_launches", "?"))
