"""Single GPU: the simulator's global model after two ResNet-56 rounds on different execution paths (native tf32 with and
without CUDA graphs / fused SGD / fused head, native bf16, plain PyTorch fp32 ops), pairwise cosine of the update."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import fedml_b200 as fedml
from fedml_b200.arguments import Arguments
from fedml_b200.native.build import build
build()
from fedml_b200.simulation import SimulatorNCCL
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda:0")


def run(tag, **kw):
    base = dict(training_type="simulation", backend="NCCL", dataset="cifar10", model=os.environ.get("MODEL", "resnet56"), client_num_in_total=16,
                client_num_per_round=8, comm_round=2, epochs=1, batch_size=10, learning_rate=0.03, partition_method="homo",
                frequency_of_the_test=0, synthetic_train_size=16 * 20, log_level="WARNING", using_gpu=True, compute_dtype="tf32",
                random_seed=0)
    base.update(kw)
    a = fedml.init(Arguments.from_dict(base))
    ds, ncls = fedml.data.load(a)
    torch.manual_seed(0)
    model = fedml.models.create(a, ncls)
    sim = SimulatorNCCL(a, dev, ds, model).simulator
    init = sim.global_flat.clone()
    outs = []
    for r in range(int(os.environ.get("ROUNDS", "2"))):
        sim.train_one_round(r)
        torch.cuda.synchronize()
        outs.append((sim.global_flat - init)[:sim.P_train].double().clone())
    return outs


res = {}
res["torch"] = run("torch", batched_backend="torch", use_cuda_graphs=False)
res["native_graph"] = run("native_graph", use_cuda_graphs=True)
res["native_nograph"] = run("native_nograph", use_cuda_graphs=False)
res["native_nofuse"] = run("native_nofuse", use_cuda_graphs=False, fused_sgd_restage=False)
res["bf16"] = run("bf16", use_cuda_graphs=True, compute_dtype="bf16")
res["native_again"] = run("native_again", use_cuda_graphs=True)
for refname in ("torch", "native_graph"):
  ref = res[refname]
  print("reference:", refname)
  for k, v in res.items():
    line = []
    for r in range(len(v)):
        cos = float(torch.dot(v[r], ref[r]) / (v[r].norm() * ref[r].norm()))
        rel = float((v[r] - ref[r]).norm() / ref[r].norm())
        line.append(f"round{r}: cos {cos:.5f} rel {rel:.4f} |du| {float(v[r].norm()):.4f}")
    print(f"  {k:16s}", " | ".join(line))
