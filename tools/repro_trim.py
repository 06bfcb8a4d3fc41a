import sys, os, time
sys.path.insert(0, "/root/repo")
import torch
import fedml_b200 as fedml
from fedml_b200.arguments import Arguments
from fedml_b200.simulation import SimulatorNCCL
for agg in ("median", "trimmed_mean"):
    a = fedml.init(Arguments.from_dict(dict(training_type="simulation", backend="NCCL", dataset="mnist", model="lr", client_num_per_round=512,
        comm_round=2, batch_size=10, learning_rate=0.03, frequency_of_the_test=0, synthetic_train_size=20000, mnist_client_num=1024,
        log_level="ERROR", using_gpu=True, compute_dtype="fp32", robust_agg=agg, byzantine_fraction=0.1, random_seed=0)))
    ds, ncls = fedml.data.load(a)
    sim = SimulatorNCCL(a, torch.device("cuda:0"), ds, fedml.models.create(a, ncls)).simulator
    t0 = time.time()
    for r in range(2):
        sim.train_one_round(r)
    torch.cuda.synchronize()
    print(agg, "ok", time.time() - t0, float(sim.global_flat.norm()), flush=True)
