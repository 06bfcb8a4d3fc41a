"""CPU tests: simulators, algorithm zoo, the reference's two CI equivalence oracles, NCCL-simulator control flow."""
import copy
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import fedml
from fedml.simulation import SimulatorMPI, SimulatorNCCL, SimulatorSingleProcess
from fedml_b200.arguments import Arguments

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPU = torch.device("cpu")


def mk(**kw):
    base = dict(dataset="mnist", model="lr", client_num_per_round=4, comm_round=3, batch_size=10, learning_rate=0.03,
                frequency_of_the_test=2, synthetic_train_size=1500, mnist_client_num=12, log_level="WARNING", eval_client_cap=4,
                simulation_timeout_s=120)
    base.update(kw)
    return fedml.init(Arguments.from_dict(base))


def build(a):
    ds, out = fedml.data.load(a)
    return ds, fedml.model.create(a, out)


def test_client_sampling_is_the_reference_rule():
    from fedml_b200.simulation.sp.fedavg_api import client_sampling
    np.random.seed(5)
    ref = np.random.choice(range(1000), 10, replace=False).tolist()
    assert client_sampling(5, 1000, 10) == ref
    assert client_sampling(3, 7, 7) == list(range(7))


def test_fedavg_full_batch_all_clients_equals_centralized():
    """Reference CI oracle (ci/CI-script-fedavg.sh:42-47): full batch, 1 epoch, every client every round ==
    centralized gradient descent on the union (LR is linear in the averaged gradient)."""
    a = mk(batch_size=-1, client_num_per_round=12, comm_round=5, frequency_of_the_test=100)
    ds, model = build(a)
    fed_model = copy.deepcopy(model)
    sim = SimulatorSingleProcess(a, CPU, ds, fed_model)
    w_fed = sim.run()
    fedds = a._federated_dataset
    cen = copy.deepcopy(model)
    opt = torch.optim.SGD(cen.parameters(), lr=0.03)
    sizes = torch.tensor(fedds.client_sizes(), dtype=torch.float32)
    for _ in range(5):          # one FedAvg round of full-batch clients == one step on the size-weighted mean of client losses
        opt.zero_grad()
        loss = sum(sizes[c] / sizes.sum() * torch.nn.functional.cross_entropy(cen(fedds.client_slice(c)[0]), fedds.client_slice(c)[1])
                   for c in range(12))
        loss.backward(); opt.step()
    for k, v in cen.state_dict().items():
        torch.testing.assert_close(w_fed[k], v, rtol=1e-4, atol=1e-5)


def test_hierarchical_matches_flat_fedavg_oracle():
    """Reference CI oracle (:50-58): (global, group) rounds = (2,5) and (5,2) land within 1e-3 of each other in train accuracy."""
    accs = []
    for g, q in ((2, 5), (5, 2)):
        a = mk(federated_optimizer="HierarchicalFL", batch_size=-1, client_num_per_round=12, global_comm_round=g,
               group_comm_round=q, group_num=1, frequency_of_the_test=100, learning_rate=0.03, eval_client_cap=12)
        ds, model = build(a)
        torch.manual_seed(0); [p.data.normal_(0, 0.01) for p in model.parameters()]
        sim = SimulatorSingleProcess(a, CPU, ds, model); sim.run()
        accs.append(sim.fl_trainer.history[-1]["train_acc"])
    assert abs(accs[0] - accs[1]) < 1e-3


@pytest.mark.parametrize("alg,extra", [("FedAvg", {}), ("FedOpt", {"server_optimizer": "adam", "server_lr": 0.01}),
                                       ("FedProx", {"fedprox_mu": 0.1}), ("FedNova", {"momentum": 0.9}),
                                       ("FedAvg_robust", {"defense_type": "median", "byzantine_fraction": 0.25}),
                                       ("turbo_aggregate", {}), ("S_FedAvg", {}), ("HS_FedAvg", {"client_num_per_round": 3})])
def test_sp_algorithms_run_and_learn(alg, extra):
    a = mk(federated_optimizer=alg, **extra)
    ds, model = build(a)
    sim = SimulatorSingleProcess(a, CPU, ds, model)
    sim.run()
    h = sim.fl_trainer.history
    assert len(h) >= 2 and np.isfinite(h[-1]["test_loss"]) and h[-1]["test_loss"] <= h[0]["test_loss"] + 0.05


def test_fedprox_proximal_term_pulls_towards_global():
    from fedml_b200.simulation.trainers import ModelTrainerCLS
    a0, a1 = mk(epochs=3), mk(epochs=3, fedprox_mu=5.0, federated_optimizer="FedProx")
    ds, model = build(a0)
    w0 = copy.deepcopy(model.state_dict())
    d = []
    for a in (a0, a1):
        m = copy.deepcopy(model); t = ModelTrainerCLS(m, a); t.train(ds[5][0], CPU, a)
        d.append(sum(float((m.state_dict()[k] - w0[k]).norm()) for k in w0))
    assert d[1] < d[0]


def test_fednova_equals_fedavg_for_equal_steps_plain_sgd():
    """tau_i identical and no momentum/prox => FedNova's normalised average reduces to FedAvg."""
    res = []
    for alg in ("FedAvg", "FedNova"):
        a = mk(federated_optimizer=alg, batch_size=-1, comm_round=2)
        ds, model = build(a)
        torch.manual_seed(0); [p.data.normal_(0, 0.01) for p in model.parameters()]
        res.append(SimulatorSingleProcess(a, CPU, ds, model).run())
    for k in res[0]:
        torch.testing.assert_close(res[0][k].float(), res[1][k].float(), rtol=1e-4, atol=1e-5)


def test_decentralized_dsgd_and_pushsum_reach_consensus():
    for kw in ({}, {"mode": "PUSHSUM", "b_symmetric": False}):
        a = mk(federated_optimizer="decentralized_fl", client_num_in_total=8, iteration_number=150, **kw)
        ds, model = build(a)
        sim = SimulatorSingleProcess(a, CPU, ds, model); sim.run()
        t = sim.fl_trainer
        assert t.regret[-1] < t.regret[5] and t.consensus_distance() < 0.1


def test_vertical_fl_sp_learns():
    a = mk(federated_optimizer="classical_vertical", comm_round=3)
    sim = SimulatorSingleProcess(a, CPU, None, None); sim.run()
    assert sim.fl_trainer.history[-1]["test_acc"] > 0.8


def test_nccl_simulator_cpu_path_matches_sp_fedavg():
    """The multi-GPU simulator's control flow on CPU (batched clients, torch back-end) reproduces SP FedAvg."""
    a = mk(batch_size=10, comm_round=2, frequency_of_the_test=100)
    ds, model = build(a)
    w_sp = SimulatorSingleProcess(a, CPU, ds, copy.deepcopy(model)).run()
    a2 = copy.copy(a); a2.backend = "NCCL"
    sim = SimulatorNCCL(a2, CPU, ds, copy.deepcopy(model)); w_nc = sim.run()
    for k in w_sp:
        torch.testing.assert_close(w_nc[k].float(), w_sp[k].float(), rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("alg,extra", [("FedOpt", {"server_optimizer": "yogi", "server_lr": 0.01}), ("FedProx", {}), ("FedNova", {"momentum": 0.5}),
                                       ("FedAvg", {"robust_agg": "trimmed_mean", "byzantine_fraction": 0.25}),
                                       ("FedAvg", {"robust_agg": "krum", "byzantine_fraction": 0.25, "krum_m": 2}),
                                       ("FedAvg", {"dp_clip": 1.0, "dp_sigma": 0.001})])
def test_nccl_simulator_algorithm_variants_cpu(alg, extra):
    a = mk(federated_optimizer=alg, backend="NCCL", **extra)
    ds, model = build(a)
    sim = SimulatorNCCL(a, CPU, ds, model); sim.run()
    h = sim.simulator.history
    assert np.isfinite(h[-1]["test_loss"]) and h[-1]["test_acc"] >= 0.0


def test_nccl_simulator_resnet_batched_cpu_and_sequential_fallback_agree():
    kw = dict(dataset="cifar10", model="resnet56", client_num_in_total=4, client_num_per_round=2, synthetic_train_size=64,
              batch_size=8, comm_round=1, frequency_of_the_test=100, backend="NCCL", shuffle_clients=False)
    a = mk(**kw)
    ds, out = fedml.data.load(a)
    from fedml_b200.models.cv import Bottleneck, ResNet
    model = ResNet(Bottleneck, [1, 1, 1], num_classes=10)
    w_b = SimulatorNCCL(a, CPU, ds, copy.deepcopy(model)).run()
    a2 = copy.copy(a); a2.batched_clients = False
    w_s = SimulatorNCCL(a2, CPU, ds, copy.deepcopy(model)).run()
    for k in w_b:
        torch.testing.assert_close(w_b[k].float(), w_s[k].float(), rtol=5e-3, atol=5e-4, msg=lambda m, k=k: f"{k}: {m}")


def test_nccl_simulator_distilbert_batched_cpu_and_sequential_fallback_agree():
    """FedNLP text classification: the client-batched DistilBERT executor and per-client eager training produce the
    same global model (dropout off so that both paths are deterministic)."""
    kw = dict(dataset="20news", model="distilbert", client_num_in_total=4, client_num_per_round=2, synthetic_train_size=32,
              max_seq_length=16, batch_size=4, comm_round=1, frequency_of_the_test=100, backend="NCCL", shuffle_clients=False)
    a = mk(**kw)
    ds, out = fedml.data.load(a)
    from fedml_b200.models.nlp import DistilBertForSequenceClassification
    model = DistilBertForSequenceClassification(num_labels=out, seq_classif_dropout=0.0, vocab_size=30522, dim=32, n_layers=2,
                                                n_heads=4, hidden_dim=64, max_pos=32, dropout=0.0, attention_dropout=0.0)
    sim_b = SimulatorNCCL(a, CPU, ds, copy.deepcopy(model))
    assert sim_b.simulator.batched is not None and sim_b.simulator.batched.input_kind == "token_ids"
    w_b = sim_b.run()
    a2 = copy.copy(a); a2.batched_clients = False
    w_s = SimulatorNCCL(a2, CPU, ds, copy.deepcopy(model)).run()
    for k in w_b:
        torch.testing.assert_close(w_b[k].float(), w_s[k].float(), rtol=5e-3, atol=5e-5, msg=lambda m, k=k: f"{k}: {m}")


def test_nccl_simulator_femnist_cnn_batched_cpu_and_sequential_fallback_agree():
    """FEMNIST CNN_DropOut (reference model/cv/cnn.py:74-142): the client-batched executor and per-client eager training give
    the same global model when dropout is off."""
    kw = dict(dataset="femnist", model="cnn", client_num_in_total=4, client_num_per_round=2, synthetic_train_size=80, femnist_client_num=8,
              batch_size=8, comm_round=1, frequency_of_the_test=100, backend="NCCL", shuffle_clients=False)
    a = mk(**kw)
    ds, out = fedml.data.load(a)
    torch.manual_seed(0)
    model = fedml.model.create(a, out)
    model.dropout_1.p = model.dropout_2.p = 0.0
    sim_b = SimulatorNCCL(a, CPU, ds, copy.deepcopy(model))
    assert type(sim_b.simulator.batched).__name__ == "BatchedCNNDropOut"
    w_b = sim_b.run()
    a2 = copy.copy(a); a2.batched_clients = False
    w_s = SimulatorNCCL(a2, CPU, ds, copy.deepcopy(model)).run()
    for k in w_b:
        torch.testing.assert_close(w_b[k].float(), w_s[k].float(), rtol=5e-3, atol=5e-5, msg=lambda m, k=k: f"{k}: {m}")


def test_nccl_simulator_two_ranks_gloo_is_g_invariant(tmp_path):
    """world_size 2 over gloo on CPU == world_size 1 (same seeds => same sampled clients => same global model)."""
    script = tmp_path / "w.py"
    script.write_text(f"""
import os, sys, json, torch
sys.path.insert(0, {ROOT!r})
import fedml
from fedml_b200.arguments import Arguments
a = fedml.init(Arguments.from_dict(dict(dataset="mnist", model="lr", client_num_per_round=6, comm_round=2, batch_size=10, learning_rate=0.03,
    frequency_of_the_test=100, synthetic_train_size=1500, mnist_client_num=12, log_level="WARNING", backend="NCCL", federated_optimizer="FedOpt",
    server_optimizer="adam", server_lr=0.01)))
ds, out = fedml.data.load(a); torch.manual_seed(0); model = fedml.model.create(a, out)
from fedml.simulation import SimulatorNCCL
w = SimulatorNCCL(a, torch.device("cpu"), ds, model).run()
if int(os.environ.get("RANK", "0")) == 0:
    torch.save(w, sys.argv[1])
""")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    one, two = str(tmp_path / "one.pt"), str(tmp_path / "two.pt")
    subprocess.run([sys.executable, str(script), one], check=True, timeout=300, env=env)
    subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                    "--master-port", "29721", str(script), two], check=True, timeout=300, env=env)
    w1, w2 = torch.load(one), torch.load(two)
    for k in w1:
        torch.testing.assert_close(w1[k], w2[k], rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("alg,kw", [("FedAvg", {}), ("FedOpt", {"server_optimizer": "sgd", "server_lr": 1.0}), ("turbo_aggregate", {}),
                                    ("base_framework", {}), ("decentralized_fl", {"worker_num": 4, "comm_round": 4}),
                                    ("classical_vertical", {"worker_num": 2}), ("FedGAN", {"model": "gan"})])
def test_mp_simulator_algorithms_threaded(alg, kw):
    a = mk(federated_optimizer=alg, backend="MPI", worker_num=kw.pop("worker_num", 3), client_num_per_round=3, comm_round=kw.pop("comm_round", 2),
           run_id=f"mp-{alg}", ci=1, frequency_of_the_test=1, **kw)
    ds, model = build(a)
    r = SimulatorMPI(a, CPU, ds, model).run()
    assert r is not None and len(r.history) >= 1


def test_mp_fedavg_equals_sp_fedavg():
    a = mk(comm_round=2, client_num_per_round=3, frequency_of_the_test=100)
    ds, model = build(a)
    w_sp = SimulatorSingleProcess(a, CPU, ds, copy.deepcopy(model)).run()
    a2 = copy.copy(a); a2.backend, a2.worker_num, a2.run_id, a2.ci = "MPI", 3, "mp-eq", 1
    agg = SimulatorMPI(a2, CPU, ds, copy.deepcopy(model)).run()
    w_mp = agg.get_global_model_params()
    for k in w_sp:
        torch.testing.assert_close(w_mp[k].float(), w_sp[k].float(), rtol=1e-5, atol=1e-6)


def test_split_nn_fedgkt_fednas_fedseg_smoke():
    import torch.nn as nn
    mlp = nn.Sequential(nn.Linear(784, 32), nn.ReLU(), nn.Linear(32, 10))
    a = mk(federated_optimizer="split_nn", backend="MPI", worker_num=2, client_num_per_round=2, epochs=1, split_cut=2, run_id="sn")
    ds, _ = build(a)
    r = SimulatorMPI(a, CPU, ds, mlp).run()
    assert len(r.history) == 2
    for alg, kw in (("FedGKT", dict(dataset="cifar10", checkpoint_dir="/tmp/fb_gkt_t")), ("FedNAS", dict(dataset="cifar10", layers=2)),
                    ("FedSeg", dict(dataset="pascal_voc", model="segnet"))):
        a = mk(federated_optimizer=alg, backend="MPI", worker_num=2, client_num_per_round=2, client_num_in_total=2, comm_round=1,
               synthetic_train_size=32, batch_size=8, run_id=alg, ci=1, frequency_of_the_test=1, **kw)
        ds, model = build(a)
        r = SimulatorMPI(a, CPU, ds, model).run()
        assert len(r.history) >= 1
    assert os.path.exists("/tmp/fb_gkt_t/last.pth") and "state_dict" in torch.load("/tmp/fb_gkt_t/last.pth", weights_only=False)


def test_fedml_runner_and_one_line_api(tmp_path):
    y = tmp_path / "fedml_config.yaml"
    y.write_text(open(os.path.join(ROOT, "fedml_b200/config/simulation_sp/fedml_config.yaml")).read()
                 .replace("comm_round: 200", "comm_round: 2").replace("client_num_in_total: 1000", "client_num_in_total: 1000\n  synthetic_train_size: 1500\n  mnist_client_num: 12\n  eval_client_cap: 2"))
    w = fedml.run_simulation(argv=["--cf", str(y)])
    assert "linear.weight" in w
    a = mk()
    ds, model = build(a)
    assert "linear.weight" in fedml.FedMLRunner(a, CPU, ds, model).run()


def test_custom_client_trainer_and_server_aggregator_hooks():
    from fedml import ClientTrainer, ServerAggregator
    calls = {"train": 0, "agg": 0}

    class T(ClientTrainer):
        def get_model_params(self): return {k: v.clone() for k, v in self.model.state_dict().items()}
        def set_model_params(self, p): self.model.load_state_dict(p)
        def train(self, d, dev, args): calls["train"] += 1
        def test(self, d, dev, args): return {"test_correct": 1, "test_loss": 0.0, "test_total": 1}

    class S(ServerAggregator):
        def get_model_params(self): return self.model.state_dict()
        def set_model_params(self, p): self.model.load_state_dict(p)
        def test(self, d, dev, args): return {}
        def aggregate(self, lst): calls["agg"] += 1; return super().aggregate(lst)

    a = mk(comm_round=2)
    ds, model = build(a)
    fedml.FedMLRunner(a, CPU, ds, model, client_trainer=T(model, a), server_aggregator=S(model, a)).run()
    assert calls["train"] == 8 and calls["agg"] == 2


def test_fedavg_robust_backdoor_protocol():
    """Backdoor robust FedAvg (reference mpi_p2p_mp/fedavg_robust): the attacker (client 1) trains on the poisoned
    loader and is force-sampled every attack_freq rounds; the server defends (norm-diff clipping) and evaluates the raw
    task and the targeted task."""
    from fedml_b200.simulation.mp.fedavg_robust import ATTACKER_CLIENT_INDEX, FedAvgRobustAggregator, FedAvgRobustTrainer
    a = mk(federated_optimizer="FedAvg_robust", backend="MPI", worker_num=3, client_num_per_round=3, comm_round=4, attack_freq=2,
           defense_type="norm_diff_clipping", norm_bound=5.0, frequency_of_the_test=1, run_id="robust-bd", ci=1, target_label=7)
    ds, model = build(a)
    agg = SimulatorMPI(a, CPU, ds, model).run()
    assert isinstance(agg, FedAvgRobustAggregator)
    assert agg.adversary_fl_rounds == [1, 3]
    for r in range(4):
        picked = list(agg.client_sampling(r, int(a.client_num_in_total), 3))
        assert (picked[0] == ATTACKER_CLIENT_INDEX and len(picked) == 4) == (r in (1, 3))
    # dual evaluation every round: raw-task accuracy in history, targeted-task accuracy next to it
    assert len(agg.history) == 4 and all("targeted_acc" in h and "test_acc" in h for h in agg.history)
    assert [h["attacker_round"] for h in agg.attack_history] == [False, True, False, True]
    # the attacker's trainer swaps in the poisoned loader and its sample count
    from fedml_b200.data import load_poisoned_dataset_from_edge_case_examples
    ptrain, _, ttest, n_poison = load_poisoned_dataset_from_edge_case_examples(a)
    tr = FedAvgRobustTrainer(0, ds[5], ds[4], ds[6], ds[0], CPU, a, agg.trainer, poisoned_train_loader=ptrain,
                             num_dps_poisoned_dataset=n_poison)
    tr.update_dataset(ATTACKER_CLIENT_INDEX)
    assert tr.train_local is ptrain and tr.local_sample_number == n_poison
    tr.update_dataset(0)
    assert tr.train_local is ds[5][0]
    assert all(int(y.unique()[0]) == 7 and y.unique().numel() == 1 for _, y in ttest)


def test_s_fedavg_class_weighted_cross_entropy():
    """S-FedAvg trains every client with class-weighted CE: weight_c = n / (#classes present * count_c)
    (reference s_fedavg/fedavg_api.py:112-136)."""
    import torch.nn as nn
    from fedml_b200.simulation.sp.shapley_fedavg import SFedAvgAPI
    a = mk(federated_optimizer="S_FedAvg", comm_round=1, client_num_per_round=2, shapley_mode="exact", frequency_of_the_test=100)
    ds, model = build(a)
    api = SFedAvgAPI(a, CPU, ds, model)
    y = torch.tensor([0, 0, 0, 0, 0, 0, 3, 3, 7])
    w = api.calc_class_weight([(torch.zeros(5, 4), y[:5]), (torch.zeros(4, 4), y[5:])])
    assert w.shape[0] >= 8 and abs(float(w[0]) - 9 / (3 * 6)) < 1e-6 and abs(float(w[3]) - 9 / (3 * 2)) < 1e-6 \
        and abs(float(w[7]) - 9 / (3 * 1)) < 1e-6 and float(w[1]) == 0.0
    seen = []
    orig = nn.CrossEntropyLoss.__init__

    def spy(self, *args, weight=None, **kw):
        seen.append(None if weight is None else weight.clone())
        orig(self, *args, weight=weight, **kw)

    nn.CrossEntropyLoss.__init__ = spy
    try:
        api.train()
    finally:
        nn.CrossEntropyLoss.__init__ = orig
    weighted = [w_ for w_ in seen if w_ is not None]
    assert len(weighted) >= 2 and all(float(w_.sum()) > 0 for w_ in weighted)     # one weighted criterion per trained client
    assert getattr(a, "class_weight", None) is None                                # evaluation stays unweighted


def test_darts_second_order_architect_and_gdas():
    """DARTS unrolled (second-order) architecture gradient (reference model/cv/darts/architect.py) vs the exact
    hyper-gradient obtained by differentiating through the virtual weight step; GDAS space samples one op per edge and
    still trains its architecture parameters; FedNAS runs with both."""
    from torch.func import functional_call
    from fedml_b200.models.darts import Architect, Network, NetworkGDAS
    torch.manual_seed(0)
    net = Network(C=4, num_classes=4, layers=1, steps=2, multiplier=2).double()
    xt, yt = torch.randn(6, 3, 8, 8, dtype=torch.float64), torch.randint(0, 4, (6,))
    xv, yv = torch.randn(6, 3, 8, 8, dtype=torch.float64), torch.randint(0, 4, (6,))
    arch = Architect(net, momentum=0.0, weight_decay=0.0)
    eta = 0.05
    approx = arch.arch_gradient(xt, yt, xv, yv, eta, None, unrolled=True)
    first = arch.arch_gradient(xt, yt, xv, yv, eta, None, unrolled=False)
    # exact: d/da L_val(w - eta * grad_w L_train(w, a), a)
    wn, an = arch._names()
    named = dict(net.named_parameters())
    w = {n: named[n] for n in wn}; a = {n: named[n] for n in an}
    bufs = dict(net.named_buffers())
    crit = torch.nn.CrossEntropyLoss()
    ltr = crit(functional_call(net, {**w, **a, **bufs}, (xt,)), yt)
    gw = torch.autograd.grad(ltr, list(w.values()), create_graph=True, allow_unused=True)
    w2 = {n: (p - eta * g) if g is not None else p for (n, p), g in zip(w.items(), gw)}
    exact = torch.autograd.grad(crit(functional_call(net, {**w2, **a, **bufs}, (xv,)), yv), list(a.values()), allow_unused=True)
    exact = [torch.zeros_like(p) if g_ is None else g_ for g_, p in zip(exact, a.values())]
    fa, fe, ff = (torch.cat([g.reshape(-1) for g in gs]) for gs in (approx, exact, first))
    cos = float(torch.dot(fa, fe) / (fa.norm() * fe.norm()))
    assert cos > 0.98 and float((fa - fe).norm() / fe.norm()) < 0.2, cos
    assert float((fa - fe).norm()) < float((ff - fe).norm()) + 1e-12     # closer to the exact hyper-gradient than first order
    before = [p.clone() for p in net.arch_parameters()]
    arch.step(xt, yt, xv, yv, eta, None, unrolled=True)
    assert not torch.equal(before[0], net.arch_parameters()[0])
    # GDAS: hard one-hot edge weights in the forward pass, gradients reach the alphas, tau is adjustable
    g = NetworkGDAS(C=4, num_classes=4, layers=3, steps=2, multiplier=2, tau=5.0)      # one normal cell, two reduction cells
    g.train()
    wgt = g._weights(g.alphas_normal)
    assert torch.equal(wgt.sum(-1), torch.ones(wgt.shape[0])) and set(wgt.unique().tolist()) <= {0.0, 1.0}
    torch.nn.functional.cross_entropy(g(xt.float()), yt).backward()
    assert all(a_.grad is not None and float(a_.grad.abs().sum()) > 0 for a_ in g.arch_parameters())
    g.set_tau(1.0); assert g.get_tau() == 1.0
    g.eval(); assert g(xt.float()).shape == (6, 4)
    for kw in (dict(arch_order=2), dict(nas_space="gdas")):
        a_ = mk(federated_optimizer="FedNAS", backend="MPI", worker_num=2, client_num_per_round=2, client_num_in_total=2, comm_round=1,
                dataset="cifar10", layers=2, synthetic_train_size=32, batch_size=8, run_id="nas" + str(list(kw)[0]), ci=1,
                frequency_of_the_test=1, **kw)
        ds, model = build(a_)
        r = SimulatorMPI(a_, CPU, ds, model).run()
        assert len(r.history) >= 1 and r.genotype is not None


def test_nccl_simulator_secure_aggregation_matches_plain_fedavg():
    """secure_aggregation in the multi-GPU simulator (CPU path here): pairwise-masked quantised uploads sum to the same
    global model as plain FedAvg, up to the quantisation step."""
    from fedml_b200.simulation import SimulatorNCCL
    outs = []
    for sec in (False, True):
        a = mk(backend="NCCL", comm_round=2, client_num_per_round=4, frequency_of_the_test=0, secure_aggregation=sec, using_gpu=False)
        ds, model = build(a)
        torch.manual_seed(0)
        sim = SimulatorNCCL(a, CPU, ds, copy.deepcopy(model)).simulator
        for r in range(2):
            sim.train_one_round(r)
        outs.append(sim.global_flat.clone())
    assert float((outs[0] - outs[1]).abs().max()) < 4 * 4 / 2 ** 16 and not torch.equal(outs[0], outs[1])


def test_fednas_train_stage_and_derived_network():
    """After the search, FedNAS trains the derived network (reference model/cv/darts/model.py NetworkCIFAR, FedNASTrainer train
    stage): published cells build, the full DARTS-V2 network has the paper's 3.35 M parameters, and a federated round runs with
    the auxiliary head and drop path."""
    from fedml_b200.models.darts import DARTS_V1, DARTS_V2, FedNAS_V1, Network, NetworkCIFAR, count_parameters_in_MB
    assert abs(count_parameters_in_MB(NetworkCIFAR(36, 10, 20, True, DARTS_V2)) - 3.35) < 0.02
    for g in (DARTS_V1, FedNAS_V1, Network(C=8, layers=3).genotype()[0]):
        net = NetworkCIFAR(8, 10, 5, True, g)
        net.drop_path_prob = 0.3
        logits, aux = net.train()(torch.randn(3, 3, 32, 32))
        assert logits.shape == (3, 10) and aux.shape == (3, 10)
        assert net.eval()(torch.randn(2, 3, 32, 32))[1] is None
    a = mk(federated_optimizer="FedNAS", backend="MPI", worker_num=2, client_num_per_round=2, client_num_in_total=2, comm_round=2,
           synthetic_train_size=32, batch_size=8, run_id="nas-train", dataset="cifar10", stage="train", arch="FedNAS_V1", layers=3,
           init_channels=4, frequency_of_the_test=1)
    ds, model = build(a)
    r = SimulatorMPI(a, CPU, ds, model).run()
    assert len(r.history) == 2 and np.isfinite(r.history[-1]["test_loss"]) and isinstance(r.trainer.model, NetworkCIFAR)


def test_reference_style_mpi_launch_script(tmp_path):
    """The reference's launch idiom - ``comm, process_id, worker_number = FedML_init()`` then ``FedML_FedAvg_distributed(...)`` -
    run as three processes started by a launcher that exports RANK / WORLD_SIZE (no mpirun here: the star goes over TCP)."""
    import json
    import random
    import subprocess
    import sys
    script = tmp_path / "main_fedavg.py"
    script.write_text('''
import sys, json, torch
import fedml
from fedml.arguments import Arguments
from fedml.simulation.mpi_p2p_mp.fedavg.FedAvgAPI import FedML_init, FedML_FedAvg_distributed

comm, process_id, worker_number = FedML_init()
args = Arguments.from_dict(dict(dataset="mnist", model="lr", federated_optimizer="FedAvg", client_num_in_total=8, client_num_per_round=2,
                                comm_round=2, epochs=1, batch_size=16, learning_rate=0.05, frequency_of_the_test=1, synthetic_train_size=400,
                                mnist_client_num=8, backend="MPI", mp_transport="TCP", tcp_base_port=int(sys.argv[1]), random_seed=0,
                                log_level="WARNING", run_id="ref-launch", using_gpu=False))
dataset, class_num = fedml.data.load(args)
model = fedml.model.create(args, class_num)
out = FedML_FedAvg_distributed(args, process_id, worker_number, comm, torch.device("cpu"), dataset, model)
if process_id == 0:
    w = out.get_global_model_params()
    print("RESULT", json.dumps({"rounds": len(out.history) if hasattr(out, "history") else -1, "norm": float(sum(v.float().norm() for v in w.values()))}))
''')
    port = random.randint(20000, 45000)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), WORLD_SIZE="3")
    procs = [subprocess.Popen([sys.executable, str(script), str(port)], env=dict(env, RANK=str(r)), stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(3)]
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=240)[0])
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
    assert all(p.returncode == 0 for p in procs), "\n----\n".join(o[-1500:] for o in outs)
    line = [ln for ln in outs[0].splitlines() if ln.startswith("RESULT")]
    assert line and json.loads(line[0][7:])["norm"] > 0
