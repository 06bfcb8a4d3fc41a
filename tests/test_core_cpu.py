"""CPU tests: config, data contract, model zoo, partitioners, topology, scheduler, aggregation oracles, codec."""
import copy
import json
import os

import numpy as np
import pytest
import torch

import fedml
from fedml_b200.arguments import Arguments, SCHEMA, load_arguments, substitute_fedsys_vars
from fedml_b200.core import aggregation as A


def mk(**kw):
    base = dict(dataset="mnist", model="lr", client_num_per_round=4, comm_round=2, batch_size=10, learning_rate=0.03,
                frequency_of_the_test=1, synthetic_train_size=1500, mnist_client_num=12, log_level="WARNING")
    base.update(kw)
    return fedml.init(Arguments.from_dict(base))


def test_arguments_yaml_flattening_and_defaults(tmp_path):
    y = tmp_path / "c.yaml"
    y.write_text("common_args:\n  training_type: simulation\n  random_seed: 3\ntrain_args:\n  comm_round: '7'\n  my_custom_key: [1, 2]\n"
                 "comm_args:\n  backend: single_process\n")
    a = load_arguments(argv=["--cf", str(y), "--rank", "2", "--run_id", "abc"])
    assert a.random_seed == 3 and a.comm_round == 7 and a.my_custom_key == [1, 2] and a.rank == 2 and a.run_id == "abc"
    assert a.client_optimizer == SCHEMA["client_optimizer"][1]           # default filled in
    d = load_arguments("simulation", "single_process", argv=[])
    assert d.dataset == "mnist" and d.model == "lr" and d.client_num_per_round == 10   # packaged default yaml
    assert substitute_fedsys_vars("run ${FEDSYS.RUN_ID}", {"RUN_ID": 9}) == "run 9"


def test_hierarchical_overlay_yaml(tmp_path):
    (tmp_path / "server.yaml").write_text("x:\n  extra_key: server\n")
    (tmp_path / "silo1.yaml").write_text("x:\n  extra_key: silo1\n  n_proc_in_silo: 2\n")
    main = tmp_path / "m.yaml"
    main.write_text("common_args:\n  training_type: cross_silo\n  scenario: hierarchical\n"
                    f"  server_config_path: {tmp_path}/server.yaml\n  client_silo_config_paths: [{tmp_path}/silo1.yaml]\n")
    a0 = load_arguments("cross_silo", None, argv=["--cf", str(main), "--rank", "0"])
    a1 = load_arguments("cross_silo", None, argv=["--cf", str(main), "--rank", "1"])
    assert a0.extra_key == "server" and a1.extra_key == "silo1" and a1.n_proc_in_silo == 2


def test_data_contract_and_shapes():
    a = mk()
    ds, ncls = fedml.data.load(a)
    assert len(ds) == 8 and ncls == 10 and a.client_num_in_total == 12         # natural partition overrides the count
    train_num, test_num, tg, teg, nums, tl, tel, c = ds
    assert sum(nums.values()) == train_num
    x, y = tl[3][0]
    assert x.shape[1:] == (784,) and x.dtype == torch.float32 and y.dtype == torch.int64 and x.shape[0] <= 10
    assert len(set(torch.cat([b[1] for b in tl[3]]).tolist())) <= 2              # two digits per device
    a2 = mk(dataset="cifar10", model="resnet56", client_num_in_total=7, partition_method="hetero", partition_alpha=0.5,
            synthetic_train_size=700)
    ds2, n2 = fedml.data.load(a2)
    xb, yb = ds2[5][0][0]
    assert xb.shape == (10, 3, 32, 32) and xb.dtype == torch.float32 and len(ds2[5]) == 7
    assert abs(float(xb.mean())) < 3.0
    fed = a2._federated_dataset
    assert fed.x.dtype == torch.uint8 and fed.client_sizes().sum() == 700 and fed.test_offsets is None


def test_centralized_and_full_batch_switches():
    a = mk(client_num_in_total=1)
    ds, _ = fedml.data.load(a)
    assert list(ds[5].keys()) == [0] and a.client_num_in_total == 1
    from fedml_b200.data import combine_batches
    merged = combine_batches(ds[5][0])
    assert len(merged) == 1 and merged[0][0].shape[0] == ds[4][0]


@pytest.mark.parametrize("name,dataset,params,entries", [
    ("lr", "mnist", 7850, 2), ("cnn", "femnist", 1206590, 8), ("resnet56", "cifar10", 591322, 350),
    ("rnn", "shakespeare", 822570, 11), ("rnn", "stackoverflow_nwp", 4053428, 9)])
def test_model_zoo_sizes_match_reference(name, dataset, params, entries):
    out = {"mnist": 10, "femnist": 62, "cifar10": 10, "shakespeare": 90, "stackoverflow_nwp": 10004}[dataset]
    m = fedml.model.create(Arguments.from_dict(dict(model=name, dataset=dataset)), out)
    assert sum(p.numel() for p in m.parameters()) == params and len(m.state_dict()) == entries


def test_resnet_state_dict_keys_are_reference_names_and_distilbert_size():
    sd = fedml.model.create(Arguments.from_dict(dict(model="resnet56", dataset="cifar10")), 10).state_dict()
    for k in ("conv1.weight", "bn1.running_mean", "layer1.0.conv1.weight", "layer1.0.downsample.0.weight",
              "layer3.5.bn3.num_batches_tracked", "fc.bias"):
        assert k in sd
    assert sd["bn1.num_batches_tracked"].dtype == torch.int64
    r18 = fedml.model.create(Arguments.from_dict(dict(model="resnet18", dataset="imagenet")), 10)
    assert sum(p.numel() for p in r18.parameters()) == 11181642 and len(r18.state_dict()) == 122
    from fedml_b200.models import distilbert
    d = distilbert(20)
    assert sum(p.numel() for p in d.parameters()) == 66968852 and len(d.state_dict()) == 104
    assert "distilbert.transformer.layer.5.ffn.lin2.weight" in d.state_dict()


def test_lr_keeps_sigmoid_before_cross_entropy():
    m = fedml.model.create(Arguments.from_dict(dict(model="lr", dataset="mnist")), 10)
    out = m(torch.randn(4, 784) * 100)
    assert float(out.min()) >= 0 and float(out.max()) <= 1


def test_dirichlet_partition_matches_reference_procedure():
    from fedml_b200.core.non_iid_partition import non_iid_partition_with_dirichlet_distribution, record_data_stats
    labels = np.random.RandomState(0).randint(0, 10, 5000)
    np.random.seed(10)
    part = non_iid_partition_with_dirichlet_distribution(labels, 8, 10, 0.5)
    sizes = [len(v) for v in part.values()]
    assert sum(sizes) == 5000 and min(sizes) >= 10 and len(set(sum(map(list, part.values()), []))) == 5000
    np.random.seed(10)
    again = non_iid_partition_with_dirichlet_distribution(labels, 8, 10, 0.5)
    assert all(list(part[i]) == list(again[i]) for i in range(8))               # same seed, same split
    stats = record_data_stats(labels, part)
    assert sum(sum(s.values()) for s in stats.values()) == 5000


def test_topology_managers_row_stochastic_without_networkx():
    from fedml_b200.core.distributed.topology import AsymmetricTopologyManager, SymmetricTopologyManager
    t = SymmetricTopologyManager(8, 4); t.generate_topology()
    W = np.asarray(t.topology)
    assert np.allclose(W.sum(1), 1) and np.allclose(W, W.T) is not None and (np.diag(W) > 0).all()
    assert sorted(t.get_in_neighbor_idx_list(0)) == [1, 2, 6, 7]
    np.random.seed(0)
    a = AsymmetricTopologyManager(8, 2, 2); a.generate_topology()
    assert np.allclose(np.asarray(a.topology).sum(1), 1)
    assert set(a.get_out_neighbor_idx_list(3)) >= {2, 4}


def test_scheduler_reference_example_and_lpt():
    from fedml_b200.core.schedule import scheduler, schedule_clients
    s = scheduler(np.array([1, 2, 3, 5, 7, 14]), np.array([1, 5]), np.array([15, 100]))
    plan = s.DP_schedule(1)
    assigned = sorted(j for res in plan for bunch in res.values() for j in bunch)
    assert assigned == [0, 1, 2, 3, 4, 5]
    big = scheduler(np.arange(1, 101), np.ones(8), np.full(8, 1e9))            # reference recursion would blow the stack
    plan = big.DP_schedule(0)
    loads = [sum(np.arange(1, 101)[j] for b in res.values() for j in b) for res in plan]
    assert max(loads) - min(loads) <= 100 and sum(loads) == 5050
    groups = schedule_clients([50] * 100, 8)
    assert sorted(len(g) for g in groups) == [12, 12, 12, 12, 13, 13, 13, 13]


def _sd(seed, bn=True):
    torch.manual_seed(seed)
    net = torch.nn.Sequential(torch.nn.Linear(6, 4), torch.nn.BatchNorm1d(4) if bn else torch.nn.Identity())
    sd = net.state_dict()
    if bn:
        sd["1.num_batches_tracked"] = torch.tensor(seed + 1)
        sd["1.running_mean"] = torch.randn(4)
    return sd


def test_fedavg_aggregate_weights_every_entry_including_int_counters():
    locs = [(10, _sd(0)), (30, _sd(1))]
    agg = A.fedavg_aggregate(locs)
    for k in locs[0][1]:
        ref = (10 * locs[0][1][k].float() + 30 * locs[1][1][k].float()) / 40
        torch.testing.assert_close(agg[k].float(), ref)
    back = A.cast_like(agg, locs[0][1])
    assert back["1.num_batches_tracked"].dtype == torch.int64


@pytest.mark.parametrize("name", ["sgd", "adam", "adagrad"])
def test_server_optimizer_matches_torch_optim(name):
    torch.manual_seed(0)
    w = {"p": torch.randn(20)}
    opt = A.ServerOptimizer(name, lr=0.1, momentum=0.9 if name == "sgd" else 0.0, beta1=0.9, beta2=0.99, eps=1e-3)
    p = torch.nn.Parameter(w["p"].clone())
    topt = {"sgd": lambda: torch.optim.SGD([p], lr=0.1, momentum=0.9), "adam": lambda: torch.optim.Adam([p], lr=0.1, betas=(0.9, 0.99), eps=1e-3),
            "adagrad": lambda: torch.optim.Adagrad([p], lr=0.1, eps=1e-3)}[name]()
    for _ in range(3):
        avg = {"p": w["p"] - 0.1 * torch.randn(20)}
        p.grad = (p.detach() - avg["p"]).clone()             # pseudo-gradient of the reference (fedopt_api.py:173-186)
        topt.step()
        w = opt.step(w, avg, ["p"])
        torch.testing.assert_close(w["p"], p.detach(), rtol=1e-5, atol=1e-6)


def test_robust_aggregators_clip_median_trimmed_krum():
    g = _sd(0, bn=False)
    honest = [(10, {k: v + 0.01 * torch.randn_like(v) for k, v in g.items()}) for _ in range(8)]
    bad = [(10, {k: v + 50.0 for k, v in g.items()}) for _ in range(2)]
    allc = bad + honest
    clipped = A.norm_diff_clip(bad[0][1], g, 1.0)
    assert abs(float(torch.linalg.vector_norm(A.vectorize_weight(clipped) - A.vectorize_weight(g))) - 1.0) < 1e-4
    for agg in (A.coordinate_median(allc), A.trimmed_mean(allc, 0.2), A.krum(allc, f=2, m=1), A.krum(allc, f=2, m=4)):
        assert float((A.vectorize_weight(agg) - A.vectorize_weight(g)).abs().max()) < 0.1
    assert float((A.vectorize_weight(A.fedavg_aggregate(allc)) - A.vectorize_weight(g)).abs().max()) > 5   # plain mean is poisoned
    even = [(1, {"w": torch.tensor([float(i)])}) for i in (1, 2, 3, 4)]
    assert float(A.coordinate_median(even)["w"]) == 2.0            # lower median, like torch.median
    from fedml_b200.core.robustness import RobustAggregator
    ra = RobustAggregator(Arguments.from_dict(dict(defense_type="weak_dp", norm_bound=1.0, stddev=0.0)))
    out = ra.defend(allc, g)
    assert float((A.vectorize_weight(out) - A.vectorize_weight(g)).abs().max()) < 1.0


def test_message_binary_codec_and_json():
    from fedml_b200.core.distributed.communication.message import Message
    m = Message(3, 1, 0)
    m.add_params("model_params", {"a": torch.randn(3, 4), "n": torch.tensor(7), "h": torch.randn(2).to(torch.bfloat16)})
    m.add_params("arr", np.arange(5)); m.add_params("nested", [1, (2, 3), {"k": b"raw"}])
    m2 = Message.from_bytes(m.to_bytes())
    assert m2.get_type() == 3 and m2.get_sender_id() == 1
    assert torch.equal(m2.get("model_params")["a"], m.get("model_params")["a"]) and m2.get("model_params")["h"].dtype == torch.bfloat16
    assert list(m2.get("arr")) == [0, 1, 2, 3, 4] and m2.get("nested")[1] == (2, 3) and m2.get("nested")[2]["k"] == b"raw"
    m3 = Message(); m3.init_from_json_string(m.to_json())
    assert m3.get("model_params")["a"][0][0] == pytest.approx(float(m.get("model_params")["a"][0, 0]))


def test_arena_roundtrip_and_layout_regions():
    from fedml_b200.parallel.arena import ArenaLayout, bind_module_to_flat
    m = fedml.model.create(Arguments.from_dict(dict(model="resnet56", dataset="cifar10")), 10)
    lay = ArenaLayout(m)
    assert lay.P % 4 == 0 and lay.P_train % 4 == 0 and lay.P_train >= 591322
    flat = lay.flatten(m.state_dict())
    sd = lay.unflatten(flat)
    for k, v in m.state_dict().items():
        assert torch.equal(sd[k], v) and sd[k].dtype == v.dtype
    names = [s.name for s in lay.segments if s.offset < lay.P_train]
    assert all(A.is_weight_param(n) for n in names)
    small = torch.nn.Sequential(torch.nn.Linear(5, 3), torch.nn.BatchNorm1d(3))
    lay2 = ArenaLayout(small); f2 = torch.zeros(lay2.P); g2 = torch.zeros(lay2.P)
    bind_module_to_flat(small, lay2, f2, g2)
    small(torch.randn(4, 5)).sum().backward()
    assert float(g2.abs().sum()) > 0 and small[0].weight.data_ptr() == f2.data_ptr()


def test_checkpoint_format_and_resume(tmp_path):
    from fedml_b200.utils.checkpoint import load_state_dict, save_checkpoint, save_state_dict
    m = fedml.model.create(Arguments.from_dict(dict(model="resnet56", dataset="cifar10")), 10)
    p = str(tmp_path / "global_model.pt")
    save_state_dict(p, m.state_dict())
    raw = torch.load(p, weights_only=False)
    assert list(raw.keys()) == list(m.state_dict().keys())                      # plain OrderedDict, reference key names
    wrapped = str(tmp_path / "ck"); save_checkpoint({"state_dict": {"module." + k: v for k, v in m.state_dict().items()},
                                                      "optim_dict": {}, "epoch": 3, "test_accTop1": 0.5, "test_accTop5": 0.9}, True, wrapped)
    sd = load_state_dict(os.path.join(wrapped, "best.pth"))
    m.load_state_dict(sd)
    from fedml.simulation import SimulatorSingleProcess
    a = mk(federated_optimizer="FedOpt", server_optimizer="adam", server_lr=0.01, comm_round=4, checkpoint_dir=str(tmp_path / "run"),
           checkpoint_every=2, eval_client_cap=2)
    ds, out = fedml.data.load(a)
    init = fedml.model.create(a, out)
    full = SimulatorSingleProcess(a, torch.device("cpu"), ds, copy.deepcopy(init)); w_full = full.run()
    a2 = copy.copy(a); a2.comm_round = 2
    SimulatorSingleProcess(a2, torch.device("cpu"), ds, copy.deepcopy(init)).run()      # "crashes" after round 2
    a3 = copy.copy(a); a3.comm_round = 4
    res = SimulatorSingleProcess(a3, torch.device("cpu"), ds, fedml.model.create(a, out))   # fresh process, fresh init
    assert res.fl_trainer.resume(str(tmp_path / "run"))
    w_res = res.run()
    for k in w_full:
        torch.testing.assert_close(w_res[k].float(), w_full[k].float(), rtol=1e-5, atol=1e-6)   # server-opt moments + round idx restored


def test_train_time_augmentation_crop_flip_cutout():
    """RandomCrop(32, 4) + HFlip + Cutout(16) as a pure function of (seed, nonce, sample) — reference
    data/cifar10/data_loader.py:58-103 — and its hook into the train loaders."""
    import torch
    import fedml_b200 as fedml
    from fedml_b200.arguments import Arguments
    from fedml_b200.data.augment import CIFAR_MEAN_U8, augment_cifar_reference, augment_params
    g = torch.Generator().manual_seed(0)
    x = torch.randint(1, 255, (64, 32, 32, 3), dtype=torch.uint8, generator=g)
    a = augment_cifar_reference(x, seed=11, nonce=3)
    assert a.shape == x.shape and a.dtype == torch.uint8
    assert torch.equal(a, augment_cifar_reference(x, 11, 3)) and not torch.equal(a, augment_cifar_reference(x, 11, 4))
    dy, dx, flip, cy, cx = augment_params(64, 11, 3, 32, 32)
    assert dy.min() >= 0 and dy.max() <= 8 and dx.min() >= 0 and dx.max() <= 8 and set(flip.tolist()) == {0, 1}
    mean = torch.tensor(CIFAR_MEAN_U8, dtype=torch.uint8)
    for i in range(64):
        y1, y2, x1, x2 = max(0, cy[i] - 8), min(32, cy[i] + 8), max(0, cx[i] - 8), min(32, cx[i] + 8)
        assert (a[i, y1:y2, x1:x2] == mean).all()                       # Cutout square = dataset mean (0 after Normalize)
        # a pixel outside the square and inside the un-padded source area is the shifted (and mirrored) source pixel
        for oy, ox in ((0, 0), (31, 31), (5, 20), (16, 16)):
            if y1 <= oy < y2 and x1 <= ox < x2:
                continue
            sxo = 31 - ox if flip[i] else ox
            sy, sx = oy + dy[i] - 4, sxo + dx[i] - 4
            want = x[i, sy, sx] if (0 <= sy < 32 and 0 <= sx < 32) else torch.zeros(3, dtype=torch.uint8)
            assert torch.equal(a[i, oy, ox], want)
    # no augmentation parameters -> identity
    assert torch.equal(augment_cifar_reference(x, 1, 1, pad=0, cutout=0)[flip_none(x, 1, 1)], x[flip_none(x, 1, 1)])
    # loaders: train batches of a client differ between visits, test batches do not, default is off
    args = fedml.init(Arguments.from_dict(dict(training_type="simulation", backend="sp", dataset="cifar10", model="resnet56",
                                               client_num_in_total=4, client_num_per_round=2, comm_round=1, batch_size=8,
                                               synthetic_train_size=64, augment=True, log_level="ERROR", using_gpu=False)))
    ds, _ = fedml.data.load(args)
    b1, b2 = ds[5][0][0][0], ds[5][0][0][0]
    assert b1.shape == b2.shape and not torch.equal(b1, b2)
    assert torch.equal(ds[6][0][0][0], ds[6][0][0][0])
    args2 = fedml.init(Arguments.from_dict(dict(training_type="simulation", backend="sp", dataset="cifar10", model="resnet56",
                                                client_num_in_total=4, client_num_per_round=2, comm_round=1, batch_size=8,
                                                synthetic_train_size=64, log_level="ERROR", using_gpu=False)))
    ds2, _ = fedml.data.load(args2)
    assert torch.equal(ds2[5][0][0][0], ds2[5][0][0][0])


def flip_none(x, seed, nonce):
    """indices of the samples that are not mirrored for (seed, nonce)"""
    import torch
    from fedml_b200.data.augment import augment_params
    return torch.from_numpy(augment_params(x.shape[0], seed, nonce, 32, 32, pad=0)[2] == 0)


def test_real_dataset_readers_from_files(tmp_path):
    """LEAF JSON (FEMNIST, Shakespeare), CIFAR-100 pickles, image folders (ImageNet layout), Landmarks CSV: written here in
    the reference's on-disk formats, read back through fedml.data.load; a corrupt .h5 export is reported clearly."""
    import json
    import pickle
    import numpy as np
    import pytest
    import torch
    from PIL import Image
    import fedml_b200 as fedml
    from fedml_b200.arguments import Arguments
    from fedml_b200.data import readers as R
    rng = np.random.default_rng(0)

    def load(**kw):
        base = dict(training_type="simulation", backend="sp", model="lr", client_num_in_total=3, client_num_per_round=2,
                    comm_round=1, batch_size=4, log_level="ERROR", using_gpu=False, data_cache_dir=str(tmp_path))
        base.update(kw)
        a = fedml.init(Arguments.from_dict(base))
        return fedml.data.load(a), a
    # ---- FEMNIST, LEAF JSON
    for split, n in (("train", 6), ("test", 2)):
        d = tmp_path / "femnist" / split; d.mkdir(parents=True)
        blob = {"users": ["u0", "u1"], "user_data": {u: {"x": rng.random((n, 784)).tolist(), "y": rng.integers(0, 62, n).tolist()}
                                                      for u in ("u0", "u1")}}
        (d / "all_data_0.json").write_text(json.dumps(blob))
    (ds, ncls), a = load(dataset="femnist")
    assert ncls == 62 and a.client_num_in_total == 2 and ds[0] == 12 and ds[1] == 4 and a._federated_dataset.x.shape == (12, 28, 28)
    # ---- Shakespeare, LEAF JSON: 80-character windows -> indices in the reference vocabulary
    text = "the quality of mercy is not strained, it droppeth as the gentle rain from heaven upon the place beneath. " * 2
    for split, n in (("train", 5), ("test", 2)):
        d = tmp_path / "shakespeare" / split; d.mkdir(parents=True)
        blob = {"users": ["PORTIA"], "user_data": {"PORTIA": {"x": [text[i:i + 80] for i in range(n)], "y": [text[i + 80] for i in range(n)]}}}
        (d / "all_data.json").write_text(json.dumps(blob))
    (ds, ncls), a = load(dataset="shakespeare", model="rnn")
    fed = a._federated_dataset
    assert fed.x.shape == (5, 80) and ncls == R.CHAR_VOCAB_SIZE and int(fed.x[0, 0]) == R.ALL_LETTERS.find("t") \
        and int(fed.y[0]) == R.ALL_LETTERS.find(text[80])
    # ---- CIFAR-100 pickles
    d = tmp_path / "cifar-100-python"; d.mkdir()
    for name, n in (("train", 40), ("test", 10)):
        with open(d / name, "wb") as fh:
            pickle.dump({b"data": rng.integers(0, 256, (n, 3072), dtype=np.uint8), b"fine_labels": rng.integers(0, 100, n).tolist()}, fh)
    (ds, ncls), a = load(dataset="cifar100", model="resnet56", partition_method="homo")
    assert ncls == 100 and ds[0] == 40 and ds[1] == 10 and a._federated_dataset.x.dtype == torch.uint8
    # ---- ImageNet folder layout
    for split, n in (("train", 3), ("val", 1)):
        for c in ("n01", "n02"):
            d = tmp_path / split / c; d.mkdir(parents=True)
            for i in range(n):
                Image.fromarray(rng.integers(0, 256, (20, 24, 3), dtype=np.uint8)).save(d / f"{c}_{i}.png")
    (ds, ncls), a = load(dataset="imagenet", model="resnet56", image_size=16, partition_method="homo", client_num_in_total=2)
    assert ncls == 2 and ds[0] == 6 and ds[1] == 2 and a._federated_dataset.x.shape == (6, 16, 16, 3)
    # ---- Landmarks CSV
    (tmp_path / "images").mkdir()
    rows_tr, rows_te = [], []
    for i in range(6):
        Image.fromarray(rng.integers(0, 256, (18, 18, 3), dtype=np.uint8)).save(tmp_path / "images" / f"img{i}.jpg")
        (rows_tr if i < 4 else rows_te).append((str(i % 2), f"img{i}", str(i % 3)))
    for nm, rows in (("gld23k_user_dict_train.csv", rows_tr), ("gld23k_user_dict_test.csv", rows_te)):
        (tmp_path / nm).write_text("user_id,image_id,class\n" + "\n".join(",".join(r) for r in rows) + "\n")
    (ds, ncls), a = load(dataset="gld23k", model="resnet56", image_size=16)
    assert a.client_num_in_total == 2 and ds[0] == 4 and ds[1] == 2 and ncls == 3
    # ---- TFF h5 exports are read without h5py (tests/test_minihdf5.py); a file that is not HDF5 is reported as such
    (tmp_path / "fed_cifar100_train.h5").write_bytes(b"not really hdf5")
    (tmp_path / "fed_cifar100_test.h5").write_bytes(b"not really hdf5")
    with pytest.raises(Exception, match="(?i)hdf5|signature"):
        load(dataset="fed_cifar100", model="resnet18_gn")


def test_synchronized_batchnorm_equals_batchnorm_on_the_global_batch():
    """SyncBN over a 2-process gloo group == nn.BatchNorm2d on the concatenated batch: outputs, input / parameter
    gradients, running statistics, state_dict keys (reference model/cv/batchnorm_utils.py:156-266)."""
    import os
    import subprocess
    import sys
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "syncbn_worker.py")
    procs = [subprocess.Popen([sys.executable, worker], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                              env=dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port)))
             for r in range(2)]
    outs = [p.communicate(timeout=180)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs) and all(f"SYNCBN_OK {r}" in outs[r] for r in range(2)), outs
    # without a process group it is plain BatchNorm
    import torch
    import torch.nn as nn
    from fedml_b200.models.sync_batchnorm import SynchronizedBatchNorm2d
    torch.manual_seed(0)
    a, b = nn.BatchNorm2d(5), SynchronizedBatchNorm2d(5)
    x = torch.randn(4, 5, 3, 3)
    torch.testing.assert_close(b(x), a(x), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(b.running_var, a.running_var, rtol=1e-5, atol=1e-6)


def test_dirichlet_partition_terminates_with_fewer_samples_than_clients():
    """64 samples over 500 clients: the reference's redraw-until-min-size loop would never end; every sample is still assigned."""
    from fedml_b200.core.non_iid_partition import non_iid_partition_with_dirichlet_distribution
    np.random.seed(0)
    y = np.random.randint(0, 10, 64)
    part = non_iid_partition_with_dirichlet_distribution(y, 500, 10, 0.1, min_required=10)
    assert sorted(i for v in part.values() for i in v) == list(range(64))


def test_model_factory_names_of_the_nas_and_gkt_algorithms():
    import fedml
    from fedml_b200.arguments import Arguments
    mk = lambda m: fedml.model.create(Arguments.from_dict(dict(model=m, dataset="cifar10", log_level="ERROR")), 10)      # noqa: E731
    assert type(mk("darts")).__name__ == "Network" and type(mk("darts_gdas")).__name__ == "NetworkGDAS"
    client, server = mk("resnet56_gkt")
    logits, feat = client(torch.randn(2, 3, 32, 32))
    assert logits.shape == (2, 10) and server(feat).shape == (2, 10)
    assert type(mk("no_such_model")).__name__ == "LogisticRegression"          # the reference's fallback (model_hub.py, last branch)


def test_algorithm_helper_functions(tmp_path):
    """Per-algorithm utils of the reference, through its import paths."""
    from functools import partial
    from fedml.simulation.mpi_p2p_mp.fedgkt.utils import (RunningAverage, bnwd_optim_params, get_flat_params_from, get_state_dict,
                                                            save_dict_to_json, set_flat_params_to)
    from fedml.simulation.mpi_p2p_mp.fedseg.utils import (EvaluationMetricsKeeper, count_parameters, load_from_pickle_file,
                                                            save_as_pickle_file, str_to_bool)
    from fedml.simulation.single_process.fednova.comm_helpers import communicate, flatten_tensors, unflatten_tensors
    from fedml.simulation.single_process.classical_vertical_fl.party_models import sigmoid
    from fedml.simulation.single_process.classical_vertical_fl.vfl_fixture import compute_correct_prediction
    net = torch.nn.Sequential(torch.nn.Conv2d(3, 4, 3), torch.nn.BatchNorm2d(4), torch.nn.Flatten(), torch.nn.Linear(4, 2))
    flat = get_flat_params_from(net)
    assert flat.numel() == count_parameters(net) == 3 * 4 * 9 + 4 + 4 + 4 + 4 * 2 + 2
    set_flat_params_to(net, torch.arange(flat.numel(), dtype=torch.float32))
    assert float(net[0].weight.flatten()[5]) == 5.0 and float(net[3].bias[-1]) == flat.numel() - 1
    with pytest.raises(ValueError):
        set_flat_params_to(net, torch.zeros(3))
    params = list(net.parameters())
    groups = bnwd_optim_params(net, params, params)
    assert groups[0]["weight_decay"] == 0 and len(groups[0]["params"]) == 2 and len(groups[1]["params"]) == 4
    torch.optim.SGD(groups, lr=0.1, weight_decay=1e-3)
    avg = RunningAverage()
    avg.update(2)
    avg.update(4)
    assert avg.value() == 3 and avg() == 3
    torch.save({"state_dict": net.state_dict()}, tmp_path / "ck.pth")
    assert "state_dict" in get_state_dict(str(tmp_path / "ck.pth"))
    save_dict_to_json({"acc": np.float32(0.5), "n": 3}, str(tmp_path / "m.json"))
    assert json.load(open(tmp_path / "m.json")) == {"acc": 0.5, "n": 3}
    save_as_pickle_file(str(tmp_path / "x.pkl"), {"a": [1, 2]})
    assert load_from_pickle_file(str(tmp_path / "x.pkl")) == {"a": [1, 2]}
    assert str_to_bool("True") is True and str_to_bool("false") is False
    with pytest.raises(ValueError):
        str_to_bool("maybe")
    k = EvaluationMetricsKeeper(0.9, 0.8, 0.7, 0.6, 0.1)
    assert (k.acc, k.acc_class, k.mIoU, k.FWIoU, k.loss) == (0.9, 0.8, 0.7, 0.6, 0.1)
    a, b = torch.ones(2, 3), torch.full((4,), 2.0)
    f = flatten_tensors([a, b])
    assert f.tolist() == [1.0] * 6 + [2.0] * 4 and [t.shape for t in unflatten_tensors(f, [a, b])] == [a.shape, b.shape]

    def double(tensor):
        tensor.mul_(2)

    communicate([a, b], partial(double))
    assert float(a.sum()) == 12.0 and float(b.sum()) == 16.0
    assert abs(float(sigmoid(0.0)) - 0.5) < 1e-12
    y_hat, stats = compute_correct_prediction(y_targets=[1, 0, 1, 1], y_prob_preds=[0.9, 0.2, 0.4, 0.7])
    assert y_hat.tolist() == [1, 0, 0, 1] and stats == [2, 2, 3]


def test_replica_synchronized_batchnorm_threads_equal_batchnorm_on_the_whole_batch():
    """The DataParallel flavour of SyncBN (reference model/cv/batchnorm_utils.py: SyncMaster / SlavePipe / FutureResult +
    replication callbacks): three replica threads, each with a slice of the batch, normalise like one BatchNorm on all of it."""
    import threading
    from fedml_b200.models.sync_batchnorm import (FutureResult, ReplicaSynchronizedBatchNorm, SyncMaster, execute_replication_callbacks)
    # the message pipe on its own
    master = SyncMaster(lambda msgs: [(i, sum(m for _, m in msgs) * (i + 1)) for i, _ in sorted(msgs)])
    pipes = [master.register_slave(i) for i in (1, 2)]
    got = {}
    ts = [threading.Thread(target=lambda p=p: got.__setitem__(p.identifier, p.run_slave(10 * p.identifier))) for p in pipes]
    [t.start() for t in ts]
    got[0] = master.run_master(5)
    [t.join() for t in ts]
    assert got == {0: 35, 1: 70, 2: 105} and master.nr_slaves == 2
    f = FutureResult()
    threading.Timer(0.05, f.put, args=("late",)).start()
    assert f.get() == "late"
    # the layer
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Conv2d(3, 4, 3, padding=1), ReplicaSynchronizedBatchNorm(4), torch.nn.ReLU()).double()
    ref = torch.nn.Sequential(torch.nn.Conv2d(3, 4, 3, padding=1), torch.nn.BatchNorm2d(4), torch.nn.ReLU()).double()
    ref.load_state_dict(net.state_dict())
    x = torch.randn(9, 3, 6, 6, dtype=torch.float64)
    parts = [x[:2], x[2:5], x[5:]]
    replicas = []
    for _ in parts:                                     # what DataParallel.replicate does: new module objects, shared parameters
        rep = torch.nn.Sequential(*[copy.copy(m) for m in net])
        replicas.append(rep)
    for rep in replicas[1:]:                            # buffers of replica 0 are the master's own
        rep[1].running_mean, rep[1].running_var = rep[1].running_mean.clone(), rep[1].running_var.clone()
    replicas[0][1].running_mean, replicas[0][1].running_var = net[1].running_mean, net[1].running_var
    execute_replication_callbacks(replicas)
    outs = [None] * 3

    def run(i):
        outs[i] = replicas[i](parts[i])

    ts = [threading.Thread(target=run, args=(i,)) for i in range(3)]
    [t.start() for t in ts]
    [t.join(30) for t in ts]
    assert all(o is not None for o in outs)
    want = ref(x)
    torch.testing.assert_close(torch.cat(outs), want, rtol=1e-10, atol=1e-12)
    torch.cat(outs).square().sum().backward()
    want.square().sum().backward()
    for (n1, p1), (_, p2) in zip(net.named_parameters(), ref.named_parameters()):
        torch.testing.assert_close(p1.grad, p2.grad, rtol=1e-8, atol=1e-10, msg=n1)
    torch.testing.assert_close(net[1].running_mean, ref[1].running_mean, rtol=1e-10, atol=1e-12)
    torch.testing.assert_close(net[1].running_var, ref[1].running_var, rtol=1e-10, atol=1e-12)
    # outside replication / in eval mode: plain BatchNorm
    net.eval(); ref.eval()
    torch.testing.assert_close(net(x), ref(x), rtol=1e-10, atol=1e-12)


def test_gpu_util_map_from_a_silo_yaml(tmp_path):
    from fedml.device.gpu_mapping_cross_silo import get_gpu_util_map
    (tmp_path / "m.yaml").write_text("node1: [2, 0, 1]\nnode2: [1]\n")
    assert get_gpu_util_map(str(tmp_path / "m.yaml")) == {0: ("node1", 0), 1: ("node1", 0), 2: ("node1", 2), 3: ("node2", 0)}
