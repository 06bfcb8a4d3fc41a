"""CPU tests: transports, managers, cross-silo / cross-device over every backend, secure aggregation, MLOps, CLI."""
import json
import os
import shutil
import threading
import time

import numpy as np
import pytest
import torch

import fedml
from fedml_b200.arguments import Arguments
from fedml_b200.core.distributed.communication import create_comm_manager
from fedml_b200.core.distributed.communication.broker import BrokerClient, LocalBroker
from fedml_b200.core.distributed.communication.inproc import InProcHub
from fedml_b200.core.distributed.communication.message import Message
from fedml_b200.core.distributed.communication.observer import Observer

PORT = [53000]


def next_port(n=10):
    PORT[0] += n
    return PORT[0]


@pytest.fixture(scope="module")
def broker():
    b = LocalBroker()
    port = b.start()
    yield port
    b.stop()


class _Collect(Observer):
    def __init__(self):
        self.got = []
        self.ev = threading.Event()

    def receive_message(self, t, p):
        self.got.append((t, p))
        if t == 9:
            self.ev.set()


@pytest.mark.parametrize("backend", ["INPROC", "TCP", "GRPC", "MQTT", "MQTT_BIN", "MQTT_MIXED", "MQTT_S3"])
def test_transport_roundtrip_with_tensor_payload(backend, broker, tmp_path):
    kind, backend = backend, backend.split("_")[0] if backend.startswith("MQTT_") and backend != "MQTT_S3" else backend
    base = dict(run_id=f"rt-{kind}", tcp_base_port=next_port(), grpc_base_port=next_port(), mqtt_port=broker,
                object_store_dir=str(tmp_path / "store"), client_id_list="[1]")
    args = Arguments.from_dict(dict(base, mqtt_binary=kind in ("MQTT_BIN", "MQTT_MIXED")))
    InProcHub.reset(args.run_id)
    server = create_comm_manager(args, 0, 2, backend)                                  # MQTT_MIXED: binary server, JSON client
    client = create_comm_manager(Arguments.from_dict(dict(base, mqtt_binary=kind == "MQTT_BIN")), 1, 2, backend)
    cs, cc = _Collect(), _Collect()
    server.add_observer(cs); client.add_observer(cc)
    ts = [threading.Thread(target=m.handle_receive_message, daemon=True) for m in (server, client)]
    [t.start() for t in ts]
    payload = {"w": torch.arange(12, dtype=torch.float32).view(3, 4), "n": torch.tensor(3)}
    m = Message(9, 1, 0); m.add_params("model_params", payload); m.add_params("num_samples", 17)
    client.send_message(m)
    assert cs.ev.wait(20)
    t, p = [g for g in cs.got if g[0] == 9][0]
    assert p["num_samples"] == 17 and torch.equal(torch.as_tensor(p["model_params"]["w"]).float(), payload["w"])
    back = Message(9, 0, 1); back.add_params("model_params", payload)
    server.send_message(back)
    assert cc.ev.wait(20)
    server.stop_receive_message(); client.stop_receive_message()


def test_broker_last_will_and_status_manager(broker):
    from fedml_b200.core.distributed.communication.mqtt_backends import MqttS3StatusManager
    got = []
    watcher = BrokerClient("127.0.0.1", broker); watcher.on_message = lambda t, p: got.append((t, json.loads(p))); watcher.subscribe("W/topic")
    dying = BrokerClient("127.0.0.1", broker, will_topic="W/topic", will_payload={"ID": 7, "stat": "Offline"})
    dying._sock.shutdown(2)                      # abrupt death, no "bye"
    deadline = time.time() + 5
    while not got and time.time() < deadline:
        time.sleep(0.02)
    assert got and got[0][1]["ID"] == 7
    sm = MqttS3StatusManager(Arguments.from_dict(dict(mqtt_port=broker)))
    seen = []
    class O:
        def receive_message(self, t, p): seen.append((t, p))
    sm.add_observer(O()); sm.subscribe("fl_client/mlops/status"); sm.send_message_json("fl_client/mlops/status", {"edge_id": 1, "status": "TRAINING"})
    deadline = time.time() + 5
    while not seen and time.time() < deadline:
        time.sleep(0.02)
    assert seen and json.loads(seen[0][1])["status"] == "TRAINING"
    sm.stop()


_INIT_LOCK = threading.Lock()


def _run_cross_silo(backend, broker, tmp_path, **kw):
    base = dict(training_type="cross_silo", dataset="mnist", model="lr", client_num_in_total=2, client_num_per_round=2, client_id_list="[1, 2]",
                comm_round=2, batch_size=10, learning_rate=0.03, frequency_of_the_test=1, synthetic_train_size=1200, mnist_client_num=2,
                backend=backend, mqtt_port=broker, log_level="WARNING", run_id=f"cs-{backend}-{kw.get('tag', '')}", grpc_base_port=next_port(),
                tcp_base_port=next_port(), object_store_dir=str(tmp_path / "store"), using_mlops=True,
                global_model_file_path=str(tmp_path / "global_model.pt"))
    base.update(kw)
    InProcHub.reset(base["run_id"])
    res, errs = {}, []

    def role(rank):
        try:
            # The three roles are threads of ONE process here, so they share torch's global RNG (in a deployment each is
            # its own process).  fedml.init() seeds it and model construction draws from it: without the lock another
            # role's seeding / drawing can land between this role's manual_seed(0) and its weight init, and the initial
            # global model - hence the final one - then depends on the thread schedule (the two runs compared by
            # test_cross_silo_backends_agree differed by ~4e-4 in about half of the runs).
            with _INIT_LOCK:
                a = fedml.init(Arguments.from_dict(dict(base, rank=rank)))
                ds, out = fedml.data.load_cross_silo(a)
                torch.manual_seed(0)
                model = fedml.model.create(a, out)
            res[rank] = (fedml.cross_silo.Server if rank == 0 else fedml.cross_silo.Client)(a, torch.device("cpu"), ds, model).run()
        except BaseException as exc:  # noqa: BLE001
            errs.append(exc); raise

    ranks = kw.get("ranks", (0, 1, 2))
    ts = [threading.Thread(target=role, args=(r,), daemon=True) for r in ranks]
    [t.start() for t in ts]
    [t.join(90) for t in ts]
    assert not errs, errs
    return res


@pytest.mark.parametrize("backend", ["INPROC", "TCP", "GRPC", "MQTT_S3"])
def test_cross_silo_horizontal_handshake_and_rounds(backend, broker, tmp_path):
    res = _run_cross_silo(backend, broker, tmp_path)
    h = res[0].history
    assert len(h) == 2 and np.isfinite(h[-1]["test_loss"])
    sd = torch.load(str(tmp_path / "global_model.pt"), weights_only=False)      # reference checkpoint format
    assert list(sd.keys()) == ["linear.weight", "linear.bias"]


def test_cross_silo_backends_agree(broker, tmp_path):
    a = _run_cross_silo("INPROC", broker, tmp_path, tag="a")[0].get_global_model_params()
    b = _run_cross_silo("MQTT_S3", broker, tmp_path, tag="b")[0].get_global_model_params()
    for k in a:
        torch.testing.assert_close(a[k], b[k])


def test_round_deadline_partial_aggregation(broker, tmp_path):
    """A silo that never answers: the server aggregates what arrived after the deadline with re-normalised weights
    (the reference waits forever)."""
    from fedml_b200.cross_silo.horizontal import FedMLClientManager

    orig = FedMLClientManager._train_on

    def lazy(self, msg_params):
        if self.rank == 2:
            return                      # dead client: online, but never uploads
        return orig(self, msg_params)

    FedMLClientManager._train_on = lazy
    try:
        t0 = time.time()
        res = _run_cross_silo("INPROC", broker, tmp_path, tag="dl", round_deadline_s=1.0, comm_round=1)
        assert len(res[0].history) == 1 and time.time() - t0 < 60
    finally:
        FedMLClientManager._train_on = orig


def test_cross_device_server_with_file_models(broker, tmp_path):
    import shutil
    from fedml_b200.cross_device.mnn_file import read_mnn_as_tensor_dict, read_mnn_like, write_tensor_dict_to_mnn
    base = dict(training_type="cross_device", dataset="mnist", model="lenet", client_num_in_total=2, client_num_per_round=2, client_id_list="[1, 2]",
                comm_round=2, batch_size=10, synthetic_train_size=300, mnist_client_num=2, backend="MQTT_S3_MNN", mqtt_port=broker, log_level="WARNING",
                run_id="cd-t", object_store_dir=str(tmp_path / "store"), model_file_cache_folder=str(tmp_path / "cache"),
                global_model_file_path=str(tmp_path / "global.mnn"))
    args = fedml.init(Arguments.from_dict(base))
    ds, out = fedml.data.load(args)
    model = fedml.model.create(args, out)
    fedml.model.create_mnn_lenet5_model(str(tmp_path / "lenet.mnn"))
    assert len(read_mnn_as_tensor_dict(str(tmp_path / "lenet.mnn"))) == len(model.state_dict())
    srv = fedml.cross_device.ServerMNN(args, torch.device("cpu"), ds[3], model)

    class Phone(Observer):
        def __init__(s, rank):
            a = Arguments.from_dict(dict(base, rank=rank)); a.model_file_cache_folder = str(tmp_path / f"phone{rank}")
            s.rank, s.cm = rank, create_comm_manager(a, rank, 3, "MQTT_S3_MNN"); s.cm.add_observer(s)

        def receive_message(s, t, p):
            if t in (1, 2):
                # a phone trains the graph it downloaded and uploads the same graph with new parameter values
                path = str(tmp_path / f"phone{s.rank}" / "up.mnn"); shutil.copyfile(p["model_params"], path)
                write_tensor_dict_to_mnn(path, {i: v + 0.01 * s.rank for i, v in read_mnn_as_tensor_dict(path).items()})
                m = Message(3, s.rank, 0); m.add_params("model_params", path); m.add_params("num_samples", 10 * s.rank); s.cm.send_message(m)
            elif t == 6:
                s.cm.stop_receive_message()

    phones = [Phone(1), Phone(2)]
    [threading.Thread(target=p.cm.handle_receive_message, daemon=True).start() for p in phones]
    init = {k: v.clone() for k, v in model.state_dict().items()}
    agg = srv.run()
    assert len(agg.history) == 2 and agg.eval_engine == "mnn_graph" and np.isfinite(agg.history[-1]["test_loss"])
    final = list(read_mnn_like(str(tmp_path / "global.mnn")).values())
    # two rounds of (+0.01*1 weighted 10, +0.01*2 weighted 20) => +2 * 0.01 * 5/3; the last parameter of the MNN graph is fc2's bias
    torch.testing.assert_close(final[-1].reshape(-1), init["fc2.bias"] + 2 * 0.01 * 5 / 3, rtol=1e-4, atol=1e-5)
    from fedml_b200.cross_device import mnn_format
    assert [o["type"] for o in mnn_format.describe(str(tmp_path / "global.mnn"))].count(mnn_format.OP_CONV) == 2     # still the LeNet graph


def test_cross_device_server_with_the_resnet20_training_graph(broker, tmp_path):
    """The CIFAR-10 cross-device example: the global file is the 416-op ResNet-20 MNN graph; the server averages the uploaded
    files and evaluates by executing the graph (batch statistics, like a phone in training mode)."""
    import shutil
    from fedml_b200.cross_device import mnn_format as MF
    from fedml_b200.cross_device.mnn_file import read_mnn_as_tensor_dict, read_mnn_like, write_tensor_dict_to_mnn
    base = dict(training_type="cross_device", dataset="cifar10", model="resnet20", client_num_in_total=2, client_num_per_round=2, client_id_list="[1, 2]",
                comm_round=1, batch_size=32, synthetic_train_size=200, backend="MQTT_S3_MNN", mqtt_port=broker, log_level="WARNING", partition_method="homo",
                run_id="cd-r20", object_store_dir=str(tmp_path / "store"), model_file_cache_folder=str(tmp_path / "cache"),
                global_model_file_path=str(tmp_path / "global.mnn"))
    args = fedml.init(Arguments.from_dict(base))
    ds, out = fedml.data.load(args)
    model = fedml.model.create(args, out)
    srv = fedml.cross_device.ServerMNN(args, torch.device("cpu"), ds[3], model)

    class Phone(Observer):
        def __init__(s, rank):
            a = Arguments.from_dict(dict(base, rank=rank)); a.model_file_cache_folder = str(tmp_path / f"phone{rank}")
            s.rank, s.cm = rank, create_comm_manager(a, rank, 3, "MQTT_S3_MNN"); s.cm.add_observer(s)

        def receive_message(s, t, p):
            if t in (1, 2):
                path = str(tmp_path / f"phone{s.rank}" / "up.mnn"); shutil.copyfile(p["model_params"], path)
                write_tensor_dict_to_mnn(path, {i: v + 0.01 * s.rank for i, v in read_mnn_as_tensor_dict(path).items()})
                m = Message(3, s.rank, 0); m.add_params("model_params", path); m.add_params("num_samples", 10 * s.rank); s.cm.send_message(m)
            elif t == 6:
                s.cm.stop_receive_message()

    phones = [Phone(1), Phone(2)]
    [threading.Thread(target=p.cm.handle_receive_message, daemon=True).start() for p in phones]
    init = [p.detach().clone() for p in model.parameters()]
    agg = srv.run()
    assert len(agg.history) == 1 and agg.eval_engine == "mnn_graph" and np.isfinite(agg.history[-1]["test_loss"])
    assert len(MF.describe(str(tmp_path / "global.mnn"))) == 416
    final = list(read_mnn_like(str(tmp_path / "global.mnn")).values())
    assert len(final) == 65
    torch.testing.assert_close(final[0], init[0] + 0.01 * 5 / 3, rtol=1e-4, atol=1e-5)             # weighted mean of +0.01 and +0.02
    torch.testing.assert_close(final[-1].reshape(-1), init[-1] + 0.01 * 5 / 3, rtol=1e-4, atol=1e-5)


REF_MNN = "/root/reference/python/examples/cross_device/mqtt_s3_fedavg_mnist_lr_example/model_file_cache/global_model.mnn"
REF_MNN_RESNET = "/root/reference/python/examples/cross_device/mqtt_s3_fedavg_cifar10_lr_example/model_file_cache/global_model.mnn"


def test_mnn_format_builder_reader_and_in_place_writer(tmp_path):
    """``.mnn`` files without the MNN runtime (reference: cross_device/server_mnn/utils.py:11-50, model/mobile/mnn_lenet.py)."""
    from fedml_b200.cross_device import mnn_format as MF
    from fedml_b200.cross_device.mnn_file import load_into_module, read_mnn_like, write_mnn_like
    from fedml_b200.models.cv import LeNet5
    torch.manual_seed(0)
    net = LeNet5()
    path = str(tmp_path / "lenet.mnn")
    fedml.model.create_mnn_lenet5_model(path)
    write_mnn_like(path, net.state_dict())
    ops = MF.describe(path)
    assert [o["type"] for o in ops][:7] == [MF.OP_INPUT, MF.OP_CONVERT_TENSOR, MF.OP_TRAINABLE_PARAM, MF.OP_TRAINABLE_PARAM, MF.OP_CONV,
                                            MF.OP_RELU, MF.OP_POOLING]
    assert ops[4]["conv"] == {"kernel": [5, 5], "stride": [1, 1], "out": 20, "in": 1, "pad": [0, 0], "pad_mode": 1} and ops[4]["inputs"] == [1, 2, 3]
    assert all(o["outputs"] == [i] for i, o in enumerate(ops)) and ops[-1]["type"] == MF.OP_SOFTMAX
    got = MF.read_params(path)
    for (k, v), w in zip(net.state_dict().items(), got.values()):
        np.testing.assert_array_equal(v.numpy().reshape(w.shape), w)
    # positional load into a fresh module, in-place overwrite, size check
    other = load_into_module(path, LeNet5())
    for k, v in net.state_dict().items():
        torch.testing.assert_close(other.state_dict()[k], v, rtol=0, atol=0)
    size = os.path.getsize(path)
    MF.write_params(path, [w * 2 for w in got.values()])
    assert os.path.getsize(path) == size and MF.describe(path) == ops
    np.testing.assert_array_equal(list(MF.read_params(path).values())[0], list(got.values())[0] * 2)
    with pytest.raises(MF.MnnFormatError):
        MF.write_params(path, [w[..., :1] for w in got.values()])
    # logistic regression = one linear layer; other architectures = parameters-only graph keyed by state-dict names
    lr = torch.nn.Linear(784, 10)
    write_mnn_like(str(tmp_path / "lr.mnn"), lr.state_dict())
    assert [o["type"] for o in MF.describe(str(tmp_path / "lr.mnn"))] == [MF.OP_INPUT, MF.OP_RESHAPE, MF.OP_TRAINABLE_PARAM, MF.OP_MATMUL,
                                                                          MF.OP_TRAINABLE_PARAM, MF.OP_BINARY, MF.OP_SOFTMAX]
    # executing the graph = the torch module's forward (softmax on top)
    x = torch.randn(4, 1, 28, 28)
    MF.write_params(path, list(got.values()))
    with torch.no_grad():
        torch.testing.assert_close(MF.run_graph(path, x), torch.softmax(net(x), 1), rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(MF.run_graph(str(tmp_path / "lr.mnn"), x), torch.softmax(lr(x.flatten(1)), 1), rtol=1e-5, atol=1e-6)
    # ResNet-20: the full 416-op training graph; its 65 tensors load back into the torch module by position
    from fedml_b200.models.cv import BasicBlock, ResNet
    fedml.model.create_mnn_resnet20_model(str(tmp_path / "r20.mnn"))
    sd = read_mnn_like(str(tmp_path / "r20.mnn"))
    assert len(sd) == 65 and list(sd.values())[-2].shape == (10, 64) and len(MF.describe(str(tmp_path / "r20.mnn"))) == 416
    r20 = ResNet(BasicBlock, [3, 3, 3], 10, "cifar")
    write_mnn_like(str(tmp_path / "r20.mnn"), r20.state_dict())
    back = load_into_module(str(tmp_path / "r20.mnn"), ResNet(BasicBlock, [3, 3, 3], 10, "cifar"))
    for (k, v), (_, w) in zip(r20.named_parameters(), back.named_parameters()):
        torch.testing.assert_close(w, v, rtol=0, atol=0)


@pytest.mark.skipif(not os.path.exists(REF_MNN), reason="reference sample .mnn files not present")
def test_mnn_format_against_files_written_by_the_mnn_sdk(tmp_path):
    """The two ``.mnn`` files shipped with the reference's cross-device examples were written by MNN itself: the LeNet-5
    graph built here must decode to the same op list, and both files must survive read -> overwrite -> read."""
    from fedml_b200.cross_device import mnn_format as MF
    from fedml_b200.models.cv import LeNet5
    ref_ops = MF.describe(REF_MNN)
    mine = str(tmp_path / "mine.mnn")
    MF.build_lenet5(mine, [v.numpy() for v in LeNet5().state_dict().values()], ref_ops[0]["dims"])
    assert MF.describe(mine) == ref_ops
    for src, n_params in ((REF_MNN, 8), (REF_MNN_RESNET, None)):
        params = MF.read_params(src)
        assert n_params is None or len(params) == n_params
        out = str(tmp_path / "patched.mnn")
        MF.write_params(src, [np.full_like(v, 0.25) for v in params.values()], out_path=out)
        again = MF.read_params(out)
        assert list(again) == list(params) and all((v == 0.25).all() for v in again.values())
        assert os.path.getsize(out) == os.path.getsize(src) and MF.describe(out) == MF.describe(src)
    import torch
    probs = MF.run_graph(REF_MNN, torch.randn(3, 1, 28, 28))          # the LeNet file the MNN SDK wrote executes here
    assert probs.shape == (3, 10) and torch.allclose(probs.sum(1), torch.ones(3), atol=1e-5)
    # the ResNet-20 training graph MNN wrote: rebuilt op for op from its parameters, and executed against a plain restatement
    # (SAME padding with the odd cell at the bottom / right, batch statistics, biased variance)
    import torch.nn.functional as F
    params = list(MF.read_params(REF_MNN_RESNET).values())
    assert [int(np.prod(a.shape)) for a in params] == [int(np.prod(sh)) for sh in MF.resnet20_param_shapes()]
    rebuilt = MF.build_resnet20(str(tmp_path / "r20.mnn"), params)
    a, b = MF.describe(REF_MNN_RESNET), MF.describe(rebuilt)
    norm = lambda ops: [{k: (v.tolist() if hasattr(v, "tolist") else v) for k, v in o.items()} for o in ops]      # noqa: E731
    assert norm(a) == norm(b)
    it = iter(torch.from_numpy(p) for p in params)

    def conv_bn(x, stride, relu):
        w = next(it)
        k = w.shape[-1]
        need = max((-(-x.shape[-1] // stride) - 1) * stride + k - x.shape[-1], 0)
        y = F.conv2d(F.pad(x, (need // 2, need - need // 2, need // 2, need - need // 2)), w, stride=stride)
        y = F.batch_norm(y, None, None, next(it).flatten(), next(it).flatten(), training=True, eps=1e-5)
        return torch.relu(y) if relu else y

    x = torch.randn(6, 3, 32, 32)
    h = conv_bn(x, 1, True)
    cin = 16
    for cout, stride in MF.RESNET20_STAGES:
        y = conv_bn(conv_bn(h, stride, True), 1, False)
        h = torch.relu(y + (conv_bn(h, stride, False) if (stride != 1 or cin != cout) else h))
        cin = cout
    w, bias = next(it), next(it)
    want = torch.softmax(F.avg_pool2d(h, 8).flatten(1) @ w.t() + bias, 1)
    got = MF.run_graph(REF_MNN_RESNET, x)
    torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(MF.run_graph(rebuilt, x), got, rtol=0, atol=0)
    with pytest.raises(MF.MnnFormatError):                            # an op outside the supported set is reported, not skipped
        g = MF.MnnGraph()
        g._add("Interp", 99, 0, {}, [g.input([1, 3, 8, 8])])
        MF.run_graph(g.save(str(tmp_path / "odd.mnn")), torch.randn(1, 3, 8, 8))


def test_mpc_toolbox_and_secure_aggregation():
    from fedml_b200.core.security import mpc_function as M
    from fedml_b200.core.security.secagg import prg_mask, secure_weighted_mean
    p = M.DEFAULT_PRIME
    assert M.modular_inv(12345, p) * 12345 % p == 1 and M.divmod(10, 5, p) == 2
    rs = np.random.RandomState(0)
    X = rs.randint(0, p, size=(4, 5))
    sh = M.BGW_encoding(X, 7, 2, p, rs)
    assert np.array_equal(M.BGW_decoding(sh[[1, 3, 6]], [1, 3, 6], p)[0], X % p)
    assert not np.array_equal(sh[0], X % p)                                       # a single share reveals nothing useful
    Xl = rs.randint(0, p, size=(6, 3))
    enc = M.LCC_encoding(Xl, 8, 2, 0, p, rs)
    assert np.array_equal(M.LCC_decoding(enc[[0, 2]], 1, 8, 2, 0, [0, 2], p).reshape(6, 3), Xl % p)
    pts_a, pts_b = [1, 2, 3], [4, 5, 6, 7]
    e2 = M.LCC_encoding_with_points(Xl[:3], pts_a, pts_b, p)
    assert np.array_equal(M.LCC_decoding_with_points(e2[:3], pts_b[:3], pts_a, p), Xl[:3] % p)
    assert np.all(M.Gen_Additive_SS(10, 4, p, rs).sum(0) % p == 0)
    sk1, sk2 = 1234567, 7654321
    assert M.my_key_agreement(sk1, M.my_pk_gen(sk2, p, 7), p, 7) == M.my_key_agreement(sk2, M.my_pk_gen(sk1, p, 7), p, 7)
    assert np.array_equal(prg_mask(99, 1001, p), prg_mask(99, 1001, p)) and prg_mask(99, 8, p).max() < p
    ups = [torch.randn(500) for _ in range(5)]
    w = [3.0, 1.0, 2.0, 5.0, 4.0]
    ref = sum(wi * u for wi, u in zip(w, ups)) / sum(w)
    assert float((secure_weighted_mean(ups, w) - ref).abs().max()) < 1e-3
    alive = [0, 1, 3, 4]
    ref2 = sum(w[i] * ups[i] for i in alive) / sum(w[i] for i in alive)
    assert float((secure_weighted_mean(ups, w, dropped=[2]) - ref2).abs().max()) < 1e-3


def test_mlops_events_metrics_and_sysstats():
    from fedml_b200.core.mlops import MLOpsMetrics, MLOpsProfilerEvent, SysStats
    a = Arguments.from_dict(dict(using_mlops=True, run_id=5, edge_id=3))
    ev = MLOpsProfilerEvent.get_instance(a)
    ev.log_event_started("train", "0"); ev.log_event_ended("train", "0")
    sent = ev.messenger.sent
    assert [t for t, _ in sent] == ["/mlops/events"] * 2
    first, last = json.loads(sent[0][1]), json.loads(sent[1][1])
    assert first["event_name"] == "train" and "started_time" in first and "ended_time" in last and first["run_id"] == 5
    mm = MLOpsMetrics.get_instance(a)
    assert mm.report_client_training_status(3, "TRAINING")[0] == "fl_client/mlops/3/status"
    assert mm.report_server_training_metric({"accuracy": 0.5})[0] == "fl_server/mlops/training_progress_and_eval"
    assert "timestamp" in SysStats().produce_info(5, 3)


def test_cli_version_build_and_package_materialisation(tmp_path):
    from click.testing import CliRunner
    from fedml_b200.cli.cli import cli, materialise_package
    r = CliRunner().invoke(cli, ["version"])
    assert r.exit_code == 0 and "fedml version" in r.output
    src, cfg, out = tmp_path / "src", tmp_path / "cfg", tmp_path / "out"
    src.mkdir(); cfg.mkdir()
    (src / "main.py").write_text("print('client entry')\n")
    (cfg / "fedml_config.yaml").write_text("train_args:\n  client_num_in_total: 1\n  run_id: x\ncomm_args:\n  backend: MQTT_S3\n")
    r = CliRunner().invoke(cli, ["build", "-t", "client", "-sf", str(src), "-ep", "main.py", "-cf", str(cfg), "-df", str(out)])
    assert r.exit_code == 0, r.output
    z = out / "dist-packages" / "client-package.zip"
    assert z.exists()
    entry, conf, dyn = materialise_package(str(z), str(tmp_path / "work"), {"RUN_ID": 42, "CLIENT_INDEX": 2, "CLIENT_ID_LIST": "[7, 8]",
                                                                             "CLIENT_NUM": 2, "MQTT_CONFIG_PATH": "", "S3_CONFIG_PATH": "",
                                                                             "LOG_FILE_DIR": "/tmp"})
    import yaml
    user = yaml.safe_load(open(conf))
    assert os.path.exists(entry) and str(dyn["rank"]) == "2" and str(user["train_args"]["client_num_in_total"]) == "2"
    assert str(user["train_args"]["run_id"]) == "42"


def test_cli_login_docker_mode_prints_the_container_command(tmp_path, monkeypatch):
    """``fedml login <id> -d``: the agent in a container; everything the command touches lives under the (redirected) home."""
    from click.testing import CliRunner
    import fedml_b200.cli.cli as C
    monkeypatch.setenv("HOME", str(tmp_path))
    monkeypatch.setattr(C, "HOME", str(tmp_path / ".fedml_b200"))
    monkeypatch.setattr(C, "PID_FILE", str(tmp_path / ".fedml_b200" / "edge-process.id"))
    monkeypatch.setattr(C.shutil, "which", lambda name: None)          # never start a container from a test
    r = CliRunner().invoke(C.cli, ["login", "77", "-v", "dev", "-d"])
    assert r.exit_code == 0 and "docker run --rm --name fedml-client-77" in r.output and "ACCOUNT_ID=77" in r.output
    assert "fedml-b200/client-agent:dev" in r.output and (tmp_path / "fedml-client" / "fedml").is_dir()


def test_hs_fft_matches_numpy_oracle():
    from fedml_b200.simulation.sp.hs_fft import amplitude_normalize, amplitude_normalize_numpy
    x = torch.rand(4, 3, 16, 16)
    run = torch.rand(3, 16, 16)
    a, b, c = amplitude_normalize(x, run)
    a2, b2, c2 = amplitude_normalize_numpy(x.numpy(), run.numpy())
    assert np.abs(a.numpy() - a2).max() < 1e-4 and np.abs(b.numpy() - b2).max() < 1e-4


def test_secagg_keys_are_secret_masks_are_per_round_and_double_masking_recovers_dropouts():
    """private keys must not be derivable from client ids, masks must differ between rounds, and a
    server must not be able to unmask a client it (falsely) declares dropped."""
    import numpy as np
    import torch
    from fedml_b200.core.security.mpc_function import DEFAULT_PRIME as p
    from fedml_b200.core.security.secagg import (SecAggClient, SecAggServer, prg_mask, secure_weighted_mean, shamir_reconstruct,
                                                 shamir_share)
    # keys: two clients with the same id get different private keys (OS entropy), ids do not determine them
    assert SecAggClient(3).sk != SecAggClient(3).sk
    a, b = SecAggClient(0), SecAggClient(1)
    a.agree({1: b.pk}); b.agree({0: a.pk})
    assert a.seeds[1] == b.seeds[0]                                  # Diffie-Hellman from public keys alone
    # the PRG takes the round: same seed, different round -> different mask; uploads of two rounds do not share a mask
    assert not np.array_equal(prg_mask(99, 64, p, 0), prg_mask(99, 64, p, 1))
    x = torch.randn(257)
    y0 = a.mask(x, 1.0, 2 ** 16, [0, 1], round_idx=0).numpy()
    y1 = a.mask(x, 1.0, 2 ** 16, [0, 1], round_idx=1).numpy()
    assert not np.array_equal(y0, y1)                                # same update, different rounds: difference is not q(x)-q(x)=0
    # Shamir: t shares reconstruct, the secret is not among fewer
    sh = shamir_share(123456, list(range(5)), 3, p)
    assert shamir_reconstruct({i: sh[i] for i in (0, 2, 4)}, p) == 123456
    assert shamir_reconstruct({i: sh[i] for i in (1, 3)}, p) != 123456
    # a participant refuses to reveal both kinds of share for one client
    c = SecAggClient(2)
    c.receive_shares(1, {"b": 5, "sk": 6})
    with pytest.raises(ValueError):
        c.reveal_shares([0, 1], [1])
    # end to end with double masking, with and without a dropout, at a non-zero round
    ups = [torch.randn(501) for _ in range(5)]
    w = [3.0, 1.0, 2.0, 5.0, 4.0]
    ref = sum(wi * u for wi, u in zip(w, ups)) / sum(w)
    assert float((secure_weighted_mean(ups, w, round_idx=4) - ref).abs().max()) < 1e-3
    alive = [0, 1, 3, 4]
    ref2 = sum(w[i] * ups[i] for i in alive) / sum(w[i] for i in alive)
    assert float((secure_weighted_mean(ups, w, dropped=[2], round_idx=9) - ref2).abs().max()) < 1e-3
    # the server alone (no revealed shares) cannot unmask: a sum without the self masks removed is garbage
    cl = [SecAggClient(i) for i in range(3)]
    pks = {c_.cid: c_.pk for c_ in cl}
    for c_ in cl:
        c_.agree(pks)
        c_.new_round_secrets([0, 1, 2])
    masked = {i: cl[i].mask(ups[i], 1 / 3, 2 ** 16, [0, 1, 2], 0, self_mask=True) for i in range(3)}
    garbage = SecAggServer(p).aggregate(masked, 2 ** 16, [0, 1, 2])
    assert float((garbage - sum(ups[:3]) / 3).abs().max()) > 1.0


def test_grpc_wire_compatibility_with_the_reference_service():
    """A peer speaking the reference's protocol (service gRPCCommManager, CommRequest{client_id, message = pickle(Message)},
    proto grpc_comm_manager.proto:3-16 / grpc_comm_manager.py:64-88) can send to our node, and a node configured with
    grpc_wire_format=reference emits exactly that protocol."""
    import pickle
    import grpc
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    from fedml_b200.core.distributed.communication.grpc_backend import (GRPCCommManager, decode_comm_request,
                                                                         encode_comm_request, pickle_like_reference)
    # the real protobuf runtime agrees with the hand-written CommRequest codec
    fdp = descriptor_pb2.FileDescriptorProto(name="grpc_comm_manager_test.proto", syntax="proto3")
    m = fdp.message_type.add(name="CommRequest")
    m.field.add(name="client_id", number=1, type=descriptor_pb2.FieldDescriptorProto.TYPE_INT32, label=1)
    m.field.add(name="message", number=2, type=descriptor_pb2.FieldDescriptorProto.TYPE_BYTES, label=1)
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fdp)
    CommRequest = message_factory.GetMessageClass(pool.FindMessageTypeByName("CommRequest"))
    for cid, blob in ((7, b"hello" * 100), (0, b""), (300, bytes(range(256)))):
        wire = CommRequest(client_id=cid, message=blob).SerializeToString()
        assert wire == encode_comm_request(cid, blob) and decode_comm_request(wire) == (cid, blob)
    # reference-format pickle names the reference's class and unpickles where that class exists (the fedml alias package)
    msg = Message(3, 1, 0); msg.add_params("model_params", {"w": torch.arange(6.0).view(2, 3)}); msg.add_params("num_samples", 17)
    blob = pickle_like_reference(msg)
    assert b"fedml.core.distributed.communication.message\nMessage" in blob
    import fedml  # noqa: F401  (alias of fedml_b200: provides fedml.core.distributed.communication.message)
    back = pickle.loads(blob)
    assert back.get_type() == 3 and back.get_sender_id() == 1 and torch.equal(back.get("model_params")["w"], torch.arange(6.0).view(2, 3))
    # (1) reference-style client -> our server
    p0, p1 = next_port(), None
    got = []

    class Obs:
        def __init__(self): self.ev = threading.Event()
        def receive_message(self, t, p): got.append((t, p)); self.ev.set()

    server = GRPCCommManager("127.0.0.1", p0, client_id=0, client_num=1, base_port=p0, allow_pickle=True)
    obs = Obs(); server.add_observer(obs)
    th = threading.Thread(target=server.handle_receive_message, daemon=True); th.start()
    ch = grpc.insecure_channel(f"127.0.0.1:{p0}")
    stub = ch.unary_unary("/gRPCCommManager/sendMessage", request_serializer=None, response_deserializer=None)
    resp = stub(CommRequest(client_id=1, message=blob).SerializeToString(), wait_for_ready=True, timeout=20)
    assert decode_comm_request(resp) == (0, b"")
    assert obs.ev.wait(20) and got[0][0] == 3 and got[0][1]["num_samples"] == 17
    # without allow_pickle the reference service refuses pickles
    strict = GRPCCommManager("127.0.0.1", p0 + 1, client_id=1, client_num=1, base_port=p0)
    ch2 = grpc.insecure_channel(f"127.0.0.1:{p0 + 1}")
    stub2 = ch2.unary_unary("/gRPCCommManager/sendMessage", request_serializer=None, response_deserializer=None)
    with pytest.raises(grpc.RpcError):
        stub2(CommRequest(client_id=1, message=blob).SerializeToString(), wait_for_ready=True, timeout=20)
    # (2) our node in reference mode -> (our server standing in for) a reference server: same envelope on the wire
    sender = GRPCCommManager("127.0.0.1", p0 + 2, client_id=2, client_num=2, base_port=p0, wire_format="reference")
    obs.ev.clear()
    out = Message(9, 2, 0); out.add_params("k", [1, 2, 3])
    sender.send_message(out)
    assert obs.ev.wait(20) and got[-1][0] == 9 and got[-1][1]["k"] == [1, 2, 3]
    for mgr in (server, strict, sender):
        mgr.stop_receive_message()
    ch.close(); ch2.close()


def test_mqtt_wire_protocol_interoperates_with_a_standard_peer(broker):
    """The bundled broker / client speak MQTT 3.1.1 on the wire: a peer written against the spec with raw sockets (the bytes
    a paho / mosquitto client emits) connects, subscribes with a wildcard, receives what BrokerClient publishes, publishes
    at QoS 1 and gets its PUBACK; topic filters follow the spec."""
    import socket
    import struct
    from fedml_b200.core.distributed.communication.broker import topic_matches
    assert topic_matches("fedml_7_/+", "fedml_7_/3") and topic_matches("a/#", "a/b/c") and not topic_matches("a/+", "a/b/c") \
        and topic_matches("W/topic", "W/topic") and not topic_matches("W/topic", "W/topic2")
    s = socket.create_connection(("127.0.0.1", broker), timeout=10)
    # CONNECT: protocol "MQTT" level 4, clean session, keep-alive 60, client id "rawpeer"
    s.sendall(bytes.fromhex("10 13 00 04 4d 51 54 54 04 02 00 3c 00 07") + b"rawpeer")
    assert s.recv(4) == bytes.fromhex("20 02 00 00")                        # CONNACK, accepted
    # SUBSCRIBE packet id 1, filter "interop/+" QoS 0
    s.sendall(bytes.fromhex("82 0e 00 01 00 09") + b"interop/+" + b"\x00")
    assert s.recv(5) == bytes.fromhex("90 03 00 01 00")                     # SUBACK, granted QoS 0
    ours = BrokerClient("127.0.0.1", broker, client_id="ours")
    got = []
    ours.on_message = lambda t, p: got.append((t, p))
    ours.subscribe("interop/back")
    ours.publish("interop/x", b"payload-\x00\x01")
    hdr = s.recv(2)
    assert hdr[0] == 0x30                                                   # PUBLISH QoS 0
    body = s.recv(hdr[1])
    (tl,) = struct.unpack(">H", body[:2])
    assert body[2:2 + tl] == b"interop/x" and body[2 + tl:] == b"payload-\x00\x01"
    # raw peer publishes at QoS 1 (packet id 9) -> broker acknowledges and forwards to our client
    topic = b"interop/back"
    pkt = struct.pack(">H", len(topic)) + topic + struct.pack(">H", 9) + b"from-raw"
    s.sendall(bytes([0x32, len(pkt)]) + pkt)
    assert s.recv(4) == bytes.fromhex("40 02 00 09")                        # PUBACK id 9
    deadline = time.time() + 5
    while not got and time.time() < deadline:
        time.sleep(0.02)
    assert got == [("interop/back", b"from-raw")]
    echo = s.recv(2)                                                        # "interop/+" also matches its own publish
    assert echo[0] == 0x30 and s.recv(echo[1])[2:14] == b"interop/back"
    s.sendall(bytes.fromhex("c0 00"))
    assert s.recv(2) == bytes.fromhex("d0 00")                              # PINGREQ -> PINGRESP
    s.sendall(bytes.fromhex("e0 00"))                                       # DISCONNECT
    s.close(); ours.disconnect()


def test_mpi_backend_over_a_communicator(monkeypatch):
    """backend "MPI" with an mpi4py communicator uses MpiCommManager (send / recv of message dicts, as the reference's
    com_manager.py does).  mpi4py is not installable here: a queue-backed stand-in with the mpi4py call signatures
    (Get_rank, Get_size, send, recv, iprobe, Status) exercises the manager; without a communicator the factory keeps the
    socket transport."""
    import queue as _q
    import sys
    import types
    from fedml_b200.core.distributed.communication import create_comm_manager
    boxes = {0: _q.Queue(), 1: _q.Queue()}

    class FakeComm:
        def __init__(self, rank): self.rank = rank
        def Get_rank(self): return self.rank
        def Get_size(self): return 2
        def send(self, obj, dest, tag=0): boxes[dest].put(obj)
        def recv(self, source=None, tag=None, status=None): return boxes[self.rank].get()
        def iprobe(self, source=None, tag=None): return not boxes[self.rank].empty()

    mpi = types.SimpleNamespace(ANY_SOURCE=-1, ANY_TAG=-1, THREAD_MULTIPLE=3, Query_thread=lambda: 3, Status=lambda: None,
                                COMM_WORLD=FakeComm(0))
    pkg = types.ModuleType("mpi4py"); pkg.MPI = mpi
    monkeypatch.setitem(sys.modules, "mpi4py", pkg)
    a = Arguments.from_dict(dict(run_id="mpi-t"))
    m0 = create_comm_manager(a, 0, 2, "MPI", comm=FakeComm(0))
    m1 = create_comm_manager(a, 1, 2, "MPI", comm=FakeComm(1))
    from fedml_b200.core.distributed.communication.mpi_backend import MpiCommManager
    assert isinstance(m0, MpiCommManager) and isinstance(m1, MpiCommManager)
    got = []

    class Obs:
        def __init__(self): self.ev = threading.Event()
        def receive_message(self, t, p): got.append((t, p)); self.ev.set()

    o = Obs(); m0.add_observer(o)
    threading.Thread(target=m0.handle_receive_message, daemon=True).start()
    msg = Message(5, 1, 0); msg.add_params("w", torch.ones(3)); m1.send_message(msg)
    assert o.ev.wait(10) and got[0][0] == 5 and torch.equal(got[0][1]["w"], torch.ones(3))
    m0.stop_receive_message(); m1.stop_receive_message()
    from fedml_b200.core.distributed.communication.tcp import TcpCommManager
    a2 = Arguments.from_dict(dict(run_id="mpi-t2", tcp_base_port=next_port()))
    t0 = create_comm_manager(a2, 0, 1, "MPI")
    assert isinstance(t0, TcpCommManager)
    t0.stop_receive_message()


def test_trpc_round_trip_between_two_processes(tmp_path):
    """TorchRPC (TensorPipe) backend: two OS processes exchange tensor messages both ways (reference trpc_comm_manager.py)."""
    import subprocess
    import sys
    cfg = tmp_path / "trpc_master.csv"
    cfg.write_text(f"master_ip,master_port\n127.0.0.1,{next_port()}\n")
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "trpc_worker.py")
    procs = [subprocess.Popen([sys.executable, worker, str(r), str(cfg)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in (0, 1)]
    outs = [p.communicate(timeout=180)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs) and "TRPC_OK rank 0" in outs[0] and "TRPC_OK rank 1" in outs[1], outs


def test_mlops_log_uploader_device_binding_and_package_download(tmp_path):
    """MLOps side services against a local HTTP server: periodic run-log upload with retry (mlops_runtime_log.py:122-160),
    edge binding (login.py:482-497), remote broker / S3 configs (mlops_configs.py:29-68), package download over http."""
    import http.server
    import logging
    import zipfile
    from types import SimpleNamespace
    from fedml_b200.cli.edge_agent import bind_account_and_device_id, get_device_id, retrieve_package
    from fedml_b200.core.mlops import MLOpsConfigs, MLOpsRuntimeLog
    posts, fail_first = [], {"n": 1}
    pkg = tmp_path / "client-package.zip"
    with zipfile.ZipFile(pkg, "w") as z:
        z.writestr("client-package/conf/fedml.yaml", "entry_config: {}\n")

    class H(http.server.BaseHTTPRequestHandler):
        def log_message(self, *a): pass
        def do_GET(self):
            data = pkg.read_bytes()
            self.send_response(200); self.send_header("Content-Length", str(len(data))); self.end_headers(); self.wfile.write(data)
        def do_POST(self):
            body = json.loads(self.rfile.read(int(self.headers["Content-Length"])) or b"{}")
            if self.path.endswith("/logs/update") and fail_first["n"] > 0:
                fail_first["n"] -= 1
                self.send_response(500); self.end_headers(); return
            posts.append((self.path, body))
            if self.path.endswith("/edges/binding"):
                out = {"code": "SUCCESS", "data": {"id": 4711}}
            elif self.path.endswith("/configs/fetch"):
                out = {"code": "SUCCESS", "data": {"mqtt_config": {"BROKER_HOST": "h", "BROKER_PORT": 1883}, "s3_config": {"BUCKET_NAME": "b"}}}
            else:
                out = {"code": "SUCCESS"}
            data = json.dumps(out).encode()
            self.send_response(200); self.send_header("Content-Length", str(len(data))); self.end_headers(); self.wfile.write(data)

    srv = http.server.ThreadingHTTPServer(("127.0.0.1", 0), H)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    base = f"http://127.0.0.1:{srv.server_address[1]}"
    try:
        assert bind_account_and_device_id(base + "/fedmlOpsServer/edges/binding", 42, get_device_id()) == 4711
        assert posts[-1][1]["accountid"] == 42 and posts[-1][1]["deviceid"] == get_device_id()
        mq, s3 = MLOpsConfigs(SimpleNamespace(mlops_config_url=base + "/fedmlOpsServer/configs/fetch")).fetch_configs()
        assert mq["BROKER_PORT"] == 1883 and s3["BUCKET_NAME"] == "b"
        got = retrieve_package(base + "/packages/client-package.zip", str(tmp_path / "dl"))
        assert zipfile.ZipFile(got).namelist() == ["client-package/conf/fedml.yaml"]
        # run log: file handler + uploader thread; the first post fails (HTTP 500) and the same lines are sent again
        args = SimpleNamespace(run_id=77, rank=1, using_mlops=True, log_file_dir=str(tmp_path / "logs"), log_level="INFO",
                               log_server_url=base + "/fedmlLogsServer/logs/update", log_upload_interval=0.2)
        rl = MLOpsRuntimeLog.get_instance(args)
        rl.init_logs()
        logging.info("first line of the run")
        logging.info("second line of the run")
        deadline = time.time() + 10
        while not [p for p in posts if p[0].endswith("/logs/update")] and time.time() < deadline:
            time.sleep(0.05)
        rl.stop_uploader()
        ups = [b for p, b in posts if p.endswith("/logs/update")]
        assert ups and ups[0]["run_id"] == 77 and ups[0]["edge_id"] == 1 and ups[0]["created_by"] == "1"
        text = "".join(line for u in ups for line in u["logs"])
        assert text.count("first line of the run") == 1 and text.count("second line of the run") == 1 and fail_first["n"] == 0
    finally:
        srv.shutdown()
        for h in list(logging.getLogger().handlers):
            if getattr(h, "_fedml_b200", False):
                logging.getLogger().removeHandler(h)


def test_interface_helpers_context_manager_timer_and_mixing_matrix(tmp_path):
    """Helpers that sit on the reference's abstract interfaces: ``with`` support and ``notify`` on a transport, the
    tick / tock latency table, the mixing matrix of a topology."""
    from fedml_b200.core.distributed.communication import utils as U
    from fedml_b200.core.distributed.topology.topology_managers import SymmetricTopologyManager
    hub = InProcHub()
    args = Arguments.from_dict(dict(backend="INPROC", inproc_hub=hub, run_id="ctx"))
    got = []

    class Obs(Observer):
        def receive_message(self, t, p):
            got.append((t, p.get("k")))

    with create_comm_manager(args, 0, 2, "INPROC") as cm:
        cm.add_observer(Obs())
        m = Message(7, 1, 0); m.add_params("k", 42)
        cm.notify(m)
        assert got == [(7, 42)] and len(cm.observers_snapshot()) == 1
    U.latency_summary(reset=True)
    U.log_communication_tick(0, 1, timestamp=10.0)
    assert abs(U.log_communication_tock(0, 1, timestamp=10.25) - 0.25) < 1e-9 and U.log_communication_tock(0, 1) is None
    with U.comm_timer(2, 3):
        pass
    s = U.latency_summary()
    assert s[(0, 1)]["count"] == 1 and abs(s[(0, 1)]["max"] - 0.25) < 1e-9 and (2, 3) in s
    topo = SymmetricTopologyManager(6, 2); topo.generate_topology()
    w = topo.mixing_matrix()
    assert w.shape == (6, 6) and topo.is_row_stochastic() and np.allclose(w, w.T)
    assert sorted(j for j, _ in topo.in_neighbors(0)) == sorted(topo.get_in_neighbor_idx_list(0))


def test_mqtt_over_tls_with_a_self_signed_certificate(tmp_path):
    """MQTT 3.1.1 over TLS (the 8883 flavour): bundled broker with a certificate, native client verifying it against a CA bundle;
    a client that trusts another CA is refused, and the MQTT comm manager takes the options from the YAML keys."""
    import datetime
    import ssl
    from cryptography import x509
    from cryptography.hazmat.primitives import hashes, serialization
    from cryptography.hazmat.primitives.asymmetric import rsa
    from cryptography.x509.oid import NameOID
    import ipaddress

    def self_signed(prefix):
        key = rsa.generate_private_key(public_exponent=65537, key_size=2048)
        name = x509.Name([x509.NameAttribute(NameOID.COMMON_NAME, "127.0.0.1")])
        now = datetime.datetime.now(datetime.timezone.utc)
        cert = (x509.CertificateBuilder().subject_name(name).issuer_name(name).public_key(key.public_key())
                .serial_number(x509.random_serial_number()).not_valid_before(now - datetime.timedelta(minutes=5))
                .not_valid_after(now + datetime.timedelta(days=1))
                .add_extension(x509.SubjectAlternativeName([x509.IPAddress(ipaddress.ip_address("127.0.0.1"))]), critical=False)
                .add_extension(x509.BasicConstraints(ca=True, path_length=None), critical=True).sign(key, hashes.SHA256()))
        c, k = tmp_path / f"{prefix}.crt", tmp_path / f"{prefix}.key"
        c.write_bytes(cert.public_bytes(serialization.Encoding.PEM))
        k.write_bytes(key.private_bytes(serialization.Encoding.PEM, serialization.PrivateFormat.TraditionalOpenSSL, serialization.NoEncryption()))
        return str(c), str(k)

    crt, key = self_signed("broker")
    other_crt, _ = self_signed("other")
    b = LocalBroker(certfile=crt, keyfile=key)
    port = b.start()
    got = threading.Event()
    sub = BrokerClient("127.0.0.1", port, "tls-sub", ca_certs=crt)
    sub.on_message = lambda t, p: got.set() if (t, p) == ("a/b", b"secret") else None
    sub.subscribe("a/#")
    pub = BrokerClient("127.0.0.1", port, "tls-pub", ca_certs=crt)
    pub.publish("a/b", b"secret")
    assert got.wait(10)
    with pytest.raises((ssl.SSLError, ConnectionError, OSError)):
        BrokerClient("127.0.0.1", port, "tls-bad", ca_certs=other_crt)
    # through the comm manager: YAML keys mqtt_tls / mqtt_ca_certs
    base = dict(backend="MQTT", mqtt_host="127.0.0.1", mqtt_port=port, mqtt_tls=True, mqtt_ca_certs=crt, run_id="tlsrun", client_id_list="[1]",
                client_num_per_round=1)
    srv = create_comm_manager(Arguments.from_dict(dict(base, rank=0)), 0, 2, "MQTT")
    cli = create_comm_manager(Arguments.from_dict(dict(base, rank=1)), 1, 2, "MQTT")
    seen = threading.Event()

    class Obs(Observer):
        def receive_message(self, t, p):
            if t == 9:
                seen.set()

    srv.add_observer(Obs())
    threading.Thread(target=srv.handle_receive_message, daemon=True).start()
    time.sleep(0.3)
    m = Message(9, 1, 0); m.add_params("x", 1)
    cli.send_message(m)
    assert seen.wait(10)
    srv.stop_receive_message(); cli.stop_receive_message(); sub.disconnect(); pub.disconnect(); b.stop()


def test_callback_and_mailbox_observers():
    from fedml_b200.core.distributed.communication.observer import CallbackObserver, MailboxObserver, Observer
    seen = []
    cb = CallbackObserver(lambda t, p: seen.append((t, p)))
    box = MailboxObserver()
    assert isinstance(cb, Observer) and isinstance(box, Observer)
    cb.receive_message(3, {"a": 1})
    assert seen == [(3, {"a": 1})]
    threading.Timer(0.05, box.receive_message, args=(7, "late")).start()
    box.receive_message(5, "first")
    box.receive_message(5, "second")
    assert box.pending() == 2 and box.pending(5) == 2 and box.pending(7) == 0
    assert box.get(5) == "first" and box.get(7, timeout=2.0) == "late" and box.get(5, 0.1) == "second"
    with pytest.raises(TimeoutError):
        box.get(9, timeout=0.05)
    assert box.received == 3
def test_generated_code_surface_of_the_grpc_service():
    """A servicer subclass + stub calls written against grpc_comm_manager_pb2_grpc run on the hand-written classes, and talk to
    the framework's own gRPC manager (reference wire format)."""
    from concurrent import futures
    import grpc
    import socket
    from fedml.core.distributed.communication.grpc import grpc_comm_manager_pb2 as pb2, grpc_comm_manager_pb2_grpc as pb2_grpc

    got = []

    class Servicer(pb2_grpc.gRPCCommManagerServicer):
        def sendMessage(self, request, context):  # noqa: N802
            got.append((request.client_id, request.message))
            return pb2.CommResponse(client_id=7, message=b"ack:" + request.message)

    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=2))
    pb2_grpc.add_gRPCCommManagerServicer_to_server(Servicer(), server)
    server.add_insecure_port(f"127.0.0.1:{port}")
    server.start()
    try:
        with grpc.insecure_channel(f"127.0.0.1:{port}") as ch:
            stub = pb2_grpc.gRPCCommManagerStub(ch)
            r = stub.sendMessage(pb2.CommRequest(client_id=3, message=b"hello"), timeout=10)
            assert (r.client_id, r.message) == (7, b"ack:hello") and got == [(3, b"hello")]
            try:
                stub.handleReceiveMessage(pb2.CommRequest(client_id=3, message=b"x"), timeout=10)
                raise AssertionError("expected UNIMPLEMENTED")
            except grpc.RpcError as e:
                assert e.code() == grpc.StatusCode.UNIMPLEMENTED
        r = pb2_grpc.gRPCCommManager.sendMessage(pb2.CommRequest(client_id=4, message=b"once"), f"127.0.0.1:{port}", insecure=True, timeout=10)
        assert r.message == b"ack:once" and got[-1] == (4, b"once")
    finally:
        server.stop(0)
    assert "gRPCCommManager" in pb2.DESCRIPTOR.services_by_name


def test_edge_deployment_helpers(tmp_path, monkeypatch, broker):
    """MqttManager topic dispatch over the bundled broker, yaml / object-store / login bookkeeping helpers."""
    from fedml.cli.edge_deployment.mqtt_manager import MqttManager
    from fedml.cli.edge_deployment.yaml_utils import load_yaml_config
    from fedml_b200.cli import edge_deployment as E
    from fedml_b200.core.distributed.communication.broker import BrokerClient
    host, port = "127.0.0.1", broker
    mgr = MqttManager(host, port, None, None, 30, 42)
    got = []
    ev = threading.Event()
    mgr.add_message_listener("flserver_agent/42/start_train", lambda t, p: (got.append((t, json.loads(p))), ev.set()))
    mgr.add_message_listener("custom/topic", lambda t, p: (got.append((t, p)), ev.set()))
    pub = BrokerClient(host, port, client_id="server-agent")
    pub.publish("flserver_agent/42/start_train", json.dumps({"runId": 9}))
    assert ev.wait(5) and got[0] == ("flserver_agent/42/start_train", {"runId": 9})
    ev.clear()
    pub.publish("flserver_agent/43/start_train", "{}")           # another edge's topic: not delivered
    pub.publish("custom/topic", "hi")
    assert ev.wait(5) and got[1] == ("custom/topic", "hi") and len(got) == 2
    back = []
    ev2 = threading.Event()
    pub.on_message = lambda t, p: (back.append((t, p)), ev2.set())
    pub.subscribe("fl_client/mlops/status")
    mgr.send_message_json("fl_client/mlops/status", json.dumps({"edge_id": 42, "status": MqttManager.MSG_MLOPS_CLIENT_STATUS_IDLE}))
    assert ev2.wait(5) and json.loads(back[0][1])["status"] == "IDLE"
    mgr.remove_message_listener("custom/topic")
    mgr.disconnect()
    pub.disconnect()
    # yaml
    (tmp_path / "ok.yaml").write_text("a: {b: 1}\n")
    (tmp_path / "bad.yaml").write_text("a: [1, 2\n")
    assert load_yaml_config(str(tmp_path / "ok.yaml")) == {"a": {"b": 1}}
    with pytest.raises(ValueError, match="Yaml error"):
        load_yaml_config(str(tmp_path / "bad.yaml"))
    # object store round trip (directory store)
    monkeypatch.setenv("FEDML_OBJECT_STORE_DIR", str(tmp_path / "store"))
    (tmp_path / "pkg.zip").write_bytes(b"PK\x03\x04 payload")
    assert E.upload_file(str(tmp_path / "pkg.zip"), "run9/pkg.zip") is True
    E.download_file("run9/pkg.zip", str(tmp_path / "dl" / "pkg.zip"))
    assert (tmp_path / "dl" / "pkg.zip").read_bytes() == b"PK\x03\x04 payload"
    E.delete_s3_zip("run9/pkg.zip")
    assert E.upload_file(str(tmp_path / "missing.zip"), "x") is False
    with pytest.raises(FileNotFoundError):
        E.download_file("run9/pkg.zip", str(tmp_path / "dl" / "again.zip"))
    # login bookkeeping
    info = E.save_edge_infos("dev-1", 17, home=str(tmp_path / "home"))
    assert load_yaml_config(info) == {"unique_device_id": "dev-1", "edge_id": 17}
    assert E.logout(home=str(tmp_path / "home")) is True and E.logout(home=str(tmp_path / "home")) is False
    monkeypatch.setenv("HOME", str(tmp_path / "h2"))
    cmd = E.login_with_docker_mode(1234, "dev", run=False)
    assert cmd[:3] == ["docker", "run", "--rm"] and "ACCOUNT_ID=1234" in cmd and cmd[-1].endswith(":dev")


def test_hs_fft_low_frequency_window_rule_of_the_reference():
    """``process`` / ``mutate`` / ``normalize`` with the reference's semantics (only the centred window of the shifted spectrum
    takes the running amplitude; L = 0 -> the DC term), checked against a per-image restatement and against the torch form."""
    from fedml.simulation.single_process.hs_fedavg.hs_fft import extract_amp, mutate, normalize, process
    from fedml_b200.simulation.sp.hs_fft import amplitude_normalize
    rs = np.random.RandomState(0)
    x = rs.rand(5, 3, 8, 12).astype(np.float32)
    run = rs.rand(3, 8, 12).astype(np.float32) + 0.5
    for L in (0.0, 0.25):
        out, new_run = process(x.copy(), run.copy(), momentum=0.1, L=L)
        amp_avg = np.stack([extract_amp(img) for img in x]).astype(np.float32).mean(0)
        want_run = run * 0.9 + amp_avg * 0.1
        np.testing.assert_allclose(new_run, want_run, rtol=1e-6)
        b = int(np.floor(min(8, 12) * L))
        for i in range(5):
            f = np.fft.fft2(x[i].astype(np.float64), axes=(-2, -1))
            a = np.fft.fftshift(np.abs(f), axes=(-2, -1))
            t = np.fft.fftshift(want_run, axes=(-2, -1))
            a[:, 4 - b:4 + b + 1, 6 - b:6 + b + 1] = t[:, 4 - b:4 + b + 1, 6 - b:6 + b + 1]
            img = np.real(np.fft.ifft2(np.fft.ifftshift(a, axes=(-2, -1)) * np.exp(1j * np.angle(f)), axes=(-2, -1)))
            np.testing.assert_allclose(out[i], img, rtol=1e-4, atol=1e-5)
        t_out, t_run, _ = amplitude_normalize(torch.from_numpy(x), torch.from_numpy(run), 0.1, window=L)
        np.testing.assert_allclose(t_out.numpy(), out, rtol=1e-3, atol=1e-4)
        np.testing.assert_allclose(t_run.numpy(), new_run, rtol=1e-5)
    # L = 0 changes the mean of every channel (DC term) and nothing else
    out0, _ = process(x.copy(), run.copy(), L=0.0)
    d = np.fft.fft2(out0.astype(np.float64) - x, axes=(-2, -1))
    assert np.abs(d[..., 0, 0]).min() > 1e-3 and np.abs(d.reshape(5, 3, -1)[..., 1:]).max() < 1e-4
    first, first_run = process(x.copy(), np.zeros_like(run))                  # first batch: the running amplitude is the batch mean
    np.testing.assert_allclose(first_run, np.stack([extract_amp(i) for i in x]).mean(0), rtol=1e-5)
    sq, sq_run = x[0, :, :, :8], run[:, :, :8]                                 # 8 x 8: L = 0.5 -> window [0, 9) on both axes = full swap
    np.testing.assert_allclose(mutate(extract_amp(sq), sq_run, L=0.5), sq_run, rtol=1e-6)
    assert normalize(x[0], run, L=0.1).shape == x[0].shape
