"""MobileNet / MobileNetV3 / EfficientNet / VGG: the parameter tree and the computed function are the reference's.

With /root/reference mounted the reference's own module files are loaded next to ours (they are plain torch code), the weights
are copied across *by name* and the outputs compared; without it the structural facts are checked on their own."""
import importlib.util
import os
import sys
import types

import numpy as np
import pytest
import torch

from fedml_b200.models import cv_zoo as Z

REF = "/root/reference/python/fedml/model/cv/"
needs_ref = pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not mounted")


def _ref(name):
    if "refcv" not in sys.modules:
        pkg = types.ModuleType("refcv")
        pkg.__path__ = [REF]
        sys.modules["refcv"] = pkg
    full = "refcv." + name
    if full not in sys.modules:
        spec = importlib.util.spec_from_file_location(full, REF + name + ".py")
        mod = importlib.util.module_from_spec(spec)
        sys.modules[full] = mod
        spec.loader.exec_module(mod)
    return sys.modules[full]


def _same(ref, mine, x, train=False):
    a = [(k, tuple(v.shape)) for k, v in ref.state_dict().items()]
    b = [(k, tuple(v.shape)) for k, v in mine.state_dict().items()]
    assert a == b
    mine.load_state_dict(ref.state_dict())
    ref.train(train)
    mine.train(train)
    with torch.no_grad():
        torch.manual_seed(0)
        ya = ref(x)
        torch.manual_seed(0)
        yb = mine(x)
    torch.testing.assert_close(yb, ya, rtol=1e-5, atol=1e-6)


@needs_ref
@pytest.mark.parametrize("alpha,classes,hw", [(1, 10, 32), (0.5, 7, 64)])
def test_mobilenet_matches_reference(alpha, classes, hw):
    _same(_ref("mobilenet").mobilenet(alpha, classes), Z.mobilenet(alpha, classes), torch.randn(2, 3, hw, hw))


@needs_ref
@pytest.mark.parametrize("mode,mult,drop", [("LARGE", 1.0, 0.0), ("SMALL", 1.0, 0.0), ("LARGE", 0.75, 0.2)])
def test_mobilenet_v3_matches_reference(mode, mult, drop):
    r = _ref("mobilenet_v3")
    _same(r.MobileNetV3(mode, 10, mult, drop), Z.MobileNetV3(mode, 10, mult, drop), torch.randn(2, 3, 64, 64))
    assert Z.get_model_parameters(Z.MobileNetV3(mode, 10, mult)) == r.get_model_parameters(r.MobileNetV3(mode, 10, mult))


@needs_ref
@pytest.mark.parametrize("name,size,train", [("efficientnet-b0", None, False), ("efficientnet-b1", None, False), ("efficientnet-b3", None, False),
                                             ("efficientnet-b0", 32, False), ("efficientnet-b0", None, True)])
def test_efficientnet_matches_reference(name, size, train):
    _ref("efficientnet_utils")
    r = _ref("efficientnet")
    kw = {"num_classes": 10}
    if size:
        kw["image_size"] = size
    hw = size or 96
    _same(r.EfficientNet.from_name(name, **kw), Z.EfficientNet.from_name(name, **kw), torch.randn(3, 3, hw, hw), train)
    assert Z.EfficientNet.get_image_size(name) == r.EfficientNet.get_image_size(name)
    u = _ref("efficientnet_utils")
    gp = u.GlobalParams(width_coefficient=1.4, depth_coefficient=1.8, depth_divisor=8, min_depth=None)
    assert [Z.round_filters(c, 1.4) for c in (32, 16, 24, 320, 1280)] == [u.round_filters(c, gp) for c in (32, 16, 24, 320, 1280)]
    assert [Z.round_repeats(n, 1.8) for n in (1, 2, 3, 4)] == [u.round_repeats(n, gp) for n in (1, 2, 3, 4)]


@needs_ref
@pytest.mark.parametrize("fn", ["vgg11", "vgg13", "vgg16_bn", "vgg19"])
def test_vgg_matches_reference(fn):
    _same(getattr(_ref("vgg"), fn)(), getattr(Z, fn)(), torch.randn(1, 3, 64, 64))


@needs_ref
def test_vgg13_bn_is_configuration_b_here():
    """The reference's ``vgg13_bn`` builds configuration "A" (model/cv/vgg.py:163), i.e. an 11-layer net; ours builds "B"."""
    ref_convs = sum(isinstance(m, torch.nn.Conv2d) for m in _ref("vgg").vgg13_bn().features)
    assert ref_convs == 8 and sum(isinstance(m, torch.nn.Conv2d) for m in Z.vgg13_bn().features) == 10


def test_structure_without_the_reference():
    m = Z.mobilenet(1, 10)
    keys = list(m.state_dict())
    assert keys[0] == "stem.0.conv.weight" and "conv3.5.pointwise.0.bias" in keys and "conv4.1.depthwise.1.running_var" in keys
    assert sum(p.numel() for p in m.parameters()) == 3223178 and len(keys) == 177
    v3 = Z.MobileNetV3("SMALL", 10)
    assert "block.0.squeeze_block.dense.0.weight" in v3.state_dict() and "out_conv1.1.dense.2.bias" in v3.state_dict()
    assert v3(torch.randn(2, 3, 32, 32)).shape == (2, 10)
    with pytest.raises(ValueError):
        Z.MobileNetV3("MEDIUM")
    assert [Z.make_divisible(v) for v in (12, 16 * 0.75, 184 * 0.75, 7)] == [16, 16, 136, 8]
    e = Z.EfficientNet.from_name("efficientnet-b0", num_classes=10)
    sd = e.state_dict()
    assert len(e._blocks) == 16 and "_blocks.0._expand_conv.weight" not in sd and "_blocks.1._expand_conv.weight" in sd
    assert sd["_blocks.0._se_reduce.weight"].shape == (8, 32, 1, 1) and sd["_fc.weight"].shape == (10, 1280)
    assert sum(p.numel() for p in e.parameters()) == 4020358
    assert isinstance(e._conv_stem.static_padding, torch.nn.ZeroPad2d) and e._conv_stem.static_padding.padding == (0, 1, 0, 1)
    ends = e.eval().extract_endpoints(torch.randn(1, 3, 64, 64))
    assert [tuple(v.shape[1:]) for v in ends.values()] == [(16, 32, 32), (24, 16, 16), (40, 8, 8), (112, 4, 4), (320, 2, 2), (1280, 2, 2)]
    # stochastic depth: whole samples are dropped and the rest rescaled; identity in eval mode
    x = torch.ones(64, 2, 1, 1)
    torch.manual_seed(0)
    y = Z.drop_connect(x, 0.5, True)
    assert set(y.flatten().tolist()) == {0.0, 2.0} and torch.equal(Z.drop_connect(x, 0.5, False), x)
    with pytest.raises(ValueError):
        Z.EfficientNet.from_name("efficientnet-b9")
    slim = Z.VGG(16, num_classes=10, batch_norm=True, pool=1, hidden=512)
    assert slim(torch.randn(2, 3, 32, 32)).shape == (2, 10) and slim.classifier[0].in_features == 512
    assert len(Z.make_layers(Z.cfgs["E"])) == 16 * 2 + 5 and Z.vgg19_bn.__name__ == "vgg19_bn"


# ------------------------------------------------------------------------------------------------ DARTS
DREF = "/root/reference/python/fedml/model/cv/darts/"


def _dref(name):
    if "refdarts" not in sys.modules:
        pkg = types.ModuleType("refdarts")
        pkg.__path__ = [DREF]
        sys.modules["refdarts"] = pkg
    full = "refdarts." + name
    if full not in sys.modules:
        spec = importlib.util.spec_from_file_location(full, DREF + name + ".py")
        mod = importlib.util.module_from_spec(spec)
        sys.modules[full] = mod
        spec.loader.exec_module(mod)
    return sys.modules[full]


@needs_ref
def test_darts_search_network_matches_reference(capsys):
    from fedml_b200.models import darts as D
    for n in ("genotypes", "operations", "utils"):
        _dref(n)
    ms = _dref("model_search")
    crit = torch.nn.CrossEntropyLoss()
    ref = ms.Network(8, 10, 5, crit, torch.device("cpu"))
    mine = D.Network(8, 10, 5, crit, torch.device("cpu"))
    _same(ref, mine, torch.randn(3, 3, 32, 32), train=True)
    _same(ref, mine, torch.randn(3, 3, 32, 32), train=False)
    # decoding: same genotype and the same counts of parametric operations
    torch.manual_seed(3)
    for p_ref, p_mine in zip(ref.arch_parameters(), mine.arch_parameters()):
        p_ref.data.normal_()
        p_mine.data.copy_(p_ref.data)
    g_ref, n_ref, r_ref = ref.genotype()
    g_mine, n_mine, r_mine = mine.genotype()
    assert (g_mine.normal, g_mine.reduce, list(g_mine.normal_concat), n_mine, r_mine) == \
        (g_ref.normal, g_ref.reduce, list(g_ref.normal_concat), n_ref, r_ref)
    assert abs(mine.get_current_model_size() - ref.get_current_model_size()) < 1e-9
    clone = mine.new()
    assert torch.equal(clone.alphas_normal, mine.alphas_normal) and clone.alphas_normal is not mine.alphas_normal
    # GDAS: same parameter tree (the forward pass samples, so only the tree is compared)
    gd = _dref("model_search_gdas").Network_GumbelSoftmax(8, 10, 5, crit, torch.device("cpu"))
    mine_gd = D.Network_GumbelSoftmax(8, 10, 5, crit, torch.device("cpu"))
    assert [(k, tuple(v.shape)) for k, v in gd.state_dict().items()] == [(k, tuple(v.shape)) for k, v in mine_gd.state_dict().items()]
    assert mine_gd.get_tau() == gd.get_tau() == 5
    capsys.readouterr()


@needs_ref
@pytest.mark.parametrize("arch", ["DARTS_V1", "DARTS_V2", "FedNAS_V1", "NASNet", "AmoebaNet"])
def test_darts_derived_networks_match_reference(arch, capsys):
    from fedml_b200.models import darts as D
    for n in ("genotypes", "operations", "utils"):
        _dref(n)
    g_ref, g_mine = getattr(_dref("genotypes"), arch), getattr(D, arch)
    assert (list(g_ref.normal), list(g_ref.normal_concat), list(g_ref.reduce), list(g_ref.reduce_concat)) == \
        (list(g_mine.normal), list(g_mine.normal_concat), list(g_mine.reduce), list(g_mine.reduce_concat))
    m = _dref("model")
    ref, mine = m.NetworkCIFAR(8, 10, 5, True, g_ref), D.NetworkCIFAR(8, 10, 5, True, g_mine)
    ref.drop_path_prob = mine.drop_path_prob = 0.0          # the reference's drop_path needs a CUDA tensor
    a = [(k, tuple(v.shape)) for k, v in ref.state_dict().items()]
    assert a == [(k, tuple(v.shape)) for k, v in mine.state_dict().items()]
    mine.load_state_dict(ref.state_dict())
    x = torch.randn(3, 3, 32, 32)
    for mode in (True, False):
        ref.train(mode)
        mine.train(mode)
        with torch.no_grad():
            (la, aa), (lb, ab) = ref(x), mine(x)
        torch.testing.assert_close(lb, la, rtol=1e-5, atol=1e-6)
        assert (aa is None) == (ab is None)
        if aa is not None:
            torch.testing.assert_close(ab, aa, rtol=1e-5, atol=1e-6)
    if arch == "DARTS_V2":
        ri, mi = m.NetworkImageNet(16, 20, 5, True, g_ref), D.NetworkImageNet(16, 20, 5, True, g_mine)
        ri.drop_path_prob = mi.drop_path_prob = 0.0
        assert [(k, tuple(v.shape)) for k, v in ri.state_dict().items()] == [(k, tuple(v.shape)) for k, v in mi.state_dict().items()]
        mi.load_state_dict(ri.state_dict())
        xi = torch.randn(2, 3, 224, 224)
        with torch.no_grad():
            (la, aa), (lb, ab) = ri.train()(xi), mi.train()(xi)
        torch.testing.assert_close(lb, la, rtol=1e-4, atol=1e-5)
        torch.testing.assert_close(ab, aa, rtol=1e-4, atol=1e-5)
    capsys.readouterr()


def test_darts_utilities(tmp_path):
    from fedml_b200.models import darts as D
    out = torch.tensor([[0.1, 0.5, 0.4], [0.8, 0.15, 0.05], [0.2, 0.3, 0.5], [0.3, 0.6, 0.1]])
    top1, top2 = D.accuracy(out, torch.tensor([1, 1, 2, 0]), topk=(1, 2))
    assert float(top1) == 50.0 and float(top2) == 100.0
    meter = D.AvgrageMeter()
    meter.update(2.0, 3)
    meter.update(4.0, 1)
    assert meter.avg == 2.5 and meter.cnt == 4
    net = D.NetworkCIFAR(4, 10, 2, True, D.DARTS_V2)
    D.save(net, str(tmp_path / "w.pt"))
    other = D.NetworkCIFAR(4, 10, 2, True, D.DARTS_V2)
    D.load(other, str(tmp_path / "w.pt"))
    assert all(torch.equal(a, b) for a, b in zip(net.state_dict().values(), other.state_dict().values()))
    assert D.count_parameters_in_MB(net) < sum(p.numel() for p in net.parameters()) / 1e6          # auxiliary head excluded
    D.save_checkpoint({"epoch": 1}, True, str(tmp_path))
    assert (tmp_path / "model_best.pth.tar").exists()
    D.create_exp_dir(str(tmp_path / "exp"), [__file__])
    assert (tmp_path / "exp" / "scripts" / os.path.basename(__file__)).exists()
    dot = D.plot(D.DARTS_V2.normal, str(tmp_path / "normal"))
    assert (tmp_path / "normal.dot").read_text() == dot and dot.count("->") == 8 + 4 and '"c_{k-2}" -> "0" [label="sep_conv_3x3"]' in dot
    # odd feature maps go through the factorised reduction (the reference's slicing only handles even sizes)
    assert D.FactorizedReduce(4, 8)(torch.randn(2, 4, 7, 7)).shape == (2, 8, 4, 4)


def test_architect_mixed_level_rules():
    """step_v2 = Adam on (grad_a L_val + lambda * grad_a L_train); lambda = 0 is the plain first-order rule, and the reference's
    constructor form ``Architect(model, criterion, args, device)`` is accepted."""
    import copy
    from types import SimpleNamespace
    from fedml_b200.models import darts as D
    torch.manual_seed(0)
    base = D.Network(4, 4, 3, steps=2, multiplier=2).double()
    xt, yt = torch.randn(5, 3, 8, 8, dtype=torch.float64), torch.randint(0, 4, (5,))
    xv, yv = torch.randn(5, 3, 8, 8, dtype=torch.float64), torch.randint(0, 4, (5,))
    args = SimpleNamespace(arch_learning_rate=1e-2, arch_weight_decay=0.0, momentum=0.9, weight_decay=3e-4)

    def run(fn):
        net = copy.deepcopy(base)
        arch = D.Architect(net, torch.nn.CrossEntropyLoss(), args, torch.device("cpu"))
        fn(arch)
        return torch.cat([p.detach().reshape(-1) for p in net.arch_parameters()]), net

    a0 = torch.cat([p.detach().reshape(-1) for p in base.arch_parameters()])
    mixed, net = run(lambda a: a.step_v2(xt, yt, xv, yv, 1.0, 1.0))
    val_only, _ = run(lambda a: a.step_v2(xt, yt, xv, yv, 0.0, 1.0))
    first, _ = run(lambda a: a.step(xt, yt, xv, yv, 0.05, None, unrolled=False))
    aos, _ = run(lambda a: a.step_AOS(xt, yt, xv, yv))
    single, _ = run(lambda a: a.step_single_level(xt, yt))
    wa, _ = run(lambda a: a.step_wa(xt, yt, xv, yv, 1.0))
    second, _ = run(lambda a: a.step_v2_2ndorder(xt, yt, xv, yv, 0.05, None, 1.0, 1.0))
    assert torch.equal(val_only, first) and torch.equal(aos, first) and torch.equal(wa, mixed)
    assert not torch.equal(mixed, first) and not torch.equal(single, first) and not torch.equal(second, mixed)
    assert all(not torch.equal(v, a0) for v in (mixed, first, single, second))
    # weights are untouched by architecture steps
    assert all(torch.equal(p, q) for p, q in zip(net.weight_parameters(), base.weight_parameters()))
    # the first Adam step moves every coordinate by lr against the sign of the combined gradient
    crit = torch.nn.CrossEntropyLoss()
    g = [gv + gt for gv, gt in zip(torch.autograd.grad(crit(base(xv), yv), base.arch_parameters()),
                                   torch.autograd.grad(crit(base(xt), yt), base.arch_parameters()))]
    g = torch.cat([t.reshape(-1) for t in g])
    moved = g.abs() > 1e-9
    torch.testing.assert_close((mixed - a0)[moved], (-1e-2 * torch.sign(g))[moved], rtol=1e-4, atol=1e-7)


@pytest.mark.parametrize("extra", [[], ["--optimization", "MiLeNAS"], ["--arch_search_method", "GDAS", "--optimization", "AOS"]])
def test_centralised_darts_search_script(tmp_path, extra):
    from fedml_b200.models import darts as D
    from fedml_b200.models.darts_scripts import search_main
    hist = search_main(["--epochs", "2", "--layers", "3", "--init_channels", "4", "--train_size", "96", "--batch_size", "16",
                        "--data", str(tmp_path / "none"), "--save", str(tmp_path / "exp"), "--report_freq", "0"] + extra)
    assert len(hist) == 2 and hist[1]["lr"] < hist[0]["lr"] and isinstance(hist[-1]["genotype"], D.Genotype)
    assert len(hist[-1]["genotype"].normal) == 8 and hist[-1]["model_size_mb"] > 0 and (tmp_path / "exp" / "weights.pt").exists()
    assert all(np.isfinite(h["train_loss"]) and np.isfinite(h["valid_loss"]) for h in hist)


def test_centralised_darts_train_script(tmp_path):
    from fedml_b200.models.darts_scripts import train_main
    hist = train_main(["--epochs", "2", "--layers", "3", "--init_channels", "4", "--train_size", "64", "--batch_size", "16", "--arch", "FedNAS_V1",
                       "--auxiliary", "--cutout", "--data", str(tmp_path / "none"), "--save", str(tmp_path / "exp"), "--report_freq", "0"])
    assert [h["drop_path_prob"] for h in hist] == [0.0, 0.1] and all(np.isfinite(h["train_loss"]) for h in hist)
    again = train_main(["--epochs", "1", "--layers", "3", "--init_channels", "4", "--train_size", "64", "--batch_size", "16", "--arch", "FedNAS_V1",
                        "--auxiliary", "--model_path", str(tmp_path / "exp" / "weights.pt"), "--data", str(tmp_path / "none"),
                        "--save", str(tmp_path / "exp2"), "--report_freq", "0"])
    assert len(again) == 1


@needs_ref
def test_fedgkt_networks_match_reference():
    from fedml_b200.models import cv as C

    def load(rel, name):
        spec = importlib.util.spec_from_file_location(name, REF + rel)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod
        spec.loader.exec_module(mod)
        return mod

    rc, rs, rp = (load(f"resnet56_gkt/resnet_{n}.py", f"ref_gkt_{n}") for n in ("client", "server", "pretrained"))
    x, feat = torch.randn(2, 3, 32, 32), torch.randn(2, 16, 32, 32)
    for ref, mine, inp in ((rc.resnet8_56(10), C.resnet8_56(10), x), (rc.resnet5_56(10), C.resnet5_56(10), x),
                           (rs.resnet56_server(10), C.resnet56_server(10), feat), (rp.resnet32_pretrained(10), C.resnet32_pretrained(10), x),
                           (rp.resnet56_pretrained(7), C.resnet56_pretrained(7), x)):
        assert [(k, tuple(v.shape)) for k, v in ref.state_dict().items()] == [(k, tuple(v.shape)) for k, v in mine.state_dict().items()]
        mine.load_state_dict(ref.state_dict())
        ref.eval()
        mine.eval()
        with torch.no_grad():
            ya, yb = ref(inp), mine(inp)
        for p, q in zip(ya if isinstance(ya, tuple) else (ya,), yb if isinstance(yb, tuple) else (yb,)):
            torch.testing.assert_close(q, p, rtol=1e-5, atol=1e-6)
    logits, fmap = C.ResNet8Client(10)(x)
    assert logits.shape == (2, 10) and fmap.shape == (2, 16, 32, 32)
    assert C.GroupNorm3d(8, 4)(torch.randn(2, 8, 3, 4, 4)).shape == (2, 8, 3, 4, 4)
    assert sum(p.numel() for p in C.resnet101(10).parameters()) > sum(p.numel() for p in C.resnet50(10).parameters())


def test_efficientnet_utils_surface(tmp_path):
    from fedml_b200.models import efficientnet_utils as U
    strings = ["r1_k3_s11_e1_i32_o16_se0.25", "r2_k5_s22_e6_i24_o40_se0.25", "r1_k3_s11_e6_i192_o320_noskip"]
    blocks = U.BlockDecoder.decode(strings)
    assert blocks[1] == U.BlockArgs(2, 5, [2], 6, 24, 40, 0.25, True) and blocks[2].id_skip is False and blocks[2].se_ratio is None
    assert U.BlockDecoder.encode(blocks) == strings
    b, gp = U.get_model_params("efficientnet-b2", {"num_classes": 7})
    assert gp.width_coefficient == 1.1 and gp.image_size == 260 and gp.num_classes == 7 and len(b) == 7 and U.round_filters(32, gp) == 32
    if os.path.isdir(REF):
        r = _ref("efficientnet_utils")
        rb, rgp = r.get_model_params("efficientnet-b2", {"num_classes": 7})
        assert [tuple(x) for x in rb] == [tuple(x) for x in b] and tuple(rgp) == tuple(gp)
        assert r.BlockDecoder.decode(strings) == blocks
        x = torch.randn(2, 3, 33, 47)
        for k, s in ((3, 2), (5, 1), (3, 1), (1, 2)):
            a, m = r.Conv2dDynamicSamePadding(3, 4, k, s), U.Conv2dDynamicSamePadding(3, 4, k, s)
            m.load_state_dict(a.state_dict())
            torch.testing.assert_close(m(x), a(x))
            pa, pm = r.MaxPool2dDynamicSamePadding(k, s), U.MaxPool2dDynamicSamePadding(k, s)
            torch.testing.assert_close(pm(x), pa(x))
            sa, sm = r.MaxPool2dStaticSamePadding(k, s, image_size=(33, 47)), U.MaxPool2dStaticSamePadding(k, s, image_size=(33, 47))
            torch.testing.assert_close(sm(x), sa(x))
        assert r.get_same_padding_conv2d(None).__name__ == U.get_same_padding_conv2d(None).__name__
    conv = U.get_same_padding_conv2d(64)(3, 8, 3, stride=2)
    assert isinstance(conv, Z.Conv2dStaticSamePadding) and conv(torch.randn(1, 3, 64, 64)).shape == (1, 8, 32, 32)
    assert U.Conv2dDynamicSamePadding(3, 8, 3, stride=2)(torch.randn(1, 3, 65, 31)).shape == (1, 8, 33, 16)
    assert U.get_same_padding_maxPool2d(None)(3, 2)(torch.randn(1, 3, 9, 9)).shape == (1, 3, 5, 5)
    x = torch.randn(5, 3, dtype=torch.float64, requires_grad=True)
    assert torch.autograd.gradcheck(U.SwishImplementation.apply, (x,))
    torch.testing.assert_close(U.SwishImplementation.apply(x), x * torch.sigmoid(x))
    net = Z.EfficientNet.from_name("efficientnet-b0", num_classes=5)
    torch.save(net.state_dict(), tmp_path / "b0.pth")
    other = Z.EfficientNet.from_name("efficientnet-b0", num_classes=9)
    U.load_pretrained_weights(other, "efficientnet-b0", str(tmp_path / "b0.pth"), load_fc=False, verbose=False)
    assert torch.equal(other._conv_stem.weight, net._conv_stem.weight) and other._fc.weight.shape == (9, 1280)
    with pytest.raises(FileNotFoundError, match="efficientnet-b0-355c32eb"):
        U.load_pretrained_weights(other, "efficientnet-b0")
    with pytest.raises(NotImplementedError):
        U.get_model_params("resnet50")
