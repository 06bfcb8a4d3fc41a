"""MobileNet / MobileNetV3 / EfficientNet / VGG: the parameter tree and the computed function are the reference's.

With /root/reference mounted the reference's own module files are loaded next to ours (they are plain torch code), the weights
are copied across *by name* and the outputs compared; without it the structural facts are checked on their own."""
import importlib.util
import os
import sys
import types

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
