"""MobileNet, MobileNetV3, EfficientNet-B0..B8/L2 and VGG with the reference's state-dict layout.

Checkpoints are exchanged as ``state_dict``s (the global model of a cross-silo run, a pre-trained file handed to
``load_state_dict``), so what has to match the reference is the *parameter tree* - names, shapes, order - and the function it
computes.  Each network here is written from its table of stages; ``tests/test_model_zoo.py`` instantiates the reference's
module next to it (when /root/reference is mounted), copies the weights across by name and compares the outputs.

Reference files: model/cv/mobilenet.py:58-150, model/cv/mobilenet_v3.py:118-300, model/cv/efficientnet.py:41-370 +
efficientnet_utils.py (round_filters / SAME-padded convolutions / block strings), model/cv/vgg.py:22-140.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F


# ================================================================================================ MobileNet (v1)
class BasicConv2d(nn.Module):
    """conv -> BN -> ReLU (children ``conv``, ``bn``, ``relu``)."""

    def __init__(self, cin, cout, kernel_size, **kw):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, kernel_size, **kw)
        self.bn = nn.BatchNorm2d(cout)
        self.relu = nn.ReLU(inplace=True)

    def forward(self, x):
        return self.relu(self.bn(self.conv(x)))


class DepthSeperabelConv2d(nn.Module):  # (sic) the reference's spelling is part of its import surface
    """``depthwise`` = [k x k grouped conv, BN, ReLU], ``pointwise`` = [1x1 conv with bias, BN, ReLU]."""

    def __init__(self, cin, cout, kernel_size, **kw):
        super().__init__()
        self.depthwise = nn.Sequential(nn.Conv2d(cin, cin, kernel_size, groups=cin, **kw), nn.BatchNorm2d(cin), nn.ReLU(inplace=True))
        self.pointwise = nn.Sequential(nn.Conv2d(cin, cout, 1), nn.BatchNorm2d(cout), nn.ReLU(inplace=True))

    def forward(self, x):
        return self.pointwise(self.depthwise(x))


class MobileNet(nn.Module):
    """13 depthwise-separable layers in five named stages; the width multiplier thins every layer (``int(c * alpha)``)."""

    # stage name -> [(out channels, stride)]; the stem starts with a plain 3x3 convolution to 32 channels
    STAGES = (("stem", [(64, 1)]), ("conv1", [(128, 2), (128, 1)]), ("conv2", [(256, 2), (256, 1)]),
              ("conv3", [(512, 2)] + [(512, 1)] * 5), ("conv4", [(1024, 2), (1024, 1)]))

    def __init__(self, width_multiplier=1, class_num=100):
        super().__init__()
        w = lambda c: int(c * width_multiplier)                      # noqa: E731
        cin = w(32)
        for name, layers in self.STAGES:
            mods: List[nn.Module] = [BasicConv2d(3, cin, 3, padding=1, bias=False)] if name == "stem" else []
            for cout, stride in layers:
                mods.append(DepthSeperabelConv2d(cin, w(cout), 3, stride=stride, padding=1, bias=False))
                cin = w(cout)
            setattr(self, name, nn.Sequential(*mods))
        self.fc = nn.Linear(cin, class_num)
        self.avg = nn.AdaptiveAvgPool2d(1)

    def forward(self, x):
        for name, _ in self.STAGES:
            x = getattr(self, name)(x)
        return self.fc(self.avg(x).flatten(1))


def mobilenet(alpha=1, class_num=100):
    return MobileNet(alpha, class_num)


# ================================================================================================ MobileNetV3
class h_sigmoid(nn.Module):  # noqa: N801
    def __init__(self, inplace=True):
        super().__init__()
        self.inplace = inplace

    def forward(self, x):
        return F.relu6(x + 3.0, inplace=self.inplace) / 6.0


class h_swish(nn.Module):  # noqa: N801
    def __init__(self, inplace=True):
        super().__init__()
        self.inplace = inplace

    def forward(self, x):
        return x * (F.relu6(x + 3.0) / 6.0)


def make_divisible(v, divisor=8, min_value=None):
    """Nearest multiple of ``divisor`` (at least ``min_value``), bumped up when rounding lost more than 10 %."""
    new_v = max(min_value or divisor, int(v + divisor / 2) // divisor * divisor)
    return new_v + divisor if new_v < 0.9 * v else new_v


class SqueezeBlock(nn.Module):
    """Squeeze-and-excite with two Linear layers (``dense.0``, ``dense.2``) and a hard sigmoid gate."""

    def __init__(self, exp_size, divide=4):
        super().__init__()
        self.dense = nn.Sequential(nn.Linear(exp_size, exp_size // divide), nn.ReLU(inplace=True),
                                   nn.Linear(exp_size // divide, exp_size), h_sigmoid())

    def forward(self, x):
        gate = self.dense(x.mean(dim=(2, 3)))
        return x * gate[:, :, None, None]


class MobileBlock(nn.Module):
    """1x1 expand (``conv``) -> depthwise (``depth_conv``, no activation) -> optional ``squeeze_block`` -> 1x1 project with
    activation (``point_conv``); identity shortcut when the shape is kept."""

    def __init__(self, in_channels, out_channels, kernal_size, stride, nonLinear, SE, exp_size):  # noqa: N803
        super().__init__()
        self.out_channels, self.nonLinear, self.SE = out_channels, nonLinear, SE
        self.use_connect = stride == 1 and in_channels == out_channels
        act = nn.ReLU if nonLinear == "RE" else h_swish
        self.conv = nn.Sequential(nn.Conv2d(in_channels, exp_size, 1, bias=False), nn.BatchNorm2d(exp_size), act(inplace=True))
        self.depth_conv = nn.Sequential(
            nn.Conv2d(exp_size, exp_size, kernal_size, stride, (kernal_size - 1) // 2, groups=exp_size), nn.BatchNorm2d(exp_size))
        if SE:
            self.squeeze_block = SqueezeBlock(exp_size)
        self.point_conv = nn.Sequential(nn.Conv2d(exp_size, out_channels, 1), nn.BatchNorm2d(out_channels), act(inplace=True))

    def forward(self, x):
        out = self.depth_conv(self.conv(x))
        if self.SE:
            out = self.squeeze_block(out)
        out = self.point_conv(out)
        return x + out if self.use_connect else out


class MobileNetV3(nn.Module):
    # (in, out, kernel, stride, activation, squeeze-excite, expansion width)
    LARGE = ((16, 16, 3, 1, "RE", False, 16), (16, 24, 3, 2, "RE", False, 64), (24, 24, 3, 1, "RE", False, 72),
             (24, 40, 5, 2, "RE", True, 72), (40, 40, 5, 1, "RE", True, 120), (40, 40, 5, 1, "RE", True, 120),
             (40, 80, 3, 2, "HS", False, 240), (80, 80, 3, 1, "HS", False, 200), (80, 80, 3, 1, "HS", False, 184),
             (80, 80, 3, 1, "HS", False, 184), (80, 112, 3, 1, "HS", True, 480), (112, 112, 3, 1, "HS", True, 672),
             (112, 160, 5, 1, "HS", True, 672), (160, 160, 5, 2, "HS", True, 672), (160, 160, 5, 1, "HS", True, 960))
    SMALL = ((16, 16, 3, 2, "RE", True, 16), (16, 24, 3, 2, "RE", False, 72), (24, 24, 3, 1, "RE", False, 88),
             (24, 40, 5, 2, "RE", True, 96), (40, 40, 5, 1, "RE", True, 240), (40, 40, 5, 1, "RE", True, 240),
             (40, 48, 5, 1, "HS", True, 120), (48, 48, 5, 1, "HS", True, 144), (48, 96, 5, 2, "HS", True, 288),
             (96, 96, 5, 1, "HS", True, 576), (96, 96, 5, 1, "HS", True, 576))

    def __init__(self, model_mode="LARGE", num_classes=1000, multiplier=1.0, dropout_rate=0.0):
        super().__init__()
        mode = str(model_mode).upper()
        if mode not in ("LARGE", "SMALL"):
            raise ValueError("model_mode must be LARGE or SMALL")
        self.num_classes = num_classes
        d = lambda c: make_divisible(c * multiplier)                 # noqa: E731
        table = self.LARGE if mode == "LARGE" else self.SMALL
        self.init_conv = nn.Sequential(nn.Conv2d(3, d(16), 3, 2, 1), nn.BatchNorm2d(d(16)), h_swish(inplace=True))
        self.block = nn.Sequential(*[MobileBlock(d(i), d(o), k, s, nl, se, d(e)) for i, o, k, s, nl, se, e in table])
        last_in, last_out = (160, 960) if mode == "LARGE" else (96, 576)
        head: List[nn.Module] = [nn.Conv2d(d(last_in), d(last_out), 1)]
        if mode == "SMALL":
            head.append(SqueezeBlock(d(last_out)))
        self.out_conv1 = nn.Sequential(*head, nn.BatchNorm2d(d(last_out)), h_swish(inplace=True))
        self.out_conv2 = nn.Sequential(nn.Conv2d(d(last_out), d(1280), 1), h_swish(inplace=True), nn.Dropout(dropout_rate),
                                       nn.Conv2d(d(1280), num_classes, 1))
        self.apply(self._init)

    @staticmethod
    def _init(m):
        if isinstance(m, nn.Conv2d):
            nn.init.xavier_uniform_(m.weight)
            if m.bias is not None:
                nn.init.zeros_(m.bias)
        elif isinstance(m, nn.BatchNorm2d):
            nn.init.ones_(m.weight)
            nn.init.zeros_(m.bias)
        elif isinstance(m, nn.Linear):
            nn.init.normal_(m.weight, 0, 0.01)
            nn.init.zeros_(m.bias)

    def forward(self, x):
        x = self.out_conv1(self.block(self.init_conv(x)))
        return self.out_conv2(x.mean(dim=(2, 3), keepdim=True)).flatten(1)


def get_model_parameters(model) -> int:
    return sum(p.numel() for p in model.parameters())


# ================================================================================================ EfficientNet
# name -> (width coefficient, depth coefficient, native resolution, dropout)
EFFICIENTNET_SCALING = {
    "efficientnet-b0": (1.0, 1.0, 224, 0.2), "efficientnet-b1": (1.0, 1.1, 240, 0.2), "efficientnet-b2": (1.1, 1.2, 260, 0.3),
    "efficientnet-b3": (1.2, 1.4, 300, 0.3), "efficientnet-b4": (1.4, 1.8, 380, 0.4), "efficientnet-b5": (1.6, 2.2, 456, 0.4),
    "efficientnet-b6": (1.8, 2.6, 528, 0.5), "efficientnet-b7": (2.0, 3.1, 600, 0.5), "efficientnet-b8": (2.2, 3.6, 672, 0.5),
    "efficientnet-l2": (4.3, 5.3, 800, 0.5),
}
VALID_MODELS = tuple(EFFICIENTNET_SCALING)
# the B0 stages: (repeats, kernel, stride, expand ratio, in, out, squeeze-excite ratio)
EFFICIENTNET_STAGES = ((1, 3, 1, 1, 32, 16, 0.25), (2, 3, 2, 6, 16, 24, 0.25), (2, 5, 2, 6, 24, 40, 0.25), (3, 3, 2, 6, 40, 80, 0.25),
                       (3, 5, 1, 6, 80, 112, 0.25), (4, 5, 2, 6, 112, 192, 0.25), (1, 3, 1, 6, 192, 320, 0.25))


def round_filters(filters, width_coefficient, depth_divisor=8, min_depth=None):
    if not width_coefficient:
        return filters
    return int(make_divisible(filters * width_coefficient, depth_divisor, min_depth))


def round_repeats(repeats, depth_coefficient):
    return int(math.ceil(depth_coefficient * repeats)) if depth_coefficient else repeats


def drop_connect(inputs, p, training):
    """Stochastic depth: drops the whole residual branch of a sample with probability ``p`` and rescales the others."""
    if not training or not p:
        return inputs
    keep = 1.0 - p
    mask = torch.floor(keep + torch.rand([inputs.shape[0], 1, 1, 1], dtype=inputs.dtype, device=inputs.device))
    return inputs / keep * mask


class Swish(nn.Module):
    def forward(self, x):
        return x * torch.sigmoid(x)


MemoryEfficientSwish = Swish        # autograd of x * sigmoid(x) already recomputes nothing that matters at these sizes


class Conv2dStaticSamePadding(nn.Conv2d):
    """TensorFlow "SAME" convolution for a known input size: the zero padding (odd cell at the bottom / right) is fixed at
    construction in the parameter-free child ``static_padding``."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, image_size=None, **kw):
        super().__init__(in_channels, out_channels, kernel_size, stride, **kw)
        ih, iw = (image_size, image_size) if isinstance(image_size, int) else image_size
        (kh, kw_), (sh, sw) = self.weight.shape[-2:], self.stride
        ph = max((math.ceil(ih / sh) - 1) * sh + (kh - 1) * self.dilation[0] + 1 - ih, 0)
        pw = max((math.ceil(iw / sw) - 1) * sw + (kw_ - 1) * self.dilation[1] + 1 - iw, 0)
        self.static_padding = nn.ZeroPad2d((pw // 2, pw - pw // 2, ph // 2, ph - ph // 2)) if (ph or pw) else nn.Identity()

    def forward(self, x):
        return F.conv2d(self.static_padding(x), self.weight, self.bias, self.stride, 0, self.dilation, self.groups)


def calculate_output_image_size(image_size, stride):
    if image_size is None:
        return None
    ih, iw = (image_size, image_size) if isinstance(image_size, int) else image_size
    s = stride if isinstance(stride, int) else stride[0]
    return [int(math.ceil(ih / s)), int(math.ceil(iw / s))]


class MBConvBlock(nn.Module):
    """Inverted residual with squeeze-excite: ``_expand_conv/_bn0`` (absent at ratio 1), ``_depthwise_conv/_bn1``,
    ``_se_reduce/_se_expand``, ``_project_conv/_bn2``."""

    def __init__(self, cin, cout, kernel, stride, expand, se_ratio, image_size, bn_momentum=0.01, bn_eps=1e-3, id_skip=True):
        super().__init__()
        self.cin, self.cout, self.stride, self.expand, self.id_skip = cin, cout, stride, expand, id_skip
        mid = cin * expand
        if expand != 1:
            self._expand_conv = Conv2dStaticSamePadding(cin, mid, 1, image_size=image_size, bias=False)
            self._bn0 = nn.BatchNorm2d(mid, momentum=bn_momentum, eps=bn_eps)
        self._depthwise_conv = Conv2dStaticSamePadding(mid, mid, kernel, stride, image_size=image_size, groups=mid, bias=False)
        self._bn1 = nn.BatchNorm2d(mid, momentum=bn_momentum, eps=bn_eps)
        image_size = calculate_output_image_size(image_size, stride)
        self.has_se = se_ratio is not None and 0 < se_ratio <= 1
        if self.has_se:
            squeezed = max(1, int(cin * se_ratio))
            self._se_reduce = Conv2dStaticSamePadding(mid, squeezed, 1, image_size=(1, 1))
            self._se_expand = Conv2dStaticSamePadding(squeezed, mid, 1, image_size=(1, 1))
        self._project_conv = Conv2dStaticSamePadding(mid, cout, 1, image_size=image_size, bias=False)
        self._bn2 = nn.BatchNorm2d(cout, momentum=bn_momentum, eps=bn_eps)
        self._swish = Swish()

    def forward(self, inputs, drop_connect_rate=None):
        x = inputs
        if self.expand != 1:
            x = self._swish(self._bn0(self._expand_conv(x)))
        x = self._swish(self._bn1(self._depthwise_conv(x)))
        if self.has_se:
            gate = self._se_expand(self._swish(self._se_reduce(F.adaptive_avg_pool2d(x, 1))))
            x = torch.sigmoid(gate) * x
        x = self._bn2(self._project_conv(x))
        if self.id_skip and self.stride == 1 and self.cin == self.cout:
            x = drop_connect(x, drop_connect_rate, self.training) + inputs
        return x


class EfficientNet(nn.Module):
    """``EfficientNet.from_name("efficientnet-b3", num_classes=...)``; the bare constructor builds B0.  ``image_size`` fixes the
    SAME paddings (default: the variant's native resolution, like the reference)."""

    def __init__(self, model_name="efficientnet-b0", num_classes=1000, in_channels=3, image_size=None, drop_connect_rate=0.2,
                 dropout_rate=None, include_top=True):
        super().__init__()
        if model_name not in EFFICIENTNET_SCALING:
            raise ValueError("model_name should be one of: " + ", ".join(VALID_MODELS))
        width, depth, res, drop = EFFICIENTNET_SCALING[model_name]
        self.model_name, self.include_top, self.drop_connect_rate = model_name, include_top, drop_connect_rate
        size = image_size or res
        mom, eps = 0.01, 1e-3
        stem = round_filters(32, width)
        self._conv_stem = Conv2dStaticSamePadding(in_channels, stem, 3, 2, image_size=size, bias=False)
        self._bn0 = nn.BatchNorm2d(stem, momentum=mom, eps=eps)
        size = calculate_output_image_size(size, 2)
        blocks = []
        for repeats, kernel, stride, expand, cin, cout, se in EFFICIENTNET_STAGES:
            cin, cout = round_filters(cin, width), round_filters(cout, width)
            for r in range(round_repeats(repeats, depth)):
                blocks.append(MBConvBlock(cin if r == 0 else cout, cout, kernel, stride if r == 0 else 1, expand, se, size, mom, eps))
                if r == 0:
                    size = calculate_output_image_size(size, stride)
        self._blocks = nn.ModuleList(blocks)
        head = round_filters(1280, width)
        self._conv_head = Conv2dStaticSamePadding(round_filters(320, width), head, 1, image_size=size, bias=False)
        self._bn1 = nn.BatchNorm2d(head, momentum=mom, eps=eps)
        self._avg_pooling = nn.AdaptiveAvgPool2d(1)
        if include_top:
            self._dropout = nn.Dropout(drop if dropout_rate is None else dropout_rate)
            self._fc = nn.Linear(head, num_classes)
        self._swish = Swish()

    @classmethod
    def from_name(cls, model_name, in_channels=3, **override_params):
        return cls(model_name, in_channels=in_channels, **override_params)

    @classmethod
    def get_image_size(cls, model_name):
        return EFFICIENTNET_SCALING[model_name][2]

    def set_swish(self, memory_efficient=True):
        """Kept for API compatibility: there is one Swish implementation here."""

    def extract_features(self, inputs):
        x = self._swish(self._bn0(self._conv_stem(inputs)))
        for idx, block in enumerate(self._blocks):
            rate = self.drop_connect_rate * idx / len(self._blocks) if self.drop_connect_rate else None
            x = block(x, drop_connect_rate=rate)
        return self._swish(self._bn1(self._conv_head(x)))

    def extract_endpoints(self, inputs):
        """``{"reduction_k": feature map before the k-th resolution drop, ...}`` plus the head output as the last one."""
        endpoints, x = {}, self._swish(self._bn0(self._conv_stem(inputs)))
        prev = x
        for idx, block in enumerate(self._blocks):
            rate = self.drop_connect_rate * idx / len(self._blocks) if self.drop_connect_rate else None
            x = block(x, drop_connect_rate=rate)
            if prev.shape[2] > x.shape[2]:
                endpoints[f"reduction_{len(endpoints) + 1}"] = prev
            elif idx == len(self._blocks) - 1:
                endpoints[f"reduction_{len(endpoints) + 1}"] = x
            prev = x
        endpoints[f"reduction_{len(endpoints) + 1}"] = self._swish(self._bn1(self._conv_head(x)))
        return endpoints

    def forward(self, inputs):
        x = self._avg_pooling(self.extract_features(inputs))
        if self.include_top:
            x = self._fc(self._dropout(x.flatten(1)))
        return x


# ================================================================================================ VGG
cfgs = {
    "A": [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "B": [64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "D": [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"],
    "E": [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512, "M"],
}
_VGG_DEPTH = {11: "A", 13: "B", 16: "D", 19: "E"}


def make_layers(cfg: Sequence, batch_norm: bool = False) -> nn.Sequential:
    layers: List[nn.Module] = []
    cin = 3
    for v in cfg:
        if v == "M":
            layers.append(nn.MaxPool2d(2, 2))
            continue
        layers.append(nn.Conv2d(cin, int(v), 3, padding=1))
        if batch_norm:
            layers.append(nn.BatchNorm2d(int(v)))
        layers.append(nn.ReLU(inplace=True))
        cin = int(v)
    return nn.Sequential(*layers)


class VGG(nn.Module):
    """torchvision layout: ``features``, 7x7 adaptive pool, ``classifier`` 25088 -> 4096 -> 4096 -> classes.  ``features`` may
    also be a depth (11 / 13 / 16 / 19) with ``batch_norm=``; ``pool`` / ``hidden`` shrink the head for 32x32 inputs (pool=1,
    hidden=512 is the usual CIFAR variant)."""

    def __init__(self, features=11, num_classes: int = 1000, init_weights: bool = True, batch_norm: bool = False, pool: int = 7,
                 hidden: int = 4096, depth: Optional[int] = None):
        super().__init__()
        if depth is not None:
            features = depth
        if isinstance(features, int):
            features = make_layers(cfgs[_VGG_DEPTH[features]], batch_norm)
        self.features = features
        self.avgpool = nn.AdaptiveAvgPool2d((pool, pool))
        self.classifier = nn.Sequential(nn.Linear(512 * pool * pool, hidden), nn.ReLU(True), nn.Dropout(), nn.Linear(hidden, hidden),
                                        nn.ReLU(True), nn.Dropout(), nn.Linear(hidden, num_classes))
        if init_weights:
            for m in self.modules():
                if isinstance(m, nn.Conv2d):
                    nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
                    nn.init.zeros_(m.bias)
                elif isinstance(m, nn.BatchNorm2d):
                    nn.init.ones_(m.weight)
                    nn.init.zeros_(m.bias)
                elif isinstance(m, nn.Linear):
                    nn.init.normal_(m.weight, 0, 0.01)
                    nn.init.zeros_(m.bias)

    def forward(self, x):
        return self.classifier(torch.flatten(self.avgpool(self.features(x)), 1))


def _vgg(depth, bn):
    def build(num_classes: int = 1000, **kw):
        return VGG(make_layers(cfgs[_VGG_DEPTH[depth]], bn), num_classes=num_classes, **kw)
    build.__name__ = f"vgg{depth}{'_bn' if bn else ''}"
    build.__doc__ = f"VGG-{depth}{' with BatchNorm' if bn else ''} (configuration {_VGG_DEPTH[depth]!r})."
    return build


# note: the reference's vgg13_bn builds configuration "A" (a copy-paste slip in model/cv/vgg.py:163); this one builds "B"
vgg11, vgg11_bn, vgg13, vgg13_bn = _vgg(11, False), _vgg(11, True), _vgg(13, False), _vgg(13, True)
vgg16, vgg16_bn, vgg19, vgg19_bn = _vgg(16, False), _vgg(16, True), _vgg(19, False), _vgg(19, True)
