"""The helper surface around EfficientNet that scripts import next to the model (reference: model/cv/efficientnet_utils.py).

``cv_zoo.EfficientNet`` is built from a stage table; this module offers the same information in the reference's vocabulary -
``BlockArgs`` / ``GlobalParams`` tuples, the compact block strings of the TensorFlow release (``r1_k3_s11_e1_i32_o16_se0.25``), the
per-variant coefficient table - plus the SAME-padded convolution / pooling layers in their static (size known at construction) and
dynamic (size read from the input) forms, and checkpoint loading from a local file (nothing is downloaded here).
"""
from __future__ import annotations

import collections
import math
import re
from typing import List, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from .cv_zoo import (EFFICIENTNET_SCALING, EFFICIENTNET_STAGES, Conv2dStaticSamePadding, MemoryEfficientSwish, Swish,  # noqa: F401
                     calculate_output_image_size, drop_connect, round_repeats)
from .cv_zoo import round_filters as _round_filters_raw

GlobalParams = collections.namedtuple("GlobalParams", [
    "width_coefficient", "depth_coefficient", "image_size", "dropout_rate", "num_classes", "batch_norm_momentum", "batch_norm_epsilon",
    "drop_connect_rate", "depth_divisor", "min_depth", "include_top"])
BlockArgs = collections.namedtuple("BlockArgs", [
    "num_repeat", "kernel_size", "stride", "expand_ratio", "input_filters", "output_filters", "se_ratio", "id_skip"])
GlobalParams.__new__.__defaults__ = (None,) * len(GlobalParams._fields)
BlockArgs.__new__.__defaults__ = (None,) * len(BlockArgs._fields)

# where the published ImageNet checkpoints live (kept for reference; load_pretrained_weights needs a local copy)
url_map = {f"efficientnet-b{i}": f"https://github.com/lukemelas/EfficientNet-PyTorch/releases/download/1.0/efficientnet-b{i}-{h}.pth"
           for i, h in enumerate(("355c32eb", "f1951068", "8bb594d6", "5fb5a3c3", "6ed6700e", "b6417697", "c76e70fd", "dcc49843"))}
url_map_advprop = {f"efficientnet-b{i}": f"https://github.com/lukemelas/EfficientNet-PyTorch/releases/download/1.0/adv-efficientnet-b{i}-{h}.pth"
                   for i, h in enumerate(("b64d5a18", "0f3ce85a", "6e9d97e5", "cdd7c0f4", "44fb3a87", "86493f6b", "ac80338e", "4652b6dd",
                                          "22e8fe65"))}


def round_filters(filters, global_params):
    """Channel count after width scaling, rounded to the depth divisor (``global_params`` form of ``cv_zoo.round_filters``)."""
    if isinstance(global_params, (int, float)) or global_params is None:
        return _round_filters_raw(filters, global_params)
    return _round_filters_raw(filters, global_params.width_coefficient, global_params.depth_divisor or 8, global_params.min_depth)


def efficientnet_params(model_name: str) -> Tuple[float, float, int, float]:
    """(width coefficient, depth coefficient, resolution, dropout) of a variant."""
    return EFFICIENTNET_SCALING[model_name]


class BlockDecoder:
    """``r<repeats>_k<kernel>_s<stride><stride>_e<expand>_i<in>_o<out>[_se<ratio>][_noskip]`` <-> :class:`BlockArgs`."""

    @staticmethod
    def _decode_block_string(block_string: str) -> BlockArgs:
        opts = {}
        for token in block_string.split("_"):
            m = re.match(r"([a-z]+)(\d.*)", token)
            if m:
                opts[m.group(1)] = m.group(2)
        s = opts["s"]
        if not (len(s) == 1 or (len(s) == 2 and s[0] == s[1])):
            raise ValueError(f"stride {s!r} in {block_string!r}: expected one digit or two equal digits")
        return BlockArgs(num_repeat=int(opts["r"]), kernel_size=int(opts["k"]), stride=[int(s[0])], expand_ratio=int(opts["e"]),
                         input_filters=int(opts["i"]), output_filters=int(opts["o"]),
                         se_ratio=float(opts["se"]) if "se" in opts else None, id_skip="noskip" not in block_string)

    @staticmethod
    def _encode_block_string(block: BlockArgs) -> str:
        stride = block.stride if isinstance(block.stride, (list, tuple)) else [block.stride]
        s = int(stride[0])
        parts = [f"r{block.num_repeat}", f"k{block.kernel_size}", f"s{s}{s}", f"e{block.expand_ratio}", f"i{block.input_filters}",
                 f"o{block.output_filters}"]
        if block.se_ratio is not None and 0 < block.se_ratio <= 1:
            parts.append(f"se{block.se_ratio}")
        if block.id_skip is False:
            parts.append("noskip")
        return "_".join(parts)

    @staticmethod
    def decode(string_list: List[str]) -> List[BlockArgs]:
        return [BlockDecoder._decode_block_string(s) for s in string_list]

    @staticmethod
    def encode(blocks_args: List[BlockArgs]) -> List[str]:
        return [BlockDecoder._encode_block_string(b) for b in blocks_args]


def efficientnet(width_coefficient=None, depth_coefficient=None, image_size=None, dropout_rate=0.2, drop_connect_rate=0.2,
                 num_classes=1000, include_top=True):
    """``(blocks_args, global_params)`` of the B0 stage table under the given scaling."""
    blocks = [BlockArgs(r, k, [s], e, i, o, se, True) for r, k, s, e, i, o, se in EFFICIENTNET_STAGES]
    gp = GlobalParams(width_coefficient=width_coefficient, depth_coefficient=depth_coefficient, image_size=image_size,
                      dropout_rate=dropout_rate, num_classes=num_classes, batch_norm_momentum=0.99, batch_norm_epsilon=1e-3,
                      drop_connect_rate=drop_connect_rate, depth_divisor=8, min_depth=None, include_top=include_top)
    return blocks, gp


def get_model_params(model_name: str, override_params: Optional[dict] = None):
    if not model_name.startswith("efficientnet") or model_name not in EFFICIENTNET_SCALING:
        raise NotImplementedError(f"model name is not pre-defined: {model_name}")
    w, d, s, p = efficientnet_params(model_name)
    blocks, gp = efficientnet(width_coefficient=w, depth_coefficient=d, dropout_rate=p, image_size=s)
    if override_params:
        gp = gp._replace(**override_params)
    return blocks, gp


# ------------------------------------------------------------------------------------------------ SAME-padded layers
def get_width_and_height_from_size(x):
    if isinstance(x, int):
        return x, x
    if isinstance(x, (list, tuple)):
        return tuple(x)
    raise TypeError(f"size must be an int or a pair, got {type(x).__name__}")


def _same_pad(size: int, kernel: int, stride: int, dilation: int = 1) -> int:
    return max((math.ceil(size / stride) - 1) * stride + (kernel - 1) * dilation + 1 - size, 0)


class Conv2dDynamicSamePadding(nn.Conv2d):
    """TensorFlow "SAME" convolution that computes its padding from each input (odd cell at the bottom / right)."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, dilation=1, groups=1, bias=True):
        super().__init__(in_channels, out_channels, kernel_size, stride, 0, dilation, groups, bias)

    def forward(self, x):
        ph = _same_pad(x.shape[-2], self.weight.shape[-2], self.stride[0], self.dilation[0])
        pw = _same_pad(x.shape[-1], self.weight.shape[-1], self.stride[1], self.dilation[1])
        if ph or pw:
            x = F.pad(x, [pw // 2, pw - pw // 2, ph // 2, ph - ph // 2])
        return F.conv2d(x, self.weight, self.bias, self.stride, 0, self.dilation, self.groups)


def get_same_padding_conv2d(image_size=None):
    """The dynamic class when the input size is unknown, else the static class bound to ``image_size``."""
    if image_size is None:
        return Conv2dDynamicSamePadding
    from functools import partial
    return partial(Conv2dStaticSamePadding, image_size=image_size)


class MaxPool2dDynamicSamePadding(nn.MaxPool2d):
    def __init__(self, kernel_size, stride, padding=0, dilation=1, return_indices=False, ceil_mode=False):
        super().__init__(kernel_size, stride, padding, dilation, return_indices, ceil_mode)
        self._k, self._s, self._d = (get_width_and_height_from_size(v) for v in (self.kernel_size, self.stride, self.dilation))

    def forward(self, x):
        ph = _same_pad(x.shape[-2], self._k[0], self._s[0], self._d[0])
        pw = _same_pad(x.shape[-1], self._k[1], self._s[1], self._d[1])
        if ph or pw:
            x = F.pad(x, [pw // 2, pw - pw // 2, ph // 2, ph - ph // 2])
        return F.max_pool2d(x, self.kernel_size, self.stride, self.padding, self.dilation, self.ceil_mode, self.return_indices)


class MaxPool2dStaticSamePadding(nn.MaxPool2d):
    def __init__(self, kernel_size, stride, image_size=None, **kw):
        super().__init__(kernel_size, stride, **kw)
        k, s, d = (get_width_and_height_from_size(v) for v in (self.kernel_size, self.stride, self.dilation))
        ih, iw = get_width_and_height_from_size(image_size)
        ph, pw = _same_pad(ih, k[0], s[0], d[0]), _same_pad(iw, k[1], s[1], d[1])
        self.static_padding = nn.ZeroPad2d((pw // 2, pw - pw // 2, ph // 2, ph - ph // 2)) if (ph or pw) else nn.Identity()

    def forward(self, x):
        return F.max_pool2d(self.static_padding(x), self.kernel_size, self.stride, self.padding, self.dilation, self.ceil_mode,
                            self.return_indices)


def get_same_padding_maxPool2d(image_size=None):  # noqa: N802
    if image_size is None:
        return MaxPool2dDynamicSamePadding
    from functools import partial
    return partial(MaxPool2dStaticSamePadding, image_size=image_size)


class SwishImplementation(torch.autograd.Function):
    """x * sigmoid(x) that keeps only ``x`` for the backward pass (the "memory efficient" form)."""

    @staticmethod
    def forward(ctx, i):
        ctx.save_for_backward(i)
        return i * torch.sigmoid(i)

    @staticmethod
    def backward(ctx, grad_output):
        (i,) = ctx.saved_tensors
        s = torch.sigmoid(i)
        return grad_output * (s * (1 + i * (1 - s)))


def load_pretrained_weights(model, model_name, weights_path=None, load_fc=True, advprop=False, verbose=True):
    """Loads a checkpoint file into ``model``.  ``weights_path`` must be a local file (there is no download step here; the
    published files are listed in ``url_map``).  ``load_fc=False`` leaves the classifier as initialised."""
    if not weights_path:
        raise FileNotFoundError(f"load_pretrained_weights({model_name}): pass weights_path=<local .pth file>; the published file is "
                                f"{(url_map_advprop if advprop else url_map).get(model_name, '<none>')}")
    state = torch.load(weights_path, map_location="cpu")
    if not load_fc:
        state = {k: v for k, v in state.items() if not k.startswith("_fc.")}
    result = model.load_state_dict(state, strict=False)
    missing = set(result.missing_keys)
    allowed = set() if load_fc else {"_fc.weight", "_fc.bias"}
    if missing - allowed or result.unexpected_keys:
        raise RuntimeError(f"checkpoint does not fit: missing {sorted(missing - allowed)}, unexpected {list(result.unexpected_keys)}")
    if verbose:
        print(f"Loaded pretrained weights for {model_name}")
    return model
