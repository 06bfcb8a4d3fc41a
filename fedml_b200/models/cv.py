"""Vision models of the zoo with reference-exact module (= ``state_dict`` key) names.

References: ``model/linear/lr.py:4-11`` (sigmoid LR), ``model/cv/cnn.py:5-142``,
``model/cv/resnet.py:87-264`` (CIFAR ResNet-56 = Bottleneck x [6,6,6], 591,322 params),
``model/cv/resnet_gn.py:26-196`` (ImageNet-stem ResNet with BN or GroupNorm),
``model/cv/mobilenet.py:58-150``, ``model/cv/vgg.py``, ``model/cv/mnist_gan.py``,
``model/cv/resnet56_gkt/*.py`` (split ResNet-8 / ResNet-56 pair for FedGKT),
``model/mobile/mnn_lenet.py`` (LeNet-5).  One configurable ``ResNet`` class covers the CIFAR and
ImageNet stems and both norm types so that checkpoints of either reference file load by key.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F


# ------------------------------------------------------------------ linear / small CNNs
class LogisticRegression(nn.Module):
    def __init__(self, input_dim, output_dim):
        super().__init__()
        self.linear = nn.Linear(input_dim, output_dim)

    def forward(self, x):
        # sigmoid *before* CrossEntropyLoss is the reference's numerics behind its 81.9 % figure
        return torch.sigmoid(self.linear(x.reshape(x.shape[0], -1)))


class CNN_OriginalFedAvg(nn.Module):  # noqa: N801
    def __init__(self, only_digits=True):
        super().__init__()
        self.only_digits = only_digits
        self.conv2d_1 = nn.Conv2d(1, 32, kernel_size=5, padding=2)
        self.max_pooling = nn.MaxPool2d(2, stride=2)
        self.conv2d_2 = nn.Conv2d(32, 64, kernel_size=5, padding=2)
        self.flatten = nn.Flatten()
        self.linear_1 = nn.Linear(3136, 512)
        self.linear_2 = nn.Linear(512, 10 if only_digits else 62)
        self.relu = nn.ReLU()

    def forward(self, x):
        x = x.reshape(x.shape[0], 1, 28, 28)
        x = self.max_pooling(self.relu(self.conv2d_1(x)))
        x = self.max_pooling(self.relu(self.conv2d_2(x)))
        x = self.relu(self.linear_1(self.flatten(x)))
        return self.linear_2(x)


class CNN_DropOut(nn.Module):  # noqa: N801
    def __init__(self, only_digits=True):
        super().__init__()
        self.conv2d_1 = nn.Conv2d(1, 32, kernel_size=3)
        self.max_pooling = nn.MaxPool2d(2, stride=2)
        self.conv2d_2 = nn.Conv2d(32, 64, kernel_size=3)
        self.dropout_1 = nn.Dropout(0.25)
        self.flatten = nn.Flatten()
        self.linear_1 = nn.Linear(9216, 128)
        self.dropout_2 = nn.Dropout(0.5)
        self.linear_2 = nn.Linear(128, 10 if only_digits else 62)
        self.relu = nn.ReLU()

    def forward(self, x):
        x = x.reshape(x.shape[0], 1, 28, 28)
        x = self.relu(self.conv2d_1(x))
        x = self.relu(self.conv2d_2(x))
        x = self.dropout_1(self.max_pooling(x))
        x = self.dropout_2(self.relu(self.linear_1(self.flatten(x))))
        return self.linear_2(x)


class LeNet5(nn.Module):
    """LeNet-5 used by the cross-device (Beehive) server examples (mnn_lenet.py:31)."""

    def __init__(self, class_num=10):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 20, 5)
        self.conv2 = nn.Conv2d(20, 50, 5)
        self.fc1 = nn.Linear(800, 500)
        self.fc2 = nn.Linear(500, class_num)

    def forward(self, x):
        x = x.reshape(x.shape[0], 1, 28, 28)
        x = F.max_pool2d(F.relu(self.conv1(x)), 2)
        x = F.max_pool2d(F.relu(self.conv2(x)), 2)
        return self.fc2(F.relu(self.fc1(x.flatten(1))))


# ------------------------------------------------------------------ ResNet family
class GroupNorm2d(nn.GroupNorm):
    """GroupNorm with ``channels_per_group`` semantics (group_normalization.py:7-93 emulates it
    through a reshaped ``F.batch_norm``; ``nn.GroupNorm`` is the same maths in one kernel)."""

    def __init__(self, planes, channels_per_group=32, eps=1e-5):
        super().__init__(max(1, planes // max(1, channels_per_group)), planes, eps=eps)


def norm2d(planes, channels_per_group=0):
    return GroupNorm2d(planes, channels_per_group) if channels_per_group > 0 else nn.BatchNorm2d(planes)


def conv3x3(i, o, stride=1):
    return nn.Conv2d(i, o, kernel_size=3, stride=stride, padding=1, bias=False)


def conv1x1(i, o, stride=1):
    return nn.Conv2d(i, o, kernel_size=1, stride=stride, bias=False)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, downsample=None, group_norm=0):
        super().__init__()
        self.conv1 = conv3x3(inplanes, planes, stride)
        self.bn1 = norm2d(planes, group_norm)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = conv3x3(planes, planes)
        self.bn2 = norm2d(planes, group_norm)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        idt = x if self.downsample is None else self.downsample(x)
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        return self.relu(out + idt)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None, group_norm=0):
        super().__init__()
        self.conv1 = conv1x1(inplanes, planes)
        self.bn1 = norm2d(planes, group_norm)
        self.conv2 = conv3x3(planes, planes, stride)
        self.bn2 = norm2d(planes, group_norm)
        self.conv3 = conv1x1(planes, planes * 4)
        self.bn3 = norm2d(planes * 4, group_norm)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        idt = x if self.downsample is None else self.downsample(x)
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        return self.relu(out + idt)


class ResNet(nn.Module):
    def __init__(self, block, layers, num_classes=10, stem="cifar", group_norm=0, KD=False,  # noqa: N803
                 zero_init_residual=False, widths=None):
        super().__init__()
        self.stem, self.KD, self.group_norm = stem, KD, group_norm
        if stem == "cifar":
            widths = widths or (16, 32, 64)
            self.inplanes = widths[0]
            self.conv1 = nn.Conv2d(3, self.inplanes, kernel_size=3, stride=1, padding=1, bias=False)
        else:
            widths = widths or (64, 128, 256, 512)
            self.inplanes = widths[0]
            self.conv1 = nn.Conv2d(3, self.inplanes, kernel_size=7, stride=2, padding=3, bias=False)
            self.maxpool = nn.MaxPool2d(kernel_size=3, stride=2, padding=1)
        self.bn1 = norm2d(self.inplanes, group_norm)
        self.relu = nn.ReLU(inplace=True)
        for i, (w, n) in enumerate(zip(widths, layers)):
            setattr(self, f"layer{i + 1}", self._make_layer(block, w, n, stride=1 if i == 0 else 2))
        self.num_stages = len(layers)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(widths[len(layers) - 1] * block.expansion, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, (nn.BatchNorm2d, nn.GroupNorm)):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)
        if zero_init_residual:
            for m in self.modules():
                if isinstance(m, Bottleneck):
                    nn.init.constant_(m.bn3.weight, 0)
                elif isinstance(m, BasicBlock):
                    nn.init.constant_(m.bn2.weight, 0)

    def _make_layer(self, block, planes, blocks, stride=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(conv1x1(self.inplanes, planes * block.expansion, stride),
                                       norm2d(planes * block.expansion, self.group_norm))
        layers = [block(self.inplanes, planes, stride, downsample, self.group_norm)]
        self.inplanes = planes * block.expansion
        for _ in range(1, blocks):
            layers.append(block(self.inplanes, planes, group_norm=self.group_norm))
        return nn.Sequential(*layers)

    def forward(self, x):
        x = self.relu(self.bn1(self.conv1(x)))
        if self.stem != "cifar":
            x = self.maxpool(x)
        for i in range(self.num_stages):
            x = getattr(self, f"layer{i + 1}")(x)
        feat = self.avgpool(x).flatten(1)
        logits = self.fc(feat)
        return (feat, logits) if self.KD else logits


def load_pretrained(model, path):
    """``torch.load(path)['state_dict']`` with optional ``module.`` prefix (resnet.py:265-277)."""
    ckpt = torch.load(path, map_location="cpu", weights_only=False)
    sd = ckpt["state_dict"] if isinstance(ckpt, dict) and "state_dict" in ckpt else ckpt
    model.load_state_dict({k.replace("module.", "", 1) if k.startswith("module.") else k: v for k, v in sd.items()})
    return model


def resnet56(class_num=10, pretrained=False, path=None, **kw):
    m = ResNet(Bottleneck, [6, 6, 6], num_classes=class_num, stem="cifar", **kw)
    return load_pretrained(m, path) if pretrained and path else m


def resnet110(class_num=10, pretrained=False, path=None, **kw):
    m = ResNet(Bottleneck, [12, 12, 12], num_classes=class_num, stem="cifar", **kw)
    return load_pretrained(m, path) if pretrained and path else m


def resnet18(num_classes=1000, group_norm=0, **kw):
    """torchvision-style ResNet-18 with BN (11,181,642 params at 10 classes)."""
    return ResNet(BasicBlock, [2, 2, 2, 2], num_classes=num_classes, stem="imagenet", group_norm=group_norm, **kw)


def resnet18_cifar(num_classes=10, **kw):
    """ResNet-18 with a 3x3 stem and no max-pool, the usual CIFAR adaptation; same keys."""
    m = ResNet(BasicBlock, [2, 2, 2, 2], num_classes=num_classes, stem="imagenet", **kw)
    m.conv1 = nn.Conv2d(3, 64, kernel_size=3, stride=1, padding=1, bias=False)
    nn.init.kaiming_normal_(m.conv1.weight, mode="fan_out", nonlinearity="relu")
    m.maxpool = nn.Identity()
    return m


def resnet18_gn(num_classes=100, group_norm=2, **kw):
    return resnet18(num_classes=num_classes, group_norm=group_norm, **kw)


def resnet34(num_classes=1000, group_norm=0):
    return ResNet(BasicBlock, [3, 4, 6, 3], num_classes, "imagenet", group_norm)


def resnet50(num_classes=1000, group_norm=0):
    return ResNet(Bottleneck, [3, 4, 6, 3], num_classes, "imagenet", group_norm)


# ------------------------------------------------------------------ FedGKT split pair
class _GKTClient(ResNet):
    """Edge side of FedGKT (reference resnet56_gkt/resnet_client.py:138-259): the CIFAR stem plus ``layer1`` only and a small
    head; returns ``(logits, feature_map)`` where the feature map [B, 16, 32, 32] - the stem output - is what gets uploaded."""

    def __init__(self, block, blocks, class_num=10, **kw):
        super().__init__(block, [blocks], num_classes=class_num, stem="cifar", **kw)
        self.maxpool = nn.MaxPool2d(kernel_size=3, stride=2, padding=1)      # present in the reference module, unused

    def forward(self, x):
        feat = self.relu(self.bn1(self.conv1(x)))
        return self.fc(self.avgpool(self.layer1(feat)).flatten(1)), feat


class ResNet8Client(_GKTClient):
    """``resnet8_56``: two bottleneck blocks (16 -> 64 channels)."""

    def __init__(self, class_num=10, **kw):
        super().__init__(Bottleneck, 2, class_num, **kw)


class ResNet5Client(_GKTClient):
    """``resnet5_56``: one basic block."""

    def __init__(self, class_num=10, **kw):
        super().__init__(BasicBlock, 1, class_num, **kw)


class ResNet56Server(ResNet):
    """Server side of FedGKT (resnet56_gkt/resnet_server.py:139-252): the three bottleneck stages of ResNet-56 applied to the
    client's [B, 16, 32, 32] feature map.  The stem (``conv1`` / ``bn1``) is part of the parameter tree, as in the reference,
    but not of the forward pass."""

    def __init__(self, class_num=10, layers=(6, 6, 6), **kw):
        super().__init__(Bottleneck, list(layers), num_classes=class_num, stem="cifar", **kw)
        self.maxpool = nn.MaxPool2d(kernel_size=3, stride=2, padding=1)

    def forward(self, feat):
        x = self.layer3(self.layer2(self.layer1(feat)))
        return self.fc(self.avgpool(x).flatten(1))


def _with_checkpoint(model, pretrained, path):
    return load_pretrained(model, path) if pretrained and path else model


def resnet8_56(c, pretrained=False, path=None, **kw):
    return _with_checkpoint(ResNet8Client(c, **kw), pretrained, path)


def resnet5_56(c, pretrained=False, path=None, **kw):
    return _with_checkpoint(ResNet5Client(c, **kw), pretrained, path)


def resnet56_server(c, pretrained=False, path=None, **kw):
    return _with_checkpoint(ResNet56Server(c, **kw), pretrained, path)


class _GKTPretrained(ResNet):
    """Full CIFAR ResNet that also returns the stem's feature map: ``(logits, feature_map)`` (resnet56_gkt/resnet_pretrained.py)."""

    def __init__(self, block, layers, class_num=10, **kw):
        super().__init__(block, layers, num_classes=class_num, stem="cifar", **kw)
        self.maxpool = nn.MaxPool2d(kernel_size=3, stride=2, padding=1)

    def forward(self, x):
        feat = self.relu(self.bn1(self.conv1(x)))
        h = self.layer3(self.layer2(self.layer1(feat)))
        return self.fc(self.avgpool(h).flatten(1)), feat


def resnet32_pretrained(c, pretrained=False, path=None, **kw):
    """Basic blocks, [5, 5, 5] stages (resnet56_gkt/resnet_pretrained.py:261)."""
    return _with_checkpoint(_GKTPretrained(BasicBlock, [5, 5, 5], c, **kw), pretrained, path)


def resnet56_pretrained(c, pretrained=False, path=None, **kw):
    return _with_checkpoint(_GKTPretrained(Bottleneck, [6, 6, 6], c, **kw), pretrained, path)


def resnet101(num_classes=1000, group_norm=0, **kw):
    return ResNet(Bottleneck, [3, 4, 23, 3], num_classes, "imagenet", group_norm, **kw)


def resnet152(num_classes=1000, group_norm=0, **kw):
    return ResNet(Bottleneck, [3, 8, 36, 3], num_classes, "imagenet", group_norm, **kw)


class GroupNorm3d(nn.GroupNorm):
    """GroupNorm over [N, C, D, H, W] with the constructor of the reference's class (group_normalization.py:165)."""

    def __init__(self, num_features, num_groups=32, eps=1e-5, momentum=0.1, affine=True, track_running_stats=False):
        super().__init__(num_groups, num_features, eps=eps, affine=affine)

    def forward(self, x):
        if x.dim() != 5:
            raise ValueError(f"expected 5D input (got {x.dim()}D input)")
        return super().forward(x)


# ------------------------------------------------------------------ MobileNet / MobileNetV3 / EfficientNet / VGG
# live in cv_zoo.py (state-dict compatible with the reference's files); re-exported here for the older import path
from .cv_zoo import VGG, EfficientNet, MobileNet, MobileNetV3, mobilenet  # noqa: E402,F401


# ------------------------------------------------------------------ GAN (FedGAN)
class MNISTGenerator(nn.Module):
    def __init__(self, latent=100):
        super().__init__()
        self.latent = latent
        self.net = nn.Sequential(nn.Linear(latent, 256), nn.LeakyReLU(0.2), nn.Linear(256, 512), nn.LeakyReLU(0.2),
                                 nn.Linear(512, 784), nn.Tanh())

    def forward(self, z):
        return self.net(z)


class MNISTDiscriminator(nn.Module):
    def __init__(self):
        super().__init__()
        self.net = nn.Sequential(nn.Linear(784, 512), nn.LeakyReLU(0.2), nn.Linear(512, 256), nn.LeakyReLU(0.2),
                                 nn.Linear(256, 1), nn.Sigmoid())

    def forward(self, x):
        return self.net(x.reshape(x.shape[0], -1))


class MNISTGan(nn.Module):
    """Generator + discriminator under one module so FedGAN averages both dicts
    (fedgan/FedGANAggregator.py:69-99)."""

    def __init__(self, latent=100):
        super().__init__()
        self.netg, self.netd = MNISTGenerator(latent), MNISTDiscriminator()

    def forward(self, z):
        return self.netg(z)


# ------------------------------------------------------------------ segmentation (FedSeg)
class SegNetSmall(nn.Module):
    """Compact encoder-decoder for the FedSeg runner (the reference plugs DeepLab/UNet in)."""

    def __init__(self, class_num=21):
        super().__init__()
        c = lambda i, o: nn.Sequential(nn.Conv2d(i, o, 3, padding=1, bias=False), nn.BatchNorm2d(o), nn.ReLU(True))
        self.enc1, self.enc2, self.enc3 = c(3, 32), c(32, 64), c(64, 128)
        self.dec2, self.dec1 = c(128 + 64, 64), c(64 + 32, 32)
        self.head = nn.Conv2d(32, class_num, 1)

    def forward(self, x):
        e1 = self.enc1(x)
        e2 = self.enc2(F.max_pool2d(e1, 2))
        e3 = self.enc3(F.max_pool2d(e2, 2))
        d2 = self.dec2(torch.cat([F.interpolate(e3, scale_factor=2.0), e2], 1))
        d1 = self.dec1(torch.cat([F.interpolate(d2, scale_factor=2.0), e1], 1))
        return self.head(d1)
