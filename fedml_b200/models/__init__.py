"""``fedml.model`` — model zoo and the ``create(args, output_dim)`` factory.

Dispatch follows the reference (model/model_hub.py:13-53): (model, dataset) pairs pick the
dataset-specific variants, unknown names fall back to logistic regression.  Added names:
``resnet18`` (torchvision-style BN), ``resnet18_cifar``, ``distilbert`` (FedNLP), ``lenet``,
``vgg11..19(_bn)``, ``resnet110``, ``gan``, ``segnet``.
"""
import logging

from .cv import (CNN_DropOut, CNN_OriginalFedAvg, EfficientNet, LeNet5, LogisticRegression, MNISTGan, MobileNet,
                 MobileNetV3, ResNet, ResNet56Server, ResNet8Client, SegNetSmall, VGG, mobilenet, resnet110, resnet18,
                 resnet18_cifar, resnet18_gn, resnet34, resnet50, resnet56)
from .finance import DenseModel, LocalModel
from .nlp import DistilBertForSequenceClassification, RNN_OriginalFedAvg, RNN_StackOverFlow, distilbert

__all__ = ["create", "LogisticRegression", "CNN_DropOut", "CNN_OriginalFedAvg", "resnet56", "resnet110", "resnet18",
           "resnet18_gn", "resnet18_cifar", "mobilenet", "MobileNetV3", "EfficientNet", "RNN_OriginalFedAvg",
           "RNN_StackOverFlow", "distilbert", "DistilBertForSequenceClassification", "LeNet5", "VGG", "MNISTGan",
           "ResNet8Client", "ResNet56Server", "DenseModel", "LocalModel", "SegNetSmall",
           "create_mnn_lenet5_model", "create_mnn_resnet20_model"]


def create(args, output_dim):
    name = str(getattr(args, "model", "lr")).lower()
    dataset = str(getattr(args, "dataset", "mnist")).lower()
    logging.info("create_model. model_name = %s, output_dim = %s", name, output_dim)
    if name == "lr":
        in_dim = {"mnist": 784, "femnist": 784, "stackoverflow_lr": 10000}.get(dataset)
        if in_dim is None:
            in_dim = 60 if dataset.startswith("synthetic") else int(getattr(args, "input_dim", 784))
        return LogisticRegression(in_dim, output_dim)
    if name == "cnn":
        return CNN_DropOut(only_digits=(output_dim == 10))
    if name == "cnn_original_fedavg":
        return CNN_OriginalFedAvg(only_digits=(output_dim == 10))
    if name == "lenet":
        return LeNet5(output_dim)
    if name == "resnet18_gn":
        return resnet18_gn(num_classes=output_dim, group_norm=int(getattr(args, "group_norm_channels", 2)))
    if name == "resnet18":
        small = dataset in ("cifar10", "cifar100", "cinic10", "fed_cifar100")
        return resnet18_cifar(output_dim) if small else resnet18(num_classes=output_dim)
    if name == "resnet34":
        return resnet34(output_dim)
    if name == "resnet50":
        return resnet50(output_dim)
    if name == "resnet20":          # the mobile ResNet-20 of the cross-device CIFAR-10 example (model/mobile/mnn_resnet.py)
        from .cv import BasicBlock
        return ResNet(BasicBlock, [3, 3, 3], output_dim, "cifar")
    if name == "resnet56":
        return resnet56(class_num=output_dim)
    if name == "resnet110":
        return resnet110(class_num=output_dim)
    if name == "rnn":
        return RNN_StackOverFlow() if dataset == "stackoverflow_nwp" else RNN_OriginalFedAvg()
    if name == "mobilenet":
        return mobilenet(class_num=output_dim)
    if name == "mobilenet_v3":
        return MobileNetV3(model_mode=str(getattr(args, "model_mode", "LARGE")), num_classes=output_dim)
    if name == "efficientnet":
        variant = str(getattr(args, "efficientnet_variant", "efficientnet-b0"))
        return EfficientNet.from_name(variant, num_classes=output_dim, image_size=getattr(args, "image_size", None))
    if name.startswith("vgg"):
        depth = int("".join(ch for ch in name if ch.isdigit()) or 11)
        small = str(dataset).lower() in ("cifar10", "cifar100", "cinic10", "fed_cifar100", "mnist", "femnist")
        # torchvision layout (7x7 pool, 4096-wide head); 32x32 sets get the usual slim head on a 1x1 pool
        return VGG(depth, num_classes=output_dim, batch_norm=name.endswith("bn"), pool=1 if small else 7, hidden=512 if small else 4096)
    if name in ("distilbert", "distilbert-base-uncased"):
        return distilbert(num_labels=output_dim, tiny=bool(getattr(args, "tiny_model", False)))
    if name == "gan":
        return MNISTGan()
    if name in ("segnet", "deeplab", "unet"):
        return SegNetSmall(output_dim)
    if name in ("darts", "darts_search", "darts_gdas"):
        from .darts import Network, NetworkGDAS
        return (NetworkGDAS if name.endswith("gdas") else Network)(num_classes=output_dim)
    if name in ("resnet56_gkt", "resnet8_gkt", "gkt"):
        from .cv import ResNet8Client, ResNet56Server
        return ResNet8Client(output_dim), ResNet56Server(output_dim)      # (edge extractor + head, server trunk) for FedGKT
    # the reference falls back to logistic regression for any other name (model/model_hub.py, last branch); so do we, loudly
    logging.warning("create_model: unknown model %r for dataset %r, falling back to LogisticRegression(784, %s)", name, dataset, output_dim)
    return LogisticRegression(28 * 28, output_dim)


def create_mnn_lenet5_model(mnn_file_path, class_num=10):
    """Writes the initial global model file of the cross-device examples (model/mobile/mnn_lenet.py:31): a genuine MNN
    graph (Input, ConvertTensor, Convolution, ReLU, Pooling, Reshape, MatMul, BinaryOp, Softmax, TrainableParam ops) built
    by :mod:`fedml_b200.cross_device.mnn_format` - op for op the graph the reference's MNN script saves."""
    from ..cross_device.mnn_file import write_mnn_like
    write_mnn_like(mnn_file_path, LeNet5(class_num).state_dict())


def create_mnn_resnet20_model(mnn_file_path, class_num=10):
    """ResNet-20 (model/mobile/mnn_resnet.py:80-141): the 416-op training graph MNN itself saves - SAME-padded bias-free
    convolutions, training-mode BatchNorm expanded into reduction / unary / binary primitives, additions on NCHW copies - emitted
    op for op by ``mnn_format.build_resnet20`` (checked against the reference's own ``global_model.mnn``)."""
    from ..cross_device.mnn_file import write_mnn_like
    from .cv import BasicBlock
    write_mnn_like(mnn_file_path, ResNet(BasicBlock, [3, 3, 3], class_num, "cifar").state_dict())
