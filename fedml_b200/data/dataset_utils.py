"""The per-dataset helper surface of the reference's image loaders, for scripts that go below ``fedml.data.load``.

The reference repeats the same dozen helpers in every ``data/<name>/data_loader.py`` + ``datasets.py`` pair (cifar10, cifar100,
cinic10, ImageNet, Landmarks, edge_case_examples): a ``*_truncated`` ``torch.utils.data.Dataset`` restricted to a client's sample
indices, the train / test transform pair (``RandomCrop(32, 4)``, horizontal flip, normalise, ``Cutout(16)``), ``partition_data`` ->
``net_dataidx_map``, ``record_net_data_stats``, the two text-file readers for a frozen non-IID split ("hetero-fix") and
``get_dataloader`` (data/cifar10/data_loader.py:17-230, data/cifar10/datasets.py:35-91).  Here they exist once, parameterised by
the dataset name:

* arrays come from the packed store (real files under ``datadir`` when present, otherwise the synthetic set of the same shape),
  so nothing is downloaded;
* the transforms are plain tensor code (no torchvision dependency) with the reference's parameters;
* the DataLoaders are the standard torch ones - this is the *compatibility* path; the B200 path (``NCCLSimulator``) reads the
  packed store directly and augments on the GPU (``data/augment.py``).
"""
from __future__ import annotations

import ast
import logging
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch.utils import data as torch_data

from . import synthetic as S

_STATS = {
    "cifar10": (S.CIFAR10_MEAN, S.CIFAR10_STD, 10),
    "cifar100": (S.CIFAR100_MEAN, S.CIFAR100_STD, 100),
    "cinic10": ((0.47889522, 0.47227842, 0.43047404), (0.24205776, 0.23828046, 0.25874835), 10),
}


class Cutout:
    """Zeroes one ``length`` x ``length`` square centred on a uniformly drawn pixel (clipped at the border) of a CHW tensor."""

    def __init__(self, length: int, generator: Optional[torch.Generator] = None):
        self.length, self.generator = int(length), generator

    def __call__(self, img: torch.Tensor) -> torch.Tensor:
        h, w = img.shape[-2:]
        cy = int(torch.randint(h, (1,), generator=self.generator))
        cx = int(torch.randint(w, (1,), generator=self.generator))
        half = self.length // 2
        out = img.clone()
        out[..., max(0, cy - half):min(h, cy + half), max(0, cx - half):min(w, cx + half)] = 0
        return out


class TrainTransform:
    """uint8 HWC image -> float CHW: pad-and-crop, horizontal flip with probability 1/2, normalise, Cutout."""

    def __init__(self, mean, std, size: int = 32, pad: int = 4, cutout: int = 16, generator: Optional[torch.Generator] = None):
        self.mean = torch.tensor(mean, dtype=torch.float32).view(3, 1, 1)
        self.std = torch.tensor(std, dtype=torch.float32).view(3, 1, 1)
        self.size, self.pad, self.generator = size, pad, generator
        self.cutout = Cutout(cutout, generator) if cutout else None

    def __call__(self, img) -> torch.Tensor:
        x = torch.as_tensor(np.asarray(img)).permute(2, 0, 1).float() / 255.0
        if self.pad:
            x = torch.nn.functional.pad(x, (self.pad,) * 4)
            top = int(torch.randint(x.shape[1] - self.size + 1, (1,), generator=self.generator))
            left = int(torch.randint(x.shape[2] - self.size + 1, (1,), generator=self.generator))
            x = x[:, top:top + self.size, left:left + self.size]
        if float(torch.rand(1, generator=self.generator)) < 0.5:
            x = x.flip(-1)
        x = (x - self.mean) / self.std
        return self.cutout(x) if self.cutout else x


class TestTransform:
    def __init__(self, mean, std):
        self.mean = torch.tensor(mean, dtype=torch.float32).view(3, 1, 1)
        self.std = torch.tensor(std, dtype=torch.float32).view(3, 1, 1)

    def __call__(self, img) -> torch.Tensor:
        return (torch.as_tensor(np.asarray(img)).permute(2, 0, 1).float() / 255.0 - self.mean) / self.std


def data_transforms(dataset: str = "cifar10"):
    """``(train_transform, test_transform)`` with the reference's constants for ``dataset``."""
    mean, std, _ = _STATS[dataset.lower()]
    return TrainTransform(mean, std), TestTransform(mean, std)


# ------------------------------------------------------------------------------------------------ arrays
_ARRAY_CACHE: Dict[tuple, Tuple[np.ndarray, ...]] = {}
SYNTHETIC_TRAIN_SIZE = 0          # 0 = the real set's size (50,000 / 90,000); smaller values keep tests and demos light


def load_image_arrays(dataset: str, datadir: Optional[str]) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
    """``X_train [n, 32, 32, 3] uint8, y_train, X_test, y_test`` - the arrays the reference takes from torchvision's CIFAR classes
    (``cifar10_train_ds.data / .target``).  Real pickles / image folders under ``datadir`` when present, else synthetic."""
    name = dataset.lower()
    key = (name, str(datadir), SYNTHETIC_TRAIN_SIZE)
    if key not in _ARRAY_CACHE:
        real = None
        if datadir:
            from . import _try_real_cifar10, readers as R
            if name == "cifar10":
                real = _try_real_cifar10(datadir)
            elif name == "cifar100":
                real = R.load_cifar100_pickles(datadir)
            elif name == "cinic10":
                import os
                folder = R.load_image_folder(os.path.join(datadir, "cinic10"), 32, "train", "test")
                real = folder[:4] if folder is not None else None
        if real is None:
            classes = _STATS[name][2]
            n = SYNTHETIC_TRAIN_SIZE or (90000 if name == "cinic10" else 50000)
            real = S.make_image_dataset(n, n // 5, classes, 32, seed=0)
        _ARRAY_CACHE[key] = tuple(np.asarray(t) for t in real)
    return _ARRAY_CACHE[key]


def load_cifar10_data(datadir):
    return load_image_arrays("cifar10", datadir)


def load_cifar100_data(datadir):
    return load_image_arrays("cifar100", datadir)


def load_cinic10_data(datadir):
    return load_image_arrays("cinic10", datadir)


class TruncatedImageDataset(torch_data.Dataset):
    """A client's view of an image set: ``dataidxs`` selects rows; ``.data`` / ``.target`` are the selected arrays like in the
    reference's ``CIFAR10_truncated`` (data/cifar10/datasets.py:35-91)."""

    dataset_name = "cifar10"

    def __init__(self, root=None, dataidxs: Optional[Sequence[int]] = None, train: bool = True, transform=None,
                 target_transform=None, download: bool = False):
        self.root, self.dataidxs, self.train = root, dataidxs, train
        self.transform, self.target_transform = transform, target_transform
        xtr, ytr, xte, yte = load_image_arrays(self.dataset_name, root)
        x, y = (xtr, ytr) if train else (xte, yte)
        if dataidxs is not None:
            idx = np.asarray(dataidxs, dtype=np.int64)
            x, y = x[idx], y[idx]
        self.data, self.target = x, np.asarray(y)

    def truncate_channel(self, index):
        """Zeroes the green and blue channels of the given rows (the reference's helper of the same name)."""
        self.data = self.data.copy()
        for i in np.asarray(index).reshape(-1):
            self.data[i, :, :, 1:] = 0

    def __getitem__(self, index):
        img, target = self.data[index], int(self.target[index])
        if self.transform is not None:
            img = self.transform(img)
        if self.target_transform is not None:
            target = self.target_transform(target)
        return img, target

    def __len__(self):
        return len(self.data)


class CIFAR10_truncated(TruncatedImageDataset):  # noqa: N801
    dataset_name = "cifar10"


class CIFAR100_truncated(TruncatedImageDataset):  # noqa: N801
    dataset_name = "cifar100"


class CINIC10_truncated(TruncatedImageDataset):  # noqa: N801
    dataset_name = "cinic10"


_TRUNCATED = {"cifar10": CIFAR10_truncated, "cifar100": CIFAR100_truncated, "cinic10": CINIC10_truncated}


# ------------------------------------------------------------------------------------------------ partitions
def record_net_data_stats(y_train, net_dataidx_map) -> Dict[int, Dict[int, int]]:
    """``{client: {class: count}}``."""
    y = np.asarray(y_train)
    out = {}
    for client, idx in net_dataidx_map.items():
        classes, counts = np.unique(y[np.asarray(idx, dtype=np.int64)], return_counts=True)
        out[client] = {int(c): int(n) for c, n in zip(classes, counts)}
    logging.debug("Data statistics: %s", out)
    return out


def _parse_nested_text(filename: str):
    """The frozen-split files are Python / JSON dict literals spread over several lines (``{ 0: [ 1, 2, ...], ...}`` and
    ``{ 0: { 3: 17, ...}, ...}``); parse them as such instead of line by line."""
    with open(filename) as fh:
        text = fh.read()
    return ast.literal_eval(text)


def read_net_dataidx_map(filename="./data_preprocessing/non-iid-distribution/CIFAR10/net_dataidx_map.txt") -> Dict[int, List[int]]:
    return {int(k): [int(i) for i in v] for k, v in _parse_nested_text(filename).items()}


def read_data_distribution(filename="./data_preprocessing/non-iid-distribution/CIFAR10/distribution.txt") -> Dict[int, Dict[int, int]]:
    return {int(k): {int(c): int(n) for c, n in v.items()} for k, v in _parse_nested_text(filename).items()}


def write_net_dataidx_map(filename: str, net_dataidx_map) -> str:
    """Inverse of :func:`read_net_dataidx_map`, in the layout the reference's files have (one key per block)."""
    with open(filename, "w") as fh:
        fh.write("{\n")
        for k, idx in net_dataidx_map.items():
            fh.write(f"{int(k)}: [\n" + ", ".join(str(int(i)) for i in idx) + "\n],\n")
        fh.write("}\n")
    return filename


def write_data_distribution(filename: str, distribution) -> str:
    with open(filename, "w") as fh:
        fh.write("{\n")
        for k, per_class in distribution.items():
            fh.write(f"{int(k)}: {{\n" + "".join(f"{int(c)}: {int(n)},\n" for c, n in per_class.items()) + "},\n")
        fh.write("}\n")
    return filename


def partition_data(dataset, datadir, partition, n_nets, alpha, dataidx_map_file: Optional[str] = None,
                   distribution_file: Optional[str] = None):
    """``(X_train, y_train, X_test, y_test, net_dataidx_map, traindata_cls_counts)``.  ``partition``: "homo" (a permutation cut
    into ``n_nets`` parts), "hetero" (per-class Dirichlet(alpha) with the balance rule, repeated until every client has at least
    10 samples) - both under ``np.random.seed(10)`` like the reference - or "hetero-fix" (the split stored in
    ``dataidx_map_file`` / ``distribution_file``)."""
    from ..core.non_iid_partition import homo_partition, non_iid_partition_with_dirichlet_distribution
    if str(dataset).lower() not in _STATS:
        raise ValueError(f"partition_data: unknown image set {dataset!r} (known: {sorted(_STATS)})")
    x_train, y_train, x_test, y_test = load_image_arrays(dataset, datadir)
    classes = _STATS[str(dataset).lower()][2]
    if partition == "hetero-fix":
        if not dataidx_map_file:
            raise ValueError("partition 'hetero-fix' needs dataidx_map_file=<net_dataidx_map.txt>")
        net_dataidx_map = read_net_dataidx_map(dataidx_map_file)
        counts = read_data_distribution(distribution_file) if distribution_file else record_net_data_stats(y_train, net_dataidx_map)
        return x_train, y_train, x_test, y_test, net_dataidx_map, counts
    state = np.random.get_state()
    np.random.seed(10)
    try:
        if partition == "homo":
            net_dataidx_map = homo_partition(len(y_train), n_nets)
        elif partition == "hetero":
            net_dataidx_map = non_iid_partition_with_dirichlet_distribution(np.asarray(y_train), n_nets, classes, alpha)
        else:
            raise ValueError(f"unknown partition {partition!r} (homo | hetero | hetero-fix)")
    finally:
        np.random.set_state(state)
    return x_train, y_train, x_test, y_test, net_dataidx_map, record_net_data_stats(y_train, net_dataidx_map)


# ------------------------------------------------------------------------------------------------ loaders
def get_dataloader(dataset, datadir, train_bs, test_bs, dataidxs=None):
    """Train loader over ``dataidxs`` (all samples when None; shuffled, augmented, last short batch dropped) and the loader
    of the whole test set."""
    return get_dataloader_test(dataset, datadir, train_bs, test_bs, dataidxs, None)


def get_dataloader_test(dataset, datadir, train_bs, test_bs, dataidxs_train=None, dataidxs_test=None):
    cls = _TRUNCATED[str(dataset).lower()]
    t_train, t_test = data_transforms(dataset)
    train_ds = cls(datadir, dataidxs=dataidxs_train, train=True, transform=t_train)
    test_ds = cls(datadir, dataidxs=dataidxs_test, train=False, transform=t_test)
    train_dl = torch_data.DataLoader(train_ds, batch_size=train_bs, shuffle=True, drop_last=True)
    test_dl = torch_data.DataLoader(test_ds, batch_size=test_bs, shuffle=False, drop_last=True)
    return train_dl, test_dl


# ---- the naturally federated sets: the fifth argument is a client index, not a list of sample indices
_FEDERATED_CACHE: Dict[tuple, object] = {}
FED_EMNIST_FILES = ("fed_emnist_train.h5", "fed_emnist_test.h5")
FED_EMNIST_CLIENTS = (3400, 3400)
FED_CIFAR100_FILES = ("fed_cifar100_train.h5", "fed_cifar100_test.h5")
FED_CIFAR100_CLIENTS = (500, 100)
FED_IMAGE_BATCH_SIZE = 20


def _federated(name: str, data_dir):
    key = (name, str(data_dir), SYNTHETIC_TRAIN_SIZE)
    if key not in _FEDERATED_CACHE:
        from . import load_federated
        from ..arguments import Arguments
        kw = dict(dataset=name, log_level="WARNING")
        if data_dir:
            kw["data_cache_dir"] = data_dir
        if SYNTHETIC_TRAIN_SIZE:
            kw["synthetic_train_size"] = SYNTHETIC_TRAIN_SIZE
        _FEDERATED_CACHE[key] = load_federated(Arguments.from_dict(kw))
    return _FEDERATED_CACHE[key]


def federated_dataloader(name: str):
    """``get_dataloader(dataset, data_dir, train_bs, test_bs, client_idx=None)`` of a naturally federated set (reference:
    data/FederatedEMNIST/data_loader.py:30-75 and the same function in fed_cifar100 / fed_shakespeare / stackoverflow_*): the
    loaders of one client, or of the union of all clients when ``client_idx`` is None.  ``dataset`` is ignored like there."""
    def get_dataloader(dataset, data_dir, train_bs, test_bs, client_idx=None):
        fed = _federated(name, data_dir)
        if client_idx is None:
            (x, y), (xt, yt) = (fed.x, fed.y), (fed.x_test, fed.y_test)
        else:
            (x, y), (xt, yt) = fed.client_slice(int(client_idx)), fed.client_test_slice(int(client_idx))
        prep = fed.transform if fed.transform is not None else (lambda t: t)
        mk = lambda a, b, bs, sh: torch_data.DataLoader(torch_data.TensorDataset(prep(a), b), batch_size=bs, shuffle=sh) if len(a) else None  # noqa: E731
        return mk(x, y, train_bs, True), mk(xt, yt, test_bs, False)

    get_dataloader.__name__ = f"get_dataloader_{name}"
    return get_dataloader


get_dataloader_femnist = federated_dataloader("femnist")
get_dataloader_fed_cifar100 = federated_dataloader("fed_cifar100")
get_dataloader_fed_shakespeare = federated_dataloader("fed_shakespeare")
get_dataloader_stackoverflow_nwp = federated_dataloader("stackoverflow_nwp")
get_dataloader_stackoverflow_lr = federated_dataloader("stackoverflow_lr")


def _fixed(name: str, fn):
    """The reference's per-module functions take a ``dataset`` argument and ignore it; bind the module's own set instead."""
    def bound(dataset, *a, **k):
        return fn(name, *a, **k)
    bound.__name__ = f"{fn.__name__}_{name}"
    bound.__doc__ = fn.__doc__
    return bound


def get_dataloader_CIFAR10(datadir, train_bs, test_bs, dataidxs=None):  # noqa: N802
    return get_dataloader("cifar10", datadir, train_bs, test_bs, dataidxs)


def get_dataloader_CIFAR100(datadir, train_bs, test_bs, dataidxs=None):  # noqa: N802
    return get_dataloader("cifar100", datadir, train_bs, test_bs, dataidxs)


def get_dataloader_cinic10(datadir, train_bs, test_bs, dataidxs=None):
    return get_dataloader("cinic10", datadir, train_bs, test_bs, dataidxs)


def get_dataloader_test_CIFAR10(datadir, train_bs, test_bs, dataidxs_train=None, dataidxs_test=None):  # noqa: N802
    return get_dataloader_test("cifar10", datadir, train_bs, test_bs, dataidxs_train, dataidxs_test)


def get_dataloader_test_CIFAR100(datadir, train_bs, test_bs, dataidxs_train=None, dataidxs_test=None):  # noqa: N802
    return get_dataloader_test("cifar100", datadir, train_bs, test_bs, dataidxs_train, dataidxs_test)


def get_dataloader_test_cinic10(datadir, train_bs, test_bs, dataidxs_train=None, dataidxs_test=None):
    return get_dataloader_test("cinic10", datadir, train_bs, test_bs, dataidxs_train, dataidxs_test)


partition_data_cifar10, partition_data_cifar100, partition_data_cinic10 = (_fixed(n, partition_data) for n in ("cifar10", "cifar100", "cinic10"))
get_dataloader_any_cifar10, get_dataloader_any_cifar100, get_dataloader_any_cinic10 = (
    _fixed(n, get_dataloader) for n in ("cifar10", "cifar100", "cinic10"))
get_dataloader_test_any_cifar10, get_dataloader_test_any_cifar100, get_dataloader_test_any_cinic10 = (
    _fixed(n, get_dataloader_test) for n in ("cifar10", "cifar100", "cinic10"))


# ------------------------------------------------------------------------------------------------ fed_cifar100 (TFF recipe)
def cifar100_transform(img_mean, img_std, train=True, crop_size=(24, 24), generator: Optional[torch.Generator] = None):
    """Callable on one float CHW image in [0, 1]: random crop + horizontal flip (train) or centre crop (test) to ``crop_size``,
    then ``(x - img_mean) / img_std`` (reference data/fed_cifar100/utils.py:9-30; the TFF recipe standardises every image with
    its own statistics)."""
    ch, cw = crop_size

    def apply(x: torch.Tensor) -> torch.Tensor:
        h, w = x.shape[-2:]
        if train:
            top = int(torch.randint(h - ch + 1, (1,), generator=generator))
            left = int(torch.randint(w - cw + 1, (1,), generator=generator))
            x = x[..., top:top + ch, left:left + cw]
            if float(torch.rand(1, generator=generator)) < 0.5:
                x = x.flip(-1)
        else:
            top, left = int(round((h - ch) / 2.0)), int(round((w - cw) / 2.0))
            x = x[..., top:top + ch, left:left + cw]
        return (x - float(img_mean)) / float(img_std)

    return apply


def preprocess_cifar_img(img: torch.Tensor, train: bool, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """``[n, 32, 32, 3]`` pixel values 0..255 -> ``[n, 3, 24, 24]`` float: scaled to [0, 1], cropped / flipped, each image
    standardised by the mean and (unbiased) standard deviation of its own 32x32x3 values."""
    x = torch.as_tensor(img).double() / 255.0
    out = []
    for i in x:
        t = cifar100_transform(i.mean(), i.std(), train, generator=generator)
        out.append(t(i.permute(2, 0, 1)))
    return torch.stack(out).float() if out else torch.zeros((0, 3, 24, 24))
