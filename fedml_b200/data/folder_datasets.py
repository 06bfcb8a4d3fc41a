"""Lazy, file-backed ``torch.utils.data.Dataset``s for the large image sets: ImageNet folders, Google Landmarks, generic image
folders restricted to a client's indices.

The packed store (``FederatedDataset``) holds a whole data set in host memory, which is right for CIFAR-sized sets and for the
synthetic stand-ins; ImageNet (1.28 M JPEGs) and Landmarks are read from disk per item instead, like in the reference
(data/ImageNet/datasets.py:14-180, data/Landmarks/datasets.py:8-50 + data_loader.py:121-160, data/cinic10/datasets.py
``ImageFolderTruncated``).  Clients are index ranges into one sorted file list:

* ImageNet: one "client" per class - ``net_dataidx_map[class] = (begin, end)``; ``dataidxs`` is a class id, a list of them or None;
* Landmarks: one client per ``user_id`` of the mapping CSV - ``dataidxs`` is that user's ``(begin, end)`` range;
* ``ImageFolderTruncated``: ``dataidxs`` is an explicit list of sample indices.
"""
from __future__ import annotations

import collections
import csv
import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch.utils import data as torch_data

IMG_EXTENSIONS = (".jpg", ".jpeg", ".png", ".ppm", ".bmp", ".pgm", ".tif", ".tiff", ".webp")


def has_file_allowed_extension(filename: str, extensions: Sequence[str] = IMG_EXTENSIONS) -> bool:
    return filename.lower().endswith(tuple(extensions))


def find_classes(dir):  # noqa: A002
    """Sorted sub-directory names and their indices."""
    classes = sorted(d for d in os.listdir(dir) if os.path.isdir(os.path.join(dir, d)))
    return classes, {c: i for i, c in enumerate(classes)}


def make_dataset(dir, class_to_idx: Dict[str, int], extensions: Sequence[str] = IMG_EXTENSIONS):  # noqa: A002
    """``([(path, class index)], {class: count}, {class: (begin, end)})`` over ``dir/<class>/**`` in sorted order."""
    images: List[Tuple[str, int]] = []
    counts: Dict[int, int] = {}
    ranges: Dict[int, Tuple[int, int]] = {}
    root = os.path.expanduser(dir)
    for target in sorted(os.listdir(root)):
        d = os.path.join(root, target)
        if not os.path.isdir(d) or target not in class_to_idx:
            continue
        begin = len(images)
        for sub, _, fnames in sorted(os.walk(d)):
            images.extend((os.path.join(sub, f), class_to_idx[target]) for f in sorted(fnames) if has_file_allowed_extension(f, extensions))
        counts[class_to_idx[target]] = len(images) - begin
        ranges[class_to_idx[target]] = (begin, len(images))
    return images, counts, ranges


def pil_loader(path):
    from PIL import Image
    with open(path, "rb") as fh:
        return Image.open(fh).convert("RGB")


def accimage_loader(path):
    """accimage when that package exists (it is optional in torchvision), PIL otherwise."""
    try:
        import accimage
        return accimage.Image(path)
    except (ImportError, OSError):
        return pil_loader(path)


def default_loader(path):
    try:
        from torchvision import get_image_backend
        if get_image_backend() == "accimage":
            return accimage_loader(path)
    except ImportError:
        pass
    return pil_loader(path)


def _to_tensor(img) -> torch.Tensor:
    """PIL image -> float CHW in [0, 1] (used when no transform is given, so that default collation works)."""
    a = np.asarray(img, dtype=np.uint8)
    if a.ndim == 2:
        a = a[:, :, None]
    return torch.from_numpy(a.copy()).permute(2, 0, 1).float().div_(255.0)


class _FileDataset(torch_data.Dataset):
    loader = staticmethod(default_loader)

    def _item(self, path, target):
        img = self.loader(path)
        img = self.transform(img) if self.transform is not None else _to_tensor(img)
        if self.target_transform is not None:
            target = self.target_transform(target)
        return img, target


class ImageNet(_FileDataset):
    """``<data_dir>/train|val/<wnid>/*.JPEG``.  Scanning the tree takes a while on the full set: build one instance and hand it
    to :class:`ImageNet_truncated` for the per-client views."""

    def __init__(self, data_dir, dataidxs=None, train=True, transform=None, target_transform=None, download=False):
        self.dataidxs, self.train, self.transform, self.target_transform = dataidxs, train, transform, target_transform
        self.data_dir = os.path.join(data_dir, "train" if train else "val")
        self.all_data, self.data_local_num_dict, self.net_dataidx_map = self.__getdatasets__()
        self.local_data = _select(self.all_data, self.net_dataidx_map, dataidxs)

    def __getdatasets__(self):
        classes, class_to_idx = find_classes(self.data_dir)
        self.classes, self.class_to_idx = classes, class_to_idx
        return make_dataset(self.data_dir, class_to_idx, IMG_EXTENSIONS)

    def get_local_data(self):
        return self.local_data

    def get_net_dataidx_map(self):
        return self.net_dataidx_map

    def get_data_local_num_dict(self):
        return self.data_local_num_dict

    def __getitem__(self, index):
        return self._item(*self.local_data[index])

    def __len__(self):
        return len(self.local_data)


def _select(all_data, net_dataidx_map, dataidxs):
    if dataidxs is None:
        return all_data
    if isinstance(dataidxs, (int, np.integer)):
        b, e = net_dataidx_map[int(dataidxs)]
        return all_data[b:e]
    out = []
    for k in dataidxs:
        b, e = net_dataidx_map[int(k)]
        out += all_data[b:e]
    return out


class ImageNet_truncated(_FileDataset):  # noqa: N801
    """View of an :class:`ImageNet` instance restricted to the classes in ``dataidxs`` (no second directory scan)."""

    def __init__(self, imagenet_dataset: ImageNet, dataidxs, net_dataidx_map=None, train=True, transform=None, target_transform=None,
                 download=False):
        self.dataidxs, self.train, self.transform, self.target_transform = dataidxs, train, transform, target_transform
        self.net_dataidx_map = net_dataidx_map or imagenet_dataset.get_net_dataidx_map()
        self.all_data = imagenet_dataset.get_local_data()
        self.local_data = _select(self.all_data, self.net_dataidx_map, dataidxs)

    def __getitem__(self, index):
        return self._item(*self.local_data[index])

    def __len__(self):
        return len(self.local_data)


class ImageFolderTruncated(_FileDataset):
    """``root/<class>/*`` with an explicit list of sample indices (the CINIC-10 / edge-case loaders' client view)."""

    def __init__(self, root, dataidxs=None, transform=None, target_transform=None, loader=None, extensions=IMG_EXTENSIONS):
        self.root, self.transform, self.target_transform = root, transform, target_transform
        if loader is not None:
            self.loader = loader
        self.classes, self.class_to_idx = find_classes(root)
        samples, _, _ = make_dataset(root, self.class_to_idx, extensions)
        if not samples:
            raise RuntimeError(f"Found 0 files in subfolders of: {root}")
        self.dataidxs = dataidxs
        self.imgs = samples if dataidxs is None else [samples[int(i)] for i in dataidxs]
        self.samples = self.imgs

    @property
    def get_train_labels(self):
        return np.array([t for _, t in self.imgs])

    def __getitem__(self, index):
        return self._item(*self.imgs[index])

    def __len__(self):
        return len(self.imgs)


# ------------------------------------------------------------------------------------------------ Landmarks
def get_mapping_per_user(fn: str):
    """Reads ``user_id,image_id,class`` rows -> ``(rows grouped by user, {user: count}, {user: (begin, end)})``; the users keep
    the order of first appearance."""
    with open(fn, newline="") as fh:
        rows = list(csv.DictReader(fh))
    if not rows or not all(c in rows[0] for c in ("user_id", "image_id", "class")):
        raise ValueError("The mapping file must contain user_id, image_id and class columns. The existing columns are "
                         + ",".join(rows[0].keys() if rows else []))
    per_user = collections.OrderedDict()
    for r in rows:
        per_user.setdefault(r["user_id"], []).append(r)
    files, counts, ranges = [], {}, {}
    for user, items in per_user.items():
        ranges[int(user)] = (len(files), len(files) + len(items))
        counts[int(user)] = len(items)
        files += items
    return files, counts, ranges


class Landmarks(_FileDataset):
    """``allfiles``: the row list of :func:`get_mapping_per_user`; ``dataidxs``: a ``(begin, end)`` range into it or None; images
    are ``<data_dir>/<image_id>.jpg``."""

    def __init__(self, data_dir, allfiles, dataidxs=None, train=True, transform=None, target_transform=None, download=False):
        self.allfiles = allfiles
        self.local_files = allfiles if dataidxs is None else allfiles[dataidxs[0]:dataidxs[1]]
        self.data_dir, self.dataidxs, self.transform, self.target_transform = data_dir, dataidxs, transform, target_transform

    def __len__(self):
        return len(self.local_files)

    def __getitem__(self, idx):
        row = self.local_files[idx]
        return self._item(os.path.join(self.data_dir, str(row["image_id"]) + ".jpg"), int(row["class"]))


# ------------------------------------------------------------------------------------------------ transforms and loaders
class _ResizeCropNormalize:
    """PIL image -> normalised float CHW of ``size`` x ``size``: random-resized crop + flip for training, resize + centre crop for
    evaluation (the usual ImageNet recipe, plain PIL / tensor code)."""

    def __init__(self, size, train, mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225), generator: Optional[torch.Generator] = None):
        self.size, self.train, self.generator = int(size), train, generator
        self.mean, self.std = torch.tensor(mean).view(3, 1, 1), torch.tensor(std).view(3, 1, 1)

    def _rand(self):
        return float(torch.rand(1, generator=self.generator))

    def __call__(self, img):
        w, h = img.size
        if self.train:
            for _ in range(10):
                area = w * h * (0.08 + 0.92 * self._rand())
                ratio = np.exp(np.log(3 / 4) + (np.log(4 / 3) - np.log(3 / 4)) * self._rand())
                cw, ch = int(round(np.sqrt(area * ratio))), int(round(np.sqrt(area / ratio)))
                if 0 < cw <= w and 0 < ch <= h:
                    x0, y0 = int(self._rand() * (w - cw + 1)), int(self._rand() * (h - ch + 1))
                    img = img.crop((x0, y0, x0 + cw, y0 + ch))
                    break
            img = img.resize((self.size, self.size))
            t = _to_tensor(img)
            if self._rand() < 0.5:
                t = t.flip(-1)
        else:
            short = int(round(self.size / 0.875))
            scale = short / min(w, h)
            img = img.resize((max(self.size, int(round(w * scale))), max(self.size, int(round(h * scale)))))
            w, h = img.size
            x0, y0 = (w - self.size) // 2, (h - self.size) // 2
            t = _to_tensor(img.crop((x0, y0, x0 + self.size, y0 + self.size)))
        return (t - self.mean) / self.std


def image_transforms(size: int = 224):
    return _ResizeCropNormalize(size, True), _ResizeCropNormalize(size, False)


def get_dataloader_ImageNet(datadir, train_bs, test_bs, dataidxs=None, image_size: int = 224, num_workers: int = 0):  # noqa: N802
    t_train, t_test = image_transforms(image_size)
    train_ds = ImageNet(datadir, dataidxs=dataidxs, train=True, transform=t_train)
    test_ds = ImageNet(datadir, dataidxs=None, train=False, transform=t_test)
    return (torch_data.DataLoader(train_ds, batch_size=train_bs, shuffle=True, drop_last=False, num_workers=num_workers),
            torch_data.DataLoader(test_ds, batch_size=test_bs, shuffle=False, drop_last=False, num_workers=num_workers))


def get_dataloader_ImageNet_truncated(imagenet_dataset_train, imagenet_dataset_test, train_bs, test_bs, dataidxs=None,  # noqa: N802
                                      net_dataidx_map=None, image_size: int = 224, num_workers: int = 0):
    t_train, t_test = image_transforms(image_size)
    train_ds = ImageNet_truncated(imagenet_dataset_train, dataidxs, net_dataidx_map, train=True, transform=t_train)
    test_ds = ImageNet_truncated(imagenet_dataset_test, None, imagenet_dataset_test.get_net_dataidx_map(), train=False, transform=t_test)
    return (torch_data.DataLoader(train_ds, batch_size=train_bs, shuffle=True, num_workers=num_workers),
            torch_data.DataLoader(test_ds, batch_size=test_bs, shuffle=False, num_workers=num_workers))


def get_dataloader_Landmarks(datadir, train_files, test_files, train_bs, test_bs, dataidxs=None, image_size: int = 224,  # noqa: N802
                             num_workers: int = 0):
    t_train, t_test = image_transforms(image_size)
    train_ds = Landmarks(datadir, train_files, dataidxs=dataidxs, train=True, transform=t_train)
    test_ds = Landmarks(datadir, test_files, dataidxs=None, train=False, transform=t_test)
    return (torch_data.DataLoader(train_ds, batch_size=train_bs, shuffle=True, num_workers=num_workers),
            torch_data.DataLoader(test_ds, batch_size=test_bs, shuffle=False, num_workers=num_workers))


# ------------------------------------------------------------------------------------------------ ImageNet in one HDF5 file
class DatasetHDF5(torch_data.Dataset):
    """``<t>_img`` (uint8 ``[n, H, W, 3]``) and ``<t>_labels`` of one HDF5 file, ``t`` = "train" or "val"; images are read one row
    at a time (reference data/ImageNet/datasets_hdf5.py:11-35).  Opened with h5py when installed, else with ``minihdf5``."""

    def __init__(self, hdf5fn, t, transform=None, target_transform=None):
        from .readers import open_h5
        self.hf, self.t = open_h5(hdf5fn), t
        self.d = self.hf[f"{t}_img"]
        self.n_images = self.d.shape[0]
        self.dlabel = np.asarray(self.hf[f"{t}_labels"][...])

    def __getitem__(self, index):
        return np.asarray(self.d[int(index)]), np.int64(self.dlabel[index])

    def __len__(self):
        return self.n_images


class ImageNet_hdf5(_FileDataset):  # noqa: N801
    """Per-class clients over a :class:`DatasetHDF5`: ``net_dataidx_map[class] = [sample indices]``."""

    def __init__(self, data_dir, dataidxs=None, train=True, transform=None, target_transform=None, download=False):
        self.dataidxs, self.train, self.transform, self.target_transform = dataidxs, train, transform, target_transform
        self.hdf5fn = os.path.join(data_dir)
        self.all_data_hdf5 = DatasetHDF5(self.hdf5fn, "train" if train else "val")
        self.net_dataidx_map: Dict[int, List[int]] = {}
        for i, label in enumerate(self.all_data_hdf5.dlabel):
            self.net_dataidx_map.setdefault(int(label), []).append(i)
        self.data_local_num_dict = {k: len(v) for k, v in self.net_dataidx_map.items()}
        self.all_data_idx = range(len(self.all_data_hdf5))
        self.local_data_idx = _select_indices(self.all_data_idx, self.net_dataidx_map, dataidxs)

    def get_net_dataidx_map(self):
        return self.net_dataidx_map

    def get_data_local_num_dict(self):
        return self.data_local_num_dict

    def _decode(self, img, target):
        from PIL import Image
        img = Image.fromarray(np.asarray(img, dtype=np.uint8))
        img = self.transform(img) if self.transform is not None else _to_tensor(img)
        if self.target_transform is not None:
            target = self.target_transform(target)
        return img, int(target)

    def __getitem__(self, index):
        return self._decode(*self.all_data_hdf5[self.local_data_idx[index]])

    def __len__(self):
        return len(self.local_data_idx)


def _select_indices(all_idx, net_dataidx_map, dataidxs):
    if dataidxs is None:
        return all_idx
    if isinstance(dataidxs, (int, np.integer)):
        return net_dataidx_map[int(dataidxs)]
    out: List[int] = []
    for k in dataidxs:
        out += net_dataidx_map[int(k)]
    return out


class ImageNet_truncated_hdf5(ImageNet_hdf5):  # noqa: N801
    """View of an :class:`ImageNet_hdf5` restricted to the classes in ``dataidxs`` (shares the open file and the class map)."""

    def __init__(self, imagenet_dataset: ImageNet_hdf5, dataidxs, net_dataidx_map=None, train=True, transform=None,
                 target_transform=None, download=False):
        self.dataidxs, self.train, self.transform, self.target_transform = dataidxs, train, transform, target_transform
        self.all_data_hdf5 = imagenet_dataset.all_data_hdf5
        self.net_dataidx_map = net_dataidx_map or imagenet_dataset.net_dataidx_map
        self.data_local_num_dict = imagenet_dataset.data_local_num_dict
        self.all_data_idx = range(len(self.all_data_hdf5))
        self.local_data_idx = _select_indices(self.all_data_idx, self.net_dataidx_map, dataidxs)
