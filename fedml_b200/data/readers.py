"""Readers for the real dataset files of the reference benchmarks, producing the packed :class:`FederatedDataset`.

Formats (reference loaders in brackets):
  * LEAF JSON ``{users, user_data{user: {x, y}}}`` — MNIST (data/MNIST/data_loader.py:32-72), FEMNIST in its LEAF form,
    Shakespeare (data/shakespeare/data_loader.py:20-89; x = 80-character strings, y = next character, vocabulary
    data/shakespeare/language_utils.py:10-17);
  * TFF ``.h5`` exports — fed_emnist, fed_cifar100, fed_shakespeare, stackoverflow
    (data/FederatedEMNIST/data_loader.py:133, fed_cifar100/data_loader.py:132, fed_shakespeare/data_loader.py:117,
    stackoverflow_*/data_loader.py): need ``h5py``, which is not installable in this image — the reader raises an
    ImportError that names the package and the file it was asked for;
  * CIFAR-100 python pickles (data/cifar100/data_loader.py), image folders ``<root>/<split>/<class>/*.jpg|png`` — ImageNet
    (data/ImageNet/data_loader.py:273, one client per group of classes or a homo split) and CINIC-10;
  * Landmarks CSV ``user_id,image_id,class`` + an image directory (data/Landmarks/data_loader.py:267).
Every reader returns None when its files are not there, so ``load_federated`` falls back to the synthetic generator of the
same shape (there is no network to download anything).
"""
from __future__ import annotations

import csv
import json
import os
import pickle
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import synthetic as S
from .federated_dataset import FederatedDataset

# vocabulary of the reference's character models (language_utils.py:10-17)
CHAR_VOCAB = list("dhlptx@DHLPTX $(,048cgkoswCGKOSW[_#'/37;?bfjnrvzBFJNRVZ\"&*.26:\naeimquyAEIMQUY]!%)-159\r")
ALL_LETTERS = "".join(CHAR_VOCAB)
CHAR_VOCAB_SIZE = len(ALL_LETTERS) + 4


def _offsets(parts: Sequence) -> np.ndarray:
    return np.concatenate([[0], np.cumsum([len(p) for p in parts])]).astype(np.int64)


def need_h5py(path: str):
    try:
        import h5py  # noqa: F401
        return h5py
    except ImportError as exc:
        raise ImportError(f"reading {path} needs the 'h5py' package (TFF .h5 export), which is not installed; install h5py or "
                          "provide the dataset in LEAF JSON / pickle / image-folder form") from exc


# ---------------------------------------------------------------- LEAF JSON
def read_leaf_dir(d: str) -> Tuple[List[str], Dict[str, dict]]:
    users, data = [], {}
    for f in sorted(os.listdir(d)):
        if f.endswith(".json"):
            with open(os.path.join(d, f)) as fh:
                blob = json.load(fh)
            users.extend(blob["users"])
            data.update(blob["user_data"])
    return users, data


def load_leaf_images(root: str, sub: str, class_num: int, name: str, shape=None) -> Optional[FederatedDataset]:
    """MNIST / FEMNIST LEAF files under ``root/sub/{train,test}``: x rows are flattened float images."""
    tr, te = os.path.join(root, sub, "train"), os.path.join(root, sub, "test")
    if not (os.path.isdir(tr) and os.path.isdir(te)):
        return None
    users, trd = read_leaf_dir(tr)
    _, ted = read_leaf_dir(te)
    users = [u for u in users if u in ted] or users
    xs = [np.asarray(trd[u]["x"], dtype=np.float32) for u in users]
    ys = [np.asarray(trd[u]["y"], dtype=np.int64) for u in users]
    txs = [np.asarray(ted[u]["x"], dtype=np.float32) for u in users if u in ted]
    tys = [np.asarray(ted[u]["y"], dtype=np.int64) for u in users if u in ted]
    x, xt = np.concatenate(xs), np.concatenate(txs)
    if shape is not None:
        x, xt = x.reshape((-1,) + tuple(shape)), xt.reshape((-1,) + tuple(shape))
    return FederatedDataset(torch.from_numpy(x), torch.from_numpy(np.concatenate(ys)), _offsets(ys), torch.from_numpy(xt),
                            torch.from_numpy(np.concatenate(tys)), _offsets(tys), class_num, None, name)


def word_to_indices(word: str) -> List[int]:
    return [ALL_LETTERS.find(c) for c in word]


def load_leaf_shakespeare(root: str) -> Optional[FederatedDataset]:
    """LEAF Shakespeare: every sample is (80 characters, next character); one client per speaking role."""
    tr, te = os.path.join(root, "shakespeare", "train"), os.path.join(root, "shakespeare", "test")
    if not (os.path.isdir(tr) and os.path.isdir(te)):
        return None
    users, trd = read_leaf_dir(tr)
    _, ted = read_leaf_dir(te)
    enc = lambda rows: np.asarray([word_to_indices(r) for r in rows], dtype=np.int64)      # noqa: E731
    lab = lambda rows: np.asarray([ALL_LETTERS.find(c) for c in rows], dtype=np.int64)      # noqa: E731
    xs = [enc(trd[u]["x"]) for u in users]
    ys = [lab(trd[u]["y"]) for u in users]
    txs = [enc(ted[u]["x"]) for u in users if u in ted]
    tys = [lab(ted[u]["y"]) for u in users if u in ted]
    return FederatedDataset(torch.from_numpy(np.concatenate(xs)), torch.from_numpy(np.concatenate(ys)), _offsets(ys),
                            torch.from_numpy(np.concatenate(txs)), torch.from_numpy(np.concatenate(tys)), _offsets(tys),
                            CHAR_VOCAB_SIZE, None, "shakespeare", task="next_char")


# ---------------------------------------------------------------- TFF h5 exports
def load_tff_h5(train_path: str, test_path: str, x_key: str, y_key: str, class_num: int, name: str, transform=None,
                to_uint8: bool = False) -> Optional[FederatedDataset]:
    """``examples/<client>/<x_key|y_key>`` groups (fed_emnist: pixels/label, fed_cifar100: image/label)."""
    if not (os.path.isfile(train_path) and os.path.isfile(test_path)):
        return None
    h5py = need_h5py(train_path)
    with h5py.File(train_path, "r") as ftr, h5py.File(test_path, "r") as fte:
        ids = sorted(ftr["examples"].keys())
        xs = [np.asarray(ftr["examples"][c][x_key][()]) for c in ids]
        ys = [np.asarray(ftr["examples"][c][y_key][()]).astype(np.int64).reshape(-1) for c in ids]
        tids = [c for c in ids if c in fte["examples"]]
        txs = [np.asarray(fte["examples"][c][x_key][()]) for c in tids]
        tys = [np.asarray(fte["examples"][c][y_key][()]).astype(np.int64).reshape(-1) for c in tids]
    x, xt = np.concatenate(xs), np.concatenate(txs)
    if to_uint8 and x.dtype != np.uint8:
        x, xt = x.astype(np.uint8), xt.astype(np.uint8)
    return FederatedDataset(torch.from_numpy(x), torch.from_numpy(np.concatenate(ys)), _offsets(ys), torch.from_numpy(xt),
                            torch.from_numpy(np.concatenate(tys)), _offsets(tys) if len(tids) == len(ids) else None,
                            class_num, transform, name)


# ---------------------------------------------------------------- pickles / image folders
def load_cifar100_pickles(root: str) -> Optional[Tuple[torch.Tensor, ...]]:
    d = os.path.join(root, "cifar-100-python")
    if not os.path.isdir(d):
        return None

    def rd(name):
        with open(os.path.join(d, name), "rb") as fh:
            blob = pickle.load(fh, encoding="bytes")
        x = np.asarray(blob[b"data"], dtype=np.uint8).reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1)
        return np.ascontiguousarray(x), np.asarray(blob[b"fine_labels"], dtype=np.int64)

    (xtr, ytr), (xte, yte) = rd("train"), rd("test")
    return tuple(torch.from_numpy(a) for a in (xtr, ytr, xte, yte))


_IMG_EXT = (".jpg", ".jpeg", ".png", ".bmp", ".ppm")


def _read_image(path: str, size: int) -> np.ndarray:
    from PIL import Image
    with Image.open(path) as im:
        im = im.convert("RGB")
        if im.size != (size, size):
            im = im.resize((size, size))
        return np.asarray(im, dtype=np.uint8)


def load_image_folder(root: str, size: int = 224, train_dir: str = "train", val_dir: str = "val",
                      max_per_class: int = 0) -> Optional[Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor, List[str]]]:
    """``root/train/<class>/*`` and ``root/val/<class>/*`` (ImageNet / CINIC-10 layout) -> uint8 NHWC tensors."""
    tr, va = os.path.join(root, train_dir), os.path.join(root, val_dir)
    if not os.path.isdir(tr):
        return None
    classes = sorted(d for d in os.listdir(tr) if os.path.isdir(os.path.join(tr, d)))
    if not classes:
        return None

    def read_split(base):
        xs, ys = [], []
        for ci, cname in enumerate(classes):
            cdir = os.path.join(base, cname)
            if not os.path.isdir(cdir):
                continue
            files = sorted(f for f in os.listdir(cdir) if f.lower().endswith(_IMG_EXT))
            if max_per_class:
                files = files[:max_per_class]
            for f in files:
                xs.append(_read_image(os.path.join(cdir, f), size))
                ys.append(ci)
        if not xs:
            return torch.zeros((0, size, size, 3), dtype=torch.uint8), torch.zeros((0,), dtype=torch.int64)
        return torch.from_numpy(np.stack(xs)), torch.tensor(ys, dtype=torch.int64)

    xtr, ytr = read_split(tr)
    xva, yva = read_split(va) if os.path.isdir(va) else (xtr[:0], ytr[:0])
    return xtr, ytr, xva, yva, classes


def load_landmarks_csv(image_dir: str, train_csv: str, test_csv: str, size: int = 224) -> Optional[FederatedDataset]:
    """Google Landmarks federated split: CSV rows ``user_id,image_id,class`` (data/Landmarks/data_loader.py:20-22, 267),
    images ``<image_dir>/<image_id>.jpg``; one client per user_id."""
    if not (os.path.isfile(train_csv) and os.path.isfile(test_csv) and os.path.isdir(image_dir)):
        return None

    def rows(path):
        with open(path, newline="") as fh:
            return list(csv.DictReader(fh))

    per_user: Dict[str, List[dict]] = {}
    for r in rows(train_csv):
        per_user.setdefault(r["user_id"], []).append(r)
    users = sorted(per_user, key=lambda u: (len(u), u))
    img = lambda r: _read_image(os.path.join(image_dir, r["image_id"] + ".jpg"), size)      # noqa: E731
    xs = [np.stack([img(r) for r in per_user[u]]) for u in users]
    ys = [np.asarray([int(r["class"]) for r in per_user[u]], dtype=np.int64) for u in users]
    te = rows(test_csv)
    xt = np.stack([img(r) for r in te]) if te else np.zeros((0, size, size, 3), dtype=np.uint8)
    yt = np.asarray([int(r["class"]) for r in te], dtype=np.int64)
    class_num = int(max(int(np.concatenate(ys).max()), int(yt.max()) if len(yt) else 0)) + 1
    return FederatedDataset(torch.from_numpy(np.concatenate(xs)), torch.from_numpy(np.concatenate(ys)), _offsets(ys),
                            torch.from_numpy(xt), torch.from_numpy(yt), None, class_num,
                            S.ImageNormalize(S.IMAGENET_MEAN, S.IMAGENET_STD), "landmarks")
