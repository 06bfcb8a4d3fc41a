"""Writes a :class:`FederatedDataset` in the TensorFlow-Federated ``.h5`` layout the reference's loaders open.

The other direction of ``readers.load_tff_*``: with no network there is nothing to download, so benchmark files of the right
*shape* are produced from the synthetic generators - for the reference arm (its loaders only take these files:
data/FederatedEMNIST/data_loader.py:14-23, data/fed_cifar100/data_loader.py:15-24, data/fed_shakespeare/data_loader.py:12-20,
data/stackoverflow_nwp/data_loader.py:13-17) and for round-trip tests of the readers.

Layout: ``/examples/<client id>/<field>`` with one data set per field; strings are variable-length UTF-8.
    fed_emnist_{train,test}.h5        pixels float32 [n, 28, 28], label int32 [n]
    fed_cifar100_{train,test}.h5      image  uint8   [n, 32, 32, 3] (24x24 crops are taken at batch time), label int64 [n]
    shakespeare_{train,test}.h5       snippets  str [n]
    stackoverflow_{train,test}.h5     tokens, title, tags  str [n]   + stackoverflow.word_count / stackoverflow.tag_count
"""
from __future__ import annotations

import json
import os
from typing import Dict, List, Sequence

import numpy as np

from .federated_dataset import FederatedDataset
from .minihdf5 import write_hdf5
from .readers import CHAR_VOCAB


def client_ids(n: int, prefix: str = "f") -> List[str]:
    """Zero-padded so that the B-tree (byte) order is the numeric order."""
    width = max(4, len(str(max(0, n - 1))))
    return [f"{prefix}{i:0{width}d}" for i in range(n)]


def _per_client(ds: FederatedDataset, test: bool):
    for c in range(ds.client_num):
        if test:
            if ds.test_offsets is None:                          # one global test set: deal it out round-robin
                yield c, ds.x_test[c::ds.client_num], ds.y_test[c::ds.client_num]
            else:
                yield (c,) + tuple(ds.client_test_slice(c))
        else:
            yield (c,) + tuple(ds.client_slice(c))


def export_images(ds: FederatedDataset, out_dir: str, stem: str, x_key: str, y_key: str, x_dtype, y_dtype, x_shape=None) -> List[str]:
    ids = client_ids(ds.client_num)
    paths = []
    for split in ("train", "test"):
        tree: Dict[str, dict] = {}
        for c, x, y in _per_client(ds, split == "test"):
            xa = x.numpy()
            if x_shape is not None:
                xa = xa.reshape((-1,) + tuple(x_shape))
            tree[ids[c]] = {x_key: xa.astype(x_dtype), y_key: y.numpy().astype(y_dtype)}
        paths.append(write_hdf5(os.path.join(out_dir, f"{stem}_{split}.h5"), {"examples": tree}))
    return paths


def export_fed_emnist(ds: FederatedDataset, out_dir: str) -> List[str]:
    return export_images(ds, out_dir, "fed_emnist", "pixels", "label", np.float32, np.int32, (28, 28))


def export_fed_cifar100(ds: FederatedDataset, out_dir: str) -> List[str]:
    return export_images(ds, out_dir, "fed_cifar100", "image", "label", np.uint8, np.int64)


def ids_to_snippet(row: Sequence[int]) -> str:
    """Inverse of the character table for ids 1..86; padding / begin / end / out-of-vocabulary ids are dropped."""
    return "".join(CHAR_VOCAB[i - 1] for i in row if 1 <= i <= len(CHAR_VOCAB))


def export_fed_shakespeare(ds: FederatedDataset, out_dir: str) -> List[str]:
    """Each 80-id training row (plus its last target id) becomes one snippet."""
    ids = client_ids(ds.client_num, "PLAY_")
    paths = []
    for split in ("train", "test"):
        tree = {}
        for c, x, y in _per_client(ds, split == "test"):
            last = y[:, -1:] if y.dim() == 2 else y.reshape(-1, 1)
            rows = np.concatenate([x.numpy(), last.numpy()], axis=1)
            tree[ids[c]] = {"snippets": [ids_to_snippet(r) for r in rows] or [""]}
        paths.append(write_hdf5(os.path.join(out_dir, f"shakespeare_{split}.h5"), {"examples": tree}))
    return paths


def export_stackoverflow(ds: FederatedDataset, out_dir: str, vocab_size: int = 10000, tag_size: int = 500, seed: int = 0) -> List[str]:
    """From a next-word data set (ids as in ``readers.stackoverflow_nwp_ids``): sentence = words of ids 1..V, the title is the
    first three words, the tags are drawn from the sentence's word ids modulo ``tag_size`` (so they are learnable from the bag of
    words).  Also writes the two count tables."""
    words = [f"w{i:05d}" for i in range(vocab_size)]
    tags = [f"tag{i:03d}" for i in range(tag_size)]
    with open(os.path.join(out_dir, "stackoverflow.word_count"), "w", encoding="utf8") as fh:
        for i, w in enumerate(words):
            fh.write(f"{w} {10 * vocab_size - i}\n")
    with open(os.path.join(out_dir, "stackoverflow.tag_count"), "w", encoding="utf8") as fh:
        json.dump({t: 10 * tag_size - i for i, t in enumerate(tags)}, fh)
    ids = client_ids(ds.client_num, "")
    paths = []
    for split in ("train", "test"):
        tree = {}
        for c, x, _ in _per_client(ds, split == "test"):
            toks, titles, tg = [], [], []
            for row in x.numpy():
                ws = [int(i) - 1 for i in row if 1 <= int(i) <= vocab_size]
                toks.append(" ".join(words[i] for i in ws))
                titles.append(" ".join(words[i] for i in ws[:3]))
                tg.append("|".join(sorted({tags[i % tag_size] for i in ws[:2]})))
            tree[ids[c]] = {"tokens": toks or [""], "title": titles or [""], "tags": tg or [""]}
        paths.append(write_hdf5(os.path.join(out_dir, f"stackoverflow_{split}.h5"), {"examples": tree}))
    return paths


def export(name: str, ds: FederatedDataset, out_dir: str) -> List[str]:
    os.makedirs(out_dir, exist_ok=True)
    fn = {"femnist": export_fed_emnist, "fed_emnist": export_fed_emnist, "fed_cifar100": export_fed_cifar100,
          "fed_shakespeare": export_fed_shakespeare, "stackoverflow_nwp": export_stackoverflow,
          "stackoverflow_lr": export_stackoverflow}.get(name)
    if fn is None:
        raise ValueError(f"{name}: no TFF .h5 layout (known: femnist, fed_cifar100, fed_shakespeare, stackoverflow_nwp)")
    return fn(ds, out_dir)
