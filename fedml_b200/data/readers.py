"""Readers for the real dataset files of the reference benchmarks, producing the packed :class:`FederatedDataset`.

Formats (reference loaders in brackets):
  * LEAF JSON ``{users, user_data{user: {x, y}}}`` — MNIST (data/MNIST/data_loader.py:32-72), FEMNIST in its LEAF form,
    Shakespeare (data/shakespeare/data_loader.py:20-89; x = 80-character strings, y = next character, vocabulary
    data/shakespeare/language_utils.py:10-17);
  * TFF ``.h5`` exports - fed_emnist, fed_cifar100 (``examples/<client>/{pixels|image, label}``), fed_shakespeare
    (``examples/<client>/snippets``) and stackoverflow (``examples/<client>/{tokens, title, tags}`` plus the
    ``stackoverflow.word_count`` / ``stackoverflow.tag_count`` tables) - data/FederatedEMNIST/data_loader.py:133,
    fed_cifar100/data_loader.py:132, fed_shakespeare/data_loader.py:117, stackoverflow_{nwp,lr}/data_loader.py.  Opened with
    h5py when it is importable, otherwise with the bundled pure-Python reader (``minihdf5``);
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


def open_h5(path: str):
    """h5py.File when h5py is installed, else the bundled reader (same ``f[group][name][()]`` access)."""
    try:
        import h5py
        return h5py.File(path, "r")
    except ImportError:
        from . import minihdf5
        return minihdf5.File(path)


def _examples(f):
    """The per-client group: TFF names it ``examples`` (some re-exports: ``examples.md``)."""
    for k in ("examples", "examples.md"):
        if k in f:
            return f[k]
    raise KeyError(f"no 'examples' group in {getattr(f, 'filename', f)}")


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
    with open_h5(train_path) as ftr, open_h5(test_path) as fte:
        gtr, gte = _examples(ftr), _examples(fte)
        ids = sorted(gtr.keys())
        test_ids = set(gte.keys())
        xs = [np.asarray(gtr[c][x_key][()]) for c in ids]
        ys = [np.asarray(gtr[c][y_key][()]).astype(np.int64).reshape(-1) for c in ids]
        tids = [c for c in ids if c in test_ids]
        txs = [np.asarray(gte[c][x_key][()]) for c in tids]
        tys = [np.asarray(gte[c][y_key][()]).astype(np.int64).reshape(-1) for c in tids]
    x, xt = np.concatenate(xs), np.concatenate(txs)
    if to_uint8 and x.dtype != np.uint8:
        x, xt = x.astype(np.uint8), xt.astype(np.uint8)
    return FederatedDataset(torch.from_numpy(x), torch.from_numpy(np.concatenate(ys)), _offsets(ys), torch.from_numpy(xt),
                            torch.from_numpy(np.concatenate(tys)), _offsets(tys) if len(tids) == len(ids) else None,
                            class_num, transform, name)


# ---------------------------------------------------------------- TFF text exports
_PAD, _BOS, _EOS = "<pad>", "<bos>", "<eos>"


def shakespeare_char_ids() -> Dict[str, int]:
    """``<pad>`` = 0, the 86 characters, ``<bos>``, ``<eos>``; anything else maps to the one out-of-vocabulary id (= 89)
    (fed_shakespeare/utils.py:22-46)."""
    return {w: i for i, w in enumerate([_PAD] + CHAR_VOCAB + [_BOS, _EOS])}


def shakespeare_sequences(snippets: Sequence[str], seq_len: int = 80) -> np.ndarray:
    """Every snippet becomes ``<bos> chars <eos>`` padded to a multiple of ``seq_len + 1`` and cut into rows of that length
    (fed_shakespeare/utils.py:49-77); returns int64 ``[rows, seq_len + 1]``."""
    table = shakespeare_char_ids()
    oov, width = len(table), seq_len + 1
    rows = []
    for text in snippets:
        ids = [table[_BOS]] + [table.get(ch, oov) for ch in text] + [table[_EOS]]
        ids += [table[_PAD]] * (-len(ids) % width)
        rows.append(np.asarray(ids, dtype=np.int64).reshape(-1, width))
    return np.concatenate(rows) if rows else np.zeros((0, width), dtype=np.int64)


def _decode_all(ds) -> List[str]:
    return [b.decode("utf8") if isinstance(b, bytes) else str(b) for b in np.asarray(ds[()]).reshape(-1)]


def load_tff_shakespeare_h5(train_path: str, test_path: str, max_clients: int = 0) -> Optional[FederatedDataset]:
    """fed_shakespeare: x = tokens[:-1], y = tokens[1:] of every 81-id row (fed_shakespeare/utils.py:80-84)."""
    if not (os.path.isfile(train_path) and os.path.isfile(test_path)):
        return None
    with open_h5(train_path) as ftr, open_h5(test_path) as fte:
        gtr, gte = _examples(ftr), _examples(fte)
        ids = list(gtr.keys())
        if max_clients:
            ids = ids[:max_clients]
        test_ids = set(gte.keys())
        tr = [shakespeare_sequences(_decode_all(gtr[c]["snippets"])) for c in ids]
        te = [shakespeare_sequences(_decode_all(gte[c]["snippets"])) if c in test_ids else np.zeros((0, 81), np.int64) for c in ids]
    a, b = np.concatenate(tr), np.concatenate(te)
    return FederatedDataset(torch.from_numpy(a[:, :-1].copy()), torch.from_numpy(a[:, 1:].copy()), _offsets(tr),
                            torch.from_numpy(b[:, :-1].copy()), torch.from_numpy(b[:, 1:].copy()), _offsets(te),
                            len(shakespeare_char_ids()) + 1, None, "fed_shakespeare", task="nwp")


def stackoverflow_vocab(data_dir: str, vocab_size: int = 10000) -> List[str]:
    """First ``vocab_size`` words of ``stackoverflow.word_count`` (one ``word count`` pair per line, most frequent first)."""
    path = os.path.join(data_dir, "stackoverflow.word_count")
    if not os.path.isfile(path):
        raise FileNotFoundError(f"{path}: the stackoverflow loaders need the TFF word-count table next to the .h5 files")
    words = []
    with open(path, encoding="utf8") as fh:
        for line in fh:
            parts = line.split()
            if parts:
                words.append(parts[0])
            if len(words) == vocab_size:
                break
    return words


def stackoverflow_tags(data_dir: str, tag_size: int = 500) -> List[str]:
    """First ``tag_size`` keys of the JSON object in ``stackoverflow.tag_count``."""
    path = os.path.join(data_dir, "stackoverflow.tag_count")
    if not os.path.isfile(path):
        raise FileNotFoundError(f"{path}: stackoverflow_lr needs the TFF tag-count table next to the .h5 files")
    with open(path, encoding="utf8") as fh:
        return list(json.load(fh).keys())[:tag_size]


def stackoverflow_nwp_ids(sentences: Sequence[str], word_id: Dict[str, int], seq_len: int = 20) -> np.ndarray:
    """``<bos> w1..wk [<eos>] <pad>..`` with at most ``seq_len`` words, ``seq_len + 1`` ids per sentence; ids: ``<pad>`` 0, words
    1..V, ``<bos>`` V+1, ``<eos>`` V+2, out-of-vocabulary V+3 (stackoverflow_nwp/utils.py:32-82; ``word_id`` holds words only)."""
    v = len(word_id)
    pad, bos, eos, oov = 0, v + 1, v + 2, v + 3
    out = np.full((len(sentences), seq_len + 1), pad, dtype=np.int64)
    for r, s in enumerate(sentences):
        words = s.split(" ")[:seq_len]
        ids = [word_id[w] + 1 if w in word_id else oov for w in words]      # word k of the table has id k + 1
        if len(ids) < seq_len:
            ids.append(eos)
        ids = [bos] + ids
        out[r, :len(ids)] = ids
    return out


def stackoverflow_bag_of_words(sentences: Sequence[str], word_id: Dict[str, int]) -> np.ndarray:
    """Mean one-hot over the words of ``tokens + " " + title`` restricted to the vocabulary (stackoverflow_lr/utils.py:118-136)."""
    v = len(word_id)
    out = np.zeros((len(sentences), v), dtype=np.float32)
    for r, s in enumerate(sentences):
        words = s.split(" ")
        for w in words:
            k = word_id.get(w)
            if k is not None:
                out[r, k] += 1.0
        out[r] /= max(1, len(words))
    return out


def stackoverflow_tag_targets(tags: Sequence[str], tag_id: Dict[str, int]) -> np.ndarray:
    """Multi-hot over the ``|``-separated tags that are in the table (stackoverflow_lr/utils.py:139-156)."""
    out = np.zeros((len(tags), len(tag_id)), dtype=np.float32)
    for r, t in enumerate(tags):
        for name in t.split("|"):
            k = tag_id.get(name)
            if k is not None:
                out[r, k] += 1.0
    return out


def load_tff_stackoverflow_h5(data_dir: str, task: str, max_clients: int = 1000, seq_len: int = 20) -> Optional[FederatedDataset]:
    """``task``: "nwp" (next-word prediction, ids ``[n, 20]`` -> shifted ids) or "lr" (bag of words ``[n, 10000]`` -> tags
    ``[n, 500]``).  The full export has 342,477 training clients / 135 M sentences; ``max_clients`` bounds what is tokenised
    (0 = everything)."""
    train_path, test_path = os.path.join(data_dir, "stackoverflow_train.h5"), os.path.join(data_dir, "stackoverflow_test.h5")
    if not (os.path.isfile(train_path) and os.path.isfile(test_path)):
        return None
    word_id = {w: i for i, w in enumerate(stackoverflow_vocab(data_dir))}
    tag_id = {t: i for i, t in enumerate(stackoverflow_tags(data_dir))} if task == "lr" else None
    with open_h5(train_path) as ftr, open_h5(test_path) as fte:
        gtr, gte = _examples(ftr), _examples(fte)
        ids = list(gtr.keys())
        if max_clients:
            ids = ids[:max_clients]
        test_ids = set(gte.keys())

        def one(group):
            tokens = _decode_all(group["tokens"])
            if task == "nwp":
                seq = stackoverflow_nwp_ids(tokens, word_id, seq_len)
                return seq[:, :-1], seq[:, 1:]
            text = [" ".join(p) for p in zip(tokens, _decode_all(group["title"]))]
            return stackoverflow_bag_of_words(text, word_id), stackoverflow_tag_targets(_decode_all(group["tags"]), tag_id)

        tr = [one(gtr[c]) for c in ids]
        empty = (tr[0][0][:0], tr[0][1][:0])
        te = [one(gte[c]) if c in test_ids else empty for c in ids]
    cat = lambda parts, k: torch.from_numpy(np.ascontiguousarray(np.concatenate([p[k] for p in parts])))      # noqa: E731
    if task == "nwp":
        return FederatedDataset(cat(tr, 0), cat(tr, 1), _offsets([p[0] for p in tr]), cat(te, 0), cat(te, 1),
                                _offsets([p[0] for p in te]), len(word_id) + 4, None, "stackoverflow_nwp", task="nwp")
    return FederatedDataset(cat(tr, 0), cat(tr, 1), _offsets([p[0] for p in tr]), cat(te, 0), cat(te, 1),
                            _offsets([p[0] for p in te]), len(tag_id), None, "stackoverflow_lr", task="tag_prediction")


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
