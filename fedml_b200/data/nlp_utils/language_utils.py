"""Character / word helpers of the LEAF Shakespeare and Sent140 loaders (reference: data/shakespeare/language_utils.py:10-120,
data/shakespeare/data_loader.py:20-60)."""
from __future__ import annotations

import json
import os
import re
from typing import Dict, List, Sequence

import numpy as np
import torch

from ..readers import ALL_LETTERS, CHAR_VOCAB, read_leaf_dir, word_to_indices  # noqa: F401  (re-exported)

VOCAB_SIZE = len(ALL_LETTERS) + 4          # + padding, out-of-vocabulary, begin, end
_TOKEN = re.compile(r"[\w']+|[.,!?;]")


def letter_to_index(letter: str) -> int:
    """Position in the 86-character table, -1 when absent."""
    return ALL_LETTERS.find(letter)


def letter_to_vec(letter: str) -> List[int]:
    """One-hot list of length ``VOCAB_SIZE`` (a letter that is not in the table sets the last slot, like index -1 does)."""
    vec = [0] * VOCAB_SIZE
    vec[letter_to_index(letter)] = 1
    return vec


def split_line(line: str) -> List[str]:
    """Words (with apostrophes) and the punctuation marks ``. , ! ? ;`` as separate tokens."""
    return _TOKEN.findall(line)


def line_to_indices(line: str, word2id: Dict[str, int], max_words: int = 25) -> List[int]:
    """First ``max_words`` tokens as ids; unknown words and the padding use ``len(word2id)``."""
    unk = len(word2id)
    ids = [word2id.get(w, unk) for w in split_line(line)[:max_words]]
    return ids + [unk] * (max_words - len(ids))


def bag_of_words(line: str, vocab: Dict[str, int]) -> List[int]:
    bag = [0] * len(vocab)
    for w in split_line(line):
        k = vocab.get(w)
        if k is not None:
            bag[k] += 1
    return bag


def process_x(raw_x_batch: Sequence[str]) -> List[List[int]]:
    return [word_to_indices(w) for w in raw_x_batch]


def process_y(raw_y_batch: Sequence[str]) -> List[int]:
    return [letter_to_index(c) for c in raw_y_batch]


def read_data(train_data_dir: str, test_data_dir: str):
    """LEAF directories -> ``(clients, groups, train_data, test_data)``."""
    clients, train = read_leaf_dir(train_data_dir)
    _, test = read_leaf_dir(test_data_dir)
    return sorted(set(clients)), [], train, test


def batch_data(data: dict, batch_size: int, seed: int = 100):
    """Shuffles a client's ``{"x": [...], "y": [...]}`` (x and y with the same permutation) and cuts it into tensor batches:
    text rows (Shakespeare) become character ids, numeric rows (MNIST, synthetic) float features with int64 labels."""
    x, y = list(data["x"]), list(data["y"])
    order = np.random.RandomState(seed).permutation(len(x))
    text = bool(x) and isinstance(x[0], str)
    out = []
    for i in range(0, len(x), batch_size):
        idx = order[i:i + batch_size]
        xb, yb = [x[j] for j in idx], [y[j] for j in idx]
        if text:
            out.append((torch.tensor(process_x(xb), dtype=torch.int64), torch.tensor(process_y(yb), dtype=torch.int64)))
        else:
            out.append((torch.tensor(np.asarray(xb, dtype=np.float32)), torch.tensor(np.asarray(yb, dtype=np.int64))))
    return out


def load_leaf_stats(name: str, data_dir: str = ".") -> dict:
    """The merged ``{users, num_samples, user_data}`` of ``<data_dir>/<name>/data/all_data`` or of a train directory."""
    for sub in (os.path.join(data_dir, name, "data", "all_data"), os.path.join(data_dir, name, "train"), data_dir):
        if os.path.isdir(sub) and any(f.endswith(".json") for f in os.listdir(sub)):
            users, num = [], []
            for f in sorted(os.listdir(sub)):
                if f.endswith(".json"):
                    with open(os.path.join(sub, f)) as fh:
                        blob = json.load(fh)
                    users += blob["users"]
                    num += blob.get("num_samples", [len(blob["user_data"][u]["y"]) for u in blob["users"]])
            return {"users": users, "num_samples": num}
    raise FileNotFoundError(f"no LEAF json files for {name!r} under {data_dir}")


def print_dataset_stats(name: str, data_dir: str = ".") -> dict:
    """Users, samples, mean / std / skewness of the samples per user (the numbers LEAF's stats.py prints)."""
    d = load_leaf_stats(name, data_dir)
    n = np.asarray(d["num_samples"], dtype=np.float64)
    mu, sd = float(n.mean()), float(n.std())
    stats = {"users": len(d["users"]), "samples": int(n.sum()), "mean": mu, "std": sd, "std_over_mean": sd / mu if mu else 0.0,
             "skewness": float(((n - mu) ** 3).mean() / sd ** 3) if sd else 0.0}
    print(f"DATASET: {name}\n{stats['users']} users\n{stats['samples']} samples (total)\n{mu:.2f} samples per user (mean)\n"
          f"num_samples (std): {sd:.2f}\nnum_samples (std/mean): {stats['std_over_mean']:.2f}\nnum_samples (skewness): {stats['skewness']:.2f}")
    return stats
