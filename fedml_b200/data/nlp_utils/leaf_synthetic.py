"""LEAF / FedProx Synthetic(alpha, beta) as files: the generator script of the reference's ``data/synthetic_*`` folders
(generate_synthetic.py:16-97) - per-user lists and the ``train/mytrain.json`` + ``test/mytest.json`` pair that
``load_partition_data_federated_synthetic_1_1`` and LEAF tools read.  The in-memory form is ``synthetic.make_synthetic_ab``."""
from __future__ import annotations

import json
import os
from typing import List, Tuple

import numpy as np

from ..synthetic import make_synthetic_ab

NUM_USER = 30


def softmax(x):
    x = np.asarray(x, dtype=np.float64)
    e = np.exp(x - x.max())
    return e / e.sum()


def generate_synthetic(alpha, beta, iid, num_user: int = NUM_USER, seed: int = 0) -> Tuple[List[list], List[list]]:
    """``(X_split, y_split)``: per user a list of 60-dimensional samples and a list of labels 0..9 (train and test together)."""
    ds = make_synthetic_ab(alpha, beta, num_user, seed=seed, iid=bool(iid))
    xs, ys = [], []
    for c in range(ds.client_num):
        x, y = ds.client_slice(c)
        xt, yt = ds.client_test_slice(c)
        xs.append(np.concatenate([x.numpy(), xt.numpy()]).tolist())
        ys.append(np.concatenate([y.numpy(), yt.numpy()]).astype(float).tolist())
    return xs, ys


def write_leaf_files(out_dir: str = ".", alpha: float = 1.0, beta: float = 1.0, iid: int = 0, seed: int = 0) -> Tuple[str, str]:
    """Writes ``<out_dir>/train/mytrain.json`` and ``<out_dir>/test/mytest.json`` (users ``f_00000`` ..., 90 / 10 split after a
    per-user shuffle)."""
    xs, ys = generate_synthetic(alpha, beta, iid, seed=seed)
    rng = np.random.RandomState(seed)
    train = {"users": [], "user_data": {}, "num_samples": []}
    test = {"users": [], "user_data": {}, "num_samples": []}
    for i, (x, y) in enumerate(zip(xs, ys)):
        order = rng.permutation(len(y))
        x, y = [x[j] for j in order], [y[j] for j in order]
        cut = int(0.9 * len(y))
        name = f"f_{i:05d}"
        for blob, xa, ya in ((train, x[:cut], y[:cut]), (test, x[cut:], y[cut:])):
            blob["users"].append(name)
            blob["user_data"][name] = {"x": xa, "y": ya}
            blob["num_samples"].append(len(ya))
    paths = []
    for sub, fname, blob in (("train", "mytrain.json", train), ("test", "mytest.json", test)):
        os.makedirs(os.path.join(out_dir, sub), exist_ok=True)
        paths.append(os.path.join(out_dir, sub, fname))
        with open(paths[-1], "w") as fh:
            json.dump(blob, fh)
    return tuple(paths)
