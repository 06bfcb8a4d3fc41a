"""The published edge-case backdoor sets, read from their original files (Wang et al., "Attack of the Tails", NeurIPS 2020).

Reference: data/edge_case_examples/data_loader.py:393-640 (``load_poisoned_dataset``) and datasets.py:142-185 (``get_ardis_dataset``,
``get_southwest_dataset``).  Two attacks:

* **Southwest Airlines -> "truck"** on CIFAR-10.  ``<root>/southwest_cifar10/`` holds pickled uint8 arrays ``[n, 32, 32, 3]``:
  ``southwest_images_new_{train,test}.pkl`` (``attack_case: edge-case``) or ``southwest_images_adv_p_percent_edge_case.pkl`` +
  ``southwest_images_p_percent_edge_case_test.pkl`` (``normal-case`` / ``almost-edge-case``).  The attacker's training set is
  M = 400 clean CIFAR-10 samples plus N = 100 airplane photos labelled 9; the targeted-task test set is every test photo
  labelled 9.
* **ARDIS sevens -> "1"** on MNIST / EMNIST digits.  ``<root>/ARDIS/ARDIS_{train,test}_2828.csv`` (one flattened 28x28 image per
  row) with one-hot ``ARDIS_{train,test}_labels.csv``; the sevens are taken, labelled 1 for training and evaluated against 1.

Everything returns plain arrays; ``data.load_poisoned_dataset_from_edge_case_examples`` turns them into batches with the clean
data set's transform.  Without the files that function keeps using its synthetic trigger-patch stand-in.
"""
from __future__ import annotations

import os
import pickle
from typing import Optional, Tuple

import numpy as np

SOUTHWEST_TARGET, ARDIS_SOURCE_DIGIT, ARDIS_TARGET = 9, 7, 1
NUM_POISONED, NUM_CLEAN = 100, 400          # N and M of the reference's mixed attacker set


def _unpickle(path: str) -> np.ndarray:
    with open(path, "rb") as fh:
        arr = pickle.load(fh)
    return np.asarray(arr)


def get_southwest_dataset(root: str, attack_case: str = "edge-case") -> Optional[Tuple[np.ndarray, np.ndarray]]:
    """``(train photos, test photos)`` as uint8 ``[n, 32, 32, 3]``, or None when the files are not under ``root``."""
    d = os.path.join(root, "southwest_cifar10")
    if attack_case == "edge-case":
        names = ("southwest_images_new_train.pkl", "southwest_images_new_test.pkl")
    elif attack_case in ("normal-case", "almost-edge-case"):
        names = ("southwest_images_adv_p_percent_edge_case.pkl", "southwest_images_p_percent_edge_case_test.pkl")
    else:
        raise NotImplementedError(f"attack_case {attack_case!r} (edge-case | normal-case | almost-edge-case)")
    paths = [os.path.join(d, n) for n in names]
    if not all(os.path.isfile(p) for p in paths):
        return None
    train, test = (_unpickle(p).astype(np.uint8) for p in paths)
    return train, test


def get_ardis_dataset(root: str, split: str = "train", digit: int = ARDIS_SOURCE_DIGIT) -> Optional[np.ndarray]:
    """The ARDIS images of ``digit`` as uint8 ``[n, 28, 28]``, or None when the CSV files are not under ``root``."""
    d = os.path.join(root, "ARDIS")
    img_path, lab_path = os.path.join(d, f"ARDIS_{split}_2828.csv"), os.path.join(d, f"ARDIS_{split}_labels.csv")
    if not (os.path.isfile(img_path) and os.path.isfile(lab_path)):
        return None
    images = np.loadtxt(img_path, dtype="float32", ndmin=2)
    labels = np.loadtxt(lab_path, dtype="float32", ndmin=2)
    picked = images[np.where(labels[:, digit] == 1)[0]]
    return picked.reshape(-1, 28, 28).clip(0, 255).astype(np.uint8)


def mix_attacker_set(clean_x: np.ndarray, clean_y: np.ndarray, poison_x: np.ndarray, target: int, rng: np.random.RandomState,
                     num_poisoned: int = NUM_POISONED, num_clean: int = NUM_CLEAN, sample_poison: bool = True):
    """M clean samples + N poisoned ones labelled ``target`` (data_loader.py:545-600).  With fewer than N photos all are used."""
    if sample_poison and len(poison_x) > num_poisoned:
        poison_x = poison_x[rng.choice(len(poison_x), num_poisoned, replace=False)]
    keep = rng.choice(len(clean_x), min(num_clean, len(clean_x)), replace=False)
    x = np.concatenate([clean_x[keep], poison_x.astype(clean_x.dtype)], axis=0)
    y = np.concatenate([np.asarray(clean_y)[keep], np.full(len(poison_x), target, dtype=np.asarray(clean_y).dtype)])
    return x, y


def load_edge_case_arrays(dataset: str, root: str, clean_x: np.ndarray, clean_y: np.ndarray, poison_type: str = "southwest",
                          attack_case: str = "edge-case", seed: int = 0):
    """``(x_attacker, y_attacker, x_targeted_test, y_targeted_test)`` or None when the real files are absent.  ``clean_x`` must
    be in the stored layout of the clean set (CIFAR-10: uint8 NHWC; MNIST: float ``[n, 784]`` in [0, 1])."""
    rng = np.random.RandomState(seed)
    name = dataset.lower()
    if name == "cifar10" and poison_type.startswith("southwest"):
        got = get_southwest_dataset(root, attack_case)
        if got is None:
            return None
        train, test = got
        x, y = mix_attacker_set(clean_x, clean_y, train, SOUTHWEST_TARGET, rng, sample_poison=(attack_case == "edge-case"))
        return x, y, test, np.full(len(test), SOUTHWEST_TARGET, dtype=np.int64)
    if name in ("mnist", "emnist", "femnist") and poison_type == "ardis":
        train, test = get_ardis_dataset(root, "train"), get_ardis_dataset(root, "test")
        if train is None or test is None:
            return None
        conv = (lambda a: a.reshape(len(a), -1).astype(np.float32) / 255.0) if clean_x.ndim == 2 else \
            (lambda a: (a.astype(np.float32) / 255.0).reshape((len(a),) + clean_x.shape[1:]))
        x, y = mix_attacker_set(clean_x, clean_y, conv(train), ARDIS_TARGET, rng)
        return x, y, conv(test), np.full(len(test), ARDIS_TARGET, dtype=np.int64)
    return None
