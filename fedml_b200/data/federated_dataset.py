"""Packed, client-major dataset container.

The reference keeps one Python list of ``(x, y)`` batches or one ``DataLoader`` per client
(data/data_loader.py:312-322).  That layout forces a Python loop and a host->device copy per
batch.  Here all samples of all clients live in two contiguous host tensors (optionally
pinned) plus an offsets table, so that (a) the 8-tuple contract is a cheap *view* over it and
(b) the multi-GPU simulator can gather the sampled clients of a round with one indexed copy
per rank and ship them with a single H2D transfer.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch


class FederatedDataset:
    def __init__(self, x: torch.Tensor, y: torch.Tensor, client_offsets: np.ndarray,
                 x_test: torch.Tensor, y_test: torch.Tensor, test_offsets: Optional[np.ndarray],
                 class_num: int, transform=None, name: str = "", task: str = "classification"):
        """``x[client_offsets[c]:client_offsets[c+1]]`` are client c's training samples.
        ``test_offsets`` None means every client evaluates on the whole global test set
        (CIFAR behaviour, cifar10/data_loader.py:207)."""
        self.x, self.y = x, y
        self.client_offsets = np.asarray(client_offsets, dtype=np.int64)
        self.x_test, self.y_test = x_test, y_test
        self.test_offsets = None if test_offsets is None else np.asarray(test_offsets, dtype=np.int64)
        self.class_num = int(class_num)
        self.transform = transform      # callable(x_raw_batch) -> float tensor, applied at batch time
        self.train_augment = None       # callable(uint8 NHWC batch) -> uint8 NHWC batch, TRAIN batches only (data/augment.py)
        self.name, self.task = name, task

    # ------------------------------------------------------------------ basic facts
    @property
    def client_num(self) -> int:
        return len(self.client_offsets) - 1

    def client_size(self, cid: int) -> int:
        return int(self.client_offsets[cid + 1] - self.client_offsets[cid])

    def client_sizes(self) -> np.ndarray:
        return np.diff(self.client_offsets)

    def client_slice(self, cid: int) -> Tuple[torch.Tensor, torch.Tensor]:
        a, b = int(self.client_offsets[cid]), int(self.client_offsets[cid + 1])
        return self.x[a:b], self.y[a:b]

    def client_test_slice(self, cid: int) -> Tuple[torch.Tensor, torch.Tensor]:
        if self.test_offsets is None:
            return self.x_test, self.y_test
        a, b = int(self.test_offsets[cid]), int(self.test_offsets[cid + 1])
        return self.x_test[a:b], self.y_test[a:b]

    def pin(self) -> "FederatedDataset":
        if torch.cuda.is_available():
            self.x, self.y = self.x.pin_memory(), self.y.pin_memory()
        return self

    # ------------------------------------------------------------------ batching
    def _prep(self, xb: torch.Tensor) -> torch.Tensor:
        return self.transform(xb) if self.transform is not None else xb

    def batches(self, xs: torch.Tensor, ys: torch.Tensor, batch_size: int, drop_last: bool = False, train: bool = False,
                ) -> List[Tuple[torch.Tensor, torch.Tensor]]:
        aug = self.train_augment if (train and xs.dtype == torch.uint8) else None
        prep = (lambda xb: self._prep(aug(xb))) if aug is not None else self._prep
        n = xs.shape[0]
        if batch_size <= 0 or batch_size >= n:
            return [(prep(xs), ys)] if n else []
        out = []
        for i in range(0, n, batch_size):
            if drop_last and i + batch_size > n:
                break
            out.append((prep(xs[i:i + batch_size]), ys[i:i + batch_size]))
        return out

    def as_tuple(self, batch_size: int, drop_last: bool = False, test_batch_size: int = 0) -> list:
        """The reference's 8-item dataset list (data/data_loader.py:312-322)."""
        tbs = test_batch_size or batch_size
        train_local_num = {c: self.client_size(c) for c in range(self.client_num)}
        train_local = _LazyClientBatches(self, batch_size, drop_last, train=True)
        test_local = _LazyClientBatches(self, tbs, False, train=False)
        train_global = _LazyGlobalBatches(self, batch_size, train=True)
        test_global = _LazyGlobalBatches(self, tbs, train=False)
        return [int(self.x.shape[0]), int(self.x_test.shape[0]), train_global, test_global,
                train_local_num, train_local, test_local, self.class_num]


class _LazyGlobalBatches:
    """List-like view over the whole train/test split, materialised batch by batch."""

    def __init__(self, ds: FederatedDataset, batch_size: int, train: bool):
        self.ds, self.bs, self.train = ds, batch_size, train
        xs = ds.x if train else ds.x_test
        n = xs.shape[0]
        self._n = 1 if (batch_size <= 0 or batch_size >= n) else (n + batch_size - 1) // batch_size
        self.dataset = range(n)  # len(loader.dataset) is used by some callers

    def __len__(self):
        return self._n

    def __iter__(self):
        xs, ys = (self.ds.x, self.ds.y) if self.train else (self.ds.x_test, self.ds.y_test)
        return iter(self.ds.batches(xs, ys, self.bs)) if self._n <= 64 else self._gen(xs, ys)

    def _gen(self, xs, ys):
        for i in range(0, xs.shape[0], self.bs):
            yield self.ds._prep(xs[i:i + self.bs]), ys[i:i + self.bs]

    def __getitem__(self, i):
        xs, ys = (self.ds.x, self.ds.y) if self.train else (self.ds.x_test, self.ds.y_test)
        if self._n == 1:
            return self.ds._prep(xs), ys
        return self.ds._prep(xs[i * self.bs:(i + 1) * self.bs]), ys[i * self.bs:(i + 1) * self.bs]


class _LazyClientBatches(dict):
    """``{client_id: [ (x,y) batches ]}`` built on first access; behaves like the reference's
    ``train_data_local_dict`` without materialising 1000 clients up front."""

    def __init__(self, ds: FederatedDataset, batch_size: int, drop_last: bool, train: bool):
        super().__init__()
        self.ds, self.bs, self.drop_last, self.train = ds, batch_size, drop_last, train

    def __missing__(self, cid):
        if not (0 <= int(cid) < self.ds.client_num):
            raise KeyError(cid)
        xs, ys = self.ds.client_slice(cid) if self.train else self.ds.client_test_slice(cid)
        val = self.ds.batches(xs, ys, self.bs, self.drop_last, train=self.train)
        if not (self.train and self.ds.train_augment is not None):
            self[cid] = val          # augmented train batches are NOT cached: every visit of a client is a fresh random pass
        return val

    def __contains__(self, cid):
        return isinstance(cid, (int, np.integer)) and 0 <= int(cid) < self.ds.client_num

    def keys(self):
        return range(self.ds.client_num)

    def __len__(self):
        return self.ds.client_num

    def __iter__(self):
        return iter(range(self.ds.client_num))

    def items(self):
        return ((c, self[c]) for c in range(self.ds.client_num))

    def values(self):
        return (self[c] for c in range(self.ds.client_num))


def merge_to_single_client(dataset8: list) -> list:
    """Centralised switch: ``client_num_in_total == 1`` merges every client's batches into
    client 0 (reference: data/data_loader.py:276-297)."""
    (train_num, test_num, train_g, test_g, local_num, train_local, test_local, class_num) = dataset8
    keys = sorted(train_local.keys())
    local_num = {0: sum(local_num[c] for c in keys)}
    train_local = {0: [b for c in keys for b in train_local[c]]}
    test_local = {0: [b for c in keys for b in test_local[c]]}
    return [train_num, test_num, train_g, test_g, local_num, train_local, test_local, class_num]


def combine_batches(batches) -> list:
    """Full-batch switch (``batch_size <= 0``; reference: data/data_loader.py:33-39)."""
    batches = list(batches)
    if not batches:
        return []
    return [(torch.cat([b[0] for b in batches], 0), torch.cat([b[1] for b in batches], 0))]
