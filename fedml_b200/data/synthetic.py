"""Synthetic datasets "of the named shape".

There is no network on the build/bench boxes, so the reference's download paths
(MNIST/data_loader.py:17-29, torchvision ``download=True``) cannot run.  Each generator below
reproduces the *shape, dtype, client count, partition law and normalisation* of the
corresponding reference loader with class-conditional content that is actually
learnable (class prototype + noise), so accuracy-based oracles stay meaningful.
"""
from __future__ import annotations

import numpy as np
import torch

from ..core.non_iid_partition import (
    homo_partition,
    non_iid_partition_with_dirichlet_distribution,
    power_law_sizes,
)
from .federated_dataset import FederatedDataset

CIFAR10_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR10_STD = (0.2470, 0.2435, 0.2616)   # reference: data/cifar10/data_loader.py:80-103
CIFAR100_MEAN = (0.5071, 0.4865, 0.4409)
CIFAR100_STD = (0.2673, 0.2564, 0.2762)
IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


class ImageNormalize:
    """uint8 NHWC -> float NCHW normalised; the batch-time transform of image datasets."""

    def __init__(self, mean, std):
        self.mean = torch.tensor(mean, dtype=torch.float32).view(1, -1, 1, 1)
        self.std = torch.tensor(std, dtype=torch.float32).view(1, -1, 1, 1)

    def __call__(self, x_u8: torch.Tensor) -> torch.Tensor:
        x = x_u8.permute(0, 3, 1, 2).to(torch.float32).div_(255.0)
        return (x - self.mean.to(x.device)) / self.std.to(x.device)


def _offsets_from_sizes(sizes) -> np.ndarray:
    return np.concatenate([[0], np.cumsum(np.asarray(sizes, dtype=np.int64))])


def _pack_by_partition(x, y, part: dict):
    order = np.concatenate([np.asarray(part[c], dtype=np.int64) for c in range(len(part))])
    sizes = [len(part[c]) for c in range(len(part))]
    idx = torch.from_numpy(order)
    return x[idx].contiguous(), y[idx].contiguous(), _offsets_from_sizes(sizes)


# ----------------------------------------------------------------------------- MNIST-like
def make_mnist_like(client_num=1000, total_train=61664, dim=784, class_num=10, seed=0,
                    classes_per_client=2, test_fraction=0.1, name="mnist") -> FederatedDataset:
    """FedProx/LEAF MNIST split: ``client_num`` devices, each with samples of only two digits,
    power-law sizes; x float32 [n, 784] in [0, 1], y int64 (data/MNIST/README.md,
    MNIST/data_loader.py:32-98).  Train and test users are identical."""
    rng = np.random.default_rng(seed)
    protos = rng.random((class_num, dim)).astype(np.float32)
    protos = (protos > 0.75).astype(np.float32) * 0.9      # sparse "strokes"
    sizes = power_law_sizes(client_num, total_train, rng, min_size=10)
    xs, ys, txs, tys, tsizes = [], [], [], [], []
    for c in range(client_num):
        labels = rng.choice(class_num, size=min(classes_per_client, class_num), replace=False)
        n = int(sizes[c])
        nt = max(1, int(round(n * test_fraction)))
        yc = rng.choice(labels, size=n + nt)
        xc = protos[yc] + rng.normal(0, 0.35, size=(n + nt, dim)).astype(np.float32)
        np.clip(xc, 0.0, 1.0, out=xc)
        xs.append(xc[:n]); ys.append(yc[:n]); txs.append(xc[n:]); tys.append(yc[n:]); tsizes.append(nt)
    x = torch.from_numpy(np.concatenate(xs)); y = torch.from_numpy(np.concatenate(ys).astype(np.int64))
    xt = torch.from_numpy(np.concatenate(txs)); yt = torch.from_numpy(np.concatenate(tys).astype(np.int64))
    return FederatedDataset(x, y, _offsets_from_sizes(sizes), xt, yt, _offsets_from_sizes(tsizes),
                            class_num, None, name)


def make_femnist_like(client_num=3400, seed=0, **kw) -> FederatedDataset:
    """Federated EMNIST: 28x28 grey, 62 classes, 3400 natural clients (FederatedEMNIST/data_loader.py:133)."""
    ds = make_mnist_like(client_num=client_num, total_train=kw.get("total_train", 200 * client_num // 2),
                         class_num=62, seed=seed, classes_per_client=62, name="femnist")
    ds.x = ds.x.view(-1, 28, 28); ds.x_test = ds.x_test.view(-1, 28, 28)
    return ds


# ----------------------------------------------------------------------------- CIFAR-like
def make_image_dataset(n_train, n_test, class_num, hw=32, seed=0):
    """uint8 NHWC images: per-class low-frequency colour pattern + pixel noise."""
    rng = np.random.default_rng(seed)
    coarse = rng.integers(40, 216, size=(class_num, 4, 4, 3)).astype(np.float32)
    protos = np.repeat(np.repeat(coarse, hw // 4, axis=1), hw // 4, axis=2)

    def draw(n, r):
        y = r.integers(0, class_num, size=n)
        x = np.empty((n, hw, hw, 3), dtype=np.uint8)
        step = 8192
        for i in range(0, n, step):
            yy = y[i:i + step]
            img = protos[yy] + r.normal(0, 48.0, size=(len(yy), hw, hw, 3)).astype(np.float32)
            x[i:i + step] = np.clip(img, 0, 255).astype(np.uint8)
        return torch.from_numpy(x), torch.from_numpy(y.astype(np.int64))

    xtr, ytr = draw(n_train, rng)
    xte, yte = draw(n_test, rng)
    return xtr, ytr, xte, yte


def make_cifar_like(client_num=10, partition_method="homo", partition_alpha=0.5, class_num=10,
                    n_train=50000, n_test=10000, hw=32, seed=0, name="cifar10",
                    mean=CIFAR10_MEAN, std=CIFAR10_STD) -> FederatedDataset:
    """CIFAR-10-shaped: 50k/10k uint8 HWC; split ``homo`` (array_split of a permutation) or
    ``hetero`` (Dirichlet alpha) under ``np.random.seed(10)`` exactly as the reference
    (cifar10/data_loader.py:122-166); every client tests on the global test set (:207)."""
    xtr, ytr, xte, yte = make_image_dataset(n_train, n_test, class_num, hw, seed)
    state = np.random.get_state()
    np.random.seed(10)
    try:
        if partition_method == "homo" or client_num == 1:
            part = homo_partition(n_train, client_num)
        else:
            need = 10 if n_train // max(client_num, 1) >= 20 else 1
            part = non_iid_partition_with_dirichlet_distribution(
                ytr.numpy(), client_num, class_num, partition_alpha, min_required=need)
    finally:
        np.random.set_state(state)
    x, y, offs = _pack_by_partition(xtr, ytr, part)
    return FederatedDataset(x, y, offs, xte, yte, None, class_num, ImageNormalize(mean, std), name)


# ----------------------------------------------------------------------------- synthetic(alpha, beta)
def make_synthetic_ab(alpha=1.0, beta=1.0, client_num=30, dim=60, class_num=10, seed=0,
                      iid=False) -> FederatedDataset:
    """FedProx Synthetic(alpha, beta): per-client softmax-linear model with W_k, b_k ~ N(u_k, 1),
    u_k ~ N(0, alpha); x ~ N(v_k, Sigma) with v_k ~ N(B_k, 1), B_k ~ N(0, beta),
    Sigma_jj = j^-1.2; log-normal client sizes (synthetic_1_1/generate_synthetic.py:16-60)."""
    rng = np.random.default_rng(seed)
    sizes = (rng.lognormal(4, 2, client_num).astype(int) + 50).clip(max=5000)
    diag = np.array([(j + 1) ** -1.2 for j in range(dim)])
    W_g, b_g = rng.normal(0, 1, (dim, class_num)), rng.normal(0, 1, class_num)
    xs, ys, txs, tys, tsz, trsz = [], [], [], [], [], []
    for k in range(client_num):
        u, B = rng.normal(0, alpha), rng.normal(0, beta)
        W = W_g if iid else rng.normal(u, 1, (dim, class_num))
        b = b_g if iid else rng.normal(u, 1, class_num)
        v = np.zeros(dim) if iid else rng.normal(B, 1, dim)
        xk = rng.normal(v, np.sqrt(diag), size=(int(sizes[k]), dim))
        yk = np.argmax(xk @ W + b, axis=1)
        ntr = int(0.9 * len(yk))
        xs.append(xk[:ntr]); ys.append(yk[:ntr]); txs.append(xk[ntr:]); tys.append(yk[ntr:])
        trsz.append(ntr); tsz.append(len(yk) - ntr)
    f = lambda a, dt: torch.from_numpy(np.concatenate(a).astype(dt))
    return FederatedDataset(f(xs, np.float32), f(ys, np.int64), _offsets_from_sizes(trsz),
                            f(txs, np.float32), f(tys, np.int64), _offsets_from_sizes(tsz),
                            class_num, None, f"synthetic_{alpha}_{beta}")


# ----------------------------------------------------------------------------- sequences / text
def make_char_lm(client_num=715, seq_len=80, vocab=90, per_client=64, seed=0, name="shakespeare",
                 next_token_only=True) -> FederatedDataset:
    """Shakespeare-shaped: int64 [n, 80] char ids; label = next char (LEAF) or the shifted
    sequence (fed_shakespeare).  A per-dataset order-1 Markov chain makes it learnable."""
    rng = np.random.default_rng(seed)
    # small alphabets: a dense transition matrix; word-sized vocabularies (stackoverflow: 10004): eight successors per token,
    # so that a step costs O(n) instead of O(n * vocab) (the dense form took 42 s and 800 MB for 10004 words)
    K = vocab if vocab <= 512 else 8
    succ = None if K == vocab else rng.integers(1, vocab, size=(vocab, K))
    trans = rng.dirichlet(np.full(K, 0.05 if K == vocab else 0.5), size=vocab)
    cdf = np.cumsum(trans, axis=1)

    def gen(n):
        seq = np.empty((n, seq_len + 1), dtype=np.int64)
        seq[:, 0] = rng.integers(1, vocab, size=n)
        u = rng.random((n, seq_len))
        for t in range(seq_len):
            j = np.minimum((u[:, t, None] > cdf[seq[:, t]]).sum(axis=1), K - 1)
            seq[:, t + 1] = j if succ is None else succ[seq[:, t], j]
        return seq

    sizes = np.maximum(8, rng.poisson(per_client, client_num))
    tsz = np.maximum(2, sizes // 5)
    tr, te = gen(int(sizes.sum())), gen(int(tsz.sum()))
    lab = (lambda s: s[:, -1]) if next_token_only else (lambda s: s[:, 1:])
    return FederatedDataset(torch.from_numpy(tr[:, :seq_len].copy()), torch.from_numpy(lab(tr).copy()),
                            _offsets_from_sizes(sizes), torch.from_numpy(te[:, :seq_len].copy()),
                            torch.from_numpy(lab(te).copy()), _offsets_from_sizes(tsz), vocab, None, name,
                            task="nwp")


def make_stackoverflow_lr(client_num=1000, vocab=10000, tags=500, per_client=32, seed=0) -> FederatedDataset:
    """stackoverflow_lr-shaped: bag-of-words float [n, 10000] -> multi-hot tags float [n, 500]
    (stackoverflow_lr/data_loader.py:133); real set has 342,477 clients — capped for memory."""
    rng = np.random.default_rng(seed)
    W = (rng.random((vocab, tags)) < 0.002).astype(np.float32)
    sizes = np.maximum(4, rng.poisson(per_client, client_num)); tsz = np.maximum(2, sizes // 4)

    def gen(n):
        x = (rng.random((n, vocab)) < 0.003).astype(np.float32)
        y = ((x @ W) > 0).astype(np.float32)
        return torch.from_numpy(x), torch.from_numpy(y)

    x, y = gen(int(sizes.sum())); xt, yt = gen(int(tsz.sum()))
    return FederatedDataset(x, y, _offsets_from_sizes(sizes), xt, yt, _offsets_from_sizes(tsz), tags,
                            None, "stackoverflow_lr", task="tag_prediction")


def make_text_classification(client_num=256, seq_len=128, vocab=30522, class_num=4, per_client=32,
                             seed=0, name="fednlp_text") -> FederatedDataset:
    """FedNLP-shaped text classification (absent in the reference; north-star config): token ids
    int64 [n, S] with [CLS]=101 first, [SEP]=102 last real token, 0 padding; labels int64.
    Class-specific keyword tokens make the task learnable."""
    rng = np.random.default_rng(seed)
    kw = rng.integers(1000, vocab, size=(class_num, 64))
    sizes = np.maximum(4, rng.poisson(per_client, client_num)); tsz = np.maximum(2, sizes // 4)

    def gen(n):
        y = rng.integers(0, class_num, size=n)
        ids = rng.integers(1000, vocab, size=(n, seq_len))
        pick = rng.random((n, seq_len)) < 0.25
        kws = kw[y][np.arange(n)[:, None], rng.integers(0, 64, size=(n, seq_len))]
        ids = np.where(pick, kws, ids)
        lens = rng.integers(seq_len // 4, seq_len, size=n)
        pos = np.arange(seq_len)[None, :]
        ids = np.where(pos < lens[:, None], ids, 0)
        ids[:, 0] = 101
        ids[np.arange(n), lens - 1] = 102
        return torch.from_numpy(ids.astype(np.int64)), torch.from_numpy(y.astype(np.int64))

    x, y = gen(int(sizes.sum())); xt, yt = gen(int(tsz.sum()))
    return FederatedDataset(x, y, _offsets_from_sizes(sizes), xt, yt, _offsets_from_sizes(tsz), class_num,
                            None, name, task="text_classification")


# ----------------------------------------------------------------------------- vertical FL (tabular)
def make_vertical_tabular(n=4000, party_dims=(32, 24, 24), seed=0, name="lending_club"):
    """Feature-partitioned binary task (NUS-WIDE / lending-club / UCI shapes): returns
    ([x_party0, x_party1, ...], y) train and test, where party 0 (guest) also owns the labels."""
    rng = np.random.default_rng(seed)
    d = int(sum(party_dims))
    w = rng.normal(0, 1, d)
    x = rng.normal(0, 1, (n, d)).astype(np.float32)
    y = ((x @ w + rng.normal(0, 0.5, n)) > 0).astype(np.float32)
    cuts = np.cumsum(party_dims)[:-1]
    ntr = int(0.8 * n)
    split = lambda a: [torch.from_numpy(p.copy()) for p in np.split(a, cuts, axis=1)]
    return (split(x[:ntr]), torch.from_numpy(y[:ntr])), (split(x[ntr:]), torch.from_numpy(y[ntr:]))


# ----------------------------------------------------------------------------- segmentation
def make_segmentation(client_num=4, n_train=64, n_test=16, class_num=21, hw=64, seed=0, name="pascal_voc"):
    """Images float [n,3,H,W] with per-pixel int64 masks [n,H,W] (FedSeg shapes); mask = argmax of a
    coarse class field so that a conv net can learn it."""
    rng = np.random.default_rng(seed)

    def gen(n):
        field = rng.integers(0, class_num, size=(n, hw // 8, hw // 8))
        mask = np.repeat(np.repeat(field, 8, axis=1), 8, axis=2)
        palette = rng.random((class_num, 3)).astype(np.float32)
        img = palette[mask].transpose(0, 3, 1, 2) + rng.normal(0, 0.1, (n, 3, hw, hw)).astype(np.float32)
        return torch.from_numpy(img.astype(np.float32)), torch.from_numpy(mask.astype(np.int64))

    x, y = gen(n_train); xt, yt = gen(n_test)
    sizes = [len(p) for p in np.array_split(np.arange(n_train), client_num)]
    return FederatedDataset(x, y, _offsets_from_sizes(sizes), xt, yt, None, class_num, None, name,
                            task="segmentation")
