"""Train-time augmentation of the reference's CIFAR-family loaders: RandomCrop(32, padding=4) + RandomHorizontalFlip +
Normalize + Cutout(16)  (data/cifar10/data_loader.py:58-103, same in cifar100 / cinic10).

Here it is a pure function of (seed, nonce, sample index): every sample's crop offset, flip bit and Cutout centre come from
one Philox4x32-10 block, so the CPU oracle below and the CUDA kernel (csrc/augment.cu, ``ops.augment_cifar``) produce the
same bytes, the client-batched GPU path needs no host RNG, and a run is reproducible regardless of how clients are
sharded over GPUs.  Images stay uint8 NHWC: the crop padding is black (0, as PIL pads before ToTensor) and the Cutout
square takes the per-channel dataset mean (what "0 after Normalize" is in pixel space); normalisation itself stays in the
input kernel of the model.
"""
from __future__ import annotations

import numpy as np
import torch

_M0, _M1, _W0, _W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
AUG_STREAM_TAG = 0xA06A06

CIFAR_MEAN_U8 = (125, 123, 114)       # round(255 * (0.4914, 0.4822, 0.4465))


def _philox(c0, c1, c2, c3, k0, k1):
    mask = np.uint64(0xFFFFFFFF)
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) for c in (c0, c1, c2, c3))
    k0, k1 = np.uint64(k0), np.uint64(k1)
    for _ in range(10):
        p0, p1 = np.uint64(_M0) * c0, np.uint64(_M1) * c2
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & mask, p1 >> np.uint64(32), p1 & mask
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & mask, lo1, (hi0 ^ c3 ^ k1) & mask, lo0
        k0, k1 = (k0 + np.uint64(_W0)) & mask, (k1 + np.uint64(_W1)) & mask
    return c0, c1, c2, c3


def augment_params(n: int, seed: int, nonce: int, H: int, W: int, pad: int = 4, first_index: int = 0):
    """Per-sample (dy, dx, flip, cy, cx): crop offsets in [0, 2*pad], flip bit, Cutout centre in [0,H) x [0,W)."""
    idx = np.arange(first_index, first_index + n, dtype=np.uint64)
    r0, r1, r2, r3 = _philox(idx & np.uint64(0xFFFFFFFF), idx >> np.uint64(32), np.full(n, AUG_STREAM_TAG, dtype=np.uint64),
                             np.full(n, int(nonce) & 0xFFFFFFFF, dtype=np.uint64), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    span = np.uint64(2 * pad + 1)
    dy = (r0 % span).astype(np.int64)
    dx = ((r0 >> np.uint64(16)) % span).astype(np.int64)
    flip = (r1 & np.uint64(1)).astype(np.int64)
    cy = (r2 % np.uint64(H)).astype(np.int64)
    cx = (r3 % np.uint64(W)).astype(np.int64)
    return dy, dx, flip, cy, cx


def augment_cifar_reference(x: torch.Tensor, seed: int, nonce: int, pad: int = 4, cutout: int = 16,
                            mean_u8=CIFAR_MEAN_U8, first_index: int = 0) -> torch.Tensor:
    """CPU oracle.  x uint8 [N, H, W, C] -> augmented uint8 [N, H, W, C]."""
    N, H, W, C = x.shape
    dy, dx, flip, cy, cx = augment_params(N, seed, nonce, H, W, pad, first_index)
    out = torch.empty_like(x)
    fill = torch.tensor(list(mean_u8)[:C] + [0] * max(0, C - len(mean_u8)), dtype=torch.uint8)
    for i in range(N):
        padded = torch.zeros(H + 2 * pad, W + 2 * pad, C, dtype=torch.uint8)
        padded[pad:pad + H, pad:pad + W] = x[i]
        img = padded[dy[i]:dy[i] + H, dx[i]:dx[i] + W]
        if flip[i]:
            img = img.flip(1)
        img = img.clone()
        if cutout > 0:
            y1, y2 = max(0, cy[i] - cutout // 2), min(H, cy[i] + cutout // 2)
            x1, x2 = max(0, cx[i] - cutout // 2), min(W, cx[i] + cutout // 2)
            img[y1:y2, x1:x2] = fill
        out[i] = img
    return out


def augment_cifar(x: torch.Tensor, seed: int, nonce, pad: int = 4, cutout: int = 16, mean_u8=CIFAR_MEAN_U8,
                  first_index: int = 0) -> torch.Tensor:
    """uint8 NHWC batch -> augmented batch; CUDA tensors use the native kernel (``nonce`` may then be a device int64
    tensor, so a captured CUDA graph sees a new value every replay), CPU tensors the oracle."""
    lead = x.shape[:-3]
    flat = x.reshape(-1, *x.shape[-3:])
    if flat.is_cuda:
        from .. import ops
        return ops.augment_cifar(flat, seed, nonce, pad, cutout, mean_u8, first_index).view(*lead, *x.shape[-3:])
    n = int(nonce.item()) if torch.is_tensor(nonce) else int(nonce)
    return augment_cifar_reference(flat, seed, n, pad, cutout, mean_u8, first_index).view(*lead, *x.shape[-3:])


class TrainAugment:
    """Callable hooked into ``FederatedDataset`` for the loaders of the single-process / worker-per-process simulators:
    every call (one batch of one client) draws a fresh nonce, like a new pass of the reference's random transforms."""

    def __init__(self, seed: int, pad: int = 4, cutout: int = 16, mean_u8=CIFAR_MEAN_U8):
        self.seed, self.pad, self.cutout, self.mean_u8, self.calls = int(seed), pad, cutout, mean_u8, 0

    def __call__(self, x_u8: torch.Tensor) -> torch.Tensor:
        self.calls += 1
        return augment_cifar(x_u8, self.seed, self.calls, self.pad, self.cutout, self.mean_u8)
