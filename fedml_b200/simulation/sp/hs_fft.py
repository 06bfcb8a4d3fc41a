"""Amplitude normalisation of HS-FedAvg (HarmoFL style; reference: single_process/hs_fedavg/hs_fft.py:8-84).

For every image channel: 2-D FFT -> amplitude / phase; the batch's mean amplitude updates a running amplitude
(momentum 0.1); the image is rebuilt from the *running* amplitude and its own phase.  The reference does this
in NumPy complex128 on the CPU per batch; on the GPU it is the hand-written shared-memory FFT kernel K14
(``csrc/fft.cu``, ``ops.amp_normalize``); CPU tensors and sizes the kernel does not cover go through ``torch.fft``.
The NumPy version below is kept as the oracle."""
import numpy as np
import torch


def amplitude_normalize(x: torch.Tensor, amp_running: torch.Tensor = None, momentum: float = 0.1):
    """x [B,C,H,W] float -> (x_normalised, new_running_amp [C,H,W], batch_mean_amp [C,H,W])."""
    if x.is_cuda:
        from ... import ops
        if ops.amp_normalize_supported(x):
            return ops.amp_normalize(x, amp_running, momentum)
    f = torch.fft.fft2(x.to(torch.float32), dim=(-2, -1))
    amp, pha = f.abs(), torch.angle(f)
    batch_amp = amp.mean(dim=0)
    amp_new = batch_amp if amp_running is None else (1 - momentum) * amp_running + momentum * batch_amp
    rebuilt = torch.polar(amp_new.unsqueeze(0).expand_as(amp), pha)
    return torch.fft.ifft2(rebuilt, dim=(-2, -1)).real.to(x.dtype), amp_new, batch_amp


def amplitude_normalize_numpy(x: np.ndarray, amp_running: np.ndarray = None, momentum: float = 0.1):
    f = np.fft.fft2(x.astype(np.float64), axes=(-2, -1))
    amp, pha = np.abs(f), np.angle(f)
    batch_amp = amp.mean(axis=0)
    amp_new = batch_amp if amp_running is None else (1 - momentum) * amp_running + momentum * batch_amp
    out = np.fft.ifft2(amp_new[None] * np.exp(1j * pha), axes=(-2, -1)).real
    return out, amp_new, batch_amp
