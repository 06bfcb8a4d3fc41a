"""Amplitude normalisation of HS-FedAvg (HarmoFL style; reference: single_process/hs_fedavg/hs_fft.py:8-84).

For every image channel: 2-D FFT -> amplitude / phase; the batch's mean amplitude updates a running amplitude
(momentum 0.1); the image is rebuilt from the *running* amplitude and its own phase.  The reference does this
in NumPy complex128 on the CPU per batch; on the GPU it is the hand-written shared-memory FFT kernel K14
(``csrc/fft.cu``, ``ops.amp_normalize``); CPU tensors and sizes the kernel does not cover go through ``torch.fft``.
The NumPy version below is kept as the oracle."""
import numpy as np
import torch


def amplitude_normalize(x: torch.Tensor, amp_running: torch.Tensor = None, momentum: float = 0.1, window=None):
    """x [B,C,H,W] float -> (x_normalised, new_running_amp [C,H,W], batch_mean_amp [C,H,W]).  ``window``: None swaps the whole
    amplitude spectrum for the running one (K14 kernel on the GPU); a float L applies the reference's low-frequency window rule
    (``amplitude_normalize_window``; L = 0 is what the reference's ``process`` uses)."""
    if window is not None:
        return amplitude_normalize_window(x, amp_running, momentum, float(window))
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


# ------------------------------------------------------------------------------------------------ low-frequency window rule
# The reference's ``process`` does not swap the whole amplitude spectrum: ``normalize(..., L=0)`` -> ``mutate`` replaces the centred
# window of half-width ``floor(min(H, W) * L)`` of the shifted spectrum, i.e. with L = 0 only the DC term takes the running value
# (hs_fft.py:15-33, 76).  ``amplitude_normalize(..., window=L)`` / the NumPy functions below implement that rule; ``window=None``
# (default of ``amplitude_normalize``) swaps every frequency, which is what the K14 kernel does.
def extract_amp(img_np: np.ndarray) -> np.ndarray:
    """|FFT2| over the last two axes."""
    return np.abs(np.fft.fft2(img_np, axes=(-2, -1)))


def _window(h: int, w: int, L: float):
    b = int(np.floor(min(h, w) * L))
    ch, cw = h // 2, w // 2
    return slice(ch - b, ch + b + 1), slice(cw - b, cw + b + 1)


def mutate(amp_src: np.ndarray, amp_trg: np.ndarray, L: float = 0.1) -> np.ndarray:
    """``amp_src`` with its low-frequency window (centre of the shifted spectrum) taken from ``amp_trg``."""
    a_src = np.fft.fftshift(amp_src, axes=(-2, -1)).copy()
    a_trg = np.fft.fftshift(amp_trg, axes=(-2, -1))
    hs, ws = _window(a_src.shape[-2], a_src.shape[-1], L)
    a_src[..., hs, ws] = a_trg[..., hs, ws]
    return np.fft.ifftshift(a_src, axes=(-2, -1))


def normalize(src_img: np.ndarray, amp_trg: np.ndarray, L: float = 0.1) -> np.ndarray:
    """The image rebuilt from its own phase and the window-mutated amplitude."""
    f = np.fft.fft2(src_img, axes=(-2, -1))
    return np.real(np.fft.ifft2(mutate(np.abs(f), amp_trg, L) * np.exp(1j * np.angle(f)), axes=(-2, -1)))


def process(x: np.ndarray, running_amp: np.ndarray, momentum: float = 0.1, fix_amp: bool = False, L: float = 0.0):
    """Batch step of the reference: update the running amplitude with the batch mean (first batch: take it), then normalise
    every image against it with window ``L``.  Returns ``(x_normalised, running_amp)``; ``x`` is not modified."""
    x = np.asarray(x)
    if not fix_amp:
        amp_avg = extract_amp(x).astype(np.float32).mean(axis=0)
        running_amp = amp_avg if np.sum(running_amp) == 0 else running_amp * (1 - momentum) + amp_avg * momentum
    out = normalize(x.astype(np.float64), np.asarray(running_amp)[None, :x.shape[1]], L).astype(x.dtype)
    return out, running_amp


def amplitude_normalize_window(x: torch.Tensor, amp_running: torch.Tensor = None, momentum: float = 0.1, window: float = 0.0):
    """Torch form of :func:`process` (any device): ``(x_normalised, new_running_amp, batch_mean_amp)``."""
    f = torch.fft.fft2(x.to(torch.float32), dim=(-2, -1))
    amp, pha = f.abs(), torch.angle(f)
    batch_amp = amp.mean(dim=0)
    amp_new = batch_amp if amp_running is None else (1 - momentum) * amp_running + momentum * batch_amp
    hs, ws = _window(x.shape[-2], x.shape[-1], window)
    shifted = torch.fft.fftshift(amp, dim=(-2, -1)).clone()
    shifted[..., hs, ws] = torch.fft.fftshift(amp_new, dim=(-2, -1))[..., hs, ws]
    mixed = torch.fft.ifftshift(shifted, dim=(-2, -1))
    return torch.fft.ifft2(torch.polar(mixed, pha), dim=(-2, -1)).real.to(x.dtype), amp_new, batch_amp
