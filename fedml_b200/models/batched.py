"""Client-batched ("stacked") model execution: C co-resident clients advance one SGD step together.

The reference trains the sampled clients one after another, each step being ~170 cuDNN/cuBLAS
launches on tiny tensors (launch-bound).  Here the C clients that share a GPU are
executed as ONE network whose every layer has a leading client dimension: activations are
``[C, B, H, W, ch]`` (NHWC, bf16 on the GPU), parameters are ``[C, *shape]`` strided views into the
stacked fp32 arena (``ArenaLayout.stacked_views``), convolutions become grouped GEMMs (1x1: directly,
3x3: im2col), BatchNorm statistics are per (client, channel).  Forward *and* backward are written
out explicitly (no autograd): every tensor is a static buffer, so a whole step is CUDA-graph
capturable and every primitive can be a hand-written kernel.

Primitive back-ends: :class:`TorchBackend` (plain PyTorch, any device/dtype — the CPU oracle path
used by the tests) and :class:`NativeBackend` (sm_100a kernels from ``csrc/``: fused BN(+ReLU+residual)
forward/backward, im2col/col2im, tcgen05 grouped GEMM).  Module/parameter names are taken from the
reference-compatible ``nn.Module`` so arena segments line up with ``state_dict`` keys
(model/cv/resnet.py:87-264, model/linear/lr.py:4-11).
"""
from __future__ import annotations

import math
import os
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from .cv import BasicBlock, Bottleneck, LogisticRegression, ResNet


# ============================================================================ back-ends
class TorchBackend:
    """Reference implementation of the batched primitives in plain PyTorch."""
    name = "torch"

    def __init__(self, dtype=torch.float32):
        self.dtype = dtype

    # ---- GEMMs (client-batched) -------------------------------------------------
    def gemm_nt(self, a, b, out_dtype=None, relu=False, stats=None):
        """a [C,M,K] x b [C,N,K]^T -> [C,M,N].  ``stats`` (fp32 [C,2,N], zeroed): back-ends that can, add the column sums
        and sums of squares of the result into it and set ``self.stats_done``; this one leaves it alone."""
        self.stats_done = False
        y = torch.bmm(a, b.transpose(1, 2))
        if relu:
            y = torch.relu_(y)
        return y if out_dtype is None else y.to(out_dtype)

    want_transposed_weights = False

    def gemm_nn(self, a, b, out_dtype=None, b_t=None):
        """a [C,M,N] x b [C,N,K] -> [C,M,K]   (``b_t`` = b^T [C,K,N], used by back-ends that want K-major operands)"""
        y = torch.bmm(a, b)
        return y if out_dtype is None else y.to(out_dtype)

    grads_prezeroed = False

    def conv3x3_ok(self, x, cout, stride):
        """True when the back-end has an implicit-GEMM 3x3 convolution for this shape (else im2col + GEMM)."""
        return False

    def conv3x3_wgrad_ok(self, x, dy, stride):
        return False

    # ---- side branch for work that is off the critical path of the backward chain (weight gradients) ----------
    def side(self, *keep):
        """Context manager: run the enclosed ops concurrently with the caller's stream when the back-end has a side
        stream (``keep``: tensors produced on the main stream that the branch reads).  No-op here."""
        import contextlib
        return contextlib.nullcontext()

    def join(self):
        """Make the caller's stream wait for everything issued under :meth:`side`."""

    def gemm_tn(self, a, b, out_dtype=torch.float32, out=None, out_is_zeroed=False):
        """a [C,M,N]^T x b [C,M,K] -> [C,N,K]  (weight gradients; fp32 result, optionally written into ``out``)"""
        r = torch.bmm(a.transpose(1, 2), b).to(out_dtype)
        if out is not None:
            out.copy_(r)
            return out
        return r

    # ---- im2col / col2im for 3x3, pad 1 (NHWC) ------------------------------------
    def im2col3x3(self, x, stride):
        C, B, H, W, ch = x.shape
        Ho, Wo = (H - 1) // stride + 1, (W - 1) // stride + 1
        xp = F.pad(x, (0, 0, 1, 1, 1, 1))
        win = xp.unfold(2, 3, stride).unfold(3, 3, stride)          # [C,B,Ho,Wo,ch,3,3]
        return win.permute(0, 1, 2, 3, 5, 6, 4).reshape(C, B * Ho * Wo, 9 * ch)

    def col2im3x3(self, dcols, B, H, W, ch, stride):
        C = dcols.shape[0]
        Ho, Wo = (H - 1) // stride + 1, (W - 1) // stride + 1
        d = dcols.view(C, B, Ho, Wo, 3, 3, ch)
        dxp = torch.zeros(C, B, H + 2, W + 2, ch, dtype=dcols.dtype, device=dcols.device)
        for kh in range(3):
            for kw in range(3):
                dxp[:, :, kh:kh + stride * Ho:stride, kw:kw + stride * Wo:stride, :] += d[:, :, :, :, kh, kw, :]
        return dxp[:, :, 1:H + 1, 1:W + 1, :].contiguous()

    # ---- BatchNorm (+residual, +ReLU) ---------------------------------------------
    def bn_fwd(self, x, res, gamma, beta, rm, rv, nbt, active, momentum, eps, relu, training, stats=None):
        """x [C,B,H,W,ch]; gamma/beta/rm/rv [C,ch] fp32 (views, rm/rv updated in place); nbt [C] fp32."""
        C = x.shape[0]
        ch = x.shape[-1]
        xf = x.float().reshape(C, -1, ch)
        M = xf.shape[1]
        if training:
            mean = xf.mean(dim=1)
            var = xf.var(dim=1, unbiased=False)
            a = active.view(C, 1) * momentum
            rm.add_(a * (mean - rm))
            rv.add_(a * (var * (M / max(M - 1, 1)) - rv))
            nbt.add_(active)
        else:
            mean, var = rm, rv
        rstd = torch.rsqrt(var + eps)
        y = (xf - mean[:, None, :]) * (rstd * gamma)[:, None, :] + beta[:, None, :]
        y = y.view(x.shape)
        if res is not None:
            y = y + res.float()
        if relu:
            y = torch.relu(y)
        return y.to(x.dtype), mean, rstd

    def bn_bwd(self, dy, x, y, gamma, mean, rstd, relu, want_dres, beta=None, had_res=True, dgamma_out=None,
               dbeta_out=None):
        """``dy`` may be a pair of tensors (gradient arriving over the main and the residual path): they are summed."""
        C = x.shape[0]
        ch = x.shape[-1]
        dyf = (dy[0].float() + dy[1].float()) if isinstance(dy, tuple) else dy.float()
        if relu:
            dyf = dyf * (y > 0)
        dres = dyf.to(x.dtype) if want_dres else None
        d2 = dyf.reshape(C, -1, ch)
        M = d2.shape[1]
        xhat = (x.float().reshape(C, -1, ch) - mean[:, None, :]) * rstd[:, None, :]
        dbeta = d2.sum(dim=1)
        dgamma = (d2 * xhat).sum(dim=1)
        dx = (gamma * rstd)[:, None, :] * (d2 - dbeta[:, None, :] / M - xhat * dgamma[:, None, :] / M)
        if dgamma_out is not None:
            dgamma_out.copy_(dgamma); dbeta_out.copy_(dbeta)
        return dx.view(x.shape).to(x.dtype), dres, dgamma, dbeta

    # ---- GroupNorm (+residual, +ReLU): statistics per (client, sample, group of cpg channels) ------------
    def gn_fwd(self, x, res, gamma, beta, cpg, eps, relu):
        C, B, ch = x.shape[0], x.shape[1], x.shape[-1]
        G = ch // cpg
        xf = x.float().reshape(C, B, -1, G, cpg)
        mean = xf.mean(dim=(2, 4))                                     # [C,B,G]
        var = xf.var(dim=(2, 4), unbiased=False)
        rstd = torch.rsqrt(var + eps)
        xhat = (xf - mean[:, :, None, :, None]) * rstd[:, :, None, :, None]
        y = xhat.reshape(C, B, -1, ch) * gamma[:, None, None, :] + beta[:, None, None, :]
        y = y.view(x.shape)
        if res is not None:
            y = y + res.float()
        if relu:
            y = torch.relu(y)
        rep = lambda t: t.repeat_interleave(cpg, dim=2).reshape(C * B, ch)   # noqa: E731
        return y.to(x.dtype), rep(mean), rep(rstd)

    def gn_bwd(self, dy, x, y, gamma, mean, rstd, cpg, relu, want_dres, beta=None, dgamma_out=None, dbeta_out=None):
        C, B, ch = x.shape[0], x.shape[1], x.shape[-1]
        dyf = (dy[0].float() + dy[1].float()) if isinstance(dy, tuple) else dy.float()
        if relu:
            dyf = dyf * (y > 0)
        dres = dyf.to(x.dtype) if want_dres else None
        d = dyf.reshape(C, B, -1, ch)
        mu, rs = mean.view(C, B, 1, ch), rstd.view(C, B, 1, ch)
        xhat = (x.float().reshape(C, B, -1, ch) - mu) * rs
        dbeta = d.sum(dim=(1, 2))
        dgamma = (d * xhat).sum(dim=(1, 2))
        dxh = d * gamma[:, None, None, :]
        G = ch // cpg
        gm = lambda t: t.reshape(C, B, -1, G, cpg).mean(dim=(2, 4), keepdim=True).expand(C, B, t.shape[2], G, cpg).reshape(t.shape)  # noqa: E731
        dx = rs * (dxh - gm(dxh) - xhat * gm(dxh * xhat))
        if dgamma_out is not None:
            dgamma_out.copy_(dgamma); dbeta_out.copy_(dbeta)
        return dx.view(x.shape).to(x.dtype), dres, dgamma, dbeta

    # ---- loss ---------------------------------------------------------------------
    def xent(self, logits, labels, pre_sigmoid=False, row_scale=None):
        """logits [C,B,ncls] fp32, labels [C,B] (-1 = padding row) -> (loss per client [C], dlogits [C,B,ncls]).
        ``row_scale`` [C] = 1 / (valid rows of that client); None means 1/B."""
        C, B, ncls = logits.shape
        lg = logits.detach().clone().requires_grad_(True)
        z = torch.sigmoid(lg) if pre_sigmoid else lg
        rows = F.cross_entropy(z.view(C * B, ncls), labels.reshape(-1), reduction="none", ignore_index=-1).view(C, B)
        scale = torch.full((C,), 1.0 / B, dtype=torch.float32, device=logits.device) if row_scale is None else row_scale
        loss = rows.sum(dim=1) * scale
        loss.sum().backward()
        return loss.detach(), lg.grad

    def head_fwd_bwd(self, h, fc_w, fc_b, labels, row_scale, g_w, g_b):
        """Classifier head, forward and backward: average pool -> fc -> cross-entropy; writes the fc gradients into
        ``g_w`` / ``g_b`` and returns (loss per client [C], d loss / d h).  Reference composition of the primitives."""
        C, B, H, W_, ch = h.shape
        feat = h.to(fc_w.dtype).mean(dim=(2, 3))
        logits = torch.bmm(feat, fc_w.transpose(1, 2)) + fc_b[:, None, :]
        loss, dlog = TorchBackend.xent(self, logits, labels, row_scale=row_scale)
        g_w.copy_(torch.bmm(dlog.transpose(1, 2), feat))
        g_b.copy_(dlog.sum(dim=1))
        dfeat = torch.bmm(dlog, fc_w) / (H * W_)
        return loss, dfeat.to(h.dtype)[:, :, None, None, :].expand(C, B, H, W_, ch).contiguous()

    def normalize_images(self, x_u8, mean, std):
        """uint8 [C,B,H,W,3] -> normalised activations in the compute dtype."""
        return ((x_u8.float() / 255.0 - mean) / std).to(self.dtype)

    # ---- transformer primitives (models/batched_nlp.py) -----------------------------------
    def mm(self, mode, a, b, out=None, out_dtype=None, bias=None, relu=False):
        """Grouped product over any leading dims: nt = a b^T, nn = a b, tn = a^T b (fp32 accumulate)."""
        af, bf = a.float(), b.float()
        if mode == "nt":
            y = af @ bf.transpose(-1, -2)
        elif mode == "nn":
            y = af @ bf
        else:
            y = af.transpose(-1, -2) @ bf
        if bias is not None:
            y = y + bias[..., None, :]
        if relu:
            y = torch.relu(y)
        if out is not None:
            out.copy_(y)
            return out
        return y.to(out_dtype or a.dtype)

    def embed_fwd(self, ids, word, pos, S):
        C, R = ids.shape
        D = word.shape[2]
        w = torch.gather(word, 1, ids[..., None].expand(C, R, D))
        return (w + pos[:, :S].repeat(1, R // S, 1)).to(self.dtype)

    def embed_bwd(self, ids, dz, gword, gpos, S, word_scale=None):
        """gword += word_scale[c] * scatter(dz) (gradient rows: zero on entry; or the parameter rows with
        word_scale = -lr, the fused sparse SGD update); gpos is overwritten."""
        C, R = ids.shape
        D = gword.shape[2]
        src = dz.float() if word_scale is None else dz.float() * word_scale[:, None, None]
        gword.scatter_add_(1, ids[..., None].expand(C, R, D), src)
        gpos.zero_()
        gpos[:, :S] = dz.float().view(C, R // S, S, D).sum(dim=1)

    def ln_fwd(self, a, b, gamma, beta, eps):
        z = a if b is None else (a.float() + b.float()).to(a.dtype)
        zf = z.float()
        mean = zf.mean(dim=-1)
        rstd = torch.rsqrt(zf.var(dim=-1, unbiased=False) + eps)
        y = (zf - mean[..., None]) * rstd[..., None] * gamma[:, None, :] + beta[:, None, :]
        return y.to(a.dtype), z, mean, rstd

    def ln_bwd(self, dy, dy2, z, mean, rstd, gamma, dgamma_out, dbeta_out):
        d = dy.float() if dy2 is None else dy.float() + dy2.float()
        xh = (z.float() - mean[..., None]) * rstd[..., None]
        dgamma_out.copy_((d * xh).sum(dim=1))
        dbeta_out.copy_(d.sum(dim=1))
        g = d * gamma[:, None, :]
        dz = rstd[..., None] * (g - g.mean(dim=-1, keepdim=True) - xh * (g * xh).mean(dim=-1, keepdim=True))
        return dz.to(z.dtype)

    def _drop_mask(self, shape, p, seed, site, step, device):
        g = torch.Generator(device="cpu").manual_seed((int(seed) * 1000003 + int(site) * 7919 + int(step)) % (2 ** 63))
        return (torch.rand(shape, generator=g) >= p).to(device)

    def dropout(self, x, p, seed, site, step_dev=None, x2=None):
        if x2 is not None:
            x = (x.float() + x2.float()).to(x.dtype)
        if p <= 0:
            return x
        step = int(step_dev.item()) if step_dev is not None else 0
        return (x.float() * self._drop_mask(x.shape, p, seed, site, step, x.device) / (1.0 - p)).to(x.dtype)

    def softmax_fwd(self, scores, key_mask, H, scale, p=0.0, seed=0, site=0, step_dev=None):
        """scores fp32 [CB,H,S,S]; key_mask [CB,S] (1 = attend) -> (P, dropout(P)) in the compute dtype."""
        s = scores.float() * scale
        if key_mask is not None:
            s = s.masked_fill(key_mask[:, None, None, :] == 0, float("-inf"))
        P = torch.softmax(s, dim=-1).to(self.dtype)
        return P, (self.dropout(P, p, seed, site, step_dev) if p > 0 else P)

    def softmax_bwd(self, dP, P, scale, p=0.0, seed=0, site=0, step_dev=None):
        g = dP.float()
        if p > 0:
            step = int(step_dev.item()) if step_dev is not None else 0
            g = g * self._drop_mask(g.shape, p, seed, site, step, g.device) / (1.0 - p)
        Pf = P.float()
        return (scale * Pf * (g - (g * Pf).sum(dim=-1, keepdim=True))).to(self.dtype)

    def act_fwd(self, x, kind):
        return (F.gelu(x.float()) if kind == "gelu" else torch.relu(x.float())).to(x.dtype)

    def act_bwd(self, dy, x, kind):
        xf = x.float()
        if kind == "gelu":
            d = 0.5 * (1 + torch.erf(xf * 0.7071067811865476)) + xf * torch.exp(-0.5 * xf * xf) * 0.3989422804014327
        else:
            d = (xf > 0).float()
        return (dy.float() * d).to(x.dtype)

    def colsum(self, x, out):
        out.copy_(x.float().sum(dim=1))
        return out

    def maxpool2x2_fwd(self, x, off=(0, 0), out_hw=None):
        """x [C,B,H,W,ch] -> [C,B,Ho,Wo,ch], windows starting at ``off`` (the centre crop of a same-padded conv output)."""
        C, B, H, W, ch = x.shape
        Ho, Wo = out_hw or ((H - off[0]) // 2, (W - off[1]) // 2)
        xw = x[:, :, off[0]:off[0] + 2 * Ho, off[1]:off[1] + 2 * Wo, :]
        return xw.reshape(C, B, Ho, 2, Wo, 2, ch).amax(dim=(3, 5))

    def maxpool2x2_bwd(self, dy, x, off=(0, 0), relu_mask=False):
        """dy to the first maximal element of every window (scan order), zero elsewhere and outside the windows."""
        C, B, H, W, ch = x.shape
        Ho, Wo = dy.shape[2], dy.shape[3]
        xw = x[:, :, off[0]:off[0] + 2 * Ho, off[1]:off[1] + 2 * Wo, :].reshape(C, B, Ho, 2, Wo, 2, ch)
        win = xw.permute(0, 1, 2, 4, 6, 3, 5).reshape(C, B, Ho, Wo, ch, 4).float()
        idx = win.argmax(dim=-1, keepdim=True)                          # first occurrence
        d = dy.float()
        if relu_mask:
            d = d * (win.gather(-1, idx).squeeze(-1) > 0)
        dwin = torch.zeros_like(win).scatter_(-1, idx, d.unsqueeze(-1))
        dxw = dwin.reshape(C, B, Ho, Wo, ch, 2, 2).permute(0, 1, 2, 5, 3, 6, 4).reshape(C, B, 2 * Ho, 2 * Wo, ch)
        dx = torch.zeros(C, B, H, W, ch, dtype=x.dtype, device=x.device)
        dx[:, :, off[0]:off[0] + 2 * Ho, off[1]:off[1] + 2 * Wo, :] = dxw.to(x.dtype)
        return dx


class NativeBackend(TorchBackend):
    """sm_100a kernels for the bandwidth-bound primitives; GEMMs go to the tcgen05 grouped GEMM when
    the shape is supported by the kernel and to cuBLAS ``bmm`` otherwise (recorded in ``fallbacks``)."""
    name = "native"

    def __init__(self, dtype=torch.bfloat16, use_tc_gemm=True):
        super().__init__(dtype)
        from .. import ops
        self.ops = ops
        self.use_tc_gemm = use_tc_gemm and hasattr(ops, "grouped_gemm")
        self.fallbacks: Dict[str, int] = {}

    def _fb(self, what):
        self.fallbacks[what] = self.fallbacks.get(what, 0) + 1

    use_implicit_conv = os.environ.get("FB_IMPLICIT_CONV", "1") != "0"

    def conv3x3_ok(self, x, cout, stride):
        return self.use_tc_gemm and self.use_implicit_conv and self.ops.conv3x3_supported(x, cout, stride)

    def conv3x3(self, x, wm, stride, stats=None):
        self.stats_done = stats is not None
        return self.ops.conv3x3(x, wm, stride, stats=stats)

    # implicit weight gradients pay off when a pixel's channel vector fills a 128-byte line (cin >= 64); for narrower
    # layers the 32/64-byte TMA gathers cost more than the im2col buffer they avoid (measured, profiles/README.md)
    implicit_wgrad_min_cin = int(os.environ.get("FB_IMPLICIT_WGRAD_MIN_CIN", "64"))

    def conv3x3_wgrad_ok(self, x, dy, stride):
        return self.use_tc_gemm and self.use_implicit_conv and x.shape[-1] >= self.implicit_wgrad_min_cin \
            and self.ops.conv3x3_wgrad_supported(x, dy, stride)

    def conv3x3_wgrad(self, x, dy, stride, out=None, out_is_zeroed=False):
        return self.ops.conv3x3_wgrad(x, dy, stride, out=out, out_is_zeroed=out_is_zeroed)

    def conv3x3_dgrad(self, dy, wt):
        return self.ops.conv3x3_dgrad(dy.contiguous(), wt)

    # weight-gradient branch: a second stream (a parallel branch of the captured CUDA graph).  At the small client
    # counts a GPU sees when a round is sharded over 8 GPUs the kernels do not fill the machine, so the im2col +
    # wgrad GEMM + scatter of a layer overlap the dgrad / BN chain of the layers below it.
    use_side_stream = os.environ.get("FB_SIDE_STREAM", "1") != "0"

    def side(self, *keep):
        import contextlib
        if not (self.use_side_stream and torch.cuda.is_available()):
            return contextlib.nullcontext()
        if getattr(self, "_sides", None) is None:
            n = max(1, int(os.environ.get("FB_SIDE_STREAMS", "1")))
            self._sides, self._keep, self._forked, self._side_rr = [torch.cuda.Stream() for _ in range(n)], [], False, 0
        side = self._sides[self._side_rr % len(self._sides)]      # round-robin over the branch streams
        self._side_rr += 1
        side.wait_stream(torch.cuda.current_stream())
        self._keep.extend(keep)          # main-stream tensors the branch reads stay allocated until join()
        self._forked = True
        return torch.cuda.stream(side)

    def join(self):
        if getattr(self, "_forked", False):
            for side in self._sides:
                torch.cuda.current_stream().wait_stream(side)
            self._keep.clear()
            self._forked = False
            self._side_rr = 0

    @staticmethod
    def _bn_ok(ch):
        return 8 <= ch <= 2048 and 2048 % ch == 0

    # BatchNorm batch statistics produced by the epilogue of the convolution GEMM instead of a statistics pass over the
    # activation.  OPT-IN (FB_EPILOGUE_STATS=1): correct (tests/test_gpu_tf32.py) but measured SLOWER on the headline bench
    # (tf32 113.9 -> 115.4 ms/round, bf16 69.3 -> 73.4 with the statistics variant held to 2 CTAs per SM): these skinny GEMMs (N = 16..64, K = 16..256) are bound by their
    # four epilogue warps, so 32 shared-memory reads + 64 FMAs per 32 x 32 output box cost more there than the saved
    # read of the activation costs in the BatchNorm kernel.  The atomic column sums also make runs non-reproducible.
    epilogue_stats = os.environ.get("FB_EPILOGUE_STATS", "0") == "1"

    def gemm_nt(self, a, b, out_dtype=None, relu=False, stats=None):
        if self.use_tc_gemm and self.ops.grouped_gemm_supported(a, b):
            n = b.shape[1]
            ok = stats is not None and self.epilogue_stats and (n * (4 if a.dtype == torch.float32 else (out_dtype or a.dtype).itemsize)) % 16 == 0
            self.stats_done = ok
            return self.ops.grouped_gemm(a, b, out_dtype or a.dtype, relu=relu, stats=stats if ok else None)
        self._fb("gemm_nt")
        return super().gemm_nt(a, b, out_dtype, relu)

    want_transposed_weights = True

    def gemm_nn(self, a, b, out_dtype=None, b_t=None):
        # [C,M,N] x [C,N,K]: the tensor-core kernel wants K-major operands -> a x (b^T)^T with the b^T copy
        # the layer prepared together with the weights
        # [C,M,N] x [C,N,K]: the NN operand form feeds b MN-major straight from the [out][in] weights, so no transposed
        # staging copy is needed; a prepared b^T (K-major) is used when the NN form does not cover the layout
        if self.use_tc_gemm and b_t is not None and os.environ.get("FB_DGRAD_NN", "1") == "0" \
                and self.ops.grouped_gemm_supported(a, b_t):
            return self.ops.grouped_gemm(a, b_t, out_dtype or a.dtype)
        if self.use_tc_gemm and self._tf().mm_supported("nn", a, b):
            return self._tf().mm("nn", a, b, out_dtype=out_dtype or a.dtype)
        if self.use_tc_gemm and b_t is not None and self.ops.grouped_gemm_supported(a, b_t):
            return self.ops.grouped_gemm(a, b_t, out_dtype or a.dtype)
        self._fb("gemm_nn")
        return super().gemm_nn(a, b, out_dtype)

    def gemm_tn(self, a, b, out_dtype=torch.float32, out=None, out_is_zeroed=False):
        if self.use_tc_gemm and out_dtype == torch.float32 and self.ops.grouped_gemm_tn_supported(a, b):
            r = self.ops.grouped_gemm_tn(a, b, out=out, out_is_zeroed=out_is_zeroed)
            if out is not None and r is not out:
                out.copy_(r)
                return out
            return r
        self._fb("gemm_tn")
        return super().gemm_tn(a, b, out_dtype, out)

    # split-K weight-gradient slices reduce-add into a gradient arena zeroed once per step (FB_PREZERO_GRADS=0:
    # fp32 partial buffers + a separate sum, for A/B timing)
    grads_prezeroed = os.environ.get("FB_PREZERO_GRADS", "1") != "0"

    _ACT = (torch.bfloat16, torch.float32)      # activation dtypes the native kernels are instantiated for

    def im2col3x3(self, x, stride):
        if x.dtype in self._ACT and x.shape[-1] % 8 == 0:
            return self.ops.im2col3x3(x, stride)
        self._fb("im2col")
        return super().im2col3x3(x, stride)

    def col2im3x3(self, dcols, B, H, W, ch, stride):
        if dcols.dtype in self._ACT and ch % 8 == 0:
            return self.ops.col2im3x3(dcols, B, H, W, ch, stride)
        self._fb("col2im")
        return super().col2im3x3(dcols, B, H, W, ch, stride)

    def bn_fwd(self, x, res, gamma, beta, rm, rv, nbt, active, momentum, eps, relu, training, stats=None):
        if x.dtype in self._ACT and self._bn_ok(x.shape[-1]):
            return self.ops.bn_fwd_stacked(x, res, gamma, beta, rm, rv, nbt, active, momentum, eps, relu, training, stats=stats)
        self._fb("bn_fwd")
        return super().bn_fwd(x, res, gamma, beta, rm, rv, nbt, active, momentum, eps, relu, training)

    def bn_bwd(self, dy, x, y, gamma, mean, rstd, relu, want_dres, beta=None, had_res=True, dgamma_out=None,
               dbeta_out=None):
        if x.dtype in self._ACT and self._bn_ok(x.shape[-1]):
            skip_y = relu and not had_res and beta is not None   # mask recomputed from x: one stream less
            d1, d2 = dy if isinstance(dy, tuple) else (dy, None)
            return self.ops.bn_bwd_stacked(d1, x, None if skip_y else y, gamma, mean, rstd, relu, want_dres, beta=beta,
                                           dgamma_out=dgamma_out, dbeta_out=dbeta_out, dy2=d2)   # grads land in the arena
        self._fb("bn_bwd")
        return super().bn_bwd(dy, x, y, gamma, mean, rstd, relu, want_dres, dgamma_out=dgamma_out, dbeta_out=dbeta_out)

    def gn_fwd(self, x, res, gamma, beta, cpg, eps, relu):
        if x.dtype in self._ACT and self._bn_ok(x.shape[-1]) and x.shape[0] * x.shape[1] <= 65535 and cpg >= 2:
            return self.ops.gn_fwd_stacked(x, res, gamma, beta, cpg, eps, relu)
        self._fb("gn_fwd")
        return super().gn_fwd(x, res, gamma, beta, cpg, eps, relu)

    def gn_bwd(self, dy, x, y, gamma, mean, rstd, cpg, relu, want_dres, beta=None, dgamma_out=None, dbeta_out=None):
        if x.dtype in self._ACT and self._bn_ok(x.shape[-1]) and x.shape[0] * x.shape[1] <= 65535 and cpg >= 2:
            d1, d2 = dy if isinstance(dy, tuple) else (dy, None)
            return self.ops.gn_bwd_stacked(d1, x, y, gamma, mean, rstd, cpg, relu, want_dres, beta=beta, dgamma_out=dgamma_out,
                                           dbeta_out=dbeta_out, dy2=d2)
        self._fb("gn_bwd")
        return super().gn_bwd(dy, x, y, gamma, mean, rstd, cpg, relu, want_dres, beta, dgamma_out, dbeta_out)

    fused_head = os.environ.get("FB_FUSED_HEAD", "1") != "0"

    def head_fwd_bwd(self, h, fc_w, fc_b, labels, row_scale, g_w, g_b):
        C, B = h.shape[0], h.shape[1]
        ch, ncls = h.shape[-1], fc_w.shape[1]
        if self.fused_head and h.dtype in self._ACT and h.is_contiguous() and ch % 4 == 0 and B <= 512 and ncls <= 512 \
                and (B * (ch + ncls + 1)) * 4 <= 200 * 1024 and fc_w.stride(1) == ch and g_w.stride(1) == ch:
            return self.ops.head_fwd_bwd(h, fc_w, fc_b, labels, row_scale, g_w, g_b)
        self._fb("head")
        return super().head_fwd_bwd(h, fc_w, fc_b, labels, row_scale, g_w, g_b)

    def xent(self, logits, labels, pre_sigmoid=False, row_scale=None):
        C, B, ncls = logits.shape
        scale = torch.full((C,), 1.0 / B, dtype=torch.float32, device=logits.device) if row_scale is None else row_scale
        loss_rows, grad = self.ops.softmax_xent(logits.reshape(C * B, ncls).contiguous(), labels.reshape(-1),
                                                ignore_index=-1, pre_sigmoid=pre_sigmoid, row_scale=scale,
                                                rows_per_group=B)
        return loss_rows.view(C, B).sum(dim=1) * scale, grad.view(C, B, ncls)

    # ---- transformer primitives: csrc/transformer.cu + the NT/NN/TN grouped GEMM ------------
    def _tf(self):
        t = getattr(self, "_tf_ops", None)
        if t is None:
            from ..ops import transformer as t
            self._tf_ops = t
        return t

    def mm(self, mode, a, b, out=None, out_dtype=None, bias=None, relu=False):
        t = self._tf()
        if self.use_tc_gemm and t.mm_supported(mode, a, b, out) and (bias is None or a.dim() == 3):
            return t.mm(mode, a, b, out=out, out_dtype=out_dtype or a.dtype, bias=bias, relu=relu)
        self._fb("mm_" + mode)
        return super().mm(mode, a, b, out, out_dtype, bias, relu)

    def embed_fwd(self, ids, word, pos, S):
        if self.dtype == torch.bfloat16 and word.shape[2] % 8 == 0:
            return self._tf().embed_gather(ids, word, pos, S)
        self._fb("embed_fwd")
        return super().embed_fwd(ids, word, pos, S)

    def embed_bwd(self, ids, dz, gword, gpos, S, word_scale=None):
        if dz.dtype == torch.bfloat16 and gword.shape[2] % 8 == 0:
            return self._tf().embed_bwd(ids, dz, gword, gpos, S, word_scale)
        self._fb("embed_bwd")
        return super().embed_bwd(ids, dz, gword, gpos, S, word_scale)

    @staticmethod
    def _ln_ok(x):
        return x.dtype == torch.bfloat16 and x.shape[-1] % 8 == 0 and x.shape[-1] <= 1024

    def ln_fwd(self, a, b, gamma, beta, eps):
        if self._ln_ok(a):
            return self._tf().add_ln_fwd(a, b, gamma, beta, eps)
        self._fb("ln_fwd")
        return super().ln_fwd(a, b, gamma, beta, eps)

    def ln_bwd(self, dy, dy2, z, mean, rstd, gamma, dgamma_out, dbeta_out):
        if self._ln_ok(z):
            return self._tf().ln_bwd(dy, dy2, z, mean, rstd, gamma, dgamma_out, dbeta_out)
        self._fb("ln_bwd")
        return super().ln_bwd(dy, dy2, z, mean, rstd, gamma, dgamma_out, dbeta_out)

    def dropout(self, x, p, seed, site, step_dev=None, x2=None):
        if p <= 0 and x2 is None:
            return x
        if x.dtype in self._ACT and x.numel() % 8 == 0:
            return self._tf().dropout(x, p, seed, site, step_dev, x2)
        self._fb("dropout")
        return super().dropout(x, p, seed, site, step_dev, x2)

    def softmax_fwd(self, scores, key_mask, H, scale, p=0.0, seed=0, site=0, step_dev=None):
        S = scores.shape[-1]
        if self.dtype == torch.bfloat16 and scores.dtype == torch.float32 and S % 8 == 0 and S <= 512:
            return self._tf().attn_softmax_fwd(scores.contiguous(), key_mask, H, scale, p, seed, site, step_dev)
        self._fb("softmax_fwd")
        return super().softmax_fwd(scores, key_mask, H, scale, p, seed, site, step_dev)

    def softmax_bwd(self, dP, P, scale, p=0.0, seed=0, site=0, step_dev=None):
        S = P.shape[-1]
        if P.dtype == torch.bfloat16 and dP.dtype == torch.float32 and S % 8 == 0 and S <= 512:
            return self._tf().attn_softmax_bwd(dP.contiguous(), P, scale, p, seed, site, step_dev)
        self._fb("softmax_bwd")
        return super().softmax_bwd(dP, P, scale, p, seed, site, step_dev)

    def act_fwd(self, x, kind):
        if x.dtype in self._ACT and x.numel() % 8 == 0:
            return self._tf().act_fwd(x, kind)
        self._fb("act_fwd")
        return super().act_fwd(x, kind)

    def act_bwd(self, dy, x, kind):
        if x.dtype in self._ACT and x.numel() % 8 == 0 and dy.dtype == x.dtype:
            return self._tf().act_bwd(dy, x, kind)
        self._fb("act_bwd")
        return super().act_bwd(dy, x, kind)

    def colsum(self, x, out):
        if x.dtype in self._ACT and x.shape[-1] % 8 == 0:
            return self._tf().colsum(x, out)
        self._fb("colsum")
        return super().colsum(x, out)

    def maxpool2x2_fwd(self, x, off=(0, 0), out_hw=None):
        if x.dtype in self._ACT and x.shape[-1] % 8 == 0:
            return self.ops.maxpool2x2_fwd(x, off, out_hw)
        self._fb("maxpool_fwd")
        return super().maxpool2x2_fwd(x, off, out_hw)

    def maxpool2x2_bwd(self, dy, x, off=(0, 0), relu_mask=False):
        if x.dtype in self._ACT and x.shape[-1] % 8 == 0 and dy.dtype == x.dtype:
            return self.ops.maxpool2x2_bwd(dy, x, off, relu_mask)
        self._fb("maxpool_bwd")
        return super().maxpool2x2_bwd(dy, x, off, relu_mask)


# ============================================================================ layers
class _Conv:
    def __init__(self, name, k, stride, cin, cout, need_dx=True, cin_pad=None):
        self.name, self.k, self.stride, self.cout, self.need_dx = name, k, stride, cout, need_dx
        self.cin_real = cin
        self.cin = cin_pad or cin            # channels of the activation tensor this conv consumes
        self.wname = name + ".weight"

    def prep(self, W, dtype, transposed=False):
        """fp32 master [C,cout,cin,k,k] -> compute-dtype matrix [C, cout, k*k*cin] in (kh,kw,cin) order
        (+ its transpose [C, k*k*cin, cout] when the back-end wants K-major dgrad operands)."""
        w = W[self.wname]
        C = w.shape[0]
        if self.cin != self.cin_real:
            w = F.pad(w, (0, 0, 0, 0, 0, self.cin - self.cin_real))
        if self.k == 1:
            wm = w.reshape(C, self.cout, self.cin).to(dtype)
        else:
            wm = w.permute(0, 1, 3, 4, 2).reshape(C, self.cout, self.k * self.k * self.cin).to(dtype)
        wt = wm.transpose(1, 2).contiguous() if (transposed and self.wants_transposed()) else None
        return wm, wt

    def wants_transposed(self):
        """Only the implicit stride-1 3x3 data gradient reads the transposed copy [k*k*cin, cout]; every other data
        gradient uses the NN GEMM form on the forward operand."""
        if os.environ.get("FB_DGRAD_NN", "1") == "0":          # A/B switch: K-major transposed copies for every data gradient
            return self.need_dx
        return self.need_dx and self.k == 3 and self.stride == 1

    def _cols(self, be, x):
        C, B, H, W_, _ = x.shape
        s = self.stride
        if self.k == 1:
            xs = x if s == 1 else x[:, :, ::s, ::s, :]
            Ho, Wo = xs.shape[2], xs.shape[3]
            return xs.reshape(C, B * Ho * Wo, self.cin), Ho, Wo
        Ho, Wo = (H - 1) // s + 1, (W_ - 1) // s + 1
        return be.im2col3x3(x, s), Ho, Wo

    def fwd(self, be, x, Wm, stats=None):
        """``stats``: zeroed fp32 [C,2,cout]; a back-end that can fills it with the output's column sums / sums of squares
        (``be.stats_done`` says whether it did), which spares the BatchNorm that follows its statistics pass."""
        C, B = x.shape[0], x.shape[1]
        be.stats_done = False
        if self.k == 3 and be.conv3x3_ok(x, self.cout, self.stride):
            if stats is not None and getattr(be, "epilogue_stats", False):
                return be.conv3x3(x, Wm[0], self.stride, stats=stats)
            return be.conv3x3(x, Wm[0], self.stride)                  # implicit GEMM: no im2col buffer
        cols, Ho, Wo = self._cols(be, x)
        if stats is not None:
            return be.gemm_nt(cols, Wm[0], stats=stats).view(C, B, Ho, Wo, self.cout)
        return be.gemm_nt(cols, Wm[0]).view(C, B, Ho, Wo, self.cout)

    def _wgrad(self, be, dy2, x, gW, dy=None):
        C = x.shape[0]
        pz = getattr(self, "_pz", be.grads_prezeroed)                 # gradient arena / scratch zeroed for this step?
        scratch = getattr(self, "_dw_scratch", None)
        if scratch is not None and scratch.shape[0] != C:
            scratch = None
        defer = scratch is not None and getattr(self, "_defer_scatter", False)   # one un-staging launch per step instead
        if self.k == 3 and dy is not None and be.conv3x3_wgrad_ok(x, dy, self.stride):
            dWm = be.conv3x3_wgrad(x, dy, self.stride, out=scratch, out_is_zeroed=pz and scratch is not None)
            if not defer:                                             # implicit GEMM: shifted windows gathered by TMA
                gW.copy_(dWm.view(C, self.cout, 3, 3, self.cin)[..., :self.cin_real].permute(0, 1, 4, 2, 3))
            return
        cols, _, _ = self._cols(be, x)
        if self.k == 1 and self.cin == self.cin_real:
            # straight into the gradient arena (split-K slices reduce-add in place when it was zeroed for the step)
            be.gemm_tn(dy2, cols, out=gW.view(C, self.cout, self.cin), out_is_zeroed=pz)
            return
        if scratch is not None:
            dWm = be.gemm_tn(dy2, cols, out=scratch, out_is_zeroed=pz)
        else:
            dWm = be.gemm_tn(dy2, cols)                               # [C, cout, kk*cin] fp32
        if defer:
            return
        if self.k == 1:
            gW.copy_(dWm.view(C, self.cout, self.cin, 1, 1)[:, :, :self.cin_real])
        else:
            gW.copy_(dWm.view(C, self.cout, self.k, self.k, self.cin)[..., :self.cin_real].permute(0, 1, 4, 2, 3))

    def bwd(self, be, dy, x, Wm, gW):
        C, B, H, W_, _ = x.shape
        s = self.stride
        Ho, Wo = (H - 1) // s + 1, (W_ - 1) // s + 1
        dy2 = dy.reshape(C, B * Ho * Wo, self.cout)
        with be.side(dy2, x):                                         # im2col + wgrad GEMM + scatter: off the dgrad chain
            self._wgrad(be, dy2, x, gW, dy if dy.is_contiguous() else None)
        if not self.need_dx:
            return None
        if self.k == 3 and s == 1 and Wm[1] is not None and be.conv3x3_ok(dy, self.cin, 1):
            return be.conv3x3_dgrad(dy, Wm[1])                        # implicit GEMM with flipped taps: no col2im
        dcols = be.gemm_nn(dy2, Wm[0], b_t=Wm[1])                     # [C, M, kk*cin]
        if self.k == 1:
            if self.stride == 1:
                return dcols.view(C, B, H, W_, self.cin)
            dx = torch.zeros(C, B, H, W_, self.cin, dtype=dcols.dtype, device=dcols.device)
            dx[:, :, ::self.stride, ::self.stride, :] = dcols.view(C, B, Ho, Wo, self.cin)
            return dx
        return be.col2im3x3(dcols, B, H, W_, self.cin, self.stride)


class _BN:
    """BatchNorm (cpg = 0) or GroupNorm with ``cpg`` channels per group, each with fused residual add and ReLU."""

    def __init__(self, name, ch, momentum=0.1, eps=1e-5, cpg=0):
        self.name, self.ch, self.momentum, self.eps, self.cpg = name, ch, momentum if momentum is not None else 0.1, eps, cpg

    @classmethod
    def of(cls, name, mod):
        if isinstance(mod, nn.GroupNorm):
            return cls(name, mod.num_channels, None, mod.eps, cpg=mod.num_channels // mod.num_groups)
        return cls(name, mod.num_features, mod.momentum, mod.eps)

    def fwd(self, be, x, W, active, relu, res=None, training=True, stats=None):
        n = self.name
        if self.cpg:
            return be.gn_fwd(x, res, W[n + ".weight"], W[n + ".bias"], self.cpg, self.eps, relu)
        if stats is not None:
            return be.bn_fwd(x, res, W[n + ".weight"], W[n + ".bias"], W[n + ".running_mean"], W[n + ".running_var"],
                             W[n + ".num_batches_tracked"], active, self.momentum, self.eps, relu, training, stats=stats)
        return be.bn_fwd(x, res, W[n + ".weight"], W[n + ".bias"], W[n + ".running_mean"], W[n + ".running_var"],
                         W[n + ".num_batches_tracked"], active, self.momentum, self.eps, relu, training)

    def bwd(self, be, dy, x, y, W, G, mean, rstd, relu, want_dres):
        n = self.name
        if self.cpg:
            dx, dres, _, _ = be.gn_bwd(dy, x, y, W[n + ".weight"], mean, rstd, self.cpg, relu, want_dres, beta=W[n + ".bias"],
                                       dgamma_out=G[n + ".weight"], dbeta_out=G[n + ".bias"])
            return dx, dres
        dx, dres, _, _ = be.bn_bwd(dy, x, y, W[n + ".weight"], mean, rstd, relu, want_dres, beta=W[n + ".bias"],
                                   had_res=want_dres, dgamma_out=G[n + ".weight"], dbeta_out=G[n + ".bias"])
        return dx, dres


class _Block:
    """BasicBlock (2 convs) or Bottleneck (3 convs) with optional conv+BN downsample."""

    def __init__(self, prefix: str, mod: nn.Module):
        self.prefix = prefix
        self.convs: List[_Conv] = []
        self.bns: List[_BN] = []
        n = 3 if isinstance(mod, Bottleneck) else 2
        for i in range(1, n + 1):
            c = getattr(mod, f"conv{i}")
            self.convs.append(_Conv(f"{prefix}.conv{i}", c.kernel_size[0], c.stride[0], c.in_channels, c.out_channels))
            self.bns.append(_BN.of(f"{prefix}.bn{i}", getattr(mod, f"bn{i}")))
        self.ds = None
        if mod.downsample is not None:
            c, b = mod.downsample[0], mod.downsample[1]
            self.ds = (_Conv(f"{prefix}.downsample.0", 1, c.stride[0], c.in_channels, c.out_channels),
                       _BN.of(f"{prefix}.downsample.1", b))

    def all_convs(self):
        return self.convs + ([self.ds[0]] if self.ds else [])

    def fwd(self, be, x, W, Wm, active, training, tape):
        saved = {"x": x}
        h = x
        n = len(self.convs)
        st = getattr(self, "_stats", None) if training else None      # {bn name: zeroed [C,2,ch] slot}, set per step
        if self.ds is not None:
            dconv, dbn = self.ds
            slot = st.get(dbn.name) if st else None
            a = dconv.fwd(be, x, Wm[dconv.name], stats=slot)
            idt, m, r = dbn.fwd(be, a, W, active, relu=False, training=training, stats=slot if be.stats_done else None)
            saved["ds"] = (a, idt, m, r)
        else:
            idt = x
        for i, (conv, bn) in enumerate(zip(self.convs, self.bns)):
            slot = st.get(bn.name) if st else None
            a = conv.fwd(be, h, Wm[conv.name], stats=slot)
            last = i == n - 1
            y, m, r = bn.fwd(be, a, W, active, relu=True, res=idt if last else None, training=training,
                             stats=slot if be.stats_done else None)
            saved[i] = (h, a, y, m, r)
            h = y
        if tape is not None:
            tape.append(saved)
        return h

    def bwd(self, be, dy, W, Wm, G, saved):
        n = len(self.convs)
        d = dy
        d_idt = None
        for i in reversed(range(n)):
            conv, bn = self.convs[i], self.bns[i]
            h, a, y, m, r = saved[i]
            last = i == n - 1
            da, dres = bn.bwd(be, d, a, y, W, G, m, r, relu=True, want_dres=last)
            if last:
                d_idt = dres
            d = conv.bwd(be, da, h, Wm[conv.name], G[conv.wname])
        if self.ds is not None:
            dconv, dbn = self.ds
            a, idt, m, r = saved["ds"]
            da, _ = dbn.bwd(be, d_idt, a, idt, W, G, m, r, relu=False, want_dres=False)
            d = (d, dconv.bwd(be, da, saved["x"], Wm[dconv.name], G[dconv.wname]))
        else:
            d = (d, d_idt)
        return d          # (main path, residual path): the consumer (a BN backward) sums them while loading


# ============================================================================ models
class BatchedModel:
    """Interface the NCCL simulator drives."""
    input_kind = "flat"

    def prepare(self, W):                       # per-step weight staging (dtype casts / re-layouts)
        raise NotImplementedError

    def train_step(self, x, y, W, G, active, row_scale=None):   # -> per-client mean loss [C]; grads written into G
        raise NotImplementedError

    def logits(self, x, W):                     # eval-mode forward
        raise NotImplementedError


class BatchedLogisticRegression(BatchedModel):
    """C logistic-regression clients; sigmoid-then-CE numerics of the reference (lr.py:10)."""

    def __init__(self, module: LogisticRegression, backend: TorchBackend):
        self.be = backend
        self.in_dim, self.out_dim = module.linear.in_features, module.linear.out_features
        self.pre_sigmoid = True

    def prepare(self, W):
        return {"w": W["linear.weight"].to(self.be.dtype)}

    def _fwd(self, x, W, Wm):
        C = x.shape[0]
        xin = x.reshape(C, x.shape[1], -1).to(self.be.dtype)
        return xin, self.be.gemm_nt(xin, Wm["w"], out_dtype=torch.float32) + W["linear.bias"][:, None, :]

    def train_step(self, x, y, W, G, active, row_scale=None):
        Wm = self.prepare(W)
        xin, logits = self._fwd(x, W, Wm)
        loss, dlog = self.be.xent(logits, y, pre_sigmoid=True, row_scale=row_scale)
        G["linear.weight"].copy_(self.be.gemm_tn(dlog.to(self.be.dtype), xin))
        G["linear.bias"].copy_(dlog.sum(dim=1))
        return loss

    def logits(self, x, W):
        return self._fwd(x, W, self.prepare(W))[1]


class BatchedResNet(BatchedModel):
    """CIFAR-stem / 3x3-stem ResNets (ResNet-56 Bottleneck, ResNet-18 BasicBlock) over C clients."""
    input_kind = "image_u8_nhwc"
    uses_batch_statistics = True      # BatchNorm: padding rows of a ragged batch would enter the statistics

    def __init__(self, module: ResNet, backend: TorchBackend, mean=None, std=None):
        self.be = backend
        if getattr(module, "stem", "cifar") != "cifar" and not isinstance(getattr(module, "maxpool", None), nn.Identity):
            raise NotImplementedError("batched ResNet supports 3x3 stems without max-pool (CIFAR-sized inputs)")
        c1 = module.conv1
        pad = 8 if (isinstance(backend, NativeBackend) and c1.in_channels % 8) else None
        self.stem_conv = _Conv("conv1", c1.kernel_size[0], c1.stride[0], c1.in_channels, c1.out_channels, need_dx=False,
                               cin_pad=pad)
        self.stem_bn = _BN.of("bn1", module.bn1)
        self.uses_batch_statistics = not self.stem_bn.cpg          # GroupNorm statistics are per sample: ragged batches are fine
        self.blocks: List[_Block] = []
        for li in range(1, module.num_stages + 1):
            for bi, blk in enumerate(getattr(module, f"layer{li}")):
                self.blocks.append(_Block(f"layer{li}.{bi}", blk))
        self.feat_dim, self.ncls = module.fc.in_features, module.fc.out_features
        self.convs = [self.stem_conv] + [c for b in self.blocks for c in b.all_convs()]
        self.mean = torch.tensor(mean if mean is not None else (0.4914, 0.4822, 0.4465), dtype=torch.float32)
        self.std = torch.tensor(std if std is not None else (0.2470, 0.2435, 0.2616), dtype=torch.float32)

    def prepare(self, W):
        dt = self.be.dtype
        tr = self.be.want_transposed_weights
        if isinstance(self.be, NativeBackend) and dt in NativeBackend._ACT and getattr(W, "arena", None) is not None:
            return self._prepare_staged(W)
        Wm = {c.name: c.prep(W, dt, tr) for c in self.convs}
        Wm["fc"] = W["fc.weight"].to(dt)
        return Wm

    def _prepare_staged(self, W):
        """All conv / fc weights of all clients -> GEMM operands in the compute dtype with ONE kernel launch.
        With fp32 (TF32) compute the 1x1-convolution and fc weights already ARE the GEMM operands ([cout][cin] rows of
        the arena, 16-byte aligned segments): they are used in place and only k x k / channel-padded layers are staged."""
        arena, layout = W.arena, W.layout
        C = arena.shape[0]
        key = (arena.data_ptr(), C)
        st = getattr(self, "_stage", None)
        if st is None or st["key"] != key:
            rows, off, plan = [], 0, {}
            al = lambda v: (v + 7) // 8 * 8   # noqa: E731  16-byte aligned segments for TMA
            fp32 = self.be.dtype == torch.float32
            in_place = {}
            for c in self.convs:
                seg = layout.by_name[c.wname]
                if fp32 and c.k == 1 and c.cin == c.cin_real and arena.stride(0) % 4 == 0:
                    in_place[c.name] = (W[c.wname].view(C, c.cout, c.cin), None)
                    continue
                n = c.cout * c.k * c.k * c.cin
                dst = off; off = al(off + n)
                dstT = -1
                if c.wants_transposed():
                    dstT = off; off = al(off + n)
                rows.append([seg.offset, c.cout, c.cin_real, c.cin, c.k, dst, dstT, n])
                plan[c.name] = (dst, dstT, c.cout, c.k * c.k * c.cin)
            seg = layout.by_name["fc.weight"]
            n = self.ncls * self.feat_dim
            if fp32 and arena.stride(0) % 4 == 0:
                in_place["fc"] = W["fc.weight"]
            else:
                rows.append([seg.offset, self.ncls, self.feat_dim, self.feat_dim, 1, off, -1, n])
                plan["fc"] = (off, -1, self.ncls, self.feat_dim)
                off = al(off + n)
            stage = torch.zeros((C, off), dtype=self.be.dtype, device=arena.device)
            table = torch.tensor(rows, dtype=torch.int64).to(arena.device)
            views = {}
            for name, (dst, dstT, cout, kdim) in plan.items():
                wm = stage[:, dst:dst + cout * kdim].view(C, cout, kdim)
                wt = stage[:, dstT:dstT + cout * kdim].view(C, kdim, cout) if dstT >= 0 else None
                views[name] = wm if name == "fc" else (wm, wt)
            views.update(in_place)
            st = self._stage = {"key": key, "stage": stage, "table": table, "views": views, "n": len(rows),
                                "max_n": max(r[7] for r in rows)}
        if not (getattr(self, "_stage_fresh", False) and st.get("fresh_key") == key):
            self.be.ops.stage_conv_weights(arena, st["stage"], st["table"], st["n"], st["max_n"])
        return st["views"]

    # ---- fused optimizer step + re-staging (plain SGD): see fb_sgd_restage ---------------------------------------
    supports_fused_sgd = True
    fused_lr = None

    def set_fused_sgd(self, lr_dev):
        """``lr_dev`` [C] (device; 0 freezes a client): the training step ends with ONE launch that applies
        w -= lr * g to every trainable segment and refreshes the bf16 GEMM operands, so the next step needs no staging
        pass and the k x k weight gradients need no un-staging pass.  None restores the separate optimizer step."""
        self.fused_lr = lr_dev

    def restage(self, W):
        """Stage the current weights now (round start, after the global model was written into the arena) and mark
        the staged copy as current; it stays current while only the fused step modifies the weights."""
        self._stage_fresh = False
        if isinstance(self.be, NativeBackend) and self.be.dtype in NativeBackend._ACT and getattr(W, "arena", None) is not None:
            self._prepare_staged(W)
            self._stage["fresh_key"] = self._stage["key"]
            self._stage_fresh = True

    def invalidate_stage(self):
        self._stage_fresh = False

    def _fused_plan(self, W, G):
        """Segment table of the fused step: staged conv / fc weights (gradients in the arena or, for k x k, in the
        GEMM-layout scratch) plus every other trainable segment as a plain row."""
        st = getattr(self, "_stage", None)
        big = getattr(self, "_dw_big", None)
        if st is None or G.arena is None:
            return None
        key = (st["key"], None if big is None else big.data_ptr())
        fp = getattr(self, "_fused", None)
        if fp is not None and fp["key"] == key:
            return fp
        layout = W.layout
        srow = {r[5]: r for r in getattr(self, "_unstage_rows", [])}      # arena offset -> un-staging row
        rows, staged_offsets = [], set()
        for r in st["table"].cpu().tolist():
            w_off, cout, cin, cinp, k, dst, dstT, n = r
            staged_offsets.add(w_off)
            if w_off in srow:
                rows.append([w_off, cout, cin, cinp, k, dst, dstT, n, srow[w_off][0], 1])
            elif k == 1 and cin == cinp:
                rows.append([w_off, cout, cin, cinp, k, dst, dstT, n, 0, 0])
            else:
                return None                                               # gradient layout not covered: keep the split path
        for sg in layout.segments:
            if sg.kind == "param" and sg.offset not in staged_offsets:
                rows.append([sg.offset, sg.numel, 1, 1, 1, -1, -1, sg.numel, 0, 0])
        fp = self._fused = {"key": key, "table": torch.tensor(rows, dtype=torch.int64).to(W.arena.device), "n": len(rows),
                            "max_n": min(sum(r[7] for r in rows), 2 ** 31 - 1)}
        return fp

    def _input(self, x):
        if x.dtype == torch.uint8:
            if self.mean.device != x.device:     # one-time move; must not happen inside a CUDA-graph capture
                self.mean, self.std = self.mean.to(x.device), self.std.to(x.device)
            x = self.be.normalize_images(x, self.mean, self.std)
        else:
            x = x.to(self.be.dtype)
        if self.stem_conv.cin != x.shape[-1]:
            x = F.pad(x, (0, self.stem_conv.cin - x.shape[-1]))       # zero channels so that K = 9*8 feeds TMA
        return x

    def _forward(self, x, W, Wm, active, training, tape, want_logits=True):
        be = self.be
        x = self._input(x)
        st = self._bn_stat_slots(x.shape[0], x.device) if (training and getattr(be, "epilogue_stats", False)) else None
        for blk in self.blocks:
            blk._stats = st
        slot = st.get(self.stem_bn.name) if st else None
        a = self.stem_conv.fwd(be, x, Wm["conv1"], stats=slot)
        h, m, r = self.stem_bn.fwd(be, a, W, active, relu=True, training=training, stats=slot if be.stats_done else None)
        if tape is not None:
            tape.append((x, a, h, m, r))
        for blk in self.blocks:
            h = blk.fwd(be, h, W, Wm, active, training, tape)
        if not want_logits:
            return None, None, h
        C, B = h.shape[0], h.shape[1]
        feat = h.float().mean(dim=(2, 3)).to(be.dtype)                 # [C,B,feat]
        logits = be.gemm_nt(feat, Wm["fc"], out_dtype=torch.float32) + W["fc.bias"][:, None, :]
        return logits, feat, h

    def _bn_stat_slots(self, C, device):
        """One zeroed fp32 buffer per step holding a [C, 2, ch] slot for every BatchNorm layer: the convolution that feeds
        the layer adds its output's per-channel sums / sums of squares in its epilogue (one memset per step)."""
        cache = getattr(self, "_bn_stats", None)
        if cache is None or cache[0].shape[0] != C or cache[0].device != device:
            bns = [self.stem_bn] + [b for blk in self.blocks for b in (blk.bns + ([blk.ds[1]] if blk.ds else []))]
            total = sum(2 * b.ch for b in bns)
            buf = torch.zeros((len(bns) and 1, C * total), dtype=torch.float32, device=device)
            slots, off = {}, 0
            for b in bns:
                slots[b.name] = buf[0, off:off + C * 2 * b.ch].view(C, 2, b.ch)
                off += C * 2 * b.ch
            cache = self._bn_stats = (buf, slots)
        cache[0].zero_()
        return cache[1]

    def _zero_grads(self, G):
        """Per-step gradient bookkeeping of the native back-end.  Weight-gradient GEMMs of k x k (and channel-padded)
        convolutions write the GEMM layout into one scratch stack; one un-staging launch per step scatters it into the
        gradient arena.  For models whose arena is small the arena and the scratch are zeroed (two memsets) so that
        split-K slices can reduce-add in place; for big models (ResNet-18 x 64 clients = 2.9 GB) zeroing would cost
        more than the partial-buffer reduction it avoids."""
        arena = getattr(G, "arena", None)
        if arena is None or not isinstance(self.be, NativeBackend):
            return
        C = arena.shape[0]
        P_train = G.layout.P_train
        pz = bool(self.be.grads_prezeroed) and C * P_train * 4 <= (1 << 30)
        big = getattr(self, "_dw_big", None)
        need = [c for c in self.convs if not (c.k == 1 and c.cin == c.cin_real)]
        if big is None or big.shape[0] != C:
            sizes = [(c.cout * c.k * c.k * c.cin + 3) // 4 * 4 for c in need]
            big = self._dw_big = torch.zeros((C, sum(sizes)), dtype=torch.float32, device=arena.device)
            off, rows = 0, []
            for c, n in zip(need, sizes):
                c._dw_scratch = big[:, off:off + c.cout * c.k * c.k * c.cin].view(C, c.cout, c.k * c.k * c.cin)
                seg = G.layout.by_name[c.wname]
                rows.append([off, c.cout, c.cin_real, c.cin, c.k, seg.offset, 0, 0])
                c._defer_scatter = c.k * c.k <= 9
                off += n
            self._unstage = None
            self._unstage_rows = rows
            if rows and all(c._defer_scatter for c in need):
                self._unstage = (torch.tensor(rows, dtype=torch.int64).to(arena.device), len(rows),
                                 max(((r[1] + 31) // 32) * ((r[2] + 31) // 32) for r in rows))
            else:
                for c in need:
                    c._defer_scatter = False
        elif pz:
            big.zero_()
        if pz:
            arena[:, :P_train].zero_()
        for c in self.convs:
            c._pz = pz

    def _scatter_grads(self, G):
        """GEMM-layout weight gradients -> parameter layout in the arena (one launch; on the weight-gradient branch)."""
        un = getattr(self, "_unstage", None)
        if un is not None:
            self.be.join()               # every weight-gradient GEMM of the branch streams has finished
            self.be.ops.unstage_conv_grads(self._dw_big, G.arena, un[0], un[1], un[2])

    def train_step(self, x, y, W, G, active, row_scale=None):
        be = self.be
        Wm = self.prepare(W)
        self._zero_grads(G)
        tape: list = []
        _, _, h_last = self._forward(x, W, Wm, active, True, tape, want_logits=False)
        # pool -> fc -> loss -> fc gradients -> d(h_last): one launch on the native back-end
        loss, d = be.head_fwd_bwd(h_last.contiguous(), W["fc.weight"], W["fc.bias"], y, row_scale, G["fc.weight"], G["fc.bias"])
        for blk, saved in zip(reversed(self.blocks), reversed(tape[1:])):
            d = blk.bwd(be, d, W, Wm, G, saved)
        x0, a, hstem, m, r = tape[0]
        da, _ = self.stem_bn.bwd(be, d, a, hstem, W, G, m, r, relu=True, want_dres=False)
        self.stem_conv.bwd(be, da, x0, Wm["conv1"], G["conv1.weight"])
        fp = self._fused_plan(W, G) if (self.fused_lr is not None and getattr(self, "_unstage", None) is not None) else None
        if fp is not None:
            be.join()                                                 # weight gradients complete
            be.ops.sgd_restage(W.arena, G.arena, self._dw_big, self._stage["stage"], fp["table"], fp["n"], fp["max_n"],
                               self.fused_lr)
            self._stage["fresh_key"] = self._stage["key"]
            self._stage_fresh = True
            self.fused_applied = True
        else:
            self._scatter_grads(G)
            be.join()
            self.fused_applied = False
        return loss

    def logits(self, x, W):
        C = x.shape[0]
        active = torch.zeros(C, dtype=torch.float32, device=x.device)
        return self._forward(x, W, self.prepare(W), active, False, None)[0]


def make_batched(module: nn.Module, backend: TorchBackend) -> Optional[BatchedModel]:
    """Batched executor for ``module`` or None when the family has no batched implementation yet
    (the simulator then falls back to per-client eager training on arena-bound parameters)."""
    if isinstance(module, LogisticRegression):
        return BatchedLogisticRegression(module, backend)
    if isinstance(module, ResNet) and not module.KD:
        try:
            return BatchedResNet(module, backend)
        except NotImplementedError:
            return None
    from .cv import CNN_DropOut
    if isinstance(module, CNN_DropOut):
        from .batched_cnn import BatchedCNNDropOut
        return BatchedCNNDropOut(module, backend)
    from .nlp import DistilBertForSequenceClassification
    if isinstance(module, DistilBertForSequenceClassification):
        from .batched_nlp import BatchedDistilBert
        try:
            return BatchedDistilBert(module, backend)
        except NotImplementedError:
            return None
    return None
