"""Client-batched ``CNN_DropOut`` (the FEMNIST / federated-EMNIST benchmark model, reference model/cv/cnn.py:74-142).

C clients train as one network with a leading client dimension, explicit forward / backward, NHWC activations:

    conv 3x3 (1 -> 32, valid) ReLU | conv 3x3 (32 -> 64, valid) ReLU | max-pool 2 | dropout .25 | fc 9216 -> 128 ReLU |
    dropout .5 | fc 128 -> classes | softmax cross-entropy

Convolutions are im2col + grouped GEMM with the bias / ReLU epilogue; a valid convolution is the centre crop of the
same-padded one, so conv 2 keeps its 26x26 same-padded output and the pooling kernel reads the 24x24 interior through a window
offset (``maxpool2x2_fwd(off=(1, 1))``), its backward writes the 26x26 gradient with a zero border and the ReLU mask already
applied.  The single input channel is padded to 8 (16-byte vectors), the class dimension to a multiple of 16.  ``linear_1`` is
staged in (h, w, c) column order because activations are NHWC.  Dropout masks are Philox-keyed by (seed, device step
counter, site): a captured CUDA graph draws new masks on every replay and the backward pass re-derives the forward's mask.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .batched import BatchedModel, TorchBackend


class BatchedCNNDropOut(BatchedModel):
    input_kind = "flat"

    def __init__(self, module, backend: TorchBackend, seed: int = 0):
        self.be = backend
        self.ncls = module.linear_2.out_features
        self.ncls_pad = (self.ncls + 15) // 16 * 16
        self.p1, self.p2 = float(module.dropout_1.p), float(module.dropout_2.p)
        self.seed = int(seed)
        self._step_dev = None
        self._bias_pad = None

    # ---- weights: fp32 master views [C, ...] of the arena -> GEMM operands in the compute dtype
    def prepare(self, W):
        dt = self.be.dtype
        C = W["conv2d_1.weight"].shape[0]
        w1 = F.pad(W["conv2d_1.weight"], (0, 0, 0, 0, 0, 7)).permute(0, 1, 3, 4, 2).reshape(C, 32, 72)       # (kh, kw, cin padded to 8)
        w2 = W["conv2d_2.weight"].permute(0, 1, 3, 4, 2).reshape(C, 64, 288)
        f1 = W["linear_1.weight"].view(C, 128, 64, 12, 12).permute(0, 1, 3, 4, 2).reshape(C, 128, 9216)      # columns in (h, w, c) order
        f2 = F.pad(W["linear_2.weight"], (0, 0, 0, self.ncls_pad - self.ncls))
        if self._bias_pad is None or self._bias_pad.shape[0] != C or self._bias_pad.device != f2.device:
            self._bias_pad = torch.zeros(C, self.ncls_pad, dtype=torch.float32, device=f2.device)
        self._bias_pad[:, :self.ncls].copy_(W["linear_2.bias"])
        return {"w1": w1.to(dt).contiguous(), "w2": w2.to(dt).contiguous(), "f1": f1.to(dt).contiguous(), "f2": f2.to(dt).contiguous(),
                "b1": W["conv2d_1.bias"], "b2": W["conv2d_2.bias"], "fb1": W["linear_1.bias"], "fb2": self._bias_pad}

    def _step_counter(self, device):
        if self._step_dev is None or self._step_dev.device != device:
            self._step_dev = torch.zeros(1, dtype=torch.int32, device=device)
        return self._step_dev

    def _input(self, x):
        C, B = x.shape[0], x.shape[1]
        x8 = torch.zeros(C, B, 28, 28, 8, dtype=self.be.dtype, device=x.device)
        x8[..., 0] = x.reshape(C, B, 28, 28).to(self.be.dtype)
        return x8

    def _forward(self, x, Wm, training):
        be = self.be
        C, B = x.shape[0], x.shape[1]
        x8 = self._input(x)
        cols1 = be.im2col3x3(x8, 1).reshape(C, B * 784, 72)
        a1f = be.mm("nt", cols1, Wm["w1"], bias=Wm["b1"], relu=True).view(C, B, 28, 28, 32)
        a1 = a1f[:, :, 1:27, 1:27, :].contiguous()                                   # valid 26x26
        cols2 = be.im2col3x3(a1, 1).reshape(C, B * 676, 288)
        a2f = be.mm("nt", cols2, Wm["w2"], bias=Wm["b2"], relu=True).view(C, B, 26, 26, 64)     # valid part = [1:25, 1:25]
        pooled = be.maxpool2x2_fwd(a2f, (1, 1), (12, 12))
        d1 = be.dropout(pooled, self.p1 if training else 0.0, self.seed, 1, self._step_dev)
        hin = d1.reshape(C, B, 9216)
        h1 = be.mm("nt", hin, Wm["f1"], bias=Wm["fb1"], relu=True)
        d2 = be.dropout(h1, self.p2 if training else 0.0, self.seed, 2, self._step_dev)
        logits = be.mm("nt", d2, Wm["f2"], bias=Wm["fb2"], out_dtype=torch.float32)[..., :self.ncls]
        return logits, (cols1, a1, cols2, a2f, hin, h1, d2)

    def train_step(self, x, y, W, G, active, row_scale=None):
        be = self.be
        C, B = x.shape[0], x.shape[1]
        self._step_counter(x.device).add_(1)                       # fresh dropout masks on every (graph-replayed) step
        Wm = self.prepare(W)
        logits, (cols1, a1, cols2, a2f, hin, h1, d2) = self._forward(x, Wm, True)
        loss, dlog = be.xent(logits.contiguous(), y, row_scale=row_scale)
        dt = be.dtype
        dlp = F.pad(dlog, (0, self.ncls_pad - self.ncls)).to(dt)                     # [C, B, ncls_pad]
        # fc 2
        G["linear_2.weight"].copy_(be.mm("tn", dlp, d2, out_dtype=torch.float32)[:, :self.ncls])
        G["linear_2.bias"].copy_(dlog.sum(dim=1))
        dd2 = be.mm("nn", dlp, Wm["f2"])                                             # [C, B, 128]
        dh1 = be.act_bwd(be.dropout(dd2, self.p2, self.seed, 2, self._step_dev), h1, "relu")
        # fc 1 (gradient columns back from (h, w, c) to the module's (c, h, w) order)
        g1 = be.mm("tn", dh1, hin, out_dtype=torch.float32)                          # [C, 128, 9216]
        G["linear_1.weight"].copy_(g1.view(C, 128, 12, 12, 64).permute(0, 1, 4, 2, 3).reshape(C, 128, 9216))
        be.colsum(dh1, G["linear_1.bias"])
        dhin = be.mm("nn", dh1, Wm["f1"]).view(C, B, 12, 12, 64)
        dpool = be.dropout(dhin, self.p1, self.seed, 1, self._step_dev)
        # pool + ReLU of conv 2 (26x26 gradient map, zero border)
        da2 = be.maxpool2x2_bwd(dpool, a2f, (1, 1), relu_mask=True).reshape(C, B * 676, 64)
        g2 = be.mm("tn", da2, cols2, out_dtype=torch.float32)                        # [C, 64, 288]
        G["conv2d_2.weight"].copy_(g2.view(C, 64, 3, 3, 32).permute(0, 1, 4, 2, 3))
        be.colsum(da2, G["conv2d_2.bias"])
        dcols2 = be.mm("nn", da2, Wm["w2"])                                          # [C, B*676, 288]
        da1 = be.act_bwd(be.col2im3x3(dcols2, B, 26, 26, 32, 1), a1, "relu")
        da1f = F.pad(da1, (0, 0, 1, 1, 1, 1)).reshape(C, B * 784, 32)                 # back onto the 28x28 same-padded grid
        g1c = be.mm("tn", da1f, cols1, out_dtype=torch.float32)                      # [C, 32, 72]
        G["conv2d_1.weight"].copy_(g1c.view(C, 32, 3, 3, 8)[..., :1].permute(0, 1, 4, 2, 3))
        be.colsum(da1f, G["conv2d_1.bias"])
        return loss

    def logits(self, x, W):
        self._step_counter(x.device)
        return self._forward(x, self.prepare(W), False)[0]
