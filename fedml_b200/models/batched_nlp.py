"""Client-batched DistilBERT sequence classifier: explicit forward / backward over C co-resident clients.

Reference behaviour: FedNLP text classification trains one HuggingFace ``DistilBertForSequenceClassification`` per
client, one after another, through autograd (``python/app/fednlp`` trainers on top of
``simulation/nccl/base_framework`` / ``my_model_trainer_classification.py:21-70``).  Here every sampled client on
the GPU advances through the same step together:

* all projections (fused QKV, attention output, both FFN layers, pre-classifier, classifier) and all of their
  data / weight gradients are grouped GEMMs with one group per client (``be.mm`` -> tcgen05 kernel, K9);
* attention runs as grouped GEMMs with two-level (client*batch, head) groups that address the Q / K / V heads in
  place inside the fused ``[C, B*S, 3*D]`` projection output and write the context / gradient heads back in place:
  no split-heads or merge-heads copies in either direction;
* embeddings, LayerNorm (+residual), softmax (+key mask, +attention dropout), GELU, dropout and bias gradients are
  the single-pass K12 kernels (``csrc/transformer.cu``);
* weight gradients land directly in the rows of the stacked gradient arena.

The same code runs on :class:`TorchBackend` (fp32, CPU) where it is checked against autograd of
``fedml_b200.models.nlp.DistilBertForSequenceClassification``.
"""
from __future__ import annotations

import math
from typing import Dict, List

import torch

from .batched import BatchedModel, NativeBackend, TorchBackend
from .nlp import DistilBertForSequenceClassification


class BatchedDistilBert(BatchedModel):
    input_kind = "token_ids"

    def __init__(self, module: DistilBertForSequenceClassification, backend: TorchBackend, seed: int = 0):
        self.be = backend
        emb = module.distilbert.embeddings
        blk0 = module.distilbert.transformer.layer[0]
        self.vocab, self.D = emb.word_embeddings.num_embeddings, emb.word_embeddings.embedding_dim
        self.max_pos = emb.position_embeddings.num_embeddings
        self.H = blk0.attention.n_heads
        self.dh = self.D // self.H
        self.F = blk0.ffn.lin1.out_features
        self.L = len(module.distilbert.transformer.layer)
        self.ncls = module.classifier.out_features
        self.eps = emb.LayerNorm.eps
        self.p_hidden = float(emb.dropout.p)
        self.p_attn = float(blk0.attention.p)
        self.p_cls = float(module.dropout.p)
        self.seed = int(seed)
        self.pad_id = 0
        self._step_dev = None
        self.sparse_lr = None
        if self.D % 8 or self.dh % 8 or self.F % 8:
            raise NotImplementedError("batched DistilBERT needs dim, head dim and FFN dim to be multiples of 8")

    # ------------------------------------------------------------------ names
    @staticmethod
    def _ln(i):
        return f"distilbert.transformer.layer.{i}."

    def _linears(self):
        """(staging key, [parameter names concatenated along the output dim])"""
        out = []
        for i in range(self.L):
            p = self._ln(i)
            out.append((f"qkv{i}", [p + "attention.q_lin", p + "attention.k_lin", p + "attention.v_lin"]))
            out.append((f"o{i}", [p + "attention.out_lin"]))
            out.append((f"f1{i}", [p + "ffn.lin1"]))
            out.append((f"f2{i}", [p + "ffn.lin2"]))
        out.append(("pre", ["pre_classifier"]))
        out.append(("cls", ["classifier"]))
        return out

    # ------------------------------------------------------------------ weight staging
    def prepare(self, W):
        if isinstance(self.be, NativeBackend) and self.be.dtype == torch.bfloat16 and getattr(W, "arena", None) is not None:
            Wm = self._prepare_staged(W)
        else:
            dt = self.be.dtype
            Wm = {key: torch.cat([W[n + ".weight"] for n in names], dim=1).to(dt) if len(names) > 1
                  else W[names[0] + ".weight"].to(dt) for key, names in self._linears()}
        for i in range(self.L):                                   # fused fp32 bias of the QKV projection
            p = self._ln(i) + "attention."
            Wm[f"bqkv{i}"] = torch.cat([W[p + "q_lin.bias"], W[p + "k_lin.bias"], W[p + "v_lin.bias"]], dim=1)
        return Wm

    def _prepare_staged(self, W):
        """fp32 arena rows -> bf16 GEMM operands of every linear layer of every client with ONE launch
        (the k = 1 case of the conv-weight staging kernel; Q, K and V land next to each other)."""
        arena, layout = W.arena, W.layout
        C = arena.shape[0]
        key = (arena.data_ptr(), C)
        st = getattr(self, "_stage", None)
        if st is None or st["key"] != key:
            rows, off, plan = [], 0, {}
            for skey, names in self._linears():
                start, n_out, n_in = off, 0, None
                for n in names:
                    seg = layout.by_name[n + ".weight"]
                    o, i_ = seg.shape
                    rows.append([seg.offset, o, i_, i_, 1, off, -1, o * i_])
                    off += o * i_
                    n_out += o
                    n_in = i_
                plan[skey] = (start, n_out, n_in)
                off = (off + 7) // 8 * 8
            stage = torch.zeros((C, off), dtype=torch.bfloat16, device=arena.device)
            table = torch.tensor(rows, dtype=torch.int64).to(arena.device)
            views = {k: stage[:, s:s + o * i_].view(C, o, i_) for k, (s, o, i_) in plan.items()}
            st = self._stage = {"key": key, "stage": stage, "table": table, "views": views, "n": len(rows),
                                "max_n": max(r[7] for r in rows)}
        if not (getattr(self, "_stage_fresh", False) and st.get("fresh_key") == key):
            self.be.ops.stage_conv_weights(arena, st["stage"], st["table"], st["n"], st["max_n"])
        return dict(st["views"])

    # ---- fused optimizer step + re-staging (plain SGD), as in BatchedResNet -------------------------------------
    # available (and tested), but not selected by the simulator: with only 1x1 layouts the separate SGD kernel (5.8 TB/s) plus
    # the vectorised staging pass are faster than the tile-scheduled fused kernel (39.6 vs 40.5 ms/round, 8 clients)
    supports_fused_sgd = False
    fused_lr = None
    fused_applied = False

    def set_fused_sgd(self, lr_dev):
        self.fused_lr = lr_dev

    def restage(self, W):
        self._stage_fresh = False
        if isinstance(self.be, NativeBackend) and self.be.dtype == torch.bfloat16 and getattr(W, "arena", None) is not None:
            self._prepare_staged(W)
            self._stage["fresh_key"] = self._stage["key"]
            self._stage_fresh = True

    def invalidate_stage(self):
        self._stage_fresh = False

    def _fused_plan(self, W):
        st = getattr(self, "_stage", None)
        if st is None:
            return None
        sparse = self.sparse_lr is not None
        fp = getattr(self, "_fused", None)
        if fp is not None and fp["key"] == (st["key"], sparse):
            return fp
        rows, staged = [], set()
        for r in st["table"].cpu().tolist():
            w_off, cout, cin, cinp, k, dst, dstT, n = r
            staged.add(w_off)
            rows.append([w_off, cout, cin, cinp, k, dst, dstT, n, 0, 0])
        for sg in W.layout.segments:
            if sg.kind != "param" or sg.offset in staged:
                continue
            if sparse and sg.name == self.sparse_sgd_segment:
                continue                                   # updated row by row inside the embedding backward
            rows.append([sg.offset, sg.numel, 1, 1, 1, -1, -1, sg.numel, 0, 0])
        fp = self._fused = {"key": (st["key"], sparse), "table": torch.tensor(rows, dtype=torch.int64).to(W.arena.device),
                            "n": len(rows), "max_n": min(sum(r[7] for r in rows), 2 ** 31 - 1)}
        return fp

    # ------------------------------------------------------------------ helpers
    def _heads(self, t, C, B, S, which=None):
        """[C, B*S, n*D] buffer -> [C*B, H, S, dh] strided view of block ``which`` (no copy)."""
        n = t.shape[-1] // self.D
        v = t.view(C * B, S, n, self.H, self.dh)
        return v[:, :, which or 0].permute(0, 2, 1, 3)

    def _step_counter(self, device):
        if self._step_dev is None or self._step_dev.device != device:
            self._step_dev = torch.zeros(1, dtype=torch.int32, device=device)
        return self._step_dev

    def _drop(self, x, p, site, training, x2=None):
        if not training or p <= 0:
            if x2 is None:
                return x
            return self.be.dropout(x, 0.0, self.seed, site, self._step_dev, x2)
        return self.be.dropout(x, p, self.seed, site, self._step_dev, x2)

    # ------------------------------------------------------------------ forward
    def _forward(self, ids, W, Wm, training, tape):
        be = self.be
        C, B, S = ids.shape
        if S > self.max_pos:
            raise ValueError(f"sequence length {S} exceeds max_position_embeddings {self.max_pos}")
        R, D, H = B * S, self.D, self.H
        ids2 = ids.reshape(C, R)
        mask = (ids2 != self.pad_id).to(torch.float32).view(C * B, S).contiguous()
        scale = 1.0 / math.sqrt(self.dh)
        pa = self.p_attn if training else 0.0
        e = be.embed_fwd(ids2, W["distilbert.embeddings.word_embeddings.weight"],
                         W["distilbert.embeddings.position_embeddings.weight"], S)
        h, z0, m0, r0 = be.ln_fwd(e, None, W["distilbert.embeddings.LayerNorm.weight"],
                                  W["distilbert.embeddings.LayerNorm.bias"], self.eps)
        h = self._drop(h, self.p_hidden, 0, training)
        if tape is not None:
            tape.append((ids2, mask, z0, m0, r0))
        for i in range(self.L):
            p = self._ln(i)
            qkv = be.mm("nt", h, Wm[f"qkv{i}"], bias=Wm[f"bqkv{i}"])                       # [C,R,3D]
            q, k, v = (self._heads(qkv, C, B, S, j) for j in range(3))                       # [CB,H,S,dh] views
            scores = be.mm("nt", q, k, out_dtype=torch.float32)                              # [CB,H,S,S]
            P, Pd = be.softmax_fwd(scores, mask, H, scale, pa, self.seed, 16 * (i + 1) + 1, self._step_dev)
            del scores
            ctx = torch.empty((C, R, D), dtype=h.dtype, device=h.device)
            be.mm("nn", Pd, v, out=self._heads(ctx, C, B, S))                                # heads merged in place
            ao = be.mm("nt", ctx, Wm[f"o{i}"], bias=W[p + "attention.out_lin.bias"])
            h1, z1, m1, r1 = be.ln_fwd(ao, h, W[p + "sa_layer_norm.weight"], W[p + "sa_layer_norm.bias"], self.eps)
            u = be.mm("nt", h1, Wm[f"f1{i}"], bias=W[p + "ffn.lin1.bias"])                   # [C,R,F]
            a = be.act_fwd(u, "gelu")
            f = be.mm("nt", a, Wm[f"f2{i}"], bias=W[p + "ffn.lin2.bias"])
            f = self._drop(f, self.p_hidden, 16 * (i + 1) + 2, training)
            h2, z2, m2, r2 = be.ln_fwd(f, h1, W[p + "output_layer_norm.weight"], W[p + "output_layer_norm.bias"],
                                       self.eps)
            if tape is not None:
                tape.append((h, qkv, P, Pd, ctx, h1, z1, m1, r1, u, a, z2, m2, r2))
            h = h2
        pooled = h.view(C, B, S, D)[:, :, 0]                                                 # [C,B,D] strided view
        pr = be.mm("nt", pooled, Wm["pre"], bias=W["pre_classifier.bias"], relu=True)
        prd = self._drop(pr, self.p_cls, 3, training)
        logits = be.mm("nt", prd, Wm["cls"], out_dtype=torch.float32, bias=W["classifier.bias"])   # [C,B,ncls]
        return logits, (pooled, pr, prd), h

    # ------------------------------------------------------------------ training step
    def train_step(self, x, y, W, G, active, row_scale=None):
        be = self.be
        ids = x
        C, B, S = ids.shape
        R, D = B * S, self.D
        step_dev = self._step_counter(ids.device)
        step_dev.add_(1)                                       # fresh dropout masks on every (graph-replayed) step
        Wm = self.prepare(W)
        tape: List[tuple] = []
        logits, (pooled, pr, prd), h_last = self._forward(ids, W, Wm, True, tape)
        loss, dlog = be.xent(logits, y, row_scale=row_scale)
        dt = h_last.dtype
        # ---- head
        npad = (self.ncls + 7) // 8 * 8
        dlog_c = torch.zeros((C, B, npad), dtype=dt, device=ids.device)[:, :, :self.ncls]    # 16-byte aligned rows
        dlog_c.copy_(dlog)
        be.mm("tn", dlog_c, prd, out=G["classifier.weight"])
        G["classifier.bias"].copy_(dlog.sum(dim=1))
        dprd = be.mm("nn", dlog_c, Wm["cls"])
        dpr = self._drop(dprd, self.p_cls, 3, True)
        dpc = be.act_bwd(dpr, pr, "relu")
        be.mm("tn", dpc, pooled, out=G["pre_classifier.weight"])
        be.colsum(dpc, G["pre_classifier.bias"])
        dH = torch.zeros((C, B, S, D), dtype=dt, device=ids.device)
        be.mm("nn", dpc, Wm["pre"], out=dH[:, :, 0])                                         # only the [CLS] rows
        d, d2 = dH.view(C, R, D), None
        scale = 1.0 / math.sqrt(self.dh)
        # ---- encoder layers
        for i in reversed(range(self.L)):
            p = self._ln(i)
            h, qkv, P, Pd, ctx, h1, z1, m1, r1, u, a, z2, m2, r2 = tape[i + 1]
            q, k, v = (self._heads(qkv, C, B, S, j) for j in range(3))
            dz2 = be.ln_bwd(d, d2, z2, m2, r2, W[p + "output_layer_norm.weight"], G[p + "output_layer_norm.weight"],
                            G[p + "output_layer_norm.bias"])
            df = self._drop(dz2, self.p_hidden, 16 * (i + 1) + 2, True)
            with be.side(df):                      # parameter gradients: off the data-gradient chain (parallel graph branch)
                be.mm("tn", df, a, out=G[p + "ffn.lin2.weight"])
                be.colsum(df, G[p + "ffn.lin2.bias"])
            da = be.mm("nn", df, Wm[f"f2{i}"])
            du = be.act_bwd(da, u, "gelu")
            with be.side(du):
                be.mm("tn", du, h1, out=G[p + "ffn.lin1.weight"])
                be.colsum(du, G[p + "ffn.lin1.bias"])
            dh1 = be.mm("nn", du, Wm[f"f1{i}"])
            dz1 = be.ln_bwd(dh1, dz2, z1, m1, r1, W[p + "sa_layer_norm.weight"], G[p + "sa_layer_norm.weight"],
                            G[p + "sa_layer_norm.bias"])
            with be.side(dz1):
                be.mm("tn", dz1, ctx, out=G[p + "attention.out_lin.weight"])
                be.colsum(dz1, G[p + "attention.out_lin.bias"])
            dctx = be.mm("nn", dz1, Wm[f"o{i}"])
            dctx_h = self._heads(dctx, C, B, S)
            dqkv = torch.empty_like(qkv)
            dq, dk, dv = (self._heads(dqkv, C, B, S, j) for j in range(3))
            be.mm("tn", Pd, dctx_h, out=dv)                                                  # dV = Pd^T dO
            dP = be.mm("nt", dctx_h, v, out_dtype=torch.float32)                             # dO V^T
            dS = be.softmax_bwd(dP, P, scale, self.p_attn, self.seed, 16 * (i + 1) + 1, self._step_dev)
            del dP
            be.mm("nn", dS, k, out=dq)                                                       # dQ = dS K
            be.mm("tn", dS, q, out=dk)                                                       # dK = dS^T Q
            with be.side(dqkv):
                for j, nm in enumerate(("q_lin", "k_lin", "v_lin")):
                    blk = dqkv[:, :, j * D:(j + 1) * D]
                    be.mm("tn", blk, h, out=G[p + f"attention.{nm}.weight"])
                    be.colsum(blk, G[p + f"attention.{nm}.bias"])
            d = be.mm("nn", dqkv, Wm[f"qkv{i}"])
            d2 = dz1
        # ---- embeddings
        ids2, _, z0, m0, r0 = tape[0]
        if self.p_hidden > 0:
            d, d2 = self._drop(d, self.p_hidden, 0, True, x2=d2), None
        de = be.ln_bwd(d, d2, z0, m0, r0, W["distilbert.embeddings.LayerNorm.weight"],
                       G["distilbert.embeddings.LayerNorm.weight"], G["distilbert.embeddings.LayerNorm.bias"])
        wname = "distilbert.embeddings.word_embeddings.weight"
        if self.sparse_lr is not None:
            # plain SGD: apply w[token] -= lr * dz row by row; the dense vocab x D gradient is never built and the
            # optimizer kernel skips the segment (``sparse_sgd_segment``)
            be.embed_bwd(ids2, de, W[wname], G["distilbert.embeddings.position_embeddings.weight"], S,
                         word_scale=-self.sparse_lr)
        else:
            G[wname].zero_()
            be.embed_bwd(ids2, de, G[wname], G["distilbert.embeddings.position_embeddings.weight"], S)
        be.join()                                  # parameter-gradient branch complete
        self.fused_applied = False
        if self.fused_lr is not None and isinstance(be, NativeBackend) and getattr(W, "arena", None) is not None:
            fp = self._fused_plan(W)
            if fp is not None:
                # optimizer step of every other segment + bf16 re-staging of the linear layers in one launch
                be.ops.sgd_restage(W.arena, G.arena, None, self._stage["stage"], fp["table"], fp["n"], fp["max_n"], self.fused_lr)
                self._stage["fresh_key"] = self._stage["key"]
                self._stage_fresh = True
                self.fused_applied = True
        return loss

    # ------------------------------------------------------------------ sparse embedding update
    sparse_sgd_segment = "distilbert.embeddings.word_embeddings.weight"

    def set_sparse_sgd(self, lr_dev):
        """``lr_dev`` [C] (device, per-client step size, 0 = frozen client) switches the word-embedding update to the
        fused sparse form; None restores dense gradients.  Only valid for plain SGD (no momentum / decay / prox)."""
        self.sparse_lr = lr_dev

    def logits(self, x, W):
        return self._forward(x, W, self.prepare(W), False, None)[0]
