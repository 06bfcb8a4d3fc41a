"""Python front-end of the K12 transformer kernels (``csrc/transformer.cu``) and of the generalised grouped
GEMM (NT / NN / TN operand forms, two-level groups) they are built around.

Tensors are CUDA bf16 activations ``[C, R, D]`` and fp32 rows of the stacked parameter / gradient arenas.
No eager fallback: unsupported layouts raise :class:`NativeError` and the caller (``NativeBackend``) decides.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from .native import FbGroupedGemm, NativeError, check, count_launch, lib, ptr, stream_ptr

f32, i64, i32, u32, u64 = C.c_float, C.c_int64, C.c_int32, C.c_uint32, C.c_uint64
vp = C.c_void_p

_MODES = {"nt": "fb_grouped_gemm_bf16", "tn": "fb_grouped_gemm_tn_bf16", "nn": "fb_grouped_gemm_nn_bf16"}
# fp32 operands run on the same kernel with tcgen05 kind::tf32 (the reference's cuDNN convolution precision)
_MODES_TF32 = {"nt": "fb_grouped_gemm_tf32", "tn": "fb_grouped_gemm_tn_tf32", "nn": "fb_grouped_gemm_nn_tf32"}
_SCRATCH = {}


def _scratch(device, n_floats: int) -> torch.Tensor:
    key = (device.type, device.index, torch.cuda.current_stream(device).cuda_stream)   # per stream: see ops._scratch_for
    buf = _SCRATCH.get(key)
    if buf is None or buf.numel() < n_floats:
        buf = _SCRATCH[key] = torch.empty(max(n_floats, 1 << 16), dtype=torch.float32, device=device)
    return buf


def _as4(t: torch.Tensor) -> torch.Tensor:
    """[G, r, c] -> [1, G, r, c]; [go, gi, r, c] unchanged (views, no copies)."""
    if t.dim() == 3:
        return t.unsqueeze(0)
    if t.dim() != 4:
        raise NativeError("mm: operands must be 3-D [G, rows, cols] or 4-D [outer, inner, rows, cols]")
    return t


def _operand_ok(t: torch.Tensor, es: int) -> bool:
    al = 16 // es
    if not (t.is_cuda and t.stride(3) == 1 and t.data_ptr() % 16 == 0 and t.stride(2) % al == 0):
        return False
    if t.shape[1] > 1 and t.stride(1) % al:
        return False
    if t.shape[0] > 1 and t.stride(0) % al:
        return False
    return True


def mm_supported(mode: str, a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None) -> bool:
    try:
        a, b = _as4(a), _as4(b)
    except NativeError:
        return False
    if a.dtype != b.dtype or a.dtype not in (torch.bfloat16, torch.float32) or a.shape[:2] != b.shape[:2]:
        return False
    es = a.element_size()
    if not (_operand_ok(a, es) and _operand_ok(b, es)):
        return False
    K = a.shape[3] if mode != "tn" else a.shape[2]
    Kb = b.shape[3] if mode == "nt" else b.shape[2]
    if K != Kb:
        return False
    if out is not None:
        o = _as4(out)
        if o.dtype not in (torch.bfloat16, torch.float32) or not _operand_ok(o, o.element_size()):
            return False
        if es == 4 and o.dtype != torch.float32:
            return False
    return True


def mm(mode: str, a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None, out_dtype=torch.bfloat16,
       bias: Optional[torch.Tensor] = None, relu: bool = False) -> torch.Tensor:
    """Grouped matrix product on the tcgen05 tensor cores.

    ``nt``: d = a @ b^T   (a [..,M,K], b [..,N,K])      forward projections, Q K^T, dO V^T
    ``nn``: d = a @ b     (a [..,M,K], b [..,K,N])      data gradients (dY W), P V, dS K
    ``tn``: d = a^T @ b   (a [..,K,M], b [..,K,N])      weight gradients, P^T dO, dS^T Q

    Leading dims are one or two group levels with arbitrary (16-byte aligned) strides, so attention heads are read
    and written in place inside ``[C*B, S, heads*dh]`` buffers.  ``out`` may be a strided view as well (e.g. a
    weight-gradient slot of the gradient arena); ``bias`` is fp32 ``[G, N]`` (one-level groups only)."""
    if not mm_supported(mode, a, b, out):
        raise NativeError(f"mm[{mode}]: unsupported operand layout")
    a4, b4 = _as4(a), _as4(b)
    go, gi = a4.shape[0], a4.shape[1]
    if mode == "nt":
        M, K, N = a4.shape[2], a4.shape[3], b4.shape[2]
    elif mode == "nn":
        M, K, N = a4.shape[2], a4.shape[3], b4.shape[3]
    else:
        K, M, N = a4.shape[2], a4.shape[3], b4.shape[3]
    if out is None:
        shape = tuple(a.shape[:-2]) + (M, N)
        out = torch.empty(shape, dtype=torch.float32 if a.dtype == torch.float32 else out_dtype, device=a.device)
    o4 = _as4(out)
    if tuple(o4.shape) != (go, gi, M, N):
        raise NativeError(f"mm[{mode}]: output shape {tuple(o4.shape)} != {(go, gi, M, N)}")
    p = FbGroupedGemm()
    p.a, p.b, p.d = a4.data_ptr(), b4.data_ptr(), o4.data_ptr()
    p.G, p.M, p.N, p.K = go * gi, M, N, K
    p.lda, p.ldb, p.ldd = a4.stride(2), b4.stride(2), o4.stride(2)
    p.a_gs, p.b_gs, p.d_gs = a4.stride(1), b4.stride(1), o4.stride(1)
    p.a_gs2, p.b_gs2, p.d_gs2 = a4.stride(0), b4.stride(0), o4.stride(0)
    p.group_inner = gi if go > 1 else 0
    if bias is not None:
        if go > 1 or bias.dtype != torch.float32 or bias.dim() != 2 or bias.stride(1) != 1:
            raise NativeError("mm: bias must be fp32 [G, N] with one-level groups")
        p.bias, p.bias_gs = bias.data_ptr(), bias.stride(0)
    p.d_fp32, p.relu = int(out.dtype == torch.float32), int(relu)
    table = _MODES_TF32 if a.dtype == torch.float32 else _MODES
    check(getattr(lib(), table[mode])(C.byref(p), vp(stream_ptr(a.device))), f"mm[{mode}]")
    count_launch()
    return out


# --------------------------------------------------------------------------- embeddings
def embed_gather(ids: torch.Tensor, word: torch.Tensor, pos: torch.Tensor, S: int) -> torch.Tensor:
    """ids int64 [C, R]; word [C, vocab, D] / pos [C, maxpos, D] fp32 arena views -> bf16 [C, R, D]."""
    Cn, R = ids.shape
    vocab, D = word.shape[1], word.shape[2]
    ld = word.stride(0) if Cn > 1 else 4
    if Cn > 1 and pos.stride(0) != ld:
        raise NativeError("embed_gather: word and position tables must share the arena row stride")
    out = torch.empty((Cn, R, D), dtype=torch.bfloat16, device=ids.device)
    check(lib().fb_embed_gather(vp(ptr(ids.contiguous())), vp(ptr(word)), vp(ptr(pos)), i64(ld), vp(ptr(out)), i32(Cn), i64(R),
                                i32(S), i32(D), i32(vocab), vp(stream_ptr(ids.device))), "embed_gather")
    count_launch()
    return out


def embed_bwd(ids: torch.Tensor, dz: torch.Tensor, gword: torch.Tensor, gpos: torch.Tensor, S: int,
              word_scale: Optional[torch.Tensor] = None):
    """Scatter-add ``word_scale[c] * dz`` into the word rows (pre-zeroed gradient rows, or the parameter rows with
    word_scale = -lr for the fused sparse SGD update); position gradients are written."""
    Cn, R = ids.shape
    vocab, D = gword.shape[1], gword.shape[2]
    ld = gword.stride(0) if Cn > 1 else 4
    if Cn > 1 and gpos.stride(0) != ld:
        raise NativeError("embed_bwd: word and position gradients must share the arena row stride")
    if gpos.shape[1] > S:
        gpos[:, S:].zero_()
    check(lib().fb_embed_bwd(vp(ptr(ids.contiguous())), vp(ptr(dz.contiguous())), vp(ptr(gword)), vp(ptr(gpos)), i64(ld),
                             i32(Cn), i64(R), i32(S), i32(D), i32(vocab), vp(ptr(word_scale)), vp(stream_ptr(ids.device))),
          "embed_bwd")
    count_launch(2)


# --------------------------------------------------------------------------- LayerNorm
def add_ln_fwd(a: torch.Tensor, b: Optional[torch.Tensor], gamma: torch.Tensor, beta: torch.Tensor, eps: float):
    """y = LN(a + b); returns (y, z, mean, rstd) with z = a + b (or ``a`` itself when b is None)."""
    Cn, R, D = a.shape
    a = a.contiguous()
    b = b.contiguous() if b is not None else None
    y = torch.empty_like(a)
    z = torch.empty_like(a) if b is not None else a
    mean = torch.empty((Cn, R), dtype=torch.float32, device=a.device)
    rstd = torch.empty_like(mean)
    ld_p = gamma.stride(0) if Cn > 1 else 4
    if Cn > 1 and beta.stride(0) != ld_p:
        raise NativeError("add_ln_fwd: gamma and beta must share the arena row stride")
    check(lib().fb_add_ln_fwd(vp(ptr(a)), vp(ptr(b)), vp(ptr(z) if b is not None else None), vp(ptr(y)), vp(ptr(mean)),
                              vp(ptr(rstd)), vp(ptr(gamma)), vp(ptr(beta)), i64(ld_p), i32(Cn), i64(R), i32(D), f32(eps),
                              vp(stream_ptr(a.device))), "add_ln_fwd")
    count_launch()
    return y, z, mean, rstd


def ln_bwd(dy: torch.Tensor, dy2: Optional[torch.Tensor], z: torch.Tensor, mean: torch.Tensor, rstd: torch.Tensor,
           gamma: torch.Tensor, dgamma: torch.Tensor, dbeta: torch.Tensor) -> torch.Tensor:
    Cn, R, D = z.shape
    dy = dy.contiguous()
    dy2 = dy2.contiguous() if dy2 is not None else None
    dz = torch.empty_like(z)
    fn = lib().fb_ln_bwd_scratch_floats
    fn.restype = C.c_int64
    scratch = _scratch(z.device, int(fn(i32(Cn), i64(R), i32(D))))
    ld_p = gamma.stride(0) if Cn > 1 else 4
    ld_g = dgamma.stride(0) if Cn > 1 else max(dgamma.stride(0), D)
    if Cn > 1 and dbeta.stride(0) != ld_g:
        raise NativeError("ln_bwd: dgamma and dbeta must share a row stride")
    check(lib().fb_ln_bwd(vp(ptr(dy)), vp(ptr(dy2)), vp(ptr(z)), vp(ptr(mean)), vp(ptr(rstd)), vp(ptr(gamma)), i64(ld_p),
                          vp(ptr(dz)), vp(ptr(dgamma)), vp(ptr(dbeta)), i64(ld_g), vp(ptr(scratch)), i32(Cn), i64(R), i32(D),
                          vp(stream_ptr(z.device))), "ln_bwd")
    count_launch(2)
    return dz


# --------------------------------------------------------------------------- attention softmax
def _key(seed: int):
    return u64(int(seed) & (2 ** 64 - 1))


def attn_softmax_fwd(scores: torch.Tensor, key_mask: Optional[torch.Tensor], H: int, scale: float, p: float = 0.0,
                     seed: int = 0, site: int = 0, step_dev: Optional[torch.Tensor] = None):
    """scores fp32 [CB, H, S, S] (contiguous) -> (P, Pdrop) bf16 of the same shape; Pdrop is P when p == 0."""
    S = scores.shape[-1]
    groups = scores.numel() // (S * S)
    P = torch.empty(scores.shape, dtype=torch.bfloat16, device=scores.device)
    Pd = torch.empty_like(P) if p > 0 else None
    check(lib().fb_attn_softmax_fwd(vp(ptr(scores)), vp(ptr(P)), vp(ptr(Pd)), vp(ptr(key_mask)), i64(groups), i32(S), i32(H),
                                    f32(scale), f32(p), _key(seed), u32(site), vp(ptr(step_dev)),
                                    vp(stream_ptr(scores.device))), "attn_softmax_fwd")
    count_launch()
    return P, (Pd if Pd is not None else P)


def attn_softmax_bwd(dP: torch.Tensor, P: torch.Tensor, scale: float, p: float = 0.0, seed: int = 0, site: int = 0,
                     step_dev: Optional[torch.Tensor] = None) -> torch.Tensor:
    S = P.shape[-1]
    groups = P.numel() // (S * S)
    dS = torch.empty_like(P)
    check(lib().fb_attn_softmax_bwd(vp(ptr(dP)), vp(ptr(P)), vp(ptr(dS)), i64(groups), i32(S), f32(scale), f32(p), _key(seed),
                                    u32(site), vp(ptr(step_dev)), vp(stream_ptr(P.device))), "attn_softmax_bwd")
    count_launch()
    return dS


# --------------------------------------------------------------------------- activations, dropout, bias gradients
_ACT = {"relu": 0, "gelu": 1}


def act_fwd(x: torch.Tensor, kind: str) -> torch.Tensor:
    x = x.contiguous()
    y = torch.empty_like(x)
    fn = lib().fb_act_fwd_f32 if x.dtype == torch.float32 else lib().fb_act_fwd
    check(fn(vp(ptr(x)), vp(ptr(y)), i64(x.numel()), i32(_ACT[kind]), vp(stream_ptr(x.device))), "act_fwd")
    count_launch()
    return y


def act_bwd(dy: torch.Tensor, x: torch.Tensor, kind: str) -> torch.Tensor:
    dy, x = dy.contiguous(), x.contiguous()
    dx = torch.empty_like(x)
    fn = lib().fb_act_bwd_f32 if x.dtype == torch.float32 else lib().fb_act_bwd
    check(fn(vp(ptr(dy)), vp(ptr(x)), vp(ptr(dx)), i64(x.numel()), i32(_ACT[kind]), vp(stream_ptr(x.device))), "act_bwd")
    count_launch()
    return dx


def dropout(x: torch.Tensor, p: float, seed: int, site: int, step_dev: Optional[torch.Tensor] = None,
            x2: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Inverted dropout of ``x (+ x2)``; the same call on a gradient is the backward pass."""
    x = x.contiguous()
    x2 = x2.contiguous() if x2 is not None else None
    y = torch.empty_like(x)
    fn = lib().fb_dropout_f32 if x.dtype == torch.float32 else lib().fb_dropout_bf16
    check(fn(vp(ptr(x)), vp(ptr(x2)), vp(ptr(y)), i64(x.numel()), f32(p), _key(seed), u32(site & 0xFFFFFFFF),
             vp(ptr(step_dev)), vp(stream_ptr(x.device))), "dropout")
    count_launch()
    return y


def colsum(x: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """out[c, :] = x[c].sum(0) for bf16 x [C, R, N] (rows may be strided, e.g. a column block of a wider buffer);
    ``out`` fp32 [C, N] (may be rows of the gradient arena)."""
    Cn, R, Nn = x.shape
    al = 16 // x.element_size()
    if x.stride(2) != 1 or x.stride(1) % al or (Cn > 1 and x.stride(0) % al) or x.data_ptr() % 16:
        x = x.contiguous()
    ld_o = out.stride(0) if Cn > 1 else max(out.stride(0), Nn)
    cs = x.stride(0) if Cn > 1 else R * x.stride(1)
    fn = lib().fb_colsum_f32 if x.dtype == torch.float32 else lib().fb_colsum_bf16
    check(fn(vp(ptr(x)), i32(Cn), i64(R), i32(Nn), i64(x.stride(1)), i64(cs), vp(ptr(out)), i64(ld_o), vp(stream_ptr(x.device))), "colsum")
    count_launch()
    return out
