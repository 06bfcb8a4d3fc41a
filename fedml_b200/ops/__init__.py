"""Python front-end of the hand-written sm_100a kernels (``csrc/*.cu``).

Every function takes CUDA tensors, validates shapes/dtypes, and launches on the current
stream.  There is no eager fallback here: CPU tensors must use the oracles in
``fedml_b200.core.aggregation`` — a CUDA tensor with a missing library raises.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from . import native as N
from .native import FB_OPT, FbRoundParams, NativeError, check, count_launch, lib, ptr, stream_ptr

f32, i64, i32, u32, u64 = C.c_float, C.c_int64, C.c_int32, C.c_uint32, C.c_uint64
vp = C.c_void_p


def _req(t: torch.Tensor, dtype, name: str):
    if not t.is_cuda:
        raise NativeError(f"{name}: expected a CUDA tensor")
    if t.dtype != dtype:
        raise NativeError(f"{name}: expected dtype {dtype}, got {t.dtype}")
    return t


def _rows(stack: torch.Tensor):
    """[C, ld-strided] 2-D view -> (ptr, ld, C)."""
    if stack.dim() != 2 or stack.stride(1) != 1:
        raise NativeError("stacked tensor must be 2-D with unit inner stride")
    return ptr(stack), stack.stride(0) if stack.shape[0] > 1 else max(stack.stride(0), stack.shape[1]), stack.shape[0]


# --------------------------------------------------------------------------- K4 / broadcast
def weighted_accumulate(clients: torch.Tensor, weights: torch.Tensor, out: torch.Tensor, scale: float = 1.0,
                        beta: float = 0.0):
    """out = beta*out + scale * sum_c weights[c] * clients[c]   (clients fp32 or bf16 [C, P])."""
    p, ld, Cn = _rows(clients)
    P = out.numel()
    _req(weights, torch.float32, "weights"); _req(out, torch.float32, "out")
    fn = lib().fb_weighted_accumulate if clients.dtype == torch.float32 else lib().fb_weighted_accumulate_bf16
    check(fn(vp(p), i64(ld), i32(Cn), vp(ptr(weights)), f32(scale), f32(beta), vp(ptr(out)), i64(P),
             vp(stream_ptr(out.device))), "weighted_accumulate")
    count_launch()
    return out


def broadcast_rows(src: torch.Tensor, rows: torch.Tensor):
    p, ld, Cn = _rows(rows)
    check(lib().fb_broadcast_rows(vp(ptr(_req(src, torch.float32, "src"))), vp(p), i64(ld), i32(Cn), i64(src.numel()),
                                  vp(stream_ptr(src.device))), "broadcast_rows")
    count_launch()
    return rows


# --------------------------------------------------------------------------- K3
def server_opt(w, avg, m, v, opt: str, lr, momentum=0.0, beta1=0.9, beta2=0.99, eps=1e-3, step=1, nesterov=False,
               grad_scale=1.0):
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    check(lib().fb_server_opt(vp(ptr(w)), vp(ptr(avg)), vp(ptr(m)), vp(ptr(v)), i64(w.numel()), i32(FB_OPT[opt.lower()]),
                              i32(int(nesterov)), f32(lr), f32(momentum), f32(beta1), f32(beta2), f32(eps), f32(bc1),
                              f32(bc2), f32(grad_scale), vp(stream_ptr(w.device))), "server_opt")
    count_launch()
    return w


# --------------------------------------------------------------------------- K2 pieces
def sqnorm_diff(clients: torch.Tensor, ref: torch.Tensor, n: Optional[int] = None) -> torch.Tensor:
    p, ld, Cn = _rows(clients)
    n = ref.numel() if n is None else n
    out = torch.empty(Cn, dtype=torch.float32, device=clients.device)
    check(lib().fb_sqnorm_diff(vp(p), i64(ld), i32(Cn), vp(ptr(ref)), i64(n), vp(ptr(out)),
                               vp(stream_ptr(out.device))), "sqnorm_diff")
    count_launch(2)
    return out


def philox_normal(n: int, seed: int, stream_id: int, substream: int, device) -> torch.Tensor:
    out = torch.empty(n, dtype=torch.float32, device=device)
    check(lib().fb_philox_normal(vp(ptr(out)), i64(n), u64(seed), u32(stream_id), u32(substream),
                                 vp(stream_ptr(device))), "philox_normal")
    count_launch()
    return out


# --------------------------------------------------------------------------- K1 fused round
def round_grid(P: int, G: int, emulated: int = 1) -> int:
    return int(lib().fb_round_grid(i64(P), i32(G), i32(emulated)))


def round_flag_bytes(G: int) -> int:
    return int(lib().fb_round_flag_bytes(i32(G)))


def round_launch(params: FbRoundParams, device=None):
    check(lib().fb_round_launch(C.byref(params), vp(stream_ptr(device))), "round_launch")
    count_launch()


def round_launch_emulated(params: Sequence[FbRoundParams], P: int, device):
    """All ranks in one cooperative launch on one GPU (tests; see B200_PROFILING "kernels that wait
    on one another must not be separate launches on one GPU")."""
    G = len(params)
    arr = (FbRoundParams * G)(*params)
    nbytes = C.sizeof(arr)
    host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
    dev = host.to(device)
    check(lib().fb_round_launch_emulated(vp(ptr(dev)), i32(G), i64(P), vp(stream_ptr(device))),
          "round_launch_emulated")
    count_launch()
    torch.cuda.current_stream(device).synchronize()   # keep `dev` alive until the kernel has read it
    return nbytes


def peer_barrier(peer_flags_dev: torch.Tensor, rank: int, G: int, epoch: int, slot: int = 0):
    check(lib().fb_peer_barrier(vp(ptr(peer_flags_dev)), i32(rank), i32(G), u32(epoch), i32(slot),
                                vp(stream_ptr(peer_flags_dev.device))), "peer_barrier")
    count_launch()


# --------------------------------------------------------------------------- K5 / K6
def ptr_table(ptrs: Sequence[int], device) -> torch.Tensor:
    return torch.tensor(list(ptrs), dtype=torch.int64).to(device)


def robust_select(rows_table: torch.Tensor, K: int, j0: int, j1: int, mode: str, trim: int,
                  outs_table: torch.Tensor, n_out: int):
    m = {"median": 0, "trimmed_mean": 1}[mode]
    check(lib().fb_robust_select(vp(ptr(rows_table)), i32(K), i64(j0), i64(j1), i32(m), i32(trim),
                                 vp(ptr(outs_table)), i32(n_out), vp(stream_ptr(rows_table.device))), "robust_select")
    count_launch()


def robust_sort_runs(rows_table: torch.Tensor, K: int, run_len: int, j0: int, j1: int):
    """Sort every run of ``run_len`` consecutive rows in place, per coordinate (step 1 of the many-client path)."""
    check(lib().fb_robust_sort_runs(vp(ptr(rows_table)), i32(K), i32(run_len), i64(j0), i64(j1),
                                    vp(stream_ptr(rows_table.device))), "robust_sort_runs")
    count_launch()


def robust_select_runs(rows_table: torch.Tensor, K: int, K_valid: int, run_len: int, j0: int, j1: int, mode: str, trim: int,
                       outs_table: torch.Tensor, n_out: int):
    m = {"median": 0, "trimmed_mean": 1}[mode]
    check(lib().fb_robust_select_runs(vp(ptr(rows_table)), i32(K), i32(K_valid), i32(run_len), i64(j0), i64(j1), i32(m), i32(trim),
                                      vp(ptr(outs_table)), i32(n_out), vp(stream_ptr(rows_table.device))), "robust_select_runs")
    count_launch()


def run_length_for(rows_per_rank: int) -> int:
    """Largest run length <= 128 that divides the rows a rank holds (runs must not straddle ranks); 0: use the plain path."""
    # runs of at most 64 rows first: the 64-input sorting network stays in registers, the 128-input one spills
    for L in list(range(min(64, rows_per_rank), 15, -1)) + list(range(min(128, rows_per_rank), 64, -1)):
        if rows_per_rank % L == 0:
            return L
    return 0


def coordinate_median(stack: torch.Tensor, out: Optional[torch.Tensor] = None, inplace: bool = False) -> torch.Tensor:
    """Lower median over dim 0 of a [K, P] fp32 stack (rows may be strided).  K > 128 takes the sorted-run path, which
    permutes the values inside each run of rows: on a copy unless ``inplace``."""
    return _select_local(stack, "median", 0, out, inplace)


def trimmed_mean(stack: torch.Tensor, trim: int, out: Optional[torch.Tensor] = None, inplace: bool = False) -> torch.Tensor:
    return _select_local(stack, "trimmed_mean", trim, out, inplace)


def _select_local(stack, mode, trim, out, inplace=False):
    _req(stack, torch.float32, "stack")
    K, P = stack.shape
    out = torch.empty(P, dtype=torch.float32, device=stack.device) if out is None else out
    outs = ptr_table([out.data_ptr()], stack.device)
    L = run_length_for(K) if K > 128 else 0
    if L:
        work = stack if inplace else stack.clone()        # the sorted-run path reorders values inside each run
        rows = ptr_table([work.data_ptr() + k * work.stride(0) * 4 for k in range(K)], stack.device)
        robust_sort_runs(rows, K, L, 0, P)
        robust_select_runs(rows, K, K, L, 0, P, mode, trim, outs, 1)
        return out
    rows = ptr_table([stack.data_ptr() + k * stack.stride(0) * 4 for k in range(K)], stack.device)
    robust_select(rows, K, 0, P, mode, trim, outs, 1)
    return out


# --------------------------------------------------------------------------- K7
def krum_pack(rows_table: torch.Tensor, K: int, ref: Optional[torch.Tensor], j0: int, j1: int, ld: int,
              split: bool = True):
    dev = rows_table.device
    hi = torch.empty((K, ld), dtype=torch.bfloat16, device=dev)
    lo = torch.empty((K, ld), dtype=torch.bfloat16, device=dev) if split else None
    check(lib().fb_krum_pack(vp(ptr(rows_table)), i32(K), vp(ptr(ref)), i64(j0), i64(j1), vp(ptr(hi)), vp(ptr(lo)),
                             i64(ld), vp(stream_ptr(dev))), "krum_pack")
    count_launch()
    return hi, lo


def krum_scores(gram: torch.Tensor, f: int) -> torch.Tensor:
    K = gram.shape[0]
    scores = torch.empty(K, dtype=torch.float32, device=gram.device)
    check(lib().fb_krum_scores(vp(ptr(_req(gram.contiguous(), torch.float32, "gram"))), i32(K), i32(f),
                               vp(ptr(scores)), vp(stream_ptr(gram.device))), "krum_scores")
    count_launch()
    return scores


# --------------------------------------------------------------------------- K8
def ff_quantize(x: torch.Tensor, scale: float, p: int) -> torch.Tensor:
    q = torch.empty(x.shape, dtype=torch.int32, device=x.device)
    check(lib().fb_ff_quantize(vp(ptr(_req(x.contiguous(), torch.float32, "x"))), i64(x.numel()), f32(scale), u32(p),
                               vp(ptr(q)), vp(stream_ptr(x.device))), "ff_quantize")
    count_launch()
    return q


def ff_dequantize(q: torch.Tensor, scale: float, p: int) -> torch.Tensor:
    x = torch.empty(q.shape, dtype=torch.float32, device=q.device)
    check(lib().fb_ff_dequantize(vp(ptr(q)), i64(q.numel()), f32(1.0 / scale), u32(p), vp(ptr(x)),
                                 vp(stream_ptr(q.device))), "ff_dequantize")
    count_launch()
    return x


def ff_matmul(U: torch.Tensor, X: torch.Tensor, p: int) -> torch.Tensor:
    """(U @ X) mod p for int32 field elements; U [n_out, n_in], X [n_in, d]."""
    n_out, n_in = U.shape
    d = X.shape[1]
    out = torch.empty((n_out, d), dtype=torch.int32, device=X.device)
    check(lib().fb_ff_matmul(vp(ptr(U.contiguous())), i32(n_out), i32(n_in), vp(ptr(X)), i64(d), i64(X.stride(0)), u32(p),
                             vp(ptr(out)), i64(d), vp(stream_ptr(X.device))), "ff_matmul")
    count_launch()
    return out


def ff_mask(x: torch.Tensor, seeds: Sequence[int], signs: Sequence[int], p: int, nonce: int = 0) -> torch.Tensor:
    dev = x.device
    y = torch.empty_like(x)
    s = torch.tensor([int(v) - (1 << 64) if int(v) >= (1 << 63) else int(v) for v in seeds], dtype=torch.int64).to(dev)
    g = torch.tensor(list(signs), dtype=torch.int32).to(dev)
    check(lib().fb_ff_mask(vp(ptr(x)), i64(x.numel()), vp(ptr(s)), vp(ptr(g)), i32(len(seeds)), u32(p), u32(int(nonce) & 0xFFFFFFFF), vp(ptr(y)),
                           vp(stream_ptr(dev))), "ff_mask")
    count_launch()
    return y


def ff_add_(acc: torch.Tensor, x: torch.Tensor, p: int) -> torch.Tensor:
    check(lib().fb_ff_add(vp(ptr(acc)), vp(ptr(x)), i64(acc.numel()), u32(p), vp(stream_ptr(acc.device))), "ff_add")
    count_launch()
    return acc


# --------------------------------------------------------------------------- K11 / K15
def sgd_step_stacked(w: torch.Tensor, g: torch.Tensor, n: int, lr: torch.Tensor, weight_decay=0.0, momentum=0.0,
                     dampening=0.0, nesterov=False, mom_buf=None, first_step=False, mu=0.0, anchor=None,
                     cum_grad=None):
    p, ld, Cn = _rows(w)
    if g.dtype == torch.bfloat16:
        check(lib().fb_sgd_step_stacked_bf16g(vp(p), vp(ptr(g)), i64(ld), i64(g.stride(0)), i32(Cn), i64(n),
                                              vp(ptr(lr)), f32(weight_decay), f32(mu), vp(ptr(anchor)),
                                              vp(stream_ptr(w.device))), "sgd_step_stacked_bf16g")
    else:
        if g.stride(0) != w.stride(0) and Cn > 1:
            raise NativeError("sgd_step_stacked: w and g must share the row stride")
        check(lib().fb_sgd_step_stacked(vp(p), vp(ptr(g)), i64(ld), i32(Cn), i64(n), vp(ptr(lr)), f32(weight_decay),
                                        f32(momentum), f32(dampening), i32(int(nesterov)), vp(ptr(mom_buf)),
                                        i32(int(first_step)), f32(mu), vp(ptr(anchor)), vp(ptr(cum_grad)),
                                        vp(stream_ptr(w.device))), "sgd_step_stacked")
    count_launch()
    return w


def softmax_xent(logits: torch.Tensor, labels: torch.Tensor, ignore_index=-100, pre_sigmoid=False,
                 row_scale: Optional[torch.Tensor] = None, rows_per_group: int = 1, want_loss=True, want_grad=True):
    """logits [R, ncls] fp32 contiguous; returns (loss_rows or None, dlogits or None)."""
    _req(logits, torch.float32, "logits"); _req(labels, torch.int64, "labels")
    R, ncls = logits.shape
    loss = torch.empty(R, dtype=torch.float32, device=logits.device) if want_loss else None
    grad = torch.empty_like(logits) if want_grad else None
    check(lib().fb_softmax_xent(vp(ptr(logits)), vp(ptr(labels)), i64(R), i32(ncls), i64(ignore_index),
                                i32(int(pre_sigmoid)), vp(ptr(row_scale)), i64(rows_per_group), vp(ptr(loss)),
                                vp(ptr(grad)), vp(stream_ptr(logits.device))), "softmax_xent")
    count_launch()
    return loss, grad


def augment_cifar(x: torch.Tensor, seed: int, nonce, pad: int = 4, cutout: int = 16, mean_u8=(125, 123, 114),
                  first_index: int = 0) -> torch.Tensor:
    """uint8 [N, H, W, C] -> randomly cropped / flipped / cut-out copy (csrc/augment.cu).  ``nonce``: int or a device
    int64 tensor of one element (read by the kernel, so CUDA-graph replays differ)."""
    _req(x, torch.uint8, "x")
    if x.dim() != 4 or not x.is_contiguous() or x.shape[3] > 4:
        raise NativeError("augment_cifar: expected contiguous uint8 [N, H, W, C<=4]")
    N, H, W, Cc = x.shape
    y = torch.empty_like(x)
    fill = 0
    for c in range(Cc):
        fill |= (int(mean_u8[c]) & 0xFF if c < len(mean_u8) else 0) << (8 * c)
    dev_nonce = nonce if torch.is_tensor(nonce) else None
    if dev_nonce is not None and (dev_nonce.dtype != torch.int64 or not dev_nonce.is_cuda):
        raise NativeError("augment_cifar: a tensor nonce must be a CUDA int64 tensor")
    check(lib().fb_augment_cifar(vp(ptr(x)), vp(ptr(y)), i64(N), i32(H), i32(W), i32(Cc), u64(int(seed) & (2 ** 64 - 1)),
                                 u32(0 if dev_nonce is not None else int(nonce) & 0xFFFFFFFF), vp(ptr(dev_nonce)), i32(pad),
                                 i32(cutout), u32(fill), i64(first_index), vp(stream_ptr(x.device))), "augment_cifar")
    count_launch()
    return y


def head_fwd_bwd(h, fc_w, fc_b, labels, row_scale, g_w, g_b):
    """Fused classifier head (csrc/train.cu head_fwd_bwd_kernel).  h [C,B,H,W,ch] bf16/fp32 contiguous; fc_w [C,ncls,ch],
    fc_b [C,ncls], g_w, g_b: fp32 views into the stacked arenas; labels int64 [C,B] (-1 = padding); row_scale fp32 [C] or
    None.  Returns (loss [C], dh like h)."""
    Cn, B = h.shape[0], h.shape[1]
    ch = h.shape[-1]
    HW = h.numel() // (Cn * B * ch)
    ncls = fc_w.shape[1]
    for t, nm in ((fc_w, "fc_w"), (g_w, "g_w")):
        if t.dtype != torch.float32 or t.stride(2) != 1 or t.stride(1) != ch:
            raise NativeError(f"head_fwd_bwd: {nm} must be fp32 [C, ncls, ch] with contiguous [ncls, ch] blocks")
    for t, nm in ((fc_b, "fc_b"), (g_b, "g_b")):
        if t.dtype != torch.float32 or t.stride(1) != 1:
            raise NativeError(f"head_fwd_bwd: {nm} must be fp32 [C, ncls] with unit inner stride")
    if h.dtype not in (torch.float32, torch.bfloat16) or not h.is_contiguous():
        raise NativeError("head_fwd_bwd: h must be contiguous bf16 / fp32")
    _req(labels, torch.int64, "labels")
    labels = labels.contiguous()
    loss = torch.empty(Cn, dtype=torch.float32, device=h.device)
    dh = torch.empty_like(h)
    ld = lambda t: t.stride(0) if Cn > 1 else max(t.stride(0), t[0].numel())   # noqa: E731
    check(lib().fb_head_fwd_bwd(vp(ptr(h)), i32(int(h.dtype == torch.float32)), i32(Cn), i32(B), i32(HW), i32(ch), vp(ptr(fc_w)),
                                i64(ld(fc_w)), vp(ptr(fc_b)), i64(ld(fc_b)), i32(ncls), vp(ptr(labels)), vp(ptr(row_scale)),
                                vp(ptr(loss)), vp(ptr(dh)), vp(ptr(g_w)), i64(ld(g_w)), vp(ptr(g_b)), i64(ld(g_b)),
                                vp(stream_ptr(h.device))), "head_fwd_bwd")
    count_launch()
    return loss, dh


def eval_counts(logits: torch.Tensor, labels: torch.Tensor, counts: torch.Tensor, loss_sum: torch.Tensor,
                pre_sigmoid=False):
    """Accumulates [correct, total, tp.., fp.., fn..] (int32 view of uint32) and the loss sum."""
    R, ncls = logits.shape
    check(lib().fb_eval_counts(vp(ptr(_req(logits, torch.float32, "logits"))), vp(ptr(labels)), i64(R), i32(ncls),
                               i32(int(pre_sigmoid)), vp(ptr(counts)), vp(ptr(loss_sum)),
                               vp(stream_ptr(logits.device))), "eval_counts")
    count_launch()


# --------------------------------------------------------------------------- K10 helpers (client-stacked conv nets)
_bn_scratch = {}


import os as _os

_BN_FUSED_ENV = _os.environ.get("FB_BN_FUSED")


def _bn_fused(x) -> bool:
    """Mirror of bn_fused_enabled() in csrc/conv.cu (stats + apply in one launch unless FB_BN_FUSED=0)."""
    if _BN_FUSED_ENV is not None:
        return _BN_FUSED_ENV != "0"
    return True


def _scratch_for(device, n_floats):
    """Partial-sum scratch, one buffer per (device, stream): client groups that train concurrently on several
    streams (fedml_b200/simulation/nccl) must not share it."""
    key = (str(device), torch.cuda.current_stream(device).cuda_stream)
    buf = _bn_scratch.get(key)
    if buf is None or buf.numel() < n_floats:
        buf = _bn_scratch[key] = torch.empty(max(n_floats, 1 << 16), dtype=torch.float32, device=device)
    return buf


def bn_fwd_stacked(x, res, gamma, beta, rm, rv, nbt, active, momentum, eps, relu, training, stats=None):
    """x [C,B,H,W,ch] bf16 / fp32 contiguous -> (y, save_mean [C,ch], save_rstd [C,ch]); running stats updated in place.
    ``stats`` fp32 [C, 2, ch] contiguous: per-client sums of x and x^2 already produced by the epilogue of the convolution
    that wrote x (``grouped_gemm(..., stats=)`` / ``conv3x3(..., stats=)``): the statistics pass over x is skipped."""
    Cn, ch = x.shape[0], x.shape[-1]
    M = x.numel() // (Cn * ch)
    x = x.contiguous()
    y = torch.empty_like(x)
    mean = torch.empty((Cn, ch), dtype=torch.float32, device=x.device)
    rstd = torch.empty((Cn, ch), dtype=torch.float32, device=x.device)
    use_stats = training and stats is not None
    if use_stats and not (stats.dtype == torch.float32 and stats.is_contiguous() and stats.numel() == Cn * 2 * ch):
        raise NativeError("bn_fwd_stacked: stats must be contiguous fp32 [C, 2, ch]")
    scratch = stats if use_stats else _scratch_for(x.device, Cn * 128 * 2 * ch)
    fn = lib().fb_bn_fwd_stacked_f32 if x.dtype == torch.float32 else lib().fb_bn_fwd_stacked
    check(fn(vp(ptr(x)), vp(ptr(res.contiguous()) if res is not None else 0), vp(ptr(y)), i32(Cn),
                                  i64(M), i32(ch), vp(ptr(gamma)), vp(ptr(beta)), i64(gamma.stride(0)), vp(ptr(rm)),
                                  vp(ptr(rv)), i64(rm.stride(0)), vp(ptr(active)), f32(momentum), f32(eps),
                                  i32(int(relu)), i32(2 if use_stats else int(training)), vp(ptr(mean)), vp(ptr(rstd)), vp(ptr(scratch)),
                                  vp(ptr(nbt)), i64(nbt.stride(0) if nbt is not None and nbt.dim() else 1),
                                  vp(stream_ptr(x.device))), "bn_fwd_stacked")
    count_launch(1 if (not training or use_stats or _bn_fused(x)) else 2)
    return y, mean, rstd


def bn_bwd_stacked(dy, x, y, gamma, mean, rstd, relu, want_dres, beta=None, dgamma_out=None, dbeta_out=None, dy2=None):
    """``y=None`` (+ ``beta``) recomputes the ReLU mask from x — only valid when the forward had no residual.
    ``dgamma_out`` / ``dbeta_out``: fp32 [C, ch] strided views (e.g. rows of the gradient arena) written in place."""
    Cn, ch = x.shape[0], x.shape[-1]
    M = x.numel() // (Cn * ch)
    dy = dy.contiguous()
    dy2 = dy2.contiguous() if dy2 is not None else None
    dx = torch.empty_like(x)
    dres = torch.empty_like(x) if want_dres else None
    direct = dgamma_out is not None and dbeta_out is not None and dgamma_out.stride(0) == dbeta_out.stride(0)
    dgamma = dgamma_out if direct else torch.empty((Cn, ch), dtype=torch.float32, device=x.device)
    dbeta = dbeta_out if direct else torch.empty((Cn, ch), dtype=torch.float32, device=x.device)
    ld_g = dgamma.stride(0) if Cn > 1 else max(dgamma.stride(0), ch)
    scratch = _scratch_for(x.device, Cn * 128 * 2 * ch)
    fn = lib().fb_bn_bwd_stacked_f32 if x.dtype == torch.float32 else lib().fb_bn_bwd_stacked
    check(fn(vp(ptr(dy)), vp(ptr(dy2)), vp(ptr(x)), vp(ptr(y)), vp(ptr(dx)), vp(ptr(dres)), i32(Cn), i64(M),
                                  i32(ch), vp(ptr(gamma)), vp(ptr(beta)), i64(gamma.stride(0)), vp(ptr(mean)),
                                  vp(ptr(rstd)), i32(int(relu)), vp(ptr(dgamma)), vp(ptr(dbeta)), i64(ld_g), vp(ptr(scratch)),
                                  vp(stream_ptr(x.device))), "bn_bwd_stacked")
    count_launch(1 if _bn_fused(x) else 2)
    return dx, dres, dgamma, dbeta


def gn_fwd_stacked(x, res, gamma, beta, cpg, eps, relu):
    """GroupNorm(+residual, +ReLU): x [C,B,H,W,ch] bf16 / fp32, gamma / beta fp32 [C, ch] arena views, ``cpg`` channels per
    group -> (y, save_mean [C*B, ch], save_rstd [C*B, ch])."""
    Cn, B, ch = x.shape[0], x.shape[1], x.shape[-1]
    rows = Cn * B
    M = x.numel() // (rows * ch)
    x = x.contiguous()
    y = torch.empty_like(x)
    mean = torch.empty((rows, ch), dtype=torch.float32, device=x.device)
    rstd = torch.empty((rows, ch), dtype=torch.float32, device=x.device)
    scratch = _scratch_for(x.device, rows * 128 * 2 * ch)
    check(lib().fb_gn_fwd_stacked(vp(ptr(x)), vp(ptr(res.contiguous()) if res is not None else 0), vp(ptr(y)),
                                  i32(int(x.dtype == torch.float32)), i32(rows), i64(M), i32(ch), vp(ptr(gamma)), vp(ptr(beta)),
                                  i64(gamma.stride(0)), i32(B), i32(cpg), f32(eps), i32(int(relu)), vp(ptr(mean)), vp(ptr(rstd)),
                                  vp(ptr(scratch)), vp(stream_ptr(x.device))), "gn_fwd_stacked")
    count_launch(2)
    return y, mean, rstd


def gn_bwd_stacked(dy, x, y, gamma, mean, rstd, cpg, relu, want_dres, beta=None, dgamma_out=None, dbeta_out=None, dy2=None):
    """Backward of :func:`gn_fwd_stacked`; dgamma / dbeta [C, ch] are the per-row results summed over a client's samples
    (written into ``dgamma_out`` / ``dbeta_out`` when given)."""
    Cn, B, ch = x.shape[0], x.shape[1], x.shape[-1]
    rows = Cn * B
    M = x.numel() // (rows * ch)
    dy = dy.contiguous()
    dy2 = dy2.contiguous() if dy2 is not None else None
    dx = torch.empty_like(x)
    dres = torch.empty_like(x) if want_dres else None
    dg_rows = torch.empty((rows, ch), dtype=torch.float32, device=x.device)
    db_rows = torch.empty((rows, ch), dtype=torch.float32, device=x.device)
    scratch = _scratch_for(x.device, rows * 128 * 2 * ch)
    check(lib().fb_gn_bwd_stacked(vp(ptr(dy)), vp(ptr(dy2)), vp(ptr(x)), vp(ptr(y)), vp(ptr(dx)), vp(ptr(dres)),
                                  i32(int(x.dtype == torch.float32)), i32(rows), i64(M), i32(ch), vp(ptr(gamma)), vp(ptr(beta)),
                                  i64(gamma.stride(0)), i32(B), i32(cpg), vp(ptr(mean)), vp(ptr(rstd)), i32(int(relu)),
                                  vp(ptr(dg_rows)), vp(ptr(db_rows)), vp(ptr(scratch)), vp(stream_ptr(x.device))), "gn_bwd_stacked")
    count_launch(2)
    dg, db = dg_rows.view(Cn, B, ch).sum(dim=1), db_rows.view(Cn, B, ch).sum(dim=1)
    if dgamma_out is not None:
        dgamma_out.copy_(dg); dbeta_out.copy_(db)
    return dx, dres, dg, db


def im2col3x3(x, stride):
    """x [C,B,H,W,ch] bf16 (ch % 8 == 0) -> cols [C, B*Ho*Wo, 9*ch]"""
    Cn, B, H, W, ch = x.shape
    if ch % 8:
        raise NativeError("im2col3x3: ch must be a multiple of 8")
    Ho, Wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    x = x.contiguous()
    cols = torch.empty((Cn, B * Ho * Wo, 9 * ch), dtype=x.dtype, device=x.device)
    fn = lib().fb_im2col3x3_nhwc_f32 if x.dtype == torch.float32 else lib().fb_im2col3x3_nhwc_bf16
    check(fn(vp(ptr(x)), vp(ptr(cols)), i32(Cn * B), i32(H), i32(W), i32(ch), i32(stride),
                                       vp(stream_ptr(x.device))), "im2col3x3")
    count_launch()
    return cols


def col2im3x3(dcols, B, H, W, ch, stride):
    Cn = dcols.shape[0]
    dcols = dcols.contiguous()
    dx = torch.empty((Cn, B, H, W, ch), dtype=dcols.dtype, device=dcols.device)
    fn = lib().fb_col2im3x3_nhwc_f32 if dcols.dtype == torch.float32 else lib().fb_col2im3x3_nhwc_bf16
    check(fn(vp(ptr(dcols)), vp(ptr(dx)), i32(Cn * B), i32(H), i32(W), i32(ch), i32(stride),
                                       vp(stream_ptr(dcols.device))), "col2im3x3")
    count_launch()
    return dx


def maxpool2x2_fwd(x, off=(0, 0), out_hw=None):
    """x [C,B,H,W,ch] bf16 / fp32 -> [C,B,Ho,Wo,ch]; windows start at ``off`` (see fb_maxpool2x2_fwd)."""
    Cn, B, H, W, ch = x.shape
    Ho, Wo = out_hw or ((H - off[0]) // 2, (W - off[1]) // 2)
    x = x.contiguous()
    y = torch.empty((Cn, B, Ho, Wo, ch), dtype=x.dtype, device=x.device)
    check(lib().fb_maxpool2x2_fwd(vp(ptr(x)), vp(ptr(y)), i32(int(x.dtype == torch.float32)), i32(Cn * B), i32(H), i32(W), i32(ch),
                                  i32(off[0]), i32(off[1]), i32(Ho), i32(Wo), vp(stream_ptr(x.device))), "maxpool2x2_fwd")
    count_launch()
    return y


def maxpool2x2_bwd(dy, x, off=(0, 0), relu_mask=False):
    """Gradient of :func:`maxpool2x2_fwd` w.r.t. the whole map x (zero outside the pooled windows)."""
    Cn, B, H, W, ch = x.shape
    Ho, Wo = dy.shape[2], dy.shape[3]
    dy, x = dy.contiguous(), x.contiguous()
    dx = torch.empty_like(x)
    check(lib().fb_maxpool2x2_bwd(vp(ptr(dy)), vp(ptr(x)), vp(ptr(dx)), i32(int(x.dtype == torch.float32)), i32(Cn * B), i32(H), i32(W),
                                  i32(ch), i32(off[0]), i32(off[1]), i32(Ho), i32(Wo), i32(int(relu_mask)), vp(stream_ptr(x.device))),
          "maxpool2x2_bwd")
    count_launch()
    return dx


# --------------------------------------------------------------------------- K9 grouped GEMM (tcgen05 / TMEM / TMA)
from .native import FbGroupedGemm  # noqa: E402


def _al(t):
    """elements per 16 bytes"""
    return 16 // t.element_size()


def _gemm_operand_ok(t):
    return (t.is_cuda and t.dtype in (torch.bfloat16, torch.float32) and t.dim() == 3 and t.stride(2) == 1
            and t.stride(1) % _al(t) == 0 and (t.shape[0] == 1 or t.stride(0) % _al(t) == 0) and t.data_ptr() % 16 == 0)


def grouped_gemm_supported(a, b) -> bool:
    """NT form: a [G,M,K], b [G,N,K] bf16 (or fp32 -> TF32 tensor cores), K-contiguous, 16-byte aligned rows."""
    return _gemm_operand_ok(a) and _gemm_operand_ok(b) and a.dtype == b.dtype and a.shape[2] == b.shape[2] \
        and a.shape[2] % _al(a) == 0 and a.shape[0] == b.shape[0]


def grouped_gemm_tn_supported(a, b) -> bool:
    """TN form: a [G,K,M], b [G,K,N] bf16 / fp32, M/N-contiguous, 16-byte multiples."""
    return _gemm_operand_ok(a) and _gemm_operand_ok(b) and a.dtype == b.dtype and a.shape[1] == b.shape[1] \
        and a.shape[0] == b.shape[0] and a.shape[2] % _al(a) == 0 and b.shape[2] % _al(a) == 0


def _launch_gemm(fn, a, b, G, M, N, K, out_dtype, relu, bias, name, out=None, accumulate=False, group_div=1, stats=None):
    d = torch.empty((G, M, N), dtype=out_dtype, device=a.device) if out is None else out
    p = FbGroupedGemm()
    p.a, p.b, p.d = a.data_ptr(), b.data_ptr(), d.data_ptr()
    p.bias = bias.data_ptr() if bias is not None else None
    p.a_gs, p.b_gs, p.d_gs = a.stride(0), b.stride(0), (M * N if out is None else (d.stride(0) if d.shape[0] > 1 else M * N))
    p.bias_gs = bias.stride(0) if bias is not None else 0
    p.G, p.M, p.N, p.K = G, M, N, K
    p.lda, p.ldb, p.ldd = a.stride(1), b.stride(1), N
    p.d_fp32, p.relu, p.accumulate, p.out_group_div = int(out_dtype == torch.float32), int(relu), int(accumulate), group_div
    if stats is not None:               # fp32 [G, 2, N] (zeroed by the caller): column sums / sums of squares of d
        if stats.dtype != torch.float32 or stats.shape != (G, 2, N) or stats.stride(2) != 1 or stats.stride(1) != N:
            raise NativeError(f"{name}: stats must be fp32 [G, 2, N] with contiguous [2, N] blocks")
        p.stats, p.stats_gs = stats.data_ptr(), stats.stride(0)
    check(fn(C.byref(p), vp(stream_ptr(a.device))), name)
    count_launch()
    return d


def grouped_gemm(a, b, out_dtype=torch.bfloat16, relu=False, bias=None, stats=None):
    """d[g] = a[g] @ b[g]^T   (a [G,M,K], b [G,N,K]) on the tcgen05 tensor cores (fp32 operands: kind::tf32, fp32 out)."""
    if not grouped_gemm_supported(a, b):
        raise NativeError("grouped_gemm: unsupported operand layout")
    G, M, K = a.shape
    tf32 = a.dtype == torch.float32
    fn = lib().fb_grouped_gemm_tf32 if tf32 else lib().fb_grouped_gemm_bf16
    return _launch_gemm(fn, a, b, G, M, b.shape[1], K, torch.float32 if tf32 else out_dtype, relu, bias, "grouped_gemm",
                        stats=stats)


def grouped_gemm_tn(a, b, out_dtype=torch.float32, out=None, out_is_zeroed=False):
    """d[g] = a[g]^T @ b[g]   (a [G,K,M], b [G,K,N]): weight gradients without transposing activations.
    The output is tiny and the reduction long, so when there are fewer tiles than CTA slots the reduction
    dimension is cut into slices that run as extra groups (split-K) and the fp32 partials are summed."""
    if not grouped_gemm_tn_supported(a, b):
        raise NativeError("grouped_gemm_tn: unsupported operand layout")
    G, K, M = a.shape
    N = b.shape[2]
    tiles = G * ((M + 127) // 128) * ((N + 255) // 256)
    S = 1
    tf32 = a.dtype == torch.float32
    BK = 32 if tf32 else 64
    fn_tn = lib().fb_grouped_gemm_tn_tf32 if tf32 else lib().fb_grouped_gemm_tn_bf16
    if tiles < 296 and a.is_contiguous() and b.is_contiguous() and K % BK == 0:
        kb = K // BK
        for cand in (16, 10, 8, 5, 4, 2):
            if kb % cand == 0 and kb // cand >= 4 and tiles * cand <= 1184:
                S = cand
                break
    # ``out``: fp32 [G, M, N] view with contiguous [M, N] blocks (e.g. a 1x1-conv gradient inside the arena)
    if out is not None and not (out.dtype == torch.float32 and out.stride(2) == 1 and out.stride(1) == N
                                and out.data_ptr() % 16 == 0 and (G == 1 or out.stride(0) % 4 == 0)):
        out = None
    if S == 1:
        return _launch_gemm(fn_tn, a, b, G, M, N, K, out_dtype, False, None, "grouped_gemm_tn",
                            out=out if out_dtype == torch.float32 else None)
    if out is not None and out_is_zeroed:
        # split-K slices reduce-add straight into the (pre-zeroed) output with TMA reduce-stores: no partial buffer
        return _launch_gemm(fn_tn, a.view(G * S, K // S, M), b.view(G * S, K // S, N), G * S, M, N,
                            K // S, torch.float32, False, None, "grouped_gemm_tn", out=out, accumulate=True, group_div=S)
    part = _launch_gemm(fn_tn, a.view(G * S, K // S, M), b.view(G * S, K // S, N), G * S, M, N,
                        K // S, torch.float32, False, None, "grouped_gemm_tn")
    if out is not None:
        return torch.sum(part.view(G, S, M, N), dim=1, out=out)
    return part.view(G, S, M, N).sum(dim=1).to(out_dtype)


def gram_bf16(hi: torch.Tensor, lo: Optional[torch.Tensor] = None, split: int = 0) -> torch.Tensor:
    """K x K Gram matrix of the rows of ``hi (+ lo)`` [K, n] bf16 with fp32 accumulation.  Rows must be padded
    with zeros to a multiple of 64 columns (``krum_pack`` does).  The long reduction dimension is cut into
    ``split`` slices that run as groups of the grouped GEMM (split-K across CTAs); the fp32 partials are summed
    by the streaming accumulate kernel.  With ``lo`` the product is hi*hi^T + hi*lo^T + lo*hi^T (bf16 hi/lo
    split of fp32 inputs; the lo*lo^T term is below fp32 round-off)."""
    K, n = hi.shape
    ld = hi.stride(0)
    if n % 64 or ld % 8:
        raise NativeError("gram_bf16: rows must be zero-padded to a multiple of 64 columns")
    if split <= 0:
        tiles = ((K + 127) // 128) * ((K + 255) // 256)
        split = max(1, min(n // 2048, (2 * 148 + tiles - 1) // tiles))
    chunk = ((n + split - 1) // split + 63) // 64 * 64
    split = (n + chunk - 1) // chunk
    tail = n - (split - 1) * chunk
    dev = hi.device

    def run(x, y):
        d = torch.empty((split, K, K), dtype=torch.float32, device=dev)
        p = FbGroupedGemm()
        p.a, p.b, p.d = x.data_ptr(), y.data_ptr(), d.data_ptr()
        p.a_gs = p.b_gs = chunk
        p.d_gs = K * K
        p.G, p.M, p.N, p.K = split, K, K, chunk
        p.lda = p.ldb = ld
        p.ldd, p.d_fp32 = K, 1
        if tail != chunk:                      # shorter last slice: its own launch with a smaller K
            if split > 1:
                p.G = split - 1
                check(lib().fb_grouped_gemm_bf16(C.byref(p), vp(stream_ptr(dev))), "gram_bf16")
                count_launch()
            q = FbGroupedGemm.from_buffer_copy(p)
            q.a, q.b = x.data_ptr() + (split - 1) * chunk * 2, y.data_ptr() + (split - 1) * chunk * 2
            q.d = d.data_ptr() + (split - 1) * K * K * 4
            q.G, q.K = 1, tail
            check(lib().fb_grouped_gemm_bf16(C.byref(q), vp(stream_ptr(dev))), "gram_bf16 tail")
        else:
            check(lib().fb_grouped_gemm_bf16(C.byref(p), vp(stream_ptr(dev))), "gram_bf16")
        count_launch()
        if split == 1:
            return d[0]
        if (K * K) % 4:
            return d.sum(0)
        out = torch.empty(K * K, dtype=torch.float32, device=dev)
        weighted_accumulate(d.view(split, K * K), torch.ones(split, dtype=torch.float32, device=dev), out)
        return out.view(K, K)

    g = run(hi, hi)
    if lo is not None:
        hl = run(hi, lo)
        g = g + hl + hl.t()
    return g


def stage_conv_weights(arena: torch.Tensor, stage: torch.Tensor, table: torch.Tensor, n_layers: int, max_n: int):
    """arena fp32 [C, ld] -> stage bf16 [C, stage_ld] following the int64 [n_layers, 8] device table."""
    fn = lib().fb_stage_conv_weights_f32 if stage.dtype == torch.float32 else lib().fb_stage_conv_weights
    check(fn(vp(ptr(arena)), i64(arena.stride(0)), i32(arena.shape[0]), vp(ptr(stage)),
                                      i64(stage.stride(0)), vp(ptr(table)), i32(n_layers), i32(max_n),
                                      vp(stream_ptr(arena.device))), "stage_conv_weights")
    count_launch()


def sgd_restage(arena: torch.Tensor, grads: torch.Tensor, scratch: Optional[torch.Tensor], stage: Optional[torch.Tensor],
                table: torch.Tensor, n_layers: int, max_n: int, lr: torch.Tensor):
    """Plain-SGD step on every listed segment fused with re-staging the bf16 GEMM operands (csrc/train.cu)."""
    fn = lib().fb_sgd_restage_f32 if (stage is not None and stage.dtype == torch.float32) else lib().fb_sgd_restage
    check(fn(vp(ptr(arena)), vp(ptr(grads)), i64(arena.stride(0)), i32(arena.shape[0]), vp(ptr(scratch)),
                               i64(scratch.stride(0) if scratch is not None else 0), vp(ptr(stage)),
                               i64(stage.stride(0) if stage is not None else 0), vp(ptr(table)), i32(n_layers), i32(max_n),
                               vp(ptr(lr)), vp(stream_ptr(arena.device))), "sgd_restage")
    count_launch()


def unstage_conv_grads(scratch: torch.Tensor, grads: torch.Tensor, table: torch.Tensor, n_layers: int, max_tiles: int):
    """GEMM-layout weight gradients (rows of ``scratch`` fp32 [C, *]) -> parameter layout inside the gradient arena."""
    check(lib().fb_unstage_conv_grads(vp(ptr(scratch)), i64(scratch.stride(0)), vp(ptr(grads)), i64(grads.stride(0)),
                                      i32(grads.shape[0]), vp(ptr(table)), i32(n_layers), i32(max_tiles),
                                      vp(stream_ptr(grads.device))), "unstage_conv_grads")
    count_launch()


# --------------------------------------------------------------------------- K14 amplitude normalisation (hand-written FFT)
def amp_normalize_supported(x: torch.Tensor) -> bool:
    ok = lambda v: 8 <= v <= 512 and (v & (v - 1)) == 0   # noqa: E731
    return x.is_cuda and x.dim() == 4 and ok(x.shape[2]) and ok(x.shape[3]) and x.shape[0] * x.shape[1] <= 65535


def amp_normalize(x: torch.Tensor, amp_running: Optional[torch.Tensor] = None, momentum: float = 0.1):
    """x [B, ch, H, W] -> (x rebuilt from the running amplitude and its own phase, new running amplitude [ch,H,W],
    batch-mean amplitude [ch,H,W]); shared-memory radix-2 FFTs, fp32 complex."""
    if not amp_normalize_supported(x):
        raise NativeError("amp_normalize: expected CUDA [B, ch, H, W] with power-of-two H, W in [8, 512]")
    B, ch, H, W = x.shape
    xf = x.to(torch.float32).contiguous()
    y = torch.empty_like(xf)
    amp_new = amp_running.to(torch.float32).clone().contiguous() if amp_running is not None else \
        torch.empty((ch, H, W), dtype=torch.float32, device=x.device)
    batch = torch.empty((ch, H, W), dtype=torch.float32, device=x.device)
    ws = torch.empty((B * ch * H * W * 2,), dtype=torch.float32, device=x.device)
    check(lib().fb_amp_normalize(vp(ptr(xf)), vp(ptr(y)), i32(B), i32(ch), i32(H), i32(W), vp(ptr(amp_new)), vp(ptr(batch)),
                                 f32(momentum), i32(1), i32(int(amp_running is not None)), vp(ptr(ws)),
                                 vp(stream_ptr(x.device))), "amp_normalize")
    count_launch(5)
    return y.to(x.dtype), amp_new, batch


# --------------------------------------------------------------------------- implicit 3x3 convolution (tcgen05, TMA-gathered taps)
def conv3x3_supported(x: torch.Tensor, cout: int, stride: int) -> bool:
    """x [C, B, H, W, cin] bf16 (or fp32: TF32 tensor cores) NHWC."""
    if not (x.is_cuda and x.dtype in (torch.bfloat16, torch.float32) and x.dim() == 5 and x.is_contiguous()
            and stride in (1, 2)):
        return False
    _, _, H, W, cin = x.shape
    Ho, Wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    pow2 = lambda v: v >= 1 and (v & (v - 1)) == 0   # noqa: E731
    row = 128 // x.element_size()                     # channels per 128-byte k-block
    ok_c = (cin * x.element_size() in (32, 64)) or (cin >= row and cin % row == 0)
    return ok_c and cout % 8 == 0 and cout >= 8 and pow2(Ho) and pow2(Wo) and Wo <= 128 and Wo * stride <= 256


def _conv3x3(x, w, cout, stride, dgrad, stats=None):
    from .native import FbConv3x3
    Cn, B, H, W, cin = x.shape
    Ho, Wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    y = torch.empty((Cn, B, Ho, Wo, cout), dtype=x.dtype, device=x.device)
    p = FbConv3x3()
    p.x, p.w, p.y = x.data_ptr(), w.data_ptr(), y.data_ptr()
    p.w_gs = w.stride(0)
    p.G, p.B, p.H, p.W, p.cin, p.cout, p.stride, p.dgrad = Cn, B, H, W, cin, cout, stride, int(dgrad)
    if stats is not None:
        if stats.dtype != torch.float32 or stats.shape != (Cn, 2, cout) or stats.stride(2) != 1 or stats.stride(1) != cout:
            raise NativeError("conv3x3: stats must be fp32 [C, 2, cout] with contiguous [2, cout] blocks")
        p.stats, p.stats_gs = stats.data_ptr(), stats.stride(0)
    fn = lib().fb_conv3x3_tf32 if x.dtype == torch.float32 else lib().fb_conv3x3_bf16
    check(fn(C.byref(p), vp(stream_ptr(x.device))), "conv3x3")
    count_launch()
    return y


def conv3x3(x: torch.Tensor, wm: torch.Tensor, stride: int = 1, stats: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Implicit-GEMM 3x3 convolution, pad 1: x [C,B,H,W,cin] bf16, wm [C, cout, 9*cin] ((kh,kw,cin) columns)."""
    cout = wm.shape[1]
    if not conv3x3_supported(x, cout, stride) or wm.shape[2] != 9 * x.shape[-1] or wm.stride(2) != 1 \
            or wm.stride(1) != wm.shape[2] or wm.dtype != x.dtype:
        raise NativeError("conv3x3: unsupported operand layout")
    return _conv3x3(x, wm, cout, stride, False, stats)


def conv3x3_dgrad(dy: torch.Tensor, wt: torch.Tensor) -> torch.Tensor:
    """Input gradient of a stride-1 3x3 convolution: dy [C,B,H,W,cout], wt [C, 9*cin, cout] (transposed staging copy)."""
    cin = wt.shape[1] // 9
    if not conv3x3_supported(dy, cin, 1) or wt.shape[2] != dy.shape[-1] or wt.stride(2) != 1 or wt.stride(1) != wt.shape[2] \
            or wt.dtype != dy.dtype:
        raise NativeError("conv3x3_dgrad: unsupported operand layout")
    return _conv3x3(dy, wt, cin, 1, True)


def conv3x3_wgrad_supported(x: torch.Tensor, dy: torch.Tensor, stride: int) -> bool:
    if not (x.is_cuda and x.dtype in (torch.bfloat16, torch.float32) and dy.dtype == x.dtype and x.dim() == 5
            and x.is_contiguous() and dy.is_contiguous() and stride in (1, 2)):
        return False
    _, B, H, W, cin = x.shape
    cout = dy.shape[-1]
    Ho, Wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    pow2 = lambda v: v >= 1 and (v & (v - 1)) == 0   # noqa: E731
    row = 128 // x.element_size()
    ok_c = (cin * x.element_size() in (32, 64) and x.dtype == torch.bfloat16) or (row <= cin <= 256 and cin % row == 0) \
        or (cin > 256 and cin % 256 == 0)       # fp32 operands fed MN-major need whole 128-byte channel blocks ...
    ok_c = ok_c or (x.dtype == torch.float32 and cin == 16)      # ... except 16 channels: one zero-padded 32-column tile per tap
    return ok_c and cin % 16 == 0 and cout % 8 == 0 and pow2(Ho) and pow2(Wo) and Wo <= 64 and (B * Ho * Wo) % row == 0 \
        and tuple(dy.shape[:4]) == (x.shape[0], B, Ho, Wo)


def conv3x3_wgrad(x: torch.Tensor, dy: torch.Tensor, stride: int = 1, out: Optional[torch.Tensor] = None,
                  out_is_zeroed: bool = False) -> torch.Tensor:
    """Weight gradient of the implicit 3x3 convolution: x [C,B,H,W,cin], dy [C,B,Ho,Wo,cout] ->
    fp32 [C, cout, 9*cin] ((kh,kw,cin) columns).  The pixel reduction is split across CTAs when there are few tiles;
    the slices reduce-add into ``out`` (zeroed here unless the caller already did)."""
    from .native import FbConv3x3Wgrad
    if not conv3x3_wgrad_supported(x, dy, stride):
        raise NativeError("conv3x3_wgrad: unsupported operand layout")
    Cn, B, H, W, cin = x.shape
    cout = dy.shape[-1]
    Ho, Wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    tpt = 9 if 9 * cin <= 256 else (3 if 3 * cin <= 256 else 1)     # taps per tile, as in the kernel launcher
    if x.dtype == torch.float32 and cin == 16:
        tpt = 1
    tiles = Cn * ((cout + 127) // 128) * (9 // tpt) * (cin // min(cin, 256))
    kb = B * Ho * Wo // (128 // x.element_size())
    S = 1
    if tiles < 296:
        for cand in (16, 10, 8, 5, 4, 2):
            if kb % cand == 0 and kb // cand >= 4 and tiles * cand <= 1184:
                S = cand
                break
    if out is None:
        out = torch.zeros((Cn, cout, 9 * cin), dtype=torch.float32, device=x.device) if S > 1 else \
            torch.empty((Cn, cout, 9 * cin), dtype=torch.float32, device=x.device)
    elif S > 1 and not out_is_zeroed:
        out.zero_()
    if not (out.dtype == torch.float32 and out.stride(2) == 1 and out.stride(1) == 9 * cin and out.data_ptr() % 16 == 0
            and (Cn == 1 or out.stride(0) % 4 == 0)):
        raise NativeError("conv3x3_wgrad: output must be fp32 [C, cout, 9*cin] with contiguous blocks")
    p = FbConv3x3Wgrad()
    p.x, p.dy, p.dw = x.data_ptr(), dy.data_ptr(), out.data_ptr()
    p.dw_gs = out.stride(0) if Cn > 1 else cout * 9 * cin
    p.G, p.B, p.H, p.W, p.cin, p.cout, p.stride, p.split, p.accumulate = Cn, B, H, W, cin, cout, stride, S, 0
    fn = lib().fb_conv3x3_wgrad_tf32 if x.dtype == torch.float32 else lib().fb_conv3x3_wgrad_bf16
    check(fn(C.byref(p), vp(stream_ptr(x.device))), "conv3x3_wgrad")
    count_launch()
    return out
