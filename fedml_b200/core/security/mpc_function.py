"""Finite-field toolbox for secure aggregation in GF(p), p prime < 2^31.

Function names and argument orders follow the reference toolbox
(simulation/single_process/turboaggregate/mpc_function.py:4-281) so TurboAggregate-style code ports over:
modular inverse / division, Lagrange coefficient matrices, BGW (Shamir) encode/decode, Lagrange coded
computing (LCC) encode/decode (+ with given randomness / points), additive secret sharing and DH-style key
agreement.  Implementation differs: inverses use Fermat (``pow(a, p-2, p)``), the Lagrange matrix is built
with exact Python integers, and the matrix-times-data products — the only O(N*d) work — go through
``ff_matmul`` which runs on the GPU kernel ``fb_ff_matmul`` (csrc/secagg.cu) for CUDA tensors and on int64
NumPy otherwise."""
from __future__ import annotations

import numpy as np

DEFAULT_PRIME = 2 ** 31 - 1   # Mersenne prime: every element < 2^31 as the kernels require


def modular_inv(a, p):
    a = int(a) % int(p)
    if a == 0:
        raise ZeroDivisionError("0 has no inverse in GF(p)")
    return pow(a, int(p) - 2, int(p))


def divmod(_num, _den, _p):  # noqa: A001 - reference name
    return (int(_num) % int(_p)) * modular_inv(_den, _p) % int(_p)


def PI(vals, p):  # noqa: N802 - reference name
    acc = 1
    for v in vals:
        acc = acc * (int(v) % int(p)) % int(p)
    return acc


def gen_Lagrange_coeffs(alpha_s, beta_s, p, is_K1=0):  # noqa: N802,N803
    """U[i][j] = prod_{o != beta_j} (alpha_i - o) / (beta_j - o)  mod p."""
    alpha_s, beta_s = [int(a) for a in alpha_s], [int(b) for b in beta_s]
    n_alpha = 1 if is_K1 == 1 else len(alpha_s)
    U = np.zeros((n_alpha, len(beta_s)), dtype=np.int64)
    for j, bj in enumerate(beta_s):
        others = [o for o in beta_s if o != bj]
        den_inv = modular_inv(PI([bj - o for o in others], p), p)
        for i in range(n_alpha):
            U[i, j] = PI([alpha_s[i] - o for o in others], p) * den_inv % p
    return U


def ff_matmul(U, X, p):
    """(U @ X) mod p with exact arithmetic.  U [n_out, n_in] int64 < p; X [n_in, d]."""
    try:
        import torch
        if torch.is_tensor(X) and X.is_cuda:
            from ... import ops
            Ut = torch.as_tensor(np.asarray(U) % p, dtype=torch.int64).to(torch.int32).to(X.device)
            return ops.ff_matmul(Ut, X.to(torch.int32).contiguous(), int(p))
    except ImportError:
        pass
    U = np.asarray(U, dtype=np.int64) % p
    X = np.asarray(X, dtype=np.int64) % p
    out = np.zeros((U.shape[0],) + X.shape[1:], dtype=np.int64)
    for j in range(U.shape[1]):          # products < 2^62: reduce after every term
        out = (out + (U[:, j].reshape((-1,) + (1,) * (X.ndim - 1)) * X[j][None]) % p) % p
    return out


def BGW_encoding(X, N, T, p, rng=None):  # noqa: N802,N803
    """Shamir shares: share_i = sum_t R_t * alpha_i^t with R_0 = X, alpha_i = i+1."""
    rng = rng or np.random
    X = np.asarray(X, dtype=np.int64) % p
    R = np.concatenate([X[None], rng.randint(p, size=(T,) + X.shape).astype(np.int64)], axis=0)
    V = np.array([[pow(a, t, p) for t in range(T + 1)] for a in range(1, N + 1)], dtype=np.int64)
    return ff_matmul(V, R.reshape(T + 1, -1), p).reshape((N,) + X.shape)


def gen_BGW_lambda_s(alpha_s, p):  # noqa: N802
    alpha_s = [int(a) for a in alpha_s]
    lam = np.zeros((1, len(alpha_s)), dtype=np.int64)
    for i, a in enumerate(alpha_s):
        others = [o for o in alpha_s if o != a]
        lam[0, i] = divmod(PI([0 - o for o in others], p), PI([a - o for o in others], p), p)
    return lam


def BGW_decoding(f_eval, worker_idx, p):  # noqa: N802
    alpha_eval = [(int(i) + 1) % p for i in worker_idx]
    lam = gen_BGW_lambda_s(alpha_eval, p)
    f_eval = np.asarray(f_eval, dtype=np.int64)
    return ff_matmul(lam, f_eval.reshape(len(alpha_eval), -1), p).reshape((1,) + f_eval.shape[1:])


def _lcc_points(N, K, T, p):
    n_beta = K + T
    stt_b, stt_a = -int(np.floor(n_beta / 2)), -int(np.floor(N / 2))
    beta_s = np.mod(np.arange(stt_b, stt_b + n_beta), p).astype(np.int64)
    alpha_s = np.mod(np.arange(stt_a, stt_a + N), p).astype(np.int64)
    return alpha_s, beta_s


def LCC_encoding_w_Random(X, R_, N, K, T, p):  # noqa: N802,N803
    X = np.asarray(X, dtype=np.int64) % p
    m = X.shape[0]
    sub = [X[i * m // K:(i + 1) * m // K] for i in range(K)] + [np.asarray(R_[t], dtype=np.int64) % p for t in range(T)]
    X_sub = np.stack(sub)
    alpha_s, beta_s = _lcc_points(N, K, T, p)
    U = gen_Lagrange_coeffs(alpha_s, beta_s, p)
    return ff_matmul(U, X_sub.reshape(K + T, -1), p).reshape((N,) + X_sub.shape[1:])


def LCC_encoding(X, N, K, T, p, rng=None):  # noqa: N802,N803
    rng = rng or np.random
    X = np.asarray(X)
    m = X.shape[0]
    R_ = rng.randint(p, size=(T, m // K) + X.shape[1:]).astype(np.int64)
    return LCC_encoding_w_Random(X, R_, N, K, T, p)


def LCC_encoding_w_Random_partial(X, R_, N, K, T, p, worker_idx):  # noqa: N802,N803
    return LCC_encoding_w_Random(X, R_, N, K, T, p)[list(worker_idx)]


def LCC_decoding(f_eval, f_deg, N, K, T, worker_idx, p):  # noqa: N802,N803
    alpha_s, _ = _lcc_points(N, K, T, p)
    stt_b = -int(np.floor(K / 2))
    beta_s = np.mod(np.arange(stt_b, stt_b + K), p).astype(np.int64)
    # NB: decoding targets are the first K interpolation points of the *encoder* only when T == 0 (reference
    # behaviour, mpc_function.py:201-217); for T > 0 use LCC_decoding_with_points with the encoder's beta_s.
    U_dec = gen_Lagrange_coeffs(beta_s, [alpha_s[i] for i in worker_idx], p)
    f_eval = np.asarray(f_eval, dtype=np.int64)
    return ff_matmul(U_dec, f_eval.reshape(len(worker_idx), -1), p).reshape((K,) + f_eval.shape[1:])


def LCC_encoding_with_points(X, alpha_s, beta_s, p):  # noqa: N802,N803
    """Interpolate through (alpha_j, X_j) and evaluate at beta_s."""
    U = gen_Lagrange_coeffs(beta_s, alpha_s, p)
    X = np.asarray(X, dtype=np.int64)
    return ff_matmul(U, X.reshape(len(alpha_s), -1), p).reshape((len(beta_s),) + X.shape[1:])


def LCC_decoding_with_points(f_eval, eval_points, target_points, p):  # noqa: N802
    U = gen_Lagrange_coeffs(target_points, eval_points, p)
    f_eval = np.asarray(f_eval, dtype=np.int64)
    return ff_matmul(U, f_eval.reshape(len(eval_points), -1), p).reshape((len(target_points),) + f_eval.shape[1:])


def Gen_Additive_SS(d, n_out, p, rng=None):  # noqa: N802
    rng = rng or np.random
    temp = rng.randint(0, p, size=(n_out - 1, d)).astype(np.int64)
    last = np.mod(-temp.sum(axis=0), p).reshape(1, d)
    return np.concatenate([temp, last], axis=0)


def my_pk_gen(my_sk, p, g):
    return int(my_sk) if g == 0 else pow(int(g), int(my_sk), int(p))


def my_key_agreement(my_sk, u_pk, p, g):
    return int(my_sk) * int(u_pk) % int(p) if g == 0 else pow(int(u_pk), int(my_sk), int(p))


def quantize(x, scale, p):
    """float -> GF(p): round(x*scale), negatives wrap to p - |.| (same rule as csrc/secagg.cu)."""
    q = np.rint(np.asarray(x, dtype=np.float64) * scale).astype(np.int64)
    return np.mod(q, p)


def dequantize(q, scale, p):
    q = np.asarray(q, dtype=np.int64)
    signed = np.where(q > p // 2, q - p, q)
    return signed.astype(np.float64) / scale
