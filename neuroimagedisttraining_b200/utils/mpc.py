"""Finite-field building blocks for secure aggregation (TurboAggregate): modular inverse, Lagrange coefficients,
Shamir/BGW and Lagrange-coded (LCC) encoding/decoding, additive secret sharing, Diffie-Hellman style key agreement and
fixed-point quantisation of model updates.  Same function surface as the reference's
``fedml_api/standalone/turboaggregate/mpc_function.py:4-275`` (whose trainer never calls them); here they are exercised by
:class:`fl.turboaggregate.TurboAggregateAPI` and tests.  All arithmetic is in Python ints / int64 numpy modulo a prime ``p``
(< 2^31 so products fit int64).
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np

DEFAULT_PRIME = 2 ** 31 - 1      # Mersenne prime: products of two residues fit in int64


def modular_inv(a: int, p: int) -> int:
    """Multiplicative inverse of ``a`` modulo the prime ``p``."""
    a = int(a) % p
    if a == 0:
        raise ZeroDivisionError("0 has no inverse")
    return pow(a, p - 2, p)


def divmod(_num, _den, _p):  # noqa: A001 - name kept for API parity
    """``num / den`` modulo the prime ``p``."""
    return (int(_num) % _p) * modular_inv(int(_den) % _p, _p) % _p


def PI(vals: Sequence[int], p: int) -> int:
    """Product of the inputs modulo ``p``."""
    acc = 1
    for v in vals:
        acc = acc * (int(v) % p) % p
    return acc


def gen_Lagrange_coeffs(alpha_s: Sequence[int], beta_s: Sequence[int], p: int, is_K1: int = 0) -> np.ndarray:
    """``U[i][j] = prod_{o != beta_j} (alpha_i - o) / (beta_j - o)``: evaluates at ``alpha_i`` the polynomial interpolated at ``beta_s``."""
    n_alpha = 1 if is_K1 == 1 else len(alpha_s)
    U = np.zeros((n_alpha, len(beta_s)), dtype=np.int64)
    for i in range(n_alpha):
        for j, bj in enumerate(beta_s):
            den = PI([bj - o for o in beta_s if o != bj], p)
            num = PI([alpha_s[i] - o for o in beta_s if o != bj], p)
            U[i][j] = divmod(num, den, p)
    return U


def _matmul_mod(U: np.ndarray, X: np.ndarray, p: int) -> np.ndarray:
    """``U @ X mod p`` without int64 overflow (accumulates one rank-1 term at a time)."""
    out = np.zeros((U.shape[0],) + X.shape[1:], dtype=np.int64)
    for j in range(U.shape[1]):
        out = (out + (U[:, j].reshape((-1,) + (1,) * (X.ndim - 1)) * X[j][None]) % p) % p
    return out


def BGW_encoding(X: np.ndarray, N: int, T: int, p: int, rng: np.random.RandomState = None) -> np.ndarray:
    """Shamir shares of ``X`` (shape [m, d]) for ``N`` parties with threshold ``T``: ``f(a) = X + R_1 a + .. + R_T a^T`` at a = 1..N."""
    rng = rng or np.random.RandomState()
    X = np.mod(np.asarray(X, dtype=np.int64), p)
    coeffs = [X] + [rng.randint(0, p, size=X.shape).astype(np.int64) for _ in range(T)]
    out = np.zeros((N,) + X.shape, dtype=np.int64)
    for i in range(N):
        a, acc, ak = i + 1, np.zeros_like(X), 1
        for c in coeffs:
            acc = (acc + c * ak) % p
            ak = ak * a % p
        out[i] = acc
    return out


def gen_BGW_lambda_s(alpha_s: Sequence[int], p: int) -> np.ndarray:
    """Lagrange weights that reconstruct ``f(0)`` from evaluations at ``alpha_s``."""
    lam = np.zeros((1, len(alpha_s)), dtype=np.int64)
    for i, ai in enumerate(alpha_s):
        num = PI([0 - o for o in alpha_s if o != ai], p)
        den = PI([ai - o for o in alpha_s if o != ai], p)
        lam[0][i] = divmod(num, den, p)
    return lam


def BGW_decoding(f_eval: np.ndarray, worker_idx: Sequence[int], p: int) -> np.ndarray:
    """Secret ``f(0)`` from T+1 (or more) shares; ``worker_idx`` are 0-based party indices of the shares."""
    alpha = [int(i) + 1 for i in worker_idx]
    lam = gen_BGW_lambda_s(alpha, p)
    return _matmul_mod(lam, np.asarray(f_eval, dtype=np.int64), p)


def _lcc_points(N: int, K: int, T: int, p: int) -> Tuple[List[int], List[int]]:
    """Interpolation points beta (K data + T random blocks) and evaluation points alpha (N workers): the reference's centred
    ranges reduced mod p (``mpc_function.py:121-125``), so shares are interchangeable with its encoder / decoder."""
    n_beta = K + T
    stt_b, stt_a = -(n_beta // 2), -(N // 2)
    beta = [int(v % p) for v in range(stt_b, stt_b + n_beta)]
    alpha = [int(v % p) for v in range(stt_a, stt_a + N)]
    return alpha, beta


def LCC_encoding_with_points(X: np.ndarray, alpha_s: Sequence[int], beta_s: Sequence[int], p: int) -> np.ndarray:
    """Evaluate at ``alpha_s`` the polynomial that takes the blocks ``X[j]`` at ``beta_s[j]``."""
    U = gen_Lagrange_coeffs(alpha_s, beta_s, p)
    return _matmul_mod(U, np.mod(np.asarray(X, dtype=np.int64), p), p)


def LCC_encoding_w_Random(X: np.ndarray, R_: np.ndarray, N: int, K: int, T: int, p: int) -> np.ndarray:
    """LCC shares of ``X`` ([m, d], split into K row blocks) with the given T random blocks ``R_`` ([T, m/K, d])."""
    X = np.mod(np.asarray(X, dtype=np.int64), p)
    m = X.shape[0]
    assert m % K == 0, "rows must split evenly into K blocks"
    blocks = np.concatenate([X.reshape(K, m // K, *X.shape[1:]), np.mod(np.asarray(R_, dtype=np.int64), p)], axis=0)
    alpha, beta = _lcc_points(N, K, T, p)
    return LCC_encoding_with_points(blocks, alpha, beta, p)


def LCC_encoding(X: np.ndarray, N: int, K: int, T: int, p: int, rng: np.random.RandomState = None) -> np.ndarray:
    rng = rng or np.random.RandomState()
    X = np.asarray(X)
    R_ = rng.randint(0, p, size=(T, X.shape[0] // K) + X.shape[1:]).astype(np.int64)
    return LCC_encoding_w_Random(X, R_, N, K, T, p)


def LCC_encoding_w_Random_partial(X, R_, N, K, T, p, worker_idx: Sequence[int]) -> np.ndarray:
    """Shares for a subset of workers only."""
    return LCC_encoding_w_Random(X, R_, N, K, T, p)[np.asarray(worker_idx)]


def LCC_decoding_with_points(f_eval: np.ndarray, eval_points: Sequence[int], target_points: Sequence[int], p: int) -> np.ndarray:
    U = gen_Lagrange_coeffs(target_points, eval_points, p)
    return _matmul_mod(U, np.mod(np.asarray(f_eval, dtype=np.int64), p), p)


def LCC_decoding(f_eval: np.ndarray, f_deg: int, N: int, K: int, T: int, worker_idx: Sequence[int], p: int) -> np.ndarray:
    """Recover the K data blocks of a degree-``f_deg`` function of LCC shares from ``f_deg*(K+T-1)+1`` worker results."""
    need = f_deg * (K + T - 1) + 1
    assert len(worker_idx) >= need, f"need {need} evaluations, got {len(worker_idx)}"
    alpha, beta = _lcc_points(N, K, T, p)
    pts = [alpha[int(i)] for i in worker_idx[:need]]
    return LCC_decoding_with_points(np.asarray(f_eval)[:need], pts, beta[:K], p)


def Gen_Additive_SS(d: int, n_out: int, p: int, rng: np.random.RandomState = None) -> np.ndarray:
    """``n_out`` random vectors of length ``d`` that sum to zero modulo ``p`` (additive masks)."""
    rng = rng or np.random.RandomState()
    shares = rng.randint(0, p, size=(n_out - 1, d)).astype(np.int64)
    last = np.mod(-np.sum(shares, axis=0), p).reshape(1, d)
    return np.concatenate([shares, last], axis=0)


def my_pk_gen(my_sk: int, p: int, g: int) -> int:
    """Public key ``g^sk mod p`` (``g == 0`` returns the secret itself, as the reference does for testing)."""
    return int(my_sk) if g == 0 else pow(int(g), int(my_sk), p)


def my_key_agreement(my_sk: int, u_pk: int, p: int, g: int) -> int:
    """Shared key ``pk_other^sk mod p`` (``g == 0``: product, test mode)."""
    return int(my_sk) * int(u_pk) % p if g == 0 else pow(int(u_pk), int(my_sk), p)


# ------------------------------------------------------------------ fixed-point <-> field
def quantize(x: np.ndarray, scale: int, p: int) -> np.ndarray:
    """Real vector -> field elements: round(x*scale), negatives wrap to ``p - |v|``."""
    v = np.rint(np.asarray(x, dtype=np.float64) * scale).astype(np.int64)
    return np.mod(v, p)


def dequantize(v: np.ndarray, scale: int, p: int) -> np.ndarray:
    """Field elements -> reals, interpreting the upper half of the field as negatives."""
    v = np.asarray(v, dtype=np.int64)
    signed = np.where(v > p // 2, v - p, v)
    return signed.astype(np.float64) / scale
