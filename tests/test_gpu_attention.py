"""The tcgen05 flash-attention kernels (forward, dQ, dK/dV) against an fp64 numpy restatement of keras MultiHeadAttention's
softmax(QK^T)V (models/CvT(Par).py:185) on the shapes of the three stages and on ragged ones.
Tolerance: TF32 operands (10-bit mantissa) -> max|a-b| / max|b| <= 2e-3 per tensor."""
import ctypes as C

import numpy as np
import pytest

from transformer_stm_b200 import _lib

pytestmark = pytest.mark.gpu


def ref_attention(Q, K, V, dO, heads):
    B, T, D = Q.shape
    q, k, v, do = (a.astype(np.float64).reshape(B, T, heads, 64) for a in (Q, K, V, dO))
    S = np.einsum("bthd,bshd->bhts", q, k)
    m = S.max(-1, keepdims=True)
    E = np.exp(S - m)
    l = E.sum(-1, keepdims=True)
    P = E / l
    O = np.einsum("bhts,bshd->bthd", P, v)
    lse = (m + np.log(l))[..., 0]
    dV = np.einsum("bhts,bthd->bshd", P, do)
    dP = np.einsum("bthd,bshd->bhts", do, v)
    dS = P * (dP - (dP * P).sum(-1, keepdims=True))
    dQ = 0.125 * np.einsum("bhts,bshd->bthd", dS, k)
    dK = np.einsum("bhts,bthd->bshd", dS, q)
    return O.reshape(B, T, D), lse.reshape(B * heads, T), dQ.reshape(B, T, D), dK.reshape(B, T, D), dV.reshape(B, T, D)


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


@pytest.mark.parametrize("B,T,heads", [(2, 1024, 1), (3, 256, 2), (2, 65, 4), (1, 485, 3), (1, 64, 1), (2, 129, 2)])
def test_flash_attention_forward_backward(B, T, heads):
    lib = _lib.load()
    D = heads * 64
    rng = np.random.default_rng(B * 1000 + T + heads)
    Q = (0.125 * 2.0 * rng.standard_normal((B, T, D))).astype(np.float32)      # pre-scaled queries, logits of a few units
    K = (2.0 * rng.standard_normal((B, T, D))).astype(np.float32)
    V = rng.standard_normal((B, T, D)).astype(np.float32)
    dO = rng.standard_normal((B, T, D)).astype(np.float32)
    O = np.zeros_like(Q); dQ = np.zeros_like(Q); dK = np.zeros_like(Q); dV = np.zeros_like(Q)
    LSE = np.zeros((B * heads, T), np.float32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    _lib.check(lib.cvt_debug_attention(B, T, heads, p(Q), p(K), p(V), p(dO), p(O), p(LSE), p(dQ), p(dK), p(dV)), "cvt_debug_attention")
    rO, rL, rQ, rK, rV = ref_attention(Q, K, V, dO, heads)
    errs = dict(O=rel(O, rO), LSE=rel(LSE, rL), dQ=rel(dQ, rQ), dK=rel(dK, rK), dV=rel(dV, rV))
    assert all(e < 2e-3 for e in errs.values()), errs


# (40, 320, 2), (10, 1100, 2), (3, 1892, 3): odd tile counts (3, 9, 15) in the persistent 128-wide backward kernels - a CTA that walks several
# work items then starts every other item on the opposite buffer / barrier parity; 2100 tokens: the dK/dV kernel's tile-by-tile statistics path
@pytest.mark.parametrize("B,T,heads", [(64, 1024, 1), (64, 256, 2), (64, 65, 4), (160, 1024, 1), (40, 320, 2), (10, 1100, 2), (3, 1892, 3), (12, 2100, 1)])
def test_flash_attention_many_ctas_per_sm(B, T, heads):
    """Grids of several hundred CTAs (two forward CTAs share an SM and its tensor pipe, several waves): the shapes of the benched batch-64
    step.  Every image carries the same Q/K/V/dO, so each must reproduce the fp64 result of image 0.  Round 2 found a race here that the
    small-grid cases above cannot see: a forward row warp read O before the last two PV products had landed."""
    lib = _lib.load()
    D = heads * 64
    rng = np.random.default_rng(T + heads)
    one = [(s * rng.standard_normal((1, T, D))).astype(np.float32) for s in (0.25, 2.0, 1.0, 1.0)]
    rO, rL, rQ, rK, rV = ref_attention(*one, heads)
    Q, K, V, dO = (np.ascontiguousarray(np.repeat(a, B, 0)) for a in one)
    O = np.zeros_like(Q); dQ = np.zeros_like(Q); dK = np.zeros_like(Q); dV = np.zeros_like(Q)
    LSE = np.zeros((B * heads, T), np.float32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    for rep in range(3):
        _lib.check(lib.cvt_debug_attention(B, T, heads, p(Q), p(K), p(V), p(dO), p(O), p(LSE), p(dQ), p(dK), p(dV)), "cvt_debug_attention")
        for name, got, ref in (("O", O, rO), ("dQ", dQ, rQ), ("dK", dK, rK), ("dV", dV, rV)):
            worst = max(rel(got[b], ref[0]) for b in range(B))
            assert worst < 2e-3, (name, rep, worst)
