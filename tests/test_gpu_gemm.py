"""The tcgen05 (TF32) contraction kernel against the fp32 FFMA kernel and numpy, for every operand major-ness."""
import ctypes as C

import numpy as np
import pytest

from transformer_stm_b200 import _lib

pytestmark = pytest.mark.gpu


def run(a_mode, b_mode, M, N, K, splitk, use_tc, A, B):
    lib = _lib.load()
    out = np.zeros((M, N), np.float32)
    _lib.check(lib.cvt_debug_gemm(a_mode, b_mode, M, N, K, splitk, use_tc, A.ctypes.data_as(C.c_void_p), B.ctypes.data_as(C.c_void_p),
                                  out.ctypes.data_as(C.c_void_p)), "cvt_debug_gemm")
    return out


def tf32(x):
    """round-to-nearest (ties away) to 10 mantissa bits, what cvt.rna.tf32.f32 does"""
    u = x.astype(np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0x1000) & 0xFFFFE000
    return u.astype(np.uint32).view(np.float32)


CASES = [(128, 64, 32), (128, 64, 64), (256, 128, 96), (300, 256, 64), (128, 64, 49), (1000, 192, 256), (64, 1024, 128), (4160, 256, 256)]


@pytest.mark.parametrize("a_mode,b_mode", [(0, 1), (0, 0), (1, 1), (1, 0)])
@pytest.mark.parametrize("M,N,K", CASES)
def test_tc_gemm_matches_ffma(a_mode, b_mode, M, N, K):
    rng = np.random.default_rng(M * 7 + N * 3 + K + a_mode * 2 + b_mode)
    A = rng.standard_normal((M, K)).astype(np.float32)
    B = rng.standard_normal((K, N)).astype(np.float32)
    Ah = np.ascontiguousarray(A if a_mode == 0 else A.T)
    Bh = np.ascontiguousarray(B if b_mode == 0 else B.T)
    ref = tf32(A).astype(np.float64) @ tf32(B).astype(np.float64)
    simt = run(a_mode, b_mode, M, N, K, 1, 0, Ah, Bh)
    assert np.abs(simt - A.astype(np.float64) @ B.astype(np.float64)).max() < 1e-3 * np.sqrt(K)
    got = run(a_mode, b_mode, M, N, K, 1, 1, Ah, Bh)
    err = np.abs(got - ref).max()
    assert err < 2e-4 * np.sqrt(K), (a_mode, b_mode, M, N, K, err, np.abs(got).max())


@pytest.mark.parametrize("a_mode,b_mode", [(1, 0), (0, 1)])
def test_tc_gemm_splitk_atomic(a_mode, b_mode):
    M, N, K = 192, 128, 4096
    rng = np.random.default_rng(5)
    A = rng.standard_normal((M, K)).astype(np.float32); B = rng.standard_normal((K, N)).astype(np.float32)
    Ah = np.ascontiguousarray(A if a_mode == 0 else A.T); Bh = np.ascontiguousarray(B if b_mode == 0 else B.T)
    ref = tf32(A).astype(np.float64) @ tf32(B).astype(np.float64)
    got = run(a_mode, b_mode, M, N, K, 8, 1, Ah, Bh)
    assert np.abs(got - ref).max() < 2e-4 * np.sqrt(K)


def tf32_trunc(x):
    """what the tensor core does to raw fp32 bits under kind::tf32: the low 13 mantissa bits are ignored"""
    return (x.astype(np.float32).view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)


@pytest.mark.parametrize("M,N,K", [(128, 64, 32), (128, 64, 64), (256, 128, 96), (300, 256, 64), (1000, 192, 256), (64, 1024, 128),
                                   (4160, 256, 256), (65536, 64, 64), (20000, 256, 64), (16384, 128, 512), (4160, 256, 1024)])
def test_tma_gemm_matches_truncated_tf32(M, N, K):
    """persistent TMA-fed kernel (gemm_tma.cu): operands reach the tensor core unrounded, i.e. truncated to TF32"""
    rng = np.random.default_rng(M + N * 3 + K)
    A = rng.standard_normal((M, K)).astype(np.float32)
    B = rng.standard_normal((K, N)).astype(np.float32)
    ref = tf32_trunc(A).astype(np.float64) @ tf32_trunc(B).astype(np.float64)
    got = run(0, 1, M, N, K, 1, 4, A, np.ascontiguousarray(B.T))
    err = np.abs(got - ref).max()
    assert err < 2e-4 * np.sqrt(K), (M, N, K, err, np.abs(got).max())


@pytest.mark.parametrize("Kin,N,tokens", [(64, 64, 256), (64, 64, 65536), (64, 256, 8192), (256, 64, 8192), (128, 128, 16384), (128, 512, 4096),
                                          (512, 128, 4096), (256, 256, 4160), (256, 1024, 4160), (1024, 256, 4160), (192, 192, 1000)])
def test_tma_wgrad_matches_truncated_tf32(Kin, N, tokens):
    """TMA-fed split-K weight-gradient kernel (gemm_wgrad.cu): dW = X^T dY with both operands read MN-major"""
    rng = np.random.default_rng(Kin + N * 3 + tokens)
    X = rng.standard_normal((tokens, Kin)).astype(np.float32)
    dY = rng.standard_normal((tokens, N)).astype(np.float32)
    ref = tf32_trunc(X).astype(np.float64).T @ tf32_trunc(dY).astype(np.float64)
    got = run(1, 0, Kin, N, tokens, 1, 8, X, dY)
    err = np.abs(got - ref).max()
    assert err < 2e-4 * np.sqrt(tokens), (Kin, N, tokens, err, np.abs(got).max())


@pytest.mark.parametrize("M,N,K", [(128, 64, 32), (256, 128, 96), (300, 256, 64), (1000, 192, 256), (65536, 64, 64), (16384, 128, 512),
                                   (4160, 256, 1024), (4160, 1024, 256)])
def test_tma_gemm_x3_is_fp32_grade(M, N, K):
    """The error-compensated variant (CVT_PREC_TF32X3): operand tiles split into high and low parts in shared memory, three tcgen05 MMAs per
    k-step.  Against the fp64 product of the UNROUNDED operands: at least 50x closer than one TF32 pass."""
    rng = np.random.default_rng(M + N * 3 + K + 1)
    A = rng.standard_normal((M, K)).astype(np.float32)
    B = rng.standard_normal((K, N)).astype(np.float32)
    ref = A.astype(np.float64) @ B.astype(np.float64)
    one = run(0, 1, M, N, K, 1, 4, A, np.ascontiguousarray(B.T))
    got = run(0, 1, M, N, K, 1, 4 | 16, A, np.ascontiguousarray(B.T))
    e1, e3 = np.abs(one - ref).max(), np.abs(got - ref).max()
    # what is left is the tensor core's own accumulation (one-sided rounding of the running sum in TMEM, ~1e-6 per accumulated k): linear in K
    assert e3 < 2e-6 * K + 2e-5 and e3 < e1 / 50, (M, N, K, e1, e3)


@pytest.mark.parametrize("Kin,N,tokens", [(64, 64, 65536), (64, 256, 8192), (256, 64, 8192), (128, 512, 4096), (256, 1024, 4160), (192, 192, 1000)])
def test_tma_wgrad_x3_is_fp32_grade(Kin, N, tokens):
    rng = np.random.default_rng(Kin + N * 3 + tokens + 1)
    X = rng.standard_normal((tokens, Kin)).astype(np.float32)
    dY = rng.standard_normal((tokens, N)).astype(np.float32)
    ref = X.astype(np.float64).T @ dY.astype(np.float64)
    one = run(1, 0, Kin, N, tokens, 1, 8, X, dY)
    got = run(1, 0, Kin, N, tokens, 1, 8 | 16, X, dY)
    e1, e3 = np.abs(one - ref).max(), np.abs(got - ref).max()
    assert e3 < 1e-6 * tokens + 2e-5 and e3 < e1 / 50, (Kin, N, tokens, e1, e3)      # split-K partials are added with fp32 atomics
