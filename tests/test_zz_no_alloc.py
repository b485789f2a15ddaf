"""cvt_attention_backward_static: the allocation-free variant of the flash-attention backward (scratch in a static __device__ array)
against the existing entry point cvt_debug_attention, through the C-ABI, on two calls in a row.  Collected last on purpose;
it must pass alone in a fresh process and after every other test.
Tolerance: that of tests/test_gpu_attention.py (max|a-b| / max|b| <= 2e-3 per tensor); both entry points launch the same kernels, so the
results are also expected to agree to the last bit."""
import ctypes as C

import numpy as np
import pytest

from transformer_stm_b200 import _lib

pytestmark = pytest.mark.gpu

MAX_ROWS = 1 << 20      # CVT_STATIC_ATTN_MAX_ROWS (include/cvt_b200.h)


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def to_tf32(x):
    """Round to the nearest TF32 value (ties away from zero), what the producers of Q / K / V / dO do in the model."""
    bits = np.ascontiguousarray(x, np.float32).view(np.uint32)
    return ((bits + np.uint32(0x1000)) & np.uint32(0xFFFFE000)).view(np.float32)


def existing_entry_point(lib, B, T, heads, Q, K, V, dO):
    O = np.zeros_like(Q); dQ = np.zeros_like(Q); dK = np.zeros_like(Q); dV = np.zeros_like(Q)
    LSE = np.zeros((B * heads, T), np.float32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    _lib.check(lib.cvt_debug_attention(B, T, heads, p(Q), p(K), p(V), p(dO), p(O), p(LSE), p(dQ), p(dK), p(dV)), "cvt_debug_attention")
    return O, LSE, dQ, dK, dV


def test_static_scratch_attention_backward_matches_the_allocating_entry_point():
    import torch
    lib = _lib.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    ptr = lambda t: C.c_void_p(t.data_ptr())
    # two calls in a row on different shapes: the second finds the first one's contents in the shared scratch array
    for call, (B, T, heads) in enumerate([(3, 485, 2), (2, 1024, 1)]):
        D = heads * 64
        rng = np.random.default_rng(100 + call)
        Q, K, V, dO = (to_tf32(s * rng.standard_normal((B, T, D))) for s in (0.25, 2.0, 1.0, 1.0))
        O, LSE, rQ, rK, rV = existing_entry_point(lib, B, T, heads, Q, K, V, dO)
        d = {k: torch.from_numpy(v).to(dev) for k, v in dict(Q=Q, K=K, V=V, O=O, LSE=LSE, dO=dO).items()}
        g = {k: torch.full((B, T, D), float("nan"), dtype=torch.float32, device=dev) for k in ("dQ", "dK", "dV")}
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]      # cudaMemGetInfo: sees any cudaMalloc, not only torch's
        rc = lib.cvt_attention_backward_static(B, T, heads, ptr(d["Q"]), ptr(d["K"]), ptr(d["V"]), ptr(d["O"]), ptr(d["LSE"]), ptr(d["dO"]),
                                               ptr(g["dQ"]), ptr(g["dK"]), ptr(g["dV"]), None)
        _lib.check(rc, "cvt_attention_backward_static")
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info(dev)[0] >= free_before, "the allocation-free entry point allocated device memory"
        for name, ref in (("dQ", rQ), ("dK", rK), ("dV", rV)):
            got = g[name].cpu().numpy()
            assert np.isfinite(got).all(), (call, name)
            assert rel(got, ref) < 2e-3, (call, name, rel(got, ref))
            assert np.array_equal(got, ref), (call, name, rel(got, ref))


def test_static_scratch_limit_is_enforced():
    import torch
    lib = _lib.load()
    x = torch.zeros(64, device="cuda:0")
    p = C.c_void_p(x.data_ptr())
    B, T, heads = 1025, 1024, 1      # 1025 * 1024 rows > CVT_STATIC_ATTN_MAX_ROWS: refused before anything is launched
    assert B * T * heads > MAX_ROWS
    rc = lib.cvt_attention_backward_static(B, T, heads, p, p, p, p, p, p, p, p, p, None)
    assert rc == -1      # CVT_ERR_INVALID
    assert b"exceeds the static scratch" in lib.cvt_last_error()
    assert lib.cvt_attention_backward_static(1, 64, 1, None, p, p, p, p, p, p, p, p, None) < 0
