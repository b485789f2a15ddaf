"""Prints the clock64 timeline of one CTA of the dQ attention kernel (CVT_ATTN_TRACE=1): python tools/attn_trace.py"""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["CVT_ATTN_TRACE"] = "1"
from transformer_stm_b200 import _lib
lib = _lib.load()
B, T, heads = 64, 1024, 1
D = heads * 64
rng = np.random.default_rng(0)
Q = (0.25 * rng.standard_normal((B, T, D))).astype(np.float32); K = (2 * rng.standard_normal((B, T, D))).astype(np.float32)
V = rng.standard_normal((B, T, D)).astype(np.float32); dO = rng.standard_normal((B, T, D)).astype(np.float32)
O = np.zeros_like(Q); dQ = np.zeros_like(Q); dK = np.zeros_like(Q); dV = np.zeros_like(Q); LSE = np.zeros((B * heads, T), np.float32)
p = lambda a: a.ctypes.data_as(C.c_void_p)
_lib.check(lib.cvt_debug_attention(B, T, heads, p(Q), p(K), p(V), p(dO), p(O), p(LSE), p(dQ), p(dK), p(dV)), "cvt_debug_attention")
