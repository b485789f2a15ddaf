"""clock64 timeline of CTA 0 of the TMA GEMM kernel on a few reference shapes (CVT_GEMM_TRACE=1): python tools/gemm_trace.py"""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["CVT_GEMM_TRACE"] = "1"
from transformer_stm_b200 import _lib
lib = _lib.load()
p = lambda a: a.ctypes.data_as(C.c_void_p)
for M, N, K in ((65536, 64, 64), (65536, 256, 64), (65536, 64, 256)):
    rng = np.random.default_rng(0)
    A = rng.standard_normal((M, K)).astype(np.float32); B = rng.standard_normal((N, K)).astype(np.float32); Cc = np.zeros((M, N), np.float32)
    _lib.check(lib.cvt_debug_gemm(0, 1, M, N, K, 1, 4, p(A), p(B), p(Cc)), "cvt_debug_gemm")
    sys.stdout.flush()
