"""Diagnostic: the tcgen05 attention kernels at large grids (many CTAs per SM / several waves).  Every image carries the same Q/K/V/dO,
so every image's result must equal the fp64 reference of image 0; prints the per-image error so that failing CTAs can be located."""
import ctypes as C
import sys

import numpy as np

sys.path.insert(0, ".")
from tests.test_gpu_attention import ref_attention, rel
from transformer_stm_b200 import _lib


def run(B, T, heads, same=True, bwd=True, qs=0.25):
    lib = _lib.load()
    D = heads * 64
    rng = np.random.default_rng(T + heads)
    nb = 1 if same else B
    Q = (qs * rng.standard_normal((nb, T, D))).astype(np.float32)
    K = (2.0 * rng.standard_normal((nb, T, D))).astype(np.float32)
    V = rng.standard_normal((nb, T, D)).astype(np.float32)
    dO = rng.standard_normal((nb, T, D)).astype(np.float32)
    ref = ref_attention(Q, K, V, dO, heads)
    if same:
        Q, K, V, dO = (np.ascontiguousarray(np.repeat(a, B, 0)) for a in (Q, K, V, dO))
    O = np.zeros_like(Q); dQ = np.zeros_like(Q); dK = np.zeros_like(Q); dV = np.zeros_like(Q)
    LSE = np.zeros((B * heads, T), np.float32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    _lib.check(lib.cvt_debug_attention(B, T, heads, p(Q), p(K), p(V), p(dO) if bwd else None, p(O), p(LSE), p(dQ), p(dK), p(dV)), "dbg")
    out = {}
    for name, got, r in (("O", O, ref[0]), ("dQ", dQ, ref[2]), ("dK", dK, ref[3]), ("dV", dV, ref[4])):
        if not bwd and name != "O":
            continue
        errs = np.array([rel(got[b], r[0 if same else b]) for b in range(B)])
        out[name] = errs
    print("B=%d T=%d heads=%d:" % (B, T, heads), {k: "max %.2e, bad images %d/%d" % (v.max(), int((v > 2e-3).sum()), B) for k, v in out.items()}, flush=True)
    for k, v in out.items():
        if (v > 2e-3).any():
            print("   ", k, "bad image indices:", np.nonzero(v > 2e-3)[0][:40].tolist(), flush=True)
            b = int(np.nonzero(v > 2e-3)[0][0])
            got = {"O": O, "dQ": dQ, "dK": dK, "dV": dV}[k][b]
            r = ref[{"O": 0, "dQ": 2, "dK": 3, "dV": 4}[k]][0 if same else b]
            rows = np.abs(got - r).max(-1) / np.abs(r).max()
            print("    image %d bad rows: %d of %d, first %s" % (b, int((rows > 2e-3).sum()), T, np.nonzero(rows > 2e-3)[0][:24].tolist()), flush=True)


if __name__ == "__main__":
    import os
    qs = float(os.environ.get("QS", "0.25"))
    for B, T, h in ((32, 1024, 1), (64, 1024, 1), (64, 256, 2)):
        run(B, T, h, bwd=False, qs=qs)
