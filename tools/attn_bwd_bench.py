"""Times the flash-attention backward kernels alone (cvt_attention_backward_static: attn_dd + dK/dV + dQ, one stream, serial) at the
stage shapes of the benched configuration and checks them against a float64 numpy reference on a few (image, head) pairs.
  python tools/attn_bwd_bench.py                 # the default kernel choice
  CVT_ATTN_BWD_TILE=64 python tools/attn_bwd_bench.py   # force the 64-wide score tiles (128: force the wide ones)"""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from transformer_stm_b200 import _lib

lib = _lib.load()
dev = torch.device("cuda:0")
ptr = lambda t: C.c_void_p(t.data_ptr())


def to_tf32(x):
    b = x.contiguous().view(torch.int32)
    return ((b + 0x1000) & ~0x1FFF).view(torch.float32)


def reference(Q, K, V, dO):
    """float64, one (image, head): Q pre-scaled"""
    S = Q @ K.T
    m = S.max(1, keepdims=True)
    P = np.exp(S - m); P /= P.sum(1, keepdims=True)
    O = P @ V
    dV = P.T @ dO
    dP = dO @ V.T
    dS = P * (dP - (dO * O).sum(1, keepdims=True))
    return O, m[:, 0] + np.log(np.exp(S - m).sum(1)), dS @ K, dS.T @ Q, dV


SHAPES = [(64, 1024, 1), (64, 256, 2), (64, 65, 4), (8, 7395, 1), (3, 485, 2)]
if "--only" in sys.argv:
    SHAPES = [SHAPES[int(sys.argv[sys.argv.index("--only") + 1])]]
for (B, T, heads) in SHAPES:
    D = heads * 64
    g = torch.Generator(device=dev); g.manual_seed(B * 1000 + T)
    Q = to_tf32(0.25 * torch.randn((B, T, D), device=dev, generator=g)); K = to_tf32(2.0 * torch.randn((B, T, D), device=dev, generator=g))
    V = to_tf32(torch.randn((B, T, D), device=dev, generator=g)); dO = to_tf32(torch.randn((B, T, D), device=dev, generator=g))
    # O and LSE from a float64 torch forward on the device (per head)
    Qh = Q.view(B, T, heads, 64).permute(0, 2, 1, 3).double(); Kh = K.view(B, T, heads, 64).permute(0, 2, 1, 3).double(); Vh = V.view(B, T, heads, 64).permute(0, 2, 1, 3).double()
    if B * heads * T * T * 8 <= (8 << 30):
        S = Qh @ Kh.transpose(-1, -2)
        LSE = torch.logsumexp(S, -1)
        O = (torch.softmax(S, -1) @ Vh)
        del S
    else:
        LSE = torch.empty((B, heads, T), device=dev, dtype=torch.float64); O = torch.empty((B, heads, T, 64), device=dev, dtype=torch.float64)
        for b in range(B):
            S = Qh[b] @ Kh[b].transpose(-1, -2); LSE[b] = torch.logsumexp(S, -1); O[b] = torch.softmax(S, -1) @ Vh[b]
    O32 = O.permute(0, 2, 1, 3).reshape(B, T, D).float().contiguous(); LSE32 = LSE.reshape(B * heads, T).float().contiguous()
    dQ, dK, dV = (torch.full((B, T, D), float("nan"), device=dev) for _ in range(3))
    call = lambda: _lib.check(lib.cvt_attention_backward_static(B, T, heads, ptr(Q), ptr(K), ptr(V), ptr(O32), ptr(LSE32), ptr(dO), ptr(dQ), ptr(dK), ptr(dV), None), "static")
    for _ in range(3):
        call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 10
    e0.record()
    for _ in range(n):
        call()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / n
    flops = 5 * 2.0 * B * heads * T * T * 64
    # check (image, head) pairs against float64 numpy
    worst = {}
    for (b, h) in {(0, 0), (B - 1, heads - 1)}:
        sl = slice(h * 64, h * 64 + 64)
        q, k, v, do = (x[b, :, sl].double().cpu().numpy() for x in (Q, K, V, dO))
        _, _, rq, rk, rv = reference(q, k, v, do)
        for name, got, ref in (("dQ", dQ, rq * 0.125), ("dK", dK, rk), ("dV", dV, rv)):
            gg = got[b, :, sl].double().cpu().numpy()
            err = float(np.abs(gg - ref).max() / np.abs(ref).max()) if np.isfinite(gg).all() else float("nan")
            worst[name] = max(worst.get(name, 0.0), err) if err == err else float("nan")
    print("B=%d T=%d heads=%d: backward %.1f us (%.0f TFLOP/s on 5 products) | max err vs float64: %s" % (B, T, heads, us, flops / us / 1e6, {k: "%.2e" % v for k, v in worst.items()}), flush=True)
