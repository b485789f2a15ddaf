import sys, numpy as np, torch
sys.path.insert(0, "/root/repo")
from tests.parity_utils import make_pair, compare_grads
from oracle import cvt_oracle as O
for prec, b in (("tf32", 1), ("tf32", 3), ("fp32", 1)):
    spec, eng, params, img, proc, y = make_pair("dw_bn", True, 5, 1, (340, 345), dims=(64, 192, 384), heads=(1, 3, 6), batch=b, precision=prec, drop_rate=0.0)
    eng.set_dropout("off")
    loss, mae = eng.train_step(img, proc, y, 1e-3)
    r = O.loss_and_grads(spec, params, img, proc, y, None, dtype=torch.float32)
    bad, report = compare_grads(eng, r["grads"], tol=1e-3)
    print(prec, b, loss, r["loss"], sorted(report, key=lambda kv: -kv[1])[:6])
    eng.close()
