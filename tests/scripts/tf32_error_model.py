"""TEST INFRASTRUCTURE (CPU only): which contractions carry the TF32 error?  The oracle's forward with emulated TF32 operand rounding (fwd and bwd) per op class."""
import sys, math, re
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from oracle import cvt_oracle as O
import inspect

def r(x):
    if x.dtype != torch.float32: return x
    i = x.contiguous().view(torch.int32)
    return ((i + 0x1000) & ~0x1fff).view(torch.float32)

class MM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, on):
        ctx.on = on; ctx.save_for_backward(a, b)
        return (r(a) @ r(b)) if on else a @ b
    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        if ctx.on:
            return r(g) @ r(b).transpose(-1, -2), r(a).transpose(-1, -2) @ r(g), None
        return g @ b.transpose(-1, -2), a.transpose(-1, -2) @ g, None

FLAGS = dict(dense=False, attn=False, conv=False)
def mm(a, b, kind):
    on = FLAGS[kind]
    # broadcast b over batch dims of a
    if a.dim() > 2 and b.dim() == 2:
        sh = a.shape; out = MM.apply(a.reshape(-1, sh[-1]), b, on); return out.reshape(*sh[:-1], b.shape[-1])
    return MM.apply(a, b, on)

src = inspect.getsource(O.forward_torch)
# dense products:  X @ P[...]
src = re.sub(r"(\w+(?:\[j\])?) @ (P\[f?\"[^\"]+\"\])", r"mm(\1, \2, 'dense')", src)
src = src.replace('torch.einsum("btd,dnh->btnh", z, P[f"{t}/attn/mha/{nm}/kernel"])', 'mm(z, P[f"{t}/attn/mha/{nm}/kernel"].reshape(d, d), "dense").reshape(bsz, -1, nh, hd)')
src = src.replace('torch.einsum("btnh,nhd->btd", O, P[f"{t}/attn/mha/attention_output/kernel"])', 'mm(O.reshape(bsz, -1, d), P[f"{t}/attn/mha/attention_output/kernel"].reshape(d, d), "dense")')
src = src.replace('S = torch.einsum("btnh,bsnh->bnts", Q, K)', 'S = mm(Q.permute(0, 2, 1, 3), K.permute(0, 2, 3, 1), "attn")')
src = src.replace('O = torch.einsum("bnts,bsnh->btnh", Pm, V)', 'O = mm(Pm, V.permute(0, 2, 1, 3), "attn").permute(0, 2, 1, 3)')
src = src.replace("def forward_torch(", "def forward_sim(")
ns = dict(O.__dict__); ns["mm"] = mm
exec(src, ns)
forward_sim = ns["forward_sim"]
assert "mm(" in src and src.count("mm(") >= 10, src.count("mm(")

def run(flags, dtype, spec, params, img, proc, y):
    FLAGS.update(flags)
    P = O._tp(params, dtype, True)
    pred, taps, _ = forward_sim(spec, P, torch.tensor(img, dtype=dtype), torch.tensor(proc, dtype=dtype), True, None)
    loss = ((pred - torch.tensor(y, dtype=dtype).reshape(pred.shape)) ** 2).mean()
    loss.backward()
    return pred.detach().double().numpy(), {k: v.grad.double().numpy() for k, v in P.items() if v.grad is not None}

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
spec = O.ref_spec(image_h=128, image_w=128); spec.drop_rate = 0.0
params = O.init_params(spec, 1234, "parity")
img, proc, y = O.synthetic_batch(spec, B, 7)
p64, g64 = run(dict(dense=False, attn=False, conv=False), torch.float64, spec, params, img, proc, y)
gmax = max(np.abs(v).max() for v in g64.values())
for name, fl in (("fp32 all", dict(dense=False, attn=False)), ("tf32 dense only", dict(dense=True, attn=False)), ("tf32 attention only", dict(dense=False, attn=True)), ("tf32 all", dict(dense=True, attn=True))):
    p, g = run(dict(fl, conv=False), torch.float32, spec, params, img, proc, y)
    perr = np.abs(p - p64).max() / np.abs(p64).max()
    per = {k: np.abs(g[k] - v).max() / max(np.abs(v).max(), 1e-3 * gmax) for k, v in g64.items()}
    num = sum(((g[k] - v) ** 2).sum() for k, v in g64.items()); den = sum((v ** 2).sum() for v in g64.values())
    vals = sorted(per.values())
    print("%-22s pred %.2e  grad worst %.2e  median %.2e  L2 %.2e  (>1e-3: %d/%d)" % (name, perr, vals[-1], vals[len(vals)//2], (num/den)**0.5, sum(v > 1e-3 for v in vals), len(vals)), flush=True)
