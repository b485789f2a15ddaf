#!/usr/bin/env python
"""Data-parallel correctness on real GPUs: run under torchrun with N >= 2 ranks.

  torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/scripts/dp_check.py [--precision tf32]

1. one cvt_train_step with lr = 0 on a global batch split contiguously over the ranks: the gradient arena every rank holds after the
   NCCL all-reduce must equal the MEAN of the per-shard oracle gradients (per-replica BatchNorm statistics, MirroredStrategy semantics);
2. five real steps: the weight checksums of all ranks must stay identical (replica_divergence == 0.0).
Prints one JSON line on rank 0."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--precision", default="tf32")
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--hw", type=int, default=64)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    import transformer_stm_b200 as T
    from oracle import cvt_oracle as O          # checker only
    from transformer_stm_b200 import dp
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    B = a.batch
    spec = O.ref_spec(image_h=a.hw, image_w=a.hw)
    spec.drop_rate = 0.0
    cfg = T.CvtConfig.reference(image_h=a.hw, image_w=a.hw, precision=a.precision, max_batch=B, device=local, drop_rate=0.0)
    eng = T.CvtEngine(cfg)
    eng.set_params(O.init_params(spec, 500 + rank, "parity"))          # ranks start different; init_data_parallel broadcasts rank 0's
    dp.init_data_parallel(eng, dist, rank, world, dev, sync_grads=True)      # sync_grads: every rank's arena ends up holding the reduced gradient
    params = {k: v.astype(np.float64) for k, v in eng.get_params().items()}
    img, proc, y = O.synthetic_batch(spec, B * world, seed=11)
    lo, hi = dp.shard_bounds(B * world, rank, world)
    d = [torch.tensor(x[lo:hi], device=dev) for x in (img, proc, y)]
    torch.cuda.synchronize()
    eng.set_dropout("off")
    eng.train_step_dev(d[0], d[1], d[2], B, 0.0)
    eng.synchronize()
    # the library keeps the SUM over ranks in its gradient arena (the 1 / world factor is folded into the Adam update)
    got = {k: v / world for k, v in eng.get_grads().items()}
    out = {"world": world, "path": "nccl" if os.environ.get("CVT_DP_NCCL") else "peer-memory kernel", "precision": a.precision, "per_rank_batch": B, "image": a.hw}
    if rank == 0:
        shards = [O.loss_and_grads(spec, params, img[r * B:(r + 1) * B], proc[r * B:(r + 1) * B], y[r * B:(r + 1) * B], None)["grads"] for r in range(world)]
        mean = {k: sum(s[k] for s in shards) / world for k in shards[0]}
        gmax = max(float(np.abs(v).max()) for v in mean.values())
        errs = {k: float(np.abs(got[k] - v).max()) / max(float(np.abs(v).max()), 1e-3 * gmax) for k, v in mean.items()}
        worst = max(errs, key=errs.get)
        single = O.loss_and_grads(spec, params, img[:B], proc[:B], y[:B], None)["grads"]
        out.update(max_err_vs_mean_of_shard_gradients=errs[worst], worst_tensor=worst,
                   err_if_it_were_rank0_shard_only=max(float(np.abs(got[k] - v).max()) / max(float(np.abs(v).max()), 1e-3 * gmax) for k, v in single.items()))
    for _ in range(5):
        eng.train_step_dev(d[0], d[1], d[2], B, 1e-3)
    eng.synchronize()
    div = dp.replica_divergence(eng, dist, world, dev)
    div_bn = dp.replica_divergence(eng, dist, world, dev, trainable=False)
    if rank == 0:
        out["replica_divergence_after_5_steps"] = div
        out["bn_moving_stat_divergence"] = div_bn       # per-replica BatchNorm: expected to be non-zero when the shards differ
        tol = 1e-3 if a.precision == "fp32" else 2e-2
        out["ok"] = bool(out["max_err_vs_mean_of_shard_gradients"] <= tol and div == 0.0)
        print(json.dumps(out))
    eng.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
