"""Batch-sharded data parallelism around libcvt_b200: one process per GPU, `torch.distributed` for the plumbing
(rendezvous, the NCCL unique id, the initial weight broadcast), the gradient all-reduce itself inside the library (cvt_comm_init).

The reference trains on one device; its only multi-GPU code is tf.distribute.MirroredStrategy in dead old_codes/BayConvT(Par)(Muti).py:17,
whose semantics are kept: the global batch is split contiguously across replicas, BatchNorm statistics are per replica, gradients are
averaged, every replica applies the same Adam step.  Inference shards by batch with no collective."""
from typing import Dict, Tuple

import numpy as np


def shard_bounds(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous partition of n samples over `world` ranks; the first n % world ranks take one extra sample."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("rank %d outside world %d" % (rank, world))
    q, r = divmod(n, world)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def _flat(params: Dict[str, np.ndarray]):
    import torch
    return torch.cat([torch.as_tensor(np.ascontiguousarray(v, dtype=np.float32)).reshape(-1) for v in params.values()])


def broadcast_weights(engine, dist, device=None, src: int = 0):
    """Every rank ends up with rank `src`'s weights (trainable and BatchNorm moving statistics)."""
    params = engine.get_params()
    flat = _flat(params)
    if device is not None:
        flat = flat.to(device)
    dist.broadcast(flat, src)
    flat = flat.cpu().numpy()
    o, new = 0, {}
    for k, v in params.items():
        new[k] = flat[o:o + v.size].reshape(v.shape)
        o += v.size
    engine.set_params(new)


def init_data_parallel(engine, dist, rank: int, world: int, device=None, peer_memory: bool = True, sync_grads: bool = False):
    """NCCL communicator inside the library (unique id made on rank 0, broadcast by torch.distributed) + identical initial weights, then
    (inside one box, world <= 8) the peer-memory mapping that replaces the per-step ncclAllReduce."""
    import torch
    uid = torch.zeros(128, dtype=torch.uint8, device=device)
    if rank == 0:
        uid.copy_(torch.tensor(list(engine.comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(uid, 0)
    engine.comm_init(bytes(uid.cpu().tolist()), rank, world)
    broadcast_weights(engine, dist, device)
    if peer_memory and 2 <= world <= 8:
        open_peer_memory(engine, dist, rank, world, device, sync_grads)


def open_peer_memory(engine, dist, rank: int, world: int, device=None, sync_grads: bool = False) -> bool:
    """Exchanges the cudaIpc handles of the ranks' workspaces (all_gather through torch.distributed) and maps them: the train step then ends
    in ONE kernel that reduce-scatters the gradients, applies Adam to the rank's slice and all-gathers the weights over NVLink peer memory
    (csrc/dp_peer.cu) instead of ncclAllReduce + a full Adam pass.  All ranks must agree on the outcome, so the success flags are reduced:
    if any rank cannot map its peers (no P2P between the devices) every rank stays on the NCCL path (returns False).  CVT_DP_NCCL=1 forces NCCL."""
    import os
    import torch
    if os.environ.get("CVT_DP_NCCL"):
        return False
    mine = torch.tensor(list(engine.comm_peer_export()), dtype=torch.uint8, device=device)
    parts = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)
    handles = b"".join(bytes(p.cpu().tolist()) for p in parts)
    ok = torch.ones(1, dtype=torch.int32, device=device)
    why = ""
    try:
        engine.comm_peer_open(handles, sync_grads)
    except Exception as e:                 # noqa: BLE001 - decided collectively below
        ok.zero_(); why = str(e)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if int(ok.item()) == 0:                # every rank takes the same path: the ranks that did map their peers let go again
        engine.comm_peer_close()
        import sys
        print("[cvt_b200] rank %d: peer memory unavailable (%s); the gradient all-reduce stays on NCCL" % (rank, why or "another rank failed"), file=sys.stderr)
        return False
    return True


def param_checksum(engine, trainable: bool = True) -> np.ndarray:
    """[sum, sum of squares, sum of |x| weighted by position] in float64 over the trainable weights (or, trainable=False, over the BatchNorm
    moving statistics): equal on two ranks iff (practically) those weights are.  The two groups are checked apart because BatchNorm is per
    replica (Keras MirroredStrategy batch statistics): every rank sees its own shard, so the moving statistics legitimately differ between
    ranks (rank 0's are the ones saved), while the trainable weights must stay bit-identical."""
    p = engine.get_params()
    state = ("moving_mean", "moving_variance")
    keep = [v for k, v in p.items() if (k.rsplit("/", 1)[-1] in state) != trainable]
    if not keep:
        return np.zeros(3)
    flat = np.concatenate([np.asarray(v, dtype=np.float64).reshape(-1) for v in keep])
    w = 1.0 + (np.arange(flat.size, dtype=np.float64) % 977) / 977.0
    return np.array([flat.sum(), (flat * flat).sum(), (np.abs(flat) * w).sum()])


def replica_divergence(engine, dist, world: int, device=None, trainable: bool = True) -> float:
    """max over ranks and checksum components of |checksum(rank) - checksum(rank 0)|; 0.0 means the replicas hold identical weights."""
    import torch
    c = torch.as_tensor(param_checksum(engine, trainable))
    if device is not None:
        c = c.to(device)
    out = [torch.zeros_like(c) for _ in range(world)]
    dist.all_gather(out, c)
    ref = out[0]
    return float(max((o - ref).abs().max().item() for o in out))


def allreduce_mean(grads: Dict[str, np.ndarray], dist, world: int, device=None) -> Dict[str, np.ndarray]:
    """Host-driven mean all-reduce of a gradient dict in one flat message (what the library does on its flat arena with NCCL)."""
    import torch
    flat = torch.cat([torch.as_tensor(np.ascontiguousarray(v, dtype=np.float64)).reshape(-1) for v in grads.values()])
    if device is not None:
        flat = flat.to(device)
    dist.all_reduce(flat)
    flat = (flat / world).cpu().numpy()
    o, out = 0, {}
    for k, v in grads.items():
        out[k] = flat[o:o + v.size].reshape(v.shape)
        o += v.size
    return out
