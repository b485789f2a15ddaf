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


def init_data_parallel(engine, dist, rank: int, world: int, device=None):
    """NCCL communicator inside the library (unique id made on rank 0, broadcast by torch.distributed) + identical initial weights."""
    import torch
    uid = torch.zeros(128, dtype=torch.uint8, device=device)
    if rank == 0:
        uid.copy_(torch.tensor(list(engine.comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(uid, 0)
    engine.comm_init(bytes(uid.cpu().tolist()), rank, world)
    broadcast_weights(engine, dist, device)


def param_checksum(engine) -> np.ndarray:
    """[sum, sum of squares, sum of |x| weighted by position] over all weights in float64: equal on two ranks iff (practically) the weights are."""
    p = engine.get_params()
    flat = np.concatenate([np.asarray(v, dtype=np.float64).reshape(-1) for v in p.values()])
    w = 1.0 + (np.arange(flat.size, dtype=np.float64) % 977) / 977.0
    return np.array([flat.sum(), (flat * flat).sum(), (np.abs(flat) * w).sum()])


def replica_divergence(engine, dist, world: int, device=None) -> float:
    """max over ranks and checksum components of |checksum(rank) - checksum(rank 0)|; 0.0 means the replicas hold identical weights."""
    import torch
    c = torch.as_tensor(param_checksum(engine))
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
