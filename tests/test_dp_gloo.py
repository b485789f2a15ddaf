"""World-size-2 `gloo` test (CPU) of the data-parallel host logic in transformer_stm_b200/dp.py: contiguous batch shards,
weight broadcast, replica-divergence checksum, and the definition the GPU path is held to - the all-reduced gradient is the MEAN of the
per-shard gradients (each shard with its own BatchNorm statistics, Keras MirroredStrategy semantics), which for a model without BatchNorm
equals the single-process gradient of the global batch."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import cvt_oracle as O


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import transformer_stm_b200 as T
        from tests.fake_engine import OracleEngine
        from transformer_stm_b200 import dp
        res = {}
        for method in ("linear", "dw_bn"):
            cfg = T.CvtConfig.reference(method, True, 5, 1, image_h=32, image_w=32, max_batch=4, drop_rate=0.0)
            eng = OracleEngine(cfg)
            eng.set_params(O.init_params(eng.spec, 100 + rank, "parity"))          # ranks start DIFFERENT
            before = dp.replica_divergence(eng, dist, world)
            dp.broadcast_weights(eng, dist)
            after = dp.replica_divergence(eng, dist, world)
            n = 6
            img, proc, y = O.synthetic_batch(eng.spec, n, seed=3)
            lo, hi = dp.shard_bounds(n, rank, world)
            g = O.loss_and_grads(eng.spec, eng.params, img[lo:hi], proc[lo:hi], y[lo:hi], None)["grads"]
            red = dp.allreduce_mean(g, dist, world)
            res[method] = dict(before=before, after=after, lo=lo, hi=hi, red={k: v for k, v in red.items()})
        np.save(os.path.join(out_dir, "rank%d.npy" % rank), res, allow_pickle=True)
    finally:
        dist.destroy_process_group()


def test_shard_bounds_partition():
    from transformer_stm_b200 import dp
    for n in (0, 1, 7, 64, 513):
        for world in (1, 2, 3, 8):
            b = [dp.shard_bounds(n, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n and all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        dp.shard_bounds(4, 2, 2)


def test_world2_gloo_broadcast_and_gradient_mean(tmp_path):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r0 = np.load(tmp_path / "rank0.npy", allow_pickle=True).item()
    r1 = np.load(tmp_path / "rank1.npy", allow_pickle=True).item()
    import transformer_stm_b200 as T
    from tests.fake_engine import OracleEngine
    for method in ("linear", "dw_bn"):
        a, b = r0[method], r1[method]
        assert a["before"] > 0 and a["after"] == 0.0 and b["after"] == 0.0            # broadcast made the replicas identical
        assert (a["lo"], a["hi"], b["lo"], b["hi"]) == (0, 3, 3, 6)
        assert all(np.array_equal(a["red"][k], b["red"][k]) for k in a["red"])        # both ranks hold the same reduced gradient
        cfg = T.CvtConfig.reference(method, True, 5, 1, image_h=32, image_w=32, max_batch=4, drop_rate=0.0)
        eng = OracleEngine(cfg)
        eng.set_params(O.init_params(eng.spec, 100, "parity"))                        # rank 0's weights
        img, proc, y = O.synthetic_batch(eng.spec, 6, seed=3)
        g0 = O.loss_and_grads(eng.spec, eng.params, img[:3], proc[:3], y[:3], None)["grads"]
        g1 = O.loss_and_grads(eng.spec, eng.params, img[3:], proc[3:], y[3:], None)["grads"]
        for k in g0:                                                                   # all-reduce == mean of per-shard gradients
            np.testing.assert_allclose(a["red"][k], 0.5 * (g0[k] + g1[k]), rtol=1e-9, atol=1e-11, err_msg=k)
        if method == "linear":                                                         # no BatchNorm: == gradient of the global batch
            gg = O.loss_and_grads(eng.spec, eng.params, img, proc, y, None)["grads"]
            for k in gg:
                np.testing.assert_allclose(a["red"][k], gg[k], rtol=1e-9, atol=1e-11, err_msg=k)


class _PeerStub:
    """Records the peer-memory calls dp.open_peer_memory makes; rank `fail_rank` cannot map its peers."""

    def __init__(self, rank, fail_rank):
        self.rank, self.fail_rank, self.opened, self.closed, self.handles = rank, fail_rank, False, False, None

    def comm_peer_export(self):
        return bytes([self.rank + 1]) * 64

    def comm_peer_open(self, handles, sync_grads=False):
        self.handles = handles
        if self.rank == self.fail_rank:
            raise RuntimeError("cudaIpcOpenMemHandle: peer access is not supported between these two devices")
        self.opened = True

    def comm_peer_close(self):
        self.opened, self.closed = False, True


def _peer_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from transformer_stm_b200 import dp
        res = {}
        for name, fail in (("all_ok", -1), ("rank1_fails", 1)):
            eng = _PeerStub(rank, fail)
            ok = dp.open_peer_memory(eng, dist, rank, world)
            res[name] = dict(ok=ok, opened=eng.opened, closed=eng.closed, handles=eng.handles)
        np.save(os.path.join(out_dir, "peer%d.npy" % rank), res, allow_pickle=True)
    finally:
        dist.destroy_process_group()


def test_world2_gloo_peer_memory_handshake_is_collective(tmp_path):
    """dp.open_peer_memory: the cudaIpc handles are all-gathered in rank order, and when ONE rank cannot map its peers EVERY rank lets go
    again and stays on the NCCL path (a mixed job would deadlock: the peer-memory kernel waits for flags the NCCL ranks never write)."""
    world, port = 2, _free_port()
    mp.spawn(_peer_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r = [np.load(tmp_path / ("peer%d.npy" % i), allow_pickle=True).item() for i in range(world)]
    want = bytes([1]) * 64 + bytes([2]) * 64
    for i in range(world):
        assert r[i]["all_ok"]["ok"] is True and r[i]["all_ok"]["opened"] and not r[i]["all_ok"]["closed"]
        assert r[i]["all_ok"]["handles"] == want
        assert r[i]["rank1_fails"]["ok"] is False and not r[i]["rank1_fails"]["opened"] and r[i]["rank1_fails"]["closed"]
