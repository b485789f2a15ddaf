"""bench.py's timing loop on fakes (no GPU): a data-parallel train step is collective, so every rank has to issue the SAME number of steps
whether or not it carries the clock sampler (only rank 0 does).  A warm-up loop that ran on the sampling rank alone once let rank 0 run ahead
at 8 GPUs: its peer-memory tail kernel trapped after the bounded spin and the other ranks hung."""
import importlib.util
import os
import types

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_bench():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class FakeEvent:
    def __init__(self, enable_timing=False):
        pass

    def record(self, stream=None):
        pass

    def elapsed_time(self, other):
        return 1.5


class FakeTensor:
    def __init__(self, v):
        self.v = v

    def item(self):
        return self.v[0]


class FakeStream:
    def synchronize(self):
        pass


def fake_torch():
    cuda = types.SimpleNamespace(Event=FakeEvent, synchronize=lambda: None, current_stream=lambda: FakeStream())
    return types.SimpleNamespace(cuda=cuda, tensor=lambda v, device=None: FakeTensor(v))


class FakeDist:
    ReduceOp = types.SimpleNamespace(MAX="max")

    def barrier(self):
        pass

    def all_reduce(self, t, op=None):
        pass


class FakeEngine:
    def __init__(self):
        self.train_steps = 0
        self.forwards = 0

    def set_dropout(self, mode, seed=0):
        pass

    def train_step_dev(self, *a):
        self.train_steps += 1

    def forward_dev(self, *a):
        self.forwards += 1

    def synchronize(self):
        pass

    def kernel_launches(self):
        return 96 * self.train_steps + 27 * self.forwards


class FakeSampler:
    """A sample 'arrives' every few polls, like nvidia-smi -lms does while the steps run."""

    def __init__(self):
        self._polls = 0
        self._list = [["1965", "1965"]]
        self.started = False

    @property
    def samples(self):
        self._polls += 1
        if self._polls % 7 == 0:
            self._list.append(["1965", "1965"])
        return self._list

    def start(self):
        self.started = True

    def wait_first(self, timeout_s):
        pass


def make_bench(mod, world, rank):
    b = object.__new__(mod.Bench)
    b.torch, b.eng, b.B, b.NB = fake_torch(), FakeEngine(), 64, 4
    b.imgs = b.procs = b.ys = [None] * 4
    b.flush = types.SimpleNamespace(zero_=lambda: None)
    b.stream = None
    b.world, b.rank = world, rank
    b.dist = FakeDist() if world > 1 else None
    return b


@pytest.mark.parametrize("world", [2, 8])
def test_every_rank_issues_the_same_number_of_train_steps(world):
    mod = load_bench()
    counts = []
    for rank in range(world):
        b = make_bench(mod, world, rank)
        sampler = FakeSampler() if rank == 0 else None          # run_config: only rank 0 samples clocks
        r = b.timed(True, steps=20, warmup=5, sampler=sampler)
        counts.append(b.eng.train_steps)
        assert r["ms_per_step"] == pytest.approx(1.5)
        if sampler:
            assert sampler.started
    assert len(set(counts)) == 1, counts
    assert counts[0] == 2 * 4 + 5 + 20                            # graph priming of the four buffer sets, W warm-up steps, K timed steps


def test_single_gpu_run_counts_only_the_timed_steps():
    mod = load_bench()
    b = make_bench(mod, 1, 0)
    s = FakeSampler()
    r = b.timed(True, steps=20, warmup=5, sampler=s)
    assert s.started and b.eng.train_steps == 2 * 4 + 5 + 20
    assert r["value"] == pytest.approx(64 * 20 / (20 * 1.5e-3))
    assert r["launches"] == 96 * 20                                # only the timed steps are counted
    assert r["per_step_ms"]["median"] == pytest.approx(1.5) and len(r["per_step_ms"]["all"]) == 20


def test_inference_ranks_are_independent():
    mod = load_bench()
    b = make_bench(mod, 8, 0)
    r = b.timed(False, steps=10, warmup=3, sampler=FakeSampler())
    assert b.eng.train_steps == 0 and b.eng.forwards == 2 * 4 + 3 + 10
    assert r["launches"] == 27 * 10
