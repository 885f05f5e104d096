"""Bucket plan: native planner (csrc/runtime/bucket_planner.cpp through ctypes) against the Python oracle, and the
divisibility every bucket needs for ANY world size (the two-shot kernels split a bucket into `world` slices of whole
16-byte vectors; torch DDP, train.py:128, runs on any GPU count)."""
import math
import random

import pytest
import torch

from picd_b200.ops._lib import lib_available
from picd_b200.parallel.engine import ALIGN, OptimizerSpec, ParamArena, plan_buckets


def _sizes(n, seed):
    rng = random.Random(seed)
    return [(rng.choice([64, 128, 2048, 9 * 64 * 64, 512 * 2048, 1000 * 2048 + 40]) + 63) // 64 * 64 for _ in range(n)]


@pytest.mark.parametrize("world", [1, 2, 3, 5, 6, 7, 8])
def test_every_bucket_splits_evenly(world):
    unit = math.lcm(ALIGN, 4 * world)
    plan = plan_buckets(_sizes(200, world), unit=unit, native=False)
    assert plan[0][0] == 0
    for (s, e, members), nxt in zip(plan, plan[1:] + [None]):
        assert (e - s) % (4 * world) == 0 and s % unit == 0 and members
        if nxt is not None:
            assert nxt[0] == e
    assert sorted(i for _, _, m in plan for i in m) == list(range(200))


@pytest.mark.skipif(not lib_available(), reason="kernel library not built")
@pytest.mark.parametrize("world", [1, 3, 6, 8])
def test_native_planner_matches_python(world):
    unit = math.lcm(ALIGN, 4 * world)
    for seed in range(4):
        sizes = _sizes(50 + 37 * seed, seed)
        assert plan_buckets(sizes, unit=unit, native=True) == plan_buckets(sizes, unit=unit, native=False)
    assert plan_buckets([], native=True) == [] or plan_buckets([], native=True) is not None


@pytest.mark.parametrize("world", [3, 6])
def test_arena_layout_for_odd_world_sizes(world):
    """ResNet-50's parameter list at W=3 / W=6: every bucket obeys count % (4*W) == 0 (the kernel's REQUIRE) and parameter
    views do not overlap."""
    from picd_b200.models import Classifier
    model = Classifier("resnet50", 7)
    arena = ParamArena(model.parameters(), torch.device("cpu"), OptimizerSpec("sgd"), world=world, rank=0)
    for b in arena.buckets:
        assert b.numel % (4 * world) == 0
    spans = sorted((arena.offsets[i], arena.offsets[i] + p.numel()) for i, p in enumerate(arena.params))
    assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:])) and spans[-1][1] <= arena.total
    for i, p in enumerate(arena.params):
        b = arena.buckets[arena.param_bucket[i]]
        assert b.start <= arena.offsets[i] and arena.offsets[i] + p.numel() <= b.end
