"""utils.LaggedScalarReader: every pushed scalar comes back, in order, `lag` calls late (CPU path; the CUDA path differs
only in the pinned buffer and the events)."""
import torch

from picd_b200.utils import LaggedScalarReader


def test_values_come_back_in_order_one_call_late():
    r = LaggedScalarReader("cpu", lag=1)
    got = [r.push(torch.tensor([float(i)])) for i in range(5)]
    assert got == [None, 0.0, 1.0, 2.0, 3.0]
    assert r.flush() == [4.0] and r.flush() == []


def test_zero_lag_is_a_plain_read():
    r = LaggedScalarReader("cpu", lag=0)
    assert [r.push(torch.tensor(2.5)) for _ in range(3)] == [2.5, 2.5, 2.5]
