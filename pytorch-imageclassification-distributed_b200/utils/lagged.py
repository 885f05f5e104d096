"""Reading a per-step device scalar on the host without stalling the device.

The reference formats ``losses.val`` into its progress bar every step (train.py:67-68): an implicit ``.item()``, i.e. the
host waits for the step to finish, and only THEN enqueues the next step -- the GPU idles for the host's per-step work
(H2D staging, graph replay call, ...).  ``LaggedScalarReader`` keeps the read every step but takes the value one step
late: the device-to-host copy of step i is enqueued behind step i on the compute stream (pinned destination, so it is
truly asynchronous) and consumed while step i+1 is already running.  ``flush()`` returns the values still in flight."""
from __future__ import annotations

from collections import deque
from typing import Deque, List, Optional, Tuple

import torch


class LaggedScalarReader:
    def __init__(self, device, lag: int = 1, dtype=torch.float32):
        self.device = torch.device(device)
        self.lag = max(int(lag), 0)
        self.cuda = self.device.type == "cuda"
        n = self.lag + 2
        self._slots = [torch.zeros(1, dtype=dtype, pin_memory=self.cuda) for _ in range(n)]
        self._events = [torch.cuda.Event() if self.cuda else None for _ in range(n)]
        self._pending: Deque[int] = deque()
        self._next = 0

    def _read(self, slot: int) -> float:
        if self.cuda:
            self._events[slot].synchronize()
        return float(self._slots[slot][0])

    def push(self, value: torch.Tensor) -> Optional[float]:
        """Enqueue the device-to-host copy of ``value`` (a one-element tensor) and return the scalar pushed ``lag`` calls
        ago, or None while the pipeline fills."""
        slot = self._next
        self._next = (self._next + 1) % len(self._slots)
        self._slots[slot].copy_(value.detach().reshape(1), non_blocking=True)
        if self.cuda:
            self._events[slot].record(torch.cuda.current_stream(self.device))
        self._pending.append(slot)
        if len(self._pending) > self.lag:
            return self._read(self._pending.popleft())
        return None

    def flush(self) -> List[float]:
        out = []
        while self._pending:
            out.append(self._read(self._pending.popleft()))
        return out
