"""Whole-step CUDA graph: forward, loss, backward and the bucketed all-reduce + optimizer update of one training
step are recorded once and replayed for every later batch of the same shape.

The reference's step is launch-bound in places (train.py:41-73 issues ~2000 small kernels per ResNet-50 step through
the Python dispatcher); this framework launches ~400 larger kernels per step, and replaying them from a graph removes
the remaining host-side launch gaps (about 1.3 ms of a 24 ms ResNet-50 step, far more for EfficientNet).

What makes the step replayable:
  * all kernels run on the stream that is current when the op is called (the ctypes bindings pass
    ``torch.cuda.current_stream()``), tensor maps travel by value in the launch parameters, activations come from the
    graph's private memory pool, gradients / weights live in the static parameter arena;
  * the only step-varying scalars -- learning rate, Adam bias corrections and the peer-flag epoch of the fused
    all-reduce kernels -- are read from a small device array (``picd_fused_allreduce_update_dyn``) whose step-dependent
    words the graph's first kernel advances itself (``picd_dyn_advance``); the host writes it only when the learning
    rate changes or an eager step ran in between, so a replay involves no host-to-device copy besides the batch;
  * DDP's per-forward buffer broadcast rides in the last bucket kernel (peer stores), so no NCCL call surrounds a replay;
  * BatchNorm ``num_batches_tracked`` and the running statistics are updated by kernels, not by the host;
  * SyncBatchNorm's statistics exchanges (double-buffered peer-memory slots, epoch-numbered flags) are recorded on
    slots of their own with the capture-time epoch; the kernels add a replay counter read from the same device array.

Caller contract: do not keep a loss / output tensor WITH its autograd graph from an eager step alive across the call
that records the graph (this class only hands out detached tensors).

Not capturable (the step then simply runs eagerly): ``no_sync`` accumulation, batches whose shape differs from the
captured one (the ragged last batch).
"""
from __future__ import annotations

from typing import Callable, Optional

import torch

from .trainer import compute_loss


class GraphedTrainStep:
    """``step(images, labels) -> loss`` with the same effect as

        out = model(images); loss, _ = compute_loss(out, labels, loss_fn)
        optimizer.zero_grad(); loss.backward(); optimizer.step()

    The first ``warmup`` calls run eagerly (they are real training steps), the next call records the graph, every later
    call copies the batch into the graph's static input buffers and replays.  ``outputs`` holds the logits of the last
    replay (a static buffer: clone it to keep it)."""

    def __init__(self, model, loss_fn: Callable, optimizer, warmup: int = 3, enabled: bool = True):
        self.model, self.loss_fn, self.optimizer = model, loss_fn, optimizer
        self.warmup = max(int(warmup), 1)
        self.calls = 0
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.key = None
        self.launches_per_replay = 0
        self.replays = 0
        self.outputs = None
        self.enabled = enabled and self._capturable()

    def _capturable(self) -> bool:
        m = self.model
        if not hasattr(m, "arena") or m.device.type != "cuda":
            return False
        return True

    def _eager(self, images, labels):
        out = self.model(images)
        loss, main = compute_loss(out, labels, self.loss_fn)
        self.optimizer.zero_grad()
        loss.backward()
        self.optimizer.step()
        # detached: a live autograd graph from an eager step would keep its AccumulateGrad nodes (and the stream they were
        # created on) alive, and the capture would then see gradient hooks running on an uncaptured stream
        self.outputs = main.detach()
        return loss.detach()

    def _capture(self, images, labels):
        from ..ops._lib import require_lib
        import ctypes
        m = self.model
        red = m._impl
        red.enable_dyn()
        self.outputs = None
        self.static_images = images.clone()
        self.static_labels = labels.clone()
        lib = require_lib()
        lib.picd_launch_count.restype = ctypes.c_ulonglong
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        step0, count0 = m.step_index, self.optimizer.step_count
        l0 = lib.picd_launch_count()
        from ..ops import native
        m.capturing = True
        native.set_step_word(red.dyn[3:4])          # dropout masks follow the device-side step counter
        red.comm.begin_capture(red.dyn)             # SyncBN collectives: slots of their own, epoch = capture epoch + delta
        try:
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                red.advance_dyn()                   # step counter / flag epoch / bias corrections move on the device
                self.static_loss = self._eager(self.static_images, self.static_labels)
                self.static_outputs = self.outputs
        finally:
            m.capturing = False
            native.set_step_word(None)
            red.comm.end_capture()
        self.launches_per_replay = int(lib.picd_launch_count() - l0)
        # recording executes nothing: undo the host-side step bookkeeping it advanced
        m.step_index, self.optimizer.step_count = step0, count0
        self.graph = g
        self.key = (tuple(images.shape), images.dtype, tuple(labels.shape), labels.dtype)

    def __call__(self, images, labels):
        self.calls += 1
        if not self.enabled or not self.model.training or self.calls <= self.warmup:
            return self._eager(images, labels)
        key = (tuple(images.shape), images.dtype, tuple(labels.shape), labels.dtype)
        if self.graph is None:
            self._capture(images, labels)
        if key != self.key:
            return self._eager(images, labels)
        m = self.model
        self.static_images.copy_(images, non_blocking=True)
        self.static_labels.copy_(labels, non_blocking=True)
        m._impl.sync_dyn(m.step_index + 1, self.replays)     # no H2D copy unless the host's view of (lr, step) moved
        self.graph.replay()
        m._masters_stale = True
        m.step_index += 1
        self.optimizer.step_count += 1
        self.replays += 1
        self.outputs = self.static_outputs
        return self.static_loss.clone()
