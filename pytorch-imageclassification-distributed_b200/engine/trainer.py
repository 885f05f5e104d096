"""Epoch loops: ``train_epoch`` / ``val_epoch`` / ``fit`` (reference train.py:36-97,161-188).

Preserved semantics: forward -> loss (inception: ``loss + 0.4 * aux``,
train.py:48-52) -> globally averaged running loss (:59-64) -> zero_grad -> backward -> step
(:71-73); validation = per-sample top-1 x100 gathered from all ranks and averaged, padding
duplicates included (:78-97); rank 0 saves ``best_model`` on improvement and ``latest_model``
every 5th epoch (:172-188); ``MultiStepLR`` steps per epoch (:166).

Removed host syncs: the reference formats ``losses.val`` into the tqdm description every step
(an implicit ``.item()``, train.py:68) and calls ``.cpu()`` per validation sample (:89).  Here
the meter accumulates on device, the loss all-reduce is folded into the (linear) running sum
and only happens when a line is actually printed, and accuracy is computed per batch.
"""
from __future__ import annotations

import time
from typing import Callable, Optional

import numpy as np
import torch

from .. import ops
from ..parallel import dist as D
from ..utils.checkpoint import BEST, LATEST, save_checkpoint
from ..utils.meters import Accuracy, AverageMeter
from ..utils.tracing import Tracer

AUX_WEIGHT = 0.4


class WeightedCrossEntropy:
    """``nn.CrossEntropyLoss(weight=w)`` (train.py:157-158): mean = sum(w_i l_i) / sum(w_i)."""

    def __init__(self, weight: Optional[torch.Tensor] = None):
        self.weight = weight

    def __call__(self, logits, labels):
        return ops.cross_entropy(logits, labels, self.weight)


def compute_loss(outputs, labels, loss_fn):
    """Handles the inception (logits, aux_logits) pair exactly like train.py:47-56."""
    if isinstance(outputs, tuple):
        main, aux = outputs[0], outputs[1]
        return loss_fn(main, labels) + AUX_WEIGHT * loss_fn(aux, labels), main
    return loss_fn(outputs, labels), outputs


def _to_device(batch, device):
    images, labels = batch["image"], batch["label"]
    if not torch.is_tensor(labels):
        labels = torch.as_tensor(labels)
    if images.device != device:
        images = images.to(device, non_blocking=True)
        labels = labels.to(device, non_blocking=True)
    if images.dtype == torch.uint8:            # raw_u8 loader: normalise + cast on device
        images = ops.normalize_u8(images, torch.bfloat16 if device.type == "cuda" else torch.float32)
    return images, labels


def train_epoch(epoch: int, model, loss_fn, optimizer, loader, device, log_interval: int = 50,
                tracer: Optional[Tracer] = None, max_steps: Optional[int] = None,
                printer: Callable[[str], None] = print, step_fn=None, progress: bool = False) -> float:
    """``step_fn`` (an :class:`picd_b200.engine.graphed.GraphedTrainStep`) replaces the forward/backward/step sequence
    with one CUDA-graph replay per batch; the loop around it is unchanged.  ``progress``: rank 0 shows the reference's
    tqdm bar (train.py:39-42) whose description ``Epoch: e; Loss val|(avg)`` (train.py:67-68) is refreshed from the
    device-side meter every ``log_interval`` steps (every step with ``log_interval=1``, like the reference) instead of
    a printed line."""
    model.train()
    losses = AverageMeter()
    tracer = tracer or Tracer(enabled=False)
    t0, seen = time.time(), 0
    bar = None
    iterator = loader
    if progress and D.is_main():
        try:
            from tqdm import tqdm
            total = None
            try:
                total = len(loader) if max_steps is None else min(len(loader), max_steps)
            except TypeError:
                total = max_steps
            bar = tqdm(loader, total=total)
            iterator = bar
        except ImportError:
            bar = None
    for index, batch in enumerate(iterator):
        if max_steps is not None and index >= max_steps:
            break
        with tracer.range("h2d"):
            images, labels = _to_device(batch, device)
        if step_fn is not None:
            with tracer.range("graph_step"):
                loss = step_fn(images, labels)
            losses.update(loss.detach().float(), images.size(0))
        else:
            with tracer.range("forward"):
                outputs = model(images)
                loss, _ = compute_loss(outputs, labels, loss_fn)
            losses.update(loss.detach().float(), images.size(0))       # stays on device
            optimizer.zero_grad()
            with tracer.range("backward"):
                loss.backward()
            with tracer.range("step"):
                optimizer.step()
        seen += images.size(0)
        if log_interval and (index + 1) % log_interval == 0:
            # running sums are linear, so one all-reduce of (val, sum) gives the global averages
            packed = torch.stack([losses.val, losses.sum]).to(torch.float32)
            packed = D.reduce_tensor(packed, average=True)
            if D.is_main():
                val, ssum = packed.tolist()
                rate = seen * D.get_world_size() / max(time.time() - t0, 1e-9)
                if bar is not None:
                    bar.set_description(f"Epoch: {epoch}; Loss {val:.4f}|({ssum / losses.count:.4f})")
                    bar.set_postfix_str(f"{rate:.1f} img/s")
                else:
                    printer(f"Epoch: {epoch}; step {index + 1}; Loss {val:.4f}|({ssum / losses.count:.4f}); "
                            f"{rate:.1f} img/s")
    if bar is not None:
        bar.close()
    if losses.count == 0:
        return float("nan")
    avg = D.reduce_tensor(torch.as_tensor(losses.avg, dtype=torch.float32, device=device), average=True)
    return float(avg.item())


def val_epoch(epoch: int, model, loader, device, printer: Callable[[str], None] = print,
              max_steps: Optional[int] = None) -> float:
    model.eval()
    hits = []
    with torch.no_grad():
        for index, batch in enumerate(loader):
            if max_steps is not None and index >= max_steps:
                break
            images, labels = _to_device(batch, device)
            outputs = model(images)
            if isinstance(outputs, tuple):
                outputs = outputs[0]
            hits.append(Accuracy(outputs.float(), labels))        # [B] int64, on device
    local = (torch.cat(hits).cpu().numpy() * 100) if hits else np.zeros((0,), dtype=np.int64)
    gathered = D.all_gather(local)                                 # rank-ordered list (ddp_utils.py:16-56)
    comb = np.concatenate(gathered)
    acc = float(comb.mean()) if comb.size else 0.0
    if D.is_main():
        printer(f"Validation Accuracy {acc}")
    return acc


def fit(model, loss_fn, optimizer, scheduler, train_loader, val_loader, device, *, name: str,
        epochs: int = 100, start_epoch: int = 0, best_score: float = 0.0, ckpt_dir: str = "dtmodel/cp",
        save_every: int = 5, log_interval: int = 50, tracer: Optional[Tracer] = None,
        max_steps: Optional[int] = None, printer: Callable[[str], None] = print, cuda_graph: bool = False,
        progress: bool = False):
    """The reference's epoch driver (train.py:161-188) with start_epoch honoured (fix D10).  ``cuda_graph`` replays the
    training step from a captured CUDA graph where the model allows it (see engine/graphed.py)."""
    history = []
    step_fn = None
    if cuda_graph and device.type == "cuda":
        from .graphed import GraphedTrainStep
        step_fn = GraphedTrainStep(model, loss_fn, optimizer)
        if not step_fn.enabled:
            step_fn = None
    for epoch in range(start_epoch, epochs):
        if hasattr(train_loader, "sampler") and hasattr(train_loader.sampler, "set_epoch"):
            train_loader.sampler.set_epoch(epoch)                 # train.py:164
        elif hasattr(train_loader, "set_epoch"):
            train_loader.set_epoch(epoch)
        loss = train_epoch(epoch, model, loss_fn, optimizer, train_loader, device, log_interval,
                           tracer, max_steps, printer, step_fn, progress)
        scheduler.step()                                          # train.py:166
        if hasattr(model, "sync_masters"):
            model.sync_masters()      # collective: sharded fp32 masters -> full copies before anything reads state_dict()
        acc = val_epoch(epoch, model, val_loader, device, printer, max_steps)
        # the optimizer state is sharded by slice owner on the peer-memory engine: assembling it is a collective, so
        # every rank takes part even though only rank 0 writes the file
        opt_state = optimizer.state_dict() if epoch % save_every == 0 else None
        if D.is_main():
            if acc > best_score:
                printer(f"Model improved to {acc} so storing checkpoint")
                best_score = acc
                save_checkpoint(model, epoch, best_score, ckpt_dir, name, BEST)
            if epoch % save_every == 0:
                save_checkpoint(model, epoch, best_score, ckpt_dir, name, LATEST,
                                extra={"optimizer": opt_state,
                                       "scheduler": scheduler.state_dict()})
        best_score = float(D.broadcast_object(best_score, 0))
        history.append({"epoch": epoch, "loss": loss, "val_accuracy": acc})
    return history, best_score
