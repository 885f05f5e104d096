#!/usr/bin/env python
"""Multi-GPU correctness + bandwidth checks for the NVLink peer-memory engine.  Launch with torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29501 \
        benchmarks/multigpu_check.py [--sweep]

Checks (every rank; rank 0 prints JSON lines and appends them to gpurun_out/multigpu.jsonl):
  1. fused all-reduce + update (one-shot and two-shot) == dist.all_reduce + /W + torch optimizer, replicas identical
  2. SyncBatchNorm over peer memory == torch.nn.SyncBatchNorm (forward, running stats, dx, dgamma, dbeta)
  3. a few ResNet-18 training steps: replicas stay bit-identical, loss finite
  4. (--sweep) config 5 of BASELINE.json: all-reduce + update time and bus bandwidth, 1 KB .. 1 GB, vs NCCL + foreach SGD
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def emit(rec):
    if dist.get_rank() == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "multigpu.jsonl"), "a") as f:
            f.write(json.dumps(rec) + "\n")
        print(json.dumps(rec), flush=True)


def relerr(a, b):
    return float((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6))


def all_ok(flag: bool, device) -> bool:
    t = torch.tensor([1 if flag else 0], device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item())


def replicas_identical(dp, world):
    """bf16 compute copies bit-identical on every rank, and -- after the collective gather of the sharded fp32 masters --
    the masters too."""
    dp.sync_masters()
    torch.cuda.synchronize()
    same = True
    for flat in (dp.arena.shadow, dp.arena.master):
        mine = flat.clone()
        gathered = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(gathered, mine)
        bits = torch.int16 if flat.dtype == torch.bfloat16 else torch.int32      # bit-for-bit, NaN-proof
        same = same and all(torch.equal(gathered[0].view(bits), t.view(bits)) for t in gathered)
    return same


def check_fused_update(device, mode, wire="bf16"):
    from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
    rank, world = dist.get_rank(), dist.get_world_size()
    ok = True
    os.environ["PICD_GRAD_WIRE"] = wire          # read when the communicator is created; the same on every rank
    # bf16 wire: the reference below rounds the gradients the same way, but a gradient within one fp32 ulp of a bf16 rounding
    # boundary can still land on the other side (0.4 % of that element), which Adam's normalisation amplifies for
    # elements with tiny second moments (measured: SGD 1e-6 .. 5e-4, AdamW 2.5e-3 .. 4.6e-3)
    # AdamW + bf16 wire is judged on the 99.9th percentile: its first steps move a weight by lr * g / |g|, so an element whose
    # summed gradient sits at the rounding level can take a step of the opposite SIGN (seen at W = 4: one element 1.2e-2 off
    # while p99.9 stays at the SGD level); the maximum is only bounded by what sign flips can do, 2 * lr * steps / max|w|.
    tols = {"fp32": {"sgd": 2e-4, "adamw": 2e-4}, "bf16": {"sgd": 2e-3, "adamw": 2e-3}}
    for kind in ("sgd", "adamw"):
        torch.manual_seed(0)
        # smooth activations on purpose: with ReLU a pre-activation at the rounding boundary flips its derivative between the
        # engine (weights 1 ulp off after a step) and the reference, which moves that unit's gradient by 1/batch -- seen at
        # W = 3, step 2, rank 2: 3 % of one row, 1e-2 after AdamW (benchmarks/dev/w3_debug4.py); the kernel was exact
        net = torch.nn.Sequential(torch.nn.Linear(256, 512), torch.nn.Tanh(), torch.nn.Linear(512, 1024),
                                  torch.nn.Tanh(), torch.nn.Linear(1024, 8)).to(device)
        ref = torch.nn.Sequential(torch.nn.Linear(256, 512), torch.nn.Tanh(), torch.nn.Linear(512, 1024),
                                  torch.nn.Tanh(), torch.nn.Linear(1024, 8)).to(device)
        ref.load_state_dict(net.state_dict())
        lr = 0.05 if kind == "sgd" else 1e-3
        dp = FusedDataParallel(net, OptimizerSpec(kind=kind, lr=lr, momentum=0.9, weight_decay=0.01),
                               first_bucket_bytes=64 << 10, bucket_cap_bytes=1 << 20, comm=mode)
        if mode == "nvls" and not dp._impl.comm.has_nvls:
            emit({"test": f"fused_update[{mode},{kind},{wire}]", "skipped": "no NVSwitch multicast mapping on this host", "ok": True})
            del dp
            continue
        opt = (torch.optim.SGD(ref.parameters(), lr=lr, momentum=0.9, weight_decay=0.01) if kind == "sgd"
               else torch.optim.AdamW(ref.parameters(), lr=lr, weight_decay=0.01))
        for it in range(4):
            g = torch.Generator(device="cpu").manual_seed(100 * it + rank)
            x = torch.randn(32, 256, generator=g).to(device)
            dp(x).square().mean().backward()
            dp.optimizer.step()
            opt.zero_grad()
            ref(x).square().mean().backward()
            for p in ref.parameters():
                if wire == "bf16":       # the reference applies the wire format: every rank's gradient is rounded to bf16
                    p.grad.copy_(p.grad.to(torch.bfloat16).float())
                # the kernel's arithmetic, not NCCL's: contributions added in rank order from zero, then times fp32(1 / W)
                # (a ring or tree all-reduce adds in another order and `/ W` is not `* (1 / W)` for W = 3, 5, 6, 7 -- both are
                # fp32 last-bit effects that Adam's g / |g| turns into visible differences where a summed gradient nearly cancels)
                parts = [torch.empty_like(p.grad) for _ in range(world)]
                dist.all_gather(parts, p.grad.contiguous())
                tot = torch.zeros_like(p.grad)
                for part in parts:
                    tot += part
                if wire == "bf16" and dp._impl._mode(dp.arena.buckets[0]) >= 3:
                    tot = tot.to(torch.bfloat16).float()                 # the switch delivers its fp32 sum as bf16
                p.grad.copy_(tot * (1.0 / world))
            opt.step()
        identical = replicas_identical(dp, world)          # also gathers the sharded masters
        errs = [relerr(a, b) for a, b in zip(net.parameters(), ref.parameters())]
        # in-switch reductions add in an order of their own: AdamW is judged on the percentile there, whatever the wire
        robust = kind == "adamw" and (wire == "bf16" or dp._impl._mode(dp.arena.buckets[0]) >= 3)
        p999 = max(float(((a.detach().float() - b.detach().float()).abs() / (b.detach().float().abs().max() + 1e-6)).flatten()
                         .kthvalue(max(1, int(a.numel() * 0.999))).values) for a, b in zip(net.parameters(), ref.parameters()))
        flip_bound = 2.0 * lr * 4 / min(float(b.detach().abs().max()) for b in ref.parameters())
        sh_ok = relerr(dp.arena.shadow.float(), dp.arena.master) < 1e-2
        zero = dp.arena.grad.abs().max().item() == 0
        tol = tols[wire][kind]
        close = (p999 < tol and max(errs) < flip_bound) if robust else max(errs) < tol
        good = all_ok(close and identical and sh_ok and zero, device)
        emit({"test": f"fused_update[{mode},{kind},{wire}]", "max_relerr": max(errs), "p99.9_relerr": p999, "tol": tol,
              "judged_on": "p99.9 (max < sign-flip bound %.3g)" % flip_bound if robust else "max", "replicas_identical": identical,
              "shadow_ok": sh_ok, "grad_zeroed": zero, "buckets": len(dp.arena.buckets),
              "masters_sharded": dp.masters_sharded, "ok": good})
        ok &= good
        del dp
    os.environ.pop("PICD_GRAD_WIRE", None)
    return ok


def check_sync_bn(device):
    from picd_b200.ops import native
    from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
    rank = dist.get_rank()
    # a slab (and therefore a communicator) exists once any model is wrapped
    holder = FusedDataParallel(torch.nn.Linear(8, 8).to(device), OptimizerSpec("sgd", lr=0.0))
    ok = True
    # the last case gives every rank a DIFFERENT batch size (a ragged last batch through a non-padding loader): the global
    # element count then has to come from the exchange, forward and backward (torch/nn/modules/_functions.py:59-62,170-179)
    for (N, C, H) in [(4, 64, 8), (2, 256, 7), (3, 2048, 3), (2 + rank, 128, 6)]:
        g = torch.Generator().manual_seed(7 + rank)
        x = (torch.randn(N, C, H, H, generator=g) * (1 + rank) + 0.2 * rank).to(device)
        xb = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        gamma = (torch.rand(C, generator=torch.Generator().manual_seed(1)) + 0.5).to(device).requires_grad_()
        beta = torch.zeros(C, device=device).requires_grad_()
        rm, rv = torch.zeros(C, device=device), torch.ones(C, device=device)
        xn = xb.clone().requires_grad_()
        z = native.bn_act(xn, gamma, beta, rm, rv, training=True, momentum=0.1, eps=1e-5, act="relu", residual=None,
                          sync_group=True)
        dz = torch.randn(N, C, H, H, generator=g).to(device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        z.backward(dz)
        bn = torch.nn.SyncBatchNorm(C).to(device)
        with torch.no_grad():
            bn.weight.copy_(gamma)
        xr = xb.float().requires_grad_()
        zr = torch.relu(bn(xr))
        zr.backward(dz.float())
        torch.cuda.synchronize()
        res = {"fwd": relerr(z, zr), "dx": relerr(xn.grad, xr.grad), "dgamma": relerr(gamma.grad, bn.weight.grad),
               "dbeta": relerr(beta.grad, bn.bias.grad), "running_mean": relerr(rm, bn.running_mean),
               "running_var": relerr(rv, bn.running_var)}
        good = all_ok(max(res.values()) < 3e-2, device)
        emit({"test": f"sync_bn[{N}x{C}x{H}x{H}]", **res, "ok": good})
        ok &= good
    del holder
    return ok


def check_training(device):
    from picd_b200.engine.trainer import WeightedCrossEntropy, compute_loss
    from picd_b200.models import Classifier, convert_sync_batchnorm
    from picd_b200.ops import synthetic_batch
    from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
    rank, world = dist.get_rank(), dist.get_world_size()
    ok = True
    for sync in (False, True):
        torch.manual_seed(0)
        m = Classifier("resnet18", 7, image_size=64)
        if sync:
            convert_sync_batchnorm(m)
        dp = FusedDataParallel(m.to(device), OptimizerSpec("sgd", lr=0.05, momentum=0.9))
        dp.train()
        losses = []
        for it in range(5):
            img, lab = synthetic_batch(16, 3, 64, 64, 7, 0, it * world + rank, device, torch.bfloat16)
            loss, _ = compute_loss(dp(img), lab, WeightedCrossEntropy(None))
            loss.backward()
            dp.optimizer.step()
            losses.append(float(loss.item()))
        identical = replicas_identical(dp, world)
        # BN running statistics: rank 0's copy is pushed to every peer by the step's last bucket kernel (DDP's per-forward
        # buffer broadcast); with SyncBN they are identical by construction
        bufs = dp._flat_buffers.clone()
        bg = [torch.empty_like(bufs) for _ in range(world)]
        dist.all_gather(bg, bufs)
        bufs_identical = all(torch.equal(bg[0], t) for t in bg) and bool(bufs.abs().sum() > 0)
        good = identical and bufs_identical and all(l == l and l < 50 for l in losses)
        emit({"test": f"resnet18_train[sync_bn={sync}]", "losses": losses, "replicas_identical": identical,
              "buffers_identical": bufs_identical, "ok": good})
        ok &= good
        del dp
    return ok


def check_sharded_masters(device):
    """PICD_SHARD_MASTERS=1 (slice owners keep their fp32 masters to themselves, only the bf16 copies travel) must produce,
    after the collective gather, bit-for-bit the masters and compute copies of the replicated-master run.  Gradients are
    written into the arena directly, so nothing in this check reads a parameter between the updates."""
    from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
    rank, world = dist.get_rank(), dist.get_world_size()
    results = {}
    for shard in (False, True):
        os.environ["PICD_SHARD_MASTERS"] = "1" if shard else "0"
        holder = torch.nn.Module()
        torch.manual_seed(5)
        holder.w = torch.nn.Parameter(torch.randn(3 * (1 << 20) + 320, device=device) * 0.05)
        dp = FusedDataParallel(holder, OptimizerSpec("adamw", lr=1e-3, weight_decay=0.01), first_bucket_bytes=1 << 20,
                               bucket_cap_bytes=4 << 20, comm="two_shot", overlap=False)
        for it in range(3):
            g = torch.Generator(device="cpu").manual_seed(17 * it + rank)
            dp.arena.grad[: holder.w.numel()].copy_(torch.randn(holder.w.numel(), generator=g).to(device) * 1e-2)
            dp.finish_step()
        stale_guard = False
        if shard:
            try:
                dp.state_dict()
            except RuntimeError:
                stale_guard = True           # state_dict() refuses while the masters are stale
        dp.sync_masters()
        torch.cuda.synchronize()
        results[shard] = (dp.arena.master.clone(), dp.arena.shadow.clone(), dp.masters_sharded, stale_guard, len(dp.arena.buckets))
        del dp
    os.environ.pop("PICD_SHARD_MASTERS", None)
    (m0, s0, f0, _, nb), (m1, s1, f1, guard, _) = results[False], results[True]
    same = bool(torch.equal(m0.view(torch.int32), m1.view(torch.int32)) and torch.equal(s0.view(torch.int16), s1.view(torch.int16)))
    good = all_ok(same and (not f0) and f1 and guard, device)
    emit({"test": "sharded_masters[two_shot,adamw]", "equal_to_replicated": same, "sharded": [f0, f1], "stale_guard": guard,
          "buckets": nb, "ok": good})
    return good


def check_optimizer_checkpoint(device):
    """Save at step k, reload into a FRESH engine, continue: must equal the uninterrupted run bit for bit.  Two-shot buckets
    keep the AdamW moments sharded by slice owner, so this only works if state_dict() assembles the owners' slices."""
    from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
    rank, world = dist.get_rank(), dist.get_world_size()

    def build():
        torch.manual_seed(0)
        net = torch.nn.Sequential(torch.nn.Linear(256, 512), torch.nn.ReLU(), torch.nn.Linear(512, 1024),
                                  torch.nn.ReLU(), torch.nn.Linear(1024, 8)).to(device)
        return FusedDataParallel(net, OptimizerSpec(kind="adamw", lr=1e-3, weight_decay=0.01), first_bucket_bytes=64 << 10,
                                 bucket_cap_bytes=1 << 20, comm="two_shot")

    def steps(dp, lo, hi):
        for it in range(lo, hi):
            g = torch.Generator(device="cpu").manual_seed(100 * it + rank)
            dp(torch.randn(32, 256, generator=g).to(device)).square().mean().backward()
            dp.optimizer.step()
        torch.cuda.synchronize()

    a = build()
    steps(a, 0, 3)
    a.sync_masters()                                       # collective: the fp32 masters are sharded by slice owner
    weights = {k: v.detach().clone() for k, v in a.state_dict().items()}
    opt = a.optimizer.state_dict()                         # collective
    steps(a, 3, 6)
    a.sync_masters()
    want = a.arena.master.clone()
    del a
    b = build()
    b.load_state_dict(weights)
    b.optimizer.load_state_dict(opt)
    steps(b, 3, 6)
    b.sync_masters()
    same = all_ok(bool(torch.equal(b.arena.master, want)), device)
    nz = [float(t.abs().sum()) > 0 for t in opt["state"]]
    emit({"test": "optimizer_checkpoint[two_shot,adamw]", "resumed_equals_uninterrupted": same, "state_nonzero": nz,
          "ok": same and all(nz)})
    del b
    return same and all(nz)


def check_graph_training(device):
    """CUDA-graph replay of the step across ranks: peer-flag epochs, SyncBN exchange slots and the learning rate come from
    device memory, an eagerly launched ragged batch in between must not disturb the recorded slots."""
    from picd_b200.engine.graphed import GraphedTrainStep
    from picd_b200.engine.trainer import WeightedCrossEntropy
    from picd_b200.models import Classifier, convert_sync_batchnorm
    from picd_b200.ops import synthetic_batch
    from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
    rank, world = dist.get_rank(), dist.get_world_size()
    ok = True
    for sync in (False, True):
        curves = {}
        for graph in (False, True):
            torch.manual_seed(0)
            m = Classifier("resnet18", 7, image_size=64)
            if sync:
                convert_sync_batchnorm(m)
            dp = FusedDataParallel(m.to(device), OptimizerSpec("sgd", lr=0.05, momentum=0.9))
            dp.train()
            step = GraphedTrainStep(dp, WeightedCrossEntropy(None), dp.optimizer, warmup=2, enabled=graph)
            losses = []
            for it in range(10):
                bsz = 8 if it == 6 else 16                      # step 6 is a ragged batch: runs eagerly in both arms
                img, lab = synthetic_batch(bsz, 3, 64, 64, 7, 0, (it % 3) * world + rank, device, torch.bfloat16)
                losses.append(float(step(img, lab).item()))
            identical = replicas_identical(dp, world)
            bufs = dp._flat_buffers.clone()
            bg = [torch.empty_like(bufs) for _ in range(world)]
            dist.all_gather(bg, bufs)
            identical = identical and all(torch.equal(bg[0], t) for t in bg)      # running statistics follow rank 0's
            curves[graph] = (losses, identical, step.replays)
            del dp, step
        (le, ie, _), (lg, ig, replays) = curves[False], curves[True]
        close = all(abs(a - b) < 0.1 * max(1.0, abs(a)) for a, b in zip(le, lg))
        good = all_ok(ie and ig and close and replays == 7 and all(l == l for l in lg), device)
        emit({"test": f"graph_replay_train[sync_bn={sync}]", "eager": le, "graph": lg, "replays": replays,
              "replicas_identical": bool(ie and ig), "ok": good})
        ok &= good
    return ok


def sweep(device):
    """all-reduce + SGD-momentum update, fused over peer memory vs NCCL all_reduce + torch._foreach update."""
    from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
    world = dist.get_world_size()
    sizes = [1 << k for k in range(8, 29, 2)]          # elements: 1 KB .. 1 GB of fp32
    if os.environ.get("PICD_SWEEP_MIN"):
        sizes = [n for n in sizes if n * 4 >= int(os.environ["PICD_SWEEP_MIN"])]
    for n in sizes:
        p = torch.nn.Parameter(torch.zeros(n, device=device))
        holder = torch.nn.Module()
        holder.w = p
        for mode in ("one_shot", "two_shot", "nvls", "nvls_ld", "two_shot+shard", "nvls+shard"):
            if mode == "one_shot" and n * 4 > (64 << 20):
                continue
            os.environ["PICD_SHARD_MASTERS"] = "1" if mode.endswith("+shard") else "0"
            mode = mode.split("+")[0]
            dp = FusedDataParallel(holder, OptimizerSpec("sgd", lr=0.01, momentum=0.9), first_bucket_bytes=1 << 40,
                                   bucket_cap_bytes=1 << 40, comm=mode, overlap=False)
            if mode.startswith("nvls") and not dp._impl.comm.has_nvls:
                del dp
                continue
            b = dp.arena.buckets[0]

            def once():
                dp._impl.launch_bucket(b)
                dp._impl.finish()
                dp.step_index += 1
            for _ in range(3):
                once()
            torch.cuda.synchronize(); dist.barrier(device_ids=[device.index])
            a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            iters = 20 if n < (1 << 24) else 5
            a.record()
            for _ in range(iters):
                once()
            e.record(); torch.cuda.synchronize()
            ms = torch.tensor([a.elapsed_time(e) / iters], device=device)
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            bytes_ = dp.arena.total * 4
            emit({"test": "sweep", "impl": f"p2p_{mode}", "wire": "bf16" if dp._impl.comm.wire_bf16 else "fp32",
                  "masters_sharded": dp.masters_sharded, "bytes": bytes_, "ms": float(ms), "world": world,
                  "busbw_GBs": 2 * (world - 1) / world * bytes_ / (float(ms) * 1e-3) / 1e9, "ok": True})
            del dp
        # NCCL baseline: all_reduce + /W + momentum SGD (foreach) + bf16 cast, what DDP + optimizer.step amount to
        g = torch.zeros(n, device=device); w = torch.zeros(n, device=device); mom = torch.zeros(n, device=device)
        sh = torch.zeros(n, device=device, dtype=torch.bfloat16)

        def nccl_once():
            dist.all_reduce(g)
            g.div_(world)
            mom.mul_(0.9).add_(g)
            w.add_(mom, alpha=-0.01)
            sh.copy_(w)
            g.zero_()
        for _ in range(3):
            nccl_once()
        torch.cuda.synchronize(); dist.barrier(device_ids=[device.index])
        a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 20 if n < (1 << 24) else 5
        a.record()
        for _ in range(iters):
            nccl_once()
        e.record(); torch.cuda.synchronize()
        ms = torch.tensor([a.elapsed_time(e) / iters], device=device)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        emit({"test": "sweep", "impl": "nccl+foreach", "bytes": n * 4, "ms": float(ms), "world": world,
              "busbw_GBs": 2 * (world - 1) / world * n * 4 / (float(ms) * 1e-3) / 1e9, "ok": True})
        del g, w, mom, sh
        torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--skip-checks", action="store_true")
    args = ap.parse_args()
    from picd_b200.parallel import dist as D
    D.init_distributed("b200")
    device = torch.device("cuda", torch.cuda.current_device())
    ok = True
    if not args.skip_checks:
        for mode in ("one_shot", "two_shot", "nvls", "auto"):
            ok &= check_fused_update(device, mode, "bf16")
        for mode in ("two_shot", "nvls"):
            ok &= check_fused_update(device, mode, "fp32")       # full-precision exchange, replicated masters
        ok &= check_sync_bn(device)
        ok &= check_training(device)
        ok &= check_optimizer_checkpoint(device)
        ok &= check_sharded_masters(device)
        ok &= check_graph_training(device)
    if args.sweep:
        sweep(device)
    emit({"test": "SUMMARY", "world": dist.get_world_size(), "ok": bool(ok)})
    D.barrier()
    D.shutdown()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
