#!/usr/bin/env python
"""Find the first operation that invalidates a CUDA-graph capture of the training step (debug aid)."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from picd_b200.engine.trainer import WeightedCrossEntropy, compute_loss  # noqa: E402
from picd_b200.models import Classifier  # noqa: E402
from picd_b200.ops import native  # noqa: E402
from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec  # noqa: E402

rt = ctypes.CDLL("libcudart.so.12")
mode = sys.argv[1] if len(sys.argv) > 1 else "thread_local"
state = {"bad": None, "n": 0}


def status():
    st = ctypes.c_int(0)
    rc = rt.cudaStreamIsCapturing(ctypes.c_void_p(torch.cuda.current_stream().cuda_stream), ctypes.byref(st))
    return rc, st.value


orig_call = native.call


def traced(name, *a):
    orig_call(name, *a)
    state["n"] += 1
    rc, st = status()
    if (rc != 0 or st == 2) and state["bad"] is None:
        state["bad"] = (name, state["n"], rc, st)
        print("FIRST BAD after native call", state["bad"], flush=True)


native.call = traced
dev = torch.device("cuda", 0)
model = Classifier("resnet18", 7, image_size=64).to(dev)
dp = FusedDataParallel(model, OptimizerSpec("sgd", lr=0.02, momentum=0.9))
dp.train()
loss_fn = WeightedCrossEntropy(torch.tensor([3, 3, 10, 1, 4, 4, 5.0], device=dev))
x = torch.randn(32, 3, 64, 64, device=dev).to(torch.bfloat16)
y = torch.randint(0, 7, (32,), device=dev)


def step():
    out = dp(x)
    print("  after forward", status(), flush=True)
    loss, _ = compute_loss(out, y, loss_fn)
    print("  after loss", status(), flush=True)
    dp.optimizer.zero_grad()
    loss.backward()
    print("  after backward", status(), flush=True)
    dp.optimizer.step()
    print("  after step", status(), flush=True)
    return loss


for _ in range(3):
    step()
torch.cuda.synchronize()
dp._impl.enable_dyn()
dp.capturing = True
g = torch.cuda.CUDAGraph()
try:
    with torch.cuda.graph(g, capture_error_mode=mode):
        loss = step()
    print("capture OK")
    dp._impl.write_dyn(4)
    g.replay()
    torch.cuda.synchronize()
    print("replay OK", float(loss))
except Exception as e:  # noqa: BLE001
    print("capture failed:", str(e).splitlines()[0])

# ---- same thing through GraphedTrainStep
from picd_b200.engine.graphed import GraphedTrainStep  # noqa: E402
model2 = Classifier("resnet18", 7, image_size=64).to(dev)
dp2 = FusedDataParallel(model2, OptimizerSpec("sgd", lr=0.02, momentum=0.9))
dp2.train()
gs = GraphedTrainStep(dp2, loss_fn, dp2.optimizer, warmup=2)
state["bad"] = None
for i in range(5):
    try:
        l = gs(x, y)
        print("graphed call", i, float(l.detach()), "graph" if gs.graph is not None else "eager", flush=True)
    except Exception as e:  # noqa: BLE001
        import traceback
        traceback.print_exc()
        break
