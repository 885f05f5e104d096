#!/bin/bash
# usage: benchmarks/run_scaling.sh N [extra bench args]   -- both arms of bench.py on N GPUs of this node
N=$1; shift
mkdir -p gpurun_out
run() {
  if [ "$N" = "1" ]; then python bench.py --gpus 1 "$@"; else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N "$@"; fi
}
run --impl reference --steps 15 --warmup 4 "$@" 2>/dev/null | tail -1 > gpurun_out/scale_ref_$N.json; cat gpurun_out/scale_ref_$N.json | cut -c1-200
run --steps 15 --warmup 4 "$@" 2>gpurun_out/scale_b200_$N.err | tail -1 > gpurun_out/scale_b200_$N.json; cat gpurun_out/scale_b200_$N.json | cut -c1-200
