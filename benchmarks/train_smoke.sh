#!/bin/bash
# End-to-end smoke run of train.py on a GPU box: folder dataset (generated PNGs), N ranks, SyncBN, checkpoint + resume.
N=${1:-2}
set -e
rm -rf /tmp/picd_data /tmp/picd_ck
python - <<'PY'
from picd_b200.data import write_fake_dataset
write_fake_dataset("/tmp/picd_data", classes=7, per_class=12, size=80)
PY
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 train.py "$@"; }
run --datadir /tmp/picd_data --model resnet50 --image-size 64 --batchsize 8 --epochs 2 --optimizer sgd --lr 0.02 --sync-bn \
    --num-workers 2 --log-interval 2 --ckpt-dir /tmp/picd_ck --calibrate-comm --cuda-graph --trace /tmp/picd_trace.jsonl 2>&1 | grep -vE "Warning|warn|OMP|\*\*\*" | tail -12
echo "--- resume ---"
run --datadir /tmp/picd_data --model resnet50 --image-size 64 --batchsize 8 --epochs 3 --optimizer sgd --lr 0.02 --sync-bn \
    --num-workers 2 --log-interval 2 --ckpt-dir /tmp/picd_ck --cuda-graph 2>&1 | grep -vE "Warning|warn|OMP|\*\*\*" | tail -8
echo "--- inception (reference default recipe: adam, aux loss, class weights) ---"
run --datadir /tmp/picd_data --batchsize 4 --epochs 1 --num-workers 2 --log-interval 2 --ckpt-dir /tmp/picd_ck 2>&1 | grep -vE "Warning|warn|OMP|\*\*\*" | tail -6
ls -la /tmp/picd_ck/*/
