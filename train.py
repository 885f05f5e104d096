#!/usr/bin/env python
"""Entry point with the reference's name and launch recipe (README.md:6):

    python -m torch.distributed.launch --nproc_per_node=8 train.py --datadir D
    torchrun --nproc-per-node=8 train.py --model resnet50 --synthetic --optimizer sgd --lr 0.1

All logic lives in picd_b200.engine.cli; nothing runs at import time (the reference executes
its whole job at module scope, train.py:99-188)."""
import sys

from picd_b200.engine.cli import main

if __name__ == "__main__":
    sys.exit(main())
