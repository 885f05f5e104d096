"""Reference-compatible module name: ``from ddp_utils import all_gather, reduce_tensor``
(reference ddp_utils.py:8-56), plus init / rank / barrier helpers."""
from picd_b200.parallel.dist import (all_gather, barrier, broadcast_object, get_local_rank, get_rank,  # noqa: F401
                                     get_world_size, init_distributed, is_main, reduce_tensor, shutdown)
