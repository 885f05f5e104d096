"""Data-parallel machinery: process-group helpers, sharded sampler, the NVLink peer-memory
arena and the fused all-reduce + optimizer engine."""
from .dist import (all_gather, barrier, broadcast_object, comm_device, get_local_rank, get_rank,
                   get_world_size, init_distributed, is_dist, is_main, reduce_tensor, shutdown)
from .sampler import ShardedSampler

__all__ = [n for n in dir() if not n.startswith("_")]
