"""``deepspeed.comm``: the torch.distributed-shaped API the collective benchmarks call
(collectives/1d/dsccl.py:59-163, collectives/3d/dsccl.py:46-118).  Every function forwards to
torch.distributed, i.e. to the dlbb backend when the process group was created with 'ccl'/'dlbb'."""
from __future__ import annotations

import torch.distributed as _dist
from torch.distributed import ReduceOp  # noqa: F401

get_rank = _dist.get_rank
get_world_size = _dist.get_world_size
barrier = _dist.barrier
all_reduce = _dist.all_reduce
all_gather = _dist.all_gather
all_gather_into_tensor = _dist.all_gather_into_tensor
broadcast = _dist.broadcast
gather = _dist.gather
scatter = _dist.scatter
reduce = _dist.reduce
reduce_scatter = _dist.reduce_scatter
reduce_scatter_tensor = _dist.reduce_scatter_tensor
all_to_all = _dist.all_to_all
all_to_all_single = _dist.all_to_all_single
send = _dist.send
recv = _dist.recv
isend = _dist.isend
irecv = _dist.irecv
is_initialized = _dist.is_initialized
destroy_process_group = _dist.destroy_process_group
get_backend = _dist.get_backend
new_group = _dist.new_group


def get_local_rank() -> int:
    import os
    return int(os.environ.get("LOCAL_RANK", "0"))


def init_distributed(*a, **kw):
    import deepspeed
    return deepspeed.init_distributed(*a, **kw)
