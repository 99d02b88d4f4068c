"""Minimal check that ``backend='ccl'`` exists after importing the bindings (ref test/test_ccl_backend.py).
    mpirun -n 2 python test/test_ccl_backend.py"""
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (_ROOT, os.path.join(_ROOT, "compat")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import oneccl_bindings_for_pytorch  # noqa: E402,F401

os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29500")
os.environ.setdefault("RANK", "0")
os.environ.setdefault("WORLD_SIZE", "1")
dist.init_process_group(backend="ccl")
rank, world = dist.get_rank(), dist.get_world_size()
t = torch.tensor([rank], dtype=torch.float32)
dist.all_reduce(t)
assert t.item() == sum(range(world)), t
if rank == 0:
    print(f"'ccl' backend is working ({world} rank(s), all_reduce -> {t.item()})")
dist.destroy_process_group()
