"""DeepSpeed-compatible facade over the dlbb framework.

Surface used by the reference: ``deepspeed.__version__``, ``init_distributed(dist_backend=...)``
(collectives/1d/dsccl.py:48, collectives/1d/dsgloo.py:57), ``deepspeed.comm`` (all collective
benchmarks) and ``deepspeed.initialize(model=, model_parameters=, config=)`` returning an engine
with ``__call__`` / ``backward`` / ``step`` / ``local_rank`` / ``global_rank`` / ``world_size``
(test/ccl.py:92-115, test/ds_mpi_test.py:36-47, test/test.py:28-41).

The engine is a data-parallel trainer: gradients live in one flat bucket (the analogue of
DeepSpeed's contiguous gradients / oneCCL's CCL_FUSION - many small tensors, one collective
launch), ZeRO-0 all-reduces it, ZeRO-2 reduce-scatters it, updates only this rank's shard and
all-gathers the parameters; Adam runs in the framework's own kernel on CUDA.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from . import comm  # noqa: F401  (deepspeed.comm)

__version__ = "0.16.0+dlbb"
__git_hash__ = "dlbb"
__git_branch__ = "b200"


def _world_from_env():
    rank = int(os.environ.get("RANK", os.environ.get("OMPI_COMM_WORLD_RANK", os.environ.get("PMI_RANK", "0"))))
    world = int(os.environ.get("WORLD_SIZE", os.environ.get("OMPI_COMM_WORLD_SIZE", os.environ.get("PMI_SIZE", "1"))))
    return rank, world


def init_distributed(dist_backend: str | None = None, auto_mpi_discovery: bool = True, distributed_port: int = 29500,
                     verbose: bool = True, timeout=None, init_method=None, dist_init_required=None, config=None,
                     rank: int = -1, world_size: int = -1):
    """Initialise torch.distributed.  'ccl' and 'dlbb' (and None) select the native backend;
    'gloo' / 'nccl' are passed through untouched."""
    if dist.is_initialized():
        return
    env_rank, env_world = _world_from_env()
    rank = env_rank if rank < 0 else rank
    world_size = env_world if world_size < 0 else world_size
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(distributed_port))
    os.environ.setdefault("RANK", str(rank))
    os.environ.setdefault("WORLD_SIZE", str(world_size))
    os.environ.setdefault("LOCAL_RANK", str(rank))
    backend = dist_backend
    if backend in (None, "ccl", "dlbb"):
        from distributed_llm_backend_benchmark_b200.parallel.backend import register
        backend = register(backend or "dlbb")
    kw = {}
    if timeout is not None:
        kw["timeout"] = timeout
    if init_method is not None:
        kw["init_method"] = init_method
    dist.init_process_group(backend=backend, rank=rank, world_size=world_size, **kw)
    if verbose and rank == 0:
        print(f"[deepspeed-dlbb] initialised torch.distributed backend={backend} world_size={world_size}")


class _FlatAdam:
    """Adam over one flat fp32 buffer (the framework's kernel on CUDA, its PyTorch oracle on CPU)."""

    def __init__(self, flat: torch.Tensor, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        self.flat, self.lr, self.betas, self.eps, self.weight_decay = flat, lr, betas, eps, weight_decay
        self.m = torch.zeros_like(flat)
        self.v = torch.zeros_like(flat)
        self.t = 0
        self.param_groups = [{"lr": lr, "betas": betas, "eps": eps, "weight_decay": weight_decay}]

    def step_range(self, grad: torch.Tensor, lo: int, hi: int, grad_scale: float) -> None:
        from distributed_llm_backend_benchmark_b200 import ops

        if hi > lo:
            ops.adam_step(self.flat[lo:hi], grad[lo:hi], self.m[lo:hi], self.v[lo:hi], None,
                          lr=self.param_groups[0]["lr"], beta1=self.betas[0], beta2=self.betas[1], eps=self.eps,
                          weight_decay=self.weight_decay, grad_scale=grad_scale, step=self.t)

    def state_dict(self):
        return {"m": self.m, "v": self.v, "t": self.t}


class DeepSpeedEngine:
    def __init__(self, model: torch.nn.Module, model_parameters=None, config: dict | None = None, optimizer=None):
        config = dict(config or {})
        self.module = model
        self.config = config
        rank, world = _world_from_env()
        if world > 1 and not dist.is_initialized():
            init_distributed(verbose=False)
        self.global_rank = dist.get_rank() if dist.is_initialized() else rank
        self.world_size = dist.get_world_size() if dist.is_initialized() else 1
        self.local_rank = int(os.environ.get("LOCAL_RANK", self.global_rank))
        self.zero_stage = int(config.get("zero_optimization", {}).get("stage", 0))
        self.gradient_accumulation_steps = int(config.get("gradient_accumulation_steps", 1))
        micro = config.get("train_micro_batch_size_per_gpu")
        total = config.get("train_batch_size")
        if micro is None and total is not None:
            micro = max(1, int(total) // (self.world_size * self.gradient_accumulation_steps))
        self.train_micro_batch_size_per_gpu = micro
        self.train_batch_size = total if total is not None else (micro or 1) * self.world_size * self.gradient_accumulation_steps
        if config.get("fp16", {}).get("enabled", False):
            raise NotImplementedError("deepspeed shim: fp16 engines are not implemented (the reference disables fp16)")
        params = [p for p in (model_parameters if model_parameters is not None else model.parameters()) if p.requires_grad]
        self.params = params
        # every rank starts from rank 0's weights (DeepSpeed broadcasts the module at initialize())
        if self.world_size > 1:
            for p in params:
                dist.broadcast(p.data, src=0)
        # flat parameter / gradient buckets; the module's tensors become views into them
        n = sum(p.numel() for p in params)
        pad = (-n) % max(self.world_size, 1)
        dev = params[0].device if params else torch.device("cpu")
        self.flat = torch.zeros(n + pad, dtype=torch.float32, device=dev)
        self.flat_grad = torch.zeros(n + pad, dtype=torch.float32, device=dev)
        off = 0
        for p in params:
            k = p.numel()
            self.flat[off:off + k].copy_(p.data.reshape(-1).float())
            p.data = self.flat[off:off + k].view_as(p.data)
            p.grad = self.flat_grad[off:off + k].view_as(p.data)
            off += k
        self.numel = n
        opt_cfg = config.get("optimizer")
        self.optimizer = optimizer
        if optimizer is None and opt_cfg is not None:
            kind = str(opt_cfg.get("type", "Adam")).lower()
            if kind not in ("adam", "adamw"):
                raise NotImplementedError(f"deepspeed shim: optimizer type {opt_cfg.get('type')!r}")
            p_ = dict(opt_cfg.get("params", {}))
            self.optimizer = _FlatAdam(self.flat, lr=p_.get("lr", 1e-3), betas=tuple(p_.get("betas", (0.9, 0.999))),
                                       eps=p_.get("eps", 1e-8), weight_decay=p_.get("weight_decay", 0.0))
        self.micro_steps = 0
        self.global_steps = 0

    # ------------------------------------------------------------------ nn.Module facade
    def __call__(self, *a, **kw):
        return self.module(*a, **kw)

    forward = __call__

    def train(self, mode: bool = True):
        self.module.train(mode)
        return self

    def eval(self):
        self.module.eval()
        return self

    def parameters(self):
        return self.module.parameters()

    def is_gradient_accumulation_boundary(self) -> bool:
        return (self.micro_steps + 1) % self.gradient_accumulation_steps == 0

    # ------------------------------------------------------------------ training
    def backward(self, loss: torch.Tensor, **kw):
        (loss / self.gradient_accumulation_steps).backward()
        return loss

    def step(self):
        boundary = self.is_gradient_accumulation_boundary()
        self.micro_steps += 1
        if not boundary:
            return
        ws = self.world_size
        opt = self.optimizer
        if isinstance(opt, _FlatAdam):
            opt.t += 1
            if ws == 1:
                opt.step_range(self.flat_grad, 0, self.numel, 1.0)
            elif self.zero_stage >= 2:
                # ZeRO-2: reduce-scatter the bucket, update my shard, all-gather the parameters
                shard = self.flat.numel() // ws
                mine = torch.empty(shard, dtype=torch.float32, device=self.flat.device)
                dist.reduce_scatter_tensor(mine, self.flat_grad, op=dist.ReduceOp.SUM)
                lo = self.global_rank * shard
                self.flat_grad[lo:lo + shard].copy_(mine)
                opt.step_range(self.flat_grad, lo, min(lo + shard, self.numel), 1.0 / ws)
                gathered = torch.empty_like(self.flat)
                dist.all_gather_into_tensor(gathered, self.flat[lo:lo + shard].contiguous())
                self.flat.copy_(gathered)
            else:
                dist.all_reduce(self.flat_grad, op=dist.ReduceOp.SUM)
                opt.step_range(self.flat_grad, 0, self.numel, 1.0 / ws)
        elif opt is not None:  # client optimizer
            if ws > 1:
                dist.all_reduce(self.flat_grad, op=dist.ReduceOp.SUM)
                self.flat_grad.div_(ws)
            opt.step()
        elif ws > 1:
            dist.all_reduce(self.flat_grad, op=dist.ReduceOp.SUM)
        self.flat_grad.zero_()
        self.global_steps += 1

    def zero_grad(self):
        self.flat_grad.zero_()

    def save_checkpoint(self, save_dir, tag=None, client_state=None, **kw):
        os.makedirs(save_dir, exist_ok=True)
        if self.global_rank == 0:
            torch.save({"module": self.module.state_dict(), "global_steps": self.global_steps,
                        "optimizer": self.optimizer.state_dict() if hasattr(self.optimizer, "state_dict") else None,
                        "client_state": client_state}, os.path.join(save_dir, f"{tag or 'latest'}.pt"))

    def load_checkpoint(self, load_dir, tag=None, **kw):
        sd = torch.load(os.path.join(load_dir, f"{tag or 'latest'}.pt"), map_location="cpu", weights_only=False)
        self.module.load_state_dict(sd["module"])
        self.global_steps = sd.get("global_steps", 0)
        return load_dir, sd.get("client_state")


def initialize(args=None, model=None, optimizer=None, model_parameters=None, training_data=None, lr_scheduler=None,
               config=None, config_params=None, **kw):
    """Returns (engine, optimizer, training_dataloader, lr_scheduler) like deepspeed.initialize."""
    if model is None:
        raise ValueError("deepspeed.initialize: model is required")
    cfg = config if config is not None else config_params
    if isinstance(cfg, str):
        import json
        with open(cfg) as fh:
            cfg = json.load(fh)
    engine = DeepSpeedEngine(model, model_parameters, cfg, optimizer)
    return engine, engine.optimizer, None, lr_scheduler
