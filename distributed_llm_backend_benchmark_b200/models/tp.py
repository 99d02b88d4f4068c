"""Tensor-parallel wrappers with the reference's class names and constructor signatures
(ref models.py:19-335): ``ColumnParallelLinear``, ``RowParallelLinear``, ``TransformerBlock``,
``LLM``, ``create_model`` and ``create_model_from_config(config, rank, world_size, comm)``.

All of them are thin faces over :class:`LLMEngine`: parameters live in the engine's flat buffer on
the rank's GPU, matmuls are the tcgen05 GEMM, and the row-parallel SUM all-reduce
(models.py:84-98: bf16 -> fp32 -> numpy -> MPI -> bf16 on the host) is one device kernel that
reduces the bf16 partial sums over NVLink with fp32 accumulation and adds the residual
(models.py:173,188) on the way out.  ``comm`` may be the mpi4py shim's communicator, a
``SymmComm`` or None (single rank); on a CPU-only host the reduction goes through
torch.distributed (Gloo) so the plumbing can be tested anywhere.
"""
from __future__ import annotations

import os

import torch

from .. import ops
from .llm import MODEL_CONFIGS, LLMConfig, LLMEngine


def default_device(rank: int = 0) -> torch.device:
    if torch.cuda.is_available() and os.environ.get("DLBB_DEVICE", "cuda") != "cpu":
        return torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank % torch.cuda.device_count())))
    return torch.device("cpu")


class _Reducer:
    """all_reduce(partial) + residual for the row-parallel linears, on whatever transport exists."""

    def __init__(self, comm, rank: int, world_size: int, device: torch.device):
        self.world = world_size
        self.device = device
        self.symm = None
        self.partial = None
        self._symm_key, self._by_ptr, self._part, self._outs = None, {}, None, None
        if world_size == 1:
            return
        from ..parallel.symm import SymmComm

        if isinstance(comm, SymmComm):
            self.symm = comm
        elif device.type == "cuda":
            self.symm = SymmComm.from_env(device)
        else:
            import torch.distributed as dist

            if not dist.is_initialized():
                from ..parallel import backend as be
                be.init_process_group(None, rank=rank, world_size=world_size)
            self.dist = dist

    def symm_buffers(self, tokens: int, hidden: int):
        """Activation buffers the fused GEMM + all-reduce kernel can write on every rank (symmetric
        memory): two ping-pong outputs and the partial-sum matrix.  None when there is no NVLink transport."""
        if self.symm is None or os.environ.get("DLBB_TP_FUSED", "1") != "1":
            return None
        key = (tokens, hidden)
        if self._symm_key != key:
            self._part = self.symm.alloc_matrix(tokens, hidden)
            self._outs = [self.symm.alloc_matrix(tokens, hidden) for _ in range(2)]
            self._by_ptr = {m.tensor.data_ptr(): m for m in self._outs}
            self._symm_key = key
        return self._outs[0].tensor, self._outs[1].tensor

    def __call__(self, inp: torch.Tensor, w: torch.Tensor, residual: torch.Tensor | None, out: torch.Tensor):
        fused_out = self._by_ptr.get(out.data_ptr()) if self.symm is not None else None
        if fused_out is not None:
            # one kernel: tcgen05 GEMM, partial tiles reduced across ranks over NVLink while later tiles compute
            return self.symm.gemm_allreduce(inp, w, self._part, fused_out, residual=residual)
        if self.partial is None or self.partial.shape != out.shape:
            self.partial = torch.empty_like(out)
        ops.matmul(inp, w, out=self.partial)
        if self.symm is not None:
            if residual is None:
                out.copy_(self.partial)
                self.symm.all_reduce(out)
            else:
                self.symm.all_reduce_residual(self.partial, residual, out)
            return out
        acc = self.partial.float()               # the reference's fp32 wire format (models.py:84)
        self.dist.all_reduce(acc)
        if residual is not None:
            acc += residual.float()
        out.copy_(acc.to(out.dtype))
        return out


class ColumnParallelLinear:
    """y = x @ W[:, shard]  (ref models.py:19-47).  Weight ``[in, out/ws]`` bf16, unscaled randn."""

    def __init__(self, in_features, out_features, rank, world_size, device=None):
        self.in_features, self.out_features = in_features, out_features
        self.rank, self.world_size = rank, world_size
        self.out_features_per_rank = out_features // world_size
        self.device = device or default_device(rank)
        self.weight = torch.randn(in_features, self.out_features_per_rank, device=self.device).bfloat16()

    def forward(self, x):
        lead = x.shape[:-1]
        y = ops.matmul(x.reshape(-1, self.in_features).to(self.device), self.weight)
        return y.view(*lead, self.out_features_per_rank)

    __call__ = forward

    def parameters(self):
        return [self.weight]


class RowParallelLinear:
    """y = all_reduce_sum(x_shard @ W[shard, :])  (ref models.py:50-100)."""

    def __init__(self, in_features, out_features, rank, world_size, comm, device=None):
        self.in_features, self.out_features = in_features, out_features
        self.rank, self.world_size, self.comm = rank, world_size, comm
        self.in_features_per_rank = in_features // world_size
        self.device = device or default_device(rank)
        self.weight = torch.randn(self.in_features_per_rank, out_features, device=self.device).bfloat16()
        self._reduce = _Reducer(comm, rank, world_size, self.device)

    def forward(self, x):
        lead = x.shape[:-1]
        x2 = x.reshape(-1, self.in_features_per_rank).to(self.device)
        out = torch.empty(x2.shape[0], self.out_features, dtype=torch.bfloat16, device=self.device)
        if self.world_size == 1:
            ops.matmul(x2, self.weight, out=out)
        else:
            self._reduce(x2, self.weight, None, out)
        return out.view(*lead, self.out_features)

    __call__ = forward

    def parameters(self):
        return [self.weight]


class LLM:
    """The N-layer model (ref models.py:197-245) as one engine; ``model(batch)`` is the timed call of
    run_mpi.py:179-180.  A CPU batch is copied to the GPU and the result copied back, so the
    reference driver works unchanged; pass a CUDA batch to keep everything resident."""

    def __init__(self, hidden_size, num_layers, num_heads, ffn_intermediate, rank, world_size, comm, device=None,
                 seed=None):
        self.hidden_size, self.num_layers, self.num_heads = hidden_size, num_layers, num_heads
        self.ffn_intermediate = ffn_intermediate
        self.rank, self.world_size, self.comm = rank, world_size, comm
        self.device = device or default_device(rank)
        cfg = LLMConfig(hidden_size, num_layers, num_heads, ffn_intermediate)
        for name, dim in (("3*hidden_size", 3 * hidden_size), ("hidden_size", hidden_size), ("ffn_intermediate", ffn_intermediate)):
            if dim % world_size:
                raise ValueError(f"{name}={dim} is not divisible by world_size={world_size}")
        self._reduce = _Reducer(comm, rank, world_size, self.device) if world_size > 1 else None
        self.engine = LLMEngine(cfg, self.device, tp_rank=rank, tp_size=world_size, seed=seed, allreduce=self._reduce)

    def forward(self, x):
        on_host = not x.is_cuda and self.device.type == "cuda"
        y = self.engine.forward(x.to(self.device, non_blocking=True) if on_host else x)
        return y.cpu() if on_host else y

    __call__ = forward

    def eval(self):
        return self

    def parameters(self):
        return [self.engine.params]

    def get_num_parameters(self):
        return self.engine.get_num_parameters()

    def get_memory_footprint(self):
        return self.engine.get_memory_footprint()


class TransformerBlock(LLM):
    """One block (ref models.py:107-190) = a one-layer engine without the final LayerNorm."""

    def __init__(self, hidden_size, num_heads, ffn_intermediate, rank, world_size, comm, device=None, seed=None):
        super().__init__(hidden_size, 1, num_heads, ffn_intermediate, rank, world_size, comm, device, seed)

    def forward(self, x):
        # ln_final has gamma=1 / beta=0 at init but still normalises, so run the layer alone
        eng = self.engine
        on_host = not x.is_cuda and self.device.type == "cuda"
        xd = x.to(self.device) if on_host else x
        H, tp = self.hidden_size, self.world_size
        tokens = xd.numel() // H
        cur = xd.reshape(tokens, H).contiguous()
        lv = eng.layers[0]
        h, _, _ = ops.layernorm_fwd(cur, lv.ln1_g, lv.ln1_b, eng.eps, save_stats=False)
        qkv = ops.matmul(h, lv.wqkv)
        x2 = eng._row_parallel(qkv[:, :H // tp], lv.wo, cur, torch.empty_like(cur))
        h, _, _ = ops.layernorm_fwd(x2, lv.ln2_g, lv.ln2_b, eng.eps, save_stats=False)
        act = ops.matmul(h, lv.wup, epilogue="gelu")
        y = eng._row_parallel(act, lv.wdown, x2, torch.empty_like(cur)).view(x.shape)
        return y.cpu() if on_host else y

    __call__ = forward


def create_model(model_size, rank, world_size, comm, **kw):
    if model_size not in MODEL_CONFIGS:
        raise ValueError(f"Invalid model size: {model_size}. Choose from {list(MODEL_CONFIGS.keys())}")
    c = MODEL_CONFIGS[model_size]
    return LLM(c["hidden_size"], c["num_layers"], c["num_heads"], c["ffn_intermediate"], rank, world_size, comm, **kw)


def create_model_from_config(config, rank, world_size, comm, **kw):
    m = config["model"]
    return LLM(m["hidden_size"], m["num_layers"], m["num_heads"], m["ffn_intermediate"], rank, world_size, comm, **kw)
