"""The reference's transformer (``models.py``) as a device-resident engine with a hand-scheduled
forward and backward - no autograd graph, no per-step allocation.

Architecture parity (ref models.py:107-245): pre-LN block ``ln1 -> qkv_proj (column parallel) ->
"attention" = first H/ws columns of the QKV output (models.py:166-167) -> out_proj (row parallel,
SUM all-reduce) -> +residual -> ln2 -> ffn_up (column parallel) -> exact GELU -> ffn_down (row
parallel, all-reduce) -> +residual``, ``ln_final`` after the last block; weights are ``[in, out]``
bf16 drawn from unscaled ``randn`` (models.py:35-38,68-71); LayerNorm is affine with eps 1e-5.

B200 mapping
  * every matmul is the tcgen05/TMEM/TMA kernel (ops.matmul) with the GELU / residual-add /
    GELU-backward work fused into its epilogue; transposed operands for dgrad / wgrad are
    expressed as UMMA-descriptor majors, never as copies; the QKV slice is a strided view;
  * parameters, gradients and activations live in flat preallocated buffers: one contiguous
    parameter/gradient *bucket per layer* so the data-parallel engine can launch one fused
    reduce+Adam kernel per layer while the backward of the next layer is still running;
  * tensor parallel forward (``run_mpi.py`` path) writes GEMM partial sums into symmetric memory
    and reduces them with the in-house all-reduce (residual add fused into the reduction).

Training is new semantics (the reference model is forward-only): the loss is
``mse(ln_final(x_L), target)`` with a fixed synthetic target.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Callable

import torch

from .. import ops

MODEL_CONFIGS = {
    "1B": dict(hidden_size=2048, num_layers=24, num_heads=16, ffn_intermediate=8192),
    "7B": dict(hidden_size=4096, num_layers=32, num_heads=32, ffn_intermediate=16384),
    "13B": dict(hidden_size=5120, num_layers=40, num_heads=40, ffn_intermediate=20480),
}


@dataclass
class LLMConfig:
    hidden_size: int
    num_layers: int
    num_heads: int
    ffn_intermediate: int

    @staticmethod
    def from_size(size: str) -> "LLMConfig":
        if size not in MODEL_CONFIGS:
            raise ValueError(f"Invalid model size: {size}. Choose from {list(MODEL_CONFIGS)}")
        return LLMConfig(**MODEL_CONFIGS[size])

    @staticmethod
    def from_dict(cfg: dict) -> "LLMConfig":
        m = cfg["model"] if "model" in cfg else cfg
        return LLMConfig(int(m["hidden_size"]), int(m["num_layers"]), int(m["num_heads"]),
                         int(m["ffn_intermediate"]))

    def params_per_layer(self, tp: int = 1) -> int:
        H, F = self.hidden_size, self.ffn_intermediate
        return H * (3 * H // tp) + (H // tp) * H + H * (F // tp) + (F // tp) * H + 4 * H

    def num_parameters(self, tp: int = 1) -> int:
        return self.num_layers * self.params_per_layer(tp) + 2 * self.hidden_size

    def forward_flops(self, tokens: int) -> float:
        H, F = self.hidden_size, self.ffn_intermediate
        return 2.0 * tokens * self.num_layers * (3 * H * H + H * H + 2 * H * F)


class _LayerView:
    """Named views into one layer's slice of a flat buffer (parameters or gradients)."""

    __slots__ = ("wqkv", "wo", "wup", "wdown", "ln1_g", "ln1_b", "ln2_g", "ln2_b", "flat")

    def __init__(self, flat: torch.Tensor, H: int, F: int, tp: int):
        self.flat = flat
        o = 0

        def take(r, c):
            nonlocal o
            v = flat[o:o + r * c].view(r, c)
            o += r * c
            return v

        self.wqkv = take(H, 3 * H // tp)
        self.wo = take(H // tp, H)
        self.wup = take(H, F // tp)
        self.wdown = take(F // tp, H)
        self.ln1_g = take(1, H).view(H)
        self.ln1_b = take(1, H).view(H)
        self.ln2_g = take(1, H).view(H)
        self.ln2_b = take(1, H).view(H)
        assert o == flat.numel()


class LLMEngine:
    """Parameters + forward / backward of the reference LLM on one device.

    ``tp_rank`` / ``tp_size`` shard the linears exactly like models.py:33,66 (integer division of
    the output / input features).  ``allreduce`` is a callable ``(partial, residual, out) -> out``
    used after the row-parallel GEMMs when ``tp_size > 1``.
    """

    def __init__(self, cfg: LLMConfig, device: torch.device | str = "cpu", tp_rank: int = 0, tp_size: int = 1,
                 seed: int | None = None, allreduce: Callable | None = None, training: bool = False,
                 eps: float = 1e-5, param_buffer: torch.Tensor | None = None,
                 grad_buffer: torch.Tensor | None = None):
        self.cfg = cfg
        self.device = torch.device(device)
        self.tp_rank, self.tp_size = tp_rank, tp_size
        self.allreduce = allreduce
        self.eps = eps
        self.training = training
        self.dense_qkv_backward = os.environ.get("DLBB_DENSE_QKV_BWD", "0") == "1"
        # Gradient buckets handed to the data-parallel engine: "matrix" = one per weight matrix, announced right after
        # that matrix's wgrad GEMM has been enqueued (the exchange of dW_down starts while dW_up is still being computed);
        # "layer" = one per layer.  None = let the data-parallel engine choose (it takes "layer": measured faster on 8 GPUs).
        self.bucket_mode = os.environ.get("DLBB_DP_BUCKETS")
        H, F, L = cfg.hidden_size, cfg.ffn_intermediate, cfg.num_layers
        self.layer_numel = cfg.params_per_layer(tp_size)
        self.total_numel = L * self.layer_numel + 2 * H
        # Flat parameter buffer: L layer buckets then the final LayerNorm.
        # (The data-parallel engine passes buffers that live in symmetric memory so peers can
        # read gradients and write parameters directly.)
        if param_buffer is not None:
            assert param_buffer.dtype == torch.bfloat16 and param_buffer.numel() >= self.total_numel
            self.params = param_buffer[:self.total_numel]
        else:
            self.params = torch.empty(self.total_numel, dtype=torch.bfloat16, device=self.device)
        self.layers = [_LayerView(self.params[i * self.layer_numel:(i + 1) * self.layer_numel], H, F, tp_size)
                       for i in range(L)]
        self.lnf_g = self.params[L * self.layer_numel:L * self.layer_numel + H]
        self.lnf_b = self.params[L * self.layer_numel + H:]
        self.grads = None
        self.glayers = None
        self._init_weights(seed)
        self._act = None
        self._act_key = None
        if training:
            if tp_size != 1:
                raise NotImplementedError("training runs data-parallel (tp_size must be 1); the tensor-parallel "
                                          "path is the forward benchmark of run_mpi.py")
            if grad_buffer is not None:
                assert grad_buffer.dtype == torch.bfloat16 and grad_buffer.numel() >= self.total_numel
                self.grads = grad_buffer[:self.total_numel]
                self.grads.zero_()
            else:
                self.grads = torch.zeros(self.total_numel, dtype=torch.bfloat16, device=self.device)
            self.glayers = [_LayerView(self.grads[i * self.layer_numel:(i + 1) * self.layer_numel], H, F, 1)
                            for i in range(L)]
            self.glnf_g = self.grads[L * self.layer_numel:L * self.layer_numel + H]
            self.glnf_b = self.grads[L * self.layer_numel + H:]

    # ------------------------------------------------------------------ parameters
    def _init_weights(self, seed: int | None) -> None:
        """Unscaled randn weights (ref models.py:35-38,68-71), LayerNorm gamma=1 / beta=0.

        Generated layer by layer on the target device; with a seed every TP rank draws its own
        stream (seed + rank) the way independent ``torch.randn`` calls in N processes would.
        """
        gen = None
        if seed is not None:
            gen = torch.Generator(device=self.device)
            gen.manual_seed(seed + self.tp_rank)
        for lv in self.layers:
            for w in (lv.wqkv, lv.wo, lv.wup, lv.wdown):
                w.copy_(torch.randn(w.shape, dtype=torch.float32, device=self.device, generator=gen))
            lv.ln1_g.fill_(1.0)
            lv.ln1_b.zero_()
            lv.ln2_g.fill_(1.0)
            lv.ln2_b.zero_()
        self.lnf_g.fill_(1.0)
        self.lnf_b.zero_()

    def bucket_slices(self) -> list[tuple[int, int]]:
        """(offset, numel) of each gradient bucket in backward-completion order: the final LayerNorm first, then
        layers L-1 .. 0 - as one bucket per layer, or (bucket_mode "matrix") as W_down, W_up, W_o, W_qkv and the
        layer's four LayerNorm vectors, each in the order in which backward completes them."""
        L, H, F = self.cfg.num_layers, self.cfg.hidden_size, self.cfg.ffn_intermediate
        out = [(L * self.layer_numel, 2 * H)]
        for i in reversed(range(L)):
            base = i * self.layer_numel
            if self.bucket_mode == "matrix":
                n_qkv, n_o, n_up, n_down = H * 3 * H, H * H, H * F, F * H
                o_qkv, o_o, o_up, o_down = base, base + n_qkv, base + n_qkv + n_o, base + n_qkv + n_o + n_up
                out += [(o_down, n_down), (o_up, n_up), (o_o, n_o), (o_qkv, n_qkv), (o_down + n_down, 4 * H)]
            else:
                out.append((base, self.layer_numel))
        return out

    def get_num_parameters(self) -> int:
        """Total across all TP ranks (ref models.py:239-241)."""
        return self.total_numel * self.tp_size

    def get_memory_footprint(self) -> int:
        """Bytes of parameters on this rank (ref models.py:243-245)."""
        return self.total_numel * self.params.element_size()

    # ------------------------------------------------------------------ buffers
    def _buffers(self, tokens: int):
        key = (tokens, self.training)
        if self._act_key == key:
            return self._act
        H, F, L, tp = self.cfg.hidden_size, self.cfg.ffn_intermediate, self.cfg.num_layers, self.tp_size
        dev, bf = self.device, torch.bfloat16

        def buf(*shape, dtype=bf):
            return torch.empty(*shape, dtype=dtype, device=dev)

        a = {}
        if self.training:
            a["x"] = [buf(tokens, H) for _ in range(L + 1)]
            a["h1"] = [buf(tokens, H) for _ in range(L)]
            a["qkv"] = [buf(tokens, 3 * H) for _ in range(L)]
            a["x2"] = [buf(tokens, H) for _ in range(L)]
            a["h2"] = [buf(tokens, H) for _ in range(L)]
            a["z"] = [buf(tokens, F) for _ in range(L)]
            a["a"] = [buf(tokens, F) for _ in range(L)]
            a["st1"] = [None] * L
            a["st2"] = [None] * L
            a["y"] = buf(tokens, H)
            a["dy"] = buf(tokens, H)
            a["dxa"] = buf(tokens, H)
            a["dxb"] = buf(tokens, H)
            a["dh"] = buf(tokens, H)
            a["dattn"] = buf(tokens, H)
            a["dz"] = buf(tokens, F)
            a["loss"] = torch.zeros(1, dtype=torch.float32, device=dev)
            if self.dense_qkv_backward:
                a["dqkv"] = torch.zeros(tokens, 3 * H, dtype=bf, device=dev)
        else:
            symm = self.allreduce.symm_buffers(tokens, H) if hasattr(self.allreduce, "symm_buffers") else None
            a["xa"], a["xb"] = symm if symm is not None else (buf(tokens, H), buf(tokens, H))
            a["h"] = buf(tokens, H)
            a["qkv"] = buf(tokens, 3 * H // tp)
            a["act"] = buf(tokens, F // tp)
            a["part"] = buf(tokens, H) if tp > 1 and self.allreduce is None else None
            a["y"] = buf(tokens, H)
        self._act, self._act_key = a, key
        return a

    # ------------------------------------------------------------------ inference forward
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """x: [batch, seq, H] bf16 -> [batch, seq, H] bf16 (ref models.py:224-237)."""
        H, tp = self.cfg.hidden_size, self.tp_size
        shape = x.shape
        tokens = x.numel() // H
        a = self._buffers(tokens)
        cur = x.reshape(tokens, H)
        nxt = a["xa"]
        for lv in self.layers:
            h, _, _ = ops.layernorm_fwd(cur, lv.ln1_g, lv.ln1_b, self.eps, out=a["h"], save_stats=False)
            qkv = ops.matmul(h, lv.wqkv, out=a["qkv"])
            attn = qkv[:, :H // tp]                       # the reference's "attention" (models.py:166-167)
            x2 = self._row_parallel(attn, lv.wo, cur, nxt)
            other = a["xb"] if nxt is a["xa"] else a["xa"]
            h, _, _ = ops.layernorm_fwd(x2, lv.ln2_g, lv.ln2_b, self.eps, out=a["h"], save_stats=False)
            act = ops.matmul(h, lv.wup, epilogue="gelu", out=a["act"])
            x3 = self._row_parallel(act, lv.wdown, x2, other)
            cur, nxt = x3, (a["xa"] if other is a["xb"] else a["xb"])
        y, _, _ = ops.layernorm_fwd(cur, self.lnf_g, self.lnf_b, self.eps, out=a["y"], save_stats=False)
        return y.view(shape)

    __call__ = forward

    def _row_parallel(self, inp: torch.Tensor, w: torch.Tensor, residual: torch.Tensor, out: torch.Tensor):
        """out = all_reduce(inp @ w) + residual  (ref models.py:73-100 followed by :173 / :188)."""
        if self.tp_size == 1:
            return ops.matmul(inp, w, epilogue="add", aux=residual, out=out)
        if self.allreduce is None:
            raise RuntimeError("tp_size > 1 needs an allreduce callable")
        return self.allreduce(inp, w, residual, out)

    # ------------------------------------------------------------------ training step pieces
    def forward_backward(self, x: torch.Tensor, target: torch.Tensor, grad_scale: float = 1.0,
                         on_bucket_ready: Callable[[int], None] | None = None) -> torch.Tensor:
        """One forward + backward.  Gradients land in ``self.grads`` (bf16, flat, per-layer buckets).

        ``on_bucket_ready(i)`` is called right after the i-th bucket of ``bucket_slices()`` is
        complete (kernels enqueued on the current stream), so a data-parallel engine can start
        reducing it while earlier layers are still in backward.  Returns the fp32 loss tensor [1].
        """
        assert self.training
        cfg = self.cfg
        H, L = cfg.hidden_size, cfg.num_layers
        tokens = x.numel() // H
        a = self._buffers(tokens)
        xs = a["x"]
        xs[0].copy_(x.reshape(tokens, H))
        # ---------------- forward, saving what backward needs
        for i, lv in enumerate(self.layers):
            _, m1, r1 = ops.layernorm_fwd(xs[i], lv.ln1_g, lv.ln1_b, self.eps, out=a["h1"][i])
            a["st1"][i] = (m1, r1)
            ops.matmul(a["h1"][i], lv.wqkv, out=a["qkv"][i])
            attn = a["qkv"][i][:, :H]
            ops.matmul(attn, lv.wo, epilogue="add", aux=xs[i], out=a["x2"][i])
            _, m2, r2 = ops.layernorm_fwd(a["x2"][i], lv.ln2_g, lv.ln2_b, self.eps, out=a["h2"][i])
            a["st2"][i] = (m2, r2)
            ops.matmul(a["h2"][i], lv.wup, epilogue="gelu", aux=a["z"][i], out=a["a"][i])
            ops.matmul(a["a"][i], lv.wdown, epilogue="add", aux=a["x2"][i], out=xs[i + 1])
        _, mf, rf = ops.layernorm_fwd(xs[L], self.lnf_g, self.lnf_b, self.eps, out=a["y"])
        # ---------------- loss and its gradient
        ops.mse_loss(a["y"], target.reshape(tokens, H), grad_scale=grad_scale, loss_out=a["loss"], dy=a["dy"])
        # ---------------- backward
        dx, _, _ = ops.layernorm_bwd(a["dy"], xs[L], mf, rf, self.lnf_g, None, dx=a["dxa"], dgamma_out=self.glnf_g,
                                     dbeta_out=self.glnf_b)   # parameter gradients land in the bucket as bf16
        fine = self.bucket_mode == "matrix"
        ready = 0   # index into bucket_slices() of the next bucket to announce

        def announce(only_if_fine: bool = False):
            nonlocal ready
            if only_if_fine and not fine:
                return
            if on_bucket_ready:
                on_bucket_ready(ready)
            ready += 1

        announce()
        spare = a["dxb"]
        for n, i in enumerate(reversed(range(L))):
            lv, gv = self.layers[i], self.glayers[i]
            # FFN
            # (a matrix's bucket is announced only after the LAST kernel of this step that reads the matrix's parameters -
            #  its dgrad GEMM - has been enqueued: the bucket kernel overwrites the parameters, on every rank)
            ops.matmul(a["a"][i], dx, transpose_a=True, out=gv.wdown)                       # dW_down = a^T dx3
            ops.matmul(dx, lv.wdown, transpose_b=True, epilogue="mul_dgelu", aux=a["z"][i], out=a["dz"])
            announce(True)
            ops.matmul(a["h2"][i], a["dz"], transpose_a=True, out=gv.wup)                    # dW_up = h2^T dz
            ops.matmul(a["dz"], lv.wup, transpose_b=True, out=a["dh"])                       # dh2
            announce(True)
            m2, r2 = a["st2"][i]
            dx2, _, _ = ops.layernorm_bwd(a["dh"], a["x2"][i], m2, r2, lv.ln2_g, dx, dx=spare, dgamma_out=gv.ln2_g,
                                          dbeta_out=gv.ln2_b)
            # "attention" + projections
            attn = a["qkv"][i][:, :H]
            ops.matmul(attn, dx2, transpose_a=True, out=gv.wo)                               # dW_o
            ops.matmul(dx2, lv.wo, transpose_b=True, out=a["dattn"])                         # d attn
            announce(True)
            if self.dense_qkv_backward:
                # what autograd would run: zero-padded d(qkv) through the full 3H-wide GEMMs
                a["dqkv"][:, :H].copy_(a["dattn"])
                ops.matmul(a["h1"][i], a["dqkv"], transpose_a=True, out=gv.wqkv)
                ops.matmul(a["dqkv"], lv.wqkv, transpose_b=True, out=a["dh"])
                announce(True)
            else:
                # d(qkv) is structurally zero outside the first H columns, so only that block of
                # dW_qkv is non-zero (the rest of the bucket stays at its initial zeros).
                ops.matmul(a["h1"][i], a["dattn"], transpose_a=True, out=gv.wqkv[:, :H])
                ops.matmul(a["dattn"], lv.wqkv[:, :H], transpose_b=True, out=a["dh"])
                announce(True)
            m1, r1 = a["st1"][i]
            dx1, _, _ = ops.layernorm_bwd(a["dh"], xs[i], m1, r1, lv.ln1_g, dx2, dx=dx, dgamma_out=gv.ln1_g,
                                          dbeta_out=gv.ln1_b)
            dx, spare = dx1, dx2
            announce()   # the layer bucket, or (matrix mode) the layer's LayerNorm vectors
        return a["loss"]
