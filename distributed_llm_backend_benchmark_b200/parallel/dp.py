"""Data-parallel training engine for the LLM (BASELINE.json config 4).

The reference only trains toy models through DeepSpeed (ZeRO-0/2 + Adam: test/ccl.py:74-115,
test/ds_mpi_test.py:13-47).  Here the same contract - average gradients over ranks, Adam step,
identical parameters everywhere afterwards - runs on the real model with one kernel per gradient
bucket that does *all* of: cross-GPU reduction of the bucket (peer loads over NVLink), 1/N scale,
Adam on this rank's shard (fp32 moments + fp32 master, sharded 1/N per rank) and broadcast of the
updated bf16 parameters into every peer's parameter buffer.  Buckets are the model's per-layer
flat slices; each is launched on a side stream as soon as the layer's backward has been enqueued,
so communication + optimizer overlap the backward GEMMs of the earlier layers.

With one GPU the same schedule degenerates to a bucketed Adam overlapped with backward.
"""
from __future__ import annotations

import torch

from .. import ops
from ..utils.trace import nvtx_range
from ..models.llm import LLMEngine


class DataParallelTrainer:
    def __init__(self, engine: LLMEngine, lr: float = 1e-4, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, comm=None, overlap: bool = True, master_weights: bool = True):
        assert engine.training, "engine must be created with training=True"
        self.engine = engine
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.comm = comm
        self.world = 1 if comm is None else comm.world_size
        self.rank = 0 if comm is None else comm.rank
        self.step_count = 0
        dev = engine.device
        self.cuda = dev.type == "cuda"
        import os as _os
        # Overlap pays when the bucket kernel also communicates (world > 1: NVLink time hides under the
        # backward GEMMs).  On one GPU the optimizer and the GEMMs compete for the same HBM bandwidth and
        # power budget and co-running gains nothing (profiles/optimizer_overlap.md), so it runs after backward.
        default_overlap = "1" if (comm is not None and comm.world_size > 1) else "0"
        self.overlap = overlap and self.cuda and _os.environ.get("DLBB_DP_OVERLAP", default_overlap) == "1"
        self._overlap_ctas = int(_os.environ.get("DLBB_ADAM_OVERLAP_CTAS", "148"))
        self.skip_optimizer = _os.environ.get("DLBB_DEBUG_SKIP_OPT", "0") == "1"   # debugging aid, invalidates results
        if engine.bucket_mode is None:
            # One bucket per layer.  Per-matrix buckets (DLBB_DP_BUCKETS=matrix: the exchange of dW_down starts while dW_up
            # is still being computed) were measured and lost on 8 GPUs: every bucket kernel begins and ends with a barrier
            # across all ranks, 161 instead of 33 of them per step cost more in waiting CTAs beside the GEMMs than the
            # earlier start gains (139.7 ms / step vs 131.2; on 2 GPUs 141.7 vs ~143.5) - profiles/step_breakdown.md.
            engine.bucket_mode = "layer"
        self.buckets = engine.bucket_slices()
        # Optimizer state is sharded: rank r owns elements [r*shard, (r+1)*shard) of every bucket
        # (buckets are padded up to a multiple of world*8 elements on the comm side).
        self.shards = []
        for off, n in self.buckets:
            per = self._shard_len(n)
            lo = min(self.rank * per, n)
            hi = min(lo + per, n)
            self.shards.append((off + lo, hi - lo))
        total = sum(n for _, n in self.shards)
        self.m = torch.zeros(total, dtype=torch.float32, device=dev)
        self.v = torch.zeros(total, dtype=torch.float32, device=dev)
        self.master = None
        if master_weights:
            self.master = torch.empty(total, dtype=torch.float32, device=dev)
            o = 0
            for off, n in self.shards:
                self.master[o:o + n].copy_(engine.params[off:off + n])
                o += n
        self.state_off = []
        o = 0
        for _, n in self.shards:
            self.state_off.append(o)
            o += n
        if self.cuda:
            # low priority: at every GEMM kernel boundary the GEMM's CTAs are placed first and the
            # optimizer fills what is left (one small CTA per SM beside the resident GEMM CTA)
            self.side = torch.cuda.Stream(device=dev, priority=0)
            self.sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
            self.ready = [torch.cuda.Event() for _ in self.buckets]
            self.done = torch.cuda.Event()
        if comm is not None:
            comm.register_dp_buffers(engine.params, engine.grads)

    def _shard_len(self, n: int) -> int:
        if self.world == 1:
            return n
        q = self.world * 8
        padded = (n + q - 1) // q * q
        return padded // self.world

    # ------------------------------------------------------------------ optimizer on one bucket
    def _update_bucket(self, i: int) -> None:
        if self.skip_optimizer:
            return
        eng = self.engine
        off, n = self.buckets[i]
        soff, sn = self.shards[i]
        so = self.state_off[i]
        b1, b2 = self.betas
        if self.comm is None or self.world == 1:
            ops.adam_step(eng.params[off:off + n], eng.grads[off:off + n], self.m[so:so + n], self.v[so:so + n],
                          None if self.master is None else self.master[so:so + n], lr=self.lr, beta1=b1, beta2=b2,
                          eps=self.eps, weight_decay=self.weight_decay, grad_scale=1.0, step=self.step_count)
        else:
            self.comm.fused_reduce_adam(off, n, self.m[so:so + sn], self.v[so:so + sn],
                                        None if self.master is None else self.master[so:so + sn], lr=self.lr,
                                        beta1=b1, beta2=b2, eps=self.eps, weight_decay=self.weight_decay,
                                        grad_scale=1.0 / self.world, step=self.step_count)

    def _on_bucket_ready(self, i: int) -> None:
        if not self.overlap:
            return
        cur = torch.cuda.current_stream(self.engine.device)
        self.ready[i].record(cur)
        last = i == len(self.buckets) - 1
        with torch.cuda.stream(self.side):
            self.side.wait_event(self.ready[i])
            if self.comm is None or self.world == 1:
                from .. import _lib
                # while backward is still running: one CTA per SM; the final bucket has the GPU to itself
                _lib.load().dlbb_adam_set_max_ctas(0 if last else int(self._overlap_ctas))
                self._update_bucket(i)
                _lib.load().dlbb_adam_set_max_ctas(0)
            else:
                self._update_bucket(i)

    # ------------------------------------------------------------------ public step
    def step(self, x: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        """forward + backward + gradient averaging + Adam.  Returns the local loss (fp32 [1], on device)."""
        self.step_count += 1
        eng = self.engine
        with nvtx_range(f"dp_step_{self.step_count}"):
            loss = eng.forward_backward(x, target, on_bucket_ready=self._on_bucket_ready if self.overlap else None)
        if self.overlap:
            self.done.record(self.side)
            torch.cuda.current_stream(eng.device).wait_event(self.done)
        else:
            for i in range(len(self.buckets)):
                self._update_bucket(i)
        return loss

    def state_dict(self) -> dict:
        return {"step": self.step_count, "m": self.m, "v": self.v, "master": self.master,
                "params": self.engine.params}

    def load_state_dict(self, sd: dict) -> None:
        self.step_count = int(sd["step"])
        self.m.copy_(sd["m"])
        self.v.copy_(sd["v"])
        if self.master is not None and sd.get("master") is not None:
            self.master.copy_(sd["master"])
        self.engine.params.copy_(sd["params"])
