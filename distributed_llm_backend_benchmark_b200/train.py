"""Public training API: ``Trainer(config).fit(steps)``.

This is the call a user makes to train the reference's model data-parallel on B200s (BASELINE.json
config 4).  One process per GPU; rank / world size come from the launcher's environment (``RANK``,
``LOCAL_RANK``, ``WORLD_SIZE``) or are passed explicitly.  Every ``fit`` step pulls the batch from
pinned host memory through the double-buffered ``DeviceFeeder`` (host->device copy inside the step)
and reads the scalar loss back to the host.
"""
from __future__ import annotations

import os

import torch

from .models.llm import LLMConfig, LLMEngine
from .parallel.dp import DataParallelTrainer
from .utils.data import DeviceFeeder, create_dataset_from_config


class Trainer:
    def __init__(self, config: dict, device=None, rank: int | None = None, world_size: int | None = None,
                 backend: str = "native", lr: float = 1e-4, overlap: bool = True, seed: int | None = 1234):
        self.config = config
        self.rank = int(os.environ.get("RANK", "0")) if rank is None else rank
        self.world = int(os.environ.get("WORLD_SIZE", "1")) if world_size is None else world_size
        if device is None:
            device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))) if torch.cuda.is_available() \
                else torch.device("cpu")
        self.device = torch.device(device)
        self.backend = backend
        self.model_cfg = LLMConfig.from_dict(config)
        # every data-parallel rank holds the same weights (same seed) and sees different data
        self.dataset = create_dataset_from_config(config, rank_offset=self.rank, with_target=True,
                                                  pin_memory=self.device.type == "cuda")
        self.batch_size = config["input"]["batch_size"]
        if backend == "native":
            comm, pbuf, gbuf = None, None, None
            if self.world > 1:
                from .parallel.symm import SymmComm
                comm = SymmComm.from_env(self.device)
                pbuf, gbuf = comm.alloc_dp_buffers(self.model_cfg.num_parameters(1))
            self.comm = comm
            self.engine = LLMEngine(self.model_cfg, self.device, seed=seed, training=True, param_buffer=pbuf,
                                    grad_buffer=gbuf)
            self.dp = DataParallelTrainer(self.engine, lr=lr, comm=comm, overlap=overlap)
            self._step = self.dp.step
            if comm is not None:
                # Ranks finish weight initialisation at different times (a cold page cache can cost one of them
                # tens of seconds); meet on the host so that the first device-side flag wait starts with every
                # rank present - those waits have a watchdog.
                torch.cuda.synchronize(self.device)
                comm.host_barrier()
        elif backend == "torch-nccl":
            import importlib.util
            import pathlib
            p = pathlib.Path(__file__).resolve().parent.parent / "baseline" / "torch_nccl.py"
            spec = importlib.util.spec_from_file_location("dlbb_baseline_torch_nccl", p)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            self.ref = mod.TorchTrainer(self.model_cfg, self.device, self.world, lr=lr, seed=seed)
            self._step = self.ref.step
        else:
            raise ValueError(f"unknown backend {backend!r}")
        self.feeder = None

    # ------------------------------------------------------------------ device-resident path
    def resident_batch(self):
        """The rank's batch and target already on the device (for kernel-only timing)."""
        return (self.dataset.get_batch().to(self.device, non_blocking=True),
                self.dataset.get_target().to(self.device, non_blocking=True))

    def step(self, x: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        """One optimizer step on device tensors; returns the device loss tensor (no host sync)."""
        return self._step(x, target)

    # ------------------------------------------------------------------ end-to-end path
    def fit(self, steps: int) -> dict:
        """Train for ``steps`` steps from host data.  Returns per-step byte counts and the losses."""
        if self.feeder is None:
            self.feeder = DeviceFeeder(self.dataset, self.device, with_target=True)
        losses = []
        for _ in range(steps):
            x, t = self.feeder.next()           # H2D of this step's inputs (prefetched one step ahead)
            loss = self._step(x, t)
            losses.append(float(loss.item()))    # D2H read of the step's result
        return {"losses": losses, "last_loss": losses[-1] if losses else None,
                "h2d_bytes_per_step": self.feeder.bytes_per_step, "d2h_bytes_per_step": 4,
                "samples_per_step": self.batch_size * self.world}

    # ------------------------------------------------------------------ checkpoint / resume
    def save_checkpoint(self, path: str) -> None:
        """Every rank writes its own file (`<path>.rank<r>`): the optimizer state is sharded over the ranks."""
        if self.backend != "native":
            raise NotImplementedError("checkpointing is implemented for the native backend")
        sd = self.dp.state_dict()
        torch.save({k: (v.cpu() if isinstance(v, torch.Tensor) else v) for k, v in sd.items()}
                   | {"world_size": self.world, "config": self.config}, f"{path}.rank{self.rank}")

    def load_checkpoint(self, path: str) -> None:
        sd = torch.load(f"{path}.rank{self.rank}", map_location=self.device, weights_only=False)
        if sd.get("world_size") != self.world:
            raise ValueError(f"checkpoint was written by {sd.get('world_size')} ranks, this job has {self.world}")
        self.dp.load_state_dict(sd)
        if self.comm is not None:
            self.comm.barrier()

    def flops_per_step(self) -> float:
        """FLOPs actually executed per rank per step (forward + backward without the structurally
        zero 2/3 of the QKV backward, unless DLBB_DENSE_QKV_BWD=1)."""
        c = self.model_cfg
        tokens = self.batch_size * self.config["input"]["sequence_length"]
        fwd = c.forward_flops(tokens)
        bwd = 2.0 * fwd
        if self.backend == "native" and not self.engine.dense_qkv_backward:
            bwd -= 2 * 2.0 * tokens * c.num_layers * c.hidden_size * 2 * c.hidden_size
        return fwd + bwd
