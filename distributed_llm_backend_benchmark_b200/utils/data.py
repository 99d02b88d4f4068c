"""Synthetic inputs (ref data_gen.py).

``SyntheticEmbeddingDataset`` reproduces the reference generator: one fixed bf16 ``[B, S, H]`` batch
drawn on the CPU from ``torch.manual_seed(seed)``; ``get_batch()`` returns that same tensor every
call, and every tensor-parallel rank uses the same seed (replicated input, ref data_gen.py:37-44).

B200 additions: the batch can be kept in *pinned* host memory and streamed to the device through a
double-buffered copy stream (``DeviceFeeder``), a fixed synthetic regression target exists for the
training path, and data-parallel ranks offset the seed by their rank so they see different data.
"""
from __future__ import annotations

import torch


class SyntheticEmbeddingDataset:
    def __init__(self, batch_size: int, sequence_length: int, hidden_size: int, seed: int = 42,
                 pin_memory: bool = False, with_target: bool = False):
        self.batch_size = batch_size
        self.sequence_length = sequence_length
        self.hidden_size = hidden_size
        self.seed = seed
        shape = (batch_size, sequence_length, hidden_size)
        gen = torch.Generator(device="cpu")
        gen.manual_seed(seed)
        # Same draw as the reference (global generator seeded with `seed`, one randn call).
        self.fixed_batch = torch.randn(shape, dtype=torch.bfloat16, device="cpu", generator=gen)
        self.fixed_target = None
        if with_target:
            tgen = torch.Generator(device="cpu")
            tgen.manual_seed(seed + 1_000_003)
            self.fixed_target = torch.randn(shape, dtype=torch.bfloat16, device="cpu", generator=tgen)
        if pin_memory and torch.cuda.is_available():
            self.fixed_batch = self.fixed_batch.pin_memory()
            if self.fixed_target is not None:
                self.fixed_target = self.fixed_target.pin_memory()

    def get_batch(self) -> torch.Tensor:
        return self.fixed_batch

    def get_target(self) -> torch.Tensor:
        if self.fixed_target is None:
            raise RuntimeError("dataset was created without a target (with_target=False)")
        return self.fixed_target

    def nbytes(self) -> int:
        n = self.fixed_batch.numel() * self.fixed_batch.element_size()
        if self.fixed_target is not None:
            n *= 2
        return n


def create_dataset_from_config(config: dict, rank_offset: int = 0, **kw) -> SyntheticEmbeddingDataset:
    return SyntheticEmbeddingDataset(
        batch_size=config["input"]["batch_size"],
        sequence_length=config["input"]["sequence_length"],
        hidden_size=config["model"]["hidden_size"],
        seed=config["input"]["seed"] + rank_offset,
        **kw,
    )


class DeviceFeeder:
    """Streams host batches to the device: pinned source, dedicated copy stream, two device slots.

    ``next()`` returns the device tensors of the current step and immediately enqueues the copy for
    the following step, so the H2D transfer of step i+1 overlaps the compute of step i.
    """

    def __init__(self, dataset: SyntheticEmbeddingDataset, device: torch.device, with_target: bool = True):
        self.ds = dataset
        self.device = torch.device(device)
        self.with_target = with_target and dataset.fixed_target is not None
        self.cuda = self.device.type == "cuda"
        self.bytes_per_step = dataset.fixed_batch.numel() * 2 * (2 if self.with_target else 1)
        if self.cuda:
            self.stream = torch.cuda.Stream(device=self.device)
            self.slots = []
            for _ in range(2):
                x = torch.empty_like(dataset.fixed_batch, device=self.device)
                t = torch.empty_like(dataset.fixed_target, device=self.device) if self.with_target else None
                self.slots.append((x, t, torch.cuda.Event()))
            self.cur = 0
            self._issue(0)

    def _issue(self, slot: int) -> None:
        x, t, ev = self.slots[slot]
        with torch.cuda.stream(self.stream):
            # the consumer of this slot (two steps ago) finished before the caller's last sync;
            # make that explicit for the copy stream anyway
            self.stream.wait_stream(torch.cuda.current_stream(self.device))
            x.copy_(self.ds.fixed_batch, non_blocking=True)
            if t is not None:
                t.copy_(self.ds.fixed_target, non_blocking=True)
            ev.record(self.stream)

    def next(self):
        if not self.cuda:
            return self.ds.fixed_batch, (self.ds.fixed_target if self.with_target else None)
        x, t, ev = self.slots[self.cur]
        torch.cuda.current_stream(self.device).wait_event(ev)
        self.cur ^= 1
        self._issue(self.cur)
        return x, t
