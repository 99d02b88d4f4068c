"""DeepSpeed-style engine on Gloo: ZeRO-0 + Adam, one step (ref test/ds_mpi_test.py).
    mpirun -n 2 python test/ds_mpi_test.py"""
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (_ROOT, os.path.join(_ROOT, "compat")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import deepspeed  # noqa: E402


class Regressor(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.linear = torch.nn.Linear(10, 1)

    def forward(self, x):
        return self.linear(x)


CONFIG = {"train_micro_batch_size_per_gpu": 1, "gradient_accumulation_steps": 1,
          "zero_optimization": {"stage": 0}, "optimizer": {"type": "Adam", "params": {"lr": 1e-3}}}


def main():
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("WORLD_SIZE", "1")
    dist.init_process_group(backend="gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    print(f"hello from rank {rank} / {world}")
    model = Regressor()
    engine, _, _, _ = deepspeed.initialize(model=model, model_parameters=model.parameters(), config=CONFIG)
    before = model.linear.weight.detach().clone()
    loss = engine(torch.randn(1, 10)).mean()
    engine.backward(loss)
    engine.step()
    assert not torch.equal(before, model.linear.weight), "step did not update the parameters"
    print(f"rank {rank}: step completed")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
