"""One data-parallel optimizer step through the DeepSpeed-style engine on the Gloo process group
(counterpart of the reference's test/ds_mpi_test.py: ZeRO stage 0, Adam, a 10 -> 1 linear model).

    mpirun -n 2 python test/ds_mpi_test.py        # or plain `python test/ds_mpi_test.py` for one rank

Checks more than the reference does: the step must move the weights, and after it every rank must hold the
same weights (the gradients were averaged before Adam ran)."""
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [p for p in (os.path.dirname(_HERE), os.path.join(os.path.dirname(_HERE), "compat")) if p not in sys.path]

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import deepspeed  # noqa: E402

DS_CONFIG = {
    "train_micro_batch_size_per_gpu": 1,
    "gradient_accumulation_steps": 1,
    "optimizer": {"type": "Adam", "params": {"lr": 1e-3}},
    "zero_optimization": {"stage": 0},
}


class TinyRegressor(torch.nn.Module):
    def __init__(self, features: int = 10):
        super().__init__()
        self.linear = torch.nn.Linear(features, 1)

    def forward(self, x):
        return self.linear(x)


def run_one_step() -> None:
    for key, default in (("MASTER_ADDR", "127.0.0.1"), ("MASTER_PORT", "29500"), ("RANK", "0"), ("WORLD_SIZE", "1")):
        os.environ.setdefault(key, default)
    dist.init_process_group(backend="gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    print(f"[ds_mpi_test] rank {rank} of {world} is up")
    try:
        torch.manual_seed(0)                       # same initial weights everywhere
        net = TinyRegressor()
        engine, _optimizer, _, _ = deepspeed.initialize(model=net, model_parameters=net.parameters(), config=DS_CONFIG)
        w0 = net.linear.weight.detach().clone()
        torch.manual_seed(100 + rank)              # different data per rank
        loss = engine(torch.randn(1, 10)).mean()
        engine.backward(loss)
        engine.step()
        w1 = net.linear.weight.detach()
        assert not torch.equal(w0, w1), "the optimizer step left the weights unchanged"
        gathered = [torch.empty_like(w1) for _ in range(world)]
        dist.all_gather(gathered, w1.clone())
        assert all(torch.allclose(g, gathered[0]) for g in gathered), "ranks diverged after one averaged step"
        print(f"[ds_mpi_test] rank {rank}: one ZeRO-0 + Adam step done, weights agree on {world} rank(s)")
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    run_one_step()
