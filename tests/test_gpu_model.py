"""The device engine (native kernels) against the CPU engine (PyTorch oracle) on the same weights."""
import pytest
import torch

from distributed_llm_backend_benchmark_b200 import _lib
from distributed_llm_backend_benchmark_b200.models.llm import LLMConfig, LLMEngine
from distributed_llm_backend_benchmark_b200.parallel.dp import DataParallelTrainer

pytestmark = pytest.mark.gpu


def _pair(training, H=256, L=2, F=512):
    cfg = LLMConfig(hidden_size=H, num_layers=L, num_heads=4, ffn_intermediate=F)
    cpu = LLMEngine(cfg, "cpu", seed=0, training=training)
    for lv in cpu.layers:
        for w in (lv.wqkv, lv.wo, lv.wup, lv.wdown):
            w.mul_(0.03)
    gpu = LLMEngine(cfg, "cuda:0", seed=0, training=training)
    gpu.params.copy_(cpu.params)
    return cfg, cpu, gpu


def _err(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


def test_forward_matches_cpu_oracle(native_lib):
    cfg, cpu, gpu = _pair(False)
    x = torch.randn(2, 96, cfg.hidden_size).bfloat16()
    n0 = _lib.launch_count()
    y = gpu.forward(x.cuda())
    assert _lib.launch_count() > n0, "forward did not run on the native kernels"
    assert _err(y, cpu.forward(x)) < 2e-2


def test_backward_matches_cpu_oracle(native_lib):
    cfg, cpu, gpu = _pair(True)
    x = torch.randn(2, 96, cfg.hidden_size).bfloat16()
    t = torch.randn(2, 96, cfg.hidden_size).bfloat16()
    lg = gpu.forward_backward(x.cuda(), t.cuda())
    lc = cpu.forward_backward(x, t)
    assert abs(lg.item() - lc.item()) / lc.item() < 2e-2
    assert _err(gpu.grads, cpu.grads) < 5e-2
    for g, c in zip(gpu.glayers, cpu.glayers):
        for k in ("wqkv", "wo", "wup", "wdown", "ln1_g", "ln2_b"):
            assert _err(getattr(g, k), getattr(c, k)) < 6e-2, k


@pytest.mark.parametrize("overlap", [False, True])
def test_training_steps_match_cpu(native_lib, overlap):
    cfg, cpu, gpu = _pair(True)
    x = torch.randn(2, 64, cfg.hidden_size).bfloat16()
    t = (torch.randn(2, 64, cfg.hidden_size) * 0.1).bfloat16()
    tg = DataParallelTrainer(gpu, lr=1e-3, overlap=overlap)
    tc = DataParallelTrainer(cpu, lr=1e-3)
    xg, tgt = x.cuda(), t.cuda()
    lg = [tg.step(xg, tgt).item() for _ in range(4)]
    lc = [tc.step(x, t).item() for _ in range(4)]
    assert lg[-1] < lg[0]
    for a, b in zip(lg, lc):
        assert abs(a - b) / b < 5e-2
