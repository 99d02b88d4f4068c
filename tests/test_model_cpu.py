"""The hand-scheduled forward / backward of the LLM engine against an independent autograd
implementation of the reference math (models.py:147-190, 224-237) on the CPU."""
import torch

from distributed_llm_backend_benchmark_b200.models.llm import LLMConfig, LLMEngine, MODEL_CONFIGS
from distributed_llm_backend_benchmark_b200.parallel.dp import DataParallelTrainer


def reference_forward(x, eng, H):
    """Plain-PyTorch restatement of the reference forward in fp64 with autograd."""
    F = torch.nn.functional
    params = []
    for lv in eng.layers:
        p = {k: getattr(lv, k).double().clone().requires_grad_(True)
             for k in ("wqkv", "wo", "wup", "wdown", "ln1_g", "ln1_b", "ln2_g", "ln2_b")}
        params.append(p)
    lnf_g = eng.lnf_g.double().clone().requires_grad_(True)
    lnf_b = eng.lnf_b.double().clone().requires_grad_(True)
    h = x.double()
    for p in params:
        r = h
        h = F.layer_norm(h, (H,), p["ln1_g"], p["ln1_b"], 1e-5)
        qkv = h @ p["wqkv"]
        h = qkv[..., :H] @ p["wo"] + r
        r = h
        h = F.layer_norm(h, (H,), p["ln2_g"], p["ln2_b"], 1e-5)
        h = F.gelu(h @ p["wup"]) @ p["wdown"] + r
    y = F.layer_norm(h, (H,), lnf_g, lnf_b, 1e-5)
    return y, params, (lnf_g, lnf_b)


def small_engine(training, scale=0.05, seed=0):
    cfg = LLMConfig(hidden_size=64, num_layers=2, num_heads=4, ffn_intermediate=128)
    eng = LLMEngine(cfg, "cpu", seed=seed, training=training)
    # the reference's unscaled randn makes bf16 comparisons meaningless on a tiny model; shrink
    for lv in eng.layers:
        for w in (lv.wqkv, lv.wo, lv.wup, lv.wdown):
            w.mul_(scale)
    return cfg, eng


def test_presets_match_reference():
    assert MODEL_CONFIGS["7B"] == dict(hidden_size=4096, num_layers=32, num_heads=32, ffn_intermediate=16384)
    assert LLMConfig.from_size("7B").num_parameters() == 32 * (4096 * 12288 + 4096 * 4096 + 2 * 4096 * 16384 + 4 * 4096) + 2 * 4096
    assert abs(LLMConfig.from_size("7B").forward_flops(4096) - 52.8e12) < 0.1e12


def test_forward_matches_reference_math():
    cfg, eng = small_engine(training=False)
    x = torch.randn(2, 8, cfg.hidden_size).bfloat16()
    y = eng.forward(x)
    yr, _, _ = reference_forward(x, eng, cfg.hidden_size)
    assert y.shape == x.shape and y.dtype == torch.bfloat16
    err = (y.double() - yr).norm() / yr.norm()
    assert err < 2e-2, err


def test_backward_matches_autograd():
    cfg, eng = small_engine(training=True)
    H = cfg.hidden_size
    x = torch.randn(2, 8, H).bfloat16()
    t = torch.randn(2, 8, H).bfloat16()
    ready = []
    loss = eng.forward_backward(x, t, on_bucket_ready=ready.append)
    assert ready == list(range(len(eng.bucket_slices())))   # every bucket announced once, in completion order
    yr, params, (lg, lb) = reference_forward(x, eng, H)
    lr = ((yr - t.double()) ** 2).mean()
    lr.backward()
    assert abs(loss.item() - lr.item()) / lr.item() < 2e-2

    def close(got, ref, tol=6e-2):
        e = (got.double() - ref).norm() / (ref.norm() + 1e-12)
        assert e < tol, e

    for gv, p in zip(eng.glayers, params):
        close(gv.wdown, p["wdown"].grad)
        close(gv.wup, p["wup"].grad)
        close(gv.wo, p["wo"].grad)
        close(gv.wqkv, p["wqkv"].grad)           # includes the structurally-zero 2/3
        assert gv.wqkv[:, H:].abs().max() == 0
        for k in ("ln1_g", "ln1_b", "ln2_g", "ln2_b"):
            close(getattr(gv, k), p[k].grad)
    close(eng.glnf_g, lg.grad)
    close(eng.glnf_b, lb.grad)


def test_dense_qkv_backward_is_identical(monkeypatch):
    torch.manual_seed(3)
    cfg, eng = small_engine(training=True)
    x = torch.randn(1, 8, cfg.hidden_size).bfloat16()
    t = torch.randn(1, 8, cfg.hidden_size).bfloat16()
    eng.forward_backward(x, t)
    g_sparse = eng.grads.clone()
    monkeypatch.setenv("DLBB_DENSE_QKV_BWD", "1")
    cfg2, eng2 = small_engine(training=True)
    assert eng2.dense_qkv_backward
    eng2.forward_backward(x, t)
    # same math, different summation order inside the wider GEMMs
    err = (g_sparse.float() - eng2.grads.float()).norm() / g_sparse.float().norm()
    assert err < 1e-2, err


def test_trainer_reduces_loss_and_matches_torch_adam():
    cfg, eng = small_engine(training=True)
    H = cfg.hidden_size
    x = torch.randn(2, 8, H).bfloat16()
    t = (torch.randn(2, 8, H) * 0.1).bfloat16()
    tr = DataParallelTrainer(eng, lr=1e-3)
    p0 = eng.params.float().clone()
    l0 = tr.step(x, t).item()
    # first Adam step moves every parameter with a non-zero gradient by ~lr
    g = eng.grads.float()
    moved = (eng.params.float() - p0)
    nz = g != 0
    big = g.abs() > 1e-4   # well above eps: |update| == lr on the first step
    assert torch.allclose(moved[big].abs(), torch.full_like(moved[big], 1e-3), rtol=0.3, atol=2e-3)
    assert moved[~nz].abs().max() == 0
    losses = [l0] + [tr.step(x, t).item() for _ in range(5)]
    assert losses[-1] < losses[0]


def test_trainer_checkpoint_roundtrip(tmp_path):
    from distributed_llm_backend_benchmark_b200.train import Trainer
    cfg = {"experiment": {"name": "t", "output_dir": str(tmp_path)},
           "model": {"size": "t", "hidden_size": 64, "num_layers": 2, "num_heads": 4, "ffn_intermediate": 128},
           "parallelism": {"world_size": 1, "cores_per_rank": 1}, "input": {"batch_size": 2, "sequence_length": 8, "seed": 3},
           "execution": {"warmup_iterations": 1, "benchmark_iterations": 1}, "system": {}}
    a = Trainer(cfg, device="cpu", rank=0, world_size=1, lr=1e-3)
    a.fit(steps=2)
    a.save_checkpoint(str(tmp_path / "ck"))
    b = Trainer(cfg, device="cpu", rank=0, world_size=1, lr=1e-3, seed=999)    # different init
    b.load_checkpoint(str(tmp_path / "ck"))
    la, lb = a.fit(steps=2)["losses"], b.fit(steps=2)["losses"]
    assert la == lb and b.dp.step_count == 4
