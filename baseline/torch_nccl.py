"""Context baseline: the reference model's math in stock PyTorch (cuBLAS GEMMs, ATen LayerNorm /
GELU, autograd, torch.optim.Adam) with DistributedDataParallel over NCCL - the "reference on the
NCCL/cuBLAS stack" bar of SURVEY.md 6.4.  None of this framework's kernels run on this path.

It is NOT the `--impl reference` arm (that must be the unmodified reference, which cannot run here:
see DESIGN.md); it exists so that every number of ours has a same-box library number beside it.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class _Block(nn.Module):
    def __init__(self, H: int, Fi: int):
        super().__init__()
        bf = torch.bfloat16
        self.H = H
        self.ln1 = nn.LayerNorm(H, dtype=bf)
        self.wqkv = nn.Parameter(torch.randn(H, 3 * H, dtype=bf))
        self.wo = nn.Parameter(torch.randn(H, H, dtype=bf))
        self.ln2 = nn.LayerNorm(H, dtype=bf)
        self.wup = nn.Parameter(torch.randn(H, Fi, dtype=bf))
        self.wdown = nn.Parameter(torch.randn(Fi, H, dtype=bf))

    def forward(self, x):
        r = x
        qkv = torch.matmul(self.ln1(x), self.wqkv)
        x = torch.matmul(qkv[:, :, :self.H], self.wo) + r
        r = x
        x = torch.matmul(F.gelu(torch.matmul(self.ln2(x), self.wup)), self.wdown) + r
        return x


class TorchLLM(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.layers = nn.ModuleList([_Block(cfg.hidden_size, cfg.ffn_intermediate) for _ in range(cfg.num_layers)])
        self.ln_final = nn.LayerNorm(cfg.hidden_size, dtype=torch.bfloat16)

    def forward(self, x):
        for layer in self.layers:
            x = layer(x)
        return self.ln_final(x)


class TorchTrainer:
    def __init__(self, cfg, device, world: int, lr: float = 1e-4, seed: int | None = 1234):
        if seed is not None:
            torch.manual_seed(seed)
        with torch.device(device):
            self.model = TorchLLM(cfg)
        self.world = world
        self.net = self.model
        if world > 1:
            from torch.nn.parallel import DistributedDataParallel as DDP
            self.net = DDP(self.model, device_ids=[device.index], bucket_cap_mb=400, gradient_as_bucket_view=True)
        # DLBB_TORCH_STATE=bf16 (default): torch.optim.Adam straight on the bf16 parameters - moments in bf16, no master
        # copy (half the optimizer bytes of the native engine).  DLBB_TORCH_STATE=fp32: the native engine's contract in
        # stock PyTorch - fp32 master weights and fp32 moments, stepped one layer at a time so that the fp32 copy of the
        # gradients never exists for more than one layer (the whole model would not fit beside the activations).
        import os
        self.state = os.environ.get("DLBB_TORCH_STATE", "bf16")
        if self.state == "fp32":
            self.groups = []
            mods = list(self.model.layers) + [self.model.ln_final]
            for mod in mods:
                ps = list(mod.parameters())
                masters = [p.detach().float().clone() for p in ps]
                self.groups.append((ps, masters, torch.optim.Adam(masters, lr=lr, fused=True)))
        else:
            self.opt = torch.optim.Adam(self.model.parameters(), lr=lr, fused=True)

    def step(self, x, target):
        self.model.zero_grad(set_to_none=True)
        y = self.net(x)
        loss = F.mse_loss(y.float(), target.float())
        loss.backward()
        if self.state == "fp32":
            for ps, masters, opt in self.groups:
                for p, m in zip(ps, masters):
                    m.grad = p.grad.float()
                opt.step()
                torch._foreach_copy_(ps, masters)
                for m in masters:
                    m.grad = None
        else:
            self.opt.step()
        return loss.detach().reshape(1)
