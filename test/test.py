"""Single-process engine smoke test: ZeRO-0, no optimizer section, one step (ref test/test.py).
    python test/test.py"""
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (_ROOT, os.path.join(_ROOT, "compat")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import deepspeed  # noqa: E402

model = torch.nn.Linear(10, 1)
config = {"train_batch_size": 1, "gradient_accumulation_steps": 1, "fp16": {"enabled": False},
          "zero_optimization": {"stage": 0}}
engine, optimizer, _, _ = deepspeed.initialize(model=model, model_parameters=model.parameters(), config=config)
loss = engine(torch.randn(1, 10)).mean()
engine.backward(loss)
engine.step()
print("SUCCESS: engine initialised and ran a training step in a single process")
