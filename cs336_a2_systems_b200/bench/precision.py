"""Mixed-precision demonstrations (reference ``cs336_systems/precision.py`` and
``mixed_precision_testing.py``; handout §"mixed precision"):

* ``accumulation_demo``  — summing 1000 × 0.01 with fp32 / fp16 accumulators and fp16 addends shows why
  accumulators (and our kernels' softmax / norm / GEMM accumulation) stay in fp32;
* ``autocast_dtypes``    — a toy MLP + LayerNorm run under ``torch.autocast`` reports the dtype of
  parameters, intermediate activations, logits, loss and gradients.

    python -m cs336_a2_systems_b200.bench.precision
"""
from __future__ import annotations

import json

import torch
import torch.nn as nn


def accumulation_demo() -> dict:
    out = {}
    s = torch.tensor(0, dtype=torch.float32)
    for _ in range(1000):
        s += torch.tensor(0.01, dtype=torch.float32)
    out["fp32_acc_fp32_add"] = float(s)
    s = torch.tensor(0, dtype=torch.float16)
    for _ in range(1000):
        s += torch.tensor(0.01, dtype=torch.float16)
    out["fp16_acc_fp16_add"] = float(s)
    s = torch.tensor(0, dtype=torch.float32)
    for _ in range(1000):
        s += torch.tensor(0.01, dtype=torch.float16)
    out["fp32_acc_fp16_add"] = float(s)
    s = torch.tensor(0, dtype=torch.float32)
    for _ in range(1000):
        s += torch.tensor(0.01, dtype=torch.float16).type(torch.float32)
    out["fp32_acc_fp16_add_upcast"] = float(s)
    return out


class ToyModel(nn.Module):
    def __init__(self, in_features: int, out_features: int):
        super().__init__()
        self.fc1 = nn.Linear(in_features, 10, bias=False)
        self.ln = nn.LayerNorm(10)
        self.fc2 = nn.Linear(10, out_features, bias=False)
        self.relu = nn.ReLU()

    def forward(self, x, record=None):
        a = self.relu(self.fc1(x))
        b = self.ln(a)
        c = self.fc2(b)
        if record is not None:
            record.update(fc1_out=str(a.dtype), layernorm_out=str(b.dtype), logits=str(c.dtype))
        return c


def autocast_dtypes(device: str | None = None, dtype=torch.bfloat16) -> dict:
    device = device or ("cuda" if torch.cuda.is_available() else "cpu")
    m = ToyModel(16, 4).to(device)
    x = torch.randn(8, 16, device=device)
    y = torch.randint(0, 4, (8,), device=device)
    rec = {}
    with torch.autocast(device_type=device, dtype=dtype):
        logits = m(x, rec)
        loss = nn.functional.cross_entropy(logits, y)
    loss.backward()
    rec.update(parameters=str(m.fc1.weight.dtype), loss=str(loss.dtype), gradients=str(m.fc1.weight.grad.dtype))
    return rec


def main():
    print(json.dumps({"accumulation": accumulation_demo(), "autocast": autocast_dtypes()}, indent=1))


if __name__ == "__main__":
    main()
