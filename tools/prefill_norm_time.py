#!/usr/bin/env python
"""Times bcg_add_rmsnorm at a prefill shape (dev tool): T rows x d, one fp32 delta slab."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from byzantine_consensus_llm_agents_b200 import lib as L  # noqa: E402

T, d = int(sys.argv[1]) if len(sys.argv) > 1 else 12400, int(sys.argv[2]) if len(sys.argv) > 2 else 5120
lib = L.load()
dev = "cuda"
res = torch.randn(T, d, device=dev)
delta = torch.randn(T, d, device=dev)
w = torch.ones(d, device=dev)
xn = torch.empty(T, d, device=dev, dtype=torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def run():
    L.check(lib.bcg_add_rmsnorm(res.data_ptr(), delta.data_ptr(), 1, 1, T * d, w.data_ptr(), xn.data_ptr(), None,
                                T, d, 1e-6, L.stream_ptr()))


for _ in range(3):
    run()
ts = []
for _ in range(10):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run()
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e3)
us = sorted(ts)[len(ts) // 2]
nbytes = T * d * (4 + 4 + 4 + 2)
print(f"add_rmsnorm T={T} d={d}: {us:.1f} us, {nbytes / us / 1e3:.0f} GB/s (res r/w + delta r + xn w)")
