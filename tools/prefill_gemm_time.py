#!/usr/bin/env python
"""Times the prefill GEMM shapes of a model (dev tool): M tokens x the four projections."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from byzantine_consensus_llm_agents_b200 import lib as L  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 12288
d, ff, qkv = 5120, 17408, 7168
dev = "cuda"
shapes = [("qkv", qkv, d, L.EPI_F32), ("o", d, d, L.EPI_F32), ("gate_up+swiglu", 2 * ff, d, L.EPI_SWIGLU),
          ("down", d, ff, L.EPI_F32)]
tot_ms, tot_fl = 0.0, 0.0
for name, N, K, epi in shapes:
    x = (torch.randn(M, K, device=dev) * 0.1).to(torch.bfloat16)
    w = (torch.randn(N, K, device=dev) * 0.02).to(torch.bfloat16)
    out = torch.empty(M, N // 2 if epi == L.EPI_SWIGLU else N, device=dev,
                      dtype=torch.bfloat16 if epi == L.EPI_SWIGLU else torch.float32)
    for _ in range(3):
        L.gemm(x, w, out, epi)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        L.gemm(x, w, out, epi)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    fl = 2.0 * M * N * K
    tot_ms += ms
    tot_fl += fl
    print(f"{name:16s} M={M} N={N} K={K}: {ms:7.3f} ms  {fl / ms / 1e9:7.1f} TFLOP/s")
print(f"layer total {tot_ms:.3f} ms  {tot_fl / tot_ms / 1e9:.1f} TFLOP/s")
