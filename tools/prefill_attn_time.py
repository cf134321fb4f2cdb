#!/usr/bin/env python
"""Prefill-attention throughput (dev tool): bcg_prefill_attn_paged on the bench's prefill group shape
(7 prompts x ~1,770 tokens, 40 query / 8 KV heads) and a few others, useful causal FLOPs / time.
    python tools/prefill_attn_time.py
"""
import json
import math
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from byzantine_consensus_llm_agents_b200 import lib as L  # noqa: E402

DEV = "cuda:0"


def run_case(lib, H, HKV, lens, cached=0, tile=128, reps=10):
    page = 16
    tot = [cached + n for n in lens]
    blocks = [(n + page - 1) // page for n in tot]
    cache = torch.randn(sum(blocks) + 1, 2, HKV, page, 128, device=DEV, dtype=torch.bfloat16)
    bt = torch.zeros(len(lens), max(blocks), dtype=torch.int32)
    nxt = 0
    for b, nb in enumerate(blocks):
        bt[b, :nb] = torch.arange(nxt, nxt + nb, dtype=torch.int32)
        nxt += nb
    bt = bt.to(DEV)
    T = sum(lens)
    q = torch.randn(T, H * 128, device=DEV, dtype=torch.bfloat16)
    o = torch.empty(T, H * 128, device=DEV, dtype=torch.bfloat16)
    tiles, row = [], 0
    for b, n in enumerate(lens):
        for r0 in range(0, n, tile):
            tiles.append([b, row + r0, cached + r0, min(tile, n - r0)])
        row += n
    ti = torch.tensor(tiles, device=DEV, dtype=torch.int32)

    def launch():
        L.check(lib.bcg_prefill_attn_paged(q.data_ptr(), cache.data_ptr(), bt.data_ptr(), bt.shape[1],
                                           ti.data_ptr(), len(tiles), o.data_ptr(), T, H, HKV, 16,
                                           1.0 / math.sqrt(128), L.stream_ptr()))
    for _ in range(3):
        launch()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        launch()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / reps
    flops = sum(4.0 * H * 128 * (n * cached + n * (n + 1) / 2) for n in lens)
    return {"H": H, "HKV": HKV, "lens": f"{len(lens)}x{lens[0]}", "cached": cached, "tiles": len(tiles),
            "us": round(us, 1), "TFLOPs": round(flops / us / 1e6, 1)}


def main():
    import os
    lib = L.load()
    tile = 128
    out = []
    for H, HKV, lens, cached in ((40, 8, [1770] * 7, 0), (40, 8, [1296] * 10, 416), (40, 8, [4096] * 3, 0),
                                 (32, 8, [1770] * 8, 0), (48, 8, [1900] * 6, 0), (64, 8, [1770] * 4, 0)):
        r = run_case(lib, H, HKV, lens, cached, tile)
        out.append(r)
        print(json.dumps(r), flush=True)
    if len(sys.argv) > 1:
        Path(sys.argv[1]).write_text(json.dumps(out, indent=1) + "\n")


if __name__ == "__main__":
    main()
