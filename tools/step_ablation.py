#!/usr/bin/env python
"""In-pipeline cost of each kernel family of the decode step (dev tool).

Times the real decode step (CUDA graph, optional PDL) and the same step with one kernel family
left out (bcg_model_decode profiling flags); the difference is what that family costs INSIDE the
pipeline, overlap included — unlike isolated kernel timings.
    python tools/step_ablation.py [--model qwen3-14b] [--rows 10] [--ctx 1600] [--pdl 1]
"""
import argparse
import ctypes as C
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from byzantine_consensus_llm_agents_b200 import lib as L  # noqa: E402
from byzantine_consensus_llm_agents_b200.engine import Engine  # noqa: E402
from byzantine_consensus_llm_agents_b200.model import PRESETS, Weights  # noqa: E402
from byzantine_consensus_llm_agents_b200.tokenizer import get_tokenizer  # noqa: E402

SKIPS = {"attention": 0x100, "add_rmsnorm": 0x200, "qknorm_rope_kvwrite": 0x400, "qkv GEMM": 0x1000,
         "o GEMM": 0x2000, "gate_up GEMM": 0x4000, "down GEMM": 0x8000, "lm_head GEMM": 0x10000}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="qwen3-14b")
    ap.add_argument("--rows", type=int, default=10)
    ap.add_argument("--ctx", type=int, default=1600)
    ap.add_argument("--pdl", type=int, default=1)
    ap.add_argument("--sustain-steps", type=int, default=1500)
    ap.add_argument("--max-splits", type=int, default=0, help="attention grid width (default: what --ctx needs)")
    ap.add_argument("--period-only", action="store_true", help="skip the per-kernel ablation")
    ap.add_argument("--base-flags", type=lambda x: int(x, 0), default=0,
                    help="extra bcg_model_decode flags for every run, e.g. 0x20000 = norm folded into the GEMMs")
    a = ap.parse_args()
    cfg = PRESETS[a.model]
    dev = torch.device("cuda:0")
    eng = Engine(cfg, Weights.random_init(cfg, dev), get_tokenizer(cfg.family, cfg.vocab), max_model_len=4096,
                 max_rows=max(16, a.rows), prefill_tokens=4096, pdl=bool(a.pdl))
    B, r = a.rows, eng.row
    pages = (a.ctx + 15) // 16
    bt = torch.arange(B * pages, dtype=torch.int32, device=dev).view(B, pages)
    r["block_table"][:B, :pages] = bt
    max_splits = a.max_splits or (a.ctx + 255) // 256

    def reset():
        r["seq_lens"][:B] = a.ctx
        r["positions"][:B] = a.ctx - 1
        r["slot_map"][:B] = (bt[:, -1] * 16 + (a.ctx - 1) % 16)
        r["next_ids"][:B] = 17
        eng.buf["attn_counters"].zero_()

    def step(extra):
        eng.ws.max_splits = max_splits
        L.check(eng.lib.bcg_model_decode(C.byref(eng.model.struct), C.byref(eng.ws), r["next_ids"].data_ptr(),
                                         r["positions"].data_ptr(), r["slot_map"].data_ptr(),
                                         r["seq_lens"].data_ptr(), r["block_table"].data_ptr(), eng.max_blocks, B,
                                         eng.flags | extra | a.base_flags, L.stream_ptr()))

    def time_step(extra, reps=30):
        reset()
        step(extra)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            step(extra)
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    peak = json.loads((Path(__file__).resolve().parents[1] / "MEASURED_PEAKS.json").read_text())["hbm_gbs"]
    full = time_step(0)
    nbytes = cfg.decode_weight_bytes() + B * a.ctx * cfg.kv_bytes_per_token
    print(f"{a.model} rows={B} ctx={a.ctx} pdl={a.pdl}: full step {full:.3f} ms "
          f"({nbytes / full / 1e6:.0f} GB/s = {100 * nbytes / full / 1e6 / peak:.1f} % of {peak} GB/s); "
          f"weights-only floor {cfg.decode_weight_bytes() / peak / 1e6:.3f} ms")
    out = {"full_ms": full}

    # the real per-token period: decode step + sampler, k tokens per graph, graphs replayed back to back
    def token_period(k, reps=24):
        reset()
        r["grammar_id"][:B] = -1
        r["max_tokens"][:B] = 1 << 20
        r["temperature"][:B] = 0.5
        eng._decode_step(B, max_splits, False)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(k):
                eng._decode_step(B, max_splits, False)
        g.replay()
        torch.cuda.synchronize()
        reset()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps // k):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / (reps // k * k)

    for k in (1, 2, 4, 8):
        out[f"token_period_ms_k{k}"] = token_period(k)
        print(f"  step + sampler, {k} token(s) per graph, graphs back to back: {out[f'token_period_ms_k{k}']:.3f} ms per token")
    # the same, replayed for several seconds: the clocks and the memory system settle where a
    # whole game round runs (the short loops above finish inside the power-cap averaging window)
    def sustained(n):
        reset()
        eng._decode_step(B, max_splits, False)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            eng._decode_step(B, max_splits, False)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n // 24):
            reset()                      # rows stay at --ctx and inside their output buffers
            r["n_out"][:B] = 0
            r["finished"][:B] = 0
            for _ in range(24):
                g.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / (n // 24 * 24)

    out["token_period_ms_sustained"] = sustained(a.sustain_steps)
    print(f"  step + sampler, {a.sustain_steps} replays back to back: {out['token_period_ms_sustained']:.3f} ms per token")
    if a.period_only:
        print(json.dumps(out))
        return
    for name, bit in SKIPS.items():
        t = time_step(bit)
        out[name] = full - t
        per = (full - t) * 1e3 / (1 if name.startswith("lm_head") else cfg.n_layers)
        print(f"  without {name:22s} {t:7.3f} ms   -> costs {full - t:6.3f} ms/step = {per:6.2f} us/launch-group")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
