#!/usr/bin/env python
"""Per-kernel decode-step timing with CUDA events (dev tool; feeds the roofline table in DESIGN.md).

Each kernel is launched back to back over all layers' weights / KV (inputs >> L2), so the average
is a warm-pipeline, cold-data figure — the regime inside the decode CUDA graph.
    python tools/kernel_times.py [--model qwen3-14b] [--rows 10] [--ctx 1600]
"""
import argparse
import ctypes as C
import json
import math
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from byzantine_consensus_llm_agents_b200 import lib as L  # noqa: E402
from byzantine_consensus_llm_agents_b200.engine import Engine  # noqa: E402
from byzantine_consensus_llm_agents_b200.model import PRESETS, Weights  # noqa: E402
from byzantine_consensus_llm_agents_b200.tokenizer import get_tokenizer  # noqa: E402


def timeit(fn, n_layers, reps=5):
    """Average device time of fn(layer), measured over a CUDA graph of n_layers launches so that
    Python / ctypes launch cost (5-15 us, more than the small kernels take) stays out of it."""
    for l in range(min(3, n_layers)):
        fn(l)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    inner = max(1, 40 // n_layers)
    with torch.cuda.graph(g):
        for _ in range(inner):
            for l in range(n_layers):
                fn(l)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * n_layers * inner)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="qwen3-14b")
    ap.add_argument("--rows", type=int, default=10)
    ap.add_argument("--ctx", type=int, default=1600)
    a = ap.parse_args()
    peak = json.loads((Path(__file__).resolve().parents[1] / "MEASURED_PEAKS.json").read_text())["hbm_gbs"]
    cfg = PRESETS[a.model]
    dev = torch.device("cuda:0")
    w = Weights.random_init(cfg, dev)
    tk = get_tokenizer(cfg.family, cfg.vocab)
    eng = Engine(cfg, w, tk, max_model_len=4096, max_rows=max(16, a.rows), prefill_tokens=4096)
    lib, B, d, nl = eng.lib, a.rows, cfg.d_model, cfg.n_layers
    buf, r = eng.buf, eng.row
    st = L.stream_ptr
    # a plausible batch: ctx tokens per row, pages laid out row after row
    pages = (a.ctx + 15) // 16
    bt = torch.arange(B * pages, dtype=torch.int32, device=dev).view(B, pages)
    r["block_table"][:B, :pages] = bt
    r["seq_lens"][:B] = a.ctx
    r["positions"][:B] = a.ctx - 1
    r["slot_map"][:B] = (bt[:, -1] * 16 + (a.ctx - 1) % 16)
    for kv in eng.kv_layers:
        kv[:B * pages].normal_()
    buf["xn"].normal_()
    buf["attn"].normal_()
    buf["act"].normal_()
    max_splits = (a.ctx + 255) // 256
    qkv_dim, hq, hkv, ff = cfg.qkv_dim, cfg.n_heads, cfg.n_kv_heads, cfg.d_ff
    s_qkv, s_o = L.splits_for(B, qkv_dim, d), L.splits_for(B, d, hq * 128)
    s_gu, s_dn = L.splits_for(B, 2 * ff, d), L.splits_for(B, d, ff)
    P = buf["partials"]
    out = {}

    def gemm(x, wl, n, k, splits, epi, o):
        return lambda l: L.check(lib.bcg_gemm_bf16_tn(x.data_ptr(), wl[l].data_ptr(), o.data_ptr(), B, n, k, k, k,
                                                      n // 2 if epi == L.EPI_SWIGLU else n, epi, splits, 0, st()))

    rows = [
        ("qkv GEMM", gemm(buf["xn"], w.w_qkv, qkv_dim, d, s_qkv, L.EPI_F32, P), qkv_dim * d * 2, s_qkv),
        ("o GEMM", gemm(buf["attn"], w.w_o, d, hq * 128, s_o, L.EPI_F32, P), d * hq * 128 * 2, s_o),
        ("gate_up GEMM", gemm(buf["xn"], w.w_gate_up, 2 * ff, d, s_gu, L.EPI_SWIGLU if s_gu == 1 else L.EPI_F32,
                              buf["act"] if s_gu == 1 else P), 2 * ff * d * 2, s_gu),
        ("down GEMM", gemm(buf["act"], w.w_down, d, ff, s_dn, L.EPI_F32, P), d * ff * 2, s_dn),
        ("add_rmsnorm", lambda l: L.check(lib.bcg_add_rmsnorm(buf["residual"].data_ptr(), P.data_ptr(), 1, s_dn, B * d,
                                                              w.ln1[l].data_ptr(), buf["xn"].data_ptr(), None, B, d,
                                                              1e-6, st())), B * d * (8 + 2 + 4 * s_dn), s_dn),
        ("qknorm_rope_kvwrite", lambda l: L.check(lib.bcg_qknorm_rope_kvwrite(
            P.data_ptr(), 1, s_qkv, B * qkv_dim, qkv_dim, L.ptr(w.q_norm[l]) if cfg.qk_norm else None,
            L.ptr(w.k_norm[l]) if cfg.qk_norm else None, w.cos_sin.data_ptr(), r["positions"].data_ptr(),
            r["slot_map"].data_ptr(), buf["q"].data_ptr(), eng.kv_layers[l].data_ptr(), B, hq, hkv, 16, 1e-6, st())),
         B * qkv_dim * (4 * s_qkv + 2), s_qkv),
        ("decode_attn (+merge)", lambda l: L.check(lib.bcg_decode_attn_paged(
            buf["q"].data_ptr(), eng.kv_layers[l].data_ptr(), r["block_table"].data_ptr(), eng.max_blocks,
            r["seq_lens"].data_ptr(), buf["attn"].data_ptr(), buf["so"].data_ptr(), buf["sml"].data_ptr(),
            buf["attn_counters"].data_ptr(), B, hq, hkv,
            16, max_splits, 1.0 / math.sqrt(128), st())), B * a.ctx * hkv * 128 * 2 * 2, max_splits),
    ]
    total = 0.0
    print(f"{a.model}  rows={B} ctx={a.ctx}   (ideal = algorithmic bytes / {peak} GB/s)")
    for name, fn, nbytes, splits in rows:
        us = timeit(fn, nl)
        ideal = nbytes / peak / 1e3
        mult = 2 if name == "add_rmsnorm" else 1
        total += us * mult
        out[name] = {"us": round(us, 2), "ideal_us": round(ideal, 2), "GBs": round(nbytes / us / 1e3, 1), "splits": splits}
        print(f"  {name:24s} {us:8.2f} us  ideal {ideal:7.2f} us  {nbytes / us / 1e3:8.1f} GB/s  "
              f"({100 * ideal / us:5.1f} %)  splits/chunks={splits}")
    lm = timeit(lambda l: L.check(lib.bcg_gemm_bf16_tn(buf["xn_last"].data_ptr(), w.lm_head.data_ptr(),
                                                        buf["logits"].data_ptr(), B, cfg.vocab, d, d, d, cfg.vocab,
                                                        L.EPI_F32, 1, 0, st())), 1, reps=20)
    lm_bytes = cfg.vocab * d * 2
    print(f"  {'lm_head GEMM':24s} {lm:8.2f} us  ideal {lm_bytes / peak / 1e3:7.2f} us  {lm_bytes / lm / 1e3:8.1f} GB/s")
    r["grammar_id"][:B] = -1
    r["max_tokens"][:B] = 1 << 20
    r["temperature"][:B] = 0.5
    samp = timeit(lambda l: eng._sample(B, False), 1, reps=20)
    print(f"  {'mask_sample (+count)':24s} {samp:8.2f} us")
    step = total * nl + lm + samp
    print(f"  sum over a step (isolated kernels, no overlap): {step / 1e3:.3f} ms; "
          f"weights-only floor {cfg.decode_weight_bytes() / peak / 1e6:.3f} ms")
    out["lm_head"] = {"us": round(lm, 2)}
    out["sampler"] = {"us": round(samp, 2)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
