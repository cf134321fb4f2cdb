#!/usr/bin/env python
"""Times the sm_100 incumbents the reference reaches through vLLM (SURVEY.md §2.1 "incumbent"
column) next to this repo's kernels, on the bench's shapes, on the same GPU in the same process:

  K3/K9-K12  cuBLAS bf16 (torch.matmul)          vs bcg_gemm_bf16_tn     decode M=10 and prefill M=12,390
  K8         FlashInfer paged decode (trtllm-gen) vs bcg_decode_attn_paged  B=10, S=1600, 40:8 heads
  K7         FlashInfer paged prefill (trtllm-gen) vs bcg_prefill_attn_paged  7 x 1,770 tokens, causal
  K13-K15    xgrammar apply_token_bitmask_inplace + torch softmax / exponential-race argmax
                                                   vs bcg_mask_sample      [10, 151936]

Timing: CUDA events around a CUDA graph of back-to-back launches over DISTINCT weight / KV buffers
(decode shapes: inputs >> L2), after warm-up.  A section that cannot run here (FlashInfer needs a
JIT build or a cubin that is not in the image, ...) is recorded as {"unavailable": "<error text>"}.

    python tools/incumbent_kernels.py --out gpurun_out/incumbent_kernels.json
"""
import argparse
import json
import math
import signal
import sys
import time
import traceback
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from byzantine_consensus_llm_agents_b200 import lib as L  # noqa: E402

DEV = "cuda:0"


class _Timeout(Exception):
    pass


def _alarm(_sig, _frm):
    raise _Timeout("section exceeded its time limit")


def graph_time(fn, n_distinct, reps=5, inner=None):
    """Average device microseconds of fn(i), i cycling over n_distinct inputs, from a CUDA graph."""
    for i in range(min(3, n_distinct)):
        fn(i)
    torch.cuda.synchronize()
    inner = inner or max(1, 40 // n_distinct)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(inner):
            for i in range(n_distinct):
                fn(i)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * n_distinct * inner)


def eager_time(fn, n_distinct, reps=3):
    for i in range(min(2, n_distinct)):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        for i in range(n_distinct):
            fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * n_distinct)


def section(results, name, limit_s=240):
    def deco(fn):
        signal.signal(signal.SIGALRM, _alarm)
        signal.alarm(limit_s)
        t0 = time.time()
        try:
            results[name] = fn()
        except BaseException as e:  # noqa: BLE001 - record and go on
            results[name] = {"unavailable": f"{type(e).__name__}: {e}"[:600],
                             "trace": traceback.format_exc()[-800:]}
        finally:
            signal.alarm(0)
        results[name]["wall_s"] = round(time.time() - t0, 1)
        print(name, json.dumps(results[name])[:400], flush=True)
        try:
            torch.cuda.synchronize()
            torch.cuda.empty_cache()
        except Exception:
            pass
        return fn
    return deco


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/incumbent_kernels.json")
    ap.add_argument("--skip-flashinfer", action="store_true")
    a = ap.parse_args()
    torch.cuda.set_device(0)
    lib = L.load()
    res = {"gpu": torch.cuda.get_device_name(0), "torch": torch.__version__}
    d, ff, V = 5120, 17408, 151936
    shapes = {"qkv": (7168, d), "o": (d, d), "gate_up": (2 * ff, d), "down": (d, ff)}

    # ---------------------------------------------------------------- GEMMs, decode (M = 10)
    @section(res, "gemm_decode_M10")
    def _():
        out = {}
        M = 10
        for name, (N, K) in list(shapes.items()) + [("lm_head", (V, d))]:
            n_w = 2 if name == "lm_head" else max(2, int(2.0e9 // (N * K * 2)))   # ~2 GB of weights >> L2
            ws = [torch.randn(N, K, device=DEV, dtype=torch.bfloat16) * 0.02 for _ in range(n_w)]
            x = torch.randn(M, K, device=DEV, dtype=torch.bfloat16)
            y = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
            t_cublas = graph_time(lambda i: torch.matmul(x, ws[i].t(), out=y), n_w)
            splits = L.splits_for(M, N, K)
            if splits == 1:
                ours_out = torch.empty(M, N, device=DEV, dtype=torch.float32 if name == "lm_head" else torch.bfloat16)
                epi = L.EPI_F32 if name == "lm_head" else L.EPI_BF16
            else:
                ours_out = torch.empty(splits, M, N, device=DEV, dtype=torch.float32)
                epi = L.EPI_F32
            t_ours = graph_time(lambda i: L.gemm(x, ws[i], ours_out, epi, splits=splits), n_w)
            gb = N * K * 2 / 1e9
            out[name] = {"N": N, "K": K, "splits": splits, "cublas_us": round(t_cublas, 2),
                         "bcg_us": round(t_ours, 2), "bcg_over_cublas": round(t_ours / t_cublas, 3),
                         "cublas_GBs": round(gb / t_cublas * 1e6, 0), "bcg_GBs": round(gb / t_ours * 1e6, 0),
                         "note": "bcg writes fp32 split-K partials that its next kernel sums; cuBLAS writes bf16"
                         if splits > 1 else ""}
            del ws
        return out

    # ---------------------------------------------------------------- GEMMs, prefill (M = 12,390)
    @section(res, "gemm_prefill_M12390")
    def _():
        out = {}
        M = 12390
        for name, (N, K) in shapes.items():
            w = torch.randn(N, K, device=DEV, dtype=torch.bfloat16) * 0.02
            x = torch.randn(M, K, device=DEV, dtype=torch.bfloat16)
            y = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
            # power-capped clocks drift over the ~0.1 s each arm runs, so the arms are interleaved
            # (cuBLAS, ours, ours, cuBLAS) and each reports its mean
            f_c = lambda i: torch.matmul(x, w.t(), out=y)
            f_o = lambda i: L.gemm(x, w, y, L.EPI_BF16)
            tc1, to1 = graph_time(f_c, 1, inner=5), graph_time(f_o, 1, inner=5)
            to2, tc2 = graph_time(f_o, 1, inner=5), graph_time(f_c, 1, inner=5)
            t_cublas, t_ours = 0.5 * (tc1 + tc2), 0.5 * (to1 + to2)
            fl = 2.0 * M * N * K
            out[name] = {"N": N, "K": K, "cublas_us": round(t_cublas, 1), "bcg_us": round(t_ours, 1),
                         "order": "cublas, bcg, bcg, cublas", "runs_us": [round(v, 1) for v in (tc1, to1, to2, tc2)],
                         "bcg_over_cublas": round(t_ours / t_cublas, 3),
                         "cublas_TFLOPs": round(fl / t_cublas / 1e6, 0), "bcg_TFLOPs": round(fl / t_ours / 1e6, 0)}
            del w, x, y
        return out

    # ---------------------------------------------------------------- paged KV shared by K7 / K8
    H, HKV, page = 40, 8, 16
    scale = 1.0 / math.sqrt(128)

    def paged(lens, n_copies):
        blocks = [(n + page - 1) // page for n in lens]
        per = sum(blocks)
        caches = [torch.randn(per + 1, 2, HKV, page, 128, device=DEV, dtype=torch.bfloat16) for _ in range(n_copies)]
        bt = torch.zeros(len(lens), max(blocks), dtype=torch.int32)
        nxt = 0
        for b, nb in enumerate(blocks):
            bt[b, :nb] = torch.arange(nxt, nxt + nb, dtype=torch.int32)
            nxt += nb
        return caches, bt.to(DEV)

    # ---------------------------------------------------------------- K8 decode attention
    dec_lens = [1600] * 10

    @section(res, "decode_attn_bcg")
    def _():
        caches, bt = paged(dec_lens, 16)          # 16 layers' worth of KV: 1.05 GB >> L2
        B = len(dec_lens)
        q = torch.randn(B, H * 128, device=DEV, dtype=torch.bfloat16)
        seq = torch.tensor(dec_lens, device=DEV, dtype=torch.int32)
        ms = (max(dec_lens) + 255) // 256
        so = torch.empty(B * H * ms * 128, device=DEV)
        sml = torch.empty(B * H * ms * 2, device=DEV)
        cnt = torch.zeros(B * HKV, device=DEV, dtype=torch.int32)
        o = torch.empty(B, H * 128, device=DEV, dtype=torch.bfloat16)

        def run(i):
            L.check(lib.bcg_decode_attn_paged(q.data_ptr(), caches[i].data_ptr(), bt.data_ptr(), bt.shape[1],
                                              seq.data_ptr(), o.data_ptr(), so.data_ptr(), sml.data_ptr(),
                                              cnt.data_ptr(), B, H, HKV, 16, ms, scale, L.stream_ptr()))
        t = graph_time(run, len(caches))
        kv_bytes = sum(dec_lens) * HKV * 128 * 2 * 2
        return {"us": round(t, 2), "GBs": round(kv_bytes / t / 1e3, 0), "kv_bytes": kv_bytes}

    if not a.skip_flashinfer:
        @section(res, "decode_attn_flashinfer_trtllm")
        def _():
            import flashinfer
            caches, bt = paged(dec_lens, 16)
            B = len(dec_lens)
            q = torch.randn(B, H, 128, device=DEV, dtype=torch.bfloat16)
            seq = torch.tensor(dec_lens, device=DEV, dtype=torch.int32)
            wsb = torch.zeros(256 << 20, device=DEV, dtype=torch.uint8)
            o = torch.empty(B, H, 128, device=DEV, dtype=torch.bfloat16)

            def run(i):
                flashinfer.decode.trtllm_batch_decode_with_kv_cache(
                    q, caches[i], wsb, bt, seq, max(dec_lens), bmm1_scale=scale, bmm2_scale=1.0, out=o,
                    kv_layout="HND")
            run(0)
            torch.cuda.synchronize()
            try:
                t = graph_time(run, len(caches))
                how = "cuda graph"
            except Exception:
                torch.cuda.synchronize()
                t = eager_time(run, len(caches))
                how = "eager launches"
            kv_bytes = sum(dec_lens) * HKV * 128 * 2 * 2
            return {"us": round(t, 2), "GBs": round(kv_bytes / t / 1e3, 0), "timed_as": how,
                    "flashinfer": flashinfer.__version__}

    # ---------------------------------------------------------------- K7 prefill attention
    pf_lens = [1770] * 7
    pf_flops = sum(4.0 * H * 128 * n * (n + 1) / 2 for n in pf_lens)

    @section(res, "prefill_attn_bcg")
    def _():
        caches, bt = paged(pf_lens, 1)
        T = sum(pf_lens)
        q = torch.randn(T, H * 128, device=DEV, dtype=torch.bfloat16)
        o = torch.empty(T, H * 128, device=DEV, dtype=torch.bfloat16)
        tiles, row = [], 0
        for b, n in enumerate(pf_lens):
            for r0 in range(0, n, 128):
                tiles.append([b, row + r0, r0, min(128, n - r0)])
            row += n
        ti = torch.tensor(tiles, device=DEV, dtype=torch.int32)

        def run(i):
            L.check(lib.bcg_prefill_attn_paged(q.data_ptr(), caches[0].data_ptr(), bt.data_ptr(), bt.shape[1],
                                               ti.data_ptr(), len(tiles), o.data_ptr(), T, H, HKV, 16, scale,
                                               L.stream_ptr()))
        t = graph_time(run, 1, inner=5)
        return {"us": round(t, 1), "TFLOPs_causal": round(pf_flops / t / 1e6, 0), "tokens": T}

    if not a.skip_flashinfer:
        @section(res, "prefill_attn_flashinfer_trtllm")
        def _():
            import flashinfer
            caches, bt = paged(pf_lens, 1)
            T, B = sum(pf_lens), len(pf_lens)
            q = torch.randn(T, H, 128, device=DEV, dtype=torch.bfloat16)
            o = torch.empty(T, H, 128, device=DEV, dtype=torch.bfloat16)
            seq = torch.tensor(pf_lens, device=DEV, dtype=torch.int32)
            cum = torch.tensor([0] + list(torch.tensor(pf_lens).cumsum(0)), device=DEV, dtype=torch.int32)
            wsb = torch.zeros(256 << 20, device=DEV, dtype=torch.uint8)

            def run(i):
                flashinfer.prefill.trtllm_batch_context_with_kv_cache(
                    q, caches[0], wsb, bt, seq, max(pf_lens), max(pf_lens), scale, 1.0, B, cum, cum, out=o,
                    kv_layout="HND")
            run(0)
            torch.cuda.synchronize()
            t = eager_time(run, 1, reps=5)
            return {"us": round(t, 1), "TFLOPs_causal": round(pf_flops / t / 1e6, 0),
                    "timed_as": "eager launches", "flashinfer": flashinfer.__version__}

    # ---------------------------------------------------------------- K13-K15 mask + sample
    @section(res, "mask_sample_torch_xgrammar")
    def _():
        import xgrammar as xgr
        B = 10
        logits = torch.randn(B, V, device=DEV)
        bitmask = torch.randint(-2 ** 31, 2 ** 31 - 1, (B, (V + 31) // 32), device=DEV, dtype=torch.int32)
        work = torch.empty_like(logits)
        qexp = torch.empty_like(logits)

        def run(i):
            work.copy_(logits)
            xgr.apply_token_bitmask_inplace(work, bitmask)
            probs = torch.softmax(work / 0.5, dim=-1, dtype=torch.float32)
            qexp.exponential_()
            return probs.div_(qexp).argmax(dim=-1)
        t = eager_time(run, 1, reps=20)
        return {"us": round(t, 1), "what": "copy + xgrammar apply_token_bitmask_inplace + softmax + "
                                           "exponential race argmax (vLLM's sampler ops), eager launches",
                "xgrammar": getattr(xgr, "__version__", "?")}

    @section(res, "mask_sample_bcg")
    def _():
        import numpy as np
        from byzantine_consensus_llm_agents_b200.engine import Engine, Request
        from byzantine_consensus_llm_agents_b200.model import ModelConfig, Weights
        from byzantine_consensus_llm_agents_b200.tokenizer import get_tokenizer
        cfg = ModelConfig("sampler-probe", n_layers=1, d_model=256, n_heads=2, n_kv_heads=1, d_ff=512, vocab=V)
        tk = get_tokenizer(cfg.family, cfg.vocab)
        eng = Engine(cfg, Weights.random_init(cfg, torch.device(DEV)), tk, max_model_len=512, max_rows=16,
                     prefill_tokens=1024, kv_pages=400)
        rng = np.random.default_rng(0)
        lo, hi = tk.regular_range
        # unguided rows so that no row finishes while the sampler is replayed (a guided vote closes
        # its JSON after ~10 tokens); the mask lookup of a guided row adds one 19 KB table read
        eng.generate([Request("", None, 0.5, 4, prompt_ids=rng.integers(lo, hi, size=20).tolist())
                      for _ in range(10)])
        eng.row["finished"][:10] = 0
        eng.row["n_out"][:10] = 0
        eng.row["max_tokens"][:10] = 500
        t = graph_time(lambda i: eng._sample(10, False), 1, inner=20)
        eng.close()
        return {"us": round(t, 2), "what": "bcg_mask_sample (+ active-row count), 10 rows, T=0.5, fp32 logits [10, 151936]"}

    Path(a.out).parent.mkdir(parents=True, exist_ok=True)
    Path(a.out).write_text(json.dumps(res, indent=1) + "\n")
    print("wrote", a.out)


if __name__ == "__main__":
    main()
