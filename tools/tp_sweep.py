#!/usr/bin/env python
"""BASELINE.json configs[4]: Qwen3-32B, tensor parallel over the GPUs of one box, agent-count sweep
N in {4, 8, 16, 32, 64} with floor(0.2 N) Byzantine agents — ONE model load, one JSON line per N
(dev / evidence tool; bench.py stays the contract-shaped benchmark).

    python tools/tp_sweep.py --tp 8 --model qwen3-32b [--counts 4,8,16,32,64] [--out file.jsonl]

Per N: one warm-up round and `--rounds` timed rounds of the faithful agent step (reference retry
schedule, recorded prompts of that agent count), then one round of the fixed schedule (all rows to
max_tokens) whose decode ms/step is the per-rank roofline figure: bytes per step and rank =
(W + rows * context * KV bytes per token) / tp.
"""
import argparse
import json
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import bench as B  # noqa: E402

SPLIT = {4: (4, 0), 8: (7, 1), 10: (8, 2), 16: (13, 3), 20: (16, 4), 32: (26, 6), 64: (52, 12)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tp", type=int, default=8)
    ap.add_argument("--model", default="qwen3-32b")
    ap.add_argument("--counts", default="4,8,16,32,64")
    ap.add_argument("--rounds", type=int, default=1)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    import torch
    from byzantine_consensus_llm_agents_b200 import vllm_agent as VA
    counts = [int(x) for x in a.counts.split(",")]
    hbm_peak, _, _ = B.load_peaks()
    fam = "mistral" if a.model.startswith("mistral") else "qwen3"
    mc = {"max_model_len": 8192, "gpu_memory_utilization": 0.9, "tensor_parallel_size": a.tp, "max_num_seqs": 4,
          "disable_qwen3_thinking": True, "device": "cuda:0", "seed": 1000, "engine_max_rows": max(counts),
          "pdl": True}
    t0 = time.perf_counter()
    agent = VA.VLLMAgent("agent_0", B.HF_NAME.get(a.model, a.model), mc)
    engine = agent.llm.engine
    torch.cuda.synchronize()
    load_s = time.perf_counter() - t0
    cfg = engine.cfg
    lines = []
    for n in counts:
        h, b = SPLIT[n]
        rounds = B.load_rounds(h, b, fam)
        cs = dict(calls=0, rows=0, ok=0, abstain=0)
        B.agent_step(agent, rounds[0], cs)                               # warm-up: graphs of these shapes
        torch.cuda.synchronize()
        cs = dict(calls=0, rows=0, ok=0, abstain=0)
        s0 = dict(engine.stats)
        t0 = time.perf_counter()
        for i in range(a.rounds):
            B.agent_step(agent, rounds[(1 + i) % len(rounds)], cs)
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        d = {k: engine.stats[k] - s0[k] for k in s0}
        line = {"model": a.model, "tp": a.tp, "agents": n, "honest": h, "byzantine": b, "rounds": a.rounds,
                "agent_steps_per_s_e2e": round(n * a.rounds / wall, 4),
                "agent_steps_per_s_device": round(n * a.rounds / max(1e-9, d["prefill_s"] + d["decode_s"]), 4),
                "engine_calls": cs["calls"], "rows": cs["rows"], "valid": cs["ok"],
                "decode_steps": d["decode_steps"], "decode_ms_per_step": round(1e3 * d["decode_s"] / max(1, d["decode_steps"]), 3),
                "rows_per_step": round(d["generated_tokens"] / max(1, d["decode_steps"]), 2),
                "prefill_s": round(d["prefill_s"], 3), "prefill_tokens": d["prefill_tokens"],
                "cached_tokens": d["cached_tokens"]}
        try:
            f0 = dict(engine.stats)
            B.fixed_round(agent, engine, rounds[0])
            torch.cuda.synchronize()
            fd = {k: engine.stats[k] - f0[k] for k in f0}
            step_s = fd["decode_s"] / max(1, fd["decode_steps"])
            ctx = fd["prompt_tokens"] / max(1, fd["prompts"]) + 0.5 * fd["generated_tokens"] / max(1, fd["prompts"])
            per_rank_bytes = (cfg.decode_weight_bytes() + n * ctx * cfg.kv_bytes_per_token) / a.tp
            line["fixed"] = {"decode_ms_per_step": round(1e3 * step_s, 3), "rows_per_step": n,
                             "avg_context": int(ctx), "bytes_per_step_per_rank": int(per_rank_bytes),
                             "GBs_per_rank": round(per_rank_bytes / step_s / 1e9, 1),
                             "frac_of_measured_hbm_peak": round(per_rank_bytes / step_s / 1e9 / hbm_peak, 4),
                             "agent_steps_per_s_device": round(n / max(1e-9, fd["prefill_s"] + fd["decode_s"]), 4)}
        except Exception as e:  # noqa: BLE001
            line["fixed"] = {"error": str(e)[:300]}
        print(json.dumps(line), flush=True)
        lines.append(line)
    print(json.dumps({"model_load_s": round(load_s, 1)}), flush=True)
    if a.out:
        Path(a.out).write_text("\n".join(json.dumps(x) for x in lines) + "\n")
    VA.VLLMAgent.shutdown()


if __name__ == "__main__":
    main()
