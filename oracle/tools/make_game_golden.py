"""Generates tests/golden/game_*.json.gz: a full seeded greedy game of the UNMODIFIED reference
played against the CPU oracle engine (transformers fp32 + xgrammar + oracle sampler), recorded
call by call at the `LLM.generate` boundary.

Because the game logic is deterministic given the LLM outputs, replaying the recorded calls in
order through another engine and getting the same outputs implies the same transcript
(proposed values, votes, termination round).  The GPU parity test does exactly that
(tests/test_gpu_game_replay.py) — the reference itself is not available on the GPU box.

    python -m oracle.tools.make_game_golden --model qwen3-0.6b --honest 4 --byzantine 0 \
        --rounds 5 --awareness none_exist --out tests/golden/game_qwen3_0p6b_4h0b.json.gz
"""
from __future__ import annotations

import argparse
import gzip
import json
import time
from pathlib import Path

from byzantine_consensus_llm_agents_b200.model import PRESETS, ModelConfig
from byzantine_consensus_llm_agents_b200.tokenizer import get_tokenizer
from oracle import ref_harness as H
from oracle.engine import OracleEngine
from oracle.model import build_hf_model

HF_NAME = {"qwen3-0.6b": "Qwen/Qwen3-0.6B", "qwen3-8b": "Qwen/Qwen3-8B", "qwen3-14b": "Qwen/Qwen3-14B",
           "mistral-22b": "mistralai/Mistral-Small-Instruct-2409"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="qwen3-0.6b")
    ap.add_argument("--layers", type=int, default=None, help="reduced depth (full width)")
    ap.add_argument("--honest", type=int, default=4)
    ap.add_argument("--byzantine", type=int, default=0)
    ap.add_argument("--rounds", type=int, default=5)
    ap.add_argument("--awareness", default="none_exist")
    ap.add_argument("--game-seed", type=int, default=1234)
    ap.add_argument("--weight-seed", type=int, default=0)
    ap.add_argument("--temperature", default="greedy", help="'greedy' or 'shipped' (0.5 / 0.3)")
    ap.add_argument("--max-tokens-decide", type=int, default=None,
                    help="override LLM_CONFIG['max_tokens_decide'] (config.py:52-58; data change only)")
    ap.add_argument("--max-tokens-vote", type=int, default=None)
    ap.add_argument("--out", required=True)
    a = ap.parse_args()

    cfg = PRESETS[a.model]
    if a.layers:
        cfg = ModelConfig(**{**cfg.__dict__, "n_layers": a.layers, "name": f"{cfg.name}-L{a.layers}"})
    tk = get_tokenizer(cfg.family, cfg.vocab)
    t0 = time.time()
    hf = build_hf_model(cfg, seed=a.weight_seed)
    eng = OracleEngine(hf, tk, base_seed=0)
    print(f"model built in {time.time() - t0:.1f}s", flush=True)
    H.install_vllm_shim(eng)
    ref = H.import_reference()
    t0 = time.time()
    stats, transcript, sim = H.run_game(ref, a.honest, a.byzantine, a.rounds, seed=a.game_seed,
                                        greedy=(a.temperature == "greedy"), awareness=a.awareness,
                                        model_name=HF_NAME.get(a.model, a.model),
                                        llm_overrides={k: v for k, v in (("max_tokens_decide", a.max_tokens_decide),
                                                                         ("max_tokens_vote", a.max_tokens_vote))
                                                       if v is not None})
    wall = time.time() - t0
    # de-duplicate identical (prompt, schema, max_tokens, temperature) generations
    uniq, calls = {}, []
    for ci, c in enumerate(H.ShimLLM.calls):
        rows = []
        for ri, (p, text, extra) in enumerate(zip(c.prompts, c.texts, c.extra)):
            # greedy generations are a function of the request; sampled ones also of their seed
            key = json.dumps([p, c.schema, c.max_tokens, c.temperature,
                              None if c.temperature < 1e-5 else [ci, ri]])
            if key not in uniq:
                uniq[key] = {"id": len(uniq), "prompt": p, "schema": c.schema, "max_tokens": c.max_tokens,
                             "temperature": c.temperature, "text": text, "tokens": extra["tokens"],
                             "top1": [round(x, 5) for x in extra["top1"]],
                             "top2": [round(x, 5) for x in extra["top2"]],
                             "finish": extra["finish"], "prompt_tokens": extra["prompt_ids"],
                             "seed": extra["seed"]}
            rows.append(uniq[key]["id"])
        calls.append({"rows": rows, "temperature": c.temperature, "max_tokens": c.max_tokens})
    keep = ["total_rounds", "consensus_reached", "final_consensus_value", "honest_agents_count",
            "byzantine_agents_count", "termination_reason", "consensus_outcome"]
    out = {
        "meta": {"model": a.model, "layers": cfg.n_layers, "vocab": cfg.vocab, "honest": a.honest,
                 "byzantine": a.byzantine, "rounds": a.rounds, "awareness": a.awareness,
                 "game_seed": a.game_seed, "weight_seed": a.weight_seed, "temperature": a.temperature,
                 "max_tokens_decide": a.max_tokens_decide, "max_tokens_vote": a.max_tokens_vote,
                 "oracle_wall_s": round(wall, 1), "oracle_stats": eng.stats,
                 "generator": "oracle/tools/make_game_golden.py"},
        "requests": sorted(uniq.values(), key=lambda r: r["id"]),
        "calls": calls,
        "transcript": transcript,
        "statistics": {k: stats.get(k) for k in keep if k in stats},
    }
    Path(a.out).parent.mkdir(parents=True, exist_ok=True)
    with gzip.open(a.out, "wt", compresslevel=9) as f:
        json.dump(out, f, separators=(",", ":"), default=str)
    print(f"wrote {a.out}: {len(calls)} calls, {len(uniq)} unique generations, game wall {wall:.0f}s; "
          f"oracle stats {eng.stats}")


if __name__ == "__main__":
    main()
