#!/usr/bin/env python
"""Dev tool: run one phase's recorded prompts (forced "{" + blanks, 3 steps) cold and again from the
prefix cache, eagerly with the per-step logits kept, and report rows whose logits are not finite or
whose first step ends 'dead'.     python tools/debug_dead_rows.py --model mistral-22b --honest 16 --byzantine 4 [--layers 4]
"""
import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import bench as B  # noqa: E402
from byzantine_consensus_llm_agents_b200 import vllm_agent as VA  # noqa: E402
from byzantine_consensus_llm_agents_b200.engine import Request  # noqa: E402
from byzantine_consensus_llm_agents_b200.model import PRESETS, ModelConfig, Weights  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="mistral-22b")
    ap.add_argument("--honest", type=int, default=16)
    ap.add_argument("--byzantine", type=int, default=4)
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--tp", type=int, default=1)
    a = ap.parse_args()
    fam = "mistral" if a.model.startswith("mistral") else "qwen3"
    rounds = B.load_rounds(a.honest, a.byzantine, fam)
    name = B.HF_NAME.get(a.model, a.model)
    if a.layers:
        full = PRESETS[a.model]
        cfg = ModelConfig(**{**full.__dict__, "n_layers": a.layers, "name": f"{a.model}-L{a.layers}"})
        name = f"debug/{cfg.name}-{fam}" + ("-mistral" if fam == "mistral" else "-qwen3")
        VA.register_model(name, cfg, lambda c, dev: Weights.random_init(c, dev, seed=0))
    mc = {"max_model_len": 8192, "gpu_memory_utilization": 0.9, "tensor_parallel_size": a.tp, "max_num_seqs": 4,
          "device": "cuda:0", "seed": 1000, "engine_max_rows": a.honest + a.byzantine, "pdl": True}
    agent = VA.VLLMAgent("agent_0", name, mc)
    eng = agent.llm.engine
    gen = agent.llm.tp.generate if agent.llm.tp is not None else eng.generate
    tk = eng.tk
    forced = tk.encode("{", False) + tk.encode(" ", False) * 2
    for phase in ("decide", "vote"):
        for rnd in (0, 1, 0):          # cold, another round, then the first again (prefix-cache hits)
            reqs = [Request(agent._format_chat_prompt(s, u), sc, 0.5, 3, forced=forced)
                    for s, u, sc in rounds[rnd][phase]["prompts"]]
            eng.debug_logits = []
            res = gen(reqs)
            lg = torch.stack(eng.debug_logits)                   # [steps, B, V]
            eng.debug_logits = None
            bad = (~torch.isfinite(lg)).flatten(2).any(-1)       # [steps, B]
            print(phase, "round", rnd, "tokens", [len(r.token_ids) for r in res], "finish",
                  sorted(set(r.finish for r in res)), "cached", [r.cached_tokens for r in res][:6], flush=True)
            for st in range(lg.shape[0]):
                rows = bad[st].nonzero().flatten().tolist()
                if rows:
                    print(f"   step {st}: non-finite logits in rows {rows}", flush=True)
            for b, r in enumerate(res):
                if r.finish == "dead" or len(r.token_ids) != 3:
                    row = lg[0, b]
                    print(f"   row {b}: finish {r.finish} tokens {r.token_ids} chosen {r.chosen} prompt {r.prompt_tokens} "
                          f"cached {r.cached_tokens} logits[0] min {row.min().item():.3f} max {row.max().item():.3f} "
                          f"nan {int(torch.isnan(row).sum())}", flush=True)
    VA.VLLMAgent.shutdown()


if __name__ == "__main__":
    main()
