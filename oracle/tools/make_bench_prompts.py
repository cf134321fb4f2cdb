"""Generates tests/golden/bench_prompts_*.json.gz: the (system, user, schema) triples the
UNMODIFIED reference hands to `batch_generate_json` (main.py:305, 412) in steady-state rounds.

bench.py replays them — the reference itself cannot travel to the GPU box.  The game is driven
through the vllm shim by a scripted backend that answers with schema-valid JSON of the length
the reference's truncation rules cap at (400-char strategy, 600-char reasoning),
always votes "continue" so the game runs to its round limit, and never triggers retries; the
prompts are therefore exactly what a well-behaved model would make the reference build.

    python -m oracle.tools.make_bench_prompts --honest 8 --byzantine 2 --rounds 22 --first 6
"""
from __future__ import annotations

import argparse
import gzip
import json
import random
import re
from pathlib import Path

from oracle import ref_harness as H

MODEL = {"qwen3": "Qwen/Qwen3-14B", "mistral": "mistralai/Mistral-Small-Instruct-2409"}


class Scripted:
    def __init__(self, seed):
        self.rng = random.Random(seed)

    def _text(self, prompt, chars):
        words = re.findall(r"[A-Za-z']+", prompt)
        out = []
        n = 0
        while n < chars:
            w = self.rng.choice(words)
            out.append(w)
            n += len(w) + 1
        return (" ".join(out))[:chars].capitalize()

    def __call__(self, prompts, sp):
        schema = sp.guided_decoding.json
        outs = []
        for p in prompts:
            if "decision" in schema["properties"]:
                outs.append(json.dumps({"decision": "continue"}))
            else:
                outs.append(json.dumps({"internal_strategy": self._text(p, 420),
                                        "value": self.rng.randint(0, 50),
                                        "public_reasoning": self._text(p, 620)}))
        return outs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--honest", type=int, default=8)
    ap.add_argument("--byzantine", type=int, default=2)
    ap.add_argument("--rounds", type=int, default=22)
    ap.add_argument("--first", type=int, default=6, help="first recorded round (steady state from 6)")
    ap.add_argument("--family", default="qwen3")
    ap.add_argument("--awareness", default="may_exist")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()

    H.install_vllm_shim(Scripted(7))
    ref = H.import_reference()
    captured = []
    orig = ref.vllm_agent.VLLMAgent.batch_generate_json

    def spy(self, prompts, temperature=0.8, max_tokens=512):
        captured.append({"prompts": [[s, u, sc] for s, u, sc in prompts], "temperature": temperature,
                         "max_tokens": max_tokens})
        return orig(self, prompts, temperature=temperature, max_tokens=max_tokens)

    ref.vllm_agent.VLLMAgent.batch_generate_json = spy
    H.run_game(ref, a.honest, a.byzantine, a.rounds, seed=1234, awareness=a.awareness,
               model_name=MODEL[a.family])
    # two batched calls per round (decide, vote) because the scripted model never fails
    assert len(captured) == 2 * a.rounds, (len(captured), a.rounds)
    rounds = []
    for r in range(a.first - 1, a.rounds):
        rounds.append({"round": r + 1, "decide": captured[2 * r], "vote": captured[2 * r + 1]})
    out = a.out or f"tests/golden/bench_prompts_{a.family}_{a.honest}h{a.byzantine}b.json.gz"
    meta = {"honest": a.honest, "byzantine": a.byzantine, "family": a.family, "model_name": MODEL[a.family],
            "awareness": a.awareness, "first_round": a.first, "rounds": len(rounds),
            "generator": "oracle/tools/make_bench_prompts.py"}
    Path(out).parent.mkdir(parents=True, exist_ok=True)
    with gzip.open(out, "wt", compresslevel=9) as f:
        json.dump({"meta": meta, "rounds": rounds}, f, separators=(",", ":"))
    chars = [len(s) + len(u) for rd in rounds for s, u, _ in rd["decide"]["prompts"]]
    print(f"wrote {out}: {len(rounds)} rounds, decide prompt chars min/mean/max = "
          f"{min(chars)}/{sum(chars) // len(chars)}/{max(chars)}")


if __name__ == "__main__":
    main()
