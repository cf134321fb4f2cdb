"""ORACLE tool (test infrastructure): what does the reference's OWN model deviate by in bf16?

north_star states "logits within 2e-2 abs in bf16 against an fp32 CPU forward".  That bound holds
at the 0.6B architecture; at the widths of the other BASELINE configs the random-init logits are
2-2.5x larger (std 0.02*sqrt(d_model)) and so is every bf16 rounding.  To give the per-width
tolerances an independent footing this script runs transformers' own model — the oracle's model,
unchanged code (modeling_qwen3.py / modeling_mistral.py) — twice on the CPU with the SAME
bf16-valued weights: once in fp32 (the oracle) and once after `model.to(torch.bfloat16)` (what the
reference's vLLM computes in, on a different backend).  The |dlogit| between the two is what any
bf16 execution of the reference is entitled to; the engine must not exceed it
(tests/test_gpu_model.py).  Nothing here knows where the engine rounds.

    python -m oracle.tools.make_bf16_floor            # writes tests/golden/bf16_floor.json
"""
from __future__ import annotations

import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

from byzantine_consensus_llm_agents_b200.model import PRESETS, ModelConfig
from byzantine_consensus_llm_agents_b200.tokenizer import get_tokenizer
from oracle.model import build_hf_model, full_logits

ROOT = Path(__file__).resolve().parents[2]
# (preset, layers, prompt lengths, forced steps, rng seed) — the rows tests/test_gpu_model.py runs
CASES = [
    ("qwen3-0.6b", 28, [700, 130], 6, 5),
    ("qwen3-8b", 2, [520, 77], 5, 11),
    ("qwen3-14b", 4, [520, 77], 5, 11),
    ("qwen3-32b", 2, [520, 77], 5, 11),
    ("mistral-22b", 2, [600, 77], 5, 11),
]


def case_inputs(tk, lens, n_new, seed):
    """Prompt ids and teacher-forced ids of a case (shared with the GPU test)."""
    rng = np.random.default_rng(seed)
    lo, hi = tk.regular_range
    prompts = [rng.integers(lo, hi, size=n).tolist() for n in lens]
    forced = [rng.integers(lo, hi, size=n_new).tolist() for _ in prompts]
    return prompts, forced


def main():
    out = {"generator": "oracle/tools/make_bf16_floor.py",
           "what": "transformers CPU forward in bf16 (model.to(bfloat16)) vs the same model in fp32, "
                   "identical bf16-valued weights; |dlogit| over the positions the GPU test compares",
           "torch": torch.__version__, "cases": {}}
    only = set(sys.argv[1:])
    path = ROOT / "tests" / "golden" / "bf16_floor.json"
    if only and path.exists():
        out["cases"] = json.loads(path.read_text())["cases"]
    for preset, layers, lens, n_new, seed in CASES:
        if only and preset not in only:
            continue
        full = PRESETS[preset]
        cfg = full if layers == full.n_layers else ModelConfig(
            **{**full.__dict__, "n_layers": layers, "name": f"{preset}-L{layers}"})
        tk = get_tokenizer(cfg.family, cfg.vocab)
        t0 = time.time()
        hf = build_hf_model(cfg, seed=0)
        prompts, forced = case_inputs(tk, lens, n_new, seed)
        refs = [full_logits(hf, p + f[:-1])[len(p) - 1:] for p, f in zip(prompts, forced)]
        hf = hf.to(torch.bfloat16)
        rows, sq, n, worst = [], 0.0, 0, 0.0
        for p, f, ref in zip(prompts, forced, refs):
            got = full_logits(hf, p + f[:-1])[len(p) - 1:]
            d = (got - ref).abs()
            top2 = ref.topk(2, dim=-1)
            flips = int((got.argmax(-1) != top2.indices[:, 0]).sum())
            rows.append({"prompt_tokens": len(p), "rms": round(d.pow(2).mean().sqrt().item(), 6),
                         "max": round(d.max().item(), 5), "logit_std": round(ref.std().item(), 4),
                         "argmax_flips": flips, "positions": int(ref.shape[0])})
            sq += d.pow(2).sum().item()
            n += d.numel()
            worst = max(worst, d.max().item())
        std = float(np.mean([r["logit_std"] for r in rows]))
        out["cases"][cfg.name] = {"preset": preset, "layers": layers, "prompt_lens": lens, "n_new": n_new,
                                  "seed": seed, "rms": round((sq / n) ** 0.5, 6), "max": round(worst, 5),
                                  "logit_std": round(std, 4), "rows": rows}
        print(f"{cfg.name}: bf16-vs-fp32 rms {(sq / n) ** 0.5:.5f} max {worst:.4f} logit std {std:.3f} "
              f"({time.time() - t0:.0f}s)", flush=True)
        del hf, refs
    path.write_text(json.dumps(out, indent=1) + "\n")
    print("wrote", path)


if __name__ == "__main__":
    main()
