#!/usr/bin/env python
"""Small engine run for ncu captures of the prefill-sized glue kernels and the sampler (dev tool):
a 2-layer Qwen3-14B-width model, one prefill group of 7 x 1,770 tokens, then 4 decode steps at
V = 151,936.    ncu --set full -k regex:'mask_sample|add_rmsnorm_row|qknorm_rope|prefill_attn' ...
"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from byzantine_consensus_llm_agents_b200.engine import Engine, Request  # noqa: E402
from byzantine_consensus_llm_agents_b200.model import PRESETS, ModelConfig, Weights  # noqa: E402
from byzantine_consensus_llm_agents_b200.tokenizer import get_tokenizer  # noqa: E402


def main():
    full = PRESETS["qwen3-14b"]
    cfg = ModelConfig(**{**full.__dict__, "n_layers": 2, "name": "qwen3-14b-L2"})
    dev = torch.device("cuda:0")
    w = Weights.random_init(cfg, dev)
    tk = get_tokenizer(cfg.family, cfg.vocab)
    eng = Engine(cfg, w, tk, max_model_len=2048, max_rows=16, prefill_tokens=16384, kv_pages=1400,
                 use_cuda_graph=False)
    rng = np.random.default_rng(0)
    lo, hi = tk.regular_range
    n_rows = int(sys.argv[1]) if len(sys.argv) > 1 else 7          # e.g. `10 1600`: the bench's decode shape
    n_tok = int(sys.argv[2]) if len(sys.argv) > 2 else 1770
    prompts = [rng.integers(lo, hi, size=n_tok).tolist() for _ in range(n_rows)]
    res = eng.generate([Request("", None, 0.5, 5, prompt_ids=p) for p in prompts])
    torch.cuda.synchronize()
    print("ok", [len(r.token_ids) for r in res])
    eng.close()


if __name__ == "__main__":
    main()
