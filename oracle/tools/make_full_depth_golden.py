"""ORACLE tool (test infrastructure): fp32 CPU logits of a FULL-DEPTH model that does not fit as one
transformers object in host memory (Qwen3-14B: 14.8 B parameters = 59 GB of fp32).

The weights are the device-independent hashed init (byzantine_consensus_llm_agents_b200.model
.hashed_normal: a pure integer function of (seed, tensor, row, col), so this process regenerates,
one layer at a time, exactly the bf16 values the GPU engine holds).  The arithmetic is
transformers' own code, unchanged: ONE `Qwen3DecoderLayer` / `MistralDecoderLayer` module
(modeling_qwen3.py:305-334 and everything below it: RMSNorm 50-64, MLP 70-83, RoPE 151-181, eager
attention with per-head q/k norm 252-291) is instantiated in fp32 and re-loaded with each layer's
weights in turn; rotary embedding, final norm and lm_head are the model classes' own modules.
For small models the streamed forward is checked against `oracle.model.full_logits` of the
ordinary whole-model object (tests/test_oracle_full_depth.py), which pins the streaming itself.

    python -m oracle.tools.make_full_depth_golden --model qwen3-14b --out tests/golden/full_depth_qwen3_14b.npz

The fixture keeps, for every compared position, the fp32 logits quantised to fp16 plus the exact
fp32 top-8 (ids and values); fp16 storage error (<= 2^-11 relative, ~2e-3 at |logit| 6) is an
order of magnitude below the tolerance it is used with and is added to it in the test.
"""
from __future__ import annotations

import argparse
import time

import numpy as np
import torch

from byzantine_consensus_llm_agents_b200 import model as M
from byzantine_consensus_llm_agents_b200.tokenizer import get_tokenizer
from oracle.model import hf_config


def case_inputs(tk, prompt_len: int, n_new: int, seed: int):
    rng = np.random.default_rng(seed)
    lo, hi = tk.regular_range
    return rng.integers(lo, hi, size=prompt_len).tolist(), rng.integers(lo, hi, size=n_new).tolist()


@torch.no_grad()
def streamed_logits(cfg: M.ModelConfig, ids, first_pos: int, weight_seed: int, log=None, also_bf16=False):
    """fp32 logits [len(ids) - first_pos, V] for positions first_pos.. of one sequence, the decoder
    layers streamed through a single transformers layer module.  With also_bf16 a second copy of
    the same modules, cast with `.to(torch.bfloat16)`, runs beside it on the same weights and the
    pair (fp32 logits, bf16-run logits as fp32) is returned: the reference's own model executed in
    bf16, i.e. the deviation from fp32 that any bf16 execution of it is entitled to at this depth."""
    hc = hf_config(cfg, n_layers=1)
    hc._attn_implementation = "eager"
    if cfg.qk_norm:
        from transformers.models.qwen3 import modeling_qwen3 as mod
        layer_cls, rope_cls, norm_cls = mod.Qwen3DecoderLayer, mod.Qwen3RotaryEmbedding, mod.Qwen3RMSNorm
    else:
        from transformers.models.mistral import modeling_mistral as mod
        layer_cls, rope_cls, norm_cls = mod.MistralDecoderLayer, mod.MistralRotaryEmbedding, mod.MistralRMSNorm
    dtypes = [torch.float32] + ([torch.bfloat16] if also_bf16 else [])
    layers = [layer_cls(hc, 0).to(dt).eval() for dt in dtypes]
    rope = rope_cls(config=hc)
    norms = [norm_cls(cfg.d_model, eps=cfg.rms_eps).to(dt).eval() for dt in dtypes]   # weight = 1
    T = len(ids)
    t_ids = torch.tensor([list(ids)], dtype=torch.long)
    emb = M.hashed_embed(cfg, weight_seed, "cpu")[t_ids[0]][None]                      # bf16 [1, T, d]
    hs = [emb.to(dt) for dt in dtypes]
    pos = torch.arange(T)[None]
    cos_sin = [rope(h, pos) for h in hs]
    masks = [torch.full((T, T), torch.finfo(dt).min, dtype=dt).triu(1)[None, None] for dt in dtypes]
    for l in range(cfg.n_layers):
        t0 = time.time()
        sd = M.hashed_state_dict_layer(cfg, l, weight_seed, "cpu")
        for i, (layer, dt) in enumerate(zip(layers, dtypes)):
            for name, w in sd.items():
                a, b = name.split(".")
                getattr(getattr(layer, a), b).weight.copy_(w.to(dt))
            h = layer(hs[i], attention_mask=masks[i], position_ids=pos, position_embeddings=cos_sin[i])
            hs[i] = h[0] if isinstance(h, tuple) else h
        del sd
        if log:
            log(f"layer {l + 1}/{cfg.n_layers} ({time.time() - t0:.0f}s) |h| rms "
                f"{hs[0].pow(2).mean().sqrt().item():.4f}")
    head = M.hashed_embed(cfg, weight_seed, "cpu", "embed" if cfg.tie_embeddings else "lm_head")
    outs = []
    for h, norm, dt in zip(hs, norms, dtypes):
        hn = norm(h[0, first_pos:])
        out = torch.empty(hn.shape[0], cfg.vocab)
        for v0 in range(0, cfg.vocab, 16384):                  # the head in slices of its own dtype
            out[:, v0:v0 + 16384] = (hn @ head[v0:v0 + 16384].to(dt).T).float()
        outs.append(out)
    return (outs[0], outs[1]) if also_bf16 else outs[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="qwen3-14b")
    ap.add_argument("--layers", type=int, default=None)
    ap.add_argument("--prompt-len", type=int, default=500)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--seed", type=int, default=7)
    ap.add_argument("--weight-seed", type=int, default=0)
    ap.add_argument("--out", required=True)
    a = ap.parse_args()
    cfg = M.PRESETS[a.model]
    if a.layers:
        cfg = M.ModelConfig(**{**cfg.__dict__, "n_layers": a.layers, "name": f"{cfg.name}-L{a.layers}"})
    tk = get_tokenizer(cfg.family, cfg.vocab)
    prompt, forced = case_inputs(tk, a.prompt_len, a.steps, a.seed)
    t0 = time.time()
    logits, logits_bf16 = streamed_logits(cfg, prompt + forced[:-1], len(prompt) - 1, a.weight_seed,
                                          log=lambda m: print(m, flush=True), also_bf16=True)
    dev = (logits_bf16 - logits).abs()
    floor = np.array([dev.pow(2).mean().sqrt().item(), dev.max().item(), logits.std().item()], np.float64)
    print(f"transformers bf16 run vs fp32 at this depth: rms {floor[0]:.4f} max {floor[1]:.4f}")
    top = logits.topk(8, dim=-1)
    np.savez_compressed(a.out, logits_f16=logits.numpy().astype(np.float16),
                        top_ids=top.indices.numpy().astype(np.int32), top_vals=top.values.numpy(),
                        prompt=np.array(prompt, np.int32), forced=np.array(forced, np.int32),
                        meta=np.array([cfg.n_layers, a.weight_seed, a.seed, a.prompt_len, a.steps], np.int64),
                        bf16_floor=floor,   # [rms, max, logit std] of transformers' own bf16 run vs fp32
                        model=np.array(a.model))
    print(f"wrote {a.out}: logits {tuple(logits.shape)} std {logits.std().item():.3f} in {time.time() - t0:.0f}s")


if __name__ == "__main__":
    main()
