"""ORACLE (test infrastructure only — never imported by the product path).

The model-forward oracle: transformers' own fp32 CPU implementation of the architectures the
reference names (Qwen3ForCausalLM / MistralForCausalLM, transformers 5.5.0 —
modeling_qwen3.py:50-64 RMSNorm, 70-83 MLP, 151-181 RoPE, 252-291 attention with per-head q/k
norm before RoPE, 305-334 decoder layer), built from an explicit config with
`torch.manual_seed(seed)` and transformers' default init (N(0, 0.02), norm weights 1).
Weights are rounded to bf16 values (kept in fp32 storage) so that oracle and engine hold
IDENTICAL parameters; all arithmetic here is fp32.
"""
from __future__ import annotations

import torch


def hf_config(cfg, n_layers=None):
    n_layers = n_layers or cfg.n_layers
    common = dict(vocab_size=cfg.vocab, hidden_size=cfg.d_model, intermediate_size=cfg.d_ff,
                  num_hidden_layers=n_layers, num_attention_heads=cfg.n_heads,
                  num_key_value_heads=cfg.n_kv_heads, head_dim=cfg.head_dim,
                  max_position_embeddings=cfg.max_pos, rms_norm_eps=cfg.rms_eps,
                  rope_theta=cfg.rope_theta, tie_word_embeddings=cfg.tie_embeddings,
                  attention_dropout=0.0)
    if cfg.qk_norm:
        from transformers import Qwen3Config
        return Qwen3Config(attention_bias=False, use_sliding_window=False, **common)
    from transformers import MistralConfig
    return MistralConfig(sliding_window=None, **common)


def build_hf_model(cfg, seed: int = 0, n_layers=None):
    from transformers import MistralForCausalLM, Qwen3ForCausalLM
    hc = hf_config(cfg, n_layers)
    hc._attn_implementation = "eager"
    torch.manual_seed(seed)
    cls = Qwen3ForCausalLM if cfg.qk_norm else MistralForCausalLM
    model = cls(hc).eval()
    with torch.no_grad():
        for p in model.parameters():
            p.copy_(p.to(torch.bfloat16).float())
    for p in model.parameters():
        p.requires_grad_(False)
    return model


@torch.no_grad()
def full_logits(model, ids) -> torch.Tensor:
    """fp32 logits [T, V] for one sequence (no cache)."""
    out = model(input_ids=torch.tensor([list(ids)], dtype=torch.long), use_cache=False)
    return out.logits[0].float()


def build_hf_model_hashed(cfg, seed: int = 0):
    """Whole transformers model holding the device-independent hashed init
    (byzantine_consensus_llm_agents_b200.model.hashed_normal) — the same parameter values
    `Weights.hashed_init` puts on the GPU; norm weights stay 1.  For models that fit in host memory;
    larger ones are streamed layer by layer (oracle/tools/make_full_depth_golden.py)."""
    from byzantine_consensus_llm_agents_b200 import model as M
    model = build_hf_model(cfg, seed=0)
    with torch.no_grad():
        model.model.embed_tokens.weight.copy_(M.hashed_embed(cfg, seed, "cpu").float())
        if not cfg.tie_embeddings:
            model.lm_head.weight.copy_(M.hashed_embed(cfg, seed, "cpu", "lm_head").float())
        for l, layer in enumerate(model.model.layers):
            for name, w in M.hashed_state_dict_layer(cfg, l, seed, "cpu").items():
                a, b = name.split(".")
                getattr(getattr(layer, a), b).weight.copy_(w.float())
    return model
