"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement of what `vllm.LLM.generate` does for the reference's calls
(vllm_agent.py:195,331,430,502), assembled from the real third-party pieces where they run
on a CPU and restated where they do not:

  tokenize(prompt, add_special_tokens=True)        synthetic tokenizer (no HF files offline)
  model forward, fp32, KV cache                     transformers (oracle/model.py)
  grammar bitmask per step                          xgrammar 0.2.0 (oracle/grammar_xgr.py)
  mask -> temperature -> sample                     oracle/sampler.py (restates vLLM's sampler)
  stop on EOS / stop ids / max_tokens; text = decode(ids, skip specials)
                                                    vLLM semantics; the reference strips the text

Pinning status: the reference ships NO golden vectors, known-answer tests or
fixtures for this path, and its own engine (vLLM, `requirements.txt:1`, un-pinned) cannot run in
this container — against the reference's own outputs the oracle is therefore "parity unpinned".
What it IS pinned against: the real third-party libraries vLLM delegates to, run here —
transformers 5.5 (`Qwen3ForCausalLM` / `MistralForCausalLM`, fp32) for the forward, xgrammar 0.2.0
for every mask, Random123's Philox known-answer vectors for the sampler's generator — and the
unmodified reference game code for prompts, schemas, retries and transcripts (oracle/ref_harness.py).
The same statement is in DESIGN.md §4.

It is a backend for oracle/ref_harness.py's vllm shim.  Greedy generation is a deterministic
function of (prompt, schema, max_tokens); identical requests are memoised, which is exact.
Every generation also records, per step, the masked top-1 / top-2 logits so GPU parity tests
can tell a genuine mismatch from a bf16-level near-tie (north_star: "wherever top-1 logit
margins exceed the tolerance").
"""
from __future__ import annotations

import json
import time

import numpy as np
import torch

from oracle import sampler as S
from oracle.grammar_xgr import XgrOracle


def row_seed(base_seed: int, call_index: int, row: int):
    """Per-row Philox key; mirrors byzantine_consensus_llm_agents_b200.engine.row_seed."""
    x = (base_seed * 0x9E3779B97F4A7C15 + call_index * 0xBF58476D1CE4E5B9 + row * 0x94D049BB133111EB
         + 0x2545F4914F6CDD1D) & 0xFFFFFFFFFFFFFFFF
    x ^= x >> 30
    x = (x * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
    x ^= x >> 27
    x = (x * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
    x ^= x >> 31
    return x & 0xFFFFFFFF, x >> 32


class OracleEngine:
    def __init__(self, model, tokenizer, base_seed: int = 0, memoise: bool = True):
        self.model = model
        self.tk = tokenizer
        self.xgr = XgrOracle(tokenizer)
        self.base_seed = base_seed
        self.calls = 0
        self.memo = {} if memoise else None
        self.stats = {"prefill_tokens": 0, "decode_steps": 0, "prefill_s": 0.0, "decode_s": 0.0}

    @torch.no_grad()
    def _generate_one(self, ids, schema, temperature, max_tokens, seed):
        matcher = self.xgr.matcher(schema) if schema is not None else None
        bm = self.xgr.xgr.allocate_token_bitmask(1, self.tk.vocab_size) if matcher else None
        t0 = time.perf_counter()
        out = self.model(input_ids=torch.tensor([ids], dtype=torch.long), use_cache=True)
        self.stats["prefill_s"] += time.perf_counter() - t0
        self.stats["prefill_tokens"] += len(ids)
        cache = out.past_key_values
        logits = out.logits[0, -1].float().numpy()
        toks, top1, top2, first_logits = [], [], [], logits.copy()
        finish = "length"
        for step in range(max_tokens):
            mask = None
            if matcher is not None:
                matcher.fill_next_token_bitmask(bm, 0)
                mask = bm[0].numpy()
            # same arithmetic as S.sample(), kept open here so the decisive margin can be recorded:
            # masked logits when greedy, masked race scores (logit/T - log q) when sampling
            allowed = np.ones(logits.shape[0], bool) if mask is None else S.unpack_mask(mask, logits.shape[0])
            if not allowed.any():
                finish = "dead"
                break
            if temperature < 1e-5:
                score = logits.astype(np.float32)
            else:
                score, _ = S.race_scores(logits, temperature, step, seed[0], seed[1])
            ml = np.where(allowed, score, np.float32(-np.inf))
            tok = int(np.argmax(ml))
            assert tok == S.sample(logits, mask, temperature, step, seed[0], seed[1])
            part = np.partition(ml, -2)[-2:] if allowed.sum() > 1 else np.array([-np.inf, ml.max()])
            top1.append(float(part[1]))
            top2.append(float(part[0]))
            toks.append(int(tok))
            if tok in self.tk.stop_ids:
                finish = "stop"
                break
            if matcher is not None and not matcher.accept_token(int(tok)):
                finish = "dead"
                break
            if len(toks) >= max_tokens:
                break
            t0 = time.perf_counter()
            out = self.model(input_ids=torch.tensor([[tok]], dtype=torch.long), past_key_values=cache,
                             use_cache=True)
            self.stats["decode_s"] += time.perf_counter() - t0
            self.stats["decode_steps"] += 1
            cache = out.past_key_values
            logits = out.logits[0, -1].float().numpy()
        text = self.tk.decode([t for t in toks if t not in self.tk.stop_ids])
        return text, {"prompt_ids": len(ids), "tokens": toks, "top1": top1, "top2": top2,
                      "finish": finish, "first_logits_max": float(first_logits.max()),
                      "seed": [int(seed[0]), int(seed[1])]}

    def __call__(self, prompts, sp):
        """vllm-shim backend: list of (text, extra)."""
        gd = getattr(sp, "guided_decoding", None)
        schema = gd.json if gd is not None else None
        call_index = self.calls
        self.calls += 1
        res = []
        for row, p in enumerate(prompts):
            ids = self.tk.encode(p, add_special_tokens=True)
            greedy = sp.temperature < 1e-5
            key = (p, json.dumps(schema), sp.max_tokens) if (greedy and self.memo is not None) else None
            if key is not None and key in self.memo:
                res.append(self.memo[key])
                continue
            seed = row_seed(self.base_seed, call_index, row)
            r = self._generate_one(ids, schema, float(sp.temperature), int(sp.max_tokens), seed)
            if key is not None:
                self.memo[key] = r
            res.append(r)
        return res
