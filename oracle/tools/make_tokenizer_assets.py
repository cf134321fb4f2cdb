"""Generates byzantine_consensus_llm_agents_b200/assets/bpe_merges.json.

No tokenizer or model files exist offline, so the engine ships a deterministic
synthetic byte-level BPE tokenizer.  Its merges are trained here, once, on the prompts the
UNMODIFIED reference builds (driven through the vllm shim with a scripted JSON backend), so
that the reference's prompts tokenise at a realistic ~4 chars/token.  Run from the repo root:

    python -m oracle.tools.make_tokenizer_assets
"""
from __future__ import annotations

import json
import random
import re
from pathlib import Path

from tokenizers import Tokenizer, models, pre_tokenizers, trainers

from oracle import ref_harness as H

OUT = Path(__file__).resolve().parents[2] / "byzantine_consensus_llm_agents_b200" / "assets" / "bpe_merges.json"
N_MERGES = 7936  # 256 byte tokens + 7936 merges = 8192 learned tokens


class ScriptedBackend:
    """Emits schema-valid JSON built from words of the prompt itself."""

    def __init__(self, seed=0):
        self.rng = random.Random(seed)

    def _text(self, prompt, n):
        words = re.findall(r"[A-Za-z']+", prompt)
        return " ".join(self.rng.choice(words) for _ in range(n)).capitalize() + "."

    def __call__(self, prompts, sp):
        outs = []
        schema = sp.guided_decoding.json if sp.guided_decoding else None
        for p in prompts:
            if schema is None:
                outs.append(self._text(p, 20))
            elif "decision" in schema["properties"]:
                outs.append(json.dumps({"decision": self.rng.choice(schema["properties"]["decision"]["enum"])}))
            else:
                d = {"internal_strategy": self._text(p, self.rng.randint(20, 60)),
                     "value": self.rng.randint(0, 50),
                     "public_reasoning": self._text(p, self.rng.randint(20, 80))}
                outs.append(json.dumps(d))
        return outs


def collect_corpus():
    corpus = []
    for (h, b, rounds, aware, model) in [(8, 2, 8, "may_exist", "Qwen/Qwen3-14B"),
                                         (4, 0, 5, "none_exist", "Qwen/Qwen3-0.6B"),
                                         (8, 0, 6, "may_exist", "Qwen/Qwen3-8B"),
                                         (16, 4, 4, "may_exist", "mistralai/Mistral-Small-Instruct-2409")]:
        H.install_vllm_shim(ScriptedBackend(seed=h * 100 + b))
        ref = H.import_reference()
        H.run_game(ref, h, b, rounds, seed=1234, awareness=aware, model_name=model)
        for c in H.ShimLLM.calls:
            corpus.extend(c.prompts)
            corpus.extend(c.texts)
        H.uninstall()
    return corpus


def main():
    corpus = collect_corpus()
    print(f"corpus: {len(corpus)} documents, {sum(map(len, corpus))} chars")
    tok = Tokenizer(models.BPE())
    tok.pre_tokenizer = pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=True)
    tr = trainers.BpeTrainer(vocab_size=256 + N_MERGES, initial_alphabet=pre_tokenizers.ByteLevel.alphabet(),
                             special_tokens=[], show_progress=False)
    tok.train_from_iterator(corpus, tr)
    merges = json.loads(tok.to_str())["model"]["merges"]
    OUT.parent.mkdir(parents=True, exist_ok=True)
    OUT.write_text(json.dumps(merges, ensure_ascii=True, separators=(",", ":")))
    print(f"wrote {len(merges)} merges to {OUT}")
    sample = corpus[0]
    print("chars/token on first prompt:", len(sample) / len(tok.encode(sample).ids))


if __name__ == "__main__":
    main()
