"""SURVEY.md §8f rank 3 — JSON validity is decided on the device: the row's DFA state at the stop
token says whether a complete object of the schema was emitted.  Pinned here over a few thousand
sampled rows at the shipped temperature (T = 0.5) for the reference's four schemas
(bcg_agents.py:590-599, 651-658, 1083-1092, 1155-1162):

  * a row that stops in an ACCEPTING state must `json.loads` and validate against its schema;
  * a row may also stop in a non-accepting state — only inside a string, where the reference's
    grammar (xgrammar as vLLM drives it, which this engine matches bit for bit) admits the stop
    token's literal text "<|im_end|>" as string content while the engine, like vLLM, ends the
    generation on that token id.  Such rows are unparseable in the reference too (its
    `json_parse_failed` path); the test checks that this is the ONLY way to stop incomplete.

(Detokenisation and json.loads stay on the host: they are 1 % of a round)"""
import json

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from byzantine_consensus_llm_agents_b200.engine import Engine, Request  # noqa: E402
from byzantine_consensus_llm_agents_b200.grammar import grammar_token_bytes  # noqa: E402
from byzantine_consensus_llm_agents_b200.model import ModelConfig, Weights  # noqa: E402
from byzantine_consensus_llm_agents_b200.tokenizer import get_tokenizer  # noqa: E402

STR = {"type": "string"}
HONEST_DECIDE = {"type": "object", "properties": {"internal_strategy": STR,
                                                  "value": {"type": "integer", "minimum": 0, "maximum": 50},
                                                  "public_reasoning": STR},
                 "required": ["internal_strategy", "value", "public_reasoning"], "additionalProperties": False}
BYZ_DECIDE = {"type": "object", "properties": {"internal_strategy": STR,
                                               "value": {"anyOf": [{"type": "integer", "minimum": 0, "maximum": 50},
                                                                   {"type": "string", "enum": ["abstain"]}]},
                                               "public_reasoning": STR},
              "required": ["internal_strategy", "value"], "additionalProperties": False}
HONEST_VOTE = {"type": "object", "properties": {"decision": {"type": "string", "enum": ["stop", "continue"]}},
               "required": ["decision"], "additionalProperties": False}
BYZ_VOTE = {"type": "object", "properties": {"decision": {"type": "string", "enum": ["stop", "continue", "abstain"]}},
            "required": ["decision"], "additionalProperties": False}


def _valid_value(v, spec) -> bool:
    if "anyOf" in spec:
        return any(_valid_value(v, s) for s in spec["anyOf"])
    if spec["type"] == "integer":
        return isinstance(v, int) and not isinstance(v, bool) and spec["minimum"] <= v <= spec["maximum"]
    if spec["type"] == "string":
        return isinstance(v, str) and ("enum" not in spec or v in spec["enum"])
    return False


def _validates(obj, schema) -> bool:
    if not isinstance(obj, dict) or set(obj) - set(schema["properties"]):
        return False
    if any(k not in obj for k in schema["required"]):
        return False
    if list(obj) != [k for k in schema["properties"] if k in obj]:     # xgrammar emits properties in order
        return False
    return all(_valid_value(v, schema["properties"][k]) for k, v in obj.items())


def test_every_stopped_guided_row_is_a_valid_object_of_its_schema():
    cfg = ModelConfig("json-validity", n_layers=1, d_model=256, n_heads=2, n_kv_heads=1, d_ff=512, vocab=8192)
    dev = torch.device("cuda:0")
    tk = get_tokenizer(cfg.family, cfg.vocab)
    eng = Engine(cfg, Weights.random_init(cfg, dev, seed=5), tk, max_model_len=512, max_rows=64,
                 prefill_tokens=4096, kv_pages=2200, base_seed=77)
    rng = np.random.default_rng(0)
    lo, hi = tk.regular_range
    schemas = [HONEST_DECIDE, BYZ_DECIDE, HONEST_VOTE, BYZ_VOTE]
    stopped = {i: 0 for i in range(4)}
    incomplete = {i: 0 for i in range(4)}
    total = 0
    tb = grammar_token_bytes(tk)
    gobj = [eng.grammar_objs[eng.grammar_index(sc)] for sc in schemas]
    for call in range(48):                                  # 48 calls x 64 rows = 3,072 rows
        reqs, which = [], []
        for b in range(64):
            si = (b + call) % 4
            which.append(si)
            # a short random prompt per row; 160 new tokens leave room for the three-field objects
            reqs.append(Request("", schemas[si], 0.5, 160, prompt_ids=rng.integers(lo, hi, size=12).tolist()))
        for si, r in zip(which, eng.generate(reqs)):
            total += 1
            assert r.finish in ("stop", "length"), r.finish      # never a grammar dead end
            if r.finish == "stop":
                g, st = gobj[si], 0
                assert r.token_ids[-1] in tk.stop_ids
                for t in r.token_ids[:-1]:
                    st = g.walk(st, tb[t])
                    assert st >= 0                                # every token was inside the grammar
                if g.accept[st]:
                    stopped[si] += 1
                    obj = json.loads(r.text)                          # must parse ...
                    assert _validates(obj, schemas[si]), (si, r.text)  # ... and be an object of ITS schema
                else:
                    incomplete[si] += 1
                    # only possible where the mask admits the stop token as TEXT: inside a string
                    word, bit = divmod(r.token_ids[-1], 32)
                    assert (int(g.masks[st][word]) >> bit) & 1
                    assert g.walk(st, b"<|im_end|>") >= 0 or g.walk(st, b"<|endoftext|>") >= 0
                    with pytest.raises(ValueError):
                        json.loads(r.text)
            else:
                assert len(r.token_ids) == 160
    print(f"{total} sampled rows; stopped complete (parsed + validated) per schema: {stopped}; "
          f"stopped inside a string on the stop token's literal text: {incomplete}")
    # the vote schemas have no free-form string: they can only stop complete
    assert incomplete[2] == 0 and incomplete[3] == 0
    assert stopped[2] + stopped[3] > 0.25 * total / 2
    assert sum(stopped.values()) > 500
    eng.close()
