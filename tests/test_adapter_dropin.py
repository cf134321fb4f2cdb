"""Boundary B1: byzantine_consensus_llm_agents_b200.vllm_agent must be indistinguishable from the
reference's vllm_agent.py at its public surface.

* against tests/golden/adapter_golden.json (recorded from the unmodified reference by
  oracle/tools/make_adapter_golden.py): signatures, defaults, every chat-template branch, the
  JSON-salvage ladder and the error-dict conventions;
* when /root/reference is present (build container): whole games of the UNMODIFIED reference
  (main.py / bcg_agents.py) played once on the reference's own adapter and once on the drop-in,
  both fed by the same scripted engine — transcripts and engine call traces must be identical
  when schemas are uniform, and per-row identical when they are mixed.
No GPU involved: the engine object behind the adapter is replaced by a scripted fake.
"""
import inspect
import json
import random
import re
from pathlib import Path

import pytest

from byzantine_consensus_llm_agents_b200 import vllm_agent as VA

GOLD = json.loads((Path(__file__).parent / "golden" / "adapter_golden.json").read_text())
HAVE_REF = Path("/root/reference/byzantine_consensus_game/vllm_agent.py").exists()


class FakeLLM:
    """Stands where B200LLM stands: run(prompts, schemas, temperature, max_tokens) -> texts."""

    def __init__(self, fn):
        self.fn = fn
        self.calls = []

    def run(self, prompts, schemas, temperature, max_tokens):
        self.calls.append({"prompts": list(prompts), "schemas": list(schemas), "temperature": temperature,
                           "max_tokens": max_tokens})
        return self.fn(prompts, schemas, temperature, max_tokens)

    def close(self):
        pass


@pytest.fixture
def fake_agent():
    def make(fn, model="Qwen/Qwen3-8B", cfg=None):
        llm = FakeLLM(fn)
        cfg = cfg or {"max_model_len": 64}
        VA.VLLMAgent._shared_llm = llm
        VA.VLLMAgent._shared_model_name = model
        VA.VLLMAgent._shared_model_config = dict(cfg)
        return VA.VLLMAgent("a", model, cfg), llm
    yield make
    VA.VLLMAgent._shared_llm = None
    VA.VLLMAgent._shared_model_name = None
    VA.VLLMAgent._shared_model_config = None


def _norm_sig(s: str) -> str:
    return re.sub(r"\b(typing\.)", "", s)


def test_public_surface_matches_reference():
    assert isinstance(VA.VERBOSE, bool)
    for name, want in GOLD["signatures"].items():
        got = str(inspect.signature(getattr(VA.VLLMAgent, name)))
        assert _norm_sig(got) == _norm_sig(want), name
    for attr in ("_shared_llm", "_shared_model_name", "_shared_model_config"):
        assert hasattr(VA.VLLMAgent, attr)
    assert VA._DEFAULT_MODEL_CONFIG == GOLD["default_model_config"]


def test_chat_templates_all_branches(fake_agent):
    for case in GOLD["templates"]:
        cfg = {"disable_qwen3_thinking": case["disable_qwen3_thinking"], "max_model_len": 64}
        agent, _ = fake_agent(lambda *a: [], model=case["model"], cfg=cfg)
        assert agent._format_chat_prompt(case["system"], case["user"]) == case["out"], case["model"]


def test_json_salvage_and_error_dicts(fake_agent):
    raws = [c["raw"] for c in GOLD["generate_json"]]
    schema_a = {"type": "object", "properties": {"decision": {"type": "string", "enum": ["stop", "continue"]}},
                "required": ["decision"], "additionalProperties": False}
    schema_b = json.loads(json.dumps(schema_a))
    schema_b["properties"]["decision"]["enum"].append("abstain")
    queue = []
    agent, llm = fake_agent(lambda p, s, t, m: [queue.pop(0) for _ in p])
    for c in GOLD["generate_json"]:
        queue[:] = [c["raw"]]
        assert agent.generate_json("u", schema_a, 0.0, 10, system_prompt="s") == c["out"], c["raw"]
    queue[:] = list(raws)
    assert agent.batch_generate_json([("s", "u", schema_a)] * len(raws), 0.3, 10) == GOLD["batch_same_schema"]
    queue[:] = list(raws)
    mixed = [("s", "u", schema_a if i % 2 else schema_b) for i in range(len(raws))]
    assert agent.batch_generate_json(mixed, 0.3, 10) == GOLD["batch_mixed_schema"]
    # mixed schemas still reach the engine as ONE call (the reference makes len(raws) calls)
    assert len(llm.calls[-1]["prompts"]) == len(raws)
    for c in GOLD["extract_json"]:
        assert agent._extract_json(c["raw"]) == c["out"]
    assert agent.batch_generate_json([]) == GOLD["empty_batch"] == []
    queue[:] = ["  free text \n"]
    assert agent.generate("hi", system_prompt="sys") == GOLD["generate_text"]


def test_engine_failures_become_error_dicts(fake_agent):
    def boom(*a):
        raise RuntimeError("engine exploded")
    agent, _ = fake_agent(boom)
    s = {"type": "object", "properties": {}, "additionalProperties": False}
    s2 = {"type": "object", "properties": {"a": {"type": "null"}}, "additionalProperties": False}
    assert agent.generate_json("u", s) == GOLD["error_single"]
    assert agent.batch_generate_json([("s", "u", s)] * 2) == GOLD["error_batch_same"]
    assert agent.batch_generate_json([("s", "u", s), ("s", "u", s2)]) == GOLD["error_batch_mixed"]


def test_shutdown_is_idempotent_and_safe_when_nothing_loaded():
    VA.VLLMAgent.shutdown()
    VA.VLLMAgent.shutdown()
    assert VA.VLLMAgent._shared_llm is None


def test_unsupported_sampling_options_are_loud(fake_agent):
    agent, _ = fake_agent(lambda p, s, t, m: ["x"] * len(p))
    with pytest.raises(NotImplementedError):
        agent.generate("hi", top_p=0.9)
    with pytest.raises(NotImplementedError):
        agent.batch_generate(["hi"], presence_penalty=1.0)


def test_no_gpu_means_loud_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(Exception) as e:
        VA.VLLMAgent("a", "Qwen/Qwen3-0.6B", {"max_model_len": 128})
    assert "CUDA" in str(e.value) or "cuda" in str(e.value)
    VA.VLLMAgent.shutdown()


# ---------------------------------------------------------------------------- whole games
class Scripted:
    """Deterministic text generator keyed on (prompt, schema): independent of batching."""

    def text_for(self, prompt, schema, temperature):
        rng = random.Random(hash((prompt, json.dumps(schema, sort_keys=True))) & 0xFFFFFFFF)
        words = re.findall(r"[A-Za-z']+", prompt)
        say = lambda n: " ".join(rng.choice(words) for _ in range(n)).capitalize() + "."
        roll = rng.random()
        if "decision" in schema["properties"]:
            if roll < 0.15:
                return '{"decision":     '                       # runs into the token limit
            return json.dumps({"decision": rng.choice(schema["properties"]["decision"]["enum"])})
        if roll < 0.2:
            return '{"internal_strategy": "' + say(30)           # never closes
        if roll < 0.3:
            return "Here you go: " + json.dumps({"internal_strategy": say(8), "value": rng.randint(0, 50),
                                                 "public_reasoning": say(12)})
        return json.dumps({"internal_strategy": say(10), "value": rng.randint(0, 50),
                           "public_reasoning": say(15)})


def _play(adapter_module, honest, byz, rounds, awareness):
    from oracle import ref_harness as H
    s = Scripted()
    if adapter_module is None:
        H.install_vllm_shim(lambda prompts, sp: [s.text_for(p, sp.guided_decoding.json, sp.temperature) for p in prompts])
        ref = H.import_reference()
    else:
        H.install_vllm_shim(None)
        ref = H.import_reference(adapter_module=adapter_module)
        llm = FakeLLM(lambda p, sc, t, m: [s.text_for(a, b, t) for a, b in zip(p, sc)])
        adapter_module.VLLMAgent._shared_llm = llm
        adapter_module.VLLMAgent._shared_model_name = "Qwen/Qwen3-14B"
        adapter_module.VLLMAgent._shared_model_config = dict(ref.config.VLLM_CONFIG)
    stats, transcript, sim = H.run_game(ref, honest, byz, rounds, seed=1234, greedy=True, awareness=awareness,
                                        model_name="Qwen/Qwen3-14B")
    if adapter_module is None:
        rows = [(p, json.dumps(c.schema), c.temperature, c.max_tokens) for c in H.ShimLLM.calls for p in c.prompts]
        n_calls = len(H.ShimLLM.calls)
    else:
        rows = [(p, json.dumps(sc), c["temperature"], c["max_tokens"]) for c in llm.calls
                for p, sc in zip(c["prompts"], c["schemas"])]
        n_calls = len(llm.calls)
    H.uninstall()
    keep = {k: stats.get(k) for k in ("total_rounds", "consensus_reached", "final_consensus_value",
                                      "termination_reason", "consensus_outcome")}
    return transcript, keep, rows, n_calls


@pytest.mark.skipif(not HAVE_REF, reason="reference tree not available on this machine")
@pytest.mark.parametrize("honest,byz,rounds,aware", [(4, 0, 5, "none_exist"), (8, 2, 4, "may_exist"),
                                                      (3, 1, 3, "may_exist")])
def test_whole_game_on_dropin_equals_game_on_reference_adapter(honest, byz, rounds, aware):
    t_ref, s_ref, rows_ref, calls_ref = _play(None, honest, byz, rounds, aware)
    t_new, s_new, rows_new, calls_new = _play(VA, honest, byz, rounds, aware)
    assert t_new == t_ref
    assert s_new == s_ref
    # the same (prompt, schema, T, max_tokens) rows reach the engine, in the same order ...
    assert rows_new == rows_ref
    # ... in at most as many engine calls (mixed-schema batches are no longer serialised)
    assert calls_new <= calls_ref
    if byz:
        assert calls_new < calls_ref
    VA.VLLMAgent._shared_llm = None
    VA.VLLMAgent._shared_model_name = None
    VA.VLLMAgent._shared_model_config = None
