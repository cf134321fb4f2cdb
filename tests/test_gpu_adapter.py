"""The drop-in adapter on the real engine (GPU): VLLMAgent.generate_json / batch_generate_json /
generate / batch_generate end to end on a small registered model, mixed schemas in one call,
reload on config change, shutdown — and agreement with the CPU oracle engine on a greedy call."""
import json

import pytest
import torch

pytestmark = pytest.mark.gpu

from byzantine_consensus_llm_agents_b200 import vllm_agent as VA  # noqa: E402
from byzantine_consensus_llm_agents_b200.model import ModelConfig, Weights  # noqa: E402
from oracle.model import build_hf_model  # noqa: E402

CFG = ModelConfig("adapter-test-qwen3", n_layers=2, d_model=256, n_heads=4, n_kv_heads=2, d_ff=512, vocab=8192)
NAME = "test/Qwen3-adapter"
VOTE = {"type": "object", "properties": {"decision": {"type": "string", "enum": ["stop", "continue"]}},
        "required": ["decision"], "additionalProperties": False}
BYZ_VOTE = {"type": "object", "properties": {"decision": {"type": "string", "enum": ["stop", "continue", "abstain"]}},
            "required": ["decision"], "additionalProperties": False}
DECIDE = {"type": "object", "properties": {"internal_strategy": {"type": "string"},
                                           "value": {"type": "integer", "minimum": 0, "maximum": 50},
                                           "public_reasoning": {"type": "string"}},
          "required": ["internal_strategy", "value", "public_reasoning"], "additionalProperties": False}
MC = {"max_model_len": 1024, "gpu_memory_utilization": 0.5, "tensor_parallel_size": 1, "max_num_seqs": 4,
      "kv_pages": 600, "engine_max_rows": 8, "seed": 3}


@pytest.fixture(scope="module")
def hf():
    model = build_hf_model(CFG, seed=0)
    VA.register_model(NAME, CFG, lambda cfg, dev: Weights.from_hf_state_dict(cfg, model.state_dict(), dev))
    yield model
    VA.VLLMAgent.shutdown()


def test_batch_generate_json_mixed_schemas_one_engine_call(hf):
    agents = [VA.VLLMAgent(f"agent_{i}", NAME, MC) for i in range(4)]
    assert all(a.llm is agents[0].llm for a in agents)          # one shared engine
    eng = agents[0].llm.engine
    calls0 = eng.stats["calls"]
    prompts = [("You are agent_0.", "Vote.", VOTE), ("You are agent_1.", "Vote.", BYZ_VOTE),
               ("You are agent_2.", "Decide.", DECIDE), ("You are agent_3.", "Vote.", VOTE)]
    out = agents[0].batch_generate_json(prompts, temperature=0.3, max_tokens=48)
    assert eng.stats["calls"] == calls0 + 1                     # the reference would make 4 calls
    assert len(out) == 4
    for (_, _, schema), r in zip(prompts, out):
        assert isinstance(r, dict)
        if "error" in r:
            assert r["error"] == "json_parse_failed" and "raw" in r
        elif "decision" in schema["properties"]:
            assert r["decision"] in schema["properties"]["decision"]["enum"]
        else:
            assert 0 <= r["value"] <= 50 and isinstance(r["internal_strategy"], str)
    # identical resubmission (the reference's batch retry): prompts come from the prefix cache
    cached0 = eng.stats["cached_tokens"]
    agents[0].batch_generate_json(prompts, temperature=0.3, max_tokens=8)
    assert eng.stats["cached_tokens"] > cached0


def test_generate_json_greedy_matches_cpu_oracle_engine(hf):
    from oracle import ref_harness as H
    from oracle.engine import OracleEngine
    agent = VA.VLLMAgent("agent_0", NAME, MC)
    oracle = OracleEngine(hf, agent.llm.tokenizer, base_seed=3)
    prompt = agent._format_chat_prompt("You are agent_0, an honest agent.", "Should the game stop?")
    sp = H.SamplingParams(temperature=0.0, max_tokens=40, guided_decoding=H.GuidedDecodingParams(json=VOTE))
    want_text, extra = oracle([prompt], sp)[0]
    got = agent.generate_json("Should the game stop?", VOTE, 0.0, 40, system_prompt="You are agent_0, an honest agent.")
    margins = [a - b for a, b in zip(extra["top1"], extra["top2"])]
    if min(margins) > 4e-2:          # every greedy step is clear of bf16 noise: outputs must be identical
        want = agent._parse_single(want_text)
        assert got == want
    else:                            # otherwise at least the shape of the result is the reference's
        assert isinstance(got, dict) and ("decision" in got or got.get("error") == "json_parse_failed")


def test_free_text_paths_and_reload(hf):
    agent = VA.VLLMAgent("agent_0", NAME, MC)
    t = agent.generate("Say something.", temperature=0.7, max_tokens=12, system_prompt="You are agent_0.")
    assert isinstance(t, str) and t == t.strip()
    ts = agent.batch_generate(["one", "two", "three"], temperature=0.0, max_tokens=6)
    assert len(ts) == 3 and all(isinstance(x, str) for x in ts)
    assert agent.batch_generate([]) == []
    first = VA.VLLMAgent._shared_llm
    other = VA.VLLMAgent("agent_9", NAME, {**MC, "max_model_len": 512})     # config changed -> reload
    assert VA.VLLMAgent._shared_llm is not first and other.llm is VA.VLLMAgent._shared_llm
    VA.VLLMAgent.shutdown()
    assert VA.VLLMAgent._shared_llm is None
    VA.VLLMAgent.shutdown()


def test_oversized_request_becomes_error_dict_not_exception(hf):
    agent = VA.VLLMAgent("agent_0", NAME, MC)
    r = agent.generate_json("x " * 4000, VOTE, 0.0, 16, system_prompt="s")
    assert "error" in r and "max_model_len" in r["error"]
    rs = agent.batch_generate_json([("s", "x " * 4000, VOTE), ("s", "ok", VOTE)], 0.0, 16)
    assert len(rs) == 2 and all("error" in x for x in rs)
