"""Generates tests/golden/adapter_golden.json from the UNMODIFIED reference adapter
(vllm_agent.py run through the vllm shim): method signatures, chat-template outputs for every
template branch (vllm_agent.py:211-290), and the JSON-salvage / error-dict behaviour
(vllm_agent.py:334-369, 433-448) on a fixed set of raw model outputs.

    python -m oracle.tools.make_adapter_golden
"""
from __future__ import annotations

import inspect
import json
from pathlib import Path

from oracle import ref_harness as H

OUT = Path(__file__).resolve().parents[2] / "tests" / "golden" / "adapter_golden.json"

MODEL_NAMES = ["Qwen/Qwen3-14B", "Qwen/Qwen3-8B", "qwen-3-test", "Qwen/Qwen3-4B-Instruct-2507",
               "Qwen/Qwen2.5-7B-Instruct", "meta-llama/Llama-3.1-8B-Instruct", "meta-llama/Llama-2-7b-chat-hf",
               "mistralai/Mistral-Small-Instruct-2409", "some/other-model"]
PAIRS = [("You are agent_1.\nRules: be honest.", "=== ROUND 3 ===\nYour value: 17"),
         ("", "hello"), ("sys {braces} %s", "user with \"quotes\" and \n newlines")]
RAW_OUTPUTS = [
    '{"decision": "stop"}',
    '  \n{"decision": "continue"}\n ',
    'Sure! {"internal_strategy": "abc", "value": 12, "public_reasoning": "because {nested} ok"} trailing',
    '{"a": {"b": 1}} {"c": 2}',
    '{"internal_strategy": "never closes',
    'no json at all',
    '{broken} then {"value": 3}',
    '',
    '{"decision":                       \n   \t\t\t\t ',
    '[1, 2, 3]',
    '"just a string"',
    '{"value": 1}}',
]
SCHEMA_A = {"type": "object", "properties": {"decision": {"type": "string", "enum": ["stop", "continue"]}},
            "required": ["decision"], "additionalProperties": False}
SCHEMA_B = {"type": "object", "properties": {"decision": {"type": "string", "enum": ["stop", "continue", "abstain"]}},
            "required": ["decision"], "additionalProperties": False}


class Canned:
    """vllm-shim backend that replays RAW_OUTPUTS in order."""

    def __init__(self):
        self.queue = []

    def __call__(self, prompts, sp):
        return [self.queue.pop(0) for _ in prompts]


def main():
    backend = Canned()
    H.install_vllm_shim(backend)
    ref = H.import_reference()
    VA = ref.vllm_agent
    gold = {"generator": "oracle/tools/make_adapter_golden.py"}
    gold["signatures"] = {n: str(inspect.signature(getattr(VA.VLLMAgent, n)))
                          for n in ["__init__", "generate", "generate_json", "batch_generate_json",
                                    "batch_generate", "_format_chat_prompt", "_extract_json", "shutdown"]}
    gold["default_model_config"] = VA.VLLMAgent("a", "Qwen/Qwen3-8B").model_config
    templates = []
    for name in MODEL_NAMES:
        for think in (True, False):
            agent = VA.VLLMAgent("a", name, {"disable_qwen3_thinking": think, "max_model_len": 64})
            for s, u in PAIRS:
                templates.append({"model": name, "disable_qwen3_thinking": think, "system": s, "user": u,
                                  "out": agent._format_chat_prompt(s, u)})
    gold["templates"] = templates
    agent = VA.VLLMAgent("a", "Qwen/Qwen3-8B", {"max_model_len": 64})
    single = []
    for raw in RAW_OUTPUTS:
        backend.queue = [raw]
        single.append({"raw": raw, "out": agent.generate_json("u", SCHEMA_A, 0.0, 10, system_prompt="s")})
    gold["generate_json"] = single
    backend.queue = list(RAW_OUTPUTS)
    gold["batch_same_schema"] = agent.batch_generate_json([("s", "u", SCHEMA_A)] * len(RAW_OUTPUTS), 0.3, 10)
    backend.queue = list(RAW_OUTPUTS)
    mixed = [("s", "u", SCHEMA_A if i % 2 else SCHEMA_B) for i in range(len(RAW_OUTPUTS))]
    gold["batch_mixed_schema"] = agent.batch_generate_json(mixed, 0.3, 10)
    gold["extract_json"] = [{"raw": r, "out": agent._extract_json(r)} for r in RAW_OUTPUTS]
    gold["empty_batch"] = agent.batch_generate_json([])
    backend.queue = ["  free text \n"]
    gold["generate_text"] = agent.generate("hi", system_prompt="sys")
    # engine failure -> error dicts, never an exception
    def boom(prompts, sp):
        raise RuntimeError("engine exploded")
    H.ShimLLM.backend = boom
    gold["error_single"] = agent.generate_json("u", SCHEMA_A)
    gold["error_batch_same"] = agent.batch_generate_json([("s", "u", SCHEMA_A)] * 2)
    gold["error_batch_mixed"] = agent.batch_generate_json([("s", "u", SCHEMA_A), ("s", "u", SCHEMA_B)])
    VA.VLLMAgent.shutdown()
    VA.VLLMAgent.shutdown()  # idempotent
    OUT.write_text(json.dumps(gold, indent=1, ensure_ascii=False))
    print(f"wrote {OUT}: {len(templates)} template cases, {len(RAW_OUTPUTS)} raw outputs")


if __name__ == "__main__":
    main()
