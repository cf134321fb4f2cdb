"""Loading what the reference loads (vllm_agent.py:126-144): a local HuggingFace model directory —
config.json, safetensors checkpoint, tokenizer.json — instead of a preset name with random-init
weights.  CPU part: the loaders reproduce the in-memory packing and the tokenizer's ids/bytes.
GPU part: the adapter, given the directory as `model_name`, generates what the registered
in-memory model generates."""
import json

import pytest
import torch

from byzantine_consensus_llm_agents_b200 import model as M
from byzantine_consensus_llm_agents_b200.tokenizer import LocalTokenizer, get_tokenizer
from oracle.model import build_hf_model

CFG = M.ModelConfig("local-dir-qwen3", n_layers=2, d_model=256, n_heads=2, n_kv_heads=1, d_ff=512, vocab=8192,
                    max_pos=2048)


@pytest.fixture(scope="module")
def model_dir(tmp_path_factory):
    d = tmp_path_factory.mktemp("hf_model")
    hf = build_hf_model(CFG, seed=0)
    hf.save_pretrained(str(d), safe_serialization=True)
    tk = get_tokenizer(CFG.family, CFG.vocab)
    tk._tok.save(str(d / "tokenizer.json"))
    (d / "generation_config.json").write_text(json.dumps({"eos_token_id": list(tk.stop_ids)}))
    return d, hf, tk


def test_config_checkpoint_and_tokenizer_round_trip(model_dir):
    d, hf, tk = model_dir
    cfg = M.config_from_dir(str(d))
    for k in ("n_layers", "d_model", "n_heads", "n_kv_heads", "d_ff", "vocab", "head_dim", "qk_norm",
              "tie_embeddings", "family"):
        assert getattr(cfg, k) == getattr(CFG, k), k
    assert abs(cfg.rms_eps - CFG.rms_eps) < 1e-12 and cfg.rope_theta == CFG.rope_theta
    want = M.Weights.from_hf_state_dict(CFG, hf.state_dict(), "cpu")
    got = M.Weights.from_safetensors_dir(cfg, str(d), "cpu")
    assert torch.equal(want.embed, got.embed) and torch.equal(want.lm_head, got.lm_head)
    for name in ("ln1", "ln2", "w_qkv", "w_o", "w_gate_up", "w_down", "q_norm", "k_norm"):
        a, b = getattr(want, name), getattr(got, name)
        assert len(a) == len(b) == CFG.n_layers
        assert all(torch.equal(x, y) for x, y in zip(a, b)), name
    lt = LocalTokenizer(str(d), cfg.vocab, cfg.family)
    text = '<|im_start|>system\nYou are agent_3.<|im_end|>\n{"value": 17, "r": "ünï — 合意"}'
    assert lt.encode(text) == tk.encode(text)
    assert lt.decode(lt.encode(text)) == tk.decode(tk.encode(text))
    assert lt.stop_ids == tk.stop_ids and lt.bos_id == tk.bos_id
    assert lt.token_bytes == tk.token_bytes              # the grammar compiler sees identical byte strings
    assert lt.regular_range == tk.regular_range


@pytest.mark.gpu
def test_adapter_loads_a_model_directory(model_dir):
    from byzantine_consensus_llm_agents_b200 import vllm_agent as VA
    d, hf, tk = model_dir
    mc = {"max_model_len": 512, "gpu_memory_utilization": 0.5, "tensor_parallel_size": 1, "max_num_seqs": 4,
          "kv_pages": 300, "engine_max_rows": 4, "seed": 3}
    schema = {"type": "object", "properties": {"decision": {"type": "string", "enum": ["stop", "continue"]}},
              "required": ["decision"], "additionalProperties": False}
    try:
        a = VA.VLLMAgent("agent_0", str(d), mc)
        from_dir = a.generate_json("Vote now.", schema, temperature=0.0, max_tokens=24, system_prompt="You are agent_0.")
        VA.register_model("test/local-dir", CFG, lambda cfg, dev: M.Weights.from_hf_state_dict(cfg, hf.state_dict(), dev))
        b = VA.VLLMAgent("agent_0", "test/local-dir", mc)
        in_memory = b.generate_json("Vote now.", schema, temperature=0.0, max_tokens=24, system_prompt="You are agent_0.")
        assert from_dir == in_memory
    finally:
        VA.VLLMAgent.shutdown()
