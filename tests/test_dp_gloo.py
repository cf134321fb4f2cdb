"""N > 1 path on CPU: two gloo ranks shard the rows of a batch_generate_json call over "replicas"
(scripted engines standing in for the GPU engine) and must reproduce the single-process result,
row order and per-row seeds included."""
import json
import os
import socket
import zlib

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from byzantine_consensus_llm_agents_b200 import dp


def test_shard_rows_is_a_partition():
    for n in (0, 1, 7, 10, 64):
        for world in (1, 2, 3, 8):
            seen = sorted(i for r in range(world) for i in dp.shard_rows(n, r, world))
            assert seen == list(range(n))
            sizes = [len(dp.shard_rows(n, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


class _FakeEngine:
    """Output depends on (prompt, schema, temperature, seed) only — like the real engine."""

    def __init__(self):
        self.calls = 0
        self.base_seed = 5
        self.seen = []

    def generate(self, reqs):
        from byzantine_consensus_llm_agents_b200.engine import Result, row_seed
        out = []
        for b, q in enumerate(reqs):
            seed = tuple(q.seed) if q.seed is not None else row_seed(self.base_seed, self.calls, b)
            self.seen.append(q.prompt)
            text = json.dumps({"decision": "stop" if (zlib.crc32(q.prompt.encode()) + seed[0]) % 2 else "continue",
                               "seed": list(seed)})
            out.append(Result(text, [], [], "stop", 1, 0))
        self.calls += 1
        return out


def _make_agent(data_parallel):
    from byzantine_consensus_llm_agents_b200 import vllm_agent as VA
    llm = VA.B200LLM.__new__(VA.B200LLM)
    llm.engine = _FakeEngine()
    llm.data_parallel = data_parallel
    llm.tp = None
    cfg = {"max_model_len": 64}
    VA.VLLMAgent._shared_llm, VA.VLLMAgent._shared_model_name, VA.VLLMAgent._shared_model_config = llm, "Qwen/Qwen3-8B", cfg
    return VA.VLLMAgent("a", "Qwen/Qwen3-8B", cfg), llm


def _prompts(n):
    s1 = {"type": "object", "properties": {"decision": {"type": "string", "enum": ["stop", "continue"]}},
          "required": ["decision"], "additionalProperties": False}
    s2 = json.loads(json.dumps(s1))
    s2["properties"]["decision"]["enum"].append("abstain")
    return [(f"system {i}", f"user {i}", s2 if i % 4 == 3 else s1) for i in range(n)]


def _worker(rank, world, port, n_rows, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        agent, llm = _make_agent(True)
        r1 = agent.batch_generate_json(_prompts(n_rows), temperature=0.5, max_tokens=8)
        r2 = agent.batch_generate_json(_prompts(n_rows)[:3], temperature=0.3, max_tokens=8)   # a "retry"
        stats = dp.reduce_stats({"generated_tokens": len(llm.engine.seen), "decode_s": 1.0 + rank})
        ret[rank] = (r1, r2, len(llm.engine.seen), stats)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_rows", [10, 1])
def test_two_replicas_equal_one(n_rows):
    agent, llm = _make_agent(False)
    want1 = agent.batch_generate_json(_prompts(n_rows), temperature=0.5, max_tokens=8)
    want2 = agent.batch_generate_json(_prompts(n_rows)[:3], temperature=0.3, max_tokens=8)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, port, n_rows, ret), nprocs=2, join=True)
    for rank in (0, 1):
        r1, r2, seen, stats = ret[rank]
        assert r1 == want1 and r2 == want2          # same dicts, same order, same per-row seeds
    assert ret[0][2] + ret[1][2] == n_rows + min(3, n_rows)   # every row decoded exactly once
    assert abs(ret[0][2] - ret[1][2]) <= 2
    assert ret[0][3] == {"generated_tokens": n_rows + min(3, n_rows), "decode_s": 2.0}
    from byzantine_consensus_llm_agents_b200 import vllm_agent as VA
    VA.VLLMAgent._shared_llm = VA.VLLMAgent._shared_model_name = VA.VLLMAgent._shared_model_config = None
