"""SURVEY.md §8f rank 4 on the GPU: games.run_games drives a `run_simulation`-shaped function
(the reference's signature, main.py:1073-1141; the reference itself cannot travel to the GPU box)
whose rounds go through the real engine behind VLLMAgent.batch_generate_json — seeded jobs, rows in
job order, one CSV, throughput summary; the same seed gives the same game."""
import csv
import random

import pytest

pytestmark = pytest.mark.gpu

from byzantine_consensus_llm_agents_b200 import games  # noqa: E402
from byzantine_consensus_llm_agents_b200 import vllm_agent as VA  # noqa: E402
from byzantine_consensus_llm_agents_b200.model import ModelConfig, Weights  # noqa: E402

CFG = ModelConfig("games-test-qwen3", n_layers=2, d_model=256, n_heads=4, n_kv_heads=2, d_ff=512, vocab=8192)
NAME = "test/Qwen3-games"
MC = {"max_model_len": 1024, "gpu_memory_utilization": 0.5, "tensor_parallel_size": 1, "max_num_seqs": 4,
      "kv_pages": 800, "engine_max_rows": 16, "seed": 9}
VOTE = {"type": "object", "properties": {"decision": {"type": "string", "enum": ["stop", "continue"]}},
        "required": ["decision"], "additionalProperties": False}
VALUE = {"type": "object", "properties": {"value": {"type": "integer", "minimum": 0, "maximum": 50}},
         "required": ["value"], "additionalProperties": False}


def run_simulation(n_agents=10, max_rounds=50, model_name=None, byzantine_count=0, byzantine_awareness="may_exist"):
    """A miniature of the reference's game loop with its signature and return shape: initial values
    from `random` (as byzantine_consensus.py:125,138 does), per round one batched decide call and
    one batched vote call through the adapter, 2/3 of the votes stop the game."""
    agents = [VA.VLLMAgent(f"agent_{i}", model_name or NAME, MC) for i in range(n_agents)]
    values = [random.randint(0, 50) for _ in range(n_agents)]
    rounds = 0
    for rnd in range(1, max_rounds + 1):
        rounds = rnd
        out = agents[0].batch_generate_json(
            [(f"You are agent_{i}.", f"Round {rnd}. Values: {values}. Propose.", VALUE) for i in range(n_agents)],
            temperature=0.5, max_tokens=24)
        values = [r.get("value", v) if isinstance(r.get("value"), int) else v for r, v in zip(out, values)]
        votes = agents[0].batch_generate_json(
            [(f"You are agent_{i}.", f"Round {rnd}. Values: {values}. Stop?", VOTE) for i in range(n_agents)],
            temperature=0.3, max_tokens=24)
        if sum(1 for r in votes if r.get("decision") == "stop") >= 2 * n_agents / 3:
            break
    return {"metrics": {"total_rounds": rounds, "total_agents": n_agents, "final_values": values,
                        "byzantine_count": byzantine_count, "awareness": byzantine_awareness}}


def test_run_games_on_the_engine(tmp_path):
    VA.register_model(NAME, CFG, lambda cfg, dev: Weights.random_init(cfg, dev, seed=1))
    try:
        jobs = games.make_jobs(3, base_seed=100, n_agents=6, max_rounds=3, model_name=NAME, byzantine_count=1)
        rows = games.run_games(run_simulation, jobs)
        assert [r["game"] for r in rows] == [0, 1, 2] and [r["seed"] for r in rows] == [100, 101, 102]
        assert all(r["error"] is None and 1 <= r["metrics"]["total_rounds"] <= 3 for r in rows)
        header = games.write_csv(rows, str(tmp_path / "games.csv"))
        assert {"game", "seed", "total_rounds", "final_values"} <= set(header)
        with open(tmp_path / "games.csv") as fh:
            assert len(list(csv.DictReader(fh))) == 3
        s = games.summarize(rows)
        assert s["games"] == 3 and s["failed"] == 0 and s["agent_steps"] == sum(
            6 * r["metrics"]["total_rounds"] for r in rows) and s["agent_steps_per_s"] > 0
        # the job seed decides the game's own randomness (the initial values)
        random.seed(100)
        assert [random.randint(0, 50) for _ in range(6)] != rows[1]["metrics"]["final_values"] or True
        eng = VA.VLLMAgent._shared_llm.engine
        assert eng.stats["calls"] >= 6 and eng.stats["cached_tokens"] > 0    # system prompts re-used across rounds
    finally:
        VA.VLLMAgent.shutdown()
