"""Game-level data-parallel driver: jobs sharded over replicas, seeded per
job, gathered in job order, flattened into one CSV.  CPU only: the simulation is a stub, except in
the last test, which runs the UNMODIFIED reference `main.run_simulation` on the drop-in adapter
with a scripted engine (needs /root/reference, so it is skipped on the GPU box)."""
import csv
import json
import os
import random
import socket
from pathlib import Path

import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from byzantine_consensus_llm_agents_b200 import games

HAVE_REF = Path("/root/reference/byzantine_consensus_game/main.py").exists()


def _stub_sim(n_agents=8, max_rounds=50, model_name=None, byzantine_count=0, byzantine_awareness="may_exist"):
    """Like the reference: all game randomness comes from the `random` module."""
    if n_agents == 13:
        raise ValueError("unlucky")
    rounds = random.randint(1, max_rounds)
    return {"metrics": {"total_agents": n_agents, "total_rounds": rounds,
                        "consensus_value": random.randint(0, 50),
                        "honest_initial_values": [random.randint(0, 50) for _ in range(n_agents - byzantine_count)],
                        "keyword_counts": {"byzantine": byzantine_count, "outlier": 0},
                        "byzantine_awareness": byzantine_awareness}}


def test_jobs_are_seeded_and_ordered(tmp_path):
    jobs = games.make_jobs(5, base_seed=100, n_agents=10, max_rounds=20, byzantine_count=2)
    assert [j["seed"] for j in jobs] == [100, 101, 102, 103, 104]
    rows = games.run_games(_stub_sim, jobs)
    again = games.run_games(_stub_sim, jobs)
    assert [r["game"] for r in rows] == list(range(5))
    assert [r["metrics"] for r in rows] == [r["metrics"] for r in again]      # reproducible per seed
    assert len({json.dumps(r["metrics"]) for r in rows}) > 1                  # and seeds matter
    random.seed(102)
    assert rows[2]["metrics"] == _stub_sim(n_agents=10, max_rounds=20, byzantine_count=2)["metrics"]
    header = games.write_csv(rows, str(tmp_path / "g.csv"))
    assert {"game", "replica", "seed", "wall_s", "n_agents", "total_rounds", "keyword_counts.byzantine",
            "honest_initial_values", "error"} <= set(header)
    back = list(csv.DictReader(open(tmp_path / "g.csv")))
    assert len(back) == 5 and back[3]["seed"] == "103"
    assert json.loads(back[0]["honest_initial_values"]) == rows[0]["metrics"]["honest_initial_values"]
    s = games.summarize(rows)
    assert s["games"] == 5 and s["failed"] == 0 and s["replicas"] == 1
    assert s["agent_steps"] == 10 * sum(r["metrics"]["total_rounds"] for r in rows)


def test_a_failed_game_does_not_lose_the_sweep():
    jobs = games.make_jobs(2, n_agents=4) + games.make_jobs(1, n_agents=13)
    rows = games.run_games(_stub_sim, jobs)
    assert rows[2]["error"] == "ValueError: unlucky" and rows[2]["metrics"] == {}
    assert rows[0]["error"] is None
    assert games.summarize(rows)["failed"] == 1
    with pytest.raises(TypeError):
        games.make_jobs(1, agents=4)


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        jobs = games.make_jobs(7, base_seed=7, n_agents=6, max_rounds=9, byzantine_count=1)
        ret[rank] = games.run_games(_stub_sim, jobs)
    finally:
        dist.destroy_process_group()


def test_two_replicas_equal_one():
    want = games.run_games(_stub_sim, games.make_jobs(7, base_seed=7, n_agents=6, max_rounds=9, byzantine_count=1))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
    for rank in (0, 1):
        got = ret[rank]
        assert [r["metrics"] for r in got] == [r["metrics"] for r in want]     # job order, same games
        assert [r["replica"] for r in got] == [0, 1, 0, 1, 0, 1, 0]           # round-robin, each job once
    assert games.summarize(ret[0])["replicas"] == 2


@pytest.mark.skipif(not HAVE_REF, reason="reference tree not available on this machine")
def test_reference_run_simulation_on_the_dropin_adapter():
    import sys
    from byzantine_consensus_llm_agents_b200 import vllm_agent as VA
    from oracle import ref_harness as H
    import importlib.util
    spec = importlib.util.spec_from_file_location("_adapter_dropin_helpers", Path(__file__).with_name("test_adapter_dropin.py"))
    helpers = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(helpers)
    FakeLLM, Scripted = helpers.FakeLLM, helpers.Scripted

    H.install_vllm_shim(None)
    for m in H.REFERENCE_MODULES:
        sys.modules.pop(m, None)
    run_simulation = games.load_reference_run_simulation(str(H.REFERENCE_DIR))
    s = Scripted()
    VA.VLLMAgent._shared_llm = FakeLLM(lambda p, sc, t, m: [s.text_for(a, b, t) for a, b in zip(p, sc)])
    VA.VLLMAgent._shared_model_name = "Qwen/Qwen3-14B"
    VA.VLLMAgent._shared_model_config = dict(sys.modules["config"].VLLM_CONFIG)
    cwd = os.getcwd()
    try:
        jobs = games.make_jobs(3, base_seed=1234, n_agents=4, max_rounds=2, byzantine_count=1,
                               model_name="Qwen/Qwen3-14B")
        rows = games.run_games(run_simulation, jobs)
    finally:
        os.chdir(cwd)
        VA.VLLMAgent._shared_llm = VA.VLLMAgent._shared_model_name = VA.VLLMAgent._shared_model_config = None
        H.uninstall()
    assert [r["error"] for r in rows] == [None, None, None]
    for r in rows:
        assert r["metrics"]["total_agents"] == 4 and 1 <= r["metrics"]["total_rounds"] <= 2
        assert r["metrics"]["byzantine_awareness"] == "may_exist"
    assert games.summarize(rows)["agent_steps"] == 4 * sum(r["metrics"]["total_rounds"] for r in rows)
