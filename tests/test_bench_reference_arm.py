"""The reference arm of bench.py (`--impl reference`) runs on the host cores only, so its contract can be
checked here: one JSON line with the keys the driver reads, exactly K timed samples, rank 0 alone works
under a multi-rank launch.  Uses BASELINE configs[0]'s architecture (Qwen3-0.6B, 4 + 0 agents) to stay
short; the arithmetic is the CPU oracle's (oracle/engine.py), which bench.py may execute for this arm."""
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
CMD = [sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--model", "qwen3-0.6b", "--honest", "4",
       "--byzantine", "0"]


def _run(extra, env_extra=None):
    env = dict(os.environ)
    env.update(env_extra or {})
    return subprocess.run(CMD + extra, cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)


def test_reference_arm_prints_the_contract_line():
    r = _run(["--steps", "2", "--warmup", "1"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [x for x in r.stdout.splitlines() if x.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "agent_steps_per_sec" and d["unit"] == "agent-steps/s"
    assert d["higher_is_better"] is True and d["steps"] == 2 and d["warmup"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0
    assert d["config"]["agents"] == 4 and "qwen3-0.6b" in d["config"]["workload"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["extrapolated"] is True and cb["samples"] == 2
    assert cb["cores"] >= 1 and cb["value"] == d["value"] and "oracle engine" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "agent-steps/s", "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": 0}
    # the claimed time is what ran: K samples of ms_per_step each fit the process's wall time
    assert d["steps"] * d["ms_per_step"] / 1e3 < d["wall_s"]


def test_reference_arm_other_ranks_do_nothing():
    r = _run(["--gpus", "2", "--steps", "1", "--warmup", "0"], {"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"})
    assert r.returncode == 0, r.stderr[-2000:]
    assert not [x for x in r.stdout.splitlines() if x.startswith("{")]
