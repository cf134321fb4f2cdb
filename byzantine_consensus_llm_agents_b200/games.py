"""Game-level data parallelism: many seeded games per box, one engine
replica per GPU, results gathered into one table.

Whole games are the unit that really scales with GPUs (§8e: inside one game every replica would
re-read the same weights each decode step).  The reference offers `run_simulation(n_agents,
max_rounds, model_name, byzantine_count, byzantine_awareness) -> {"metrics": stats}`
(main.py:1073-1141) for batch experiments but no driver around it; this module is that driver:

  * `run_games(run_simulation, jobs)` — job i goes to replica i % world (`dp.shard_rows`), each
    replica seeds Python's `random` with the job's seed (the reference's only source of game
    randomness: initial values and Byzantine assignment use the `random` module) and calls the
    reference function unchanged; the metric dicts travel as host objects
    (`all_gather_object`), never through a device collective; every rank returns all rows in
    job order, so rank 0 can write them out.
  * `write_csv(rows, path)` — one line per game, metric keys flattened (`a.b` for nested dicts).
  * `python -m byzantine_consensus_llm_agents_b200.games --reference-dir …` — loads the
    unmodified reference `main.py` with THIS package's `vllm_agent` in place of the
    reference's (cut line B1) and runs a sweep, under plain `python` (one GPU)
    or `torchrun --nproc-per-node N` (one replica per GPU).
"""
from __future__ import annotations

import csv
import json
import random
import time
from typing import Callable, Dict, List, Optional, Sequence

from . import dp

JOB_KEYS = ("n_agents", "max_rounds", "model_name", "byzantine_count", "byzantine_awareness")


def make_jobs(n_games: int, base_seed: int = 1234, **sim_kwargs) -> List[dict]:
    """n_games jobs that differ only in their seed (base_seed + i)."""
    unknown = set(sim_kwargs) - set(JOB_KEYS)
    if unknown:
        raise TypeError(f"unknown run_simulation arguments: {sorted(unknown)}")
    return [dict(sim_kwargs, seed=base_seed + i) for i in range(n_games)]


def _run_one(run_simulation: Callable[..., dict], job: dict, index: int, rank: int) -> dict:
    kw = {k: v for k, v in job.items() if k != "seed"}
    random.seed(job.get("seed"))
    t0 = time.perf_counter()
    try:
        out = run_simulation(**kw)
        metrics = out.get("metrics", out) if isinstance(out, dict) else {"result": out}
        err = None
    except Exception as e:  # one failed game must not lose the sweep
        metrics, err = {}, f"{type(e).__name__}: {e}"
    row = {"game": index, "replica": rank, "seed": job.get("seed"), "wall_s": time.perf_counter() - t0}
    row.update({k: v for k, v in kw.items()})
    row["error"] = err
    row["metrics"] = metrics
    return row


def run_games(run_simulation: Callable[..., dict], jobs: Sequence[dict], group=None) -> List[dict]:
    """Run every job once, sharded over the replicas of the default process group.

    Returns one row per job, in job order, on every rank:
    {"game", "replica", "seed", "wall_s", <job kwargs>, "error", "metrics"}.
    """
    rank, world = dp.dist_info()
    mine = dp.shard_rows(len(jobs), rank, world)
    local = [_run_one(run_simulation, jobs[i], i, rank) for i in mine]
    return dp.gather_rows(local, len(jobs), rank, world, group)


def _flatten(prefix: str, value, out: Dict[str, object]) -> None:
    if isinstance(value, dict):
        for k, v in value.items():
            _flatten(f"{prefix}.{k}" if prefix else str(k), v, out)
    elif isinstance(value, (list, tuple)):
        out[prefix] = json.dumps(value)
    else:
        out[prefix] = value


def flatten_row(row: dict) -> Dict[str, object]:
    flat: Dict[str, object] = {}
    for k, v in row.items():
        if k == "metrics":
            _flatten("", v, flat)
        else:
            flat[k] = v
    return flat


def write_csv(rows: Sequence[dict], path: str) -> List[str]:
    """One line per game; the header is the union of all keys in first-seen order."""
    flat = [flatten_row(r) for r in rows]
    header: List[str] = []
    for f in flat:
        for k in f:
            if k not in header:
                header.append(k)
    with open(path, "w", newline="") as fh:
        w = csv.DictWriter(fh, fieldnames=header)
        w.writeheader()
        for f in flat:
            w.writerow(f)
    return header


def summarize(rows: Sequence[dict]) -> dict:
    """Aggregate throughput of a sweep: games / slowest replica's busy time."""
    busy: Dict[int, float] = {}
    for r in rows:
        busy[r["replica"]] = busy.get(r["replica"], 0.0) + r["wall_s"]
    rounds = sum(int(r["metrics"].get("total_rounds", 0) or 0) for r in rows if r["metrics"])
    steps = sum(int(r["metrics"].get("total_rounds", 0) or 0) *
                int(r["metrics"].get("total_agents") or r.get("n_agents") or 0)
                for r in rows if r["metrics"])
    span = max(busy.values()) if busy else 0.0
    return {"games": len(rows), "failed": sum(1 for r in rows if r["error"]), "replicas": len(busy),
            "rounds": rounds, "agent_steps": steps, "span_s": span,
            "agent_steps_per_s": (steps / span) if span > 0 else None}


def load_reference_run_simulation(reference_dir: str) -> Callable[..., dict]:
    """Import the reference's main.py with this package's adapter standing in for vllm_agent.py."""
    import importlib
    import sys
    from . import vllm_agent as ours
    if reference_dir not in sys.path:
        sys.path.insert(0, reference_dir)
    sys.modules["vllm_agent"] = ours
    main = importlib.import_module("main")
    return main.run_simulation


def main(argv: Optional[Sequence[str]] = None) -> int:
    import argparse
    import os
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--reference-dir", required=True, help="directory holding the reference's main.py")
    ap.add_argument("--games", type=int, default=8)
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--n-agents", type=int, default=10)
    ap.add_argument("--byzantine", type=int, default=2)
    ap.add_argument("--max-rounds", type=int, default=50)
    ap.add_argument("--model-name", default=None)
    ap.add_argument("--awareness", default="may_exist", choices=["may_exist", "none_exist"])
    ap.add_argument("--out", default="games.csv")
    a = ap.parse_args(argv)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        import torch
        import torch.distributed as dist
        if torch.cuda.is_available():
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        # host objects only: gloo is enough and keeps NCCL out of a path with no device exchange
        dist.init_process_group("gloo")
    run_simulation = load_reference_run_simulation(a.reference_dir)
    jobs = make_jobs(a.games, a.seed, n_agents=a.n_agents, max_rounds=a.max_rounds,
                     model_name=a.model_name, byzantine_count=a.byzantine,
                     byzantine_awareness=a.awareness)
    rows = run_games(run_simulation, jobs)
    rank, _ = dp.dist_info()
    if rank == 0:
        write_csv(rows, a.out)
        print(json.dumps(summarize(rows)))
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
