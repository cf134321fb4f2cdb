"""Host side of tensor parallelism on CPU (gloo, 4 processes = 2 replicas of 2 ranks): replica
groups cut out of one torch.distributed job, the leader/worker request protocol, and the rule that
a rank's failure is reported to the leader instead of killing the rank (the engines are fakes; the
kernels run in tests/test_gpu_tp.py)."""
import os
import socket

import pytest
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class _FakeEngine:
    def __init__(self, position, fail_on=None):
        self.position, self.fail_on, self.seen = position, fail_on, []

    def generate(self, reqs):
        self.seen.append(list(reqs))
        if self.fail_on is not None and reqs and reqs[0] == self.fail_on:
            raise MemoryError("KV page pool exhausted")
        return [f"r{self.position}:{q}" for q in reqs]


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from byzantine_consensus_llm_agents_b200 import tp
        from byzantine_consensus_llm_agents_b200.lib import BcgError
        group, pos, replica, leader = tp.replica_groups(2)
        assert (pos, replica, leader) == (rank % 2, rank // 2, rank - rank % 2)
        # the worker of replica 1 fails on the request "boom"
        eng = _FakeEngine(pos, fail_on="boom" if (replica == 1 and pos == 1) else None)
        out = {"replica": replica, "pos": pos}
        if pos == 0:
            out["first"] = tp.leader_generate(eng, [f"a{replica}", "b"], group, leader)
            try:
                tp.leader_generate(eng, ["boom"], group, leader)
                out["second"] = "ok"
            except BcgError as e:
                out["second"] = str(e)
            out["third"] = tp.leader_generate(eng, ["c"], group, leader)      # the replica is still usable
            tp.broadcast_requests(tp.STOP, group, leader)
        else:
            tp.worker_loop(eng, group, leader)
        out["seen"] = eng.seen
        ret[rank] = out
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_replica_groups_and_error_reporting():
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(4, _free_port(), ret), nprocs=4, join=True)
    for rank in range(4):
        r = ret[rank]
        assert r["seen"] == [[f"a{r['replica']}", "b"], ["boom"], ["c"]]       # every rank ran every call
    assert ret[0]["first"] == ["r0:a0", "r0:b"] and ret[2]["first"] == ["r0:a1", "r0:b"]
    assert ret[0]["second"] == "ok"                                            # replica 0 never failed
    assert "rank 1" in ret[2]["second"] and "KV page pool exhausted" in ret[2]["second"]
    assert ret[0]["third"] == ["r0:c"] and ret[2]["third"] == ["r0:c"]
