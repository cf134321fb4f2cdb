"""Tensor-parallel process plumbing: communicator set-up and the worker loop.

One process per GPU.  The C library owns the NCCL communicator used on the data path
(bcg_tp_init); a torch.distributed group (gloo is enough) is only the CONTROL plane: it carries
the 128-byte NCCL id at start-up and, per call, the requests from the rank that runs the game
to the ranks that merely compute (the reference gets the same from vLLM's `mp` executor,
vllm_agent.py:141-142).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch
import torch.distributed as dist

from . import lib as L

STOP = "__bcg_tp_stop__"


def init_comm(rank: int, size: int, group=None, src: int = 0) -> None:
    """Collective over `group` (ranks of one TP replica): create the library's NCCL communicator.
    `src` is the GLOBAL rank of the replica's leader; `rank` the position inside the replica."""
    lib = L.load()
    buf = (C.c_char * 128)()
    if rank == 0:
        L.check(lib.bcg_tp_unique_id(buf), "bcg_tp_unique_id")
    box = [bytes(buf)]
    dist.broadcast_object_list(box, src=src, group=group)
    ident = (C.c_char * 128).from_buffer_copy(box[0])
    L.check(lib.bcg_tp_init(ident, rank, size), "bcg_tp_init")


def setup_symmetric(nbytes: int, rank: int, size: int, group=None) -> None:
    """Collective: every rank allocates the symmetric decode buffer and maps its peers' (cudaIpc)."""
    lib = L.load()
    mine = (C.c_char * 64)()
    L.check(lib.bcg_tp_symm_create(nbytes, mine), "bcg_tp_symm_create")
    handles = [None] * size
    dist.all_gather_object(handles, bytes(mine), group=group)
    blob = (C.c_char * (64 * size)).from_buffer_copy(b"".join(handles))
    L.check(lib.bcg_tp_symm_open(blob, rank, size), "bcg_tp_symm_open")
    dist.barrier(group=group)


def shutdown_comm() -> None:
    L.load().bcg_tp_shutdown()


def broadcast_requests(reqs, group=None, src: int = 0):
    """Leader -> workers: the requests of one generate() call (or STOP)."""
    box = [reqs]
    dist.broadcast_object_list(box, src=src, group=group)
    return box[0]


def _gather_errors(err: Optional[str], group=None) -> List[Optional[str]]:
    size = dist.get_world_size(group)
    box: List[Optional[str]] = [None] * size
    dist.all_gather_object(box, err, group=group)
    return box


def _describe(e: BaseException) -> str:
    return f"{type(e).__name__}: {e}"


def worker_loop(engine, group=None, src: int = 0) -> None:
    """Ranks 1..tp-1 of a replica: run every generate() the leader runs, in lock step.  A call that
    raises on this rank (page pool exhausted, oversized request: raised before any collective, and
    on every rank alike since the ranks share pool size and requests) is reported to the leader
    instead of killing the rank, so the replica stays usable."""
    while True:
        reqs = broadcast_requests(None, group, src)
        if isinstance(reqs, str) and reqs == STOP:
            return
        err = None
        try:
            engine.generate(reqs)
        except Exception as e:  # noqa: BLE001 - reported to the leader below
            err = _describe(e)
        _gather_errors(err, group)


def leader_generate(engine, reqs, group=None, src: int = 0):
    """The leader's half of worker_loop: ship the requests, run them, then learn whether any rank
    failed; a failure anywhere is raised here (the adapter turns it into its error dicts)."""
    broadcast_requests(reqs, group, src)
    err, out = None, None
    try:
        out = engine.generate(reqs)
    except Exception as e:  # noqa: BLE001
        err = _describe(e)
    errs = _gather_errors(err, group)
    bad = [f"rank {r}: {e}" for r, e in enumerate(errs) if e]
    if bad:
        raise L.BcgError("tensor-parallel generate failed — " + "; ".join(bad))
    return out


# ---------------------------------------------------------------------------------------------
# Single-process front end: spawn the other ranks, as vLLM's `mp` executor does for the reference
# ---------------------------------------------------------------------------------------------
def _device_for(model_config: dict, rank: int) -> torch.device:
    return torch.device(f"cuda:{int(str(model_config.get('device', 'cuda:0')).split(':')[-1]) + rank}")


def _build_engine(model_name: str, model_config: dict, rank: int, size: int):
    """Engine holding rank's shard of the model on cuda:<first_device + rank>."""
    import os

    from . import model as M
    from . import vllm_agent as VA
    from .engine import Engine
    from .tokenizer import get_tokenizer
    key = model_name.lower()
    dev_index = int(str(model_config.get("device", "cuda:0")).split(":")[-1]) + rank
    device = torch.device(f"cuda:{dev_index}")
    torch.cuda.set_device(device)
    if key in VA._REGISTRY:
        cfg, factory = VA._REGISTRY[key]
        weights = factory(cfg, device).shard(rank, size)
    else:
        cfg = M.config_for(model_name)
        seed = int(model_config.get("weight_seed", os.environ.get("BCG_B200_WEIGHT_SEED", 0)))
        weights = M.Weights.random_init_sharded(cfg, device, seed, rank, size)
    tok = get_tokenizer(cfg.family, cfg.vocab)
    eng = Engine(cfg, weights, tok,
                 max_model_len=int(model_config.get("max_model_len", 8192)),
                 gpu_memory_utilization=float(model_config.get("gpu_memory_utilization", 0.85)),
                 max_rows=int(model_config.get("engine_max_rows", 64)),
                 kv_pages=model_config.get("kv_pages"),
                 base_seed=int(model_config.get("seed", os.environ.get("BCG_B200_SEED", 0))),
                 device=str(device), use_cuda_graph=bool(model_config.get("use_cuda_graph", True)),
                 pdl=bool(model_config.get("pdl", True)), tp_rank=rank, tp_size=size,
                 fused_rope=bool(model_config.get("fused_rope", False)),
                 # decode all-reduces inside bcg_tp_add_norm_stats (pushes over NVLink) unless asked
                 # for NCCL
                 tp_fused_allreduce=bool(model_config.get("tp_fused_allreduce", True)),
                 tp_group=model_config.get("tp_group"))
    return cfg, tok, eng


def _spawned_worker(rank: int, size: int, port: int, model_name: str, model_config: dict, registry: dict):
    import os
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    from . import vllm_agent as VA
    VA._REGISTRY.update(registry)
    dist.init_process_group("gloo", rank=rank, world_size=size)
    try:
        torch.cuda.set_device(_device_for(model_config, rank))   # NCCL binds to the current device
        init_comm(rank, size)
        _, _, eng = _build_engine(model_name, model_config, rank, size)
        worker_loop(eng)
        eng.close()
        torch.cuda.synchronize()
        shutdown_comm()
    finally:
        dist.destroy_process_group()


class SpawnedTP:
    """Rank 0 of a tensor-parallel replica living in the caller's process; ranks 1.. are spawned
    (one per additional GPU) and driven over a private gloo group."""

    def __init__(self, model_name: str, model_config: dict, size: int):
        import os
        import socket

        import torch.multiprocessing as mp

        from . import vllm_agent as VA
        if dist.is_available() and dist.is_initialized():
            raise L.BcgError("tensor_parallel_size > 1 from a process that already runs torch.distributed: "
                             "launch one TP replica per process group instead (see bench.py --tp)")
        if torch.cuda.device_count() < size:
            raise L.BcgError(f"tensor_parallel_size={size} needs {size} GPUs, {torch.cuda.device_count()} visible")
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            port = s.getsockname()[1]
        ctx = mp.get_context("spawn")
        self.procs = [ctx.Process(target=_spawned_worker,
                                  args=(r, size, port, model_name, dict(model_config), dict(VA._REGISTRY)),
                                  daemon=True) for r in range(1, size)]
        for p in self.procs:
            p.start()
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        dist.init_process_group("gloo", rank=0, world_size=size)
        torch.cuda.set_device(_device_for(model_config, 0))
        init_comm(0, size)
        self.cfg, self.tokenizer, self.engine = _build_engine(model_name, model_config, 0, size)
        self.size = size

    def generate(self, reqs):
        return leader_generate(self.engine, reqs)

    def close(self):
        try:
            broadcast_requests(STOP)
            self.engine.close()
            torch.cuda.synchronize()
            shutdown_comm()
        finally:
            if dist.is_initialized():
                dist.destroy_process_group()
            for p in self.procs:
                p.join(timeout=30)
                if p.is_alive():
                    p.terminate()


# ---------------------------------------------------------------------------------------------
# Replicas inside an existing torch.distributed job (torchrun): TP composes with data parallelism
# ---------------------------------------------------------------------------------------------
def replica_groups(tp_size: int):
    """Collective over the WHOLE job: cut the world into consecutive replicas of tp_size ranks and
    create one gloo control group per replica.  Returns (this rank's group, its position in the
    replica, the replica index, the global rank of the replica's leader)."""
    world, rank = dist.get_world_size(), dist.get_rank()
    if world % tp_size:
        raise L.BcgError(f"world size {world} is not a multiple of tensor_parallel_size {tp_size}")
    mine = None
    for r0 in range(0, world, tp_size):
        g = dist.new_group(list(range(r0, r0 + tp_size)), backend="gloo")   # every rank creates every group
        if r0 <= rank < r0 + tp_size:
            mine = g
    return mine, rank % tp_size, rank // tp_size, (rank // tp_size) * tp_size


class GroupTP:
    """Leader (position 0) of a replica whose other ranks are processes of the same torch.distributed
    job running serve_worker().  `model_config` carries the replica: "tp_group" (from replica_groups),
    "tp_leader" (global rank) and "device" = the leader's GPU; rank i of the replica uses the GPU
    i ordinals above it."""

    def __init__(self, model_name: str, model_config: dict, size: int):
        self.group, self.src = model_config["tp_group"], int(model_config.get("tp_leader", dist.get_rank()))
        torch.cuda.set_device(_device_for(model_config, 0))
        init_comm(0, size, self.group, self.src)
        self.cfg, self.tokenizer, self.engine = _build_engine(model_name, model_config, 0, size)
        self.size = size

    def generate(self, reqs):
        return leader_generate(self.engine, reqs, self.group, self.src)

    def close(self):
        broadcast_requests(STOP, self.group, self.src)
        self.engine.close()
        torch.cuda.synchronize()
        shutdown_comm()


def serve_worker(model_name: str, model_config: dict, position: int, size: int) -> None:
    """Positions 1..size-1 of a GroupTP replica: build this rank's shard and serve until the leader
    closes.  Call it from the worker ranks where the leader constructs its VLLMAgent."""
    group, src = model_config["tp_group"], int(model_config["tp_leader"])
    torch.cuda.set_device(_device_for(model_config, position))
    init_comm(position, size, group, src)
    _, _, eng = _build_engine(model_name, model_config, position, size)
    worker_loop(eng, group, src)
    eng.close()
    torch.cuda.synchronize()
    shutdown_comm()
