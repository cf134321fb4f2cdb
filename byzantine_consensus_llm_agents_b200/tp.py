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


def worker_loop(engine, group=None, src: int = 0) -> None:
    """Ranks 1..tp-1 of a replica: run every generate() the leader runs, in lock step."""
    while True:
        reqs = broadcast_requests(None, group, src)
        if isinstance(reqs, str) and reqs == STOP:
            return
        engine.generate(reqs)


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
                 pdl=bool(model_config.get("pdl", False)), tp_rank=rank, tp_size=size,
                 # measured: the peer-memory reduce beats NCCL at TP=2 (8.65 vs 9.70 ms/step, Mistral-22B)
                 # and loses to NCCL's in-switch reduction at TP=8 (13.4 vs 8.9 ms/step, Qwen3-32B)
                 tp_fused_allreduce=bool(model_config.get("tp_fused_allreduce", size == 2)))
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
        broadcast_requests(reqs)
        return self.engine.generate(reqs)

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
