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
