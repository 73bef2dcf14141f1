"""Tensor-parallel bring-up for the 7B model: one process per GPU, Megatron
sharding done by weights.pack(..., tp_rank, tp_size), NCCL all-reduce inside the engine.  The
host only has to agree on the NCCL unique id: rank 0 asks the library for it and every rank
receives it over torch.distributed (any backend; gloo works)."""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from . import _lib


def exchange_unique_id(group=None) -> bytes:
    """Collective: returns the same 128-byte ncclUniqueId on every rank of `group`."""
    rank = dist.get_rank(group)
    buf = C.create_string_buffer(128)
    if rank == 0:
        _lib.check(_lib.lib.b200llm_nccl_unique_id(buf))
    box = [bytes(buf.raw) if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    return box[0]


def shard_bounds(n: int, rank: int, size: int):
    """[begin, end) of rank's contiguous share of n items (heads, ffn rows, vocab rows)."""
    assert n % size == 0
    per = n // size
    return rank * per, (rank + 1) * per


def connect_peers(engine, group=None) -> None:
    """Collective: switch `engine` (one per rank of `group`, all on one node) from ncclAllReduce to the
    one-shot all-reduce over NVLink peer memory (csrc/tp_ar.cu).  Every rank exports the CUDA IPC handle
    of its inbox region, the handles travel over torch.distributed, every rank opens its peers'."""
    size = dist.get_world_size(group)
    buf = C.create_string_buffer(64)
    _lib.check(_lib.lib.b200llm_tp_export(engine._h, buf))
    box = [None] * size
    dist.all_gather_object(box, bytes(buf.raw), group=group)
    blob = C.create_string_buffer(b"".join(box), 64 * size)
    _lib.check(_lib.lib.b200llm_tp_import(engine._h, blob, size))
    dist.barrier(group)     # nobody reduces before everybody has opened everybody's region
