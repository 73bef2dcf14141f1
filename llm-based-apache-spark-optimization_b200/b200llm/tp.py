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


# environment switches that decide which kernels a rank launches: ranks that disagree on one of them would leave the
# lock-step of the reductions (one rank in the persistent kernel, its peer in the per-phase kernels) and wait for each other
PATH_SWITCHES = ("B200LLM_NO_TCD", "B200LLM_NO_TC", "B200LLM_PERSISTENT", "B200LLM_TCD_STAGES", "B200LLM_TCD_DATAFLOW",
                 "B200LLM_TC_MIN_ROWS", "B200LLM_NO_PREFILL_ATTN", "B200LLM_PREFILL_BALANCED", "B200LLM_TC_BF16ACT")


def check_same_path(group=None, extra=()) -> None:
    """Collective: raise on every rank unless all ranks of `group` see the same path-selecting settings (and `extra`,
    e.g. the engine's use_megakernel / use_tc / max_rows)."""
    import os
    mine = tuple(os.environ.get(k) for k in PATH_SWITCHES) + tuple(extra)
    box = [None] * dist.get_world_size(group)
    dist.all_gather_object(box, mine, group=group)
    if any(b != box[0] for b in box):
        bad = [k for i, k in enumerate(PATH_SWITCHES + tuple(f"extra[{j}]" for j in range(len(extra))))
               if any(b[i] != box[0][i] for b in box)]
        raise RuntimeError(f"tensor-parallel ranks disagree on {bad}: {box}")


def connect_peers(engine, group=None) -> None:
    """Collective: switch `engine` (one per rank of `group`, all on one node) from ncclAllReduce to the
    one-shot all-reduce over NVLink peer memory (csrc/tp_ar.cu).  Every rank exports the CUDA IPC handle
    of its inbox region, the handles travel over torch.distributed, every rank opens its peers'."""
    size = dist.get_world_size(group)
    check_same_path(group, (bool(engine.cfg.use_megakernel), bool(engine.cfg.use_tc), int(engine.cfg.max_rows), int(engine.cfg.use_graphs)))
    buf = C.create_string_buffer(64)
    _lib.check(_lib.lib.b200llm_tp_export(engine._h, buf))
    box = [None] * size
    dist.all_gather_object(box, bytes(buf.raw), group=group)
    blob = C.create_string_buffer(b"".join(box), 64 * size)
    _lib.check(_lib.lib.b200llm_tp_import(engine._h, blob, size))
    dist.barrier(group)     # nobody reduces before everybody has opened everybody's region
