"""CPU, world_size 2, gloo: the host-side logic of the multi-GPU paths.
Replicas need no collective at all (bench.py only reduces the timing); tensor parallel needs the
ranks to agree on one NCCL id and to hold complementary shards."""
import os
import subprocess
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")


def _worker(rank, world, port, out_dir):
    for p in (ROOT, PKG):
        sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from b200llm import _lib, geometry, tp, weights
    from oracle import llama_oracle as lo
    # 1. every rank ends up with rank 0's 128-byte id (the library call is replaced: no GPU/NCCL here)
    real = _lib.lib.b200llm_nccl_unique_id
    class Fake:                                        # noqa: E306
        def __call__(self, buf):
            buf.raw = bytes(range(128))
            return 0
    _lib.lib.b200llm_nccl_unique_id = Fake()
    uid = tp.exchange_unique_id()
    _lib.lib.b200llm_nccl_unique_id = real
    assert uid == bytes(range(128)) and len(uid) == 128
    # 2. complementary shards: gather both ranks' packed tensors and rebuild the full matrices
    geo = geometry.resolve("tiny")
    sd = lo.make_weights(lo.GEOMETRIES["tiny"], 3)
    mine = {(k, l): t for k, l, t in weights.pack(geo, sd, rank, world).tensors}
    box = [None] * world
    dist.all_gather_object(box, {k: v.view(torch.int16).numpy() for k, v in mine.items() if k[0] in (_lib.W_O, _lib.W_DOWN,
                                                                                                     _lib.W_LM_HEAD)})
    if rank == 0:
        import numpy as np
        o = np.concatenate([b[(_lib.W_O, 1)] for b in box], 1)
        assert np.array_equal(o, sd["model.layers.1.self_attn.o_proj.weight"].view(torch.int16).numpy())
        d = np.concatenate([b[(_lib.W_DOWN, 0)] for b in box], 1)
        assert np.array_equal(d, sd["model.layers.0.mlp.down_proj.weight"].view(torch.int16).numpy())
        v = np.concatenate([b[(_lib.W_LM_HEAD, -1)] for b in box], 0)
        assert np.array_equal(v, sd["model.embed_tokens.weight"].view(torch.int16).numpy())
    assert tp.shard_bounds(geo.ffn, rank, world) == (rank * geo.ffn // world, (rank + 1) * geo.ffn // world)
    # 2b. ranks must agree on the switches that select the kernels (a rank alone in the persistent kernel would wait for
    # its peers for ever): same settings pass, a switch set on one rank only is named in the error on EVERY rank
    tp.check_same_path(extra=(True, 256))
    if rank == 1:
        os.environ["B200LLM_NO_TCD"] = "1"
    try:
        tp.check_same_path(extra=(True, 256))
        raise AssertionError("a disagreement went unnoticed")
    except RuntimeError as e:
        assert "B200LLM_NO_TCD" in str(e)
    os.environ.pop("B200LLM_NO_TCD", None)
    try:
        tp.check_same_path(extra=(True, 256 + rank))
        raise AssertionError("a disagreement went unnoticed")
    except RuntimeError as e:
        assert "extra[1]" in str(e)
    # 3. the timing reduction bench.py does: max over ranks
    t = torch.tensor([1.0 + rank], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    assert float(t) == float(world)
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")


def test_two_rank_host_logic(tmp_path):
    port = 29600 + os.getpid() % 300
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert sorted(os.listdir(tmp_path)) == ["ok0", "ok1"]


def test_reference_arm_only_rank0_prints():
    """Under torchrun the reference arm runs on rank 0 alone; other ranks exit 0 silently."""
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "0"], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""
