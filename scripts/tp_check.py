"""torchrun --nproc-per-node N scripts/tp_check.py [small] [duckdb-nsql] : tensor-parallel engines vs the oracle.

small        live CPU oracle (greedy tokens equal, near-ties flagged).
duckdb-nsql  the committed golden vectors tests/golden/duckdb-nsql.npz (config 4's 2048 + 24-id prompt): 64 greedy
             tokens equal on every rank, prefill logits at the probe ids within 3e-3 of the oracle's (each probe id is
             checked by the rank that owns its vocabulary rows), one launch per decode call at every TP degree.
Also prints TTFT and ms/step, so the same run is the TP measurement kept under profiles/.
"""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, _lib, geometry, tp, weights            # noqa: E402
from oracle import llama_oracle as lo, weightgen as wg              # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("gloo")
ok = True
for name in (["small", "duckdb-nsql"] if len(sys.argv) < 2 else sys.argv[1:]):
    geo = geometry.resolve(name)
    uid = tp.exchange_unique_id()          # one id per communicator
    golden = None
    if name == "small":
        ogeo = lo.GEOMETRIES[name]
        sd = lo.make_weights(ogeo, 0xB200, sigma=0.06)
        prompt = wg.gen_tokens(45, 9, geo.vocab)
        n_new = 16
        want, lg, margins = lo.LlamaOracle(ogeo, sd, kv_dtype=torch.bfloat16).greedy(prompt, n_new)
        w = weights.pack(geo, {k: v.cuda() for k, v in sd.items()}, rank, world)
    else:
        fname = os.path.join(ROOT, "tests", "golden", name.replace(":", "_") + ".npz")
        golden = np.load(fname)
        prompt, want, margins = golden["prompt"], golden["tokens"].tolist(), golden["margins"]
        n_new = len(want)
        # shard the device-generated logical tensors layer by layer (full model never resident twice)
        w = weights.random_init(geo, int(golden["seed"]), local, tp_rank=rank, tp_size=world, sigma=float(golden["sigma"]))
    eng = Engine(geo, w, device=local, max_lanes=2, max_ctx=len(prompt) + n_new + 64, max_rows=256 if golden is not None else 16,
                 tp_rank=rank, tp_size=world, nccl_id=uid)
    if os.environ.get("B200LLM_TP_NCCL_ONLY") != "1":
        tp.connect_peers(eng)      # one-shot all-reduce over NVLink peer memory instead of ncclAllReduce
    lane = eng.open_lane()
    eng.prefill(lane, prompt)      # (first pass: tensor maps, NCCL connections)
    eng.sync()
    eng.prefill(lane, prompt)
    first, _ = eng.read_lane(lane)
    ttft = eng.timing().prefill_ms
    logit_ok, worst = True, 0.0
    if golden is not None:         # prefill logits at the probe ids this rank owns
        vl = geo.vocab // world
        mine = eng.logits(0)
        for j, tid in enumerate(golden["probe_ids"].tolist()):
            if rank * vl <= tid < (rank + 1) * vl:
                worst = max(worst, abs(float(mine[tid - rank * vl]) - float(golden["probe_kvbf16"][0][j])) / float(golden["absmax_kvbf16"][0]))
        logit_ok = worst <= 3e-3
    out = eng.decode([lane], n_new - 1)[:, 0].tolist()
    tm = eng.timing()
    got = [first] + out
    box = [None] * world
    dist.all_gather_object(box, (got, logit_ok, worst, tm.decode_ms / (n_new - 1), ttft))
    same = all(b[0] == box[0][0] for b in box)
    bad = [i for i in range(len(want)) if got[i] != want[i]]
    # a first mismatch is tolerated only where the oracle itself is at a near-tie (top-1 minus top-2 logit below
    # 5e-3: inside the engine's 3e-3 relative logit tolerance), as in tests/test_gpu_engine.py
    near_tie = bool(bad) and float(margins[bad[0]]) < 5e-3
    match = got == want or near_tie
    msg = (f"[{name}] tp={world}: ranks agree={same} oracle match={match} logits ok={all(b[1] for b in box)} "
           f"(worst {max(b[2] for b in box):.1e}) ms/step={max(b[3] for b in box):.3f} ttft_ms={max(b[4] for b in box):.2f} "
           f"({len(prompt)} ids) launches per decode call={tm.kernel_launches}")
    if bad:
        msg += f" (first difference at step {bad[0]}: got {got[bad[0]]} want {want[bad[0]]}, oracle margin {float(margins[bad[0]]):.3e})"
    us = C.c_float()
    _lib.check(_lib.lib.b200llm_debug_tp_allreduce_us(eng._h, 1, 256, C.byref(us)))
    msg += f" all-reduce [1, d] {us.value:.2f} us"
    ok = ok and same and match and all(b[1] for b in box)
    if rank == 0:
        print(msg, got[:8], flush=True)
    eng.close()
    del eng, w
    torch.cuda.empty_cache()
dist.barrier()
if rank == 0:
    print("TP CHECK", "OK" if ok else "FAILED", flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
