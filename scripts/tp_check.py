"""torchrun --nproc-per-node 2 scripts/tp_check.py : tensor-parallel engine vs the CPU oracle."""
import os, sys
import numpy as np, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, geometry, weights, tp
from oracle import llama_oracle as lo, weightgen as wg

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("gloo")
ok = True
for name in (["small", "duckdb-nsql"] if len(sys.argv) < 2 else sys.argv[1:]):
    geo = geometry.resolve(name)
    uid = tp.exchange_unique_id()          # one id per communicator
    if name == "small":
        ogeo = lo.GEOMETRIES[name]
        sd = lo.make_weights(ogeo, 0xB200, sigma=0.06)
        prompt = wg.gen_tokens(45, 9, geo.vocab)
        want, lg, margins = lo.LlamaOracle(ogeo, sd, kv_dtype=torch.bfloat16).greedy(prompt, 16)
        w = weights.pack(geo, {k: v.cuda() for k, v in sd.items()}, rank, world)
    else:
        prompt = wg.gen_tokens(64, 9, geo.vocab)
        want = None
        # shard the device-generated logical tensors layer by layer (full model never resident twice)
        w = weights.random_init(geo, 0xB200, local, tp_rank=rank, tp_size=world)
    eng = Engine(geo, w, device=local, max_lanes=2, max_ctx=512, max_rows=16, tp_rank=rank, tp_size=world, nccl_id=uid)
    if os.environ.get("B200LLM_TP_NCCL_ONLY") != "1":
        tp.connect_peers(eng)      # one-shot all-reduce over NVLink peer memory instead of ncclAllReduce
    lane = eng.open_lane()
    eng.prefill(lane, prompt)
    first, _ = eng.read_lane(lane)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 15 if name == "small" else 64
    ev0.record(eng.stream)
    out = eng.decode([lane], n)[:, 0].tolist()
    ev1.record(eng.stream); torch.cuda.synchronize()
    got = [first] + out
    box = [None] * world
    dist.all_gather_object(box, got)
    same = all(b == box[0] for b in box)
    msg = (f"[{name}] tp={world} rank {rank}: ranks agree={same} ms/step={ev0.elapsed_time(ev1)/n:.3f} "
           f"launches per decode call={eng.timing().kernel_launches}")
    if want is not None:
        bad = [i for i in range(len(want)) if got[i] != want[i]]
        # a first mismatch is tolerated only where the oracle itself is at a near-tie (top-1 minus top-2 logit below
        # 5e-3: inside the engine's 3e-3 relative logit tolerance), as in tests/test_gpu_engine.py
        near_tie = bool(bad) and margins[bad[0]] < 5e-3
        msg += f" oracle match={got == want or near_tie}"
        if bad:
            msg += f" (first difference at step {bad[0]}: got {got[bad[0]]} want {want[bad[0]]}, oracle margin {margins[bad[0]]:.3e})"
        ok = ok and (got == want or near_tie)
    ok = ok and same
    if rank == 0:
        print(msg, got[:8], flush=True)
    eng.close()
dist.barrier()
if rank == 0:
    print("TP CHECK", "OK" if ok else "FAILED", flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
