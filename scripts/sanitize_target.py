"""Target for compute-sanitizer (memcheck / racecheck / synccheck): every hand-written kernel once, at tiny shapes.
  compute-sanitizer --tool racecheck python scripts/sanitize_target.py [tcd|batch|prefill|all]
tcd      persistent tcgen05 decode kernel (3 greedy steps + 2 sampled steps, data-flow edges on), per-phase GEMV kernels
batch    tcgen05 GEMM path: 8-lane batched decode step (gemm_tc, rmsnorm, rope_append, attn_decode, swiglu, sampler)
prefill  640-id prefill in 256-row chunks (gemm_tc + attn_prefill from 512 keys on)"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, SamplingParams, geometry, weights

what = sys.argv[1] if len(sys.argv) > 1 else "all"
geo = geometry.resolve("tiny")
w = weights.random_init(geo, 0xB200, 0, sigma=0.08)
rng = np.random.default_rng(0)
if what in ("tcd", "all"):
    eng = Engine(geo, w, max_lanes=2, max_ctx=128, max_rows=8, use_tc=False)
    lane = eng.open_lane()
    eng.prefill(lane, rng.integers(0, geo.vocab, 40, dtype=np.int32))
    print("tcd greedy", eng.decode([lane], 3)[:, 0].tolist(), "launches", eng.timing().kernel_launches)
    eng.set_sampling(lane, SamplingParams(temperature=0.8, top_k=40, top_p=0.9, repeat_penalty=1.1, seed=3))
    print("tcd sampled", eng.decode([lane], 2)[:, 0].tolist(), "launches", eng.timing().kernel_launches)
    eng.close()
if what in ("batch", "all"):
    eng = Engine(geo, w, max_lanes=8, max_ctx=128, max_rows=16, use_tc=True)
    lanes = [eng.open_lane() for _ in range(8)]
    for i, ln in enumerate(lanes):
        eng.prefill(ln, rng.integers(0, geo.vocab, 10 + 3 * i, dtype=np.int32))
    print("batch", eng.decode(lanes, 2)[-1].tolist())
    eng.close()
if what in ("prefill", "all"):
    eng = Engine(geo, w, max_lanes=1, max_ctx=704, max_rows=256, use_tc=True)
    lane = eng.open_lane()
    eng.prefill(lane, rng.integers(0, geo.vocab, 640, dtype=np.int32))
    print("prefill first token", eng.read_lane(lane)[0])
    eng.close()
