"""Decode tok/s at 1..4 lanes on the persistent kernel (one launch for all lanes) against the per-phase / GEMM paths:
  python scripts/bench_lanes.py [model] [--steps 64] [--sampled]
B200LLM_TCD_LANES=1 in the environment turns the multi-lane launch off (2 lanes: GEMV kernels, 3-4: GEMM path)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, SamplingParams, geometry, weights

args = [a for a in sys.argv[1:] if not a.startswith("--")]
geo = geometry.resolve(args[0] if args else "llama3.2")
steps = int(sys.argv[sys.argv.index("--steps") + 1]) if "--steps" in sys.argv else 64
eng = Engine(geo, weights.random_init(geo, 0xB200, 0), max_lanes=4, max_ctx=296 + 2 * steps + 64, max_rows=256)
rng = np.random.default_rng(1)
lanes = [eng.open_lane() for _ in range(4)]
for i, ln in enumerate(lanes):
    if "--sampled" in sys.argv:
        eng.set_sampling(ln, SamplingParams(temperature=0.8, top_k=40, top_p=0.9, repeat_penalty=1.1, seed=1 + i))
    eng.prefill(ln, rng.integers(0, min(geo.vocab, 128000), 296, dtype=np.int32))
for nb in (1, 2, 3, 4):
    eng.decode(lanes[:nb], 4, read=False)
    eng.sync()
    best = 1e9
    for _ in range(2):
        for ln in lanes[:nb]:
            eng.set_lane(ln, 17, 300)
        eng.decode(lanes[:nb], steps, read=False)
        eng.sync()
        best = min(best, eng.timing().decode_ms / steps)
    gbs = geo.step_bytes(nb, 300 + (steps - 1) / 2) / (best * 1e-3) / 1e9
    print(f"{geo.name} {nb} lane(s): {best:.4f} ms/step {nb * 1000 / best:.1f} tok/s {gbs:.0f} GB/s = {gbs / 6457.1:.3f} of the HBM roofline, "
          f"{eng.timing().kernel_launches} launch(es) per decode call", flush=True)
eng.close()
