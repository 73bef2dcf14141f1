"""Profiling target: one 64-lane decode step on the tcgen05 GEMM path (config 2): python scripts/run_batch_once.py [model] [lanes] [ctx]
Only the step between cudaProfilerStart/Stop is meant to be captured (ncu --profile-from-start off)."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, geometry, weights
geo = geometry.resolve(sys.argv[1] if len(sys.argv) > 1 else "llama3.2:1b")
bs = int(sys.argv[2]) if len(sys.argv) > 2 else 64
ctx = int(sys.argv[3]) if len(sys.argv) > 3 else 296
eng = Engine(geo, weights.random_init(geo, 0xB200, 0), max_lanes=bs, max_ctx=ctx + 64, max_rows=max(64, bs))
rng = np.random.default_rng(5)
lanes = [eng.open_lane() for _ in range(bs)]
for ln in lanes:
    eng.prefill(ln, rng.integers(0, min(geo.vocab, 128000), ctx, dtype=np.int32))
eng.decode(lanes, 4, read=False)     # warm (graph capture, tensor maps)
eng.sync()
torch.cuda.profiler.start()
eng.decode(lanes, 2, read=False)
eng.sync()
torch.cuda.profiler.stop()
ms = eng.timing().decode_ms / 2
gbs = geo.step_bytes(bs, ctx + 5) / (ms * 1e-3) / 1e9
print(f"{geo.name} bs{bs} ctx {ctx}: {ms:.3f} ms/step = {bs * 1000 / ms:.0f} tok/s = {gbs:.0f} GB/s")
