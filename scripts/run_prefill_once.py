"""Profiling target: one 296-id prefill of llama3.2-3B (tcgen05 GEMM path, 256-row chunks)."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, geometry, weights
geo = geometry.resolve(sys.argv[1] if len(sys.argv) > 1 else "llama3.2")
eng = Engine(geo, weights.random_init(geo, 0xB200, 0), max_lanes=2, max_ctx=512, max_rows=256, use_megakernel=False)
ids = np.arange(5, 301, dtype=np.int32)
lane = eng.open_lane()
eng.prefill(lane, ids)          # warm (tensor maps, attributes)
eng.sync()
lane2 = eng.open_lane()
torch.cuda.profiler.start()
eng.prefill(lane2, ids)
eng.sync()
torch.cuda.profiler.stop()
print("prefill ms", eng.timing().prefill_ms)
