"""Profiling target: one short persistent-kernel decode call (3 steps) of llama3.2-3B at ctx 296."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, geometry, weights
geo = geometry.resolve(sys.argv[1] if len(sys.argv) > 1 else "llama3.2")
eng = Engine(geo, weights.random_init(geo, 0xB200, 0), max_lanes=1, max_ctx=512, max_rows=64, use_megakernel=True)
lane = eng.open_lane()
eng.prefill(lane, np.arange(5, 301, dtype=np.int32))
print(eng.decode([lane], 3)[:, 0])
