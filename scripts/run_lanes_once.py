"""Profiling target: one persistent-kernel decode call (3 steps) of N lanes of llama3.2-3B at ctx 296: python scripts/run_lanes_once.py [model] [lanes]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, geometry, weights
geo = geometry.resolve(sys.argv[1] if len(sys.argv) > 1 else "llama3.2")
nb = int(sys.argv[2]) if len(sys.argv) > 2 else 4
eng = Engine(geo, weights.random_init(geo, 0xB200, 0), max_lanes=nb, max_ctx=512, max_rows=256, use_megakernel=True)
lanes = [eng.open_lane() for _ in range(nb)]
for i, ln in enumerate(lanes):
    eng.prefill(ln, np.arange(5 + i, 301 + i, dtype=np.int32))
print(eng.decode(lanes, 3))
