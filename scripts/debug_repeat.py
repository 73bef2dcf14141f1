import sys, os, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, SamplingParams, geometry, weights
geo = geometry.resolve("small")
w = weights.random_init(geo, 0xB200, 0, sigma=0.06)
eng = Engine(geo, w, max_lanes=8, max_ctx=4096, max_rows=64)
ids = np.arange(3, 90, dtype=np.int32)
for pen in (1.0, 1.1, 1.0, 1.1):
    outs = []
    for rep in range(3):
        toks, tm = eng.generate_ids(ids, 12, SamplingParams(temperature=0.0, repeat_penalty=pen), stop_ids=(2,))
        outs.append(toks)
    print(pen, outs[0] == outs[1] == outs[2], outs[0][:6], outs[1][:6], outs[2][:6])
