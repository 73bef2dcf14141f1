import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, SamplingParams, geometry, weights
g = np.load(os.path.join(ROOT, "tests/golden/llama3.2.npz"))
geo = geometry.resolve("llama3.2")
eng = Engine(geo, weights.random_init(geo, int(g["seed"]), 0, sigma=float(g["sigma"])), max_lanes=1, max_ctx=len(g["prompt"]) + 128,
             max_rows=64, use_tc=False)
sp = SamplingParams(temperature=0.8, top_k=40, top_p=0.9, repeat_penalty=1.1, seed=20250211)
prompt = g["prompt"]
def single():
    lane = eng.open_lane(); eng.set_sampling(lane, sp); eng.prefill(lane, prompt)
    toks = [eng.read_lane(lane)[0]] + [int(eng.decode([lane], 1)[0, 0]) for _ in range(32)]
    eng.close_lane(lane); return toks
def chunked(n):
    lane = eng.open_lane(); eng.set_sampling(lane, sp); eng.prefill(lane, prompt)
    toks = [eng.read_lane(lane)[0]]
    while len(toks) < 33:
        k = min(n, 33 - len(toks)); toks += eng.decode([lane], k)[:, 0].tolist()
    eng.close_lane(lane); return toks
a = single()
print("single   ", a[:12])
for n in (1, 2, 4, 8, 32, 32, 32):
    b = chunked(n)
    first = next((i for i, (x, y) in enumerate(zip(a, b)) if x != y), None)
    print(f"chunk {n:2d}: first diff {first}", b[:12])
