import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, SamplingParams, geometry, weights
from oracle import sampler
g = np.load(os.path.join(ROOT, "tests/golden/llama3.2.npz"))
geo = geometry.resolve("llama3.2")
eng = Engine(geo, weights.random_init(geo, int(g["seed"]), 0, sigma=float(g["sigma"])), max_lanes=1, max_ctx=len(g["prompt"]) + 128,
             max_rows=64, use_tc=False)
sp = SamplingParams(temperature=0.8, top_k=40, top_p=0.9, repeat_penalty=1.1, seed=20250211)
prompt = g["prompt"]; P = len(prompt)
lane = eng.open_lane(); eng.set_sampling(lane, sp); eng.prefill(lane, prompt)
toks = [eng.read_lane(lane)[0]]
for s in range(3):
    toks.append(int(eng.decode([lane], 1)[0, 0]))
print("state", toks)
outs = {}
for trial in range(40):
    eng.set_lane(lane, toks[3], P + 3)
    t3 = int(eng.decode([lane], 1)[0, 0])   # regenerates toks[3]'s successor? (positions: prefill P ids -> first tok at pos P)
    lg = eng.logits(0).copy()
    outs.setdefault(t3, []).append(lg)
for t, ls in outs.items():
    print("token", t, "x", len(ls))
ks = list(outs)
if len(ks) > 1:
    a, b = outs[ks[0]][0], outs[ks[1]][0]
    d = np.abs(a - b); i = int(d.argmax())
    print("max logit diff between outcome classes", d.max(), "at id", i, a[i], b[i], "ratio", a[i] / b[i])
    big = np.nonzero(d > 1e-3)[0]
    print("ids with |diff| > 1e-3:", big[:20], [(float(a[j]), float(b[j])) for j in big[:5]])
    for lgx in (a, b):
        want, ids, probs = sampler.sample(lgx, sp.temperature, sp.top_k, sp.top_p, 1.0, sp.seed, P + 3)
        print("oracle sampler on these logits ->", want, "keep", len(ids), "margin", sampler.top_p_margin(lgx, 0.8, 40, 0.9))
