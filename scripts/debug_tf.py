"""Debug: teacher-forced per-step logit error of the persistent kernel vs the golden oracle vectors."""
import os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, geometry, weights
from oracle import llama_oracle as lo
for mega in (True, False):
    g = np.load(os.path.join(ROOT, "tests/golden/small.npz"))
    geo = lo.GEOMETRIES["small"]
    sd = lo.make_weights(geo, int(g["seed"]), sigma=float(g["sigma"]), embed_sigma=float(g["embed_sigma"]))
    pg = geometry.resolve("small")
    eng = Engine(pg, weights.pack(pg, {k: v.cuda() for k, v in sd.items()}), max_lanes=1, max_ctx=256, max_rows=8, use_megakernel=mega)
    lane = eng.open_lane()
    prompt, toks, probe = g["prompt"], g["tokens"], g["probe_ids"]
    eng.prefill(lane, prompt)
    errs = []
    for s in range(64):
        l = eng.logits(0)
        errs.append(float(np.abs(l[probe] - g["probe_kvbf16"][s]).max() / g["absmax_kvbf16"][s]))
        if s < 63:
            eng.set_lane(lane, int(toks[s]), len(prompt) + s)
            eng.decode([lane], 1)
    print("mega", mega, "max err %.2e" % max(errs), "at step", int(np.argmax(errs)), ["%.1e" % e for e in errs[:12]])
    eng.close()
