"""Stress: the 3B golden-parity check repeated on fresh engines (looks for rare wrong results)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200"), os.path.join(ROOT, "tests")]
from b200llm import Engine, geometry, weights  # noqa: E402

fname, name = (sys.argv[1], sys.argv[2]) if len(sys.argv) > 2 else ("llama3.2", "llama3.2")
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 20
g = np.load(os.path.join(ROOT, "tests", "golden", fname + ".npz"))
geo = geometry.resolve(name)
prompt, toks, probe = g["prompt"], g["tokens"].tolist(), g["probe_ids"]
w = weights.random_init(geo, int(g["seed"]), 0, sigma=float(g["sigma"]))
bad = 0
for it in range(iters):
    eng = Engine(geo, w, max_lanes=1, max_ctx=len(prompt) + 128, max_rows=256, use_tc=(os.environ.get("B200LLM_STRESS_TC") == "1"), use_megakernel=True)
    lane = eng.open_lane()
    eng.prefill(lane, prompt)
    first, _ = eng.read_lane(lane)
    l0 = eng.logits(0)
    e0 = float(np.abs(l0[probe] - g["probe_kvbf16"][0]).max() / g["absmax_kvbf16"][0])
    got = [first] + eng.decode([lane], 63)[:, 0].tolist()
    l63 = eng.logits(0)
    e63 = float(np.abs(l63[probe] - g["probe_kvbf16"][63]).max() / g["absmax_kvbf16"][63])
    mism = [s for s in range(64) if got[s] != toks[s]]
    flag = "" if (not mism and e0 < 3e-3 and e63 < 3e-3) else "  <-- BAD"
    bad += bool(flag)
    print(f"iter {it}: prefill err {e0:.2e}, step-63 err {e63:.2e}, token mismatches at {mism[:6]}{flag}", flush=True)
    eng.close()
print(f"{bad} bad of {iters}")
