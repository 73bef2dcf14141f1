"""Stress: the SAME decode step (same lane state, same input token) many times; the logits may differ from run to run only by
the order of fp32 atomic adds (~1e-6 of the largest logit).  Anything larger is a synchronisation bug (a partial sum read
before it was complete).  Usage: python scripts/stress_logits.py [model] [iters] [ctx]   (B200LLM_TCD_DATAFLOW selects the mode)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, geometry, weights  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "llama3.2"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 400
ctx = int(sys.argv[3]) if len(sys.argv) > 3 else 296
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 4
geo = geometry.resolve(name)
eng = Engine(geo, weights.random_init(geo, 0xB200, 0), max_lanes=1, max_ctx=ctx + 64, max_rows=64, use_megakernel=True)
lane = eng.open_lane()
prompt = np.random.default_rng(1).integers(0, min(geo.vocab, 128000), ctx, dtype=np.int32)
eng.prefill(lane, prompt)
ref, worst, bad = None, 0.0, 0
for i in range(iters):
    eng.set_lane(lane, int(prompt[3]), ctx)
    eng.decode([lane], steps)                 # `steps` steps in one launch; the logits of the last one are compared
    lg = eng.logits(0)
    if ref is None:
        ref = lg.copy()
        continue
    d = float(np.abs(lg - ref).max() / np.abs(ref).max())
    worst = max(worst, d)
    if d > 1e-4:
        bad += 1
        if bad <= 5:
            print(f"iteration {i}: logits differ from the first run by {d:.2e} of the largest logit")
print(f"{name} dataflow={os.environ.get('B200LLM_TCD_DATAFLOW', 'default')}: {iters} x decode({steps}): worst deviation {worst:.2e}, "
      f"{bad} runs beyond 1e-4")
