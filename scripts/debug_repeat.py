"""Debug: repeated generate calls on one engine must give identical greedy tokens (and match the oracle)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, SamplingParams, geometry, weights  # noqa: E402
from oracle import llama_oracle as lo  # noqa: E402

geo = geometry.resolve("small")
ogeo = lo.GEOMETRIES["small"]
sd = lo.make_weights(ogeo, 0xB200, sigma=0.06)
orc = lo.LlamaOracle(ogeo, sd, kv_dtype=torch.bfloat16)
ids = np.arange(3, 90, dtype=np.int32)
want, lg, _ = orc.greedy(ids, 12)
print("oracle", want[:6])
for max_rows, mega in ((64, True), (64, False), (16, True)):
    eng = Engine(geo, weights.pack(geo, {k: v.cuda() for k, v in sd.items()}), max_lanes=8, max_ctx=4096,
                 max_rows=max_rows, use_megakernel=mega)
    for rep in range(3):
        lane = eng.open_lane()
        eng.prefill(lane, ids)
        first, pos = eng.read_lane(lane)
        l0 = eng.logits(0)
        rel = float(np.abs(l0 - lg[0].numpy()).max() / np.abs(lg[0].numpy()).max())
        out = eng.decode([lane], 11)[:, 0].tolist()
        print(f"max_rows={max_rows} mega={mega} rep={rep}: first={first} pos={pos} prefill-logit rel={rel:.2e} "
              f"tokens_ok={[first] + out == want} {([first] + out)[:6]}")
        eng.close_lane(lane)
    eng.close()
