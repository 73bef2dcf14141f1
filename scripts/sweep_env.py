#!/usr/bin/env python
"""Decode tok/s of one model under a list of environment settings the engine reads per decode call
(B200LLM_TCD_INFLIGHT, B200LLM_TCD_STAGES, ...), one engine, same prompt and context for every setting.

  python scripts/sweep_env.py llama3.2 B200LLM_TCD_INFLIGHT=0,3,4,5,6 [B200LLM_TCD_STAGES=9,8] [--steps 128] [--sampled]
"""
import itertools
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]

import numpy as np
import torch
from b200llm import Engine, SamplingParams, geometry, weights


def main():
    argv = sys.argv[1:]
    if "--steps" in argv:        # (its value is not a sweep)
        i = argv.index("--steps")
        argv = argv[:i] + argv[i + 2:]
    args = [a for a in argv if not a.startswith("--")]
    model, sweeps = args[0], [a.split("=") for a in args[1:]]
    steps = int(sys.argv[sys.argv.index("--steps") + 1]) if "--steps" in sys.argv else 128
    geo = geometry.resolve(model)
    eng = Engine(geo, weights.random_init(geo, 0xB200, 0), max_lanes=2, max_ctx=296 + steps + 64, max_rows=256)
    prompt = np.random.default_rng(1).integers(0, min(geo.vocab, 128000), 296, dtype=np.int32)
    lane = eng.open_lane()
    if "--sampled" in sys.argv:
        eng.set_sampling(lane, SamplingParams(temperature=0.8, top_k=40, top_p=0.9, repeat_penalty=1.1, seed=1))
    eng.prefill(lane, prompt)
    eng.decode([lane], 8)
    names = [k for k, _ in sweeps]
    for combo in itertools.product(*[v.split(",") for _, v in sweeps]):
        for k, v in zip(names, combo):
            if v == "-":
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
        best = 1e9
        for _ in range(3):
            eng.set_lane(lane, int(prompt[0]), 296)
            eng.decode([lane], steps, read=False)
            eng.sync()
            best = min(best, eng.timing().decode_ms / steps)
        mean_ctx = 296 + (steps - 1) / 2
        gbs = geo.step_bytes(1, mean_ctx) / (best * 1e-3) / 1e9
        print(" ".join(f"{k}={v}" for k, v in zip(names, combo)), f"{best:.4f} ms/step {1000 / best:.1f} tok/s {gbs:.0f} GB/s "
              f"{gbs / 6457.1:.3f}", flush=True)
    eng.close()


if __name__ == "__main__":
    main()
