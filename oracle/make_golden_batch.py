"""oracle/make_golden_batch.py -- TEST INFRASTRUCTURE: writes tests/golden/llama3.2_1b_bs64.npz.

BASELINE.json config 2 (Llama-3.2-1B, batch 64): 64 prompts of different lengths (24..87 ids, seeded), the first 16
greedy tokens of every one and the fp32 top1-top2 margin of every step, from oracle/llama_oracle.py in the engine's
storage model (fp32 activations, bf16 KV cache) -- the same oracle, weights stream and prompt generator as
oracle/make_golden.py.  A prompt seed is kept when its 16 steps hold a margin >= 5e-3 (up to 8 seeds are tried), so that
a batched engine that is ~1e-3 from the oracle must reproduce every token.
Run here (CPU container, ~10 min):   python -m oracle.make_golden_batch
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np
import torch

from . import llama_oracle as lo
from . import weightgen as wg
from .make_golden import OUT, SEED

N_LANES, N_NEW, MIN_MARGIN = 64, 16, 5e-3


def main(name="llama3.2:1b"):
    geo = lo.GEOMETRIES[name]
    sd = lo.make_weights(geo, SEED, sigma=0.02, embed_sigma=0.02)
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    id_mod = min(geo.vocab, 128000)
    t0 = time.time()
    lens, seeds, toks, mars, prompts = [], [], [], [], []
    for lane in range(N_LANES):
        plen = 24 + lane
        best = None
        for k in range(8):
            pseed = 5000 + 16 * lane + k
            prompt = wg.gen_tokens(plen, pseed, id_mod)
            t, _, m = orc.greedy(prompt, N_NEW)
            if best is None or min(m) > min(best[2]):
                best = (pseed, t, m, prompt)
            if min(m) >= MIN_MARGIN:
                break
        print(f"lane {lane}: {plen} ids, prompt seed {best[0]}, min margin {min(best[2]):.4f} ({time.time() - t0:.0f}s)", flush=True)
        lens.append(plen), seeds.append(best[0]), toks.append(best[1]), mars.append(best[2])
        p = np.zeros(24 + N_LANES, np.int32)
        p[:plen] = best[3]
        prompts.append(p)
    np.savez_compressed(os.path.join(OUT, name.replace(":", "_") + "_bs64.npz"), seed=SEED, sigma=0.02, embed_sigma=0.02,
                        prompt_len=np.asarray(lens, np.int32), prompt_seed=np.asarray(seeds, np.int32),
                        prompts=np.stack(prompts), tokens=np.asarray(toks, np.int32), margins=np.asarray(mars, np.float32))
    print(f"wrote golden ({time.time() - t0:.0f}s); worst margin {np.asarray(mars).min():.4f}")


if __name__ == "__main__":
    main(*sys.argv[1:])
