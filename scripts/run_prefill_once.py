"""Profiling target: one prefill (tcgen05 GEMM path, 256-row chunks): python scripts/run_prefill_once.py [model] [n ids]
(B200LLM_NO_PREFILL_ATTN=1: the decode-style attention per query row instead of attn_prefill.cu)."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, geometry, weights
geo = geometry.resolve(sys.argv[1] if len(sys.argv) > 1 else "llama3.2")
n = int(sys.argv[2]) if len(sys.argv) > 2 else 296
eng = Engine(geo, weights.random_init(geo, 0xB200, 0), max_lanes=2, max_ctx=n + 64, max_rows=256, use_megakernel=bool(os.environ.get("B200LLM_PREFILL_TM")))   # B200LLM_PREFILL_TM=1: with the tile-major copies (gemm_tc then takes its weight tiles from them)
ids = (np.arange(5, 5 + n, dtype=np.int32) * 7919) % min(geo.vocab, 128000)
lane = eng.open_lane()
eng.prefill(lane, ids)          # warm (tensor maps, attributes)
eng.sync()
lane2 = eng.open_lane()
torch.cuda.profiler.start()
eng.prefill(lane2, ids)
eng.sync()
torch.cuda.profiler.stop()
import bench
fl = bench.prefill_flops(geo, n)
ms = eng.timing().prefill_ms
print(f"{geo.name} prefill of {n} ids: {ms:.3f} ms = {fl / ms / 1e9:.1f} TFLOP/s = {fl / ms / 1e9 / bench.sustained_tflops():.3f} of sustained bf16"
      f" (attention: {'decode-style per row' if os.environ.get('B200LLM_NO_PREFILL_ATTN') else 'attn_prefill.cu'})")
