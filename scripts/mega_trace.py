"""Debug: phase timeline of the persistent decode kernel (B200LLM_MEGA_TRACE=1).
Columns: CTA 0 (owns the most rows of every matrix) and the last CTA (the fewest)."""
import ctypes as C
import os
import sys

import numpy as np

os.environ["B200LLM_MEGA_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, _lib, geometry, weights  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "llama3.2"
ctx = int(sys.argv[2]) if len(sys.argv) > 2 else 296
geo = geometry.resolve(name)
eng = Engine(geo, weights.random_init(geo, 0xB200, 0), max_lanes=1, max_ctx=ctx + 64, max_rows=64, use_megakernel=True)
lane = eng.open_lane()
eng.prefill(lane, np.arange(5, 5 + ctx, dtype=np.int32) % 1000)
eng.decode([lane], 8)
eng.decode([lane], 4)
out = np.zeros(2048, dtype=np.uint64)
_lib.check(_lib.lib.b200llm_debug_read_trace(eng._h, out.ctypes.data_as(C.POINTER(C.c_uint64)), 2048))
names = []
for l in range(2):
    for ph in ("qkv", "attn", "o", "gu", "down"):
        names += ([f"L{l}.{ph}.xld", f"L{l}.{ph}.xtab", f"L{l}.{ph}.x", f"L{l}.{ph}.ring"] if ph != "attn" else []) + ([f"L{l}.attn.q", f"L{l}.attn.kv", f"L{l}.attn.warps"] if ph == "attn" else []) + [f"L{l}.{ph}.done", f"L{l}.{ph}.bar"]
t0 = out[:1024].astype(np.int64)
t1 = out[1024:].astype(np.int64)
n = int((t0 > 0).sum())
print(f"{name} ctx {ctx}: {n} stamps; step total {(t0[n-1]-t0[0])/1e3:.1f} us (CTA 0), {(t1[n-1]-t1[0])/1e3:.1f} us (last CTA)")
print(f"{'phase':14s} {'CTA0 us':>9s} {'last us':>9s}")
sums = {}
for i, nm in enumerate(names):
    if i + 1 >= n:
        break
    d0, d1 = (t0[i + 1] - t0[i]) / 1e3, (t1[i + 1] - t1[i]) / 1e3
    print(f"{nm:14s} {d0:9.2f} {d1:9.2f}")
    if nm.startswith("L1."):
        key = nm.split(".")[2]
        sums[key] = sums.get(key, 0.0) + d0
print("layer-1 totals on CTA 0 by kind:", {k: round(v, 2) for k, v in sums.items()}, "sum", round(sum(sums.values()), 2))
fine = out[512:1024].astype(np.int64).reshape(-1, 2)
fine = fine[(fine[:, 0] > 0) & (fine[:, 1] > 0)]
if len(fine) > 2:
    wait = (fine[1:, 0] - fine[:-1, 1]) / 1e3      # arrive(i-1) -> data of stage i ready
    comp = (fine[:, 1] - fine[:, 0]) / 1e3         # data ready -> stage released
    print(f"L1.gu stages on CTA0/warp0: n={len(fine)} wait mean {wait.mean():.3f} us (max {wait.max():.3f}), "
          f"compute mean {comp.mean():.3f} us (max {comp.max():.3f}); first 12 waits {np.round(wait[:12], 2)}")
