"""Debug: phase timeline of the persistent decode kernel (B200LLM_MEGA_TRACE=1)."""
import ctypes as C, os, sys, numpy as np
os.environ["B200LLM_MEGA_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, _lib, geometry, weights
name = sys.argv[1] if len(sys.argv) > 1 else "llama3.2"
geo = geometry.resolve(name)
eng = Engine(geo, weights.random_init(geo, 0xB200, 0), max_lanes=1, max_ctx=1024, max_rows=64)
lane = eng.open_lane()
eng.prefill(lane, np.arange(5, 301, dtype=np.int32))
eng.decode([lane], 8)
eng.decode([lane], 4)
out = np.zeros(2048, dtype=np.uint64)
_lib.check(_lib.lib.b200llm_debug_read_trace(eng._h, out.ctypes.data_as(C.POINTER(C.c_uint64)), 2048))
names = []
for l in range(2):
    for ph in ("qkv", "attn", "o", "gu", "down"):
        names += ([f"L{l}.{ph}.x", f"L{l}.{ph}.ring"] if ph != "attn" else []) + [f"L{l}.{ph}.done", f"L{l}.{ph}.bar"]
for cta in (0, 1):
    t = out[cta * 1024:(cta + 1) * 1024].astype(np.int64)
    n = int((t > 0).sum())
    print(f"== CTA {'0' if cta == 0 else 'last'}: {n} stamps, step total {(t[n-1]-t[0])/1e3:.1f} us")
    prev = t[0]
    for i, nm in enumerate(names):
        if i + 1 >= n: break
        print(f"  {nm:14s} +{(t[i+1]-prev)/1e3:7.2f} us")
        prev = t[i + 1]
