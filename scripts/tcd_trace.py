"""Debug: phase timeline of the tcgen05 persistent decode kernel (B200LLM_MEGA_TRACE=1).
Needs the trace build of the library: make -C llm-based-apache-spark-optimization_b200/csrc TRACE=1
(libb200llm_trace.so beside the default build, which compiles the stamps out).
Tagged stamps of thread 0 of CTA 0 and of the last CTA; durations are attributed to the tag
that ENDS the interval."""
import ctypes as C
import os
import sys

import numpy as np

os.environ["B200LLM_MEGA_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("B200LLM_LIB", os.path.join(ROOT, "llm-based-apache-spark-optimization_b200", "b200llm", "libb200llm_trace.so"))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]
from b200llm import Engine, _lib, geometry, weights  # noqa: E402

TAGS = {1: "step", 2: "x_conv", 3: "x_ready", 4: "epilogue", 5: "barrier", 6: "att_q", 7: "att_kv", 8: "att_warps", 9: "att_done", 10: "mma_phase", 11: "full", 12: "issued", 13: "seg", 14: "bar_arrive"}
name = sys.argv[1] if len(sys.argv) > 1 else "llama3.2"
ctx = int(sys.argv[2]) if len(sys.argv) > 2 else 296
geo = geometry.resolve(name)
eng = Engine(geo, weights.random_init(geo, 0xB200, 0), max_lanes=1, max_ctx=ctx + 64, max_rows=64, use_megakernel=True)
lane = eng.open_lane()
eng.prefill(lane, np.arange(5, 5 + ctx, dtype=np.int32) % 1000)
eng.decode([lane], 8)
eng.decode([lane], 4)
full = np.zeros(4096, dtype=np.uint64)
_lib.check(_lib.lib.b200llm_debug_read_trace(eng._h, full.ctypes.data_as(C.POINTER(C.c_uint64)), 4096))
out = full[:2048]
for which, arr in (("CTA 0", out[:1024]), ("last CTA", out[1024:1536])):
    arr = arr[arr > 0]
    tag = (arr >> np.uint64(48)).astype(np.int64)
    t = (arr & np.uint64((1 << 48) - 1)).astype(np.int64)
    d = np.diff(t) / 1e3
    print(f"== {which}: {len(arr)} stamps, step {(t[-1] - t[0]) / 1e3:.1f} us")
    tot = {}
    for i in range(1, len(arr)):
        tot.setdefault(TAGS[int(tag[i])], []).append(d[i - 1])
    for k, v in tot.items():
        print(f"  {k:10s} n={len(v):4d} total {sum(v):8.1f} us  mean {np.mean(v):6.2f}  max {max(v):6.2f}")
    if which == "CTA 0":
        # layer 1 in detail
        seq = [(TAGS[int(tag[i])], d[i - 1]) for i in range(1, len(arr))]
        bars = [i for i, (k, _) in enumerate(seq) if k == "barrier"]
        if len(bars) > 10:
            print("  layer 1:", " ".join(f"{k}:{v:.2f}" for k, v in seq[bars[4] + 1: bars[9] + 1]))

arr = out[1536:]
arr = arr[arr > 0]
tag = (arr >> np.uint64(48)).astype(np.int64)
t = (arr & np.uint64((1 << 48) - 1)).astype(np.int64)
# absolute times relative to the CTA-0 compute thread's stamps of the same step
base = int(out[0] & np.uint64((1 << 48) - 1))
print("== MMA issuer of CTA 0, layer 1 (us since step start: tag)")
line = []
for i in range(len(arr)):
    line.append(f"{(t[i] - base) / 1e3:.2f}:{TAGS[int(tag[i])]}")
print(" ".join(line))
c0 = out[:1024]; c0 = c0[c0 > 0]
tg0 = (c0 >> np.uint64(48)).astype(np.int64); t0 = (c0 & np.uint64((1 << 48) - 1)).astype(np.int64)
bars = [i for i in range(len(c0)) if tg0[i] == 5]
print("== compute thread 0 of CTA 0, same window:")
print(" ".join(f"{(t0[i] - base) / 1e3:.2f}:{TAGS[int(tg0[i])]}" for i in range(bars[4], bars[9] + 1)))

print("== arrival of every CTA at the 5 barriers of layer 1 (us after the first arrival)")
for b, nm in enumerate(["qkv", "attn", "o", "gate_up", "down"]):
    a = full[2048 + b * 148: 2048 + (b + 1) * 148].astype(np.int64)
    if (a == 0).any():
        continue
    rel = (a - a.min()) / 1e3
    order = np.argsort(rel)
    print(f"  {nm:8s} median {np.median(rel):5.2f} p90 {np.percentile(rel, 90):5.2f} max {rel.max():5.2f}  "
          f"last CTAs {order[-6:].tolist()}  first CTAs {order[:4].tolist()}  by quarter of grid "
          f"{[round(float(rel[i:i + 37].mean()), 2) for i in range(0, 148, 37)]}")

acc = full[4000:4007].astype(np.int64)
if acc[6] > 0:
    tot = float(acc[6])
    print(f"== CTA 0 over the whole launch ({tot / 1.965e3:.0f} us): MMA thread waits for tile {acc[0] / tot:.1%}, for operand block "
          f"{acc[1] / tot:.1%}, for a free accumulator {acc[3] / tot:.1%}, busy {acc[2] / tot:.1%} ({acc[2] / max(acc[4], 1):.0f} clk per tile, "
          f"{acc[4]} tiles); producer waits for a free slot {acc[5] / tot:.1%}")
