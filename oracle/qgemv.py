"""oracle/qgemv.py -- TEST / BASELINE INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes face of oracle/qgemv.c: the restatement of llama.cpp's quantised CPU math (Q4_K x Q8_K for llama3.2's Q4_K_M
files, Q4_0 x Q8_0 for duckdb-nsql's), the CPU path the reference really runs behind `ollama.generate`
(/root/reference/Model_Evaluation_&_Comparision.py:113-119 times it).  tests/test_qgemv.py checks the block arithmetic
against gguf-py's de-quantiser; bench.py's cpu legs time `decode_tok_s`.

Only tests/ and bench.py's CPU legs may import this.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
FORMATS = {"q4_K": 0, "q4_0": 1}
# the quantisation the reference's two models ship in (Ollama library defaults)
MODEL_FORMAT = {"llama3.2": "q4_K", "llama3.2:1b": "q4_K", "duckdb-nsql": "q4_0", "tiny": "q4_K", "small": "q4_K"}


def _load():
    p = os.path.join(_HERE, "liboracle_qgemv.so")
    if not os.path.exists(p):
        return None
    # libgomp's default of spinning at barriers collapses when anything else wants the cores (measured here: 8 threads 27 ms
    # per GEMV spinning, 1.1 ms passive); read when libgomp initialises, i.e. at this load
    os.environ.setdefault("OMP_WAIT_POLICY", "PASSIVE")
    lib = C.CDLL(p)
    for name in ("qgemv_sizeof_q4_K", "qgemv_sizeof_q8_K", "qgemv_sizeof_q4_0", "qgemv_sizeof_q8_0", "qgemv_has_vnni"):
        getattr(lib, name).restype = C.c_int
    lib.qgemv_force_scalar.argtypes = [C.c_int]
    lib.qgemv_set_threads.argtypes = [C.c_int]
    for name in ("qgemv_quantize_q8_K", "qgemv_quantize_q8_0"):
        getattr(lib, name).argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    for name in ("qgemv_q4_K", "qgemv_q4_0"):
        getattr(lib, name).argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
    lib.qgemv_decode_bench.argtypes = [C.c_int] * 10 + [C.POINTER(C.c_double), C.POINTER(C.c_int)]
    lib.qgemv_decode_bench.restype = C.c_double
    return lib


LIB = _load()


def available() -> bool:
    return LIB is not None


def simd() -> str:
    return "avx512-vnni" if LIB.qgemv_has_vnni() else "auto-vectorised (avx2 or baseline)"


def quantize_activations(fmt: str, x: np.ndarray) -> np.ndarray:
    """fp32 [k] -> raw Q8_K / Q8_0 blocks (uint8)."""
    x = np.ascontiguousarray(x, np.float32)
    per, size = (256, LIB.qgemv_sizeof_q8_K()) if fmt == "q4_K" else (32, LIB.qgemv_sizeof_q8_0())
    out = np.zeros(x.size // per * size, np.uint8)
    (LIB.qgemv_quantize_q8_K if fmt == "q4_K" else LIB.qgemv_quantize_q8_0)(x.ctypes.data, out.ctypes.data, x.size)
    return out


def gemv(fmt: str, w_blocks: np.ndarray, xq: np.ndarray, n_rows: int, k: int, scalar: bool = False) -> np.ndarray:
    """y = W x on raw blocks: w_blocks uint8 [n_rows * k / block * block_bytes], xq from quantize_activations."""
    y = np.empty(n_rows, np.float32)
    LIB.qgemv_force_scalar(1 if scalar else 0)
    try:
        (LIB.qgemv_q4_K if fmt == "q4_K" else LIB.qgemv_q4_0)(w_blocks.ctypes.data, xq.ctypes.data, y.ctypes.data, n_rows, k)
    finally:
        LIB.qgemv_force_scalar(0)
    return y


def decode_tok_s(geo, fmt: str, n_steps: int, n_warm: int = 2, threads: int = 0):
    """Decode steps per second of the geometry's weight matrices as quantised GEMVs.  threads = 0: the thread count is
    tuned the way a llama.cpp user tunes -t -- one step each at all / half / a quarter of the host threads the process
    may use, the fastest is kept (with hyper-threads or a CPU quota "all" is often the slowest).
    Returns (tok_s, seconds, threads, quantised weight bytes per step)."""
    import os

    def run(n, steps, warm):
        LIB.qgemv_set_threads(n)
        nbytes, used = C.c_double(0), C.c_int(0)
        secs = LIB.qgemv_decode_bench(FORMATS[fmt], geo.d, geo.n_layers, geo.n_q, geo.n_kv, geo.head_dim, geo.ffn, geo.vocab,
                                      steps, warm, C.byref(nbytes), C.byref(used))
        if secs <= 0:
            raise RuntimeError(f"qgemv_decode_bench failed ({secs})")
        return secs, used.value, nbytes.value

    if threads <= 0:
        try:
            n_all = max(1, len(os.sched_getaffinity(0)))
        except (AttributeError, OSError):
            n_all = max(1, os.cpu_count() or 1)
        best = None
        for n in sorted({n_all, max(1, n_all // 2), max(1, n_all // 4)}, reverse=True):
            secs, _, _ = run(n, 2, 1)
            if best is None or secs < best[0]:
                best = (secs, n)
        threads = best[1]
    secs, used, nbytes = run(threads, n_steps, n_warm)
    return n_steps / secs, secs, used, nbytes
