"""oracle/weightgen.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Seeded random-init weights and synthetic prompt ids for the oracle
(SURVEY.md section 8(d) "Weights"; BASELINE.json north_star "same random-init
weights and prompts").  The stream is specified in oracle/weightgen.c; this
module calls that C build when present and otherwise runs the same integer
arithmetic in numpy (bit-identical, slower).

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_GOLD = np.uint64(0x9E3779B97F4A7C15)
IRWIN_HALL_STD = 37837.22765                # std of the sum of four uniform u16 lanes

# tensor kinds inside one decoder layer; tensor stream id = (layer + 1) * 16 + kind
KIND = {"q_proj": 0, "k_proj": 1, "v_proj": 2, "o_proj": 3, "gate_proj": 4, "up_proj": 5,
        "down_proj": 6, "input_layernorm": 7, "post_attention_layernorm": 8}
GLOBAL_TID = {"embed_tokens": 0, "norm": 1, "lm_head": 2}


def tensor_seed(base_seed: int, tid: int) -> int:
    return ((int(base_seed) << 20) + int(tid)) & 0xFFFFFFFFFFFFFFFF


def _load():
    p = os.path.join(_HERE, "liboracle_weightgen.so")
    if not os.path.exists(p):
        return None
    lib = ctypes.CDLL(p)
    lib.oracle_weightgen_bf16.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64,
                                          ctypes.c_uint64, ctypes.c_float, ctypes.c_float]
    lib.oracle_weightgen_bf16.restype = None
    lib.oracle_tokengen_i32.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64,
                                        ctypes.c_uint32]
    lib.oracle_tokengen_i32.restype = None
    return lib


_LIB = _load()


def _mix64(z: np.ndarray) -> np.ndarray:
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def _gen_numpy(n: int, first: int, seed: int, scale: np.float32, mean: np.float32) -> np.ndarray:
    out = np.empty(n, dtype=np.uint16)
    step = 1 << 22
    with np.errstate(over="ignore"):
        for s in range(0, n, step):
            e = min(n, s + step)
            idx = np.arange(s + first + 1, e + first + 1, dtype=np.uint64)
            z = _mix64(np.uint64(seed) + idx * _GOLD)
            m = np.uint64(0xFFFF)
            ssum = ((z & m) + ((z >> np.uint64(16)) & m) + ((z >> np.uint64(32)) & m)
                    + (z >> np.uint64(48))).astype(np.int64) - 131070
            v = ssum.astype(np.float32) * scale
            w = (mean + v).astype(np.float32)
            u = w.view(np.uint32)
            u = u + np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1))
            out[s:e] = (u >> np.uint32(16)).astype(np.uint16)
    return out


def gen_bf16(shape, seed: int, sigma: float, mean: float = 0.0, force_numpy: bool = False) -> torch.Tensor:
    """bf16 tensor of `shape` from stream `seed` (row-major element index)."""
    n = int(np.prod(shape))
    scale = np.float32(np.float32(sigma) / np.float32(IRWIN_HALL_STD))
    if _LIB is not None and not force_numpy:
        buf = np.empty(n, dtype=np.uint16)
        _LIB.oracle_weightgen_bf16(buf.ctypes.data, n, 0, seed, float(scale), float(np.float32(mean)))
    else:
        buf = _gen_numpy(n, 0, seed, scale, np.float32(mean))
    return torch.from_numpy(buf.view(np.int16)).view(torch.bfloat16).reshape(tuple(shape))


def gen_tokens(n: int, seed: int, modulo: int) -> np.ndarray:
    """Synthetic prompt ids, uniform over [0, modulo) (SURVEY.md 8(d) Config 1)."""
    if _LIB is not None:
        out = np.empty(n, dtype=np.int32)
        _LIB.oracle_tokengen_i32(out.ctypes.data, n, seed, modulo)
        return out
    with np.errstate(over="ignore"):
        idx = np.arange(1, n + 1, dtype=np.uint64)
        z = _mix64(np.uint64(seed) + idx * _GOLD)
        return ((z >> np.uint64(11)) % np.uint64(modulo)).astype(np.int32)
