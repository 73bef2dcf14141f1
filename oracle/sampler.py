"""oracle/sampler.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement of the sampler chain the decode path ends in.
The reference has no sampler code of its own (it is Ollama/llama.cpp's default chain,
un-vendored: repeat-penalty -> top-k -> top-p -> temperature -> draw), so
the chain is restated here in the order and arithmetic the engine documents
(DESIGN.md "Sampler"); parity unpinned by the reference.  numpy fp32, sequential sums.
"""
from __future__ import annotations

import numpy as np

_GOLD = 0x9E3779B97F4A7C15
_M = (1 << 64) - 1


def _mix64(z: int) -> int:
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M
    return z ^ (z >> 31)


def uniform(seed: int, ctr: int) -> np.float32:
    z = _mix64((seed + (ctr + 1) * _GOLD) & _M)
    return np.float32(z >> 40) * np.float32(1.0 / 16777216.0)


def apply_penalty(logits: np.ndarray, repeat_penalty: float, history=None) -> np.ndarray:
    """fp32 copy of the logits with each distinct recent token penalised once: logit > 0 ? logit / p : logit * p."""
    lg = logits.astype(np.float32).copy()
    V = lg.shape[0]
    if repeat_penalty != 1.0 and history is not None:
        seen = set()
        for t in history:
            t = int(t)
            if t < 0 or t >= V or t in seen:
                continue
            seen.add(t)
            pen = np.float32(repeat_penalty)
            lg[t] = lg[t] / pen if lg[t] > 0 else lg[t] * pen
    return lg


def sample(logits: np.ndarray, temperature: float, top_k: int, top_p: float, repeat_penalty: float,
           seed: int, ctr: int, history=None, cap: int = 1024):
    """Returns (token, candidate ids kept after top-k/top-p, their unnormalised probs)."""
    lg = apply_penalty(logits, repeat_penalty, history)
    V = lg.shape[0]
    if temperature <= 0:
        return int(np.argmax(lg)), None, None           # first maximal index = lowest id
    k = min(top_k if top_k > 0 else cap, cap, V)
    order = np.lexsort((np.arange(V), -lg))[:k]          # value descending, id ascending
    vals = lg[order]
    inv_t = np.float32(1.0) / np.float32(temperature)
    probs = np.exp(((vals - vals[0]) * inv_t).astype(np.float32)).astype(np.float32)
    total = np.float32(0)
    for p in probs:
        total = np.float32(total + p)
    keep = k
    if top_p < 1.0:
        cum = np.float32(0)
        for i, p in enumerate(probs):
            cum = np.float32(cum + p)
            if np.float32(cum / total) >= np.float32(top_p):
                keep = i + 1
                break
    mass = np.float32(0)
    for p in probs[:keep]:
        mass = np.float32(mass + p)
    r = np.float32(uniform(seed, ctr) * mass)
    cum = np.float32(0)
    pick = keep - 1
    for i in range(keep):
        cum = np.float32(cum + probs[i])
        if cum > r:
            pick = i
            break
    return int(order[pick]), order[:keep], probs[:keep]


def top_p_margin(logits: np.ndarray, temperature: float, top_k: int, top_p: float, cap: int = 1024) -> float:
    """How close the top-p cut is to keeping one candidate more or fewer: min |cum_i / total - top_p| over the top-k
    candidates (1.0 when top-p is off).  The kept SET, and with it the mass the uniform draw is scaled by, changes
    discontinuously there, so logits that differ by rounding noise may legitimately lead to different draws."""
    if temperature <= 0 or top_p >= 1.0:
        return 1.0
    lg = logits.astype(np.float32)
    k = min(top_k if top_k > 0 else cap, cap, lg.shape[0])
    vals = np.sort(lg)[::-1][:k].astype(np.float64)
    probs = np.exp((vals - vals[0]) / float(temperature))
    return float(np.abs(np.cumsum(probs) / probs.sum() - top_p).min())
