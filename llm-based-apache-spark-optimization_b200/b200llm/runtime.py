"""Model manager behind the `ollama` shim: resolves a model name to a loaded engine,
loading it on first use and keeping it resident (what Ollama's scheduler + keep_alive do
for the reference).  Weights are seeded random-init: the reference's GGUF
blobs are not on this box (BASELINE.json: "random-init weights").
"""
from __future__ import annotations

import os
import threading
import time
from dataclasses import dataclass
from typing import Dict

import torch

from . import _lib, geometry, weights
from .engine import Engine
from .router import Router
from .scheduler import Scheduler
from .tokenizer import TEMPLATES, ByteTokenizer

SEED = 0xB200


@dataclass
class LoadedModel:
    name: str
    engine: Engine
    tokenizer: object         # ByteTokenizer | SpmTokenizer | BpeTokenizer (same encode / decode / stop_ids)
    load_ns: int
    scheduler: object = None  # Scheduler, or a Router over one Scheduler per GPU (B200LLM_DEVICES)
    template: str = ""        # Modelfile TEMPLATE; empty for the seeded random-init models (byte tokenizer)
    engines: tuple = ()       # every replica's engine (engine = engines[0])


_models: Dict[str, LoadedModel] = {}
_checkpoints: Dict[str, str] = {}      # model name -> GGUF path (gguf_loader.register)
_lock = threading.Lock()


def register_checkpoint(name: str, path: str) -> None:
    """Serve `name` from a GGUF checkpoint instead of seeded random-init weights (loaded on first use)."""
    if not os.path.isfile(path):
        raise FileNotFoundError(path)
    with _lock:
        _checkpoints[name] = path


def get_model(name: str) -> LoadedModel:
    ckpt = _checkpoints.get(name)
    key = name if ckpt else geometry.resolve(name).name   # KeyError -> the shim turns it into ResponseError 404
    with _lock:
        m = _models.get(key)
        if m is not None:
            return LoadedModel(m.name, m.engine, m.tokenizer, 0, m.scheduler, m.template, m.engines)
        if not torch.cuda.is_available():
            raise _lib.B200LLMError(_lib.ENODEV, "no CUDA device: b200llm has no CPU fallback")
        t0 = time.perf_counter_ns()
        # B200LLM_DEVICES=0,1,..: one replica (engine + scheduler thread) per listed GPU behind a request router
        # (b200llm/router.py; B200LLM_ROUTING = least_loaded | round_robin); else the one device B200LLM_DEVICE names
        devs = [int(d) for d in os.environ.get("B200LLM_DEVICES", "").split(",") if d.strip()] or \
               [int(os.environ.get("B200LLM_DEVICE", "0"))]
        engines, tok, template = [], None, ""
        for dev in devs:
            torch.cuda.set_device(dev)
            if ckpt:
                from . import gguf_loader
                geo, sd = gguf_loader.load(ckpt, dev, name)
                w = weights.pack(geo, sd)
                del sd
                tok = tok or gguf_loader.load_tokenizer(ckpt)
            else:
                geo = geometry.resolve(name)
                w = weights.random_init(geo, SEED, dev)
            engines.append(Engine(geo, w, device=dev, max_lanes=int(os.environ.get("B200LLM_MAX_LANES", "8")),
                                  max_ctx=int(os.environ.get("B200LLM_NUM_CTX", "4096")),
                                  max_rows=int(os.environ.get("B200LLM_MAX_ROWS", "256")),   # prefill chunk = tcgen05 GEMM N limit
                                  # B200LLM_LEAN_WEIGHTS=1: tile-major matrices only (half the weight bytes; no GEMV fallback)
                                  lean_weights=os.environ.get("B200LLM_LEAN_WEIGHTS", "0") not in ("", "0")))
        # a checkpoint with tokenizer tables is prompted through its model's template (by the registered name, else
        # by the base name before ':'); without tables text goes through the byte stand-in, untemplated
        template = (TEMPLATES.get(name) or TEMPLATES.get(name.split(":")[0], "")) if tok is not None else ""
        scheds = [Scheduler(e) for e in engines]
        sched = scheds[0] if len(scheds) == 1 else Router(scheds, os.environ.get("B200LLM_ROUTING", "least_loaded"))
        m = LoadedModel(key, engines[0], tok or ByteTokenizer(geo.vocab), time.perf_counter_ns() - t0, sched, template,
                        tuple(engines))
        _models[key] = m
        return m


def unload_all() -> None:
    with _lock:
        for m in _models.values():
            m.scheduler.shutdown()
            for e in (m.engines or (m.engine,)):
                e.close()
        _models.clear()
