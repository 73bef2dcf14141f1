"""Drop-in for the `ollama` Python client on the one call the reference makes:

    res = ollama.generate(model="duckdb-nsql", system=..., prompt=...);  res.response
        Flask/app.py:102-107, FastAPI/app.py:85-90
    res = ollama.generate(model="llama3.2", system=..., prompt=...);     res.response
        Flask/app.py:160-166, FastAPI/app.py:105-111
    Model_Evaluation_&_Comparision.py:23-45 and :114-120 (same call, three models)

Put the directory that holds this package on sys.path ahead of site-packages and the
reference's app.py files run unmodified: the request goes to the in-process B200 engine
(b200llm) instead of HTTP -> Ollama -> llama.cpp.  Signature, return object and error
type follow the upstream client (>= 0.4: attribute access and subscripting both work);
arguments the engine does not support are accepted and ignored, as upstream does for
unknown options.
"""
from __future__ import annotations

import datetime
import os
import secrets
import time
from typing import Any, Dict, Optional, Sequence

_list = list          # the module exports its own `list()` (upstream name) below

__all__ = ["generate", "ResponseError", "GenerateResponse", "Client", "list"]

# Ollama's documented defaults: the reference passes no `options`, so these apply.
# No default seed: like Ollama, a call without options["seed"] draws a fresh one, so repeated identical prompts differ.
DEFAULT_OPTIONS = {"temperature": 0.8, "top_k": 40, "top_p": 0.9, "repeat_penalty": 1.1, "seed": None,
                   "num_predict": int(os.environ.get("B200LLM_NUM_PREDICT", "128")), "num_ctx": 4096, "stop": []}


class ResponseError(Exception):
    """Same shape as ollama.ResponseError: .error (str) and .status_code (int)."""

    def __init__(self, error: str, status_code: int = -1):
        super().__init__(error)
        self.error = error
        self.status_code = status_code

    def __str__(self) -> str:
        return f"{self.error} (status code: {self.status_code})"


class GenerateResponse:
    """Attribute- and key-addressable result, fields as in Ollama's /api/generate reply
    (durations in nanoseconds).  The reference reads only `.response`."""
    _FIELDS = ("model", "created_at", "response", "done", "done_reason", "context", "total_duration",
               "load_duration", "prompt_eval_count", "prompt_eval_duration", "eval_count", "eval_duration")

    def __init__(self, **kw: Any):
        for f in self._FIELDS:
            setattr(self, f, kw.get(f))

    def __getitem__(self, key: str) -> Any:
        if key not in self._FIELDS:
            raise KeyError(key)
        return getattr(self, key)

    def __contains__(self, key: str) -> bool:
        return key in self._FIELDS

    def get(self, key: str, default: Any = None) -> Any:
        return getattr(self, key, default) if key in self._FIELDS else default

    def keys(self):
        return list(self._FIELDS)

    def model_dump(self) -> Dict[str, Any]:
        return {f: getattr(self, f) for f in self._FIELDS}

    def __repr__(self) -> str:
        return "GenerateResponse(" + ", ".join(f"{f}={getattr(self, f)!r}" for f in self._FIELDS) + ")"


def generate(model: str = "", prompt: str = "", suffix: str = "", system: str = "", template: str = "",
             context: Optional[Sequence[int]] = None, stream: bool = False, raw: bool = False,
             format: Any = None, images: Any = None, options: Optional[Dict[str, Any]] = None,
             keep_alive: Any = None, **_ignored: Any):
    t_start = time.perf_counter_ns()
    if not model:
        raise ResponseError("model is required", 400)
    try:
        from b200llm import runtime
        from b200llm._lib import B200LLMError
        from b200llm.engine import SamplingParams
        from b200llm.tokenizer import render_prompt
    except ImportError as e:                      # library not built: fail loudly, no fallback
        raise ResponseError(f"b200llm engine unavailable: {e}", 500) from e
    try:
        m = runtime.get_model(model)
    except KeyError:
        raise ResponseError(f"model '{model}' not found", 404) from None
    except B200LLMError as e:
        raise ResponseError(e.message, 500) from e
    opt = dict(DEFAULT_OPTIONS)
    opt.update(options or {})
    text = render_prompt(system or "", prompt or "", template or m.template, raw)
    ids = _list(context or []) + m.tokenizer.encode(text, bos=not context)
    max_ctx = m.engine.max_pages_per_lane * 16
    num_predict = int(opt["num_predict"])
    if num_predict < 0:
        num_predict = max_ctx
    ids = ids[-(max_ctx - 1):]                    # Ollama truncates over-long prompts to num_ctx
    num_predict = max(1, min(num_predict, max_ctx - len(ids)))
    sp = SamplingParams(temperature=float(opt["temperature"]), top_k=int(opt["top_k"]), top_p=float(opt["top_p"]),
                        repeat_penalty=float(opt["repeat_penalty"]),
                        seed=int(opt["seed"]) if opt.get("seed") is not None else secrets.randbits(63))
    try:
        toks, tm = m.scheduler.generate_ids(ids, num_predict, sp, stop_ids=m.tokenizer.stop_ids)
    except B200LLMError as e:
        raise ResponseError(e.message, 500) from e
    stopped = bool(toks) and toks[-1] in m.tokenizer.stop_ids
    out_toks = toks[:-1] if stopped else toks
    response = m.tokenizer.decode(out_toks)
    for s in opt.get("stop") or []:
        cut = response.find(s)
        if cut >= 0:
            response, stopped = response[:cut], True
    res = GenerateResponse(
        model=model, created_at=datetime.datetime.now(datetime.timezone.utc).isoformat(), response=response,
        done=True, done_reason="stop" if stopped else "length", context=_list(ids) + _list(out_toks),
        total_duration=time.perf_counter_ns() - t_start, load_duration=m.load_ns, prompt_eval_count=len(ids),
        prompt_eval_duration=int(tm.prefill_ms * 1e6), eval_count=len(toks), eval_duration=int(tm.decode_ms * 1e6))
    if stream:
        return iter([res])
    return res


def list():  # noqa: A001 - upstream name
    from b200llm import geometry
    return {"models": [{"model": n, "name": n} for n in geometry.REGISTRY]}


class Client:
    """ollama.Client(host=...) look-alike; the host is ignored (the engine is in-process)."""

    def __init__(self, host: Optional[str] = None, **_kw: Any):
        self.host = host

    def generate(self, *a: Any, **kw: Any):
        return generate(*a, **kw)

    def list(self):
        return list()
