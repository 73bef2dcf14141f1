"""Continuous-batching scheduler: the thread-safe submit()/wait() surface of SURVEY.md 8(b).

The reference's Flask dev server is threaded (`app.run`, Flask/app.py:239), so several Python
threads may sit in `ollama.generate` at once; behind FastAPI the handler serialises requests
(FastAPI/app.py:63).  One scheduler thread per engine owns the GPU: it admits queued requests
(prefill, one lane each), then advances every active lane together, `chunk` decode steps per
engine call, retiring lanes at their stop token or token budget.  Batch > 1 therefore arises on
its own when callers overlap; a lone caller gets the batch-1 persistent-kernel path.
"""
from __future__ import annotations

import threading
from collections import deque
from dataclasses import dataclass, field
from typing import Deque, Dict, Iterable, List, Optional, Sequence

from ._lib import EINVAL, B200LLMError
from .engine import MAX_DECODE_STEPS, PAGE_TOKENS, Engine, GenTiming, SamplingParams
from .prefix_cache import PrefixCache


@dataclass
class Request:
    prompt_ids: Sequence[int]
    max_new: int
    sampling: SamplingParams
    stop_ids: frozenset
    tokens: List[int] = field(default_factory=list)
    timing: GenTiming = field(default_factory=GenTiming)
    error: Optional[BaseException] = None
    done: threading.Event = field(default_factory=threading.Event)
    lane: int = -1


class Scheduler:
    def __init__(self, engine: Engine, chunk: int = 16, prefix_cache: bool = True):
        self.engine = engine
        self.chunk = chunk
        # KV pages of prompt prefixes seen before (the reference repeats its system prompts): prefix_cache.py
        self.cache = PrefixCache(engine) if prefix_cache else None
        self._queue: Deque[Request] = deque()
        self._cv = threading.Condition()
        self._active: Dict[int, Request] = {}
        self._stop = False
        self._thread = threading.Thread(target=self._run, name="b200llm-scheduler", daemon=True)
        self._thread.start()

    # ---- caller side (any thread) ------------------------------------------------------
    def submit(self, prompt_ids: Sequence[int], max_new: int, sampling: Optional[SamplingParams] = None,
               stop_ids: Iterable[int] = ()) -> Request:
        req = Request(list(prompt_ids), int(max_new), sampling or SamplingParams(), frozenset(int(s) for s in stop_ids))
        with self._cv:
            if self._stop:
                raise RuntimeError("scheduler is shut down")
            self._queue.append(req)
            self._cv.notify()
        return req

    @staticmethod
    def wait(req: Request, timeout: Optional[float] = None):
        if not req.done.wait(timeout):
            raise TimeoutError("generation still running")
        if req.error is not None:
            raise req.error
        return req.tokens, req.timing

    def generate_ids(self, prompt_ids, max_new, sampling=None, stop_ids=()):
        return self.wait(self.submit(prompt_ids, max_new, sampling, stop_ids))

    def shutdown(self):
        with self._cv:
            self._stop = True
            self._cv.notify()
        self._thread.join(timeout=30)
        if self.cache is not None:
            self.cache.clear()

    # ---- scheduler thread -----------------------------------------------------------------
    def _finish(self, req: Request, error: Optional[BaseException] = None):
        if req.lane >= 0:
            self.engine.close_lane(req.lane)
            self._active.pop(req.lane, None)
            req.lane = -1
        req.error = error
        req.done.set()

    def _admit(self, req: Request):
        eng = self.engine
        try:
            # a request is checked against the lane context here, so that it alone fails (or is cut to what fits):
            # found later, inside a batched decode call, the error would take every co-batched request with it
            max_ctx = eng.max_pages_per_lane * PAGE_TOKENS
            if len(req.prompt_ids) < 1 or len(req.prompt_ids) >= max_ctx:
                raise B200LLMError(EINVAL, f"prompt of {len(req.prompt_ids)} ids does not fit the {max_ctx}-token context")
            req.max_new = max(1, min(req.max_new, max_ctx - len(req.prompt_ids)))
            req.lane = eng.open_lane()
            eng.set_sampling(req.lane, req.sampling)
            shared, pages = self.cache.lookup(req.prompt_ids) if self.cache else (0, [])
            if shared:
                eng.attach_prefix(req.lane, pages, req.prompt_ids[:shared])
                eng.prefill(req.lane, req.prompt_ids[shared:], pos0=shared)
            else:
                eng.prefill(req.lane, req.prompt_ids)
            if self.cache:
                self.cache.insert(req.prompt_ids, eng.lane_pages(req.lane))
            first, _ = eng.read_lane(req.lane)
            t = eng.timing()
            req.timing.prefill_ms, req.timing.prefill_tokens = t.prefill_ms, t.prefill_tokens
            req.tokens.append(first)
            self._active[req.lane] = req
            if first in req.stop_ids or len(req.tokens) >= req.max_new:
                self._finish(req)
        except BaseException as e:  # noqa: BLE001 - handed to the waiting caller
            self._finish(req, e)

    def _run(self):
        import torch
        torch.cuda.set_device(self.engine.device)
        eng = self.engine
        while True:
            with self._cv:
                while not self._queue and not self._active and not self._stop:
                    self._cv.wait()
                if self._stop and not self._queue and not self._active:
                    return
                admit = []
                while self._queue and len(self._active) + len(admit) < eng.max_lanes:
                    admit.append(self._queue.popleft())
            for req in admit:
                self._admit(req)
            if not self._active:
                continue
            lanes = sorted(self._active)
            budget = min(self._active[ln].max_new - len(self._active[ln].tokens) for ln in lanes)
            # a lone request with no stop ids can run to its budget in one call; otherwise step in
            # chunks so that finished lanes leave and queued requests join
            lone = len(lanes) == 1 and not self._queue and not self._active[lanes[0]].stop_ids
            n = min(MAX_DECODE_STEPS, max(1, budget if lone else min(budget, self.chunk)))
            # pages for the whole call, lane by lane: a lane that cannot get them (pool exhausted while other lanes pin
            # prefix pages) is retired with its own error, the others decode on
            for ln in list(lanes):
                try:
                    eng._ensure_pages(ln, eng.pos[ln] + n)
                except BaseException as e:  # noqa: BLE001
                    self._finish(self._active[ln], e)
                    lanes.remove(ln)
            if not lanes:
                continue
            try:
                out = eng.decode(lanes, n)
                t = eng.timing()
            except BaseException as e:  # noqa: BLE001 - an engine fault: nothing in the batch can be trusted
                for ln in list(lanes):
                    self._finish(self._active[ln], e)
                continue
            for i, ln in enumerate(lanes):
                req = self._active[ln]
                req.timing.decode_ms += t.decode_ms
                req.timing.decode_steps += n
                req.timing.kernel_launches += t.kernel_launches
                hit = False
                for tok in out[:, i].tolist():
                    req.tokens.append(tok)
                    if tok in req.stop_ids or len(req.tokens) >= req.max_new:
                        hit = True
                        break
                if hit:
                    self._finish(req)
