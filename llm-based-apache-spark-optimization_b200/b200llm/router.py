"""Request router over several engines ("host router (round-robin / least-loaded) over 8 engines").

The Llama-3.2 models are served as replicas: requests are independent, every GPU holds the full weights, KV never
crosses GPUs.  The reference's threaded Flask server (`app.run`, Flask/app.py:239) puts several callers into
`ollama.generate` at once on its error path (Flask/app.py:151-166); behind this router they spread over the GPUs
instead of queueing on one.  One `Scheduler` (one thread, one engine, one GPU) per replica; the router only picks the
replica, so it has the scheduler's caller-side surface: submit / wait / generate_ids / shutdown.

Policies:
  round_robin   replica i, i + 1, ... in submission order (what BASELINE.json's config 3 names)
  least_loaded  the replica with the fewest outstanding tokens (prompt ids still to prefill + tokens still to
                generate of its unfinished requests); ties go to the lowest index.  Better when request lengths
                are mixed, as in config 3's log-uniform 64..1024-id prompts.
"""
from __future__ import annotations

import threading
from typing import Iterable, List, Optional, Sequence

from .engine import SamplingParams
from .scheduler import Request, Scheduler


class Router:
    def __init__(self, schedulers: Sequence[Scheduler], policy: str = "least_loaded"):
        if not schedulers:
            raise ValueError("Router needs at least one scheduler")
        if policy not in ("round_robin", "least_loaded"):
            raise ValueError(f"unknown routing policy '{policy}'")
        self.schedulers: List[Scheduler] = list(schedulers)
        self.policy = policy
        self._lock = threading.Lock()
        self._next = 0
        self._inflight: List[List[Request]] = [[] for _ in self.schedulers]
        self.routed = [0] * len(self.schedulers)          # requests sent to each replica so far

    # the first replica's engine stands for "the" engine where callers only need its limits (context length, ...)
    @property
    def engine(self):
        return self.schedulers[0].engine

    def _load(self, i: int) -> int:
        """Outstanding tokens of replica i; finished requests are dropped from its list on the way."""
        live = [r for r in self._inflight[i] if not r.done.is_set()]
        self._inflight[i] = live
        return sum(max(0, len(r.prompt_ids) - r.timing.prefill_tokens) + max(0, r.max_new - len(r.tokens)) for r in live)

    def pick(self) -> int:
        """Index of the replica the next request goes to (caller holds the lock)."""
        if self.policy == "round_robin":
            i = self._next
            self._next = (self._next + 1) % len(self.schedulers)
            return i
        loads = [self._load(i) for i in range(len(self.schedulers))]
        return loads.index(min(loads))

    def submit(self, prompt_ids: Sequence[int], max_new: int, sampling: Optional[SamplingParams] = None,
               stop_ids: Iterable[int] = ()) -> Request:
        with self._lock:
            i = self.pick()
            req = self.schedulers[i].submit(prompt_ids, max_new, sampling, stop_ids)
            req.replica = i
            self._inflight[i].append(req)
            self.routed[i] += 1
        return req

    wait = staticmethod(Scheduler.wait)

    def generate_ids(self, prompt_ids, max_new, sampling=None, stop_ids=()):
        return self.wait(self.submit(prompt_ids, max_new, sampling, stop_ids))

    def shutdown(self):
        for s in self.schedulers:
            s.shutdown()
