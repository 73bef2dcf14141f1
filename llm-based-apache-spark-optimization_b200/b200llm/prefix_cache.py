"""Prefix cache: KV pages of prompt prefixes that earlier requests computed, shared with later ones.

The reference sends the SAME system prompt with every request (the NL->SQL instructions with the
schema at Flask/app.py:104 and FastAPI/app.py:87, the error-explanation instructions at
Flask/app.py:162 and FastAPI/app.py:107); only the short user part changes.  Ollama serves
that from its prompt cache; here the scheduler keeps the KV pages of the whole-page prefix of each
prompt it has prefilled and attaches them to a new lane whose prompt starts with the same tokens,
so only the rest is prefilled.  Pages are shared read-only and reference-counted by the engine
(Engine.retain_pages / release_pages / attach_prefix); entries leave in LRU order.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np

PAGE_TOKENS = 16


class PrefixCache:
    def __init__(self, engine, max_entries: int = 8, max_pages: int = 0):
        self.engine = engine
        self.max_entries = max_entries
        # at most a quarter of the KV pool is held by the cache
        self.max_pages = max_pages or max(1, engine.n_pages // 4)
        self._entries: List[Tuple[np.ndarray, List[int]]] = []     # (token ids of whole pages, pages), MRU last
        self.hits = self.misses = 0
        engine.on_page_pressure = self._on_pressure

    @staticmethod
    def _common(a: np.ndarray, b: np.ndarray) -> int:
        n = min(len(a), len(b))
        if n == 0:
            return 0
        diff = np.nonzero(a[:n] != b[:n])[0]
        return int(diff[0]) if len(diff) else n

    def lookup(self, ids: Sequence[int]) -> Tuple[int, List[int]]:
        """(shared tokens, their pages) for the longest cached whole-page prefix of `ids` that still leaves at
        least one token to prefill (the prefill produces the first output token)."""
        ids = np.asarray(ids, dtype=np.int32)
        best, best_i = 0, -1
        for i, (toks, _) in enumerate(self._entries):
            n = min(self._common(ids, toks), len(ids) - 1) // PAGE_TOKENS
            if n > best:
                best, best_i = n, i
        if best_i < 0:
            self.misses += 1
            return 0, []
        self.hits += 1
        entry = self._entries.pop(best_i)
        self._entries.append(entry)                                  # most recently used
        return best * PAGE_TOKENS, list(entry[1][:best])

    def insert(self, ids: Sequence[int], lane_pages: Sequence[int]) -> None:
        """Remember the whole-page prefix of a prompt that has just been prefilled into `lane_pages`."""
        ids = np.asarray(ids, dtype=np.int32)
        n = min(len(ids) // PAGE_TOKENS, len(lane_pages))
        if n == 0:
            return
        toks = ids[:n * PAGE_TOKENS].copy()
        for i, (t, _) in enumerate(self._entries):
            if len(t) >= len(toks) and self._common(t, toks) == len(toks):
                return                                               # an entry already covers this prefix
        # entries that this one extends are superseded
        keep = []
        for t, pg in self._entries:
            if len(t) <= len(toks) and self._common(t, toks) == len(t):
                self.engine.release_pages(pg)
            else:
                keep.append((t, pg))
        self._entries = keep
        pages = list(lane_pages[:n])
        self.engine.retain_pages(pages)
        self._entries.append((toks, pages))
        self._shrink()

    def _shrink(self) -> None:
        while self._entries and (len(self._entries) > self.max_entries or
                                 sum(len(p) for _, p in self._entries) > self.max_pages):
            _, pg = self._entries.pop(0)
            self.engine.release_pages(pg)

    def _on_pressure(self, short: int) -> None:
        """The pool cannot serve a lane: drop least recently used entries (pages still attached to a running
        lane only come back when that lane closes, so this may drop every entry)."""
        want = len(self.engine._free_pages) + short
        while self._entries and len(self.engine._free_pages) < want:
            _, pg = self._entries.pop(0)
            self.engine.release_pages(pg)

    def clear(self) -> None:
        for _, pg in self._entries:
            self.engine.release_pages(pg)
        self._entries = []
