"""Host logic of the prefix cache (b200llm/prefix_cache.py) against a fake page allocator: no GPU.

The reference repeats its system prompts on every call (Flask/app.py:104,162; FastAPI/app.py:87,107);
the GPU side of the same feature is tests/test_gpu_engine.py::test_prefix_cache_shares_system_prompt_pages.
"""
import importlib.util
import os
import sys

import pytest

_PKG = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "llm-based-apache-spark-optimization_b200")


def _load():
    # prefix_cache.py has no dependency on the CUDA library: load it without importing the package
    spec = importlib.util.spec_from_file_location("_prefix_cache_under_test", os.path.join(_PKG, "b200llm", "prefix_cache.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod
    spec.loader.exec_module(mod)
    return mod


class FakeEngine:
    """The page bookkeeping of Engine (reference counts + free list), nothing else."""

    def __init__(self, n_pages):
        self.n_pages = n_pages
        self._free_pages = list(range(n_pages - 1, -1, -1))
        self._page_ref = [0] * n_pages
        self.on_page_pressure = None

    def alloc(self, n):
        if n > len(self._free_pages) and self.on_page_pressure:
            self.on_page_pressure(n - len(self._free_pages))
        assert n <= len(self._free_pages), "pool exhausted"
        pages = [self._free_pages.pop() for _ in range(n)]
        for p in pages:
            self._page_ref[p] = 1
        return pages

    def retain_pages(self, pages):
        for p in pages:
            self._page_ref[p] += 1

    def release_pages(self, pages):
        for p in reversed(list(pages)):
            self._page_ref[p] -= 1
            assert self._page_ref[p] >= 0
            if self._page_ref[p] == 0:
                self._free_pages.append(p)


def _serve(cache, eng, ids):
    """What Scheduler._admit + _finish do with the cache for one request; returns tokens prefilled."""
    shared, pages = cache.lookup(ids)
    eng.retain_pages(pages)                                   # attach_prefix
    own = eng.alloc((len(ids) + 15) // 16 - len(pages))       # prefill of the rest
    lane_pages = list(pages) + own
    cache.insert(ids, lane_pages)
    eng.release_pages(lane_pages)                             # close_lane
    return len(ids) - shared


def test_lookup_insert_and_reference_counts():
    pc = _load()
    eng = FakeEngine(64)
    cache = pc.PrefixCache(eng, max_entries=4, max_pages=16)
    system = list(range(100, 153))                             # 53 tokens: 3 whole pages
    a, b = system + [1, 2, 3], system + [7] * 40
    assert _serve(cache, eng, a) == len(a)                     # miss
    assert _serve(cache, eng, b) == len(b) - 48                # the 3 pages of the system prompt are shared
    assert _serve(cache, eng, b) == len(b) - 80                # b itself: 93 tokens, 5 whole pages
    assert (cache.hits, cache.misses) == (2, 1)
    # the 5-page entry of b superseded the 3-page entry of a
    assert len(cache._entries) == 1 and len(cache._entries[0][1]) == 5
    assert sum(1 for r in eng._page_ref if r) == 5 and len(eng._free_pages) == 64 - 5
    cache.clear()
    assert len(eng._free_pages) == 64 and not any(eng._page_ref)


def test_a_hit_always_leaves_a_token_to_prefill():
    pc = _load()
    eng = FakeEngine(32)
    cache = pc.PrefixCache(eng)
    ids = list(range(64))                                      # exactly 4 pages
    assert _serve(cache, eng, ids) == 64
    assert _serve(cache, eng, ids) == 16                       # only 3 pages reused: the prefill must emit a token
    assert _serve(cache, eng, ids[:20]) == 4
    assert _serve(cache, eng, [5] + ids) == 65                 # shifted by one token: nothing in common
    assert _serve(cache, eng, ids[:15]) == 15                  # shorter than a page


def test_lru_and_page_budget():
    pc = _load()
    eng = FakeEngine(64)
    cache = pc.PrefixCache(eng, max_entries=3, max_pages=8)
    prompts = [[1000 * k + i for i in range(40)] for k in range(5)]      # 2 whole pages each, no common prefix
    for p in prompts[:3]:
        _serve(cache, eng, p)
    assert len(cache._entries) == 3
    _serve(cache, eng, prompts[0])                             # touch 0: entry 1 is now the oldest
    _serve(cache, eng, prompts[3])
    firsts = sorted(int(t[0]) for t, _ in cache._entries)
    assert firsts == [0, 2000, 3000]
    big = [77000 + i for i in range(16 * 7 + 3)]               # 7 pages: only fits the 8-page budget with one other gone...
    _serve(cache, eng, big)
    assert sum(len(pg) for _, pg in cache._entries) <= 8
    assert int(cache._entries[-1][0][0]) == 77000
    cache.clear()
    assert len(eng._free_pages) == 64


def test_pool_pressure_evicts_entries():
    pc = _load()
    eng = FakeEngine(8)
    cache = pc.PrefixCache(eng, max_entries=8, max_pages=8)
    _serve(cache, eng, [1] * 48)                               # 3 pages stay cached
    _serve(cache, eng, [2] * 48)                               # 3 more
    assert len(eng._free_pages) == 2
    got = eng.alloc(7)                                         # a lane that needs 7 of the 8 pages
    assert len(got) == 7 and len(cache._entries) == 0
    eng.release_pages(got)
    assert len(eng._free_pages) == 8
