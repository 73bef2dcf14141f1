"""Request router over replicas (b200llm/router.py; SURVEY.md 8(e) "host router (round-robin / least-loaded) over 8
engines").  CPU: the policies against stand-in schedulers.  -m gpu: 16 mixed-length requests from concurrent callers over
two real schedulers give every caller exactly the oracle's tokens, and both replicas get work."""
import threading

import pytest

from b200llm.engine import GenTiming, SamplingParams
from b200llm.router import Router
from b200llm.scheduler import Request


class _FakeScheduler:
    """Accepts requests and finishes them only when told to, so that load can pile up."""

    def __init__(self):
        self.reqs = []
        self.engine = object()

    def submit(self, prompt_ids, max_new, sampling=None, stop_ids=()):
        r = Request(list(prompt_ids), int(max_new), sampling or SamplingParams(), frozenset(stop_ids))
        self.reqs.append(r)
        return r

    def finish(self, r):
        r.tokens = [1] * r.max_new
        r.timing = GenTiming(prefill_tokens=len(r.prompt_ids))
        r.done.set()

    def shutdown(self):
        pass


def test_round_robin_cycles_over_the_replicas():
    scheds = [_FakeScheduler() for _ in range(3)]
    router = Router(scheds, "round_robin")
    for i in range(7):
        router.submit([0] * (10 + i), 5)
    assert router.routed == [3, 2, 2]
    assert [len(s.reqs) for s in scheds] == [3, 2, 2]


def test_least_loaded_follows_outstanding_tokens():
    scheds = [_FakeScheduler() for _ in range(2)]
    router = Router(scheds, "least_loaded")
    a = router.submit([0] * 1000, 100)       # replica 0: 1100 tokens outstanding
    b = router.submit([0] * 50, 10)          # replica 1: 60
    c = router.submit([0] * 50, 10)          # replica 1 again (120 < 1100)
    d = router.submit([0] * 50, 10)          # and again (180 < 1100)
    assert (a.replica, b.replica, c.replica, d.replica) == (0, 1, 1, 1)
    scheds[0].finish(a)                      # the long request is done: replica 0 is idle
    e = router.submit([0] * 50, 10)
    assert e.replica == 0
    # prefill progress counts: a request whose prompt is already in costs only its remaining budget
    b.timing = GenTiming(prefill_tokens=50)
    b.tokens = [1] * 9
    assert router._load(1) == 1 + 60 + 60


def test_router_rejects_bad_arguments():
    with pytest.raises(ValueError):
        Router([], "round_robin")
    with pytest.raises(ValueError):
        Router([_FakeScheduler()], "random")


@pytest.mark.gpu
@pytest.mark.parametrize("policy", ["round_robin", "least_loaded"])
def test_sixteen_mixed_requests_over_two_replicas_match_the_oracle(policy):
    import torch
    from b200llm import Engine, geometry, weights
    from b200llm.scheduler import Scheduler
    from oracle import llama_oracle as lo
    from oracle import weightgen as wg
    geo = lo.GEOMETRIES["tiny"]
    sd = lo.make_weights(geo, 0xB200, sigma=0.08)
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    g = geometry.resolve("tiny")
    # two replicas; on a one-GPU box both live on device 0 (each with its own weights, KV pool, stream and thread)
    devs = [0, 1] if torch.cuda.device_count() > 1 else [0, 0]
    engines = []
    for d in devs:
        torch.cuda.set_device(d)
        engines.append(Engine(g, weights.pack(g, {k: v.cuda(d) for k, v in sd.items()}), device=d, max_lanes=4, max_ctx=160,
                              max_rows=8, use_tc=False))
    router = Router([Scheduler(e, chunk=4) for e in engines], policy)
    lens = [5, 90, 17, 64, 33, 8, 120, 41, 12, 77, 3, 58, 100, 26, 9, 70]
    prompts = [wg.gen_tokens(n, 1200 + i, geo.vocab) for i, n in enumerate(lens)]
    budgets = [6 + (i * 5) % 19 for i in range(16)]
    want = [orc.greedy(p, n)[0] for p, n in zip(prompts, budgets)]
    got = [None] * 16

    def call(i):
        got[i] = router.generate_ids(prompts[i], budgets[i], SamplingParams(temperature=0.0))[0]
    threads = [threading.Thread(target=call, args=(i,)) for i in range(16)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=180)
    assert got == want
    assert sum(router.routed) == 16 and min(router.routed) >= 4        # both replicas worked
    router.shutdown()
    for e in engines:
        e.close()
    torch.cuda.set_device(0)        # (later tests of the session allocate on the current device)
