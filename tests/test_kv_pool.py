"""One KV pool for two engines (b200llm/kv_pool.py, b200llm_bind_kv_pool_shared; BASELINE.json config 5: the 7B and the
3B co-served on every GPU; the reference uses both models from one process, Flask/app.py:102-107 and :160-166).
CPU: the block / slot allocator.  -m gpu: two engines of different geometry decode interleaved requests out of one
pool and every request still equals the oracle; blocks travel from one engine to the other and all come back."""
import pytest

from b200llm.kv_pool import SharedKVPool


def test_slots_blocks_and_occupancy():
    pool = SharedKVPool(n_blocks=4, block_bytes=8 << 20, allocate=False)
    big = pool.view("duckdb-nsql", 8 << 20, 4096)                      # a 7B page is exactly one block
    small = pool.view("llama3.2", 28 * 2 * 8 * 16 * 128 * 2, 4096)      # 1.75 MiB pages: 4 slots of 2 MiB per block
    assert (big.k, big.slot_bytes) == (1, 8 << 20) and (small.k, small.slot_bytes) == (4, 2 << 20)
    assert len(big) == 4 and len(small) == 16 and small.n_slots == 16 and big.n_slots == 4
    a = [small.pop() for _ in range(5)]                                 # 5 small pages: two blocks
    assert len(set(a)) == 5 and pool.occupancy()["free_blocks"] == 2
    assert {p // 4 for p in a[:4]} == {a[0] // 4} and a[4] // 4 != a[0] // 4     # the first four share a block
    b = [big.pop(), big.pop()]                                          # the 7B takes the other two blocks
    assert pool.occupancy()["free_blocks"] == 0 and len(big) == 0 and len(small) == 3
    with pytest.raises(IndexError):
        big.pop()
    # slots never overlap in bytes, whoever owns them
    spans = [(p * small.slot_bytes, (p + 1) * small.slot_bytes) for p in a] + [(p * big.slot_bytes, (p + 1) * big.slot_bytes) for p in b]
    spans.sort()
    assert all(spans[i][1] <= spans[i + 1][0] for i in range(len(spans) - 1))
    small.append(a[4])                                                  # its block empties -> back to the pool -> the 7B can have it
    assert pool.occupancy()["free_blocks"] == 1
    c = big.pop()
    assert c == a[4] // 4
    for p in a[:4]:
        small.append(p)
    for p in b + [c]:
        big.append(p)
    occ = pool.occupancy()
    assert occ["free_blocks"] == 4 and occ["peak_blocks_in_use"] == 4
    assert all(e["slots_in_use"] == 0 and e["blocks_held"] == 0 for e in occ["engines"].values())


def test_view_rejects_pages_that_do_not_fit():
    pool = SharedKVPool(2, 1 << 20, allocate=False)
    with pytest.raises(ValueError):
        pool.view("too-big", (1 << 20) + 4096, 4096)
    pool.view("a", 300_000 // 4096 * 4096, 4096)
    with pytest.raises(ValueError):
        pool.view("a", 4096, 4096)          # attached already


@pytest.mark.gpu
def test_two_models_decode_out_of_one_pool():
    import threading
    import torch
    from b200llm import Engine, SamplingParams, geometry, weights
    from b200llm.scheduler import Scheduler
    from oracle import llama_oracle as lo
    from oracle import weightgen as wg
    torch.cuda.set_device(0)
    names = ["small", "tiny"]                       # different layer counts, head_dim and kv heads: different page sizes
    ogeo = {n: lo.GEOMETRIES[n] for n in names}
    sd = {n: lo.make_weights(ogeo[n], 0xB200 + i, sigma=0.07) for i, n in enumerate(names)}
    orc = {n: lo.LlamaOracle(ogeo[n], sd[n], kv_dtype=torch.bfloat16) for n in names}
    page = {n: ogeo[n].n_layers * 2 * ogeo[n].n_kv * 16 * ogeo[n].head_dim * 2 for n in names}
    pool = SharedKVPool(n_blocks=96, block_bytes=max(page.values()), device=0)
    eng, sched = {}, {}
    for n in names:
        g = geometry.resolve(n)
        eng[n] = Engine(g, weights.pack(g, {k: v.cuda() for k, v in sd[n].items()}), max_lanes=4, max_ctx=192, max_rows=16,
                        kv_pool=pool, name=n)
        sched[n] = Scheduler(eng[n], chunk=4, prefix_cache=False)
    assert eng["small"]._free_pages.k == 1 and eng["tiny"]._free_pages.k > 1
    jobs = [(names[i % 2], wg.gen_tokens(20 + 13 * i, 400 + i, ogeo[names[i % 2]].vocab), 8 + (i * 7) % 15) for i in range(12)]
    want = [orc[n].greedy(p, k)[0] for n, p, k in jobs]
    got, peak = [None] * len(jobs), []

    def call(i):
        n, p, k = jobs[i]
        got[i] = sched[n].generate_ids(p, k, SamplingParams(temperature=0.0))[0]
        peak.append(pool.occupancy()["peak_blocks_in_use"])
    ts = [threading.Thread(target=call, args=(i,)) for i in range(len(jobs))]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=180)
    assert got == want
    occ = pool.occupancy()
    assert occ["free_blocks"] == 96 and max(peak) >= 4                  # everything came back; both engines held blocks
    for n in names:
        sched[n].shutdown()
        eng[n].close()
