#!/usr/bin/env python
"""Batched synthetic benchmark harness for the B200 decode engine.

Takes the place of the reference's `Model_Evaluation_&_Comparision.py` (a sequential, batch-1
script that timed `ollama.generate` with time.time() and scored exact match / Levenshtein
distance; reference lines 19-66 and 109-158).  Those three text columns are kept -- they go
through the same `ollama.generate(model=, system=, prompt=)` call -- and the harness adds what
BASELINE.json asks for: decode tokens/s and p50 TTFT per model and batch size, as absolute
numbers and as a fraction of the measured HBM roofline, on 1 GPU or (under torchrun) on N GPUs
as request-level replicas, with the CPU path (the oracle port; the real Ollama/llama.cpp cannot
be built here) timed on the host cores in the same run.

    python "Model_Evaluation_&_Comparision.py" [--models llama3.2 duckdb-nsql] [--batch 1 8 64]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 "Model_Evaluation_&_Comparision.py"

With random-init weights and the stand-in byte tokenizer the text metrics are meaningless as
quality numbers (exact match 0, large edit distance); they are reported to keep the columns.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for _p in (ROOT, HERE):
    if _p not in sys.path:
        sys.path.insert(0, _p)

SCHEMA = ("CREATE TABLE taxi (VendorID bigint, tpep_pickup_datetime timestamp, tpep_dropoff_datetime timestamp, "
          "passenger_count double, trip_distance double, fare_amount double, extra double, tip_amount double, "
          "tolls_amount double, improvement_surcharge double, total_amount double);")
# (question, expected SQL): the four multi-query cases of the reference's second part
CASES = [
    ("Get all taxis with more than 2 passengers.", "SELECT * FROM taxi WHERE passenger_count > 2;"),
    ("Show total fare collected by each vendor.",
     "SELECT VendorID, SUM(total_amount) AS Total_Fare FROM taxi GROUP BY VendorID;"),
    ("Find the average trip distance for trips that had more than 2 passengers.",
     "SELECT AVG(trip_distance) FROM taxi WHERE passenger_count > 2;"),
    ("List all vendors ordered by total fare in descending order.",
     "SELECT VendorID, SUM(total_amount) AS Total_Fare FROM taxi GROUP BY VendorID ORDER BY Total_Fare DESC;"),
]


def edit_distance(a: str, b: str) -> int:
    """Levenshtein distance (the `Levenshtein` package is not on this image)."""
    if len(a) < len(b):
        a, b = b, a
    prev = list(range(len(b) + 1))
    for i, ca in enumerate(a, 1):
        cur = [i]
        for j, cb in enumerate(b, 1):
            cur.append(min(prev[j] + 1, cur[j - 1] + 1, prev[j - 1] + (ca != cb)))
        prev = cur
    return prev[-1]


def text_metrics(model: str, greedy: bool):
    """The reference's columns: exact match, edit distance, wall latency of ollama.generate."""
    import ollama
    rows = []
    opts = {"temperature": 0, "num_predict": 48} if greedy else {"num_predict": 48}
    for question, expected in CASES:
        t0 = time.time()
        try:
            res = ollama.generate(model=model, system="Here is the database schema that the SQL query will run on: "
                                  + SCHEMA, prompt=question, options=opts)
        except ollama.ResponseError as e:
            return {"model": model, "error": str(e)}
        latency = time.time() - t0
        got = res.response.strip()
        rows.append({"exact": int(got == expected), "edit": edit_distance(got, expected), "latency_s": latency,
                     "ttft_ms": (res.load_duration + res.prompt_eval_duration) / 1e6,
                     "tok_s": res.eval_count / max(res.eval_duration, 1) * 1e9})
    n = len(rows)
    return {"model": model, "exact_match_rate": 100.0 * sum(r["exact"] for r in rows) / n,
            "avg_edit_distance": sum(r["edit"] for r in rows) / n, "avg_latency_s": sum(r["latency_s"] for r in rows) / n,
            "queries": rows}


def engine_bench(model: str, batch_sizes, prompt_len: int, new_tokens: int, n_ttft: int, local: int):
    """decode tok/s per batch size + p50 TTFT on this rank's GPU (synthetic prompts, greedy)."""
    import numpy as np
    import torch
    from b200llm import Engine, geometry, weights
    geo = geometry.resolve(model)
    w = weights.random_init(geo, 0xB200, local)
    max_bs = max(batch_sizes)
    eng = Engine(geo, w, device=local, max_lanes=max_bs, max_ctx=prompt_len + new_tokens + 32, max_rows=256)   # prefill in 256-row chunks, decode batches up to 256 lanes
    rng = np.random.default_rng(7 + local)
    out = {"model": model, "ttft_ms_p50": None, "decode": []}
    ttfts = []
    for _ in range(n_ttft):
        ln = eng.open_lane()
        p = rng.integers(0, min(geo.vocab, 128000), prompt_len, dtype=np.int32)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        eng.prefill(ln, p)
        eng.read_lane(ln)                 # first token on the host = TTFT
        ttfts.append((time.perf_counter() - t0) * 1e3)
        eng.close_lane(ln)
    out["ttft_ms_p50"] = statistics.median(ttfts[1:] or ttfts)
    for bs in batch_sizes:
        lanes = [eng.open_lane() for _ in range(bs)]
        for ln in lanes:
            eng.prefill(ln, rng.integers(0, min(geo.vocab, 128000), prompt_len, dtype=np.int32))
        eng.decode(lanes, 4, read=False)  # warm-up / graph capture
        eng.sync()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(eng.stream)
        eng.decode(lanes, new_tokens, read=False)
        ev1.record(eng.stream)
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1) / new_tokens
        mean_ctx = prompt_len + 4 + (new_tokens - 1) / 2
        out["decode"].append({"bs": bs, "ms_per_step": ms, "tok_s": bs * 1000.0 / ms,
                              "bytes_per_step": geo.step_bytes(bs, mean_ctx)})
        for ln in lanes:
            eng.close_lane(ln)
    eng.close()
    del w
    torch.cuda.empty_cache()
    return out


def config3_sweep(model: str, device_counts, policy: str, n_requests: int, new_tokens: int, max_lanes: int):
    """BASELINE.json config 3 / SURVEY.md 8(d): `n_requests` synthetic Spark error logs, prompt lengths log-uniform in
    [64, 1024] ids (seed 3), `new_tokens` greedy tokens each, routed over one replica per GPU by b200llm.router.Router
    (one process, one scheduler thread per GPU, continuous batching of up to `max_lanes` requests per replica).
    Stands for the concurrent callers of the reference's threaded server (Flask/app.py:239) on its error path
    (Flask/app.py:151-166).  Returns one row per GPU count: aggregate tok/s, efficiency vs one GPU, routing counts."""
    import numpy as np
    import torch
    from b200llm import Engine, SamplingParams, geometry, weights
    from b200llm.router import Router
    from b200llm.scheduler import Scheduler
    geo = geometry.resolve(model)
    rng = np.random.default_rng(3)
    lens = np.exp(rng.uniform(np.log(64), np.log(1024), n_requests)).astype(int)
    prompts = [rng.integers(0, min(geo.vocab, 128000), int(n), dtype=np.int32) for n in lens]
    rows, base = [], None
    for n_gpu in device_counts:
        engines = []
        for d in range(n_gpu):
            torch.cuda.set_device(d)
            engines.append(Engine(geo, weights.random_init(geo, 0xB200, d), device=d, max_lanes=max_lanes,
                                  max_ctx=1024 + new_tokens + 32, max_rows=256))
        router = Router([Scheduler(e, chunk=16, prefix_cache=False) for e in engines], policy)
        warm = [router.submit(prompts[i % n_requests][:64], 4, SamplingParams(temperature=0.0)) for i in range(2 * n_gpu)]
        for r in warm:
            router.wait(r)
        router.routed = [0] * n_gpu
        for d in range(n_gpu):
            torch.cuda.synchronize(d)
        t0 = time.perf_counter()
        reqs = [router.submit(p, new_tokens, SamplingParams(temperature=0.0)) for p in prompts]
        n_tok = sum(len(router.wait(r)[0]) for r in reqs)
        dt = time.perf_counter() - t0
        ttft = statistics.median(r.timing.prefill_ms for r in reqs)
        row = {"gpus": n_gpu, "requests": n_requests, "prompt_ids_total": int(lens.sum()), "new_tokens": n_tok, "wall_s": dt,
               "tok_s": n_tok / dt, "prefill_ms_p50": ttft, "routed": list(router.routed), "policy": policy,
               "max_lanes_per_replica": max_lanes}
        base = base or row["tok_s"] / n_gpu
        row["efficiency_vs_1gpu"] = row["tok_s"] / (base * n_gpu)
        rows.append(row)
        router.shutdown()
        for e in engines:
            e.close()
        del engines, router
        torch.cuda.empty_cache()
    return rows


def config5(n_gpus: int, in_flight_per_model: int, new_tokens: int, pool_gib: float, policy: str):
    """BASELINE.json config 5 / SURVEY.md 8(d): duckdb-nsql-7B and Llama-3.2-3B co-resident on each of `n_gpus` GPUs, their
    KV pages drawn from ONE pool per GPU (b200llm.kv_pool.SharedKVPool, 16-token pages), `in_flight_per_model` sequences
    per model in flight over the whole box (requests routed per model over the GPUs by b200llm.router.Router), 3B prompts
    of 296 ids, 7B prompts log-uniform in [128, 2048] ids, `new_tokens` greedy tokens each.  The reference serves both
    models from one process (Flask/app.py:102-107 and :160-166).  Returns aggregate tok/s per model and pool occupancy."""
    import threading
    import numpy as np
    import torch
    from b200llm import Engine, SamplingParams, geometry, weights
    from b200llm.kv_pool import SharedKVPool
    from b200llm.router import Router
    from b200llm.scheduler import Scheduler
    models = ["duckdb-nsql", "llama3.2"]
    geos = {m: geometry.resolve(m) for m in models}
    page = {m: geos[m].n_layers * 2 * geos[m].n_kv * 16 * geos[m].head_dim * 2 for m in models}
    block = max(page.values())
    lanes = max(1, (in_flight_per_model + n_gpus - 1) // n_gpus)
    pools, engines = [], {m: [] for m in models}
    for d in range(n_gpus):
        torch.cuda.set_device(d)
        pools.append(SharedKVPool(int(pool_gib * (1 << 30)) // block, block, device=d))
        for m in models:
            engines[m].append(Engine(geos[m], weights.random_init(geos[m], 0xB200, d), device=d, max_lanes=lanes,
                                     max_ctx=2048 + new_tokens + 32, max_rows=256, kv_pool=pools[d], name=m,
                                     lean_weights=True))   # co-residency: tile-major matrices only, 13.2 + 6.4 GB instead of 26.4 + 12.9
    routers = {m: Router([Scheduler(e, chunk=16, prefix_cache=False) for e in engines[m]], policy) for m in models}
    rng = np.random.default_rng(5)
    n_req = 2 * in_flight_per_model                  # two waves per model, so that lanes are refilled while others still run
    prompts = {"llama3.2": [rng.integers(0, 128000, 296, dtype=np.int32) for _ in range(n_req)],
               "duckdb-nsql": [rng.integers(0, geos["duckdb-nsql"].vocab, int(n), dtype=np.int32)
                               for n in np.exp(rng.uniform(np.log(128), np.log(2048), n_req)).astype(int)]}
    for m in models:                                 # warm-up: tensor maps, graphs
        for r in [routers[m].submit(prompts[m][i][:128], 4, SamplingParams(temperature=0.0)) for i in range(n_gpus)]:
            routers[m].wait(r)
    for d in range(n_gpus):
        torch.cuda.synchronize(d)
    peak = [0] * n_gpus
    stop = threading.Event()

    def watch():
        while not stop.is_set():
            for d in range(n_gpus):
                occ = pools[d].occupancy()
                peak[d] = max(peak[d], occ["blocks"] - occ["free_blocks"])
            time.sleep(0.02)
    w = threading.Thread(target=watch, daemon=True)
    w.start()
    t0 = time.perf_counter()
    reqs = {m: [routers[m].submit(p, new_tokens, SamplingParams(temperature=0.0)) for p in prompts[m]] for m in models}
    done_at, n_tok = {}, {}

    def collect(m):                                  # one waiter per model: each model's wall time ends with ITS last request
        n_tok[m] = sum(len(routers[m].wait(r)[0]) for r in reqs[m])
        done_at[m] = time.perf_counter() - t0
    waiters = [threading.Thread(target=collect, args=(m,)) for m in models]
    for t in waiters:
        t.start()
    for t in waiters:
        t.join()
    stop.set()
    w.join()
    out = {"gpus": n_gpus, "in_flight_per_model": in_flight_per_model, "lanes_per_engine": lanes, "new_tokens_each": new_tokens,
           "policy": policy, "models": {}, "pool": {"block_bytes": block, "blocks_per_gpu": pools[0].n_blocks,
                                                     "gib_per_gpu": pools[0].nbytes / (1 << 30),
                                                     "peak_blocks_in_use_per_gpu": peak,
                                                     "peak_occupancy": max(peak) / pools[0].n_blocks,
                                                     "after": [p.occupancy()["free_blocks"] for p in pools]}}
    for m in models:
        out["models"][m] = {"requests": n_req, "prompt_ids_total": int(sum(len(p) for p in prompts[m])), "new_tokens": n_tok[m],
                            "wall_s": done_at[m], "tok_s": n_tok[m] / done_at[m], "routed": list(routers[m].routed),
                            "slots_per_block": engines[m][0]._free_pages.k}
    for m in models:
        routers[m].shutdown()
        for e in engines[m]:
            e.close()
    return out


def cpu_row(model: str):
    import bench
    v, dt, cores = bench.cpu_decode_tok_s(model, 64, 12, 1)
    return {"tok_s": v, "cores": cores, "sample": "12 decode steps at bs1 after a 64-id prompt, oracle port, torch CPU bf16",
            "quantised": bench.cpu_quantised_tok_s(model)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--models", nargs="+", default=["llama3.2", "duckdb-nsql"])
    ap.add_argument("--batch", nargs="+", type=int, default=[1, 8, 64])
    ap.add_argument("--prompt-len", type=int, default=296)
    ap.add_argument("--new-tokens", type=int, default=128)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-text", action="store_true")
    ap.add_argument("--config3", action="store_true", help="config 3: 512 mixed-length requests routed over 1/2/4/8 replicas "
                    "(single process; do not run under torchrun)")
    ap.add_argument("--config5", action="store_true", help="config 5: the 7B and the 3B co-served on every GPU out of one KV pool "
                    "per GPU, 128 sequences per model in flight (single process; do not run under torchrun)")
    ap.add_argument("--tp7b", action="store_true", help="config 4: duckdb-nsql-7B tensor parallel over all ranks (run under torchrun "
                    "with 2, 4 or 8 ranks): 2048 + 24 prompt ids, 64 new tokens; ms per step, tok/s, TTFT, all-reduce share")
    ap.add_argument("--in-flight", type=int, default=128)
    ap.add_argument("--pool-gib", type=float, default=48.0)
    ap.add_argument("--policy", default="least_loaded", choices=["least_loaded", "round_robin"])
    ap.add_argument("--requests", type=int, default=512)
    ap.add_argument("--lanes", type=int, default=16)
    args = ap.parse_args()
    if args.config3:
        import torch
        counts = [n for n in (1, 2, 4, 8) if n <= torch.cuda.device_count()]
        rows = config3_sweep("llama3.2", counts, args.policy, args.requests, args.new_tokens, args.lanes)
        print("=" * 100)
        print(f"Config 3: Llama-3.2-3B, {args.requests} requests, prompt lengths log-uniform 64..1024 ids (seed 3), "
              f"{args.new_tokens} new tokens each, router policy {args.policy}, <= {args.lanes} lanes per replica")
        for r in rows:
            print(f"  {r['gpus']} GPU(s): {r['tok_s']:10.1f} tok/s aggregate  ({r['wall_s']:.2f} s, efficiency vs 1 GPU "
                  f"{r['efficiency_vs_1gpu']:.3f}, p50 prefill {r['prefill_ms_p50']:.1f} ms, routed {r['routed']})")
        print("=" * 100)
        print(json.dumps({"config3": rows}))
        return
    if args.config5:
        import torch
        r = config5(torch.cuda.device_count(), args.in_flight, args.new_tokens, args.pool_gib, args.policy)
        print("=" * 100)
        print(f"Config 5: duckdb-nsql-7B + Llama-3.2-3B co-served on {r['gpus']} GPU(s), one KV pool of {r['pool']['gib_per_gpu']:.0f} GiB "
              f"per GPU ({r['pool']['blocks_per_gpu']} blocks of {r['pool']['block_bytes'] >> 20} MiB), {r['in_flight_per_model']} sequences "
              f"per model in flight ({r['lanes_per_engine']} lanes per engine), {r['new_tokens_each']} new tokens each")
        for m, v in r["models"].items():
            print(f"  {m:12s}: {v['tok_s']:10.1f} tok/s aggregate  ({v['requests']} requests, {v['prompt_ids_total']} prompt ids, "
                  f"{v['wall_s']:.2f} s, {v['slots_per_block']} page(s) per block, routed {v['routed']})")
        print(f"  pool: peak {max(r['pool']['peak_blocks_in_use_per_gpu'])} blocks in use on the fullest GPU = "
              f"{100 * r['pool']['peak_occupancy']:.1f} % of the pool; all blocks free afterwards: "
              f"{all(a == r['pool']['blocks_per_gpu'] for a in r['pool']['after'])}")
        print("=" * 100)
        print(json.dumps({"config5": r}))
        return
    import torch
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl" if args.tp7b else "gloo")   # (the tensor-parallel leg synchronises through device tensors)
    import bench
    if args.tp7b:
        if world < 2:
            raise SystemExit("--tp7b needs torchrun with 2, 4 or 8 ranks")
        r = bench.bench_tp_7b(rank, world, local)
        if rank == 0:
            print("=" * 100)
            print(f"Config 4: duckdb-nsql-7B, tensor parallel {world}, {r['prompt_ids']} prompt ids, {r['new_tokens']} new tokens (one sequence)")
            print(f"  {r['ms_per_step']:.3f} ms/step = {r['tok_s']:.1f} tok/s  TTFT {r['ttft_ms']:.1f} ms  "
                  f"{r['achieved_gbs_per_gpu']:.0f} GB/s per GPU = {100 * r['frac_per_gpu']:.1f}% of the HBM roofline  "
                  f"one [1, d] all-reduce {r['allreduce_us']:.1f} us (share of the step {r['allreduce_share']:.2f})  "
                  f"{r['gpu_launches']} launch per decode call")
            print("=" * 100)
            print(json.dumps({"config4": r}))
        dist.destroy_process_group()
        return
    peak, peak_src = bench.peaks()
    report = {"n_gpus": world, "hbm_peak_gbs": peak, "peak_source": peak_src, "models": []}
    for model in args.models:
        entry = {"model": model}
        if not args.no_text and rank == 0:
            os.environ["B200LLM_DEVICE"] = str(local)
            entry["text"] = text_metrics(model, greedy=True)
            from b200llm import runtime
            runtime.unload_all()
            torch.cuda.empty_cache()
        if world > 1:
            dist.barrier()
        mine = engine_bench(model, args.batch, args.prompt_len, args.new_tokens, 4, local)
        if world > 1:
            box = [None] * world
            dist.all_gather_object(box, mine)
        else:
            box = [mine]
        rows = []
        for i, d in enumerate(mine["decode"]):
            ms = max(b["decode"][i]["ms_per_step"] for b in box)       # replicas: slowest rank sets the time
            tok_s = world * d["bs"] * 1000.0 / ms
            gbs = d["bytes_per_step"] / (ms * 1e-3) / 1e9
            rows.append({"bs": d["bs"], "ms_per_step": ms, "tok_s_total": tok_s, "gbs_per_gpu": gbs,
                         "pct_of_measured_peak": 100 * gbs / peak, "pct_of_8000": 100 * gbs / 8000.0})
        entry["gpu"] = {"ttft_ms_p50": statistics.median(b["ttft_ms_p50"] for b in box), "decode": rows,
                        "parallelism": f"{world} replica(s)"}
        if not args.no_cpu and rank == 0:
            entry["cpu"] = cpu_row(model)
        report["models"].append(entry)
    if rank == 0:
        print("=" * 100)
        for e in report["models"]:
            print(f"Model: {e['model']}   ({report['n_gpus']} GPU(s), roofline {peak:.0f} GB/s {peak_src})")
            t = e.get("text")
            if t and "error" not in t:
                print(f"  Exact Match Rate: {t['exact_match_rate']:.2f}%   Average Edit Distance: "
                      f"{t['avg_edit_distance']:.2f}   Average Latency: {t['avg_latency_s']:.4f} sec")
            elif t:
                print(f"  text metrics: {t['error']}")
            if "cpu" in e:
                print(f"  CPU ({e['cpu']['cores']} cores, oracle port): {e['cpu']['tok_s']:.2f} tok/s")
                q = e["cpu"].get("quantised")
                if q and "value" in q:
                    print(f"  CPU ({q['cores']} threads, {q['format']} restatement of llama.cpp math, {q['simd']}): {q['value']:.2f} tok/s "
                          f"({q['weight_gb_per_step']:.2f} GB of quantised weights per step)")
            print(f"  p50 TTFT ({args.prompt_len} ids): {e['gpu']['ttft_ms_p50']:.2f} ms")
            for r in e["gpu"]["decode"]:
                print(f"  bs {r['bs']:3d}: {r['tok_s_total']:10.1f} tok/s  {r['ms_per_step']:.3f} ms/step  "
                      f"{r['gbs_per_gpu']:.0f} GB/s = {r['pct_of_measured_peak']:.1f}% of measured, "
                      f"{r['pct_of_8000']:.1f}% of 8000")
            print("=" * 100)
        print(json.dumps(report))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
