#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on its config: decode tok/s of Llama-3.2-3B at batch 1
(296 synthetic prompt ids, greedy, seeded random-init weights).

  python bench.py --gpus N --steps K --warmup W          one JSON line (rank 0)
  python bench.py --impl reference ...                    the reference's CPU path (oracle port) timed
                                                          on the host cores, same metric/config

A "step" is one decode step (one new token for every sequence of the batch).  N > 1 runs N
independent replicas (request-level data parallelism, no data-path collective; weak scaling);
the timed region is bracketed by barrier + synchronize and the max over ranks is reported.
Inputs larger than L2: every step streams the 6.4 GB of weights (L2 is 126 MB).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "decode_tok_s_llama3.2-3b_bs1"
UNIT = "tok/s"
PROMPT_LEN = 296          # 20 system + 25 wrapper + ~250 Spark-error tokens
E2E_NEW_TOKENS = 128
NCU_DRAM_BYTES_PER_STEP = (19.349388e9 + 7.412224e6) / 3   # 3-step launch under ncu --set full
NCU_SOURCE = "profiles/tcd_full_r2e_summary.txt: 19.35 GB for a 3-step launch"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=512)
    ap.add_argument("--warmup", type=int, default=16)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="llama3.2")
    ap.add_argument("--bs", type=int, default=1)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graphs", action="store_true")
    ap.add_argument("--pdl", action="store_true")
    ap.add_argument("--tp", action="store_true", help="tensor parallel over the --gpus ranks (one sequence, strong scaling; "
                    "the 7B configuration) instead of one replica per rank")
    ap.add_argument("--max-rows", type=int, default=256, help="rows per forward pass (prefill chunk size / batch limit)")
    ap.add_argument("--no-mega", action="store_true", help="multi-kernel CUDA-graph path instead of the persistent kernel")
    ap.add_argument("--attn-splits", type=int, default=0)
    ap.add_argument("--no-tp-leg", action="store_true", help="skip the tp_7b leg (duckdb-nsql-7B tensor parallel over the --gpus ranks, config 4)")
    ap.add_argument("--no-batch", action="store_true", help="skip the bs64 leg (Llama-3.2-1B, 64 sequences, config 2)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def workload_name(model, bs):
    return f"{model} decode bs{bs}, {PROMPT_LEN} synthetic prompt ids, greedy, seeded random-init bf16 weights"


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 20 ms while the GPU is under the bench load."""
    Q = ("index,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "20", "-i", str(index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def lines(self):
        try:
            return open(self.f.name).read().strip().splitlines()
        except Exception:
            return []

    def stop(self, load_from=0, region=(0, 0)):
        """Samples of the timed region when the sampler hit it (lines region[0]:region[1]); a region shorter than
        the sampling period is described by the samples of the identical load that ran right before it
        (lines load_from:region[0], the filler decode calls), and `window` says which."""
        if self.p is not None:
            self.p.terminate()
            try:
                self.p.wait(timeout=5)
            except Exception:
                self.p.kill()
        all_lines = self.lines()
        try:
            os.unlink(self.f.name)
        except Exception:
            pass

        def parse(ls):
            rows = [l.split(", ") for l in ls]
            return [r for r in rows if len(r) >= 7]
        rows, window = parse(all_lines[region[0]:region[1]]), "timed region"
        if not rows:
            rows = parse(all_lines[load_from:max(region[1], region[0])])
            window = "decode load from just before the timed region to its end (the region is shorter than one sampling period)"
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "window": "no samples"}
        sm = [float(r[1]) for r in rows]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].strip().lower() == "active" for r in rows)]
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": float(rows[0][2]), "reasons": reasons,
                "samples": len(rows), "window": window}


# ----------------------------------------------------------------------------- CPU legs (oracle)
def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except (AttributeError, OSError):
        return max(1, os.cpu_count() or 1)


def cpu_decode_tok_s(model, n_prompt, n_steps, n_warm):
    """The oracle (CPU restatement of the reference path; kind "port") timed on all host cores:
    bf16 weights and matmuls, fp32 accumulate, same seeded weights and prompt stream."""
    import torch
    from oracle import llama_oracle as lo
    from oracle import weightgen as wg
    # all host cores: torchrun exports OMP_NUM_THREADS=1 for multi-rank launches, and only rank 0 runs this leg
    torch.set_num_threads(host_threads())
    geo = lo.GEOMETRIES[model]
    sd = lo.make_weights(geo, 0xB200)
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16, compute_dtype=torch.bfloat16)
    del sd
    ids = wg.gen_tokens(n_prompt, 1, min(geo.vocab, 128000))
    st = orc.new_state()
    logits = orc.forward(torch.as_tensor(ids, dtype=torch.int64), st)
    tok = int(torch.argmax(logits))
    for _ in range(n_warm):
        tok = int(torch.argmax(orc.forward(torch.tensor([tok]), st)))
    t0 = time.perf_counter()
    for _ in range(n_steps):
        tok = int(torch.argmax(orc.forward(torch.tensor([tok]), st)))
    dt = time.perf_counter() - t0
    return n_steps / dt, dt, torch.get_num_threads()


def cpu_quantised_tok_s(model, n_steps=12):
    """The closer CPU stand-in: llama.cpp's quantised math restated in oracle/qgemv.c -- every
    weight matrix of one decode step as a Q4_K x Q8_K (llama3.2) or Q4_0 x Q8_0 (duckdb-nsql) GEMV, AVX-512 VNNI when the
    host has it, OpenMP over rows, thread count tuned -- on synthetic quantised blocks of the model's geometry.  A restatement
    of llama.cpp's arithmetic, not llama.cpp; None when the helper library is not built."""
    try:
        from oracle import qgemv
        from oracle import llama_oracle as lo
        if not qgemv.available() or model not in qgemv.MODEL_FORMAT:
            return None
        fmt = qgemv.MODEL_FORMAT[model]
        v, secs, threads, nbytes = qgemv.decode_tok_s(lo.GEOMETRIES[model], fmt, n_steps, 2)
        return {"value": v, "unit": UNIT, "cores": threads, "kind": "restatement of llama.cpp math, not llama.cpp",
                "format": {"q4_K": "Q4_K x Q8_K", "q4_0": "Q4_0 x Q8_0"}[fmt], "simd": qgemv.simd(),
                "weight_gb_per_step": nbytes / 1e9, "achieved_gbs": nbytes * v / 1e9,
                "sample": f"{n_steps} decode steps: all weight matrices of the geometry as quantised GEMVs on synthetic blocks "
                          f"(activations re-quantised per matrix, SwiGLU; attention and norms left out)"}
    except Exception as e:  # noqa: BLE001 -- a baseline leg must not take the bench line down
        return {"error": str(e)[:200]}


def run_reference(args):
    """The reference's CPU path for the same config: the full PROMPT_LEN-id prompt is prefilled (untimed, like the
    B200 arm's), then `warmup` + `steps` decode steps at the same context; a step is bounded work (one token)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    warm = max(3, args.warmup)
    v, dt, cores = cpu_decode_tok_s(args.model, PROMPT_LEN, args.steps, warm)
    sample = (f"{args.steps} decode steps at bs1 after the {PROMPT_LEN}-id prompt and {warm} warm-up steps, "
              f"torch CPU bf16, {cores} threads")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": warm, "ms_per_step": 1000.0 * dt / args.steps, "higher_is_better": True,
            "scaling": "strong" if (args.tp and args.gpus > 1) else "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(args.model, 1), "parallelism": "host cores", "l2": "inputs larger than L2 (weights streamed every step)"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                             "quantised": cpu_quantised_tok_s(args.model)},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "reference path (Ollama + llama.cpp) is not in /root/reference and cannot be built here; "
                    "this is the oracle port of it"}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- B200 arm
def _note(msg):
    if os.environ.get("B200LLM_BENCH_VERBOSE"):
        print(f"[bench {time.time():.1f}] {msg}", file=sys.stderr, flush=True)


SYSTEM_PROMPT = "You are an AI that helps troubleshoot Apache Spark errors. Provide clear, concise solutions."   # Flask/app.py:162


def reference_style_prompts(n, n_ids):
    """(system, prompt) pairs shaped like the reference's error-path call (Flask/app.py:154-166): the constant
    system prompt, the constant wrapper and a varying Spark error, sized so that the byte stand-in tokenizer
    (random-init models carry no vocabulary) yields exactly n_ids prompt ids."""
    out = []
    for i in range(n):
        head, tail = "The following Spark error occurred:\n\n", "\n\nPlease analyze this error and suggest possible solutions."
        room = n_ids - 1 - len(SYSTEM_PROMPT) - 2 - len(head) - len(tail)          # BOS + system + blank line + wrapper
        err = (f"[UNRESOLVED_COLUMN.WITH_SUGGESTION #{i}] A column or function parameter with name `col_{i}` cannot be "
               f"resolved. Did you mean one of the following? " + "x" * n_ids)[:room]
        out.append((SYSTEM_PROMPT, head + err + tail))
    return out


def prefill_flops(geo, T):
    """F(T) = 2 N_read T + 4 L n_q head_dim T^2 / 2 (causal QK^T + PV)."""
    return 2.0 * (geo.weight_bytes_per_step() / 2) * T + 4.0 * geo.n_layers * geo.n_q * geo.head_dim * T * T / 2.0


def sustained_tflops():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"])
    except Exception:
        return 1389.8


def bench_batch(model, bs, steps, device):
    """Config 2: `bs` sequences of PROMPT_LEN ids decoded as one batch on one GPU."""
    import numpy as np
    import torch
    from b200llm import Engine, geometry, weights
    geo = geometry.resolve(model)
    w = weights.random_init(geo, 0xB200, device)
    warm = 3
    eng = Engine(geo, w, device=device, max_lanes=bs, max_ctx=PROMPT_LEN + steps + warm + 32, max_rows=max(bs, 256))
    lanes = [eng.open_lane() for _ in range(bs)]
    for i, ln in enumerate(lanes):
        eng.prefill(ln, np.random.default_rng(100 + i).integers(0, min(geo.vocab, 128000), PROMPT_LEN, dtype=np.int32))
    eng.decode(lanes, warm, read=False)
    for ln in lanes:
        eng._ensure_pages(ln, eng.pos[ln] + steps)
    eng.sync()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(eng.stream)
    eng.decode(lanes, steps, read=False)
    ev1.record(eng.stream)
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / steps
    launches = eng.timing().kernel_launches
    mean_ctx = PROMPT_LEN + warm + (steps - 1) / 2.0
    gbs = geo.step_bytes(bs, mean_ctx) / (ms * 1e-3) / 1e9
    eng.close()
    del eng, w
    torch.cuda.empty_cache()
    return {"model": model, "bs": bs, "tok_s": bs / (ms * 1e-3), "ms_per_step": ms, "frac": gbs / peaks()[0], "achieved_gbs": gbs,
            "steps": steps, "mean_ctx": mean_ctx, "gpu_launches": int(launches)}


def bench_tp_7b(rank, world, local, n_new=64):
    """Config 4: duckdb-nsql-7B, 2048 + 24 synthetic prompt ids, 64 new tokens, tensor parallel over all
    `world` ranks of the node (one sequence, strong scaling).  Collective: every rank calls it."""
    import ctypes as C
    import numpy as np
    import torch
    import torch.distributed as dist
    from b200llm import Engine, _lib, geometry, tp, weights
    geo = geometry.resolve("duckdb-nsql")
    n_prompt = 2048 + 24
    w = weights.random_init(geo, 0xB200, local, tp_rank=rank, tp_size=world)
    eng = Engine(geo, w, device=local, max_lanes=2, max_ctx=n_prompt + n_new + 64, max_rows=256, tp_rank=rank, tp_size=world,
                 nccl_id=tp.exchange_unique_id())
    tp.connect_peers(eng)
    prompt = np.random.default_rng(4).integers(0, geo.vocab, n_prompt, dtype=np.int32)      # the same stream on every rank
    lane = eng.open_lane()
    eng.prefill(lane, prompt)
    eng.sync()
    eng.prefill(lane, prompt)            # timed warm (the first pass pays tensor-map creation / connection set-up)
    eng.sync()
    ttft = eng.timing().prefill_ms
    eng.decode([lane], 8, read=False)    # warm-up
    eng.set_lane(lane, int(prompt[0]), n_prompt)
    eng._ensure_pages(lane, n_prompt + n_new)
    eng.sync()
    dist.barrier()
    torch.cuda.synchronize()
    eng.decode([lane], n_new, read=False)
    eng.sync()
    tm = eng.timing()
    us = C.c_float()
    _lib.check(_lib.lib.b200llm_debug_tp_allreduce_us(eng._h, 1, 256, C.byref(us)))
    t = torch.tensor([tm.decode_ms / n_new, ttft, us.value], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ttft, ar_us = float(t[0]), float(t[1]), float(t[2])
    mean_ctx = n_prompt + (n_new - 1) / 2.0
    per_gpu_gbs = geo.step_bytes(1, mean_ctx) / world / (ms * 1e-3) / 1e9
    fl = prefill_flops(geo, n_prompt)
    out = {"model": "duckdb-nsql", "tp": world, "prompt_ids": n_prompt, "new_tokens": n_new, "ms_per_step": ms, "tok_s": 1000.0 / ms,
           "ttft_ms": ttft, "ttft_frac_of_sustained": fl / (ttft * 1e-3) / 1e12 / sustained_tflops() / world,
           "frac_per_gpu": per_gpu_gbs / peaks()[0], "achieved_gbs_per_gpu": per_gpu_gbs,
           "allreduce_us": ar_us, "allreduce_share": 2 * geo.n_layers * ar_us * 1e-3 / ms,
           "allreduce_share_what": "2 x n_layers reductions per token x the time of one [1, d] fp32 all-reduce through the engine's "
                                   "peer-memory kernel, measured back to back in this run, over the step time (an estimate: inside the "
                                   "persistent kernel the reduction is a phase between two grid barriers)",
           "gpu_launches": int(tm.kernel_launches)}
    eng.close()
    del eng, w
    torch.cuda.empty_cache()
    return out


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from b200llm import Engine, SamplingParams, geometry, weights

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    geo = geometry.resolve(args.model)
    bs = args.bs
    tp_mode = args.tp and world > 1
    warm = max(3, args.warmup)
    max_ctx = PROMPT_LEN + args.steps + warm + E2E_NEW_TOKENS + 64
    # default flags: the engine the `ollama` shim serves from (b200llm.runtime), so that the device-timed steps and the
    # end-to-end calls run on the same engine a drop-in user gets; otherwise an engine built from the flags
    via_runtime = not (tp_mode or args.no_mega or args.no_graphs or args.pdl or args.attn_splits or args.max_rows != 256)
    model_rt = None
    if tp_mode:
        from b200llm import tp
        w = weights.random_init(geo, 0xB200, local, tp_rank=rank, tp_size=world)
        eng = Engine(geo, w, device=local, max_lanes=bs + 1, max_ctx=max_ctx, max_rows=max(bs, args.max_rows),
                     tp_rank=rank, tp_size=world, nccl_id=tp.exchange_unique_id(), use_graphs=not args.no_graphs,
                     attn_splits=args.attn_splits, use_megakernel=not args.no_mega)
        tp.connect_peers(eng)       # all-reduce over NVLink peer memory; persistent kernel per rank
    elif via_runtime:
        from b200llm import runtime
        os.environ["B200LLM_DEVICE"] = str(local)
        os.environ["B200LLM_MAX_LANES"] = str(max(8, bs + 1))
        os.environ["B200LLM_NUM_CTX"] = str(max(4096, max_ctx))
        model_rt = runtime.get_model(args.model)
        eng = model_rt.engine
    else:
        w = weights.random_init(geo, 0xB200, local)
        eng = Engine(geo, w, device=local, max_lanes=bs + 1, max_ctx=max_ctx, max_rows=max(bs, args.max_rows),
                     use_graphs=not args.no_graphs, use_pdl=args.pdl, attn_splits=args.attn_splits,
                     use_megakernel=not args.no_mega)
    # synthetic prompts: uniform ids, one stream per (rank, sequence); tensor parallel: the same stream on every rank
    rng = np.random.default_rng(1 if tp_mode else 1000 * rank + 1)
    prompts = [rng.integers(0, min(geo.vocab, 128000), PROMPT_LEN, dtype=np.int32) for _ in range(bs)]
    lanes = [eng.open_lane() for _ in range(bs)]
    for ln, p in zip(lanes, prompts):
        eng.prefill(ln, p)
    eng.sync()
    eng.prefill(lanes[0], prompts[0])   # timed warm (the first pass pays tensor-map creation / NCCL connection set-up)
    eng.sync()
    ttft_ms = eng.timing().prefill_ms
    _note("prefill done")

    sampler = ClockSampler(local) if rank == 0 else None
    # warm-up: W untimed steps (at least 3), then keep the GPU under the same load until nvidia-smi has sampled it
    eng.decode(lanes, warm, read=False)
    eng.sync()
    _note("warm-up done")
    load_from = len(sampler.lines()) if sampler is not None else 0
    if tp_mode:
        # tensor parallel: every rank must make exactly the same engine calls (their kernels wait for each other),
        # so the filler is a fixed number of calls
        for _ in range(24):
            for ln in lanes:
                eng.set_lane(ln, 1, PROMPT_LEN)
            eng.decode(lanes, 32, read=False)
            eng.sync()
    elif sampler is not None and sampler.p is not None:
        t_wait = time.time()
        while len(sampler.lines()) < load_from + 3 and time.time() - t_wait < 4.0:
            for ln in lanes:
                eng.set_lane(ln, 1, PROMPT_LEN)        # rewind so the context stays at the config's length
            eng.decode(lanes, 32, read=False)
            eng.sync()
        _note("clock-sampler filler decode done")
    for i, ln in enumerate(lanes):                     # timed region starts at the same context on every run
        eng.set_lane(ln, int(prompts[i][0]), PROMPT_LEN)
        eng._ensure_pages(ln, PROMPT_LEN + args.steps)  # page tables complete before the region: no allocator work inside it
    eng.sync()

    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    profiling = os.environ.get("B200LLM_PROFILE") == "1"      # ncu --profile-from-start off
    if profiling:
        torch.cuda.profiler.start()
    region0 = len(sampler.lines()) if sampler is not None else 0
    ev0.record(eng.stream)
    eng.decode(lanes, args.steps, read=False)
    ev1.record(eng.stream)
    torch.cuda.synchronize()
    region1 = len(sampler.lines()) if sampler is not None else 0
    _note("timed region done")
    if profiling:
        torch.cuda.profiler.stop()
    if world > 1:
        dist.barrier()
    ms_outer = ev0.elapsed_time(ev1)
    tm = eng.timing()
    launches, ms_engine = tm.kernel_launches, tm.decode_ms     # the engine's own events, recorded around its launch(es)
    clocks = sampler.stop(load_from, (region0, region1)) if sampler is not None else None
    t = torch.tensor([ms_outer, ms_engine], dtype=torch.float64, device="cuda")
    per_rank = [t.clone() for _ in range(world)]
    if world > 1:
        dist.all_gather(per_rank, t)
    else:
        per_rank = [t]
    per_rank = [[float(x[0]), float(x[1])] for x in per_rank]
    ms = max(x[0] for x in per_rank)
    toks = eng.read_tokens(args.steps, bs)
    assert toks.shape == (args.steps, bs) and int(toks.min()) >= 0
    for ln in lanes:
        eng.close_lane(ln)
    # ---- end to end: the public call, host buffers in and out, copies and prefill inside the timed region ----
    def timed_calls(fn, n_calls):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n_tok = sum(fn(i) for i in range(n_calls))
        torch.cuda.synchronize()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        return (1 if tp_mode else world) * n_tok / float(te.item()), n_tok

    e2e_runs = 3
    n_pages = (PROMPT_LEN + E2E_NEW_TOKENS + 15) // 16
    h2d = 3 * 4 * PROMPT_LEN + 4 * n_pages + 24 + 4 + 4          # ids/pos/lane arrays, page table, sampling, lane ids
    d2h = 4 * E2E_NEW_TOKENS + 8
    eng_val, _ = timed_calls(lambda i: len(eng.generate_ids(prompts[0], E2E_NEW_TOKENS, SamplingParams(temperature=0.0))[0]),
                             e2e_runs)
    e2e_engine = {"value": eng_val, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                  "what": f"Engine.generate_ids: {PROMPT_LEN} host prompt ids -> prefill -> {E2E_NEW_TOKENS} greedy tokens -> host, "
                          f"x{e2e_runs}; tokens / wall time incl. prefill"}
    e2e, e2e_sampled = e2e_engine, None
    if model_rt is not None:
        import ollama
        texts = reference_style_prompts(e2e_runs + 1, PROMPT_LEN)
        stats = {}

        def call(i, options, key):
            res = ollama.generate(model=args.model, system=texts[i][0], prompt=texts[i][1], options=options)
            st = stats.setdefault(key, {"eval_count": 0, "eval_ns": 0, "prompt_ids": res.prompt_eval_count})
            st["eval_count"] += res.eval_count
            st["eval_ns"] += res.eval_duration
            return res.eval_count
        greedy_opts = {"temperature": 0, "num_predict": E2E_NEW_TOKENS}
        call(e2e_runs, greedy_opts, "warm")           # the first request of a process pays the prefix-cache miss of the system prompt
        call(e2e_runs, None, "warm")
        g_val, _ = timed_calls(lambda i: call(i, greedy_opts, "greedy"), e2e_runs)
        s_val, _ = timed_calls(lambda i: call(i, None, "sampled"), e2e_runs)
        text_bytes = len(texts[0][0]) + len(texts[0][1])
        h2d_o = 4 * PROMPT_LEN * 3 + 4 * n_pages + 24 + 8       # what the engine copies for a request of this size (upper bound: no prefix hit)
        what = (f"ollama.generate(model, system, prompt[, options]) through the shim: template + tokenizer ({text_bytes} bytes of "
                f"text -> {stats['greedy']['prompt_ids']} ids) -> scheduler thread -> prefix cache -> prefill -> decode in 16-step "
                f"calls with stop-token check -> detokenise; x{e2e_runs} different error texts with the reference's constant system "
                f"prompt; generated tokens / wall time")
        e2e = {"value": g_val, "unit": UNIT, "h2d_bytes_per_step": h2d_o, "d2h_bytes_per_step": d2h,
               "options": greedy_opts, "decode_tok_s": stats["greedy"]["eval_count"] / (stats["greedy"]["eval_ns"] * 1e-9),
               "what": what}
        e2e_sampled = {"value": s_val, "unit": UNIT, "h2d_bytes_per_step": h2d_o, "d2h_bytes_per_step": d2h,
                       "options": "none (the reference's call: Ollama defaults temperature 0.8, top_k 40, top_p 0.9, "
                                  "repeat_penalty 1.1, random seed)",
                       "decode_tok_s": stats["sampled"]["eval_count"] / (stats["sampled"]["eval_ns"] * 1e-9), "what": what}

    # four callers at once, as the reference's threaded Flask server produces them (app.run, Flask/app.py:239, on the error
    # path :151-166): the scheduler batches their lanes and one persistent-kernel launch decodes all four
    e2e_4 = None
    if model_rt is not None and rank == 0 and world == 1:
        import threading
        texts4 = reference_style_prompts(8, PROMPT_LEN)
        counts = [0] * 4

        def caller(j, base):
            r = ollama.generate(model=args.model, system=texts4[base + j][0], prompt=texts4[base + j][1], options=None)
            counts[j] = r.eval_count

        def round4(base):
            ts = [threading.Thread(target=caller, args=(j, base)) for j in range(4)]
            for t in ts:
                t.start()
            for t in ts:
                t.join()
            return sum(counts)
        round4(0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n4 = round4(4)
        dt4 = time.perf_counter() - t0
        e2e_4 = {"value": n4 / dt4, "unit": UNIT, "callers": 4, "options": "none (Ollama's default sampling)",
                 "what": "4 threads call ollama.generate at once with different error texts; generated tokens of all four / wall time"}

    tp_7b = None
    if world > 1 and not tp_mode and args.model == "llama3.2" and bs == 1 and not args.no_tp_leg:
        if model_rt is not None:
            from b200llm import runtime
            runtime.unload_all()
            model_rt = None
        else:
            eng.close()
        torch.cuda.empty_cache()
        tp_7b = bench_tp_7b(rank, world, local)

    bs64 = bs4 = None
    if rank == 0 and world == 1 and bs == 1 and args.model == "llama3.2" and not args.no_batch:
        if model_rt is not None:
            from b200llm import runtime
            runtime.unload_all()
            model_rt = None
        bs64 = bench_batch("llama3.2:1b", 64, max(8, min(args.steps, 64)), local)
        # 4 concurrent callers of the headline model: one launch of the persistent kernel decodes all four lanes
        bs4 = bench_batch("llama3.2", 4, max(8, min(args.steps, 64)), local)

    if rank == 0:
        mean_ctx = PROMPT_LEN + (args.steps - 1) / 2.0
        step_bytes = geo.step_bytes(bs, mean_ctx)
        peak, peak_src = peaks()
        achieved = step_bytes / (ms / args.steps * 1e-3) / 1e9 / (world if tp_mode else 1)   # per GPU
        fl = prefill_flops(geo, PROMPT_LEN)
        line = {
            "metric": METRIC if (args.model == "llama3.2" and bs == 1) else f"decode_tok_s_{args.model}_bs{bs}",
            "value": (1 if tp_mode else world) * bs * args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong" if tp_mode else "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(args.model, bs), "parallelism": f"tp{world}" if tp_mode else f"replicas x{world}",
                       "l2": "inputs larger than L2 (weights streamed every step)", "cuda_graph": not args.no_graphs,
                       "pdl": args.pdl, "persistent_kernel": not args.no_mega, "kv_cache": "bf16 paged (16-token pages)", "mean_ctx": mean_ctx,
                       "engine": "b200llm.runtime (the ollama shim's engine)" if via_runtime else "built from the flags"},
            "ms_per_step_engine": max(x[1] for x in per_rank) / args.steps,
            "ms_per_rank": {"outer_events": [x[0] / args.steps for x in per_rank],
                            "engine_events": [x[1] / args.steps for x in per_rank],
                            "what": "ms per step of every rank: CUDA events recorded by bench.py around Engine.decode "
                                    "(= ms_per_step, max over ranks) and by the engine around its kernel launch"},
            "ttft_ms": ttft_ms,
            "ttft": {"ms": ttft_ms, "prompt_ids": PROMPT_LEN, "tflops": fl / (ttft_ms * 1e-3) / 1e12,
                     "frac_of_sustained": fl / (ttft_ms * 1e-3) / 1e12 / sustained_tflops() / (world if tp_mode else 1)},
            "clocks": clocks,
            "gpu_launches": int(launches),
            "e2e": e2e,
            "e2e_engine": e2e_engine,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": (NCU_DRAM_BYTES_PER_STEP * args.steps
                                     if (args.model == "llama3.2" and bs == 1 and not args.no_mega) else None),
                         "peak_source": peak_src,
                         "kernel": "tcd_decode_kernel<128,3> (tcgen05 persistent kernel, one launch = all timed steps)"
                                   if (bs == 1 and not args.no_mega) else "CUDA graph of the per-phase kernels",
                         "what": "algorithmic bytes B(bs, mean ctx) of SURVEY.md 8(d) x steps / device time of the "
                                 "timed region; traffic = dram read+write bytes of the same kernel under ncu "
                                 f"({NCU_SOURCE}), scaled to the steps of this launch",
                         "bytes_per_launch": step_bytes * args.steps, "bytes_per_step": step_bytes},
        }
        if e2e_sampled is not None:
            line["e2e_sampled"] = e2e_sampled
        if e2e_4 is not None:
            line["e2e_4callers"] = e2e_4
        if bs64 is not None:
            line["bs64"] = bs64
        if bs4 is not None:
            line["bs4"] = bs4
        if tp_7b is not None:
            line["tp_7b"] = tp_7b
        if not args.no_cpu_baseline and world == 1:
            v, dt, cores = cpu_decode_tok_s(args.model, 64, 24, 2)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "24 decode steps at bs1 after a 64-id prompt, oracle port in torch CPU bf16",
                                    "quantised": cpu_quantised_tok_s(args.model)}
        print(json.dumps(line), flush=True)
    if model_rt is not None:
        from b200llm import runtime
        runtime.unload_all()
    elif not via_runtime and tp_7b is None:
        eng.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
