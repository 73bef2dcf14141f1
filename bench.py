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
NCU_DRAM_BYTES_PER_STEP = (19.343379e9 + 7.5648e6) / 3   # profiles/tcd_full_r1e_summary.txt (3-step launch)


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
    """nvidia-smi clocks/throttle reasons sampled while the GPU is under the bench load."""
    Q = ("index,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100", "-i", str(index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def lines(self):
        try:
            return open(self.f.name).read().strip().splitlines()
        except Exception:
            return []

    def stop(self, first_line=0):
        if self.p is not None:
            self.p.terminate()
            try:
                self.p.wait(timeout=5)
            except Exception:
                self.p.kill()
        rows = [l.split(", ") for l in self.lines()[first_line:]]
        rows = [r for r in rows if len(r) >= 7]
        try:
            os.unlink(self.f.name)
        except Exception:
            pass
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = [float(r[1]) for r in rows]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].strip().lower() == "active" for r in rows)]
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": float(rows[0][2]), "reasons": reasons,
                "samples": len(rows)}


# ----------------------------------------------------------------------------- CPU legs (oracle)
def cpu_decode_tok_s(model, n_prompt, n_steps, n_warm):
    """The oracle (CPU restatement of the reference path; kind "port") timed on all host cores:
    bf16 weights and matmuls, fp32 accumulate, same seeded weights and prompt stream."""
    import torch
    from oracle import llama_oracle as lo
    from oracle import weightgen as wg
    # all host cores: torchrun exports OMP_NUM_THREADS=1 for multi-rank launches, and only rank 0 runs this leg
    try:
        torch.set_num_threads(max(1, len(os.sched_getaffinity(0))))
    except (AttributeError, OSError):
        torch.set_num_threads(max(1, os.cpu_count() or 1))
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


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_prompt = 64
    v, dt, cores = cpu_decode_tok_s(args.model, n_prompt, args.steps, min(args.warmup, 4))
    sample = (f"{args.steps} decode steps at bs1 after a {n_prompt}-id prompt (bounded sample of the "
              f"{PROMPT_LEN}-id workload), torch CPU bf16, {cores} threads")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": min(args.warmup, 4), "ms_per_step": 1000.0 * dt / args.steps, "higher_is_better": True,
            "scaling": "strong" if (args.tp and args.gpus > 1) else "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(args.model, 1), "l2": "inputs larger than L2"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "reference path (Ollama + llama.cpp) is not in /root/reference and cannot be built here; "
                    "this is the oracle port of it"}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- B200 arm
def _note(msg):
    if os.environ.get("B200LLM_BENCH_VERBOSE"):
        print(f"[bench {time.time():.1f}] {msg}", file=sys.stderr, flush=True)


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
    max_ctx = PROMPT_LEN + args.steps + args.warmup + E2E_NEW_TOKENS + 64
    if tp_mode:
        from b200llm import tp
        w = weights.random_init(geo, 0xB200, local, tp_rank=rank, tp_size=world)
        eng = Engine(geo, w, device=local, max_lanes=bs + 1, max_ctx=max_ctx, max_rows=max(bs, args.max_rows),
                     tp_rank=rank, tp_size=world, nccl_id=tp.exchange_unique_id(), use_graphs=not args.no_graphs,
                     attn_splits=args.attn_splits, use_megakernel=not args.no_mega)
        tp.connect_peers(eng)       # all-reduce over NVLink peer memory; persistent kernel per rank
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
    if tp_mode:     # the first tensor-parallel pass pays NCCL's connection setup: time the prompt again, warm
        eng.prefill(lanes[0], prompts[0])
        eng.sync()
    ttft_ms = eng.timing().prefill_ms
    _note("prefill done")

    sampler = ClockSampler(local) if rank == 0 else None
    # warm-up: graph capture + W steps (at least 3), and keep the GPU loaded until nvidia-smi has sampled
    eng.decode(lanes, max(3, args.warmup), read=False)
    eng.sync()
    _note("warm-up done")
    first_line = 0
    if tp_mode:
        # tensor parallel: every rank must make exactly the same engine calls (their kernels wait for each other),
        # so the filler that keeps the GPU loaded until nvidia-smi has sampled is a fixed number of calls
        for _ in range(24):
            for ln in lanes:
                eng.set_lane(ln, 1, PROMPT_LEN)
            eng.decode(lanes, 32, read=False)
            eng.sync()
        if sampler is not None:
            first_line = max(0, len(sampler.lines()) - 2)
    elif sampler is not None and sampler.p is not None:
        t_wait = time.time()
        while len(sampler.lines()) < 2 and time.time() - t_wait < 4.0:
            for ln in lanes:
                eng.set_lane(ln, 1, PROMPT_LEN)        # rewind so the context stays at the config's length
            eng.decode(lanes, 32, read=False)
            eng.sync()
            _note("clock-sampler filler decode done")
        first_line = len(sampler.lines())
    for i, ln in enumerate(lanes):                     # timed region starts at the same context on every run
        eng.set_lane(ln, int(prompts[i][0]), PROMPT_LEN)
    eng.sync()

    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    profiling = os.environ.get("B200LLM_PROFILE") == "1"      # ncu --profile-from-start off
    if profiling:
        torch.cuda.profiler.start()
    ev0.record(eng.stream)
    eng.decode(lanes, args.steps, read=False)
    ev1.record(eng.stream)
    torch.cuda.synchronize()
    _note("timed region done")
    if profiling:
        torch.cuda.profiler.stop()
    if world > 1:
        dist.barrier()
    ms = ev0.elapsed_time(ev1)
    launches = eng.timing().kernel_launches
    clocks = sampler.stop(first_line) if sampler is not None else None
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    toks = eng.read_tokens(args.steps, bs)
    assert toks.shape == (args.steps, bs) and int(toks.min()) >= 0

    # ---- e2e: the public call (host ids in, host tokens out), copies inside the timed region ----
    for ln in lanes:
        eng.close_lane(ln)
    e2e_runs = 3
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n_e2e = 0
    for r in range(e2e_runs):
        out, tm = eng.generate_ids(prompts[0], E2E_NEW_TOKENS, SamplingParams(temperature=0.0))
        n_e2e += len(out)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_val = (1 if tp_mode else world) * n_e2e / float(te.item())
    n_pages = (PROMPT_LEN + E2E_NEW_TOKENS + 15) // 16
    h2d = 3 * 4 * PROMPT_LEN + 4 * n_pages + 24 + 4 + 4          # ids/pos/lane arrays, page table, sampling, lane ids
    d2h = 4 * E2E_NEW_TOKENS + 8

    if rank == 0:
        mean_ctx = PROMPT_LEN + (args.steps - 1) / 2.0
        step_bytes = geo.step_bytes(bs, mean_ctx)
        peak, peak_src = peaks()
        achieved = step_bytes / (ms / args.steps * 1e-3) / 1e9 / (world if tp_mode else 1)   # per GPU
        line = {
            "metric": METRIC if (args.model == "llama3.2" and bs == 1) else f"decode_tok_s_{args.model}_bs{bs}",
            "value": (1 if tp_mode else world) * bs * args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong" if tp_mode else "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(args.model, bs), "parallelism": f"tp{world}" if tp_mode else f"replicas x{world}",
                       "l2": "inputs larger than L2 (weights streamed every step)", "cuda_graph": not args.no_graphs,
                       "pdl": args.pdl, "persistent_kernel": not args.no_mega, "kv_cache": "bf16 paged (16-token pages)", "mean_ctx": mean_ctx},
            "ttft_ms": ttft_ms,
            "clocks": clocks,
            "gpu_launches": int(launches),
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "what": f"ollama-style generate: {PROMPT_LEN} host prompt ids -> prefill -> "
                            f"{E2E_NEW_TOKENS} greedy tokens -> host, x{e2e_runs}; tokens / wall time incl. prefill"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": (NCU_DRAM_BYTES_PER_STEP * args.steps
                                     if (args.model == "llama3.2" and bs == 1 and not args.no_mega) else None),
                         "peak_source": peak_src,
                         "kernel": "tcd_decode_kernel<128,3> (tcgen05 persistent kernel, one launch = all timed steps)"
                                   if (bs == 1 and not args.no_mega) else "CUDA graph of the per-phase kernels",
                         "what": "algorithmic bytes B(bs, mean ctx) of SURVEY.md 8(d) x steps / device time of the "
                                 "timed region; traffic = dram read+write bytes of the same kernel under ncu "
                                 "(profiles/tcd_full_r1e_summary.txt: 19.34 GB for a 3-step launch), scaled to "
                                 "the steps of this launch",
                         "bytes_per_launch": step_bytes * args.steps, "bytes_per_step": step_bytes},
        }
        if not args.no_cpu_baseline and world == 1:
            v, dt, cores = cpu_decode_tok_s(args.model, 64, 24, 2)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "24 decode steps at bs1 after a 64-id prompt, oracle port in torch CPU bf16"}
        print(json.dumps(line), flush=True)
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
