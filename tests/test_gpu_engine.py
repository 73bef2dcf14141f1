"""-m gpu: whole-model parity through the C ABI (b200llm_prefill / b200llm_decode /
b200llm_step_logits) against the CPU oracle.

Bar (BASELINE.json north_star): same seeded weights + same prompt ids => the first 64
greedy tokens identical (tie-break lowest id) and logits within 1e-2 relative (max|a-b| /
max|b|).  The asserted comparison is against the oracle in the engine's storage model
(fp32 activations, bf16 KV cache); the distance to the nothing-rounded fp32 oracle is
bounded by twice the oracle's own bf16-KV distance (on random-init weights that distance
alone is 0.7-2e-2, see DESIGN.md "Parity").
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import llama_oracle as lo          # noqa: E402
from oracle import weightgen as wg             # noqa: E402

LOGIT_TOL = 1e-2       # north_star tolerance
TIGHT_TOL = 3e-3       # what the fp32-activation path achieves vs the mirrored oracle (measured <= 1.4e-3:
                       # a bf16 KV element that rounds the other way than the oracle's costs 2^-9 on it)


def _engine(name, sd_cpu=None, seed=0xB200, sigma=0.02, **kw):
    from b200llm import Engine, geometry, weights
    geo = geometry.resolve(name)
    kw.setdefault("use_tc", False)      # fp32-activation path: what the kvbf16 goldens describe
    if sd_cpu is not None:
        w = weights.pack(geo, {k: v.cuda() for k, v in sd_cpu.items()})
    else:
        w = weights.random_init(geo, seed, 0, sigma=sigma)
    return Engine(geo, w, **kw)


def _rel(a, b):
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-9))


@pytest.mark.parametrize("name,use_graphs,use_pdl,mega", [("tiny", False, False, False), ("tiny", True, True, False),
                                                           ("small", True, True, False), ("small", True, False, False),
                                                           ("tiny", True, False, True), ("small", True, False, True),
                                                           ("tiny", True, False, "mma"), ("small", True, False, "mma")])
def test_live_oracle_parity(name, use_graphs, use_pdl, mega, golden_dir, monkeypatch):
    """Engine vs the oracle run live on this box's CPU, weights uploaded from the oracle's stream.
    mega: False = per-phase kernels, True = tcgen05 persistent kernel, "mma" = the mma.sync persistent
    kernel (what geometries that cannot be tiled 128 x 64 fall back to)."""
    if mega == "mma":
        monkeypatch.setenv("B200LLM_PERSISTENT", "mma")
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    geo = lo.GEOMETRIES[name]
    sd = lo.make_weights(geo, int(g["seed"]), sigma=float(g["sigma"]), embed_sigma=float(g["embed_sigma"]))
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    prompt = g["prompt"]
    toks, lg, mar = orc.greedy(prompt, 64)
    assert toks == g["tokens"].tolist()
    eng = _engine(name, sd, max_lanes=2, max_ctx=256, max_rows=16, use_graphs=use_graphs, use_pdl=use_pdl,
                  use_megakernel=bool(mega))
    lane = eng.open_lane()
    eng.prefill(lane, prompt)
    first, pos = eng.read_lane(lane)
    assert pos == len(prompt)
    l0 = eng.logits(0)
    assert _rel(l0, lg[0].numpy()) <= TIGHT_TOL
    assert first == toks[0]
    out = eng.decode([lane], 63)[:, 0].tolist()
    assert [first] + out == toks
    # the last step's logits (free-running) still match the oracle's
    assert _rel(eng.logits(0), lg[63].numpy()) <= TIGHT_TOL
    eng.close()


@pytest.mark.parametrize("name,vocab", [("tiny", 1000), ("small", 4000 + 16)])
def test_vocabulary_that_is_no_multiple_of_the_tile(name, vocab, monkeypatch):
    """CodeLlama-class checkpoints have 32016 vocabulary rows: the tile-major lm_head copy is padded with zero
    rows to whole 128-row tiles and the tcgen05 persistent kernel still runs (one launch), never storing or
    ranking the padding: tokens and logits equal the oracle's."""
    from dataclasses import replace
    from b200llm import Engine, geometry, weights
    ogeo = replace(lo.GEOMETRIES[name], vocab=vocab)
    sd = lo.make_weights(ogeo, 0xC0DE, sigma=0.08)
    orc = lo.LlamaOracle(ogeo, sd, kv_dtype=torch.bfloat16)
    prompt = wg.gen_tokens(21, 77, vocab)
    toks, lg, _ = orc.greedy(prompt, 24)
    geo = replace(geometry.resolve(name), vocab=vocab)
    pw = weights.pack(geo, {k: v.cuda() for k, v in sd.items()})
    eng = Engine(geo, pw, max_lanes=1, max_ctx=128, max_rows=8, use_tc=False, use_megakernel=True)
    lane = eng.open_lane()
    eng.prefill(lane, prompt)
    first, _ = eng.read_lane(lane)
    out = eng.decode([lane], 23)[:, 0].tolist()
    assert eng.timing().kernel_launches == 1          # the persistent kernel, not the per-phase fallback
    assert [first] + out == toks
    got = eng.logits(0)
    assert got.shape[0] == vocab and _rel(got, lg[23].numpy()) <= TIGHT_TOL
    eng.close()


@pytest.mark.parametrize("mega", [False, True])
def test_teacher_forced_logits_every_step(mega, golden_dir):
    """Per-step logits along the golden token path (b200llm_set_lane forces the input token)."""
    g = np.load(os.path.join(golden_dir, "small.npz"))
    geo = lo.GEOMETRIES["small"]
    sd = lo.make_weights(geo, int(g["seed"]), sigma=float(g["sigma"]), embed_sigma=float(g["embed_sigma"]))
    eng = _engine("small", sd, max_lanes=1, max_ctx=256, max_rows=8, use_megakernel=mega)
    lane = eng.open_lane()
    prompt, toks, probe = g["prompt"], g["tokens"], g["probe_ids"]
    eng.prefill(lane, prompt)
    worst = worst_truth = 0.0
    for s in range(64):
        l = eng.logits(0)
        worst = max(worst, float(np.abs(l[probe] - g["probe_kvbf16"][s]).max() / g["absmax_kvbf16"][s]))
        worst_truth = max(worst_truth, float(np.abs(l[probe] - g["probe_fp32"][s]).max() / g["absmax_fp32"][s]))
        assert int(np.argmax(l)) == int(toks[s])
        if s < 63:
            eng.set_lane(lane, int(toks[s]), len(prompt) + s)
            eng.decode([lane], 1)
    oracle_own = float((np.abs(g["probe_kvbf16"] - g["probe_fp32"]).max(1) / g["absmax_fp32"]).max())
    assert worst <= TIGHT_TOL
    assert worst_truth <= max(LOGIT_TOL, 2 * oracle_own)
    eng.close()


def test_chunked_prefill_ragged_and_batched_decode(golden_dir):
    """Prefill in chunks that straddle page boundaries; 3 lanes with different prompts decoded
    as one batch give the same tokens as each decoded alone."""
    g = np.load(os.path.join(golden_dir, "tiny.npz"))
    geo = lo.GEOMETRIES["tiny"]
    sd = lo.make_weights(geo, int(g["seed"]), sigma=float(g["sigma"]), embed_sigma=float(g["embed_sigma"]))
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    prompts = [g["prompt"], wg.gen_tokens(1, 77, geo.vocab), wg.gen_tokens(50, 78, geo.vocab)]
    want = [orc.greedy(p, 20)[0] for p in prompts]
    eng = _engine("tiny", sd, max_lanes=4, max_ctx=128, max_rows=7)      # 7-row chunks: 37 = 5*7 + 2
    lanes = [eng.open_lane() for _ in prompts]
    firsts = []
    for ln, p in zip(lanes, prompts):
        eng.prefill(ln, p)
        firsts.append(eng.read_lane(ln)[0])
    out = eng.decode(lanes, 19)
    for i in range(3):
        assert [firsts[i]] + out[:, i].tolist() == want[i]
    # a lane can be closed and its pages reused
    eng.close_lane(lanes[1])
    ln = eng.open_lane()
    eng.prefill(ln, prompts[2])
    assert eng.read_lane(ln)[0] == want[2][0]
    eng.close()


@pytest.mark.parametrize("mega,tc", [(False, False), (True, False), (True, True)])
@pytest.mark.parametrize("fname,name", [("llama3.2_1b", "llama3.2:1b"), ("llama3.2", "llama3.2"),
                                        ("duckdb-nsql", "duckdb-nsql")])
def test_golden_parity_full_size(fname, name, mega, tc, golden_dir):
    """BASELINE.json configs 1/2 geometries against the committed golden vectors (oracle run
    in the CPU container); weights come from the device twin of the seeded stream.  (True, True) is
    the production configuration: tcgen05 GEMM prefill in 256-row chunks + persistent decode kernel."""
    g = np.load(os.path.join(golden_dir, fname + ".npz"))
    eng = _engine(name, None, seed=int(g["seed"]), sigma=float(g["sigma"]), max_lanes=1,
                  max_ctx=len(g["prompt"]) + 128, max_rows=256 if tc else 64, use_megakernel=mega, use_tc=tc)
    lane = eng.open_lane()
    prompt, toks, probe = g["prompt"], g["tokens"].tolist(), g["probe_ids"]
    eng.prefill(lane, prompt)
    first, _ = eng.read_lane(lane)
    l0 = eng.logits(0)
    assert float(np.abs(l0[probe] - g["probe_kvbf16"][0]).max() / g["absmax_kvbf16"][0]) <= TIGHT_TOL
    got = [first] + eng.decode([lane], 63)[:, 0].tolist()
    margins = g["margins"]
    for s in range(64):
        if got[s] != toks[s]:
            # a mismatch is only tolerated at a near-tie of the oracle itself (flagged)
            assert margins[s] < 1e-3, f"step {s}: got {got[s]} want {toks[s]} margin {margins[s]}"
            eng.close()
            pytest.xfail(f"near-tie at step {s} (oracle margin {margins[s]:.2e}); tokens equal before it")
    assert got == toks
    l63 = eng.logits(0)
    d_first = float(np.abs(l0[probe] - g["probe_kvbf16"][0]).max() / g["absmax_kvbf16"][0])
    d_last = float(np.abs(l63[probe] - g["probe_kvbf16"][63]).max() / g["absmax_kvbf16"][63])
    assert d_last <= TIGHT_TOL
    assert float(np.abs(l63[g["top_ids"][63]] - g["top_vals"][63]).max() / g["absmax_kvbf16"][63]) <= TIGHT_TOL
    # distance to the nothing-rounded fp32 oracle (north_star: 1e-2), bounded by twice the oracle's own bf16-KV distance
    # where that alone exceeds it (DESIGN.md "Parity"); the measured distances go to the test log
    truth = max(float(np.abs(l0[probe] - g["probe_fp32"][0]).max() / g["absmax_fp32"][0]),
                float(np.abs(l63[probe] - g["probe_fp32"][63]).max() / g["absmax_fp32"][63]))
    oracle_own = float((np.abs(g["probe_kvbf16"] - g["probe_fp32"]).max(1) / g["absmax_fp32"]).max())
    print(f"{name} mega={mega} tc={tc}: logits vs the oracle in the engine's storage model {d_first:.1e} (step 0) / {d_last:.1e} "
          f"(step 63); vs the fp32 oracle {truth:.1e} (the oracle's own bf16-KV mode is {oracle_own:.1e} away from it)")
    assert truth <= max(LOGIT_TOL, 2 * oracle_own)
    eng.close()


@pytest.mark.parametrize("mega", [True, False])
def test_repeated_requests_on_one_engine(mega):
    """Request after request on one engine (64-row prefill chunks, then batch-1 decode, which
    use different attention split layouts of the same scratch) keeps matching the oracle."""
    from b200llm import SamplingParams
    geo = lo.GEOMETRIES["small"]
    sd = lo.make_weights(geo, 0xB200, sigma=0.06)
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    ids = wg.gen_tokens(87, 21, geo.vocab)
    want = orc.greedy(ids, 12)[0]
    eng = _engine("small", sd, max_lanes=4, max_ctx=1024, max_rows=64, use_megakernel=mega)
    for _ in range(3):
        toks, tm = eng.generate_ids(ids, 12, SamplingParams(temperature=0.0))
        assert toks == want
        assert tm.decode_steps == 11 and tm.prefill_tokens == 87
    eng.close()


@pytest.mark.parametrize("name,tile_major", [("tiny", True), ("small", True), ("tiny", False), ("small", False)])
def test_tensor_core_batch_path_matches_oracle(name, tile_major, golden_dir):
    """Prefill chunks and an 8-lane decode batch on the tcgen05 GEMM path.  Its activation operands
    are bf16 hi + lo pairs (both multiplied, ~fp32 accurate), so it is held to the same oracle and
    the same tolerance as the fp32-activation paths.  tile_major: the weight tiles come as 16 KB bulk
    copies out of the tile-major copy an engine with the persistent kernel holds; without it as 2-D
    tensor-map boxes out of the row-major matrix."""
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    geo = lo.GEOMETRIES[name]
    sd = lo.make_weights(geo, int(g["seed"]), sigma=float(g["sigma"]), embed_sigma=float(g["embed_sigma"]))
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    prompts = [wg.gen_tokens(20 + 9 * i, 300 + i, geo.vocab) for i in range(8)]
    eng = _engine(name, sd, max_lanes=8, max_ctx=256, max_rows=32, use_tc=True, use_megakernel=tile_major)
    lanes = [eng.open_lane() for _ in prompts]
    worst = 0.0
    for ln, p in zip(lanes, prompts):
        st = orc.new_state()
        ref = orc.forward(torch.as_tensor(p, dtype=torch.int64), st)
        eng.prefill(ln, p)                           # 20..83 tokens in chunks of 32 -> tensor-core passes
        eng.sync()
        worst = max(worst, _rel(eng.logits(0), ref.numpy()))
    assert worst <= TIGHT_TOL, worst
    # one batched decode step, teacher-forced: every lane gets the oracle's greedy token
    states, refs = [], []
    for i, p in enumerate(prompts):
        st = orc.new_state()
        l0 = orc.forward(torch.as_tensor(p, dtype=torch.int64), st)
        t0 = int(torch.argmax(l0))
        eng.set_lane(lanes[i], t0, len(p))
        refs.append(orc.forward(torch.tensor([t0]), st))
    eng.decode(lanes, 1)
    for i in range(8):
        assert _rel(eng.logits(i), refs[i].numpy()) <= TIGHT_TOL
    eng.close()


def test_tensor_core_path_with_fused_swiglu_epilogue():
    """A geometry whose gate/up matrix has >= 74 row tiles: the tcgen05 GEMM then runs without a k
    split and applies SwiGLU in its epilogue (bf16 straight into the down-projection operand), and
    the attention kernel writes its output as bf16 hi + lo.  Same oracle and tolerance as above."""
    from b200llm import Engine, geometry, weights
    og = lo.OracleGeometry("wide", 256, 2, 4, 2, 64, 4864, 1024, True, 5e5, True)      # 2 * ffn / 128 = 76 tiles
    geo = geometry.Geometry("wide", 256, 2, 4, 2, 64, 4864, 1024, True, 5e5, True)
    sd = lo.make_weights(og, 0xB200 + 7, sigma=0.05, embed_sigma=0.05)
    orc = lo.LlamaOracle(og, sd, kv_dtype=torch.bfloat16)
    eng = Engine(geo, weights.pack(geo, {k: v.cuda() for k, v in sd.items()}), max_lanes=4, max_ctx=256, max_rows=64,
                 use_tc=True, use_megakernel=False)
    prompts = [wg.gen_tokens(37 + 11 * i, 900 + i, geo.vocab) for i in range(4)]
    lanes = [eng.open_lane() for _ in prompts]
    for ln, p in zip(lanes, prompts):
        ref = orc.forward(torch.as_tensor(p, dtype=torch.int64), orc.new_state())
        eng.prefill(ln, p)
        eng.sync()
        assert _rel(eng.logits(0), ref.numpy()) <= TIGHT_TOL
    eng.close()


def test_scheduler_batches_concurrent_callers():
    """Threaded callers (the Flask dev server case) share one engine through the scheduler: every
    request gets exactly the tokens it gets when run alone, whatever it was batched with."""
    import threading
    from b200llm import SamplingParams
    from b200llm.scheduler import Scheduler
    geo = lo.GEOMETRIES["tiny"]
    sd = lo.make_weights(geo, 0xB200, sigma=0.08)
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    prompts = [wg.gen_tokens(10 + 7 * i, 500 + i, geo.vocab) for i in range(6)]
    budgets = [9, 20, 5, 33, 14, 1]
    want = [orc.greedy(p, n)[0] for p, n in zip(prompts, budgets)]
    eng = _engine("tiny", sd, max_lanes=4, max_ctx=128, max_rows=8)      # 6 callers, 4 lanes: two must queue
    sched = Scheduler(eng, chunk=4)
    got = [None] * 6

    def call(i):
        got[i] = sched.generate_ids(prompts[i], budgets[i], SamplingParams(temperature=0.0))[0]

    threads = [threading.Thread(target=call, args=(i,)) for i in range(6)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=120)
    assert got == want
    sched.shutdown()
    eng.close()


def test_prefix_cache_shares_system_prompt_pages():
    """The reference sends one system prompt with every request (Flask/app.py:104,162; FastAPI/app.py:87,107).
    The scheduler's prefix cache attaches the KV pages of that shared prefix to later requests: same tokens as
    the oracle's full-prompt run, fewer tokens prefilled, pages returned to the pool when the cache is cleared."""
    from b200llm import SamplingParams
    from b200llm.scheduler import Scheduler
    geo = lo.GEOMETRIES["tiny"]
    sd = lo.make_weights(geo, 0xB200, sigma=0.08)
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    system = [int(t) for t in wg.gen_tokens(53, 900, geo.vocab)]     # 3 whole pages + 5 tokens
    users = [[int(t) for t in wg.gen_tokens(n, 910 + i, geo.vocab)] for i, n in enumerate([9, 1, 30, 12])]
    prompts = [system + u for u in users] + [system[:40] + users[0]]  # the last one shares only 2 pages
    want = [orc.greedy(p, 12)[0] for p in prompts]
    eng = _engine("tiny", sd, max_lanes=2, max_ctx=160, max_rows=8)
    free0 = len(eng._free_pages)
    sched = Scheduler(eng, chunk=4)
    greedy = SamplingParams(temperature=0.0)
    prefilled = []
    for p, w in zip(prompts, want):
        toks, timing = sched.generate_ids(p, 12, greedy)
        assert toks == w
        prefilled.append(timing.prefill_tokens)
    assert prefilled[0] == len(prompts[0])                           # cold
    assert prefilled[1:4] == [len(p) - 48 for p in prompts[1:4]]     # 3 shared pages
    assert prefilled[4] == len(prompts[4]) - 32                      # 2 shared pages
    assert sched.cache.hits == 4 and sched.cache.misses == 1
    # a repeat-penalty run sees the same 64-token window with and without the cache
    pen = SamplingParams(temperature=0.0, repeat_penalty=1.5)
    with_cache = sched.generate_ids(prompts[2], 12, pen)[0]
    sched.cache.clear()
    assert len(eng._free_pages) == free0 and all(r == 0 for r in eng._page_ref)
    cold = Scheduler(eng, chunk=4, prefix_cache=False)
    assert cold.generate_ids(prompts[2], 12, pen)[0] == with_cache
    cold.shutdown()
    sched.shutdown()
    eng.close()


def test_shim_replays_reference_call_sites():
    """The six ollama.generate call patterns of the reference, unmodified, against the shim."""
    import ollama
    # Flask/app.py:160-166 and FastAPI/app.py:105-111
    err = ("[UNRESOLVED_COLUMN.WITH_SUGGESTION] A column or function parameter with name `Location_of_company` "
           "cannot be resolved.")
    prompt_text = (f"The following Spark error occurred:\n\n{err}\n\n"
                   f"Please analyze this error and suggest possible solutions.")
    response = ollama.generate(
        model="tiny",
        system="You are an AI that helps troubleshoot Apache Spark errors. Provide clear, concise solutions.",
        prompt=prompt_text)
    assert isinstance(response.response, str) and response.done and response.eval_count >= 1
    assert response["prompt_eval_count"] > 100 and response.eval_duration >= 0
    # Flask/app.py:102-107 and FastAPI/app.py:85-90
    res = ollama.generate(model="small", system="Table name is temp_view. The structure of the table is:\nA (int)",
                          prompt="count rows", options={"temperature": 0, "num_predict": 12})
    res2 = ollama.generate(model="small", system="Table name is temp_view. The structure of the table is:\nA (int)",
                           prompt="count rows", options={"temperature": 0, "num_predict": 12})
    assert res.response == res2.response and res.eval_count == 12       # greedy is deterministic
    with pytest.raises(ollama.ResponseError) as ei:                      # Model_Eval...py:69 names "mistral"
        ollama.generate(model="mistral", system="s", prompt="p")
    assert ei.value.status_code == 404
    from b200llm import runtime
    runtime.unload_all()


# ---------------------------------------------------------------------------------------------------------
# Sampling: the reference's calls pass no `options`, so Ollama's defaults apply (temperature 0.8, top-k 40,
# top-p 0.9, repeat penalty 1.1).  The persistent kernel runs that chain as its last phase.
def _oracle_sampled(orc, prompt, n_new, sp):
    """Oracle forward + oracle/sampler.py, free-running.  Returns (tokens, per step: the penalised logits, slack,
    margin).  slack = distance of the uniform draw from the nearest edge of the kept candidates' cumulative mass,
    relative to that mass; margin = top-1 minus top-2 of the penalised logits."""
    from oracle import sampler
    st = orc.new_state()
    seq = [int(t) for t in prompt]
    logits = orc.forward(torch.as_tensor(seq, dtype=torch.int64), st)
    toks, pen, slack, margin = [], [], [], []
    for _ in range(n_new):
        ctr = len(seq) - 1                       # position of the token that was fed last
        lg = sampler.apply_penalty(logits.numpy(), sp.repeat_penalty, seq[-64:])
        tok, ids, probs = sampler.sample(logits.numpy(), sp.temperature, sp.top_k, sp.top_p, sp.repeat_penalty,
                                         sp.seed, ctr, history=seq[-64:])
        top2 = np.sort(lg)[-2:]
        margin.append(float(top2[1] - top2[0]))
        if probs is not None:
            mass = float(probs.sum())
            r = float(sampler.uniform(sp.seed, ctr)) * mass
            slack.append(float(np.abs(np.cumsum(probs.astype(np.float64)) - r).min() / mass))
        else:
            slack.append(1.0)
        pen.append(lg)
        toks.append(tok)
        seq.append(tok)
        if len(toks) < n_new:
            logits = orc.forward(torch.tensor([tok], dtype=torch.int64), st)
    return toks, pen, slack, margin


@pytest.mark.parametrize("name,sigma", [("tiny", 0.08), ("small", 0.06)])
@pytest.mark.parametrize("opts", [dict(temperature=0.8, top_k=40, top_p=0.9, repeat_penalty=1.1, seed=0),      # Ollama defaults
                                  dict(temperature=1.3, top_k=64, top_p=1.0, repeat_penalty=1.0, seed=77),
                                  dict(temperature=0.0, top_k=40, top_p=0.9, repeat_penalty=1.5, seed=5),      # penalised argmax
                                  dict(temperature=0.5, top_k=1, top_p=0.5, repeat_penalty=1.3, seed=9)])
def test_sampled_decode_in_the_persistent_kernel_matches_oracle(name, sigma, opts):
    """64 sampled tokens at a fixed seed against the CPU oracle forward + oracle/sampler.py.
    (1) Along the oracle's token path (b200llm_set_lane forces its token, one launch of the persistent kernel with
        the sampler as its last phase per step): the penalised logits agree within TIGHT_TOL and every draw is the
        oracle's, except where the MEASURED logit deviation can move the draw across a boundary (the cumulative
        masses, as fractions of the kept mass, move by at most 2 (exp(2 dev / T) - 1); for the penalised argmax, a top-2 margin below 2 dev): those
        are counted and printed, and must be few.
    (2) Free-running, all 63 steps in ONE launch: the oracle's tokens up to the first such boundary case, then the
        token (1) drew there; bit-identical to 63 launches of one step.
    (3) The per-phase kernels + stand-alone sample_kernel (B200LLM_NO_TCD_SAMPLER) draw the same tokens."""
    from b200llm import SamplingParams
    geo = lo.GEOMETRIES[name]
    sd = lo.make_weights(geo, 0xB200, sigma=sigma)
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    prompt = wg.gen_tokens(90, 31, geo.vocab)
    sp = SamplingParams(**opts)
    want, pen, slack, margin = _oracle_sampled(orc, prompt, 64, sp)
    assert len(set(want)) > 4                                            # the chain really samples
    eng = _engine(name, sd, max_lanes=2, max_ctx=256, max_rows=16, use_megakernel=True)
    # (1) teacher-forced
    lane = eng.open_lane()
    eng.set_sampling(lane, sp)
    eng.prefill(lane, prompt)
    forced, cases, worst = [], [], 0.0
    for s in range(64):
        if s == 0:
            tok = eng.read_lane(lane)[0]
        else:
            eng.set_lane(lane, want[s - 1], len(prompt) + s - 1)
            tok = int(eng.decode([lane], 1)[0, 0])
            assert eng.timing().kernel_launches == 1, "sampled decode must be one launch of the persistent kernel"
        got_lg = eng.logits(0)                                           # the kernel penalises the logits in place
        dev = float(np.abs(got_lg - pen[s]).max())
        worst = max(worst, dev / float(np.abs(pen[s]).max()))
        forced.append(tok)
        if tok != want[s]:
            explained = (slack[s] < 2 * np.expm1(2 * dev / sp.temperature) + 1e-6) if sp.temperature > 0 else (margin[s] <= 2 * dev)
            assert explained, (f"step {s}: drew {tok}, oracle {want[s]}; draw slack {slack[s]:.2e}, top-2 margin "
                               f"{margin[s]:.2e}, logit deviation {dev:.2e}")
            cases.append(s)
    eng.close_lane(lane)
    print(f"{name} {opts}: {64 - len(cases)}/64 draws identical to oracle forward + oracle/sampler.py along the oracle's "
          f"path; {len(cases)} boundary case(s) at steps {cases}; penalised logits within {worst:.1e} of the oracle's")
    assert worst <= TIGHT_TOL and len(cases) <= 4
    # (2) free-running, one launch
    toks, tm = eng.generate_ids(prompt, 64, sp)
    assert tm.kernel_launches == 1
    n_ok = cases[0] if cases else 64
    assert toks[:n_ok] == want[:n_ok] and (n_ok == 64 or toks[n_ok] == forced[n_ok])
    lane = eng.open_lane()
    eng.set_sampling(lane, sp)
    eng.prefill(lane, prompt)
    single = [eng.read_lane(lane)[0]] + [int(eng.decode([lane], 1)[0, 0]) for _ in range(63)]
    eng.close_lane(lane)
    assert single == toks
    eng.close()
    # (3) the multi-kernel path (CUDA graph + stand-alone sample_kernel) draws the same tokens
    os.environ["B200LLM_NO_TCD_SAMPLER"] = "1"
    try:
        eng2 = _engine(name, sd, max_lanes=2, max_ctx=256, max_rows=16, use_megakernel=True)
        toks2, tm2 = eng2.generate_ids(prompt, 64, sp)
        eng2.close()
    finally:
        del os.environ["B200LLM_NO_TCD_SAMPLER"]
    assert tm2.kernel_launches > 1
    assert toks2[:n_ok] == toks[:n_ok]


def test_sampled_decode_full_size_one_launch(golden_dir):
    """Llama-3.2-3B (128256 logits) with Ollama's default options.  One launch per decode call, and every draw of the
    in-kernel sampler equals oracle/sampler.py run on the very logits the kernel left in memory (penalty already applied
    in place) with the same seed and counter -- the chain itself, independent of logit rounding (the golden tests cover
    the logits).  Then 32 steps in one launch give the same tokens as 32 launches."""
    from b200llm import SamplingParams
    from oracle import sampler
    g = np.load(os.path.join(golden_dir, "llama3.2.npz"))
    sp = SamplingParams(temperature=0.8, top_k=40, top_p=0.9, repeat_penalty=1.1, seed=20250211)
    prompt = g["prompt"]
    eng = _engine("llama3.2", None, seed=int(g["seed"]), sigma=float(g["sigma"]), max_lanes=1,
                  max_ctx=len(prompt) + 128, max_rows=64, use_megakernel=True)
    lane = eng.open_lane()
    eng.set_sampling(lane, sp)
    eng.prefill(lane, prompt)
    toks = [eng.read_lane(lane)[0]]
    n_close, kept = 0, [None]
    for s in range(32):
        tok = int(eng.decode([lane], 1)[0, 0])
        assert eng.timing().kernel_launches == 1
        lg = eng.logits(0)
        ctr = len(prompt) + s                       # position of the token that was fed
        want, ids, probs = sampler.sample(lg, sp.temperature, sp.top_k, sp.top_p, 1.0, sp.seed, ctr)
        mass = float(probs.sum())
        r = float(sampler.uniform(sp.seed, ctr)) * mass
        slack = float(np.abs(np.cumsum(probs.astype(np.float64)) - r).min() / mass)
        if tok != want:                             # only an expf ulp at a cumulative-mass edge may differ
            assert slack < 1e-5, f"step {s}: kernel drew {tok}, oracle sampler on the same logits {want} (slack {slack:.2e})"
            n_close += 1
        toks.append(tok)
        kept.append(set(int(i) for i in ids))
    eng.close_lane(lane)
    print(f"3B sampled: 32/32 draws equal oracle/sampler.py on the kernel's own logits ({n_close} within an expf ulp of an edge)")
    assert len(set(toks)) > 8
    # the same 32 steps as ONE launch.  Split-K partial sums meet in fp32 atomics, whose order differs from run to run, so the
    # logits agree to ~4e-5 of the largest only (scripts/stress_logits.py) -- and the sampler is discontinuous in them: two
    # candidates whose logits are that close swap ranks and with them their slices of the cumulative mass, a top-p cut moves
    # by one candidate.  (Measured: from one and the same state 37 of 40 launches drew one token and 3 another, each equal to
    # oracle/sampler.py on its own logits.)  So the sequences may part; where they do, the other draw must at least be one of
    # the candidates the chain kept at that step.
    again, tm = eng.generate_ids(prompt, 33, sp)
    assert tm.kernel_launches == 1 and again[0] == toks[0]
    first = next((i for i, (a, b) in enumerate(zip(again, toks)) if a != b), None)
    assert first is None or again[first] in kept[first], f"one launch differs from 32 launches at step {first}: {again[first]}"
    eng.close()


def test_scheduler_fails_only_the_offending_request():
    """A request that does not fit the lane context gets its own error; requests batched with it finish normally,
    and a token budget beyond the context is cut to what fits instead of failing the batch."""
    import threading
    from b200llm import SamplingParams, _lib
    from b200llm.scheduler import Scheduler
    geo = lo.GEOMETRIES["tiny"]
    sd = lo.make_weights(geo, 0xB200, sigma=0.08)
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    eng = _engine("tiny", sd, max_lanes=4, max_ctx=96, max_rows=8)
    sched = Scheduler(eng, chunk=4)
    good = [wg.gen_tokens(12 + 5 * i, 700 + i, geo.vocab) for i in range(3)]
    want = [orc.greedy(p, 10)[0] for p in good]
    too_long = wg.gen_tokens(200, 9, geo.vocab)                  # > 96-token context
    res = {}

    def call(key, prompt, n):
        try:
            res[key] = sched.generate_ids(prompt, n, SamplingParams(temperature=0.0))[0]
        except Exception as e:  # noqa: BLE001
            res[key] = e
    ts = [threading.Thread(target=call, args=(i, good[i], 10)) for i in range(3)]
    ts.append(threading.Thread(target=call, args=("long", too_long, 10)))
    ts.append(threading.Thread(target=call, args=("greedy_budget", good[0], 10_000)))   # budget >> context: cut, not failed
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=120)
    assert isinstance(res["long"], _lib.B200LLMError) and res["long"].code == _lib.EINVAL
    assert [res[i] for i in range(3)] == want
    assert isinstance(res["greedy_budget"], list) and len(res["greedy_budget"]) == 96 - len(good[0])
    assert res["greedy_budget"][:10] == want[0]
    sched.shutdown()
    eng.close()


def test_c_abi_refuses_rows_outside_the_lanes_pages():
    """The C ABI itself (below the Python wrapper, which would grow the page table first) refuses a prefill or a
    decode whose K/V rows would land outside the pages the caller put in the lane's table, a lane or step count out
    of range, and a prompt beyond the lane context -- before any launch, and the engine keeps working afterwards."""
    import ctypes as C
    from b200llm import _lib
    geo = lo.GEOMETRIES["tiny"]
    sd = lo.make_weights(geo, 0xB200, sigma=0.08)
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    eng = _engine("tiny", sd, max_lanes=2, max_ctx=64, max_rows=4)
    lib, h = _lib.lib, eng._h
    lane = eng.open_lane()
    ids = np.ascontiguousarray(wg.gen_tokens(20, 5, geo.vocab), dtype=np.int32)
    p_ids = ids.ctypes.data_as(C.POINTER(C.c_int32))
    # no pages in the lane's table yet: 20 rows have nowhere to go
    assert lib.b200llm_prefill(h, lane, p_ids, 20, 0) == _lib.EINVAL
    assert b"pages" in lib.b200llm_last_error()
    assert lib.b200llm_prefill(h, 7, p_ids, 20, 0) == _lib.EINVAL               # lane out of range
    assert lib.b200llm_prefill(h, lane, p_ids, 20, 60) == _lib.EINVAL           # 60 + 20 > 64-token context
    assert lib.b200llm_prefill(h, lane, p_ids, 0, 0) == _lib.EINVAL
    eng.prefill(lane, ids)                                                       # the wrapper sets 2 pages (32 rows), pos = 20
    one = (C.c_int32 * 1)(lane)
    assert lib.b200llm_decode(h, one, 1, 13) == _lib.EINVAL                      # 20 + 13 > 32 rows of its 2 pages
    assert b"pages" in lib.b200llm_last_error()
    bad = (C.c_int32 * 1)(5)
    assert lib.b200llm_decode(h, bad, 1, 1) == _lib.EINVAL
    assert lib.b200llm_decode(h, one, 1, 0) == _lib.EINVAL
    assert lib.b200llm_decode(h, one, 9, 1) == _lib.EINVAL                       # more lanes than max_rows
    got = [eng.read_lane(lane)[0]] + eng.decode([lane], 11)[:, 0].tolist()      # refused calls left the lane untouched
    assert got == orc.greedy(ids.tolist(), 12)[0]
    eng.close()


def test_batch64_decode_of_the_1b_matches_golden(golden_dir):
    """BASELINE.json config 2: Llama-3.2-1B, 64 lanes decoded as ONE batch on the tcgen05 GEMM path (prefill in chunks on the
    same path), 16 greedy steps, every lane's tokens equal to the oracle's (tests/golden/llama3.2_1b_bs64.npz, written by
    oracle/make_golden_batch.py: 64 prompts of 24..87 ids whose oracle margins stay >= 5e-3 where a seed allowed)."""
    g = np.load(os.path.join(golden_dir, "llama3.2_1b_bs64.npz"))
    lens, prompts, want, margins = g["prompt_len"], g["prompts"], g["tokens"], g["margins"]
    n_lanes, n_new = want.shape
    assert n_lanes == 64 and n_new == 16
    eng = _engine("llama3.2:1b", None, seed=int(g["seed"]), sigma=float(g["sigma"]), max_lanes=64, max_ctx=128,
                  max_rows=64, use_tc=True)
    lanes = [eng.open_lane() for _ in range(n_lanes)]
    first = []
    for ln, p, n in zip(lanes, prompts, lens):
        eng.prefill(ln, p[:n])
        first.append(eng.read_lane(ln)[0])
    got = np.concatenate([np.asarray(first, np.int32)[None, :], eng.decode(lanes, n_new - 1)], axis=0).T    # [lane, step]
    bad = []
    for i in range(n_lanes):
        diff = np.nonzero(got[i] != want[i])[0]
        if len(diff) and margins[i, diff[0]] >= 1e-3:        # below that the oracle itself is at a near-tie
            bad.append((i, int(diff[0]), int(got[i, diff[0]]), int(want[i, diff[0]]), float(margins[i, diff[0]])))
    assert not bad, f"(lane, step, got, want, oracle margin): {bad}"
    exact = int((got == want).all(axis=1).sum())
    print(f"batch-64 decode of the 1B: {exact}/64 lanes identical over 16 steps (worst oracle margin {margins.min():.1e})")
    assert exact >= 60
    eng.close()


def test_several_lanes_at_very_different_positions(golden_dir):
    """Lanes of one launch at 40, 150, 300 and 470 tokens: their attention runs with 1, 2, 4 and 5 splits per kv head in
    the same phase (split-T partials merged per lane by the o_proj prologue), pages beyond the first of the table; 8 greedy
    steps, every lane's tokens equal to the oracle's."""
    g = np.load(os.path.join(golden_dir, "small.npz"))
    geo = lo.GEOMETRIES["small"]
    sd = lo.make_weights(geo, int(g["seed"]), sigma=float(g["sigma"]), embed_sigma=float(g["embed_sigma"]))
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    prompts = [wg.gen_tokens(n, 2000 + n, geo.vocab) for n in (40, 150, 300, 470)]
    eng = _engine("small", sd, max_lanes=4, max_ctx=512, max_rows=64)
    lanes = [eng.open_lane() for _ in prompts]
    want = [orc.greedy(p, 9)[0] for p in prompts]
    first = []
    for ln, p in zip(lanes, prompts):
        eng.prefill(ln, p)
        first.append(eng.read_lane(ln)[0])
    got = eng.decode(lanes, 8)
    assert eng.timing().kernel_launches == 1
    for i in range(4):
        assert [first[i]] + got[:, i].tolist() == want[i], f"lane {i} ({len(prompts[i])} ids)"
    eng.close()


@pytest.mark.parametrize("name,n_lanes", [("tiny", 2), ("tiny", 4), ("small", 3), ("small", 4)])
def test_several_lanes_in_one_persistent_kernel_launch(name, n_lanes, golden_dir):
    """2..4 lanes decoded by ONE launch of the tcgen05 persistent kernel (operand rows 2b / 2b + 1 = lane b): lanes at
    different positions, 12 greedy steps, every lane's tokens equal to the oracle's and its last-step logits within the
    tolerance of the single-lane path; then the same lanes continue one by one (single-lane launches) on the KV the
    multi-lane launch appended."""
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    geo = lo.GEOMETRIES[name]
    sd = lo.make_weights(geo, int(g["seed"]), sigma=float(g["sigma"]), embed_sigma=float(g["embed_sigma"]))
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    prompts = [wg.gen_tokens(9 + 14 * i, 900 + i, geo.vocab) for i in range(n_lanes)]       # 9, 23, 37, 51 ids: different pages
    eng = _engine(name, sd, max_lanes=4, max_ctx=128, max_rows=8)
    lanes = [eng.open_lane() for _ in prompts]
    want, states = [], []
    for p in prompts:
        st = orc.new_state()
        toks = [int(torch.argmax(orc.forward(torch.as_tensor(p, dtype=torch.int64), st)))]
        lg = None
        for _ in range(12):
            lg = orc.forward(torch.tensor([toks[-1]]), st)
            toks.append(int(torch.argmax(lg)))
        want.append(toks)
        states.append((st, lg))
    first = []
    for ln, p in zip(lanes, prompts):
        eng.prefill(ln, p)
        first.append(eng.read_lane(ln)[0])
    got = eng.decode(lanes, 12)                                  # [12, n_lanes]
    assert eng.timing().kernel_launches == 1, "the lanes did not ride one persistent-kernel launch"
    for i in range(n_lanes):
        assert [first[i]] + got[:, i].tolist() == want[i], f"lane {i}"
        assert _rel(eng.logits(i), states[i][1].numpy()) <= TIGHT_TOL
    # one lane at a time on what the joint launch left (positions, KV pages, token feedback)
    for i in range(n_lanes):
        st = states[i][0]
        nxt = [int(torch.argmax(orc.forward(torch.tensor([want[i][-1]]), st)))]
        for _ in range(3):
            nxt.append(int(torch.argmax(orc.forward(torch.tensor([nxt[-1]]), st))))
        assert eng.decode([lanes[i]], 4)[:, 0].tolist() == nxt, f"lane {i} alone after the joint launch"
    eng.close()


def test_several_lanes_with_their_own_sampling_settings():
    """Lanes of one launch keep their own sampler settings (greedy, Ollama's defaults, a hot top-k 8) and seeds: the joint
    launch draws what every lane draws when it is decoded alone from the same state."""
    from b200llm import SamplingParams
    geo = lo.GEOMETRIES["small"]
    sd = lo.make_weights(geo, 0xB200, sigma=0.06)
    settings = [SamplingParams(temperature=0.0), SamplingParams(temperature=0.8, top_k=40, top_p=0.9, repeat_penalty=1.1, seed=11),
                SamplingParams(temperature=1.3, top_k=8, top_p=1.0, repeat_penalty=1.0, seed=5),
                SamplingParams(temperature=0.0, repeat_penalty=1.3)]
    prompts = [wg.gen_tokens(12 + 7 * i, 40 + i, geo.vocab) for i in range(4)]

    def run(joint):
        eng = _engine("small", sd, max_lanes=4, max_ctx=128, max_rows=8)
        lanes = [eng.open_lane() for _ in prompts]
        for ln, p, sp in zip(lanes, prompts, settings):
            eng.set_sampling(ln, sp)
            eng.prefill(ln, p)
        if joint:
            out = eng.decode(lanes, 10)
            assert eng.timing().kernel_launches == 1
            res = [out[:, i].tolist() for i in range(4)]
        else:
            res = []
            for ln in lanes:
                res.append(eng.decode([ln], 10)[:, 0].tolist())
                assert eng.timing().kernel_launches == 1
        eng.close()
        return res
    alone, joint = run(False), run(True)
    assert joint[0] == alone[0] and joint[3] == alone[3]            # the greedy / penalised-argmax lanes
    assert joint[1] == alone[1] and joint[2] == alone[2]            # the sampled lanes (same seeds, same positions)
    assert len({tuple(t) for t in joint}) == 4


def test_binding_a_buffer_of_another_gpu_is_refused():
    """Weights (or a pool, or a workspace) that live on another GPU of the process are refused at bind time with a message
    that names both GPUs, instead of faulting in the first kernel that touches them."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from b200llm import Engine, _lib, geometry, weights
    geo = geometry.resolve("tiny")
    with torch.cuda.device(1):
        w_other = weights.random_init(geo, 0xB200, 1, sigma=0.08)
    with pytest.raises(_lib.B200LLMError) as ei:
        Engine(geo, w_other, device=0, max_lanes=1, max_ctx=64, max_rows=8)
    assert ei.value.code == _lib.EINVAL and "GPU 1" in str(ei.value) and "GPU 0" in str(ei.value)
    torch.cuda.set_device(0)


@pytest.mark.parametrize("name", ["tiny", "small"])
def test_lean_engine_with_tile_major_weights_only(name, golden_dir):
    """lean_weights=True drops the row-major matrices (the GEMV kernels' layout): the persistent kernel (1-4 lanes) and the
    tcgen05 GEMM path (prefill chunks down to a single row, larger batches) both stream the tile-major copies.  Same
    tokens as the oracle on every path, about half the weight bytes, and a clear error where a GEMV kernel would be needed."""
    from b200llm import Engine, SamplingParams, _lib, geometry, weights
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    ogeo = lo.GEOMETRIES[name]
    sd = lo.make_weights(ogeo, int(g["seed"]), sigma=float(g["sigma"]), embed_sigma=float(g["embed_sigma"]))
    orc = lo.LlamaOracle(ogeo, sd, kv_dtype=torch.bfloat16)
    geo = geometry.resolve(name)
    full = weights.pack(geo, {k: v.cuda() for k, v in sd.items()})
    weights.add_tile_major(full)
    full_bytes = sum(t.numel() * t.element_size() for _, _, t in full.tensors)
    w = weights.pack(geo, {k: v.cuda() for k, v in sd.items()})
    eng = Engine(geo, w, max_lanes=6, max_ctx=160, max_rows=32, lean_weights=True)
    lean_bytes = sum(t.numel() * t.element_size() for _, _, t in w.tensors)
    assert lean_bytes < 0.7 * full_bytes
    prompts = [wg.gen_tokens(n, 3000 + n, ogeo.vocab) for n in (33, 64, 20, 47, 9, 70)]      # 33 = 32 + a 1-row tail chunk
    want = [orc.greedy(p, 9)[0] for p in prompts]
    lanes = [eng.open_lane() for _ in prompts]
    first = []
    for ln, p in zip(lanes, prompts):
        eng.prefill(ln, p)
        first.append(eng.read_lane(ln)[0])
    one = eng.decode(lanes[:1], 8)                      # persistent kernel, one lane
    assert eng.timing().kernel_launches == 1 and [first[0]] + one[:, 0].tolist() == want[0]
    three = eng.decode(lanes[1:4], 8)                   # persistent kernel, three lanes
    assert eng.timing().kernel_launches == 1
    for j in range(3):
        assert [first[1 + j]] + three[:, j].tolist() == want[1 + j]
    for ln, p in zip(lanes, prompts):                   # six lanes at once: the GEMM path
        eng.close_lane(ln)
    lanes = [eng.open_lane() for _ in prompts]
    for ln, p in zip(lanes, prompts):
        eng.prefill(ln, p)
    six = eng.decode(lanes, 8)
    for j in range(6):
        assert [first[j]] + six[:, j].tolist() == want[j], f"lane {j} on the GEMM path"
    # a sampler setting the persistent kernel does not take (top-k 200) runs the GEMM pass + the stand-alone sampler
    eng.set_sampling(lanes[0], SamplingParams(temperature=0.7, top_k=200, top_p=0.95, repeat_penalty=1.0, seed=3))
    assert eng.decode(lanes[:1], 2).shape == (2, 1) and eng.timing().kernel_launches > 1
    eng.close()
    # without the tcgen05 GEMM path nothing could run the prefill: refused at bind time
    w2 = weights.pack(geo, {k: v.cuda() for k, v in sd.items()})
    with pytest.raises(_lib.B200LLMError):
        Engine(geo, w2, max_lanes=1, max_ctx=64, max_rows=8, lean_weights=True, use_tc=False)


@pytest.mark.parametrize("name", ["tiny", "small"])
def test_prefill_with_the_row_group_attention_kernel(name, golden_dir, monkeypatch):
    """attn_rows_kernel (four rows of a prefill chunk per CTA share every K/V tile; the engine picks it for short contexts
    with many kv heads) forced on for the small geometries: ragged prompts in 32-row chunks (a 3-row tail takes the per-row
    kernel), prefill logits and 6 greedy tokens against the oracle."""
    monkeypatch.setenv("B200LLM_ROWS_ATTN", "1")
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    geo = lo.GEOMETRIES[name]
    sd = lo.make_weights(geo, int(g["seed"]), sigma=float(g["sigma"]), embed_sigma=float(g["embed_sigma"]))
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    eng = _engine(name, sd, max_lanes=2, max_ctx=256, max_rows=32, use_tc=True)
    for n in (7, 35, 64, 150, 201):
        p = wg.gen_tokens(n, 4100 + n, geo.vocab)
        st = orc.new_state()
        ref = orc.forward(torch.as_tensor(p, dtype=torch.int64), st)
        lane = eng.open_lane()
        eng.prefill(lane, p)
        eng.sync()
        assert _rel(eng.logits(0), ref.numpy()) <= TIGHT_TOL, n
        want = orc.greedy(p, 6)[0]
        assert [eng.read_lane(lane)[0]] + eng.decode([lane], 5)[:, 0].tolist() == want, n
        eng.close_lane(lane)
    eng.close()
