"""CPU tests of the host logic: registry, byte accounting, tokenizer stand-in, the shim's
call surface (the six reference call sites), weight packing."""
import inspect
import os

import numpy as np
import pytest
import torch

from b200llm import geometry
from b200llm.tokenizer import ByteTokenizer, render_prompt
from oracle import llama_oracle as lo


def test_registry_matches_oracle_geometry_and_survey_constants():
    for name, og in lo.GEOMETRIES.items():
        g = geometry.resolve(name)
        for f in ("d", "n_layers", "n_q", "n_kv", "head_dim", "ffn", "vocab", "tied", "rope_theta", "rope_llama3"):
            assert getattr(g, f) == getattr(og, f), (name, f)
        assert torch.equal(geometry.inv_freq(g), lo.inv_freq(og))
    # parameter counts and algorithmic bytes per decode step
    assert geometry.resolve("llama3.2:1b").params == 1_235_814_400
    assert geometry.resolve("llama3.2").params == 3_212_749_824
    assert geometry.resolve("duckdb-nsql").params == 6_738_415_616
    assert geometry.resolve("llama3.2:1b").weight_bytes_per_step() == 2_471_628_800
    assert geometry.resolve("llama3.2").weight_bytes_per_step() == 6_425_499_648
    assert geometry.resolve("duckdb-nsql").weight_bytes_per_step() == 13_214_687_232
    assert geometry.resolve("duckdb-nsql:codellama").weight_bytes_per_step() == 13_214_818_304
    assert [geometry.resolve(n).kv_bytes_per_token() for n in ("llama3.2:1b", "llama3.2", "duckdb-nsql")] == \
        [32_768, 114_688, 524_288]
    assert geometry.resolve("llama3.2:latest") is geometry.resolve("llama3.2")
    with pytest.raises(KeyError):
        geometry.resolve("mistral")


def test_tokenizer_roundtrip_and_prompt():
    tk = ByteTokenizer(128256)
    s = "SELECT * FROM temp_view WHERE `Location_of_company` = 'Raipur';\n"
    ids = tk.encode(s)
    assert ids[0] == 1 and tk.decode(ids) == s
    assert len(tk.decode([5000, 70000, 128255])) == 3          # any id decodes to something printable
    assert render_prompt("sys", "q") == "sys\n\nq" and render_prompt("", "q") == "q"
    assert render_prompt("s", "p", raw=True) == "p"


def test_shim_signature_is_the_upstream_one():
    import ollama
    params = list(inspect.signature(ollama.generate).parameters)
    assert params[:12] == ["model", "prompt", "suffix", "system", "template", "context", "stream", "raw", "format",
                           "images", "options", "keep_alive"]
    r = ollama.GenerateResponse(model="m", response="SELECT 1;", done=True, eval_count=3)
    assert r.response == "SELECT 1;" and r["response"] == "SELECT 1;" and "eval_count" in r
    assert r.get("nope", 7) == 7
    with pytest.raises(KeyError):
        r["nope"]
    e = ollama.ResponseError("model 'x' not found", 404)
    assert e.error.startswith("model") and e.status_code == 404 and isinstance(e, Exception)
    with pytest.raises(ollama.ResponseError):
        ollama.generate(model="")


def test_pack_layout_cpu():
    """q/k rows pair-interleaved per head, gate/up interleaved, TP slices partition the tensors."""
    from b200llm import _lib, weights
    geo = geometry.resolve("tiny")
    sd = {k: v for k, v in lo.make_weights(lo.GEOMETRIES["tiny"], 1).items()}
    full = {(k, l): t for k, l, t in weights.pack(geo, sd).tensors}
    hd, half = geo.head_dim, geo.head_dim // 2
    q = sd["model.layers.1.self_attn.q_proj.weight"]
    qkv = full[(_lib.W_QKV, 1)]
    assert qkv.shape == ((geo.n_q + 2 * geo.n_kv) * hd, geo.d)
    assert torch.equal(qkv[2 * 5], q[5]) and torch.equal(qkv[2 * 5 + 1], q[5 + half])
    assert torch.equal(qkv[hd + 2 * 7 + 1], q[hd + 7 + half])
    k = sd["model.layers.1.self_attn.k_proj.weight"]
    assert torch.equal(qkv[geo.n_q * hd + 2 * 3], k[3])
    v = sd["model.layers.1.self_attn.v_proj.weight"]
    assert torch.equal(qkv[(geo.n_q + geo.n_kv) * hd:], v)
    gu = full[(_lib.W_GATE_UP, 0)]
    assert torch.equal(gu[0::2], sd["model.layers.0.mlp.gate_proj.weight"])
    assert torch.equal(gu[1::2], sd["model.layers.0.mlp.up_proj.weight"])
    assert full[(_lib.W_LM_HEAD, -1)].data_ptr() == full[(_lib.W_EMBED, -1)].data_ptr()      # tied, tp=1
    parts = [{(k, l): t for k, l, t in weights.pack(geo, sd, r, 2).tensors} for r in range(2)]
    assert torch.equal(torch.cat([p[(_lib.W_O, 0)] for p in parts], 1), sd["model.layers.0.self_attn.o_proj.weight"])
    assert torch.equal(torch.cat([p[(_lib.W_DOWN, 0)] for p in parts], 1), sd["model.layers.0.mlp.down_proj.weight"])
    assert torch.equal(torch.cat([p[(_lib.W_LM_HEAD, -1)] for p in parts], 0), sd["model.embed_tokens.weight"])
    assert parts[0][(_lib.W_QKV, 0)].shape[0] == (geo.n_q + 2 * geo.n_kv) * hd // 2


def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` (the CPU arm the driver times beside ours) must exit 0 and print one JSON line
    with the contract's keys; run here on the tiny geometry so that it takes seconds."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--model", "tiny",
                          "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=300, cwd=root)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "tok/s" and line["value"] > 0
    for key in ("metric", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "config", "cpu_baseline", "e2e"):
        assert key in line
    assert line["cpu_baseline"]["kind"] in ("port", "reference") and line["e2e"]["h2d_bytes_per_step"] == 0
