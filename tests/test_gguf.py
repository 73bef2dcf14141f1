"""GGUF loader (second next row).  The `gguf` package that ships with llama.cpp is
the checker here: it WRITES the test files and its reference de-quantiser is compared with
b200llm.gguf_loader's; the product code never imports it."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]

gguf = pytest.importorskip("gguf")
from b200llm import geometry, gguf_loader  # noqa: E402

QT = gguf.GGMLQuantizationType


@pytest.mark.parametrize("qt", [QT.Q8_0, QT.Q4_0, QT.F16, QT.BF16, QT.F32])
def test_dequantize_matches_ggml_reference(qt):
    rng = np.random.default_rng(int(qt))
    x = (rng.standard_normal((8, 512)) * 0.05).astype(np.float32)
    x[0, :7] = [0.0, 1.0, -1.0, 0.5, -0.031, 3.0, -2.5]
    q = gguf.quants.quantize(x, qt)
    ref = gguf.quants.dequantize(q, qt).astype(np.float32).reshape(-1)
    got = gguf_loader.dequantize(np.ascontiguousarray(q).view(np.uint8).reshape(-1), int(qt))
    assert got.dtype == np.float32 and got.shape == ref.shape
    assert np.array_equal(got, ref)                      # block arithmetic is exact in fp32


@pytest.mark.parametrize("qt,block_bytes,f16_at", [(QT.Q4_K, 144, (0, 2)), (QT.Q6_K, 210, (208,))])
def test_k_quant_blocks_match_ggml_reference(qt, block_bytes, f16_at):
    """gguf-py cannot quantise the K formats, so the blocks are random bit patterns (every value of
    the packed 6-bit scales, nibbles and high bits occurs) with finite fp16 super-block scales."""
    rng = np.random.default_rng(int(qt))
    nb = 96
    raw = rng.integers(0, 256, (nb, block_bytes), dtype=np.uint8)
    for off in f16_at:
        raw[:, off:off + 2] = (rng.standard_normal(nb) * 0.01).astype(np.float16).view(np.uint8).reshape(nb, 2)
    ref = gguf.quants.dequantize(raw.reshape(1, -1), qt).astype(np.float32).reshape(-1)
    got = gguf_loader.dequantize(raw.reshape(-1), int(qt))
    assert got.shape == ref.shape == (nb * 256,)
    assert np.array_equal(got, ref)


def _hf_state_dict(geo, seed):
    g = torch.Generator().manual_seed(seed)
    qd, kd = geo.n_q * geo.head_dim, geo.n_kv * geo.head_dim
    sd = {"model.embed_tokens.weight": torch.randn(geo.vocab, geo.d, generator=g) * 0.05,
          "model.norm.weight": 1 + torch.randn(geo.d, generator=g) * 0.02}
    if not geo.tied:
        sd["lm_head.weight"] = torch.randn(geo.vocab, geo.d, generator=g) * 0.05
    for l in range(geo.n_layers):
        p = f"model.layers.{l}."
        for nm, shape in (("self_attn.q_proj", (qd, geo.d)), ("self_attn.k_proj", (kd, geo.d)),
                          ("self_attn.v_proj", (kd, geo.d)), ("self_attn.o_proj", (geo.d, qd)),
                          ("mlp.gate_proj", (geo.ffn, geo.d)), ("mlp.up_proj", (geo.ffn, geo.d)),
                          ("mlp.down_proj", (geo.d, geo.ffn))):
            sd[p + nm + ".weight"] = torch.randn(*shape, generator=g) * 0.05
        sd[p + "input_layernorm.weight"] = 1 + torch.randn(geo.d, generator=g) * 0.02
        sd[p + "post_attention_layernorm.weight"] = 1 + torch.randn(geo.d, generator=g) * 0.02
    return {k: v.to(torch.bfloat16) for k, v in sd.items()}


def _permute(w, n_head):      # llama.cpp convert_hf_to_gguf.py LlamaModel.permute
    return w.reshape(n_head, 2, w.shape[0] // n_head // 2, *w.shape[1:]).swapaxes(1, 2).reshape(w.shape)


_INV = {v: k for k, v in gguf_loader._LAYER_MAP.items()}


def write_gguf(path, geo, sd, qt=None, tok_md=None):
    """What llama.cpp's converter emits for a Llama checkpoint (names, q/k permutation, metadata; with `tok_md`
    also the byte-level BPE tokenizer tables)."""
    w = gguf.GGUFWriter(path, "llama")
    w.add_name(geo.name)
    w.add_block_count(geo.n_layers)
    w.add_embedding_length(geo.d)
    w.add_feed_forward_length(geo.ffn)
    w.add_head_count(geo.n_q)
    w.add_head_count_kv(geo.n_kv)
    w.add_key_length(geo.head_dim)
    w.add_rope_freq_base(geo.rope_theta)
    w.add_layer_norm_rms_eps(geo.rms_eps)
    if tok_md is None:
        w.add_token_list([f"<{i}>" for i in range(geo.vocab)])
    else:
        w.add_tokenizer_model(tok_md["tokenizer.ggml.model"])
        w.add_tokenizer_pre(tok_md["tokenizer.ggml.pre"])
        w.add_token_list(tok_md["tokenizer.ggml.tokens"])
        w.add_token_types(tok_md["tokenizer.ggml.token_type"])
        w.add_token_merges(tok_md["tokenizer.ggml.merges"])
        w.add_bos_token_id(tok_md["tokenizer.ggml.bos_token_id"])
        w.add_eos_token_id(tok_md["tokenizer.ggml.eos_token_id"])
    if geo.rope_llama3:
        w.add_tensor("rope_freqs.weight", np.ones(geo.head_dim // 2, np.float32))
    for name, t in sd.items():
        a = t.float().numpy()
        if name == "model.embed_tokens.weight":
            gname = "token_embd.weight"
        elif name == "model.norm.weight":
            gname = "output_norm.weight"
        elif name == "lm_head.weight":
            gname = "output.weight"
        else:
            parts = name.split(".")
            gname = f"blk.{parts[2]}.{_INV['.'.join(parts[3:-1])]}.weight"
        if gname.endswith("attn_q.weight"):
            a = _permute(a, geo.n_q)
        if gname.endswith("attn_k.weight"):
            a = _permute(a, geo.n_kv)
        if qt is not None and a.ndim == 2 and a.shape[1] % gguf.GGML_QUANT_SIZES[qt][0] == 0:
            w.add_tensor(gname, gguf.quants.quantize(a, qt), raw_shape=None, raw_dtype=qt)
        else:
            w.add_tensor(gname, a.astype(np.float32))
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()


def test_file_round_trip_names_permutation_geometry(tmp_path):
    geo = geometry.resolve("tiny")
    sd = _hf_state_dict(geo, 3)
    path = str(tmp_path / "tiny-f32.gguf")
    write_gguf(path, geo, sd)
    gf = gguf_loader.GGUFFile(path)
    assert gf.metadata["llama.block_count"] == geo.n_layers and len(gf.metadata["tokenizer.ggml.tokens"]) == geo.vocab
    g2, sd2 = gguf_loader.load(path, "cpu", "tiny-from-gguf")
    for f in ("d", "n_layers", "n_q", "n_kv", "head_dim", "ffn", "vocab", "tied", "rope_llama3"):
        assert getattr(g2, f) == getattr(geo, f), f
    # GGUF stores both as float32, which is also what the engine's config struct holds
    assert np.float32(g2.rms_eps) == np.float32(geo.rms_eps) and np.float32(g2.rope_theta) == np.float32(geo.rope_theta)
    assert set(sd2) == set(sd)
    for k in sd:
        assert torch.equal(sd2[k], sd[k]), k            # bf16 -> f32 file -> bf16 is exact, permutation undone


def test_quantised_file_loads_as_its_dequantised_values(tmp_path):
    geo = geometry.resolve("small")
    sd = _hf_state_dict(geo, 5)
    path = str(tmp_path / "small-q4_0.gguf")
    write_gguf(path, geo, sd, QT.Q4_0)
    g2, sd2 = gguf_loader.load(path, "cpu")
    assert not g2.tied and g2.n_kv == geo.n_kv
    k = "model.layers.1.mlp.down_proj.weight"
    ref = gguf.quants.dequantize(gguf.quants.quantize(sd[k].float().numpy(), QT.Q4_0), QT.Q4_0)
    assert torch.equal(sd2[k], torch.from_numpy(ref).to(torch.bfloat16))
    assert float((sd2[k].float() - sd[k].float()).abs().max()) < 0.05       # and it is close to the original


def test_rejects_other_files(tmp_path):
    p = tmp_path / "x.gguf"
    p.write_bytes(b"not a gguf file at all")
    with pytest.raises(ValueError):
        gguf_loader.GGUFFile(str(p))


@pytest.mark.gpu
def test_gguf_checkpoint_runs_like_the_same_weights_packed_directly(tmp_path):
    """Engine on weights loaded from a GGUF file == engine on the same HF tensors packed directly
    (greedy tokens), and both equal the CPU oracle on those weights."""
    from b200llm import Engine, weights
    from oracle import llama_oracle as lo
    geo = geometry.resolve("small")
    sd = _hf_state_dict(geo, 11)
    path = str(tmp_path / "small.gguf")
    write_gguf(path, geo, sd)
    g2, sd2 = gguf_loader.load(path, "cuda", "small")
    prompt = np.arange(3, 40, dtype=np.int32)
    outs = []
    for g, s in ((geo, {k: v.cuda() for k, v in sd.items()}), (g2, sd2)):
        eng = Engine(g, weights.pack(g, s), max_lanes=1, max_ctx=256, max_rows=64)
        toks, _ = eng.generate_ids(prompt, 24)
        outs.append(list(toks))
        eng.close()
    assert outs[0] == outs[1]
    oracle = lo.LlamaOracle(lo.GEOMETRIES["small"], {k: v.float() for k, v in sd.items()}, kv_dtype=torch.bfloat16)
    assert outs[0] == oracle.greedy(prompt.tolist(), 24, keep_logits=False)[0]


@pytest.mark.gpu
def test_text_in_text_out_with_the_checkpoints_tokenizer_and_template(tmp_path):
    """`ollama.generate(model=, system=, prompt=) -> .response` on a GGUF checkpoint that carries tokenizer tables
    (the reference's call, Flask/app.py:160-166): the prompt goes through the model's template and the file's
    byte-level BPE, the answer back through it.  Checkers: the `tokenizers` library for text <-> ids, the CPU
    oracle for the ids in between.  The vocabulary (trained here) is no multiple of 128 rows."""
    from dataclasses import replace
    from b200llm import runtime, tokenizer
    from oracle import llama_oracle as lo
    from test_tokenizer import train_bpe
    import ollama
    tk, tok, vocab, md = train_bpe()
    name = "llama3.2:tiny-bpe"
    geo = replace(geometry.resolve("tiny"), name=name, vocab=len(tok.tokens))
    assert geo.vocab % 128 != 0
    sd = _hf_state_dict(geo, 23)
    path = str(tmp_path / "tiny_bpe.gguf")
    write_gguf(path, geo, sd, tok_md=md)
    gguf_loader.register(name, path)
    system, prompt = "You explain errors.", "Why did the query fail near the unexpected token?"
    try:
        r = ollama.generate(model=name, system=system, prompt=prompt, options={"temperature": 0, "repeat_penalty": 1.0, "num_predict": 20})
    finally:
        runtime.unload_all()
    text = tokenizer.render_prompt(system, prompt, model="llama3.2")
    ids = [vocab["<|begin_of_text|>"]] + tk.encode(text, add_special_tokens=False).ids
    assert r.prompt_eval_count == len(ids)
    oracle = lo.LlamaOracle(replace(lo.GEOMETRIES["tiny"], vocab=geo.vocab), {k: v.float() for k, v in sd.items()},
                            kv_dtype=torch.bfloat16)
    want = oracle.greedy(ids, 20, keep_logits=False)[0]
    stops = {vocab["<|end_of_text|>"], vocab["<|eot_id|>"]}
    cut = next((i for i, t in enumerate(want) if t in stops), len(want))
    assert r.response == tk.decode(want[:cut], skip_special_tokens=True)
    assert r.done_reason == ("stop" if cut < len(want) else "length")
    assert list(r.context) == ids + want[:cut]


@pytest.mark.gpu
@pytest.mark.parametrize("qt,block_bytes,per,f16_at", [(QT.Q4_K, 144, 256, (0, 2)), (QT.Q6_K, 210, 256, (208,)), (QT.Q4_0, 18, 32, (0,)),
                                                      (QT.Q8_0, 34, 32, (0,)), (QT.F16, 2, 1, ()), (QT.BF16, 2, 1, ()), (QT.F32, 4, 1, ())])
def test_device_dequantiser_is_bit_equal_to_the_host_one(qt, block_bytes, per, f16_at):
    """csrc/dequant.cu (what `load` uses on a CUDA device) against the numpy de-quantiser above, which the CPU tests pin to
    ggml's reference: random bit patterns in every field, finite fp16 scales; the bf16 results must be the same bits."""
    rng = np.random.default_rng(100 + int(qt))
    nb = 4096 if per > 1 else 8192
    raw = rng.integers(0, 256, (nb, block_bytes), dtype=np.uint8)
    for off in f16_at:
        raw[:, off:off + 2] = (rng.standard_normal(nb) * 0.01).astype(np.float16).view(np.uint8).reshape(nb, 2)
    if qt == QT.F16:
        raw = (rng.standard_normal(nb) * 3).astype(np.float16).view(np.uint8).reshape(nb, 2)
    if qt == QT.F32:
        raw = (rng.standard_normal(nb) * 3).astype(np.float32).view(np.uint8).reshape(nb, 4)
    if qt == QT.BF16:
        raw[:, 1] &= 0x7E          # keep the exponent away from inf / nan
    want = torch.from_numpy(gguf_loader.dequantize(raw.reshape(-1), int(qt))).to(torch.bfloat16)
    got = gguf_loader.dequantize_on_device(raw.reshape(-1), int(qt), (nb * per,), "cuda")
    torch.cuda.synchronize()
    assert got.dtype == torch.bfloat16 and got.shape == want.shape
    assert torch.equal(got.cpu().view(torch.int16), want.view(torch.int16))


@pytest.mark.gpu
def test_quantised_file_expanded_on_the_device_equals_the_host_path(tmp_path, monkeypatch):
    geo = geometry.resolve("small")
    sd = _hf_state_dict(geo, 5)
    path = str(tmp_path / "small-q4_0.gguf")
    write_gguf(path, geo, sd, QT.Q4_0)
    _, on_dev = gguf_loader.load(path, "cuda")
    monkeypatch.setenv("B200LLM_GGUF_HOST_DEQUANT", "1")
    _, on_host = gguf_loader.load(path, "cuda")
    assert set(on_dev) == set(on_host)
    for k in on_dev:                                  # q / k un-permutation included
        assert torch.equal(on_dev[k].view(torch.int16), on_host[k].view(torch.int16)), k
