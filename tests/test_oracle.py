"""CPU tests of the oracle (test infrastructure): pinned against HF transformers' Llama
(the algorithm it restates) and against the committed golden fixtures."""
import os

import numpy as np
import pytest
import torch

from oracle import llama_oracle as lo
from oracle import sampler as osamp
from oracle import weightgen as wg


def test_weightgen_c_matches_numpy():
    a = wg.gen_bf16((257, 129), 0xABCDEF, 0.02)
    b = wg.gen_bf16((257, 129), 0xABCDEF, 0.02, force_numpy=True)
    assert torch.equal(a, b)
    g = wg.gen_bf16((4096,), 7, 0.02, 1.0)
    assert abs(float(g.float().mean()) - 1.0) < 2e-3 and abs(float(g.float().std()) - 0.02) < 2e-3
    assert abs(float(a.float().std()) - 0.02) < 5e-4


def test_weightgen_known_answers():
    # frozen first elements of the stream: any change to the spec breaks every golden file
    t = wg.gen_bf16((4,), wg.tensor_seed(0xB200, 0), 0.02).view(torch.int16).numpy().astype(np.uint16)
    assert t.tolist() == [0xB94F, 0xBA89, 0xBC7F, 0xBC5E]
    ids = wg.gen_tokens(8, 1, 128000)
    assert ids.tolist() == [24651, 24029, 122442, 126041, 78902, 1696, 66247, 85232]


def _hf_model(geo, sd):
    from transformers import LlamaConfig, LlamaForCausalLM
    rp = {"rope_type": "llama3", "rope_theta": geo.rope_theta, "factor": 32.0, "low_freq_factor": 1.0,
          "high_freq_factor": 4.0, "original_max_position_embeddings": 8192} if geo.rope_llama3 else \
         {"rope_type": "default", "rope_theta": geo.rope_theta}
    cfg = LlamaConfig(vocab_size=geo.vocab, hidden_size=geo.d, intermediate_size=geo.ffn,
                      num_hidden_layers=geo.n_layers, num_attention_heads=geo.n_q, num_key_value_heads=geo.n_kv,
                      head_dim=geo.head_dim, rms_norm_eps=geo.rms_eps, tie_word_embeddings=geo.tied,
                      max_position_embeddings=131072, rope_parameters=rp)
    m = LlamaForCausalLM(cfg).float().eval()
    hsd = {k: v.float() for k, v in sd.items()}
    if geo.tied:
        hsd["lm_head.weight"] = hsd["model.embed_tokens.weight"]
    m.load_state_dict(hsd, strict=True)
    return m


@pytest.mark.parametrize("name", ["tiny", "small"])
def test_oracle_matches_hf_llama(name):
    """Pins the restatement against transformers' modeling_llama on the same weights."""
    geo = lo.GEOMETRIES[name]
    sd = lo.make_weights(geo, 0xB200, sigma=0.05)
    m = _hf_model(geo, sd)
    ids = torch.from_numpy(wg.gen_tokens(45, 3, geo.vocab)).long()
    with torch.no_grad():
        ref = m(ids[None]).logits[0]
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.float32)
    mine = orc.forward(ids, orc.new_state(), all_logits=True)
    assert float((ref - mine).abs().max()) <= 2e-5 * float(ref.abs().max()) + 1e-5
    assert torch.equal(m.model.rotary_emb.inv_freq, lo.inv_freq(geo))
    # incremental decode == full forward; HF greedy == oracle greedy
    st = orc.new_state()
    l1 = orc.forward(ids[:30], st)
    for i in range(30, 45):
        l1 = orc.forward(ids[i:i + 1], st)
    assert float((l1 - mine[-1]).abs().max()) < 1e-4
    toks, _, _ = orc.greedy(ids[:20], 12)
    with torch.no_grad():
        hf = m.generate(ids[None, :20], max_new_tokens=12, do_sample=False)[0, 20:].tolist()
    assert hf == toks


@pytest.mark.parametrize("name", sorted(lo.GEOMETRIES))
def test_inv_freq_bit_equal_to_transformers_for_every_geometry(name):
    """The RoPE table of every served geometry (head_dim 64 + llama3 scaling of the 1B, head_dim 128 + llama3 of the 3B,
    head_dim 128 default of the 7B, ...) is bit-equal to LlamaRotaryEmbedding's, in the oracle AND in the product's
    host-side table (b200llm.geometry.inv_freq, what the engine uploads).  No model weights needed."""
    from transformers import LlamaConfig
    from transformers.models.llama.modeling_llama import LlamaRotaryEmbedding
    from b200llm import geometry
    geo = lo.GEOMETRIES[name]
    rp = {"rope_type": "llama3", "rope_theta": geo.rope_theta, "factor": 32.0, "low_freq_factor": 1.0,
          "high_freq_factor": 4.0, "original_max_position_embeddings": 8192} if geo.rope_llama3 else \
         {"rope_type": "default", "rope_theta": geo.rope_theta}
    cfg = LlamaConfig(vocab_size=geo.vocab, hidden_size=geo.d, intermediate_size=geo.ffn, num_hidden_layers=1,
                      num_attention_heads=geo.n_q, num_key_value_heads=geo.n_kv, head_dim=geo.head_dim,
                      max_position_embeddings=131072, rope_parameters=rp)
    ref = LlamaRotaryEmbedding(cfg).inv_freq
    assert torch.equal(ref, lo.inv_freq(geo))
    assert torch.equal(ref, geometry.inv_freq(geometry.resolve(name)))


@pytest.mark.parametrize("name", ["tiny", "small"])
def test_golden_reproduces(name, golden_dir):
    """tests/golden/*.npz are outputs of this oracle (oracle/make_golden.py)."""
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    geo = lo.GEOMETRIES[name]
    sd = lo.make_weights(geo, int(g["seed"]), sigma=float(g["sigma"]), embed_sigma=float(g["embed_sigma"]))
    orc = lo.LlamaOracle(geo, sd, kv_dtype=torch.bfloat16)
    toks, lg, mar = orc.greedy(g["prompt"], 64)
    assert toks == g["tokens"].tolist()
    assert np.allclose(lg[:, g["probe_ids"]].numpy(), g["probe_kvbf16"], atol=1e-5)
    assert len(set(toks)) >= 16 and min(mar) >= 0.004       # non-degenerate; margins >> fp32 noise (~1e-5)


def test_golden_files_present(golden_dir):
    for name in ("tiny", "small", "llama3.2_1b", "llama3.2", "duckdb-nsql"):
        g = np.load(os.path.join(golden_dir, name + ".npz"))
        assert g["tokens"].shape == (64,) and g["probe_kvbf16"].shape == (64, 256)


def test_sampler_oracle_properties():
    rng = np.random.default_rng(0)
    lg = rng.standard_normal(5000).astype(np.float32)
    assert osamp.sample(lg, 0.0, 40, 0.9, 1.0, 1, 0)[0] == int(np.argmax(lg))
    tok, ids, probs = osamp.sample(lg, 0.8, 40, 0.9, 1.0, 1, 0)
    assert tok in ids and len(ids) <= 40 and np.all(np.diff(probs) <= 0)
    # penalty pushes a repeated positive logit down
    top = int(np.argmax(lg))
    hist = np.full(64, -1); hist[3] = top
    lg2 = lg.copy(); lg2[top] = lg.max()
    t2 = osamp.sample(lg2, 0.0, 40, 0.9, 5.0, 1, 0, history=hist)[0]
    assert t2 != top
    # draws follow the kept distribution
    counts = {}
    for c in range(2000):
        t, ids, probs = osamp.sample(lg, 1.0, 5, 1.0, 1.0, 42, c)
        counts[t] = counts.get(t, 0) + 1
    p = probs / probs.sum()
    for i, t in enumerate(ids):
        assert abs(counts.get(int(t), 0) / 2000 - p[i]) < 0.05
