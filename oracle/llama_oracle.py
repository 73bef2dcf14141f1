"""oracle/llama_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement (plain torch, fp32 activations) of the hot path BASELINE.json's
north_star names: Llama-family prefill + autoregressive greedy decode, the
computation behind every `ollama.generate(model=, system=, prompt=)` call of the
reference (call sites: /root/reference/Flask/app.py:102-107,160-166;
FastAPI/app.py:85-90,105-111; Model_Evaluation_&_Comparision.py:23-45,114-120).

PARITY UNPINNED by the reference: the reference implements none of this path
(it lives in Ollama + llama.cpp, un-vendored and un-versioned; SURVEY.md
section 8c) and has no tests or golden vectors for it.  The algorithm restated
here is the published Llama architecture as implemented by HF transformers
5.5.0 `transformers/models/llama/modeling_llama.py` (LlamaRMSNorm,
rotate_half/apply_rotary_pos_emb, LlamaAttention with GQA repeat_kv, LlamaMLP)
and `modeling_rope_utils._compute_llama3_parameters`; tests/test_oracle.py pins
this file against that HF implementation on the same weights, and
tests/golden/*.npz are outputs of this file (oracle/make_golden.py).

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch

from . import weightgen


@dataclass(frozen=True)
class OracleGeometry:
    """Model-card geometry."""
    name: str
    d: int
    n_layers: int
    n_q: int
    n_kv: int
    head_dim: int
    ffn: int
    vocab: int
    tied: bool
    rope_theta: float
    rope_llama3: bool                 # factor 32, low 1, high 4, orig ctx 8192
    rms_eps: float = 1e-5


GEOMETRIES = {
    "tiny": OracleGeometry("tiny", 256, 2, 4, 2, 64, 512, 1024, True, 5e5, True),
    "small": OracleGeometry("small", 512, 3, 8, 2, 128, 1536, 4096, False, 1e4, False),
    "llama3.2:1b": OracleGeometry("llama3.2:1b", 2048, 16, 32, 8, 64, 8192, 128256, True, 5e5, True),
    "llama3.2": OracleGeometry("llama3.2", 3072, 28, 24, 8, 128, 8192, 128256, True, 5e5, True),
    "duckdb-nsql": OracleGeometry("duckdb-nsql", 4096, 32, 32, 32, 128, 11008, 32000, False, 1e4, False),
}


def inv_freq(geo: OracleGeometry) -> torch.Tensor:
    """RoPE inverse frequencies, fp32 [head_dim/2].

    Follows transformers 5.5.0 modeling_rope_utils._compute_default_rope_parameters
    and _compute_llama3_parameters (factor 32, low_freq_factor 1, high_freq_factor 4,
    original_max_position_embeddings 8192 -- the Llama-3.2 model cards)."""
    dim = geo.head_dim
    inv = 1.0 / (geo.rope_theta ** (torch.arange(0, dim, 2, dtype=torch.int64).to(torch.float) / dim))
    if not geo.rope_llama3:
        return inv
    factor, low, high, old_ctx = 32.0, 1.0, 4.0, 8192
    low_wl = old_ctx / low
    high_wl = old_ctx / high
    wavelen = 2 * math.pi / inv
    inv_l = torch.where(wavelen > low_wl, inv / factor, inv)
    smooth = (old_ctx / wavelen - low) / (high - low)
    smoothed = (1 - smooth) * inv_l / factor + smooth * inv_l
    is_medium = ~(wavelen < high_wl) * ~(wavelen > low_wl)
    return torch.where(is_medium, smoothed, inv_l)


def make_weights(geo: OracleGeometry, seed: int, sigma: float = 0.02, gain_sigma: float = 0.02,
                 embed_sigma: Optional[float] = None) -> Dict[str, torch.Tensor]:
    """HF-named bf16 state dict from the seeded stream (SURVEY.md 8(d) "Weights").

    Every matrix ~ N(0, sigma^2) (Irwin-Hall n=4 approximation), norm gains
    1 + N(0, gain_sigma^2), cast to bf16 once; oracle and engine consume the
    identical bf16 bits."""
    es = sigma if embed_sigma is None else embed_sigma
    g = weightgen
    sd: Dict[str, torch.Tensor] = {}
    sd["model.embed_tokens.weight"] = g.gen_bf16((geo.vocab, geo.d), g.tensor_seed(seed, g.GLOBAL_TID["embed_tokens"]), es)
    sd["model.norm.weight"] = g.gen_bf16((geo.d,), g.tensor_seed(seed, g.GLOBAL_TID["norm"]), gain_sigma, 1.0)
    if not geo.tied:
        sd["lm_head.weight"] = g.gen_bf16((geo.vocab, geo.d), g.tensor_seed(seed, g.GLOBAL_TID["lm_head"]), sigma)
    shapes = {
        "q_proj": (geo.n_q * geo.head_dim, geo.d), "k_proj": (geo.n_kv * geo.head_dim, geo.d),
        "v_proj": (geo.n_kv * geo.head_dim, geo.d), "o_proj": (geo.d, geo.n_q * geo.head_dim),
        "gate_proj": (geo.ffn, geo.d), "up_proj": (geo.ffn, geo.d), "down_proj": (geo.d, geo.ffn),
    }
    for l in range(geo.n_layers):
        for kind, tid in g.KIND.items():
            s = g.tensor_seed(seed, (l + 1) * 16 + tid)
            if kind in shapes:
                sub = "self_attn" if kind in ("q_proj", "k_proj", "v_proj", "o_proj") else "mlp"
                sd[f"model.layers.{l}.{sub}.{kind}.weight"] = g.gen_bf16(shapes[kind], s, sigma)
            else:
                sd[f"model.layers.{l}.{kind}.weight"] = g.gen_bf16((geo.d,), s, gain_sigma, 1.0)
    return sd


def _rms_norm(x: torch.Tensor, gain: torch.Tensor, eps: float) -> torch.Tensor:
    # modeling_llama.LlamaRMSNorm.forward
    var = x.pow(2).mean(-1, keepdim=True)
    return gain * (x * torch.rsqrt(var + eps))


def _rotate_half(x: torch.Tensor) -> torch.Tensor:
    # modeling_llama.rotate_half
    h = x.shape[-1] // 2
    return torch.cat((-x[..., h:], x[..., :h]), dim=-1)


@dataclass
class OracleState:
    k: List[torch.Tensor] = field(default_factory=list)   # per layer [n_kv, T, hd] fp32 (values maybe bf16-rounded)
    v: List[torch.Tensor] = field(default_factory=list)
    pos: int = 0


class LlamaOracle:
    """fp32 forward over bf16-rounded weights.

    kv_dtype=torch.bfloat16 models the engine's bf16 paged KV cache (SURVEY.md
    8(d): kv_tok counts 2 bytes per element): K (after RoPE) and V are rounded
    to bf16 when stored.  kv_dtype=torch.float32 is the un-rounded truth used
    for the 1e-2 logit tolerance.  act_dtype=torch.bfloat16 additionally
    rounds each matmul input to bf16 (what the tensor-core batch path does)."""

    def __init__(self, geo: OracleGeometry, sd: Dict[str, torch.Tensor], kv_dtype=torch.bfloat16,
                 act_dtype=torch.float32, compute_dtype=torch.float32):
        self.geo = geo
        self.kv_dtype = kv_dtype
        self.act_dtype = act_dtype
        self.cd = compute_dtype
        self.w = {k: v.to(compute_dtype) for k, v in sd.items()}
        if geo.tied:
            self.w["lm_head.weight"] = self.w["model.embed_tokens.weight"]
        self.inv_freq = inv_freq(geo)

    def new_state(self) -> OracleState:
        return OracleState([None] * self.geo.n_layers, [None] * self.geo.n_layers, 0)

    def _mm(self, x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
        if self.act_dtype != torch.float32:
            x = x.to(self.act_dtype)
        return (x.to(self.cd) @ w.t()).float()

    @torch.no_grad()
    def forward(self, ids: torch.Tensor, st: OracleState, all_logits: bool = False) -> torch.Tensor:
        """ids int64 [T_new]; appends to st; returns fp32 logits [vocab] of the last
        token (or [T_new, vocab])."""
        geo, w = self.geo, self.w
        T = ids.numel()
        h = w["model.embed_tokens.weight"][ids].float()                       # [T, d]
        pos = torch.arange(st.pos, st.pos + T, dtype=torch.float32)
        freqs = pos[:, None] * self.inv_freq[None, :]                         # LlamaRotaryEmbedding.forward
        emb = torch.cat((freqs, freqs), dim=-1)
        cos, sin = emb.cos()[None], emb.sin()[None]                           # [1, T, hd]
        rep = geo.n_q // geo.n_kv
        for l in range(geo.n_layers):
            p = f"model.layers.{l}."
            x = _rms_norm(h, w[p + "input_layernorm.weight"].float(), geo.rms_eps)
            q = self._mm(x, w[p + "self_attn.q_proj.weight"]).view(T, geo.n_q, geo.head_dim).transpose(0, 1)
            k = self._mm(x, w[p + "self_attn.k_proj.weight"]).view(T, geo.n_kv, geo.head_dim).transpose(0, 1)
            v = self._mm(x, w[p + "self_attn.v_proj.weight"]).view(T, geo.n_kv, geo.head_dim).transpose(0, 1)
            q = q * cos + _rotate_half(q) * sin                                # apply_rotary_pos_emb
            k = k * cos + _rotate_half(k) * sin
            k = k.to(self.kv_dtype).float()
            v = v.to(self.kv_dtype).float()
            st.k[l] = k if st.k[l] is None else torch.cat((st.k[l], k), dim=1)
            st.v[l] = v if st.v[l] is None else torch.cat((st.v[l], v), dim=1)
            K = st.k[l].repeat_interleave(rep, dim=0)                          # repeat_kv
            V = st.v[l].repeat_interleave(rep, dim=0)
            att = (q @ K.transpose(1, 2)) / math.sqrt(geo.head_dim)            # [n_q, T, Ttot]
            Ttot = K.shape[1]
            qpos = torch.arange(st.pos, st.pos + T)[:, None]
            mask = torch.arange(Ttot)[None, :] > qpos
            att = att.masked_fill(mask[None], float("-inf"))
            att = torch.softmax(att, dim=-1, dtype=torch.float32)
            o = (att @ V).transpose(0, 1).reshape(T, geo.n_q * geo.head_dim)
            h = h + self._mm(o, w[p + "self_attn.o_proj.weight"])
            x = _rms_norm(h, w[p + "post_attention_layernorm.weight"].float(), geo.rms_eps)
            gate = self._mm(x, w[p + "mlp.gate_proj.weight"])
            up = self._mm(x, w[p + "mlp.up_proj.weight"])
            h = h + self._mm(torch.nn.functional.silu(gate) * up, w[p + "mlp.down_proj.weight"])
        st.pos += T
        x = _rms_norm(h if all_logits else h[-1:], w["model.norm.weight"].float(), geo.rms_eps)
        logits = self._mm(x, w["lm_head.weight"])
        return logits if all_logits else logits[0]

    @torch.no_grad()
    def greedy(self, prompt_ids, n_new: int, keep_logits: bool = True):
        """Greedy decode, tie-break = lowest token id (torch.argmax returns the first
        maximal index).  Returns (tokens [n_new], logits [n_new, vocab] or None,
        margins [n_new] = top1 - top2 of each step)."""
        st = self.new_state()
        logits = self.forward(torch.as_tensor(prompt_ids, dtype=torch.int64), st)
        toks, lg, margins = [], [], []
        for _ in range(n_new):
            top2 = torch.topk(logits, 2).values
            margins.append(float(top2[0] - top2[1]))
            t = int(torch.argmax(logits))
            toks.append(t)
            if keep_logits:
                lg.append(logits.clone())
            if len(toks) == n_new:
                break
            logits = self.forward(torch.tensor([t], dtype=torch.int64), st)
        return toks, (torch.stack(lg) if keep_logits else None), margins
