"""Model registry: the `model=` strings of the reference's ollama.generate calls ->
Llama geometry (values from the public model cards).

Reference call sites: model="duckdb-nsql" (Flask/app.py:103, FastAPI/app.py:86),
model="llama3.2" (Flask/app.py:161, FastAPI/app.py:106); the evaluation script also
names "mistral" (Model_Evaluation_&_Comparision.py:69,83), which the apps never use and
which is not a target model.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, replace
from typing import Dict

import torch


@dataclass(frozen=True)
class Geometry:
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
    rope_llama3: bool          # llama3 frequency scaling: factor 32, low 1, high 4, original ctx 8192
    rms_eps: float = 1e-5
    eos_ids: tuple = ()        # no tokenizer assets exist offline; real ids arrive with a checkpoint

    @property
    def params(self) -> int:
        per_layer = (self.n_q + 2 * self.n_kv) * self.head_dim * self.d + self.d * self.n_q * self.head_dim \
            + 3 * self.ffn * self.d + 2 * self.d
        return self.vocab * self.d * (1 if self.tied else 2) + self.n_layers * per_layer + self.d

    def weight_bytes_per_step(self) -> int:
        """bf16 bytes every decode step must read (all layer weights + lm_head
        matrix + final norm; the embedding gather is excluded)."""
        per_layer = (self.n_q + 2 * self.n_kv) * self.head_dim * self.d + self.d * self.n_q * self.head_dim \
            + 3 * self.ffn * self.d + 2 * self.d
        return 2 * (self.n_layers * per_layer + self.vocab * self.d + self.d)

    def kv_bytes_per_token(self) -> int:
        return 2 * self.n_layers * self.n_kv * self.head_dim * 2

    def step_bytes(self, bs: int, mean_ctx: float) -> float:
        """Algorithmic HBM bytes of one decode step: B(bs, T) = weights + bs*T*kv_tok + bs*kv_tok."""
        return self.weight_bytes_per_step() + bs * mean_ctx * self.kv_bytes_per_token() + bs * self.kv_bytes_per_token()


REGISTRY: Dict[str, Geometry] = {
    "tiny": Geometry("tiny", 256, 2, 4, 2, 64, 512, 1024, True, 5e5, True),
    "small": Geometry("small", 512, 3, 8, 2, 128, 1536, 4096, False, 1e4, False),
    "llama3.2:1b": Geometry("llama3.2:1b", 2048, 16, 32, 8, 64, 8192, 128256, True, 5e5, True),
    "llama3.2": Geometry("llama3.2", 3072, 28, 24, 8, 128, 8192, 128256, True, 5e5, True),
    "duckdb-nsql": Geometry("duckdb-nsql", 4096, 32, 32, 32, 128, 11008, 32000, False, 1e4, False),
}
ALIASES = {"llama3.2:latest": "llama3.2", "llama3.2:3b": "llama3.2", "duckdb-nsql:latest": "duckdb-nsql",
           "duckdb-nsql:7b": "duckdb-nsql"}
# CodeLlama-flavoured variant of the 7B geometry (north_star: "CodeLlama-7B-class")
REGISTRY["duckdb-nsql:codellama"] = replace(REGISTRY["duckdb-nsql"], name="duckdb-nsql:codellama", vocab=32016,
                                            rope_theta=1e6)


def resolve(model: str) -> Geometry:
    key = ALIASES.get(model, model)
    if key not in REGISTRY:
        raise KeyError(model)
    return REGISTRY[key]


def inv_freq(geo: Geometry) -> torch.Tensor:
    """fp32 [head_dim/2] RoPE inverse frequencies with the llama3 wavelength-dependent
    scaling where the geometry asks for it (same table the oracle derives from HF)."""
    dim = geo.head_dim
    inv = 1.0 / (geo.rope_theta ** (torch.arange(0, dim, 2, dtype=torch.int64).to(torch.float) / dim))
    if not geo.rope_llama3:
        return inv
    factor, low, high, old_ctx = 32.0, 1.0, 4.0, 8192
    wavelen = 2 * math.pi / inv
    scaled = torch.where(wavelen > old_ctx / low, inv / factor, inv)
    smooth = (old_ctx / wavelen - low) / (high - low)
    blended = (1 - smooth) * scaled / factor + smooth * scaled
    medium = ~(wavelen < old_ctx / high) * ~(wavelen > old_ctx / low)
    return torch.where(medium, blended, scaled)
