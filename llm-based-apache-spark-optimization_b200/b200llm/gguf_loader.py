"""GGUF checkpoints -> the engine's packed bf16 weights (second next row).

The reference's models are Ollama blobs: `duckdb-nsql` (Q4_0) and `llama3.2` (Q4_K_M, i.e. Q4_K with
Q6_K for a few tensors) pulled by `ollama pull` (README of the reference, Model_Evaluation_&_
Comparision.py:69-83).  A blob is a GGUF v3 file; this module reads it without the `gguf` package
(that package is only the CHECKER in tests/test_gguf.py):

  * header, metadata key/values and tensor directory (little endian, `general.alignment`);
  * de-quantisation of F32 / F16 / BF16 / Q8_0 / Q4_0 / Q4_K / Q6_K blocks to fp32 (vectorised numpy,
    one tensor at a time), rounded to bf16 on the device -- the engine computes in bf16 x fp32, so
    a quantised checkpoint is expanded once at load time;
  * llama.cpp tensor names -> HF names, with the q/k row permutation of llama.cpp's converter
    undone (it interleaves the two rotary halves of every head; HF keeps them as halves);
  * geometry from the `llama.*` metadata (llama3 RoPE scaling is recognised by the `rope_freqs`
    tensor the converter emits for it).

`load(path, device)` returns (Geometry, HF-named state dict of device bf16 tensors) ready for
`weights.pack`; `register(name, path)` makes the model available to `ollama.generate(model=name)`.
The GGUF tokenizer tables (`tokenizer.ggml.*`) become the model's tokenizer: load_tokenizer() hands
the metadata to tokenizer.from_gguf_metadata (SentencePiece BPE or byte-level BPE).
"""
from __future__ import annotations

import struct
from dataclasses import dataclass
from typing import Any, BinaryIO, Dict, List, Tuple

import numpy as np

GGUF_MAGIC = 0x46554747  # "GGUF"
# ggml type id -> (name, elements per block, bytes per block)
GGML_TYPES = {0: ("F32", 1, 4), 1: ("F16", 1, 2), 2: ("Q4_0", 32, 18), 8: ("Q8_0", 32, 34), 12: ("Q4_K", 256, 144),
              14: ("Q6_K", 256, 210), 30: ("BF16", 1, 2)}
_SCALAR = {0: "<B", 1: "<b", 2: "<H", 3: "<h", 4: "<I", 5: "<i", 6: "<f", 7: "<?", 10: "<Q", 11: "<q", 12: "<d"}
_STRING, _ARRAY = 8, 9


@dataclass
class TensorInfo:
    name: str
    shape: Tuple[int, ...]   # row-major (slowest first), i.e. ggml's ne[] reversed
    ggml_type: int
    offset: int              # from the start of the data section

    @property
    def n_elements(self) -> int:
        return int(np.prod(self.shape, dtype=np.int64))

    @property
    def n_bytes(self) -> int:
        _, per_block, block_bytes = GGML_TYPES[self.ggml_type]
        return self.n_elements // per_block * block_bytes


def _read(f: BinaryIO, fmt: str):
    return struct.unpack(fmt, f.read(struct.calcsize(fmt)))[0]


def _read_str(f: BinaryIO) -> str:
    n = _read(f, "<Q")
    return f.read(n).decode("utf-8", "replace")


def _read_value(f: BinaryIO, vtype: int):
    if vtype in _SCALAR:
        return _read(f, _SCALAR[vtype])
    if vtype == _STRING:
        return _read_str(f)
    if vtype == _ARRAY:
        etype, n = _read(f, "<I"), _read(f, "<Q")
        if etype in _SCALAR and etype != 7:
            dt = np.dtype(_SCALAR[etype])
            return np.frombuffer(f.read(dt.itemsize * n), dtype=dt)
        return [_read_value(f, etype) for _ in range(n)]
    raise ValueError(f"GGUF: unknown metadata value type {vtype}")


class GGUFFile:
    """Parsed directory of one GGUF file; tensor bytes are memory-mapped on demand."""

    def __init__(self, path: str):
        self.path = path
        with open(path, "rb") as f:
            if _read(f, "<I") != GGUF_MAGIC:
                raise ValueError(f"{path}: not a GGUF file")
            self.version = _read(f, "<I")
            if self.version not in (2, 3):
                raise ValueError(f"{path}: GGUF version {self.version} not supported (2 and 3 are)")
            n_tensors, n_kv = _read(f, "<Q"), _read(f, "<Q")
            self.metadata: Dict[str, Any] = {}
            for _ in range(n_kv):
                key = _read_str(f)
                self.metadata[key] = _read_value(f, _read(f, "<I"))
            self.tensors: Dict[str, TensorInfo] = {}
            for _ in range(n_tensors):
                name = _read_str(f)
                nd = _read(f, "<I")
                ne = [_read(f, "<Q") for _ in range(nd)]
                gt, off = _read(f, "<I"), _read(f, "<Q")
                if gt not in GGML_TYPES:
                    raise ValueError(f"{path}: tensor {name} has ggml type {gt}, supported: "
                                     f"{sorted(v[0] for v in GGML_TYPES.values())}")
                self.tensors[name] = TensorInfo(name, tuple(reversed(ne)), gt, off)
            align = int(self.metadata.get("general.alignment", 32))
            self.data_start = (f.tell() + align - 1) // align * align
        self._map = np.memmap(path, dtype=np.uint8, mode="r")

    def raw(self, name: str) -> np.ndarray:
        t = self.tensors[name]
        a = self.data_start + t.offset
        return self._map[a:a + t.n_bytes]

    def dequantize(self, name: str) -> np.ndarray:
        t = self.tensors[name]
        return dequantize(self.raw(name), t.ggml_type).reshape(t.shape)


# ---------------------------------------------------------------- block formats (ggml-quants.c) -> fp32
def _f16(b: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(b).view(np.float16).astype(np.float32)


def _deq_q8_0(raw: np.ndarray) -> np.ndarray:
    blk = raw.reshape(-1, 34)
    d = _f16(blk[:, :2])                                      # [nb, 1]
    q = blk[:, 2:].view(np.int8).astype(np.float32)           # [nb, 32]
    return (q * d).reshape(-1)


def _deq_q4_0(raw: np.ndarray) -> np.ndarray:
    blk = raw.reshape(-1, 18)
    d = _f16(blk[:, :2])
    qs = blk[:, 2:]
    lo = (qs & 0x0F).astype(np.int8) - 8                      # elements 0..15
    hi = (qs >> 4).astype(np.int8) - 8                        # elements 16..31
    return (np.concatenate([lo, hi], axis=1).astype(np.float32) * d).reshape(-1)


def _deq_q4_k(raw: np.ndarray) -> np.ndarray:
    blk = raw.reshape(-1, 144)
    nb = blk.shape[0]
    d, dmin = _f16(blk[:, 0:2]), _f16(blk[:, 2:4])            # [nb, 1]
    sc = blk[:, 4:16].astype(np.uint8)                        # 12 bytes: 8 x (6-bit scale, 6-bit min)
    qs = blk[:, 16:]                                          # [nb, 128]
    scale = np.empty((nb, 8), np.float32)
    mn = np.empty((nb, 8), np.float32)
    for j in range(8):                                        # get_scale_min_k4
        if j < 4:
            s_, m_ = sc[:, j] & 63, sc[:, j + 4] & 63
        else:
            s_ = (sc[:, j + 4] & 0x0F) | ((sc[:, j - 4] >> 6) << 4)
            m_ = (sc[:, j + 4] >> 4) | ((sc[:, j] >> 6) << 4)
        scale[:, j], mn[:, j] = s_, m_
    scale *= d
    mn *= dmin
    q = qs.reshape(nb, 4, 32)
    out = np.empty((nb, 8, 32), np.float32)
    out[:, 0::2] = (q & 0x0F).astype(np.float32)              # sub-block 2j: low nibbles of byte group j
    out[:, 1::2] = (q >> 4).astype(np.float32)                # sub-block 2j+1: high nibbles
    return (out * scale[:, :, None] - mn[:, :, None]).reshape(-1)


def _deq_q6_k(raw: np.ndarray) -> np.ndarray:
    blk = raw.reshape(-1, 210)
    nb = blk.shape[0]
    ql = blk[:, :128].reshape(nb, 2, 64)
    qh = blk[:, 128:192].reshape(nb, 2, 32)
    sc = blk[:, 192:208].view(np.int8).astype(np.float32).reshape(nb, 2, 8)
    d = _f16(blk[:, 208:210])                                 # [nb, 1]
    l0, l1 = ql[:, :, :32], ql[:, :, 32:]
    q = np.empty((nb, 2, 4, 32), np.int16)                    # [block, half, group of 32, l]
    q[:, :, 0] = (l0 & 0x0F) | (((qh >> 0) & 3) << 4)
    q[:, :, 1] = (l1 & 0x0F) | (((qh >> 2) & 3) << 4)
    q[:, :, 2] = (l0 >> 4) | (((qh >> 4) & 3) << 4)
    q[:, :, 3] = (l1 >> 4) | (((qh >> 6) & 3) << 4)
    q = q.astype(np.float32) - 32.0
    # scale index of element (group g, l) = 2 g + l // 16
    s = sc.reshape(nb, 2, 4, 2)[:, :, :, :, None]             # [nb, half, g, l // 16, 1]
    y = q.reshape(nb, 2, 4, 2, 16) * s
    return (y * d[:, :, None, None, None]).reshape(-1)


def dequantize(raw: np.ndarray, ggml_type: int) -> np.ndarray:
    """Flat fp32 array of the tensor's elements."""
    name = GGML_TYPES[ggml_type][0]
    raw = np.asarray(raw, dtype=np.uint8)
    if name == "F32":
        return np.ascontiguousarray(raw).view(np.float32).copy()
    if name == "F16":
        return _f16(raw)
    if name == "BF16":
        return (np.ascontiguousarray(raw).view(np.uint16).astype(np.uint32) << 16).view(np.float32)
    return {"Q8_0": _deq_q8_0, "Q4_0": _deq_q4_0, "Q4_K": _deq_q4_k, "Q6_K": _deq_q6_k}[name](raw)


# ---------------------------------------------------------------- llama.cpp names / layout -> HF
_LAYER_MAP = {"attn_q": "self_attn.q_proj", "attn_k": "self_attn.k_proj", "attn_v": "self_attn.v_proj",
              "attn_output": "self_attn.o_proj", "ffn_gate": "mlp.gate_proj", "ffn_up": "mlp.up_proj",
              "ffn_down": "mlp.down_proj", "attn_norm": "input_layernorm", "ffn_norm": "post_attention_layernorm"}
_GLOBAL_MAP = {"token_embd.weight": "model.embed_tokens.weight", "output_norm.weight": "model.norm.weight",
               "output.weight": "lm_head.weight"}


def hf_name(gguf_name: str):
    """HF parameter name of a llama.cpp tensor name, or None for tensors the engine derives itself."""
    if gguf_name in _GLOBAL_MAP:
        return _GLOBAL_MAP[gguf_name]
    parts = gguf_name.split(".")
    if len(parts) == 4 and parts[0] == "blk" and parts[2] in _LAYER_MAP and parts[3] == "weight":
        return f"model.layers.{int(parts[1])}.{_LAYER_MAP[parts[2]]}.weight"
    return None   # rope_freqs.weight (recomputed from the metadata), anything unknown


def unpermute_qk(w: np.ndarray, n_head: int) -> np.ndarray:
    """Inverse of llama.cpp's convert-time permute(): rows (2i, 2i+1) of a head -> rows (i, i + hd/2)."""
    rows = w.shape[0]
    return w.reshape(n_head, rows // n_head // 2, 2, *w.shape[1:]).swapaxes(1, 2).reshape(w.shape)


def geometry_from_metadata(md: Dict[str, Any], tensors: Dict[str, TensorInfo], name: str):
    from .geometry import Geometry
    arch = md.get("general.architecture", "llama")
    if arch != "llama":
        raise ValueError(f"GGUF architecture '{arch}' is not a Llama-family model")
    g = lambda k, default=None: md.get(f"{arch}.{k}", default)  # noqa: E731
    d, n_q = int(g("embedding_length")), int(g("attention.head_count"))
    n_kv = int(g("attention.head_count_kv", n_q))
    head_dim = int(g("attention.key_length", d // n_q))
    vocab = int(tensors["token_embd.weight"].shape[0])
    return Geometry(name, d, int(g("block_count")), n_q, n_kv, head_dim, int(g("feed_forward_length")), vocab,
                    tied="output.weight" not in tensors, rope_theta=float(g("rope.freq_base", 10000.0)),
                    rope_llama3="rope_freqs.weight" in tensors,
                    rms_eps=float(g("attention.layer_norm_rms_epsilon", 1e-5)))


def dequantize_on_device(raw: np.ndarray, ggml_type: int, shape, device):
    """bf16 tensor of `shape` on the CUDA `device`: the raw blocks are uploaded as they lie in the file and expanded by
    the library's kernel (csrc/dequant.cu), bit-equal to rounding `dequantize` to bf16."""
    import torch
    from . import _lib
    n = int(np.prod(shape))
    import warnings
    with warnings.catch_warnings():   # the blocks come straight out of the read-only file mapping and are only read
        warnings.simplefilter("ignore", UserWarning)
        blocks = torch.from_numpy(np.ascontiguousarray(np.asarray(raw, dtype=np.uint8))).to(device)
    out = torch.empty(n, dtype=torch.bfloat16, device=device)
    with torch.cuda.device(out.device):
        _lib.check(_lib.lib.b200llm_op_dequant_bf16(blocks.data_ptr(), int(ggml_type), n, out.data_ptr(),
                                                    torch.cuda.current_stream(out.device).cuda_stream))
    return out.view(*shape)


def unpermute_qk_t(w, n_head: int):
    """unpermute_qk for a torch tensor (the device path of `load`)."""
    out, cols = w.shape
    return w.view(n_head, out // n_head // 2, 2, cols).transpose(1, 2).reshape(out, cols)


def load(path: str, device, name: str = "") -> Tuple[Any, Dict[str, Any]]:
    """(Geometry, {HF name: bf16 tensor on `device`}) of a Llama-family GGUF file.  On a CUDA device the tensors are
    expanded there (B200LLM_GGUF_HOST_DEQUANT=1: in numpy on the host, the checker's path)."""
    import os
    import torch
    gf = GGUFFile(path)
    geo = geometry_from_metadata(gf.metadata, gf.tensors, name or gf.metadata.get("general.name", "gguf-model"))
    sd: Dict[str, Any] = {}
    on_device = torch.device(device).type == "cuda" and not os.environ.get("B200LLM_GGUF_HOST_DEQUANT")
    for tname, info in gf.tensors.items():
        hf = hf_name(tname)
        if hf is None:
            continue
        if on_device:
            w = dequantize_on_device(gf.raw(tname), info.ggml_type, info.shape, device)
            if tname.endswith("attn_q.weight"):
                w = unpermute_qk_t(w, geo.n_q).contiguous()
            elif tname.endswith("attn_k.weight"):
                w = unpermute_qk_t(w, geo.n_kv).contiguous()
            sd[hf] = w
            continue
        w = gf.dequantize(tname)
        if tname.endswith("attn_q.weight"):
            w = unpermute_qk(w, geo.n_q)
        elif tname.endswith("attn_k.weight"):
            w = unpermute_qk(w, geo.n_kv)
        sd[hf] = torch.from_numpy(np.ascontiguousarray(w)).to(device=device, dtype=torch.bfloat16)
    missing = [k for k in expected_names(geo) if k not in sd]
    if missing:
        raise ValueError(f"{path}: missing tensors {missing[:4]}{'...' if len(missing) > 4 else ''}")
    return geo, sd


def load_tokenizer(path: str):
    """The tokenizer described by the file's `tokenizer.ggml.*` tables, or None when it has no tokenizer model
    (a weights-only file: the caller keeps its stand-in)."""
    from . import tokenizer
    md = GGUFFile(path).metadata
    if "tokenizer.ggml.model" not in md:
        return None
    return tokenizer.from_gguf_metadata(md)


def expected_names(geo) -> List[str]:
    names = ["model.embed_tokens.weight", "model.norm.weight"] + ([] if geo.tied else ["lm_head.weight"])
    for l in range(geo.n_layers):
        names += [f"model.layers.{l}.{v}.weight" for v in _LAYER_MAP.values()]
    return names


def register(model_name: str, path: str) -> None:
    """Make `ollama.generate(model=model_name)` serve this checkpoint (loaded on first use)."""
    from . import runtime
    runtime.register_checkpoint(model_name, path)
