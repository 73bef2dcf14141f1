"""Weights: the engine's packed HBM layout and seeded random initialisation.

The reference's weights are GGUF blobs inside Ollama (not in the repo, not on this box),
and BASELINE.json fixes "random-init weights of that architecture", so weights come from
a seeded integer-only stream generated ON THE DEVICE by b200llm_op_fill_weights
(csrc/misc.cu).  HF tensor names are kept for the logical tensors so a real checkpoint
can be packed by the same `pack_layer` later.

Packed layout (DESIGN.md "Data layout in HBM"), all bf16 row-major:
  W_QKV     [(n_q + 2 n_kv)/tp * hd, d]  q heads, then k heads, then v heads of this rank;
                                         inside every q and k head the rows are pair-interleaved
                                         (row 2i = dim i, row 2i+1 = dim i + hd/2) so the GEMV
                                         epilogue can apply RoPE to adjacent rows
  W_O       [d, n_q/tp * hd]             column slice of o_proj
  W_GATE_UP [2 ffn/tp, d]                row 2i = gate_i, row 2i+1 = up_i  (SwiGLU epilogue)
  W_DOWN    [d, ffn/tp]                  column slice of down_proj
  LM_HEAD   [vocab/tp, d]                row slice (aliases EMBED when tied and tp == 1)
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

import torch

from . import _lib
from .geometry import Geometry, inv_freq

IRWIN_HALL_STD = 37837.22765
KIND = {"q_proj": 0, "k_proj": 1, "v_proj": 2, "o_proj": 3, "gate_proj": 4, "up_proj": 5, "down_proj": 6,
        "input_layernorm": 7, "post_attention_layernorm": 8}
GLOBAL_TID = {"embed_tokens": 0, "norm": 1, "lm_head": 2}


def tensor_seed(base_seed: int, tid: int) -> int:
    return ((int(base_seed) << 20) + int(tid)) & 0xFFFFFFFFFFFFFFFF


def fill(shape, seed: int, sigma: float, mean: float, device) -> torch.Tensor:
    """bf16 tensor of `shape` on `device`, element i = stream(seed)[i] (device-generated)."""
    t = torch.empty(shape, dtype=torch.bfloat16, device=device)
    _lib.check(_lib.lib.b200llm_op_fill_weights(t.data_ptr(), t.numel(), seed, sigma, mean,
                                                torch.cuda.current_stream(device).cuda_stream))
    return t


@dataclass
class PackedWeights:
    geo: Geometry
    tp_rank: int
    tp_size: int
    tensors: List[Tuple[int, int, torch.Tensor]] = field(default_factory=list)   # (kind, layer, tensor)

    @property
    def ffn_local(self) -> int:
        """MLP units of this rank's shard.  Under tensor parallel the shard is padded with zero units to a multiple of
        64 (duckdb-nsql over 8 ranks: 11008 / 8 = 1376 -> 1408), so that gate/up rows come in whole 128-row tiles and
        down_proj's K in whole 64-wide blocks: the tcgen05 kernels then serve every TP degree.  A zero gate/up unit
        gives silu(0) * 0 = 0 and meets zero down_proj columns, so results are unchanged bit for bit."""
        ffn_l = self.geo.ffn // self.tp_size
        return ffn_l if self.tp_size == 1 else (ffn_l + 63) // 64 * 64

    @property
    def ffn_cfg(self) -> int:
        """The `ffn` to put into b200llm_config (the engine divides it by tp_size)."""
        return self.ffn_local * self.tp_size

    def descs(self):
        arr = (_lib.TensorDesc * len(self.tensors))()
        for i, (kind, layer, t) in enumerate(self.tensors):
            if kind >= _lib.W_QKV_ST:          # stage-major copy: the descriptor carries the logical [N, K]
                rows, cols = t.logical_shape
            else:
                rows, cols = (t.shape[0], t.shape[1]) if t.dim() == 2 else (1, t.shape[0])
            arr[i] = _lib.TensorDesc(kind, layer, t.data_ptr(), rows, cols)
        return arr

    def nbytes(self) -> int:
        seen, total = set(), 0
        for _, _, t in self.tensors:
            if t.data_ptr() not in seen:
                seen.add(t.data_ptr())
                total += t.numel() * t.element_size()
        return total


def k_parts(K: int) -> int:
    p = (K + 2047) // 2048
    while (K // 32) % p:
        p += 1
    return p


def stage_major(w: torch.Tensor) -> torch.Tensor:
    """[N, K] -> [N/8, parts, 8, kc + 8]: the persistent kernel's ring stages, each contiguous and
    with its rows already skewed by 16 bytes (see B200LLM_W_*_ST in include/b200llm.h)."""
    N, K = w.shape
    assert N % 8 == 0 and K % 32 == 0
    parts = k_parts(K)
    kc = K // parts
    out = torch.zeros(N // 8, parts, 8, kc + 8, dtype=w.dtype, device=w.device)
    out[..., :kc] = w.view(N // 8, 8, parts, kc).permute(0, 2, 1, 3)
    out.logical_shape = (N, K)
    return out


_ST_KIND = {_lib.W_QKV: _lib.W_QKV_ST, _lib.W_O: _lib.W_O_ST, _lib.W_GATE_UP: _lib.W_GATE_UP_ST,
            _lib.W_DOWN: _lib.W_DOWN_ST, _lib.W_LM_HEAD: _lib.W_LM_HEAD_ST}


def tile_major(w: torch.Tensor) -> torch.Tensor:
    """[N, K] -> [N/128, K/64, 128, 64]: the tcgen05 persistent kernel's operand tiles, each 16 KB
    tile contiguous and pre-swizzled (16-byte chunk c of row r stored at chunk c ^ (r & 7)), see
    B200LLM_W_*_TM in include/b200llm.h."""
    N, K = w.shape
    assert K % 64 == 0
    if N % 128:     # lm_head of a vocabulary that is no multiple of 128 (32016, or 32000 / 4 ranks): zero rows up to
        w = torch.cat([w, w.new_zeros(128 - N % 128, K)])   # the next tile; the kernel never stores or ranks them
        N = w.shape[0]
    t = w.view(N // 128, 128, K // 64, 8, 8).permute(0, 2, 1, 3, 4)            # [tile, kb, row, chunk, 8]
    idx = torch.arange(8, device=w.device)[None, :] ^ (torch.arange(128, device=w.device)[:, None] & 7)
    out = t[:, :, torch.arange(128, device=w.device)[:, None], idx].contiguous()  # out[r, c] = in[r, c ^ (r & 7)]
    out.logical_shape = (N, K)
    return out


def tile_major_ok(geo, tp_size: int = 1) -> bool:
    """Can every matrix of one rank's shard be cut into [128 x 64] tiles?"""
    nq, nkv, ffn, vocab = geo.n_q // tp_size, geo.n_kv // tp_size, geo.ffn // tp_size, geo.vocab // tp_size
    if tp_size > 1:
        ffn = (ffn + 63) // 64 * 64     # PackedWeights.ffn_local: the shard is padded with zero units
    qkv = (nq + 2 * nkv) * geo.head_dim
    shapes = [(qkv, geo.d), (geo.d, nq * geo.head_dim), (2 * ffn, geo.d), (geo.d, ffn)]
    # the lm_head's rows are padded to the next tile (tile_major), the other matrices must fit as they are
    return all(n % 128 == 0 and k % 64 == 0 for n, k in shapes) and vocab >= 1 and geo.d % 64 == 0


_TM_KIND = {_lib.W_QKV: _lib.W_QKV_TM, _lib.W_O: _lib.W_O_TM, _lib.W_GATE_UP: _lib.W_GATE_UP_TM,
            _lib.W_DOWN: _lib.W_DOWN_TM, _lib.W_LM_HEAD: _lib.W_LM_HEAD_TM}


def add_tile_major(pw: "PackedWeights", start: int = 0) -> None:
    """Append tile-major copies of the matrices pw.tensors[start:] (tcgen05 persistent-kernel layout)."""
    for kind, layer, t in list(pw.tensors[start:]):
        if kind in _TM_KIND:
            pw.tensors.append((_TM_KIND[kind], layer, tile_major(t)))


def drop_row_major(pw: "PackedWeights") -> None:
    """Lean engine: remove the row-major matrices that have a tile-major copy (qkv, o, gate/up, down, and the lm_head unless
    it is the embedding table itself), freeing their memory.  Needs one GPU, a vocabulary that is a multiple of 128 and all
    five kinds of tile-major copies (b200llm_bind_weights checks)."""
    have_tm = {(k, l) for k, l, _ in pw.tensors if k in _TM_KIND.values()}
    if pw.tp_size != 1 or not all((_TM_KIND[k], l) in have_tm for k, l, _ in pw.tensors if k in _TM_KIND):
        raise ValueError("drop_row_major: needs tp_size 1 and a tile-major copy of every matrix (use_megakernel on a geometry that tiles)")
    embed = next(t for k, _, t in pw.tensors if k == _lib.W_EMBED)
    keep = []
    for k, l, t in pw.tensors:
        if k in _TM_KIND and not (k == _lib.W_LM_HEAD and t.data_ptr() == embed.data_ptr()):
            continue
        keep.append((k, l, t))
    pw.tensors[:] = keep


def add_stage_major(pw: "PackedWeights", start: int = 0) -> None:
    """Append stage-major copies of the matrices pw.tensors[start:] (persistent-kernel layout)."""
    for kind, layer, t in list(pw.tensors[start:]):
        if kind in _ST_KIND and pw.tp_size == 1:
            pw.tensors.append((_ST_KIND[kind], layer, stage_major(t)))


def _pair_interleave(hd: int, device) -> torch.Tensor:
    half = hd // 2
    idx = torch.empty(hd, dtype=torch.long, device=device)
    idx[0::2] = torch.arange(half, device=device)
    idx[1::2] = torch.arange(half, device=device) + half
    return idx


def pack_layer(geo: Geometry, layer: int, sd: Dict[str, torch.Tensor], tp_rank: int, tp_size: int,
               out: PackedWeights) -> None:
    """HF-named tensors of one decoder layer -> packed tensors of this rank."""
    p = f"model.layers.{layer}."
    hd, nq_l, nkv_l, ffn_l = geo.head_dim, geo.n_q // tp_size, geo.n_kv // tp_size, geo.ffn // tp_size
    dev = sd[p + "self_attn.q_proj.weight"].device
    perm = _pair_interleave(hd, dev)

    def heads(w, n_local):
        w = w[tp_rank * n_local * hd:(tp_rank + 1) * n_local * hd].view(n_local, hd, -1)
        return w

    q = heads(sd[p + "self_attn.q_proj.weight"], nq_l)[:, perm].reshape(nq_l * hd, geo.d)
    k = heads(sd[p + "self_attn.k_proj.weight"], nkv_l)[:, perm].reshape(nkv_l * hd, geo.d)
    v = heads(sd[p + "self_attn.v_proj.weight"], nkv_l).reshape(nkv_l * hd, geo.d)
    out.tensors.append((_lib.W_QKV, layer, torch.cat((q, k, v), 0).contiguous()))
    out.tensors.append((_lib.W_O, layer,
                        sd[p + "self_attn.o_proj.weight"][:, tp_rank * nq_l * hd:(tp_rank + 1) * nq_l * hd].contiguous()))
    g = sd[p + "mlp.gate_proj.weight"][tp_rank * ffn_l:(tp_rank + 1) * ffn_l]
    u = sd[p + "mlp.up_proj.weight"][tp_rank * ffn_l:(tp_rank + 1) * ffn_l]
    dn = sd[p + "mlp.down_proj.weight"][:, tp_rank * ffn_l:(tp_rank + 1) * ffn_l]
    pad = out.ffn_local - ffn_l if tp_size > 1 else 0   # zero MLP units that round the shard up to whole tiles (PackedWeights.ffn_local)
    if pad > 0:
        g, u = torch.cat((g, g.new_zeros(pad, geo.d))), torch.cat((u, u.new_zeros(pad, geo.d)))
        dn = torch.cat((dn, dn.new_zeros(geo.d, pad)), 1)
    out.tensors.append((_lib.W_GATE_UP, layer, torch.stack((g, u), 1).reshape(2 * (ffn_l + pad), geo.d).contiguous()))
    out.tensors.append((_lib.W_DOWN, layer, dn.contiguous()))
    out.tensors.append((_lib.W_ATTN_NORM, layer, sd[p + "input_layernorm.weight"].contiguous()))
    out.tensors.append((_lib.W_MLP_NORM, layer, sd[p + "post_attention_layernorm.weight"].contiguous()))


def pack_globals(geo: Geometry, sd: Dict[str, torch.Tensor], tp_rank: int, tp_size: int, out: PackedWeights) -> None:
    emb = sd["model.embed_tokens.weight"].contiguous()
    vocab_l = geo.vocab // tp_size
    head = emb if geo.tied else sd["lm_head.weight"]
    lm = head if tp_size == 1 else head[tp_rank * vocab_l:(tp_rank + 1) * vocab_l].contiguous()
    out.tensors.append((_lib.W_EMBED, -1, emb))
    out.tensors.append((_lib.W_LM_HEAD, -1, lm.contiguous()))
    out.tensors.append((_lib.W_FINAL_NORM, -1, sd["model.norm.weight"].contiguous()))
    out.tensors.append((_lib.W_INV_FREQ, -1, inv_freq(geo).to(emb.device).contiguous()))


def pack(geo: Geometry, sd: Dict[str, torch.Tensor], tp_rank: int = 0, tp_size: int = 1,
         stage_major_copies: bool = False) -> PackedWeights:
    """Full HF-named state dict (device tensors, bf16) -> packed weights of this rank."""
    out = PackedWeights(geo, tp_rank, tp_size)
    pack_globals(geo, sd, tp_rank, tp_size, out)
    for l in range(geo.n_layers):
        pack_layer(geo, l, sd, tp_rank, tp_size, out)
    if stage_major_copies:
        add_stage_major(out)
    return out


def _layer_shapes(geo: Geometry):
    return {"q_proj": (geo.n_q * geo.head_dim, geo.d), "k_proj": (geo.n_kv * geo.head_dim, geo.d),
            "v_proj": (geo.n_kv * geo.head_dim, geo.d), "o_proj": (geo.d, geo.n_q * geo.head_dim),
            "gate_proj": (geo.ffn, geo.d), "up_proj": (geo.ffn, geo.d), "down_proj": (geo.d, geo.ffn)}


def random_init(geo: Geometry, seed: int, device, tp_rank: int = 0, tp_size: int = 1, sigma: float = 0.02,
                embed_sigma: float = 0.02, gain_sigma: float = 0.02, stage_major_copies: bool = False) -> PackedWeights:
    """Seeded random-init weights generated on the device, packed layer by layer (the
    un-packed copy of a layer is dropped before the next one is made)."""
    out = PackedWeights(geo, tp_rank, tp_size)
    sd = {"model.embed_tokens.weight": fill((geo.vocab, geo.d), tensor_seed(seed, GLOBAL_TID["embed_tokens"]),
                                            embed_sigma, 0.0, device),
          "model.norm.weight": fill((geo.d,), tensor_seed(seed, GLOBAL_TID["norm"]), gain_sigma, 1.0, device)}
    if not geo.tied:
        sd["lm_head.weight"] = fill((geo.vocab, geo.d), tensor_seed(seed, GLOBAL_TID["lm_head"]), sigma, 0.0, device)
    pack_globals(geo, sd, tp_rank, tp_size, out)
    shapes = _layer_shapes(geo)
    for l in range(geo.n_layers):
        lsd = {}
        for kind, tid in KIND.items():
            s = tensor_seed(seed, (l + 1) * 16 + tid)
            if kind in shapes:
                sub = "self_attn" if kind in ("q_proj", "k_proj", "v_proj", "o_proj") else "mlp"
                lsd[f"model.layers.{l}.{sub}.{kind}.weight"] = fill(shapes[kind], s, sigma, 0.0, device)
            else:
                lsd[f"model.layers.{l}.{kind}.weight"] = fill((geo.d,), s, gain_sigma, 1.0, device)
        pack_layer(geo, l, lsd, tp_rank, tp_size, out)
    if stage_major_copies:
        add_stage_major(out)
    return out
