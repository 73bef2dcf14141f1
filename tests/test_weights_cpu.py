"""Host-side weight packing on CPU tensors: tile-major copies and the lean (tile-major only) set."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]

from b200llm import _lib, geometry, weights          # noqa: E402
from oracle import llama_oracle as lo                # noqa: E402


def _packed(name):
    geo = geometry.resolve(name)
    sd = lo.make_weights(lo.GEOMETRIES[name], 5)
    return geo, sd, weights.pack(geo, sd)


def test_tile_major_is_a_permutation_of_the_matrix():
    """tile_major(): [N, K] -> [N/128, K/64, 128, 64] with 16-byte chunk c of row r stored at chunk c ^ (r & 7): undoing
    the swizzle and the tiling gives the matrix back (the layout tcgen05.mma and the 1-D bulk copies rely on)."""
    w = (torch.arange(256 * 128, dtype=torch.float32).reshape(256, 128) % 251).to(torch.bfloat16)
    t = weights.tile_major(w)
    assert tuple(t.shape) == (2, 2, 128, 8, 8) and t.logical_shape == (256, 128)
    idx = torch.arange(8)[None, :] ^ (torch.arange(128)[:, None] & 7)
    un = t[:, :, torch.arange(128)[:, None], idx]                       # the XOR is its own inverse
    back = un.permute(0, 2, 1, 3, 4).reshape(256, 128)
    assert torch.equal(back, w)
    padded = weights.tile_major(w[:200])                                # rows up to the next tile are zero
    assert padded.shape[0] == 2 and padded.logical_shape == (256, 128)
    assert float(padded[1, :, 72:].float().abs().max()) == 0.0


@pytest.mark.parametrize("name", ["tiny", "small"])
def test_lean_set_keeps_one_copy_of_every_matrix(name):
    geo, sd, pw = _packed(name)
    with pytest.raises(ValueError):
        weights.drop_row_major(pw)                                      # no tile-major copies yet
    weights.add_tile_major(pw)
    full = sum(t.numel() * t.element_size() for _, _, t in pw.tensors)
    weights.drop_row_major(pw)
    kinds = [k for k, _, _ in pw.tensors]
    for k in (_lib.W_QKV, _lib.W_O, _lib.W_GATE_UP, _lib.W_DOWN):
        assert k not in kinds
    for k in (_lib.W_QKV_TM, _lib.W_O_TM, _lib.W_GATE_UP_TM, _lib.W_DOWN_TM, _lib.W_LM_HEAD_TM, _lib.W_EMBED, _lib.W_FINAL_NORM,
              _lib.W_ATTN_NORM, _lib.W_MLP_NORM, _lib.W_INV_FREQ):
        assert k in kinds
    # a tied lm_head IS the embedding table (needed for the token lookup): it stays; an untied one goes
    assert (_lib.W_LM_HEAD in kinds) == bool(lo.GEOMETRIES[name].tied)
    lean = sum(t.numel() * t.element_size() for _, _, t in pw.tensors)
    assert lean < 0.75 * full
    descs = pw.descs()
    assert len(descs) == len(pw.tensors)
