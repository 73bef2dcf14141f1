"""Helpers for the -m gpu tests: call the single-kernel C-ABI entry points on torch tensors."""
import ctypes as C

import numpy as np
import torch

from b200llm import _lib


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def gemv(w, x, epi=_lib.EPI_STORE, gain=None, eps=1e-5, out=None, **kw):
    """w bf16 [N,K], x fp32 [bs,K]; returns out (fp32)."""
    N, K = w.shape
    bs = x.shape[0]
    if out is None:
        ncols = N // 2 if epi == _lib.EPI_SWIGLU else (kw["n_q"] * kw["head_dim"] if epi == _lib.EPI_QKV_ROPE else N)
        out = torch.zeros(bs, ncols, dtype=torch.float32, device=w.device)
    a = _lib.GemvArgs()
    a.w, a.x, a.x_stride = w.data_ptr(), x.data_ptr(), x.stride(0)
    a.norm_gain = gain.data_ptr() if gain is not None else None
    a.rms_eps, a.out, a.out_stride = eps, out.data_ptr(), out.stride(0)
    a.N, a.K, a.bs, a.epi = N, K, bs, epi
    keep = []
    for f in ("n_q", "n_kv", "head_dim", "page_tokens", "max_pages_per_lane", "logits_idx_offset", "grid_hint"):
        if f in kw:
            setattr(a, f, kw[f])
    for f in ("inv_freq", "row_lane", "row_pos", "page_table", "k_pages", "v_pages", "part_max", "part_idx"):
        if f in kw and kw[f] is not None:
            keep.append(kw[f])
            setattr(a, f, kw[f].data_ptr())
    _lib.check(_lib.lib.b200llm_op_gemv(C.byref(a), stream_ptr()))
    torch.cuda.synchronize()
    return out


def gemv_grid(N, K, bs, epi):
    g = C.c_int32()
    _lib.check(_lib.lib.b200llm_op_gemv_grid(N, K, bs, epi, C.byref(g)))
    return g.value


def attn_scratch(bs, n_q, n_kv, hd, n_splits, device="cuda"):
    n = C.c_size_t()
    _lib.check(_lib.lib.b200llm_op_attn_scratch_bytes(bs, n_q, n_kv, hd, n_splits, C.byref(n)))
    return torch.zeros(max(n.value, 256), dtype=torch.uint8, device=device)


def attn_decode(q, k_pages, v_pages, row_lane, row_pos, page_table, n_q, n_kv, hd, n_splits, scratch=None,
                entry="b200llm_op_attn_decode"):
    bs = q.shape[0]
    out = torch.empty_like(q)
    if scratch is None:
        scratch = attn_scratch(bs, n_q, n_kv, hd, n_splits, q.device)
    a = _lib.AttnArgs()
    a.q, a.out, a.k_pages, a.v_pages = q.data_ptr(), out.data_ptr(), k_pages.data_ptr(), v_pages.data_ptr()
    a.row_lane, a.row_pos, a.page_table = row_lane.data_ptr(), row_pos.data_ptr(), page_table.data_ptr()
    a.bs, a.n_q, a.n_kv, a.head_dim, a.page_tokens = bs, n_q, n_kv, hd, 16
    a.max_pages_per_lane, a.n_splits = page_table.shape[1], n_splits
    a.scratch, a.scratch_bytes = scratch.data_ptr(), scratch.numel()
    _lib.check(getattr(_lib.lib, entry)(C.byref(a), stream_ptr()))
    torch.cuda.synchronize()
    return out


def sample(logits, params, part_max=None, part_idx=None, history=None, step=0):
    bs, V = logits.shape
    out = torch.empty(bs, dtype=torch.int32, device=logits.device)
    arr = (_lib.Sampling * bs)(*[p for p in params])
    dparams = torch.from_numpy(np.frombuffer(bytes(arr), dtype=np.uint8).copy()).to(logits.device)
    a = _lib.SampleArgs()
    a.logits, a.bs, a.vocab = logits.data_ptr(), bs, V
    if part_max is not None:
        a.part_max, a.part_idx, a.n_part = part_max.data_ptr(), part_idx.data_ptr(), part_max.shape[1]
    a.params = dparams.data_ptr()
    a.history = history.data_ptr() if history is not None else None
    a.step, a.out_token = step, out.data_ptr()
    _lib.check(_lib.lib.b200llm_op_sample(C.byref(a), stream_ptr()))
    torch.cuda.synchronize()
    return out.cpu().numpy()


def pair_perm(hd):
    half = hd // 2
    idx = torch.empty(hd, dtype=torch.long)
    idx[0::2] = torch.arange(half)
    idx[1::2] = torch.arange(half) + half
    return idx
