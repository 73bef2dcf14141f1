"""-m gpu, collected last on purpose: the no-allocation attention entry point.

b200llm_op_attn_decode_static is for host programs that cannot give the library a scratch
buffer because they cannot allocate one (all of HBM already belongs to their own pool, or they
call from a thread that must not take the allocator lock / the device-wide synchronisation of
cudaMalloc).  It keeps the split-T partials and tickets in a 4 MiB buffer inside the library.
This test calls it through the C ABI twice in a row and compares each result with the regular
entry point (caller-provided scratch) on the same inputs, to the tolerance the attention tests
use; it also checks that an oversized request is refused with a clear message.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(seed, nq, nkv, hd, ctx):
    gen = torch.Generator().manual_seed(seed)
    bs = len(ctx)
    max_pages = (max(ctx) + 15) // 16 + 1
    total = sum((c + 15) // 16 for c in ctx) + 2
    order = torch.randperm(total, generator=gen)
    ptab = torch.zeros(bs, max_pages, dtype=torch.int32)
    nxt = 0
    for b, c in enumerate(ctx):
        for p in range((c + 15) // 16):
            ptab[b, p] = order[nxt]
            nxt += 1
    kp = torch.randn(total, nkv, 16, hd, generator=gen).to(torch.bfloat16).cuda()
    vp = torch.randn(total, nkv, 16, hd, generator=gen).to(torch.bfloat16).cuda()
    q = torch.randn(bs, nq * hd, generator=gen).cuda()
    pos = torch.tensor([c - 1 for c in ctx], dtype=torch.int32).cuda()
    lanes = torch.arange(bs, dtype=torch.int32).cuda()
    return q, kp, vp, lanes, pos, ptab.cuda()


def _torch_reference(q, kp, vp, pos, ptab, nq, nkv, hd):
    """Plain fp32 softmax(q k^T / sqrt(hd)) v over each row's paged context (GQA: q head h reads kv head h // group)."""
    out = torch.empty_like(q)
    rep = nq // nkv
    for b in range(q.shape[0]):
        n = int(pos[b]) + 1
        pages = ptab[b, :(n + 15) // 16].long()
        K = kp[pages].float().permute(1, 0, 2, 3).reshape(nkv, -1, hd)[:, :n]      # [kv head, token, dim]
        V = vp[pages].float().permute(1, 0, 2, 3).reshape(nkv, -1, hd)[:, :n]
        qb = q[b].view(nq, hd)
        for h in range(nq):
            s_ = (K[h // rep] @ qb[h]) / (hd ** 0.5)
            out[b, h * hd:(h + 1) * hd] = torch.softmax(s_, dim=0) @ V[h // rep]
    return out


def test_static_scratch_attention_matches_regular_entry_point_twice():
    from gpu_util import attn_decode
    nq, nkv, hd, splits = 24, 8, 128, 8
    for call, ctx in enumerate(([300, 77, 1024], [513, 16, 45])):       # two calls in a row, different inputs
        q, kp, vp, lanes, pos, ptab = _case(100 + call, nq, nkv, hd, ctx)
        want = attn_decode(q, kp, vp, lanes, pos, ptab, nq, nkv, hd, splits)
        got = attn_decode(q, kp, vp, lanes, pos, ptab, nq, nkv, hd, splits, entry="b200llm_op_attn_decode_static")
        scale = float(want.abs().max())
        err = float((got - want).abs().max()) / scale
        assert err <= 1e-4, f"call {call}: rel err {err:.3e}"           # tolerance of test_attn_decode_matches_oracle
        ref = _torch_reference(q, kp, vp, pos, ptab, nq, nkv, hd)        # and directly against fp32 attention in torch
        err_ref = float((got - ref).abs().max()) / float(ref.abs().max())
        assert err_ref <= 1e-4, f"call {call}: rel err vs torch fp32 {err_ref:.3e}"
    # same inputs again: the shared buffer kept the previous call's partials, the result must not change
    again = attn_decode(q, kp, vp, lanes, pos, ptab, nq, nkv, hd, splits, entry="b200llm_op_attn_decode_static")
    assert torch.equal(again, got)


def test_static_scratch_attention_refuses_oversized_requests():
    from b200llm import _lib
    from gpu_util import attn_decode
    nq, nkv, hd = 32, 8, 128
    q, kp, vp, lanes, pos, ptab = _case(7, nq, nkv, hd, [40] * 16)
    with pytest.raises(_lib.B200LLMError) as ei:                         # 16 x 32 x 64 x 130 x 4 B = 17 MB > 4 MiB
        attn_decode(q, kp, vp, lanes, pos, ptab, nq, nkv, hd, 64, entry="b200llm_op_attn_decode_static")
    assert ei.value.code == _lib.ENOMEM and "built-in buffer" in ei.value.message
    # a launch that fits still works afterwards
    small = attn_decode(q[:2], kp, vp, lanes[:2], pos[:2], ptab[:2], nq, nkv, hd, 4, entry="b200llm_op_attn_decode_static")
    ref = attn_decode(q[:2], kp, vp, lanes[:2], pos[:2], ptab[:2], nq, nkv, hd, 4)
    assert float((small - ref).abs().max()) <= 1e-4 * float(ref.abs().max())
