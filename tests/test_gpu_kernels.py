"""-m gpu: every kernel through its C-ABI entry point against the CPU oracle math
(torch fp32 on the same bf16 weights).  Integer work (weight stream, token ids, argmax)
is bit-exact; floating point is fp32-accumulated in a different order than the oracle,
tolerance 2e-4 relative to the output scale (far inside north_star's 1e-2)."""
import ctypes as C
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import llama_oracle as lo          # noqa: E402
from oracle import sampler as osamp            # noqa: E402
from oracle import weightgen as wg             # noqa: E402

TOL = 2e-4


def _close(got, ref, tol=TOL):
    scale = float(ref.abs().max()) + 1e-6
    err = float((got - ref).abs().max()) / scale
    assert err <= tol, f"rel err {err:.3e} > {tol}"


def test_device_weight_stream_is_bit_exact():
    from b200llm import weights
    for shape, seed, sigma, mean in [((1000, 333), 12345, 0.02, 0.0), ((4096,), 7, 0.02, 1.0),
                                     ((3, 5, 7), (0xB200 << 20) + 33, 0.08, 0.0)]:
        dev = weights.fill(shape, seed, sigma, mean, "cuda")
        torch.cuda.synchronize()
        ref = wg.gen_bf16(shape, seed, sigma, mean)
        assert torch.equal(dev.cpu().view(torch.int16), ref.view(torch.int16))
    assert weights.tensor_seed(0xB200, 17) == wg.tensor_seed(0xB200, 17)
    assert weights.KIND == wg.KIND and weights.GLOBAL_TID == wg.GLOBAL_TID


SHAPES = [(64, 256), (300, 512), (1000, 1024), (520, 1376), (3072, 3072), (1024, 8192), (512, 11008), (148 * 3 + 5, 2048)]


@pytest.mark.parametrize("N,K", SHAPES)
@pytest.mark.parametrize("bs", [1, 2, 3, 4])
def test_gemv_store_and_residual(N, K, bs):
    from b200llm import _lib
    from gpu_util import gemv
    w = wg.gen_bf16((N, K), N * 131 + K, 0.05)
    x = torch.randn(bs, K, generator=torch.Generator().manual_seed(K + bs))
    ref = x @ w.float().t()
    got = gemv(w.cuda(), x.cuda())
    _close(got.cpu(), ref)
    h0 = torch.randn(bs, N, generator=torch.Generator().manual_seed(1))
    out = h0.clone().cuda()
    gemv(w.cuda(), x.cuda(), epi=_lib.EPI_RESIDUAL, out=out)
    _close(out.cpu(), h0 + ref)


@pytest.mark.parametrize("N,K,bs", [(1024, 256, 1), (16384, 3072, 1), (2752, 4096, 3), (8192, 2048, 4)])
def test_gemv_rmsnorm_swiglu(N, K, bs):
    from b200llm import _lib
    from gpu_util import gemv
    w = wg.gen_bf16((N, K), 99 + N, 0.03)
    gain = wg.gen_bf16((K,), 5, 0.02, 1.0)
    x = 3.0 * torch.randn(bs, K, generator=torch.Generator().manual_seed(2))
    xn = lo._rms_norm(x, gain.float(), 1e-5)
    y = xn @ w.float().t()
    ref = torch.nn.functional.silu(y[:, 0::2]) * y[:, 1::2]
    got = gemv(w.cuda(), x.cuda(), epi=_lib.EPI_SWIGLU, gain=gain.cuda(), eps=1e-5)
    _close(got.cpu(), ref)


@pytest.mark.parametrize("name,bs", [("tiny", 1), ("small", 3), ("llama3.2", 1), ("llama3.2:1b", 4), ("duckdb-nsql", 2)])
def test_gemv_qkv_rope_kv_append(name, bs):
    """Fused RMSNorm -> QKV -> RoPE -> paged-KV append against the oracle's per-layer math."""
    from b200llm import _lib, geometry, weights
    from gpu_util import gemv, pair_perm
    geo = geometry.resolve(name)
    hd, nq, nkv, d = geo.head_dim, geo.n_q, geo.n_kv, geo.d
    g = torch.Generator().manual_seed(11)
    sd = {"model.layers.0.self_attn.q_proj.weight": wg.gen_bf16((nq * hd, d), 1, 0.03),
          "model.layers.0.self_attn.k_proj.weight": wg.gen_bf16((nkv * hd, d), 2, 0.03),
          "model.layers.0.self_attn.v_proj.weight": wg.gen_bf16((nkv * hd, d), 3, 0.03),
          "model.layers.0.self_attn.o_proj.weight": torch.zeros(d, nq * hd, dtype=torch.bfloat16),
          "model.layers.0.mlp.gate_proj.weight": torch.zeros(8, d, dtype=torch.bfloat16),
          "model.layers.0.mlp.up_proj.weight": torch.zeros(8, d, dtype=torch.bfloat16),
          "model.layers.0.mlp.down_proj.weight": torch.zeros(d, 8, dtype=torch.bfloat16),
          "model.layers.0.input_layernorm.weight": wg.gen_bf16((d,), 4, 0.02, 1.0),
          "model.layers.0.post_attention_layernorm.weight": torch.ones(d, dtype=torch.bfloat16)}
    import dataclasses
    pk = weights.PackedWeights(geo, 0, 1)
    weights.pack_layer(dataclasses.replace(geo, ffn=8), 0, sd, 0, 1, pk)
    wqkv = [t for k, l, t in pk.tensors if k == _lib.W_QKV][0].cuda()
    x = torch.randn(bs, d, generator=g)
    pos = torch.tensor([5, 16, 1000, 8191][:bs], dtype=torch.int32)
    lanes = torch.arange(bs, dtype=torch.int32)
    max_pages = 8192 // 16
    n_pages = bs * 2 + 3
    ptab = torch.zeros(bs, max_pages, dtype=torch.int32)
    for b in range(bs):
        ptab[b, int(pos[b]) // 16] = 2 + b                      # the page holding `pos`
    kp = torch.zeros(n_pages, nkv, 16, hd, dtype=torch.bfloat16, device="cuda")
    vp = torch.zeros_like(kp)
    inv = geometry.inv_freq(geo)
    q = gemv(wqkv, x.cuda(), epi=_lib.EPI_QKV_ROPE, gain=sd["model.layers.0.input_layernorm.weight"].cuda(),
             n_q=nq, n_kv=nkv, head_dim=hd, page_tokens=16, max_pages_per_lane=max_pages, inv_freq=inv.cuda(),
             row_lane=lanes.cuda(), row_pos=pos.cuda(), page_table=ptab.cuda(), k_pages=kp, v_pages=vp)
    # oracle
    xn = lo._rms_norm(x, sd["model.layers.0.input_layernorm.weight"].float(), geo.rms_eps)
    perm = pair_perm(hd)
    for b in range(bs):
        fr = float(pos[b]) * inv
        emb = torch.cat((fr, fr))
        cos, sin = emb.cos(), emb.sin()
        qr = (xn[b] @ sd["model.layers.0.self_attn.q_proj.weight"].float().t()).view(nq, hd)
        kr = (xn[b] @ sd["model.layers.0.self_attn.k_proj.weight"].float().t()).view(nkv, hd)
        vr = (xn[b] @ sd["model.layers.0.self_attn.v_proj.weight"].float().t()).view(nkv, hd)
        qr = qr * cos + lo._rotate_half(qr) * sin
        kr = kr * cos + lo._rotate_half(kr) * sin
        _close(q[b].cpu().view(nq, hd), qr[:, perm], 5e-4)       # engine keeps q/k dims pair-interleaved
        page, slot = 2 + b, int(pos[b]) % 16
        _close(kp[page, :, slot].float().cpu(), kr[:, perm].to(torch.bfloat16).float(), 1e-2)   # bf16 storage
        _close(vp[page, :, slot].float().cpu(), vr.to(torch.bfloat16).float(), 1e-2)
        # exactly-rounded in all but borderline elements
        same = (kp[page, :, slot].cpu().view(torch.int16) == kr[:, perm].to(torch.bfloat16).view(torch.int16)).float().mean()
        assert same > 0.97
    # nothing but the appended rows was written: one (page, token slot) per batch row, every kv head of it
    written = (kp != 0).any(-1).cpu()                            # [page, kv head, slot]
    assert int(written.sum()) == bs * nkv
    for b in range(bs):
        assert bool(written[2 + b, :, int(pos[b]) % 16].all())


@pytest.mark.parametrize("V,K,bs", [(1024, 256, 1), (32000, 4096, 2), (128256, 2048, 1), (128256, 3072, 4)])
def test_gemv_logits_argmax_partials(V, K, bs):
    from b200llm import _lib
    from gpu_util import gemv, gemv_grid, sample
    w = wg.gen_bf16((V, K), 1234 + V, 0.02)
    gain = wg.gen_bf16((K,), 6, 0.02, 1.0)
    x = torch.randn(bs, K, generator=torch.Generator().manual_seed(4))
    ref = lo._rms_norm(x, gain.float(), 1e-5) @ w.float().t()
    grid = gemv_grid(V, K, bs, _lib.EPI_LOGITS)
    pm = torch.full((bs, grid), float("nan"), device="cuda")
    pi = torch.full((bs, grid), -1, dtype=torch.int32, device="cuda")
    logits = torch.empty(bs, V, device="cuda")
    gemv(w.cuda(), x.cuda(), epi=_lib.EPI_LOGITS, gain=gain.cuda(), out=logits, part_max=pm, part_idx=pi)
    _close(logits.cpu(), ref)
    # partial argmax is bit-consistent with the logits the kernel wrote (lowest id on ties)
    lg = logits.cpu()
    best = pm.cpu().max(1)
    for b in range(bs):
        assert float(best.values[b]) == float(lg[b].max())
        cands = pi.cpu()[b][pm.cpu()[b] == best.values[b]]
        assert int(cands.min()) == int(torch.argmax(lg[b]))
    greedy = [_lib.Sampling(0.0, 0, 1.0, 1.0, 0)] * bs
    toks = sample(logits, greedy, pm, pi)
    assert toks.tolist() == torch.argmax(lg, 1).tolist()
    # ties resolve to the lowest id
    lg2 = torch.zeros(1, 5000, device="cuda")
    lg2[0, [4000, 77, 3000]] = 1.0
    assert sample(lg2, [greedy[0]]).tolist() == [77]


def _paged_kv(bs, ctx_lens, nkv, hd, gen, extra_pages=3):
    max_pages = (max(ctx_lens) + 15) // 16 + 1
    total = sum((c + 15) // 16 for c in ctx_lens) + extra_pages
    order = torch.randperm(total, generator=gen)
    ptab = torch.zeros(bs, max_pages, dtype=torch.int32)
    kp = torch.randn(total, nkv, 16, hd, generator=gen).to(torch.bfloat16)
    vp = torch.randn(total, nkv, 16, hd, generator=gen).to(torch.bfloat16)
    nxt = 0
    for b, c in enumerate(ctx_lens):
        for p in range((c + 15) // 16):
            ptab[b, p] = order[nxt]
            nxt += 1
    return ptab, kp, vp


@pytest.mark.parametrize("nq,nkv,hd", [(4, 2, 64), (8, 2, 128), (24, 8, 128), (32, 8, 64), (32, 32, 128), (4, 4, 128)])
@pytest.mark.parametrize("ctx,splits", [([1], 1), ([16, 17], 2), ([37, 300, 5, 1024], 4), ([2072], 16), ([513, 64, 31], 32)])
def test_attn_decode_matches_oracle(nq, nkv, hd, ctx, splits):
    from gpu_util import attn_decode
    bs = len(ctx)
    gen = torch.Generator().manual_seed(nq * 1000 + hd + sum(ctx))
    ptab, kp, vp = _paged_kv(bs, ctx, nkv, hd, gen)
    q = torch.randn(bs, nq * hd, generator=gen)
    pos = torch.tensor([c - 1 for c in ctx], dtype=torch.int32)
    lanes = torch.arange(bs, dtype=torch.int32)
    out = attn_decode(q.cuda(), kp.cuda(), vp.cuda(), lanes.cuda(), pos.cuda(), ptab.cuda(), nq, nkv, hd, splits)
    rep = nq // nkv
    for b in range(bs):
        T = ctx[b]
        K = torch.cat([kp[ptab[b, p]] for p in range((T + 15) // 16)], 1)[:, :T].float()      # [nkv, T, hd]
        V = torch.cat([vp[ptab[b, p]] for p in range((T + 15) // 16)], 1)[:, :T].float()
        qh = q[b].view(nq, 1, hd)
        att = (qh @ K.repeat_interleave(rep, 0).transpose(1, 2)) / math.sqrt(hd)
        ref = (torch.softmax(att, -1) @ V.repeat_interleave(rep, 0)).reshape(nq * hd)
        _close(out[b].cpu(), ref, 1e-4)


def test_attn_scratch_is_reusable_and_checked():
    """Tickets return to zero: the same scratch serves launch after launch; a short scratch is refused."""
    from b200llm import _lib
    from gpu_util import attn_decode, attn_scratch
    gen = torch.Generator().manual_seed(5)
    nq, nkv, hd, ctx = 8, 2, 128, [200, 90]
    ptab, kp, vp = _paged_kv(2, ctx, nkv, hd, gen)
    pos = torch.tensor([c - 1 for c in ctx], dtype=torch.int32).cuda()
    lanes = torch.arange(2, dtype=torch.int32).cuda()
    sc = attn_scratch(2, nq, nkv, hd, 8)
    outs = []
    for i in range(3):
        q = torch.randn(2, nq * hd, generator=torch.Generator().manual_seed(i)).cuda()
        o1 = attn_decode(q, kp.cuda(), vp.cuda(), lanes, pos, ptab.cuda(), nq, nkv, hd, 8, scratch=sc)
        o2 = attn_decode(q, kp.cuda(), vp.cuda(), lanes, pos, ptab.cuda(), nq, nkv, hd, 1)
        _close(o1.cpu(), o2.cpu(), 1e-5)
        outs.append(o1)
    assert int(sc[:256 * 1024].view(torch.int32).abs().sum()) == 0
    with pytest.raises(_lib.B200LLMError) as ei:
        attn_decode(q, kp.cuda(), vp.cuda(), lanes, pos, ptab.cuda(), nq, nkv, hd, 8, scratch=sc[:256 * 1024 + 512])
    assert ei.value.code == _lib.ENOMEM


def test_sampler_matches_oracle_chain():
    from b200llm import _lib
    from gpu_util import sample
    rng = np.random.default_rng(3)
    V = 32000
    agree = total = 0
    for trial in range(24):
        lg = (rng.standard_normal(V) * 2.0).astype(np.float32)
        hist = np.full(64, -1, dtype=np.int32)
        hist[:10] = rng.integers(0, V, 10)
        hist[10] = hist[2]
        temp, k, p, pen = [(0.8, 40, 0.9, 1.1), (1.0, 5, 1.0, 1.0), (0.5, 0, 0.95, 1.3), (1.5, 100, 0.5, 1.0)][trial % 4]
        sp = _lib.Sampling(temp, k, p, pen, 1000 + trial)
        got = sample(torch.from_numpy(lg)[None].cuda(), [sp], history=torch.from_numpy(hist)[None].cuda(), step=trial)
        ref, ids, probs = osamp.sample(lg, temp, k, p, pen, 1000 + trial, trial, history=hist)
        assert int(got[0]) in set(int(i) for i in ids)          # always inside the oracle's kept set
        agree += int(got[0]) == ref
        total += 1
    assert agree >= total - 1                                   # expf ulps may move one boundary draw
    # greedy + penalty goes through the full-vocab argmax
    lg = np.zeros(V, np.float32); lg[123] = 5.0; lg[77] = 4.9
    hist = np.full(64, -1, np.int32); hist[0] = 123
    got = sample(torch.from_numpy(lg)[None].cuda(), [_lib.Sampling(0.0, 0, 1.0, 2.0, 0)],
                 history=torch.from_numpy(hist)[None].cuda())
    assert got.tolist() == [77]


@pytest.mark.parametrize("N,K,T,acc", [(128, 64, 1, 0), (256, 256, 5, 1), (3072, 3072, 64, 1), (1024, 8192, 256, 1),
                                       (148 * 128, 512, 33, 0), (5120, 3072, 17, 1), (512, 11008, 100, 1)])
def test_gemm_tc_matches_fp32_matmul(N, K, T, acc):
    """tcgen05 GEMM (bf16 x bf16 -> fp32) against torch fp32 on the same bf16 values; the k-split
    partials meet with fp32 atomics, so the comparison is to 2e-4 of the output scale."""
    import ctypes as C
    from b200llm import _lib
    w = wg.gen_bf16((N, K), 31 * N + K, 0.05)
    x = (torch.randn(T, K, generator=torch.Generator().manual_seed(T)) * 0.5).to(torch.bfloat16)
    base = torch.randn(T, N, generator=torch.Generator().manual_seed(1)) if acc else torch.zeros(T, N)
    ref = base + x.float() @ w.float().t()
    out = (base.clone() if acc else torch.full((T, N), float("nan"))).cuda()
    wd, xd = w.cuda(), x.cuda()
    _lib.check(_lib.lib.b200llm_op_gemm_tc(wd.data_ptr(), xd.data_ptr(), out.data_ptr(), N, T, N, K, acc,
                                           torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    _close(out.cpu(), ref)


@pytest.mark.parametrize("N,K", [(128, 64), (512, 256), (3072, 3072), (5120, 3072), (3072, 8192), (16384, 3072),
                                 (4096, 11008), (128256, 2048)])
@pytest.mark.parametrize("k_split", [0, 1])
def test_gemv_tile_major_tcgen05_phase(N, K, k_split):
    """One phase of the tcgen05 persistent decode kernel: tile-major weights, activation split into
    bf16 hi + lo operand rows, fp32 accumulation in TMEM.  k_split = 1 is the stream-K mode (partial
    sums meet with fp32 atomics), 0 the whole-tile mode the lm_head uses."""
    from b200llm import _lib, weights
    if k_split == 0 and K // 64 * ((N // 128 + 147) // 148) > 64 and K // 64 > 64:
        pytest.skip("share of one CTA exceeds the operand buffer in whole-tile mode")
    w = wg.gen_bf16((N, K), N * 17 + K, 0.05)
    x = torch.randn(K, generator=torch.Generator().manual_seed(K + N))
    ref = w.float() @ x
    wt = weights.tile_major(w.cuda())
    xd = x.cuda()
    out = torch.full((N,), 7.0, device="cuda")
    torch.cuda.synchronize()
    _lib.check(_lib.lib.b200llm_op_gemv_tm(wt.data_ptr(), xd.data_ptr(), out.data_ptr(), N, K, k_split,
                                           torch.cuda.current_stream().cuda_stream))
    _close(out.cpu(), ref, 2e-5)


@pytest.mark.parametrize("nq,nkv,hd,pos0,rows,kvph", [(4, 2, 128, 0, 128, 0), (8, 2, 64, 0, 200, 0), (6, 2, 128, 300, 77, 0),
                                                        (4, 4, 128, 1000, 256, 0), (4, 2, 128, 37, 150, 7)])
def test_attn_prefill_matches_fp32_attention(nq, nkv, hd, pos0, rows, kvph):
    """Causal prefill attention on tcgen05 (csrc/attn_prefill.cu) against plain fp32 softmax(q k^T / sqrt(hd)) v in torch on
    the same bf16 K/V: chunks that start at position 0 and in the middle of a sequence, ragged row counts, GQA groups 1..3,
    both head sizes, scattered pages, and a page stride larger than n_kv (the shared-pool layout)."""
    import ctypes as C
    from b200llm import _lib
    gen = torch.Generator().manual_seed(1000 + rows + pos0)
    n_tok = pos0 + rows
    n_lane_pages = (n_tok + 15) // 16
    n_pages = n_lane_pages + 5
    stride = kvph or nkv
    order = torch.randperm(n_pages, generator=gen)[:n_lane_pages]
    max_pages = n_lane_pages + 3
    ptab = torch.zeros(max_pages, dtype=torch.int32)
    ptab[:n_lane_pages] = order.to(torch.int32)
    kp = torch.zeros(n_pages, stride, 16, hd, dtype=torch.bfloat16)
    vp = torch.zeros_like(kp)
    K = torch.randn(n_tok, nkv, hd, generator=gen).to(torch.bfloat16)
    V = torch.randn(n_tok, nkv, hd, generator=gen).to(torch.bfloat16)
    for t in range(n_tok):
        kp[order[t // 16], :nkv, t % 16] = K[t]
        vp[order[t // 16], :nkv, t % 16] = V[t]
    q = torch.randn(rows, nq * hd, generator=gen)
    qd, kpd, vpd, ptd = q.cuda(), kp.cuda(), vp.cuda(), ptab.cuda()
    out = torch.full_like(qd, float("nan"))
    a = _lib.AttnPrefillArgs(qd.data_ptr(), out.data_ptr(), kpd.data_ptr(), vpd.data_ptr(), ptd.data_ptr(), max_pages, rows, pos0,
                             nq, nkv, hd, n_pages, kvph)
    _lib.check(_lib.lib.b200llm_op_attn_prefill(C.byref(a), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    # reference
    rep = nq // nkv
    Kf, Vf = K.float(), V.float()
    ref = torch.empty(rows, nq, hd)
    for h in range(nq):
        s = (q.view(rows, nq, hd)[:, h] @ Kf[:, h // rep].t()) / hd ** 0.5                     # [rows, n_tok]
        mask = torch.arange(n_tok)[None, :] > (pos0 + torch.arange(rows))[:, None]
        ref[:, h] = torch.softmax(s.masked_fill(mask, float("-inf")), dim=-1) @ Vf[:, h // rep]
    got = out.cpu().view(rows, nq, hd)
    assert torch.isfinite(got).all()
    err = float((got - ref).abs().max() / ref.abs().max())
    assert err <= 2e-4, f"rel err {err:.3e}"
