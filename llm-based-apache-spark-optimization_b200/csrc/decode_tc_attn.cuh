// decode_tc_attn.cuh -- attention phase of the tcgen05 persistent decode kernel (decode_tc.cu).
//
// Same register-streamed scheme as mega_attn.cuh (all compute warps read K/V rows of the lane's
// pages straight into registers, 16 lanes per token, split-T only for long contexts), with one
// difference: the qkv phase before it leaves RAW q|k|v sums in p.qkv (its k-split partials meet
// in global memory, so no single CTA ever holds a finished row).  RoPE is therefore applied here:
// q on load, and the current token's k/v by warp 0 of the item that owns the last page, which
// rounds them to bf16, starts its running softmax with them and appends them to the KV page for
// the steps to come.  The page loop masks the current position out of the cache reads.
// Work items = (lane, kv head, split), lane after lane, each lane with its own position, page table and split count
// (tcd_att_item); item i runs on CTA i % G.  With more than one split the items
// leave (max, sum, unnormalised output) partials in part_ml / part_o and the next phase's
// prologue merges them while it builds its operand (tcd_prologue, TCD_X_ATT).
#pragma once

__device__ __forceinline__ void prefetch_l2(const void *ptr) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr));
}

// Work item `item` of an attention phase -> (lane, kv head, split).  Lane b at position pos_b has n_kv * ns_b items
// (ns_b: its split factor), lane after lane; false past the last item.
struct TcdAttItem {
    int b, hk, split, ns, pps, n_pages, pos;
};
template <int NB>
__device__ __forceinline__ bool tcd_att_item(const TcdParams &p, const TcdSmem &s, const TcdCtxT<NB> &c, int item, TcdAttItem &it) {
    const int nb = NB == 1 ? 1 : p.nb;
#pragma unroll
    for (int b = 0; b < NB; ++b) {
        if (NB > 1 && b >= nb) break;
        const int pos = tcd_pos<NB>(c, s, b);
        const int n_pages = pos / kTPT + 1;
        const int ns = min(p.n_splits, (2 * n_pages + kTW - 1) / kTW);  // one half-page (one L2 round trip) per warp while splits last
        if (item < p.n_kv * ns) {
            it.b = b, it.hk = item / ns, it.split = item % ns, it.ns = ns, it.pps = (n_pages + ns - 1) / ns, it.n_pages = n_pages, it.pos = pos;
            return true;
        }
        item -= p.n_kv * ns;
    }
    return false;
}

// Issued at the start of the qkv phase: pull the K/V pages this CTA's attention item(s) will
// read into L2 while the qkv weights stream, so the attention phase sees L2 instead of HBM
// latency (the pages were last touched a whole step ago and have been evicted since).
template <int HD, int NB>
__device__ __forceinline__ void tcd_attn_prefetch(const TcdParams &p, const TcdJob &qkv_job, const TcdSmem &s,
                                                   const TcdCtxT<NB> &c, int G, int bid, int tid) {
    constexpr int LINES = kTPT * HD * 2 / 128;  // 128-byte lines per K (or V) page tile
    const uint8_t *kbase = reinterpret_cast<const uint8_t *>(qkv_job.k_pages);
    const uint8_t *vbase = reinterpret_cast<const uint8_t *>(qkv_job.v_pages);
    TcdAttItem it;
    for (int item = bid; tcd_att_item<NB>(p, s, c, item, it); item += G) {
        const int pg_begin = it.split * it.pps, pg_end = min(it.n_pages, pg_begin + it.pps);
        for (int i = tid; i < (pg_end - pg_begin) * LINES; i += kTC) {
            const int pg = pg_begin + i / LINES, line = i % LINES;
            const int page = tcd_page<NB>(p, s, it.b, pg);
            const size_t off = ((size_t)page * p.kv_page_heads + it.hk) * (size_t)(kTPT * HD * 2) + (size_t)line * 128;
            prefetch_l2(kbase + off);
            prefetch_l2(vbase + off);
        }
    }
}

// Data-flow mode (TcdParams::dataflow bit 2): instead of a grid barrier after the qkv phase, a CTA that has an attention
// item waits until the (sum, count) pairs of its kv head's rows -- the GQ q heads, k and v -- carry the full count of
// their tile's contributors.  One row per thread and round; the polling load IS the load of q, k and v: the sums go
// to shared memory (s.qst), so the phase starts without another L2 round trip.  (One item per CTA: n_kv * splits <= G.)
template <int HD, int GQ, int NB>
__device__ __forceinline__ void tcd_attn_wait_qkv(const TcdParams &p, const TcdJob &qkv_job, const TcdSmem &s, const TcdCtxT<NB> &c, int G,
                                                  int bid, int tid) {
    const int qrows = p.n_q * HD, krows = p.n_kv * HD;
    TcdAttItem it;
    for (int item = bid; tcd_att_item<NB>(p, s, c, item, it); item += G) {
        const int hk = it.hk;
        const float2 *qkv2 = p.qkv2 + (size_t)it.b * (qrows + 2 * krows);
        for (int i = tid; i < (GQ + 2) * HD; i += kTC) {
            const int row = i < GQ * HD ? hk * GQ * HD + i : (i < (GQ + 1) * HD ? qrows + hk * HD + (i - GQ * HD) : qrows + krows + hk * HD + (i - (GQ + 1) * HD));
            const float want = tcd_tile_contributors(qkv_job.N, qkv_job.K, row / 128, G);
            float sum, cnt;
            for (;;) {
                asm volatile("ld.relaxed.gpu.global.v2.f32 {%0, %1}, [%2];" : "=f"(sum), "=f"(cnt) : "l"(qkv2 + row) : "memory");
                if (cnt == want) break;
                __nanosleep(32);
            }
            s.qst[i] = sum;
        }
    }
    bar_sync(1, kTC);
}
// four consecutive q | k | v sums of the item's head (index i0 into [GQ q heads | k | v] x HD, `row` the same position in the
// qkv buffer): staged in shared memory by tcd_attn_wait_qkv in data-flow mode, else read from the plain buffer
__device__ __forceinline__ float4 tcd_ld_sums4(const TcdParams &p, const TcdSmem &s, int i0, size_t row) {
    if (p.dataflow & 4) return *reinterpret_cast<const float4 *>(s.qst + i0);
    return __ldcg(reinterpret_cast<const float4 *>(p.qkv + row));
}

template <int HD, int GQ, int NB>
__device__ void tcd_attn_phase(const TcdParams &p, const TcdJob &qkv_job, const TcdSmem &s, TcdCtxT<NB> &c,
                                int G, int bid, int tid) {
    constexpr int EL = HD / 16;  // dims per lane: 16 lanes cover one K/V row
    const int warp = tid >> 5, lane = tid & 31;
    const int tg = lane >> 4, p16 = lane & 15;
    const uint16_t *kbase = reinterpret_cast<const uint16_t *>(qkv_job.k_pages);
    const uint16_t *vbase = reinterpret_cast<const uint16_t *>(qkv_job.v_pages);
    float *mrg = s.mrg;  // [kTW][GQ][HD + 2]
    const int qrows = p.n_q * HD, krows = p.n_kv * HD;

    TcdAttItem it;
    for (int item = bid; tcd_att_item<NB>(p, s, c, item, it); item += G) {
        const int hk = it.hk, split = it.split, ns = it.ns, pos = it.pos;
        const int pg_begin = split * it.pps, pg_end = min(it.n_pages, pg_begin + it.pps);
        const int own_split = (it.n_pages - 1) / it.pps;  // the item of each kv head that takes the current token
        const size_t lane_row0 = (size_t)it.b * (qrows + 2 * krows);   // lane b's q | k | v sums
        const float *rope = s.rope + (NB == 1 ? 0 : it.b * 128);
        float q[GQ][EL], m[GQ], l[GQ], o[GQ][EL];
#pragma unroll
        for (int g = 0; g < GQ; ++g) {
            const size_t qrow = lane_row0 + (size_t)(hk * GQ + g) * HD + p16 * EL;
#pragma unroll
            for (int e = 0; e < EL; e += 4) {
                const float4 v = tcd_ld_sums4(p, s, g * HD + p16 * EL + e, qrow + e);
                const float *rp = rope + p16 * EL + e;  // (cos, sin) of the pairs (e, e+1), (e+2, e+3)
                q[g][e] = (v.x * rp[0] - v.y * rp[1]) * p.scale_log2, q[g][e + 1] = (v.y * rp[0] + v.x * rp[1]) * p.scale_log2;
                q[g][e + 2] = (v.z * rp[2] - v.w * rp[3]) * p.scale_log2, q[g][e + 3] = (v.w * rp[2] + v.z * rp[3]) * p.scale_log2;
            }
#pragma unroll
            for (int e = 0; e < EL; ++e) o[g][e] = 0.f;
            m[g] = -INFINITY, l[g] = 0.f;
        }
        if (c.ti >= 0 && q[0][0] != 1.2345e30f) tcd_trace(p, c.ti, bid, tid, TT_ATT_Q);  // (debug) q arrived
        if (split == own_split && warp == 0) {
            // the current token: k (RoPE) and v of this kv head from the raw sums, rounded to bf16 as
            // the cache holds them; both half-warps compute it, lanes 0..15 adopt and append it
            const size_t krow = lane_row0 + (size_t)qrows + (size_t)hk * HD + p16 * EL;
            float kf[EL], vf[EL];
#pragma unroll
            for (int e = 0; e < EL; e += 4) {
                const float4 kv = tcd_ld_sums4(p, s, GQ * HD + p16 * EL + e, krow + e);
                const float4 vv = tcd_ld_sums4(p, s, (GQ + 1) * HD + p16 * EL + e, krow + krows + e);
                const float *rp = rope + p16 * EL + e;
                kf[e] = tcd_bf16_round(kv.x * rp[0] - kv.y * rp[1]), kf[e + 1] = tcd_bf16_round(kv.y * rp[0] + kv.x * rp[1]);
                kf[e + 2] = tcd_bf16_round(kv.z * rp[2] - kv.w * rp[3]), kf[e + 3] = tcd_bf16_round(kv.w * rp[2] + kv.z * rp[3]);
                vf[e] = tcd_bf16_round(vv.x), vf[e + 1] = tcd_bf16_round(vv.y), vf[e + 2] = tcd_bf16_round(vv.z), vf[e + 3] = tcd_bf16_round(vv.w);
            }
#pragma unroll
            for (int g = 0; g < GQ; ++g) {
                float sv = 0.f;
#pragma unroll
                for (int e = 0; e < EL; ++e) sv = fmaf(q[g][e], kf[e], sv);
                sv += __shfl_xor_sync(0xffffffffu, sv, 1);
                sv += __shfl_xor_sync(0xffffffffu, sv, 2);
                sv += __shfl_xor_sync(0xffffffffu, sv, 4);
                sv += __shfl_xor_sync(0xffffffffu, sv, 8);
                if (tg == 0) {
                    m[g] = sv, l[g] = 1.f;
#pragma unroll
                    for (int e = 0; e < EL; ++e) o[g][e] = vf[e];
                }
            }
            if (tg == 0) {
                const int pgi = pos / kTPT;
                const int page = tcd_page<NB>(p, s, it.b, pgi);
                const size_t at = (((size_t)page * p.kv_page_heads + hk) * kTPT + (pos % kTPT)) * HD + p16 * EL;
                uint32_t kw[EL / 2], vw[EL / 2];
#pragma unroll
                for (int e = 0; e < EL / 2; ++e) kw[e] = tcd_pack_bf16(kf[2 * e], kf[2 * e + 1]), vw[e] = tcd_pack_bf16(vf[2 * e], vf[2 * e + 1]);
                if constexpr (EL == 8) {
                    *reinterpret_cast<uint4 *>(const_cast<uint16_t *>(kbase) + at) = make_uint4(kw[0], kw[1], kw[2], kw[3]);
                    *reinterpret_cast<uint4 *>(const_cast<uint16_t *>(vbase) + at) = make_uint4(vw[0], vw[1], vw[2], vw[3]);
                } else {
                    *reinterpret_cast<uint2 *>(const_cast<uint16_t *>(kbase) + at) = make_uint2(kw[0], kw[1]);
                    *reinterpret_cast<uint2 *>(const_cast<uint16_t *>(vbase) + at) = make_uint2(vw[0], vw[1]);
                }
            }
        }
        for (int hp = 2 * pg_begin + warp; hp < 2 * pg_end; hp += kTW) {  // half-pages of this item, round-robin over the warps
            const int pg = hp >> 1;
            const int page = tcd_page<NB>(p, s, it.b, pg);
            const size_t tile = ((size_t)page * p.kv_page_heads + hk) * (size_t)(kTPT * HD);
            const uint16_t *kp = kbase + tile + p16 * EL;
            const uint16_t *vp = vbase + tile + p16 * EL;
            // a half-page = 4 x 2 tokens: all its K and V loads in flight together, its 8 scores
            // computed before one running-max update (page-at-a-time softmax, see attn.cu)
            {
                const int hb = hp & 1;
                uint4 kr[4];
#pragma unroll
                for (int it = 0; it < 4; ++it) {
                    const int t = hb * 8 + it * 2 + tg;
                    if constexpr (EL == 8) {
                        kr[it] = __ldcg(reinterpret_cast<const uint4 *>(kp + (size_t)t * HD));
                    } else {
                        const uint2 A = __ldcg(reinterpret_cast<const uint2 *>(kp + (size_t)t * HD));
                        kr[it] = make_uint4(A.x, A.y, 0u, 0u);
                    }
                }
                // V rows of the same tokens, issued together with K: one round trip per half-page
                uint4 vr[4];
#pragma unroll
                for (int it = 0; it < 4; ++it) {
                    const int t = hb * 8 + it * 2 + tg;
                    if constexpr (EL == 8) {
                        vr[it] = __ldcg(reinterpret_cast<const uint4 *>(vp + (size_t)t * HD));
                    } else {
                        const uint2 B = __ldcg(reinterpret_cast<const uint2 *>(vp + (size_t)t * HD));
                        vr[it] = make_uint4(B.x, B.y, 0u, 0u);
                    }
                }
                float sc[4][GQ];
#pragma unroll
                for (int it = 0; it < 4; ++it) {
                    const int t = hb * 8 + it * 2 + tg;
                    const bool ok = (pg * kTPT + t) < pos;  // the current token comes from registers (above)
                    const uint32_t kw[4] = {kr[it].x, kr[it].y, kr[it].z, kr[it].w};
                    float kf[EL];
#pragma unroll
                    for (int e = 0; e < EL / 2; ++e) kf[2 * e] = bf16lo(kw[e]), kf[2 * e + 1] = bf16hi(kw[e]);
#pragma unroll
                    for (int g = 0; g < GQ; ++g) {
                        float sv = 0.f;
#pragma unroll
                        for (int e = 0; e < EL; ++e) sv = fmaf(q[g][e], kf[e], sv);
                        sv += __shfl_xor_sync(0xffffffffu, sv, 1);
                        sv += __shfl_xor_sync(0xffffffffu, sv, 2);
                        sv += __shfl_xor_sync(0xffffffffu, sv, 4);
                        sv += __shfl_xor_sync(0xffffffffu, sv, 8);
                        sc[it][g] = ok ? sv : -INFINITY;
                    }
                }
#pragma unroll
                for (int g = 0; g < GQ; ++g) {
                    float mn = m[g];
#pragma unroll
                    for (int it = 0; it < 4; ++it) mn = fmaxf(mn, sc[it][g]);
                    const float corr = mn == -INFINITY ? 1.f : exp2f(m[g] - mn);
                    float ps = 0.f;
#pragma unroll
                    for (int it = 0; it < 4; ++it) {
                        sc[it][g] = mn == -INFINITY ? 0.f : exp2f(sc[it][g] - mn);
                        ps += sc[it][g];
                    }
                    l[g] = l[g] * corr + ps;
                    m[g] = mn;
#pragma unroll
                    for (int e = 0; e < EL; ++e) o[g][e] *= corr;
                }
#pragma unroll
                for (int it = 0; it < 4; ++it) {
                    const uint32_t vw[4] = {vr[it].x, vr[it].y, vr[it].z, vr[it].w};
                    float vf[EL];
#pragma unroll
                    for (int e = 0; e < EL / 2; ++e) vf[2 * e] = bf16lo(vw[e]), vf[2 * e + 1] = bf16hi(vw[e]);
#pragma unroll
                    for (int g = 0; g < GQ; ++g)
#pragma unroll
                        for (int e = 0; e < EL; ++e) o[g][e] = fmaf(sc[it][g], vf[e], o[g][e]);
                }
            }
        }
        if (c.ti >= 0 && l[0] != 1.2345e30f) tcd_trace(p, c.ti, bid, tid, TT_ATT_KV);  // (debug) this warp's pages done
        // merge the two token streams of the warp (lanes 0..15 and 16..31)
#pragma unroll
        for (int g = 0; g < GQ; ++g) {
            const float m2 = __shfl_xor_sync(0xffffffffu, m[g], 16);
            const float l2 = __shfl_xor_sync(0xffffffffu, l[g], 16);
            const float mn = fmaxf(m[g], m2);
            const float c1 = mn == -INFINITY ? 0.f : exp2f(m[g] - mn);
            const float c2 = mn == -INFINITY ? 0.f : exp2f(m2 - mn);
            l[g] = l[g] * c1 + l2 * c2;
            m[g] = mn;
#pragma unroll
            for (int e = 0; e < EL; ++e) {
                const float o2 = __shfl_xor_sync(0xffffffffu, o[g][e], 16);
                o[g][e] = o[g][e] * c1 + o2 * c2;
            }
        }
        if (tg == 0) {
#pragma unroll
            for (int g = 0; g < GQ; ++g) {
                float *d = mrg + ((size_t)warp * GQ + g) * (HD + 2);
                if (p16 == 0) d[HD] = m[g], d[HD + 1] = l[g];
#pragma unroll
                for (int e = 0; e < EL; ++e) d[p16 * EL + e] = o[g][e];
            }
        }
        bar_sync(1, kTC);
        tcd_trace(p, c.ti, bid, tid, TT_ATT_WARPS);  // (debug) all warps done
        // ---- CTA result = merge of the 16 warps; one thread per (q head, dim) ----
        for (int idx = tid; idx < GQ * HD; idx += kTC) {
            const int g = idx / HD, dd = idx % HD;
            float M = -INFINITY;
#pragma unroll
            for (int w = 0; w < kTW; ++w) M = fmaxf(M, mrg[((size_t)w * GQ + g) * (HD + 2) + HD]);
            float L = 0.f, O = 0.f;
            if (M != -INFINITY) {
#pragma unroll
                for (int w = 0; w < kTW; ++w) {
                    const float *sp = mrg + ((size_t)w * GQ + g) * (HD + 2);
                    const float cf = exp2f(sp[HD] - M);
                    L += sp[HD + 1] * cf;
                    O += sp[dd] * cf;
                }
            }
            const int hq = hk * GQ + g;
            if (p.dataflow & 2) {   // (value, tag) pairs, 8-byte stores: o_proj's prologue polls for this layer's tag
                const float tag = (float)(reinterpret_cast<volatile int *>(s.misc)[57] * p.n_layers + reinterpret_cast<volatile int *>(s.misc)[58] + 1);
                float2 *att2 = p.att2 + (size_t)it.b * p.att2_stride;
                float2 *part_o2 = att2 + (size_t)p.n_q * HD;
                if (ns == 1) {
                    att2[(size_t)hq * HD + dd] = make_float2(O / L, tag);
                } else {
                    const size_t pi = (size_t)hq * p.n_splits + split;
                    part_o2[pi * HD + dd] = make_float2(O, tag);
                    if (dd == 0) reinterpret_cast<float4 *>(part_o2 + (size_t)p.n_q * p.n_splits * HD)[pi] = make_float4(M, L, tag, 0.f);
                }
            } else if (ns == 1) {
                p.att[(size_t)it.b * qrows + (size_t)hq * HD + dd] = O / L;
            } else {
                const size_t pi = (size_t)hq * p.n_splits + split;
                p.part_o[(size_t)it.b * p.part_stride + pi * HD + dd] = O;
                if (dd == 0) p.part_ml[(size_t)it.b * p.part_stride + pi * 2] = M, p.part_ml[(size_t)it.b * p.part_stride + pi * 2 + 1] = L;
            }
        }
        // (ns > 1: the split partials are merged by the o_proj phase's prologue, which has to read the
        // attention vector from memory anyway -- no ticket, no second pass here)
        bar_sync(1, kTC);  // merge buffers are reused by the next item / phase
    }
}
