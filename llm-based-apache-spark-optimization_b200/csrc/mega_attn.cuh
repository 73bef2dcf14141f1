// mega_attn.cuh -- attention phase of the persistent decode kernel (included by decode_mega.cu).
// Work items = (kv head, split of the lane's KV pages); item i runs on CTA i % G.  Inside a
// CTA, kAttW warps stream the item's pages (one K+V tile pair per warp in flight, 1-D TMA
// bulk copies), 8 lanes per token as in attn.cu.  The item's partial (m, l, o) goes to the
// global scratch; the last CTA to arrive for a KV head (atomic ticket) merges the splits.
#pragma once

template <int HD, int GQ>
__device__ void mega_attn_phase(const MegaParams &p, const MegaJob &qkv_job, const MegaSmem &s, const StepCtx &c,
                                uint32_t &att_use, int G, int bid, int tid) {
    constexpr int EL = HD / 16;
    constexpr int TILE = kMPT * HD * 2;
    const int warp = tid >> 5, lane = tid & 31;
    const int tg = lane >> 3, p8 = lane & 7;
    const int n_pages = c.pos / kMPT + 1;
    // one page per attention warp and item until the split cap is reached: fewer, fatter items
    // keep the cross-split merge short
    const int ns = min(p.n_splits, (n_pages + kAttW - 1) / kAttW);
    const int pps = (n_pages + ns - 1) / ns;
    const int n_items = p.n_kv * ns;
    const uint8_t *kbase = reinterpret_cast<const uint8_t *>(qkv_job.k_pages);
    const uint8_t *vbase = reinterpret_cast<const uint8_t *>(qkv_job.v_pages);
    float *mrg = s.part;  // [kAttW][GQ][HD + 2]

    for (int item = bid; item < n_items; item += G) {
        const int hk = item / ns, split = item % ns;
        const int pg_begin = split * pps, pg_end = min(n_pages, pg_begin + pps);
        if (warp < kAttW) {
            float qa[GQ][EL], qb[GQ][EL], m[GQ], l[GQ], oa[GQ][EL], ob[GQ][EL];
#pragma unroll
            for (int g = 0; g < GQ; ++g) {
                const float *qp = p.q + (size_t)(hk * GQ + g) * HD;
#pragma unroll
                for (int e = 0; e < EL; ++e) {
                    qa[g][e] = __ldcg(qp + p8 * EL + e) * p.scale_log2;
                    qb[g][e] = __ldcg(qp + HD / 2 + p8 * EL + e) * p.scale_log2;
                    oa[g][e] = 0.f, ob[g][e] = 0.f;
                }
                m[g] = -INFINITY, l[g] = 0.f;
            }
            uint8_t *ks = s.xtab + (size_t)warp * 2 * TILE;
            uint8_t *vs = ks + TILE;
            for (int pg = pg_begin + warp; pg < pg_end; pg += kAttW) {
                if (lane == 0) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    const size_t off = ((size_t)p.page_table[pg] * p.n_kv + hk) * (size_t)TILE;
                    mbar_expect_tx(&s.att_full[warp], 2 * TILE);
                    tma_load_1d(ks, kbase + off, TILE, &s.att_full[warp]);
                    tma_load_1d(vs, vbase + off, TILE, &s.att_full[warp]);
                }
                mbar_wait(&s.att_full[warp], att_use & 1);
                ++att_use;
#pragma unroll
                for (int it = 0; it < kMPT / 4; ++it) {
                    const int t = it * 4 + tg;
                    const bool ok = (pg * kMPT + t) <= c.pos;
                    float ka[EL], kb[EL], va[EL], vb[EL];
                    if constexpr (EL == 8) {
                        const uint4 A = *reinterpret_cast<const uint4 *>(ks + (size_t)t * HD * 2 + p8 * 16);
                        const uint4 B = *reinterpret_cast<const uint4 *>(ks + (size_t)t * HD * 2 + HD + p8 * 16);
                        const uint4 C = *reinterpret_cast<const uint4 *>(vs + (size_t)t * HD * 2 + p8 * 16);
                        const uint4 D = *reinterpret_cast<const uint4 *>(vs + (size_t)t * HD * 2 + HD + p8 * 16);
                        const uint32_t aw[4] = {A.x, A.y, A.z, A.w}, bw[4] = {B.x, B.y, B.z, B.w};
                        const uint32_t cw[4] = {C.x, C.y, C.z, C.w}, dw[4] = {D.x, D.y, D.z, D.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            ka[2 * e] = bf16lo(aw[e]), ka[2 * e + 1] = bf16hi(aw[e]);
                            kb[2 * e] = bf16lo(bw[e]), kb[2 * e + 1] = bf16hi(bw[e]);
                            va[2 * e] = bf16lo(cw[e]), va[2 * e + 1] = bf16hi(cw[e]);
                            vb[2 * e] = bf16lo(dw[e]), vb[2 * e + 1] = bf16hi(dw[e]);
                        }
                    } else {
                        const uint2 A = *reinterpret_cast<const uint2 *>(ks + (size_t)t * HD * 2 + p8 * 8);
                        const uint2 B = *reinterpret_cast<const uint2 *>(ks + (size_t)t * HD * 2 + HD + p8 * 8);
                        const uint2 C = *reinterpret_cast<const uint2 *>(vs + (size_t)t * HD * 2 + p8 * 8);
                        const uint2 D = *reinterpret_cast<const uint2 *>(vs + (size_t)t * HD * 2 + HD + p8 * 8);
                        ka[0] = bf16lo(A.x), ka[1] = bf16hi(A.x), ka[2] = bf16lo(A.y), ka[3] = bf16hi(A.y);
                        kb[0] = bf16lo(B.x), kb[1] = bf16hi(B.x), kb[2] = bf16lo(B.y), kb[3] = bf16hi(B.y);
                        va[0] = bf16lo(C.x), va[1] = bf16hi(C.x), va[2] = bf16lo(C.y), va[3] = bf16hi(C.y);
                        vb[0] = bf16lo(D.x), vb[1] = bf16hi(D.x), vb[2] = bf16lo(D.y), vb[3] = bf16hi(D.y);
                    }
#pragma unroll
                    for (int g = 0; g < GQ; ++g) {
                        float sc = 0.f;
#pragma unroll
                        for (int e = 0; e < EL; ++e) sc = fmaf(qa[g][e], ka[e], fmaf(qb[g][e], kb[e], sc));
                        sc += __shfl_xor_sync(0xffffffffu, sc, 1);
                        sc += __shfl_xor_sync(0xffffffffu, sc, 2);
                        sc += __shfl_xor_sync(0xffffffffu, sc, 4);
                        if (ok) {
                            const float mn = fmaxf(m[g], sc);
                            const float corr = exp2f(m[g] - mn), pr = exp2f(sc - mn);
                            l[g] = l[g] * corr + pr;
#pragma unroll
                            for (int e = 0; e < EL; ++e) {
                                oa[g][e] = fmaf(pr, va[e], oa[g][e] * corr);
                                ob[g][e] = fmaf(pr, vb[e], ob[g][e] * corr);
                            }
                            m[g] = mn;
                        }
                    }
                }
                __syncwarp();
            }
            // merge the 4 token-group streams of the warp
#pragma unroll
            for (int off = 8; off <= 16; off <<= 1) {
#pragma unroll
                for (int g = 0; g < GQ; ++g) {
                    const float m2 = __shfl_xor_sync(0xffffffffu, m[g], off);
                    const float l2 = __shfl_xor_sync(0xffffffffu, l[g], off);
                    const float mn = fmaxf(m[g], m2);
                    const float c1 = mn == -INFINITY ? 0.f : exp2f(m[g] - mn);
                    const float c2 = mn == -INFINITY ? 0.f : exp2f(m2 - mn);
                    l[g] = l[g] * c1 + l2 * c2;
                    m[g] = mn;
#pragma unroll
                    for (int e = 0; e < EL; ++e) {
                        const float o2a = __shfl_xor_sync(0xffffffffu, oa[g][e], off);
                        const float o2b = __shfl_xor_sync(0xffffffffu, ob[g][e], off);
                        oa[g][e] = oa[g][e] * c1 + o2a * c2;
                        ob[g][e] = ob[g][e] * c1 + o2b * c2;
                    }
                }
            }
            if (tg == 0) {
#pragma unroll
                for (int g = 0; g < GQ; ++g) {
                    float *d = mrg + ((size_t)warp * GQ + g) * (HD + 2);
                    if (p8 == 0) d[HD] = m[g], d[HD + 1] = l[g];
#pragma unroll
                    for (int e = 0; e < EL; ++e) d[p8 * EL + e] = oa[g][e], d[HD / 2 + p8 * EL + e] = ob[g][e];
                }
            }
        }
        bar_sync(1, kMC);
        // ---- CTA partial = merge of the kAttW warps; one thread per (q head, dim) ----
        if (tid < GQ * HD) {
            const int g = tid / HD, dd = tid % HD;
            float M = -INFINITY;
#pragma unroll
            for (int w = 0; w < kAttW; ++w) M = fmaxf(M, mrg[((size_t)w * GQ + g) * (HD + 2) + HD]);
            float L = 0.f, O = 0.f;
            if (M != -INFINITY) {
#pragma unroll
                for (int w = 0; w < kAttW; ++w) {
                    const float *sp = mrg + ((size_t)w * GQ + g) * (HD + 2);
                    const float cf = exp2f(sp[HD] - M);
                    L += sp[HD + 1] * cf;
                    O += sp[dd] * cf;
                }
            }
            const int hq = hk * GQ + g;
            if (ns == 1) {
                p.att[(size_t)hq * HD + dd] = O / L;
            } else {
                const size_t pi = (size_t)hq * p.n_splits + split;
                p.part_o[pi * HD + dd] = O;
                if (dd == 0) p.part_ml[pi * 2] = M, p.part_ml[pi * 2 + 1] = L;
            }
        }
        if (ns > 1) {
            __threadfence();
            bar_sync(1, kMC);
            if (tid == 0) {
                const unsigned prev = atomicAdd(&p.tickets[hk], 1u);
                s.misc[0] = (prev == (unsigned)(ns - 1));
            }
            bar_sync(1, kMC);
            if (s.misc[0]) {
                __threadfence();
                // (m, l) of every split -> shared memory in one round trip, then one pass over part_o
                float *ml = mrg;  // [GQ][ns][2], the merge buffer is free again
                for (int i = tid; i < GQ * ns * 2; i += kMC) {
                    const int g = i / (ns * 2), r = i % (ns * 2);
                    ml[i] = __ldcg(&p.part_ml[((size_t)(hk * GQ + g) * p.n_splits) * 2 + r]);
                }
                bar_sync(1, kMC);
                if (tid < GQ * HD) {
                    const int g = tid / HD, dd = tid % HD;
                    const size_t pb = (size_t)(hk * GQ + g) * p.n_splits;
                    float M = -INFINITY;
                    for (int sp = 0; sp < ns; ++sp) M = fmaxf(M, ml[(g * ns + sp) * 2]);
                    float L = 0.f, O = 0.f;
#pragma unroll 8
                    for (int sp = 0; sp < ns; ++sp) {
                        const float ms = ml[(g * ns + sp) * 2];
                        const float cf = ms == -INFINITY ? 0.f : exp2f(ms - M);
                        L += ml[(g * ns + sp) * 2 + 1] * cf;
                        O += __ldcg(&p.part_o[(pb + sp) * HD + dd]) * cf;
                    }
                    p.att[(size_t)(hk * GQ + g) * HD + dd] = O / L;
                }
                if (tid == 0) p.tickets[hk] = 0;
            }
        }
        bar_sync(1, kMC);  // merge buffer and misc are reused by the next item / phase
    }
}
