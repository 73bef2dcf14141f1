// mega_attn.cuh -- attention phase of the persistent decode kernel (included by decode_mega.cu).
//
// Latency, not bandwidth, bounds this phase at batch 1 (a few MB of KV per layer), so the chain
// of dependent round trips is kept as short as possible: all 16 compute warps of a CTA read K/V
// rows of the lane's pages straight from L2/HBM into registers (16 lanes per token, each lane
// one 16-byte slice of the row, several tokens in flight per warp), and the split-T factor only
// grows once a warp would have to walk more than two pages, so short contexts need no
// cross-CTA merge at all.  (The standalone kernel in attn.cu, used for batches and prefill,
// stages K/V tiles with TMA instead.)
// Work items = (kv head, split); item i runs on CTA i % G.
#pragma once

__device__ __forceinline__ void prefetch_l2(const void *ptr) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr));
}

// Issued at the start of the qkv phase: pull the K/V pages this CTA's attention item(s) will
// read into L2 while the qkv weights stream, so the attention phase sees L2 instead of HBM
// latency (the pages were last touched a whole step ago and have been evicted since).
template <int HD>
__device__ __forceinline__ void mega_attn_prefetch(const MegaParams &p, const MegaJob &qkv_job, const MegaSmem &s,
                                                   const StepCtx &c, int G, int bid, int tid) {
    constexpr int LINES = kMPT * HD * 2 / 128;  // 128-byte lines per K (or V) page tile
    const int n_pages = c.pos / kMPT + 1;
    const int ns = min(p.n_splits, (n_pages + 2 * kMW - 1) / (2 * kMW));
    const int pps = (n_pages + ns - 1) / ns;
    const int n_items = p.n_kv * ns;
    const uint8_t *kbase = reinterpret_cast<const uint8_t *>(qkv_job.k_pages);
    const uint8_t *vbase = reinterpret_cast<const uint8_t *>(qkv_job.v_pages);
    for (int item = bid; item < n_items; item += G) {
        const int hk = item / ns, split = item % ns;
        const int pg_begin = split * pps, pg_end = min(n_pages, pg_begin + pps);
        for (int i = tid; i < (pg_end - pg_begin) * LINES; i += kMC) {
            const int pg = pg_begin + i / LINES, line = i % LINES;
            const int page = pg < kMaxPtab ? s.ptab[pg] : p.page_table[pg];
            const size_t off = ((size_t)page * p.kv_page_heads + hk) * (size_t)(kMPT * HD * 2) + (size_t)line * 128;
            prefetch_l2(kbase + off);
            prefetch_l2(vbase + off);
        }
    }
}

template <int HD, int GQ>
__device__ void mega_attn_phase(const MegaParams &p, const MegaJob &qkv_job, const MegaSmem &s, StepCtx &c,
                                int G, int bid, int tid) {
    constexpr int EL = HD / 16;  // dims per lane: 16 lanes cover one K/V row
    const int warp = tid >> 5, lane = tid & 31;
    const int tg = lane >> 4, p16 = lane & 15;
    const int n_pages = c.pos / kMPT + 1;
    const int ns = min(p.n_splits, (n_pages + 2 * kMW - 1) / (2 * kMW));
    const int pps = (n_pages + ns - 1) / ns;
    const int n_items = p.n_kv * ns;
    const uint16_t *kbase = reinterpret_cast<const uint16_t *>(qkv_job.k_pages);
    const uint16_t *vbase = reinterpret_cast<const uint16_t *>(qkv_job.v_pages);
    float *mrg = reinterpret_cast<float *>(s.xtab);  // [kMW][GQ][HD + 2]; the operand table is idle here

    for (int item = bid; item < n_items; item += G) {
        const int hk = item / ns, split = item % ns;
        const int pg_begin = split * pps, pg_end = min(n_pages, pg_begin + pps);
        float q[GQ][EL], m[GQ], l[GQ], o[GQ][EL];
#pragma unroll
        for (int g = 0; g < GQ; ++g) {
            const float *qp = p.q + (size_t)(hk * GQ + g) * HD + p16 * EL;
#pragma unroll
            for (int e = 0; e < EL; e += 4) {
                const float4 v = __ldcg(reinterpret_cast<const float4 *>(qp + e));
                q[g][e] = v.x * p.scale_log2, q[g][e + 1] = v.y * p.scale_log2;
                q[g][e + 2] = v.z * p.scale_log2, q[g][e + 3] = v.w * p.scale_log2;
            }
#pragma unroll
            for (int e = 0; e < EL; ++e) o[g][e] = 0.f;
            m[g] = -INFINITY, l[g] = 0.f;
        }
        if (q[0][0] != 1.2345e30f) mega_trace(p, c.ti, bid, tid);  // (debug) q arrived
        for (int pg = pg_begin + warp; pg < pg_end; pg += kMW) {
            const int page = pg < kMaxPtab ? s.ptab[pg] : p.page_table[pg];
            const size_t tile = ((size_t)page * p.kv_page_heads + hk) * (size_t)(kMPT * HD);
            const uint16_t *kp = kbase + tile + p16 * EL;
            const uint16_t *vp = vbase + tile + p16 * EL;
            // two half-pages of 4 x 2 tokens: all loads of a half-page in flight together, its 8
            // scores computed before one running-max update (page-at-a-time softmax, see attn.cu)
#pragma unroll 1
            for (int hb = 0; hb < 2; ++hb) {
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
                float sc[4][GQ];
#pragma unroll
                for (int it = 0; it < 4; ++it) {
                    const int t = hb * 8 + it * 2 + tg;
                    const bool ok = (pg * kMPT + t) <= c.pos;
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
                // V rows of the same tokens: issued now, they arrive while the softmax bookkeeping runs
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
        if (l[0] != 1.2345e30f) mega_trace(p, c.ti, bid, tid);  // (debug) this warp's pages done
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
        bar_sync(1, kMC);
        mega_trace(p, c.ti, bid, tid);  // (debug) all warps done
        // ---- CTA result = merge of the 16 warps; one thread per (q head, dim) ----
        for (int idx = tid; idx < GQ * HD; idx += kMC) {
            const int g = idx / HD, dd = idx % HD;
            float M = -INFINITY;
#pragma unroll
            for (int w = 0; w < kMW; ++w) M = fmaxf(M, mrg[((size_t)w * GQ + g) * (HD + 2) + HD]);
            float L = 0.f, O = 0.f;
            if (M != -INFINITY) {
#pragma unroll
                for (int w = 0; w < kMW; ++w) {
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
            // cross-split merge by the last CTA to arrive for this kv head (fixed split order)
            __threadfence();
            bar_sync(1, kMC);
            if (tid == 0) {
                const unsigned prev = atomicAdd(&p.tickets[hk], 1u);
                s.misc[0] = (prev == (unsigned)(ns - 1));
            }
            bar_sync(1, kMC);
            if (s.misc[0]) {
                __threadfence();
                float *ml = s.part;  // [GQ][ns][2]
                for (int i = tid; i < GQ * ns * 2; i += kMC) {
                    const int g = i / (ns * 2), r = i % (ns * 2);
                    ml[i] = __ldcg(&p.part_ml[((size_t)(hk * GQ + g) * p.n_splits) * 2 + r]);
                }
                bar_sync(1, kMC);
                for (int idx = tid; idx < GQ * HD; idx += kMC) {
                    const int g = idx / HD, dd = idx % HD;
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
        bar_sync(1, kMC);  // merge buffers are reused by the next item / phase
    }
}
