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

template <int HD, int GQ>
__device__ void mega_attn_phase(const MegaParams &p, const MegaJob &qkv_job, const MegaSmem &s, const StepCtx &c,
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
        for (int pg = pg_begin + warp; pg < pg_end; pg += kMW) {
            const size_t tile = ((size_t)p.page_table[pg] * p.n_kv + hk) * (size_t)(kMPT * HD);
            const uint16_t *kp = kbase + tile + p16 * EL;
            const uint16_t *vp = vbase + tile + p16 * EL;
#pragma unroll 4
            for (int it = 0; it < kMPT / 2; ++it) {
                const int t = it * 2 + tg;
                const bool ok = (pg * kMPT + t) <= c.pos;
                float kf[EL], vf[EL];
                if constexpr (EL == 8) {
                    const uint4 A = __ldcg(reinterpret_cast<const uint4 *>(kp + (size_t)t * HD));
                    const uint4 B = __ldcg(reinterpret_cast<const uint4 *>(vp + (size_t)t * HD));
                    const uint32_t aw[4] = {A.x, A.y, A.z, A.w}, bw[4] = {B.x, B.y, B.z, B.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        kf[2 * e] = bf16lo(aw[e]), kf[2 * e + 1] = bf16hi(aw[e]);
                        vf[2 * e] = bf16lo(bw[e]), vf[2 * e + 1] = bf16hi(bw[e]);
                    }
                } else {
                    const uint2 A = __ldcg(reinterpret_cast<const uint2 *>(kp + (size_t)t * HD));
                    const uint2 B = __ldcg(reinterpret_cast<const uint2 *>(vp + (size_t)t * HD));
                    kf[0] = bf16lo(A.x), kf[1] = bf16hi(A.x), kf[2] = bf16lo(A.y), kf[3] = bf16hi(A.y);
                    vf[0] = bf16lo(B.x), vf[1] = bf16hi(B.x), vf[2] = bf16lo(B.y), vf[3] = bf16hi(B.y);
                }
#pragma unroll
                for (int g = 0; g < GQ; ++g) {
                    float sc = 0.f;
#pragma unroll
                    for (int e = 0; e < EL; ++e) sc = fmaf(q[g][e], kf[e], sc);
                    sc += __shfl_xor_sync(0xffffffffu, sc, 1);
                    sc += __shfl_xor_sync(0xffffffffu, sc, 2);
                    sc += __shfl_xor_sync(0xffffffffu, sc, 4);
                    sc += __shfl_xor_sync(0xffffffffu, sc, 8);
                    if (ok) {
                        const float mn = fmaxf(m[g], sc);
                        const float corr = exp2f(m[g] - mn), pr = exp2f(sc - mn);
                        l[g] = l[g] * corr + pr;
#pragma unroll
                        for (int e = 0; e < EL; ++e) o[g][e] = fmaf(pr, vf[e], o[g][e] * corr);
                        m[g] = mn;
                    }
                }
            }
        }
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
        // ---- CTA result = merge of the 16 warps; one thread per (q head, dim) ----
        if (tid < GQ * HD) {
            const int g = tid / HD, dd = tid % HD;
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
        bar_sync(1, kMC);  // merge buffers are reused by the next item / phase
    }
}
