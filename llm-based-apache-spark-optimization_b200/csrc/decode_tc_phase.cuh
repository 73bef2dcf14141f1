// decode_tc_phase.cuh -- compute-warp side of one weight-streaming phase of decode_tc.cu:
// the prologue that turns the activation vector into tcgen05 B-operand blocks, and the
// epilogue that drains finished accumulators out of TMEM.
#pragma once

// ---- prologue -------------------------------------------------------------------------------
// Operand block `slot` (1 KB, 1024-byte aligned) is an [8 rows x 64 k] bf16 K-major tile with the
// 128-byte swizzle: row r at +128 r, its 16-byte chunk c at chunk position c ^ r.  Row 0 holds
// the bf16 hi parts of 64 activations, row 1 the lo parts (x - hi), rows 2..7 stay zero.
// A CTA only needs the k blocks its unit range touches: block kb goes to slot
// (kb - kb_start) mod nkb, where kb_start is the k block of the range's first unit.
// Returns the RMSNorm scale rsqrt(mean(x^2) + eps) (1 when the phase has no norm); the scale
// is applied in the epilogue (it is a scalar factor of the whole dot product).
// How many CTAs add partial sums to the rows of row tile `tile` of an [N, K] phase split stream-K over G CTAs
// (unit = tile * nkb + kb; the first `rem` CTAs take base + 1 units, the others base).
__device__ __forceinline__ float tcd_tile_contributors(int N, int K, int tile, int G) {
    const int nkb = K / 64, n_units = (N / 128) * nkb;
    const int base = n_units / G, rem = n_units % G, cut = rem * (base + 1);
    const int u0 = tile * nkb, u1 = u0 + nkb - 1;
    const int c0 = u0 < cut ? u0 / (base + 1) : rem + (u0 - cut) / base;
    const int c1 = u1 < cut ? u1 / (base + 1) : rem + (u1 - cut) / base;
    return (float)(c1 - c0 + 1);
}

template <int NB>
struct TcdScale {
    float v[NB];
};

// lane b's 8 activations of chunk `chunk` (k = 8 chunk .. 8 chunk + 7) of this phase's input, fp32
template <int NB>
__device__ __forceinline__ void tcd_load_chunk(const TcdParams &p, const TcdSmem &s, const TcdCtxT<NB> &c, int x_kind, int b, int chunk,
                                               int jidx, int G, int bid, float (&v)[8]) {
    if (x_kind == TCD_X_EMBED) {
        const uint16_t *emb = reinterpret_cast<const uint16_t *>(p.embed) + (size_t)tcd_tok<NB>(c, s, b) * p.d;
        const uint4 u = *reinterpret_cast<const uint4 *>(emb + chunk * 8);
        v[0] = bf16lo(u.x), v[1] = bf16hi(u.x), v[2] = bf16lo(u.y), v[3] = bf16hi(u.y);
        v[4] = bf16lo(u.z), v[5] = bf16hi(u.z), v[6] = bf16lo(u.w), v[7] = bf16hi(u.w);
        if ((chunk % G) == bid) {  // residual stream h <- embedding row, each element written once
            float *hb = p.h + (size_t)b * p.d;
            *reinterpret_cast<float4 *>(hb + chunk * 8) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4 *>(hb + chunk * 8 + 4) = make_float4(v[4], v[5], v[6], v[7]);
        }
    } else if (x_kind == TCD_X_SWIGLU && (p.dataflow & 1)) {
        // (sum, count) pairs of 8 gate/up pairs: poll until every count says its tile's contributors are all in
        const float want = tcd_tile_contributors(2 * p.ffn, p.d, chunk >> 3, G);
        const float2 *gp = p.gu2 + (size_t)b * 2 * p.ffn + (size_t)chunk * 16;
        float4 q[8];
        for (;;) {
#pragma unroll
            for (int e = 0; e < 8; ++e)
                asm volatile("ld.relaxed.gpu.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                             : "=f"(q[e].x), "=f"(q[e].y), "=f"(q[e].z), "=f"(q[e].w) : "l"(gp + 2 * e) : "memory");
            bool ok = true;
#pragma unroll
            for (int e = 0; e < 8; ++e) ok = ok && q[e].y == want && q[e].w == want;
            if (ok) break;
            // not there yet: watch one pair of counts (16 bytes instead of 128 per probe) until it is, then look again
            float4 w1;
            do {
                __nanosleep(64);
                asm volatile("ld.relaxed.gpu.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                             : "=f"(w1.x), "=f"(w1.y), "=f"(w1.z), "=f"(w1.w) : "l"(gp + 14) : "memory");
            } while (w1.y != want || w1.w != want);
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = (q[e].x / (1.f + expf(-q[e].x))) * q[e].z;
    } else if (x_kind == TCD_X_SWIGLU) {
        const float4 *gp = reinterpret_cast<const float4 *>(p.gu + (size_t)b * 2 * p.ffn + (size_t)chunk * 16);
        const float4 a = __ldcg(gp), bb = __ldcg(gp + 1), cc = __ldcg(gp + 2), d = __ldcg(gp + 3);
        v[0] = (a.x / (1.f + expf(-a.x))) * a.y, v[1] = (a.z / (1.f + expf(-a.z))) * a.w;
        v[2] = (bb.x / (1.f + expf(-bb.x))) * bb.y, v[3] = (bb.z / (1.f + expf(-bb.z))) * bb.w;
        v[4] = (cc.x / (1.f + expf(-cc.x))) * cc.y, v[5] = (cc.z / (1.f + expf(-cc.z))) * cc.w;
        v[6] = (d.x / (1.f + expf(-d.x))) * d.y, v[7] = (d.z / (1.f + expf(-d.z))) * d.w;
    } else if (x_kind == TCD_X_ATT && (p.dataflow & 2)) {
        // (value, tag) pairs written by the attention items of this layer: poll for the layer's tag (no grid barrier
        // between the attention phase and this one); split-T partials are merged as below
        const int att_ns = tcd_att_ns(p, tcd_pos<NB>(c, s, b));
        const float tag = (float)(reinterpret_cast<volatile int *>(s.misc)[57] * p.n_layers + (jidx >> 2) + 1);
        const int hd = p.head_dim, hq = (chunk * 8) / hd, dd = (chunk * 8) % hd;
        const float2 *att2 = p.att2 + (size_t)b * p.att2_stride;
        if (att_ns == 1) {
            const float2 *ap = att2 + (size_t)chunk * 8;
            float4 q[4];
            for (;;) {
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    asm volatile("ld.relaxed.gpu.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                                 : "=f"(q[e].x), "=f"(q[e].y), "=f"(q[e].z), "=f"(q[e].w) : "l"(ap + 2 * e) : "memory");
                bool ok = true;
#pragma unroll
                for (int e = 0; e < 4; ++e) ok = ok && q[e].y == tag && q[e].w == tag;
                if (ok) break;
                __nanosleep(32);
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) v[2 * e] = q[e].x, v[2 * e + 1] = q[e].z;
        } else {
            const float2 *part_o2 = att2 + (size_t)p.n_q * hd;
            const float4 *part_ml2 = reinterpret_cast<const float4 *>(part_o2 + (size_t)p.n_q * p.n_splits * hd);
            const size_t pb = (size_t)hq * p.n_splits;
            float M = -INFINITY, L = 0.f;
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = 0.f;
            for (int sp = 0; sp < att_ns; ++sp) {
                float4 ml, q[4];
                for (;;) {
                    asm volatile("ld.relaxed.gpu.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                                 : "=f"(ml.x), "=f"(ml.y), "=f"(ml.z), "=f"(ml.w) : "l"(part_ml2 + pb + sp) : "memory");
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        asm volatile("ld.relaxed.gpu.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                                     : "=f"(q[e].x), "=f"(q[e].y), "=f"(q[e].z), "=f"(q[e].w)
                                     : "l"(part_o2 + (pb + sp) * hd + dd + 2 * e) : "memory");
                    bool ok = ml.z == tag;
#pragma unroll
                    for (int e = 0; e < 4; ++e) ok = ok && q[e].y == tag && q[e].w == tag;
                    if (ok) break;
                    __nanosleep(32);
                }
                const float mn = fmaxf(M, ml.x);
                const float c_old = mn == -INFINITY ? 1.f : exp2f(M - mn);
                const float c_new = mn == -INFINITY ? 0.f : exp2f(ml.x - mn);
                L = L * c_old + ml.y * c_new;
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    v[2 * e] = v[2 * e] * c_old + q[e].x * c_new;
                    v[2 * e + 1] = v[2 * e + 1] * c_old + q[e].z * c_new;
                }
                M = mn;
            }
            const float inv = 1.f / L;
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] *= inv;
        }
    } else if (x_kind == TCD_X_ATT && tcd_att_ns(p, tcd_pos<NB>(c, s, b)) > 1) {
        // split-T attention partials of this head -> merged output (running max, fixed split order)
        const int att_ns = tcd_att_ns(p, tcd_pos<NB>(c, s, b));
        const int hd = p.head_dim, hq = (chunk * 8) / hd, dd = (chunk * 8) % hd;
        const size_t pb = (size_t)hq * p.n_splits;
        const float *part_o = p.part_o + (size_t)b * p.part_stride, *part_ml = p.part_ml + (size_t)b * p.part_stride;
        float M = -INFINITY, L = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = 0.f;
#pragma unroll 4
        for (int sp = 0; sp < att_ns; ++sp) {
            const float2 ml = __ldcg(reinterpret_cast<const float2 *>(part_ml + (pb + sp) * 2));
            const float4 a = __ldcg(reinterpret_cast<const float4 *>(part_o + (pb + sp) * hd + dd));
            const float4 bb = __ldcg(reinterpret_cast<const float4 *>(part_o + (pb + sp) * hd + dd + 4));
            const float mn = fmaxf(M, ml.x);
            const float c_old = mn == -INFINITY ? 1.f : exp2f(M - mn);
            const float c_new = mn == -INFINITY ? 0.f : exp2f(ml.x - mn);
            L = L * c_old + ml.y * c_new;
            v[0] = v[0] * c_old + a.x * c_new, v[1] = v[1] * c_old + a.y * c_new;
            v[2] = v[2] * c_old + a.z * c_new, v[3] = v[3] * c_old + a.w * c_new;
            v[4] = v[4] * c_old + bb.x * c_new, v[5] = v[5] * c_old + bb.y * c_new;
            v[6] = v[6] * c_old + bb.z * c_new, v[7] = v[7] * c_old + bb.w * c_new;
            M = mn;
        }
        const float inv = 1.f / L;
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] *= inv;
    } else {
        const float *xsrc = x_kind == TCD_X_ATT ? p.att + (size_t)b * p.n_q * p.head_dim : p.h + (size_t)b * p.d;
        const float4 a = __ldcg(reinterpret_cast<const float4 *>(xsrc + chunk * 8));
        const float4 bb = __ldcg(reinterpret_cast<const float4 *>(xsrc + chunk * 8 + 4));
        v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w, v[4] = bb.x, v[5] = bb.y, v[6] = bb.z, v[7] = bb.w;
    }
}

template <int NB>
__device__ TcdScale<NB> tcd_prologue(const TcdParams &p, const TcdJob &job, const TcdSmem &s, TcdCtxT<NB> &c, const TRange &r,
                                     int G, int bid, int tid) {
    const int warp = tid >> 5, lane = tid & 31;
    const int K = job.K, nkb = r.nkb;
    const int kb_start = r.g0 % nkb;
    const int n_slots = min(r.g1 - r.g0, nkb);
    const int x_kind = job.x_src & 0xff;
    const uint16_t *gain = reinterpret_cast<const uint16_t *>(job.gain);
    const int jidx = (int)(&job - s.jobs);
    const int nb = NB == 1 ? 1 : p.nb;

    {   // the RMSNorm gain vectors of the next two phases are cold (read once per step): pull them into L2
        const TcdJob *first = s.jobs, *last = s.jobs + 4 * p.n_layers;
#pragma unroll
        for (int a = 1; a <= 2; ++a) {
            const TcdJob *nj = &job + a;
            if (nj > last) nj = first + (nj - last - 1);
            if (nj->gain != nullptr && tid * 64 < nj->K)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const uint16_t *>(nj->gain) + tid * 64));
        }
    }
    if ((job.x_src & 0x400) && p.tp_size > 1) {  // tensor parallel: clear the partial-sum vector of the next row-parallel phase
        const int nd4 = (p.dataflow & 8) ? p.d / 2 : p.d / 4;   // (sum, count) pairs when the reduction is triggered by the data
        for (int i = bid * kTC + tid; i < nd4; i += G * kTC) reinterpret_cast<float4 *>(p.delta)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    // Clear the sum buffers the next qkv / gate-up phases add into, each while nobody can be reading it and a grid barrier
    // before it is added to again: gate/up's in the o_proj phase (bit 8; its readers were the previous layer's down_proj
    // prologues), qkv's in the gate/up phase (bit 9; its readers were this layer's attention items: after the barrier that
    // ends the o_proj phase every CTA is past its attention item, and the barrier after down_proj comes before the next adds).
    // (the lanes' buffers are contiguous)
    if (job.x_src & 0x100) {
        const int ng4 = nb * ((p.dataflow & 1) ? 4 * p.ffn / 4 : 2 * p.ffn / 4);
        float4 *gz = (p.dataflow & 1) ? reinterpret_cast<float4 *>(p.gu2) : reinterpret_cast<float4 *>(p.gu);
        for (int i = bid * kTC + tid; i < ng4; i += G * kTC) gz[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (job.x_src & 0x200) {
        const int nq4 = nb * ((p.n_q + 2 * p.n_kv) * p.head_dim / ((p.dataflow & 4) ? 2 : 4));
        float4 *qz = (p.dataflow & 4) ? reinterpret_cast<float4 *>(p.qkv2) : reinterpret_cast<float4 *>(p.qkv);
        for (int i = bid * kTC + tid; i < nq4; i += G * kTC) qz[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    // phases with a norm (or that seed the residual stream from the embedding) read the whole
    // vector; the others only the chunks (8 values) of the k blocks this CTA streams
    // The operand blocks below are still being READ by the previous phase's MMAs until that phase's accumulators are
    // drained.  Where a grid barrier separates the phases, the epilogue warps have passed it; the attention edges have a
    // CTA barrier of their own (tcd_attn_wait_qkv).  On the gate/up -> down edge nothing did: the other warps came
    // straight from gate/up's prologue, polled, and could overwrite blocks that late gate/up tiles of THIS CTA were
    // still multiplying (the two phases' slots overlap when a CTA's gate/up share wraps around its k blocks: the 7B; the
    // stress run with two lanes found it).  So: wait for the epilogue warps, which have seen every accumulator.
    if (x_kind == TCD_X_SWIGLU && (p.dataflow & 1)) bar_sync(1, kTC);
    const bool full_scan = gain != nullptr || x_kind == TCD_X_EMBED;
    const int n_iter = full_scan ? K / 8 : n_slots * 8;
    float ss[NB];
#pragma unroll
    for (int b = 0; b < NB; ++b) ss[b] = 0.f;
    // 8 activations of lane b -> its sum of squares, and (when the CTA streams this k block) rows 2b (hi) / 2b + 1 (lo) of the
    // operand block; chunk c8 of row r sits at chunk position c8 ^ r
    auto emit = [&](int b, int chunk, int slot, float (&v)[8], const uint4 &gn) {
        float sq = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) sq = fmaf(v[e], v[e], sq);
        if constexpr (NB == 1) {
            ss[0] += sq;
        } else {
#pragma unroll
            for (int l = 0; l < NB; ++l) ss[l] += l == b ? sq : 0.f;
        }
        if (slot >= n_slots) return;
        if (gain != nullptr) {
            v[0] *= bf16lo(gn.x), v[1] *= bf16hi(gn.x), v[2] *= bf16lo(gn.y), v[3] *= bf16hi(gn.y);
            v[4] *= bf16lo(gn.z), v[5] *= bf16hi(gn.z), v[6] *= bf16lo(gn.w), v[7] *= bf16hi(gn.w);
        }
        float hi[8], lo[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) hi[e] = tcd_bf16_round(v[e]), lo[e] = v[e] - hi[e];
        uint8_t *blk = s.bop + (size_t)slot * 1024 + b * 256;
        const int c8 = chunk & 7;
        *reinterpret_cast<uint4 *>(blk + ((c8 ^ (2 * b)) * 16)) = make_uint4(tcd_pack_bf16(hi[0], hi[1]), tcd_pack_bf16(hi[2], hi[3]),
                                                                             tcd_pack_bf16(hi[4], hi[5]), tcd_pack_bf16(hi[6], hi[7]));
        *reinterpret_cast<uint4 *>(blk + 128 + ((c8 ^ (2 * b + 1)) * 16)) =
            make_uint4(tcd_pack_bf16(lo[0], lo[1]), tcd_pack_bf16(lo[2], lo[3]), tcd_pack_bf16(lo[4], lo[5]),
                       tcd_pack_bf16(lo[6], lo[7]));
    };
    auto place = [&](int idx, int &chunk, int &slot) {   // work item -> chunk (8 values of k) and its operand block
        if (full_scan) {
            chunk = idx;
            slot = (chunk >> 3) - kb_start;
            slot += slot < 0 ? nkb : 0;
        } else {
            slot = idx >> 3;
            int kb = kb_start + slot;
            kb -= kb >= nkb ? nkb : 0;
            chunk = kb * 8 + (idx & 7);
        }
    };
    if (NB > 1 && x_kind == TCD_X_H) {
        // several lanes, plain loads of the residual stream (the norm phases scan all of it): every lane's loads of a chunk
        // are in flight together -- one L2 round trip per chunk, not one per lane
        for (int idx = tid; idx < n_iter; idx += kTC) {
            int chunk, slot;
            place(idx, chunk, slot);
            uint4 gn = make_uint4(0u, 0u, 0u, 0u);
            if (gain != nullptr) gn = *reinterpret_cast<const uint4 *>(gain + chunk * 8);
            float4 xa[NB], xb[NB];
#pragma unroll
            for (int b = 0; b < NB; ++b) {   // (a lane that is not there re-reads lane 0: predicating the loads made ptxas spill 600 bytes)
                const float *src = p.h + (size_t)(b < nb ? b : 0) * p.d + chunk * 8;
                xa[b] = __ldcg(reinterpret_cast<const float4 *>(src)), xb[b] = __ldcg(reinterpret_cast<const float4 *>(src + 4));
            }
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                if (b >= nb) break;
                float v[8] = {xa[b].x, xa[b].y, xa[b].z, xa[b].w, xb[b].x, xb[b].y, xb[b].z, xb[b].w};
                emit(b, chunk, slot, v, gn);
            }
        }
    } else {
        // work items = (lane, chunk), lane-major: with several lanes the threads take different lanes' chunks side by side
        // instead of one thread walking the lanes of a chunk one L2 round trip after the other
        const int n_items = NB == 1 ? n_iter : n_iter * nb;
#pragma unroll 2
        for (int it = tid; it < n_items; it += kTC) {
            int b = 0, idx = it;
            if constexpr (NB > 1) b = it / n_iter, idx = it - b * n_iter;
            int chunk, slot;
            place(idx, chunk, slot);
            // issued with the activation loads below: behind the `continue` it would cost a second L2 round trip
            uint4 gn = make_uint4(0u, 0u, 0u, 0u);
            if (gain != nullptr) gn = *reinterpret_cast<const uint4 *>(gain + chunk * 8);
            float v[8];
            tcd_load_chunk<NB>(p, s, c, x_kind, b, chunk, jidx, G, bid, v);
            emit(b, chunk, slot, v, gn);
        }
    }
    // generic-proxy writes above -> visible to the async proxy (tcgen05.mma reads the blocks)
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (gain != nullptr) {
#pragma unroll
        for (int b = 0; b < NB; ++b) {
            const float t = warp_sum(ss[b]);
            if (lane == 0) s.red[b * kTW + warp] = t;
        }
    }
    tcd_trace(p, c.ti, bid, tid, TT_XCONV);
    __syncwarp();
    if (lane == 0) mbar_arrive(s.xrdy);   // to the MMA issuer (which polls, moving weight stages meanwhile): this warp's operand blocks are complete
    bar_sync(1, kTC);                     // the norm partials of all warps
    TcdScale<NB> sc;
#pragma unroll
    for (int b = 0; b < NB; ++b) {
        sc.v[b] = 1.f;
        if (gain != nullptr) {
            float tot = 0.f;
#pragma unroll
            for (int w = 0; w < kTW; ++w) tot += s.red[b * kTW + w];
            sc.v[b] = rsqrtf(tot / (float)K + p.rms_eps);
        }
    }
    return sc;
}

// ---- epilogue (warps 0..3) ----------------------------------------------------------------------
// One accumulator per (row tile) segment of the CTA's range: lane i of warp w owns TMEM lane
// 32 w + i = weight row 128 t + 32 w + i; columns 0 and 1 are the hi and lo partial dot products.
template <int NB>
__device__ void tcd_epilogue(const TcdParams &p, const TcdJob &job, const TcdSmem &s, TcdCtxT<NB> &c, const TRange &r,
                             const TcdScale<NB> &sc, int warp, int lane, uint32_t tmem) {
    const int nb = NB == 1 ? 1 : p.nb;
    // lane b's rows of this phase's output: the next lane's start one output vector further
    const size_t ostride = job.epi == TCD_EPI_LOGITS ? (size_t)p.vocab : (size_t)job.N;
    for (int g = r.g0; g < r.g1;) {
        const int t = g / r.nkb;
        const int seg_end = min(r.g1, (t + 1) * r.nkb);
        const uint32_t aslot = c.as % kTAcc;
        mbar_wait(&s.acc_full[aslot], (c.as / kTAcc) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        // columns 2b / 2b + 1: hi / lo sums of lane b from the first issuer's stages, 8 + 2b / 9 + 2b: the second issuer's
        // (if the segment has > 1 stage)
        uint32_t a0[2 * NB], a1[2 * NB];
        const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + aslot * kTAccCols;
        if constexpr (NB == 1) {
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(a0[0]), "=r"(a0[1]) : "r"(taddr) : "memory");
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(a1[0]), "=r"(a1[1]) : "r"(taddr + 8) : "memory");
        } else {
            static_assert(NB == 1 || NB == 4, "one or four lanes");
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                         : "=r"(a0[0]), "=r"(a0[1]), "=r"(a0[2]), "=r"(a0[3]), "=r"(a0[4]), "=r"(a0[5]), "=r"(a0[6]), "=r"(a0[7]) : "r"(taddr) : "memory");
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                         : "=r"(a1[0]), "=r"(a1[1]), "=r"(a1[2]), "=r"(a1[3]), "=r"(a1[4]), "=r"(a1[5]), "=r"(a1[6]), "=r"(a1[7]) : "r"(taddr + 8) : "memory");
        }
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(&s.acc_empty[aslot]);
        const int n = t * 128 + warp * 32 + lane;
#pragma unroll
        for (int b = 0; b < NB; ++b) {
            if (NB > 1 && b >= nb) break;
            float v = __uint_as_float(a0[2 * b]) + __uint_as_float(a0[2 * b + 1]);
            if (kTIssuers > 1 && seg_end - g > 1) v += __uint_as_float(a1[2 * b]) + __uint_as_float(a1[2 * b + 1]);
            v *= sc.v[b];
            if (job.epi == TCD_EPI_ADD) {
                asm volatile("red.global.add.f32 [%0], %1;" ::"l"(job.out + b * ostride + n), "f"(v) : "memory");
            } else if (job.epi == TCD_EPI_ADD_CNT) {   // sum and contribution count in one 8-byte vector red
                asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(reinterpret_cast<float2 *>(job.out) + b * ostride + n), "f"(v), "f"(1.f) : "memory");
            } else if (n < p.vocab) {   // rows past the vocabulary pad the last tile
                job.out[b * ostride + n] = v;
                if (v > c.best[b]) c.best[b] = v, c.best_i[b] = n;
            }
        }
        ++c.as;
        g = seg_end;
    }
}

// ---- tensor parallel: all-reduce phase ----------------------------------------------------------------
// h += sum over ranks of delta, the exchange of tp_ar.cu as a phase of the persistent kernel (after the
// grid barrier that completes this rank's delta): CTA j pushes slice j (1024 floats) into every peer's
// inbox, publishes the sequence number, waits for the peers' flags in its own memory, adds in rank order.
__device__ __forceinline__ unsigned tcd_ld_acquire_sys(const unsigned *q) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(q) : "memory");
    return v;
}
__device__ __forceinline__ void tcd_st_release_sys(unsigned *q, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(q), "r"(v) : "memory");
}

// bounded wait for a peer's flag (see tp_wait_flag in tp_ar.cu): 2 s, then the local error word is set and the kernel goes on
__device__ __forceinline__ bool tcd_wait_flag(const unsigned *f, unsigned want, unsigned *err) {
    if (tcd_ld_acquire_sys(f) == want) return true;
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (tcd_ld_acquire_sys(f) != want) {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 > 2000000000ull) {
            atomicExch(err, 1u);
            return false;
        }
    }
    return true;
}

// With TcdParams::dataflow bit 3 there is no grid barrier in front of this phase: the row-parallel phase added (partial sum, 1)
// pairs into `delta`, and the CTA that owns slice j polls the pairs of its rows until every count equals the number of CTAs
// that contribute to the row's tile (`red_N`, `red_K`: the shape of the phase that produced the sums).
static __device__ void tcd_allreduce_phase(const TcdParams &p, const TcdSmem &s, int G, int bid, int tid, int red_N, int red_K) {
    const unsigned sq = (unsigned)s.misc[62] + 1u;
    const int phase = (int)(sq & 1u);
    const int n_slices = (p.d + 1023) / 1024;
    for (int j = bid; j < n_slices; j += G) {
        const int i = j * 1024 + tid * 4;
        const bool live = tid < 256 && i < p.d;
        float4 mine = make_float4(0.f, 0.f, 0.f, 0.f);
        if (live && (p.dataflow & 8)) {
            const float want = tcd_tile_contributors(red_N, red_K, i / 128, G);   // 4 consecutive rows lie in one tile
            const float2 *dp = reinterpret_cast<const float2 *>(p.delta) + i;
            float4 a, b;
            for (;;) {
                asm volatile("ld.relaxed.gpu.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "l"(dp) : "memory");
                asm volatile("ld.relaxed.gpu.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(dp + 2) : "memory");
                if (a.y == want && a.w == want && b.y == want && b.w == want) break;
                __nanosleep(32);
            }
            mine = make_float4(a.x, a.z, b.x, b.z);
        } else if (live) {
            mine = __ldcg(reinterpret_cast<const float4 *>(p.delta + i));
        }
        if (live)
            for (int r = 0; r < p.tp_size; ++r)
                if (r != p.tp_rank)
                    *reinterpret_cast<float4 *>(p.ar_inbox[r] + ((size_t)phase * p.tp_size + p.tp_rank) * p.ar_slot_floats + i) = mine;
        __threadfence_system();
        bar_sync(1, kTC);
        if (tid < p.tp_size && tid != p.tp_rank) {
            tcd_st_release_sys(p.ar_flags[tid] + ((size_t)phase * p.tp_size + p.tp_rank) * p.ar_max_slices + j, sq);
            const unsigned *f = p.ar_flags[p.tp_rank] + ((size_t)phase * p.tp_size + tid) * p.ar_max_slices + j;
            tcd_wait_flag(f, sq, p.ar_seq + 3);
        }
        bar_sync(1, kTC);
        if (live) {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int r = 0; r < p.tp_size; ++r) {   // rank order: every rank holds the same residual stream bit for bit
                float4 v = mine;
                if (r != p.tp_rank)
                    v = __ldcv(reinterpret_cast<const float4 *>(p.ar_inbox[p.tp_rank] + ((size_t)phase * p.tp_size + r) * p.ar_slot_floats + i));
                acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
            }
            float4 hv = __ldcg(reinterpret_cast<const float4 *>(p.h + i));
            hv.x += acc.x, hv.y += acc.y, hv.z += acc.z, hv.w += acc.w;
            *reinterpret_cast<float4 *>(p.h + i) = hv;
        }
    }
    bar_sync(1, kTC);   // every thread has read the sequence number
    if (tid == 0) s.misc[62] = (int)sq;
}

// ---- tensor parallel: every rank's best (logit, token id) -> the global greedy token ------------------
// Called by all compute threads of every CTA with this rank's winner; CTA 0 pushes it to the peers, every
// CTA waits for the peers' entries in local memory and returns the best of all (lowest id on ties).
static __device__ int tcd_token_exchange(const TcdParams &p, const TcdSmem &s, float m, int gi, int bid, int tid) {
    const unsigned sq = (unsigned)s.misc[63] + 1u;
    const int phase = (int)(sq & 1u);
    const unsigned long long mine = (unsigned long long)__float_as_uint(m) | ((unsigned long long)(unsigned)gi << 32);
    if (bid == 0 && tid < p.tp_size && tid != p.tp_rank) {
        *reinterpret_cast<volatile unsigned long long *>(p.tok_inbox[tid] + (size_t)phase * p.tp_size + p.tp_rank) = mine;
        __threadfence_system();
        tcd_st_release_sys(p.tok_flags[tid] + (size_t)phase * p.tp_size + p.tp_rank, sq);
    }
    if (tid < p.tp_size) {
        unsigned long long v = mine;
        if (tid != p.tp_rank) {
            const unsigned *f = p.tok_flags[p.tp_rank] + (size_t)phase * p.tp_size + tid;
            tcd_wait_flag(f, sq, p.ar_seq + 3);
            v = *reinterpret_cast<volatile unsigned long long *>(p.tok_inbox[p.tp_rank] + (size_t)phase * p.tp_size + tid);
        }
        s.red[48 + tid] = __uint_as_float((unsigned)(v & 0xffffffffull));
        s.misc[48 + tid] = (int)(v >> 32);
    }
    bar_sync(1, kTC);
    float bm = s.red[48];
    int bi = s.misc[48];
    for (int r = 1; r < p.tp_size; ++r)
        if (s.red[48 + r] > bm || (s.red[48 + r] == bm && s.misc[48 + r] < bi)) bm = s.red[48 + r], bi = s.misc[48 + r];
    bar_sync(1, kTC);   // red/misc are reused
    if (tid == 0) s.misc[63] = (int)sq;   // read again only after the barriers of a whole step
    return bi;
}
