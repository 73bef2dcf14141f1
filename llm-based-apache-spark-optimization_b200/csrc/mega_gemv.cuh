// mega_gemv.cuh -- one weight-streaming GEMV phase of the persistent decode kernel, consumer
// side (included by decode_mega.cu).  Tensor-core formulation of y = W x at batch 1:
//   D[m16, n8] += A[m16, k16] * B[k16, n8]   (mma.sync.m16n8k16, bf16 x bf16 -> fp32)
// One instruction covers 8 weight rows x 32 k: A rows 0..7 are the 8 weight rows over k-block
// j, A rows 8..15 the SAME weight rows over k-block j+1 (one ldmatrix.x4 from the ring); the
// activation vector is split as x = hi + lo (both bf16, together ~fp32 accurate) and B's
// columns are (hi_j, lo_j, hi_j+1, lo_j+1, 0, 0, 0, 0).  D[r][0] + D[r][1] + D[8+r][2] + D[8+r][3]
// is then the dot product of weight row r over the 32 k values; the other D entries are unused.
// (mma.sync issues at about 1 per 28 clk per SM sub-partition on sm_100, measured in
// profiles/: packing 256 weights into each instruction keeps the consumer ahead of the stream.)
//
// The phase sits on the critical path between two grid barriers, so every global round trip
// is issued as early as possible: all activation (and gain) loads first, then arithmetic; the
// residual-stream values the epilogue adds to are fetched before the weight loop.
#pragma once

__device__ __forceinline__ void ldsm_x4(uint32_t saddr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(saddr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                               uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
        "{%0, %1, %2, %3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// mbarrier arrive that is not issued before `dep` (a register produced by an earlier ldmatrix) is ready
__device__ __forceinline__ void mbar_arrive_after(uint64_t *bar, uint32_t dep) {
    asm volatile("{\n\t.reg .b32 t;\n\tmov.b32 t, %1;\n\tmbarrier.arrive.shared::cta.b64 _, [%0];\n\t}" ::"r"(smem_u32(bar)),
                 "r"(dep)
                 : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    return (uint32_t)f32_to_bf16_bits(a) | ((uint32_t)f32_to_bf16_bits(b) << 16);
}
__device__ __forceinline__ float bf16_round(float a) { return __uint_as_float((uint32_t)f32_to_bf16_bits(a) << 16); }

__device__ void mega_gemv_phase(const MegaParams &p, const MegaJob &job, const MegaSmem &s, StepCtx &c, int G,
                                int bid, int tid) {
    const int warp = tid >> 5, lane = tid & 31;
    const int K = job.K;
    const int parts = k_parts(K);
    const int kc = K / parts;
    const uint32_t row_smem = (uint32_t)kc * 2 + 16;
    const RowRange rr = cta_rows(job.N, G, bid);
    const int epi = job.epi;

    // the RMSNorm gain vectors of the next two phases are cold (read once per step): start pulling
    // them into L2 now
    {
        const MegaJob *first = s.jobs, *last = s.jobs + 4 * p.n_layers;
#pragma unroll
        for (int a = 1; a <= 2; ++a) {
            const MegaJob *nj = &job + a;
            if (nj > last) nj = first + (nj - last - 1);
            if (nj->gain != nullptr && tid * 64 < nj->K)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const uint16_t *>(nj->gain) + tid * 64));
        }
    }
    // ---- 1. activation vector -> operand table (single pass, nothing kept in registers) ----
    // item = (k16 block i, t): the four values x[16i+2t], x[16i+2t+1], x[16i+2t+8], x[16i+2t+9].
    // The RMSNorm scale is applied in the epilogue (see below), so each item can be converted and
    // stored as soon as it arrives; register arrays here used to spill, and spills are expensive
    // right after a grid barrier (the acquire polls invalidate L1, so every local-memory access
    // is an L2 round trip).
    const int n_items = K / 4;
    const uint16_t *gain = reinterpret_cast<const uint16_t *>(job.gain);
    const uint16_t *emb = reinterpret_cast<const uint16_t *>(p.embed) + (size_t)c.tok * p.d;
    const int x_kind = job.x_src & 0xff;
    const float *xsrc = x_kind == MEGA_X_H ? p.h : (x_kind == MEGA_X_ATT ? p.att : p.act);
    const bool from_embed = x_kind == MEGA_X_EMBED;
    // residual stream values this thread's epilogue will add to (single-block phases only)
    float hres = 0.f;
    const bool res_pre = epi == B200LLM_EPI_RESIDUAL && rr.nrows <= kRB;
    if (res_pre && tid < rr.nrows) hres = __ldcg(p.h + rr.row0 + tid);
    float ss = 0.f;
#pragma unroll 2
    for (int it = tid; it < n_items; it += kMC) {
        const int i16 = it >> 2, t = it & 3;
        const int k = i16 * 16 + t * 2;
        float v[4];
        if (from_embed) {
            const uint32_t u0 = *reinterpret_cast<const uint32_t *>(emb + k);
            const uint32_t u1 = *reinterpret_cast<const uint32_t *>(emb + k + 8);
            v[0] = bf16lo(u0), v[1] = bf16hi(u0), v[2] = bf16lo(u1), v[3] = bf16hi(u1);
            if ((it % G) == bid) {  // residual stream h <- embedding row, each element written once
                *reinterpret_cast<float2 *>(p.h + k) = make_float2(v[0], v[1]);
                *reinterpret_cast<float2 *>(p.h + k + 8) = make_float2(v[2], v[3]);
            }
        } else {
            const float2 v0 = __ldcg(reinterpret_cast<const float2 *>(xsrc + k));
            const float2 v1 = __ldcg(reinterpret_cast<const float2 *>(xsrc + k + 8));
            v[0] = v0.x, v[1] = v0.y, v[2] = v1.x, v[3] = v1.y;
        }
        ss += v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3];
        if (gain != nullptr) {
            const uint32_t g0 = *reinterpret_cast<const uint32_t *>(gain + k);
            const uint32_t g1 = *reinterpret_cast<const uint32_t *>(gain + k + 8);
            v[0] *= bf16lo(g0), v[1] *= bf16hi(g0), v[2] *= bf16lo(g1), v[3] *= bf16hi(g1);
        }
        float hi[4], lo[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) hi[e] = bf16_round(v[e]), lo[e] = v[e] - hi[e];
        uint2 *tab = reinterpret_cast<uint2 *>(s.xtab) + i16 * 8 + t;
        tab[0] = make_uint2(pack_bf16(hi[0], hi[1]), pack_bf16(hi[2], hi[3]));  // B column 2b   (hi)
        tab[4] = make_uint2(pack_bf16(lo[0], lo[1]), pack_bf16(lo[2], lo[3]));  // B column 2b+1 (lo)
    }
    mega_trace(p, c.ti, bid, tid);  // (debug) this thread's x values have arrived and are converted
    // Fused RMSNorm prologue: the operand table holds gain * x; the scalar rsqrt(mean(x^2) + eps)
    // multiplies the result in the epilogue, so the block reduction of sum(x^2) is off the
    // critical path (it completes under the same barrier that publishes the table).
    if (gain != nullptr) {
        ss = warp_sum(ss);
        if (lane == 0) s.red[warp] = ss;
    }
    mega_trace(p, c.ti, bid, tid);  // (debug) this thread's table entries written
    bar_sync(1, kMC);
    if (tid == 0) *reinterpret_cast<volatile int *>(s.misc + 2) = ++c.seq;  // open the producer's gate for this phase
    float sc = 1.f;
    if (gain != nullptr) {
        float tot = 0.f;
#pragma unroll
        for (int w = 0; w < kMW; ++w) tot += s.red[w];
        sc = rsqrtf(tot / (float)K + p.rms_eps);
    }
    mega_trace(p, c.ti, bid, tid);  // operand table ready

    // ---- 2. the weight loop ----
    const int gran = (epi == B200LLM_EPI_SWIGLU || epi == B200LLM_EPI_QKV_ROPE) ? 2 : 1;
    const int steps32 = kc / 32;
    const uint32_t ring_s = smem_u32(s.ring);
    const uint2 *xtab = reinterpret_cast<const uint2 *>(s.xtab);
    // matrices 0..3 of ldmatrix.x4 = (k 0-7 | k 16-23 | k 8-15 | k 24-31) of the 32-wide step
    const uint32_t lane_off = (uint32_t)(lane & 7) * row_smem + (uint32_t)(((lane >> 3) & 1) * 32 + (lane >> 4) * 16);
    const int b_ofs = (lane >> 3) * 8 + ((lane >> 2) & 1) * 4 + (lane & 3);  // lanes 0..15: operand table entry

    // (debug) per-stage stamps of warp 0 / CTA 0 for the gate/up phase of layer 1
    const bool fine = c.ti >= 0 && bid == 0 && tid == 0 && (&job == s.jobs + 6);
    int fi = 0;
    for (int blk0 = 0; blk0 < rr.nrows; blk0 += kRB) {
        const int blk_rows = min(kRB, rr.nrows - blk0);
        for (int grp = 0; grp < blk_rows / 8; ++grp) {
            float acc4[4][4];  // independent accumulators: mma.sync has a long result latency on sm_100
#pragma unroll
            for (int a = 0; a < 4; ++a) acc4[a][0] = acc4[a][1] = acc4[a][2] = acc4[a][3] = 0.f;
            for (int pt = 0; pt < parts; ++pt, ++c.gs) {
                const uint32_t slot = c.gs % (uint32_t)p.ring_slots, use = c.gs / (uint32_t)p.ring_slots;
                mbar_wait(&s.full[slot], use & 1);
                if (fine) {
                    unsigned long long t;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                    if (fi < 255) p.trace[512 + 2 * fi] = t;
                }
                const uint32_t st_base = ring_s + slot * kSlotBytes + lane_off;
                const int i16_base = (pt * kc) >> 4;
                // (a) pull this warp's weight fragments out of the stage, (b) hand the slot back to
                // the producer, (c) only then spend time in the tensor pipe
                uint32_t af[kSPW][4];
                uint2 bx[kSPW];
                uint32_t dep = 0;
#pragma unroll
                for (int j = 0; j < kSPW; ++j) {  // steps32 <= 64: kSPW per warp
                    const int st = warp + j * kMW;
                    af[j][0] = af[j][1] = af[j][2] = af[j][3] = 0u;
                    bx[j] = make_uint2(0u, 0u);
                    if (st < ((p.dbg & 1) ? 0 : steps32)) {
                        ldsm_x4(st_base + (uint32_t)st * 64, af[j][0], af[j][1], af[j][2], af[j][3]);
                        if (lane < 16) bx[j] = xtab[(i16_base + 2 * st) * 8 + b_ofs];
                        dep ^= af[j][3];
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive_after(&s.empty[slot], dep);
#pragma unroll
                for (int j = 0; j < kSPW; ++j) {
                    const int st = warp + j * kMW;
                    if (st < ((p.dbg & 1) ? 0 : steps32))
                        mma_bf16_16816(acc4[j & 3], af[j][0], af[j][1], af[j][2], af[j][3], bx[j].x, bx[j].y);
                }
                if (fine) {
                    unsigned long long t;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                    if (fi < 255) p.trace[512 + 2 * fi + 1] = t;
                    ++fi;
                }
            }
            float d[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) d[e] = (acc4[0][e] + acc4[1][e]) + (acc4[2][e] + acc4[3][e]);
            // lane (g, t): d[0], d[1] = D[g][2t], D[g][2t+1]; d[2], d[3] = D[8+g][2t], D[8+g][2t+1]
            const int t4 = lane & 3;
            float v = t4 == 0 ? d[0] + d[1] : (t4 == 1 ? d[2] + d[3] : 0.f);
            v += __shfl_xor_sync(0xffffffffu, v, 1);
            if (t4 == 0) s.part[(grp * 8 + (lane >> 2)) * kMW + warp] = v;
        }
        bar_sync(1, kMC);
        mega_trace(p, c.ti, bid, tid);  // ring consumed
        // ---- 3. combine the 16 warp partials of every row of the block + fused epilogue ----
        const int blk_units = blk_rows / gran;
        for (int u = tid; u < blk_units; u += kMC) {
            const int rl = u * gran;
            const int n = rr.row0 + blk0 + rl;
            float v0 = 0.f, v1 = 0.f;
#pragma unroll
            for (int w = 0; w < kMW; ++w) v0 += s.part[rl * kMW + w];
            if (gran == 2) {
#pragma unroll
                for (int w = 0; w < kMW; ++w) v1 += s.part[(rl + 1) * kMW + w];
            }
            v0 *= sc, v1 *= sc;
            if (epi == B200LLM_EPI_RESIDUAL) {
                p.h[n] = (res_pre ? hres : __ldcg(p.h + n)) + v0;
            } else if (epi == B200LLM_EPI_SWIGLU) {
                p.act[n >> 1] = (v0 / (1.f + expf(-v0))) * v1;
            } else if (epi == B200LLM_EPI_LOGITS) {
                p.logits[n] = v0;
                if (v0 > c.best) c.best = v0, c.best_i = n;
            } else {  // QKV_ROPE
                const int hd = p.head_dim;
                const int qrows = p.n_q * hd, krows = p.n_kv * hd;
                if (n < qrows + krows) {
                    const float cs = s.rope[(n % hd) & ~1], sn = s.rope[((n % hd) & ~1) + 1];  // cos/sin of this step
                    const float o0 = v0 * cs - v1 * sn, o1 = v1 * cs + v0 * sn;
                    v0 = o0, v1 = o1;
                }
                if (n < qrows) {
                    *reinterpret_cast<float2 *>(p.q + n) = make_float2(v0, v1);
                } else {
                    const int pgi = c.pos / kMPT;
                    const int page = pgi < kMaxPtab ? s.ptab[pgi] : p.page_table[pgi];
                    const bool is_k = n < qrows + krows;
                    const int kn = is_k ? n - qrows : n - qrows - krows;
                    uint16_t *dst = reinterpret_cast<uint16_t *>(is_k ? job.k_pages : job.v_pages) +
                                    (((size_t)page * p.kv_page_heads + kn / hd) * kMPT + (c.pos % kMPT)) * hd + kn % hd;
                    *reinterpret_cast<uint32_t *>(dst) = pack_bf16(v0, v1);
                }
            }
        }
        if (blk0 + kRB < rr.nrows) bar_sync(1, kMC);  // partials are overwritten by the next block
    }
}
