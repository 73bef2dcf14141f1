// decode_mega.cu -- the batch-1 decode step as ONE persistent kernel (one CTA per SM).
//
// At batch 1 a Llama-3.2-3B step is ~140 GEMV/attention phases of 3-15 us each; launched as
// separate kernels, ramp-up and drain cost more than the streaming itself (profiles/: 2.3 ms
// per step against a 0.98 ms HBM floor).  Here all phases of all steps of a decode call run
// inside one cooperative grid:
//   * a producer warp per CTA walks the whole weight schedule (layer after layer, step after
//     step) and streams this CTA's rows of every matrix through a 160 KB shared-memory ring
//     with 1-D TMA bulk copies; it never waits for anything but ring space, so HBM stays busy
//     across phase boundaries;
//   * 16 compute warps consume the ring (bf16 x fp32 FMA, activations in registers), combine
//     per-warp partials in shared memory and run the fused epilogues (RMSNorm prologue, RoPE +
//     paged-KV append, SwiGLU, residual, logits + argmax);
//   * phases are separated by a grid barrier (one atomic + acquire poll per CTA) on the
//     consumer side only; attention (split-T over the lane's KV pages, K/V tiles staged by TMA)
//     runs between the qkv and o_proj phases while the ring fills with o_proj's weights;
//   * greedy sampling and token feedback happen in-kernel, so n_steps tokens need one launch.
#include <cooperative_groups.h>

#include "common.cuh"
#include "kernels.h"
#include "mega.h"

namespace b200 {

constexpr int kMC = 512;                 // compute threads
constexpr int kMW = kMC / 32;            // compute warps
constexpr int kSlotBytes = 32 * 1024;
constexpr int kSlots = 5;
constexpr int kAttW = 4;                 // warps that stream KV pages in the attention phase
constexpr int kPartBytes = 17 * 1024;    // per-warp GEMV partials / attention merge buffer
constexpr int kRB = 128;                 // rows combined per partial block
constexpr int kMPT = 16;                 // tokens per KV page

struct MegaSmem {
    uint8_t *ring, *att_tiles;
    float *part;
    uint64_t *full, *empty, *att_full;
    float *red;
    int *misc;
};

__device__ __forceinline__ MegaSmem carve(uint8_t *base, int hd) {
    MegaSmem s;
    s.ring = base;
    s.att_tiles = base + kSlots * kSlotBytes;
    const int att_bytes = kAttW * 2 * kMPT * hd * 2;
    s.part = reinterpret_cast<float *>(s.att_tiles + att_bytes);
    s.full = reinterpret_cast<uint64_t *>(reinterpret_cast<uint8_t *>(s.part) + kPartBytes);
    s.empty = s.full + kSlots;
    s.att_full = s.empty + kSlots;
    s.red = reinterpret_cast<float *>(s.att_full + kAttW);
    s.misc = reinterpret_cast<int *>(s.red + 64);
    return s;
}

size_t mega_smem_bytes(int head_dim) {
    return (size_t)kSlots * kSlotBytes + (size_t)kAttW * 2 * kMPT * head_dim * 2 + kPartBytes +
           (2 * kSlots + kAttW) * 8 + 64 * 4 + 256;
}

bool mega_supported(int d, int ffn, int n_q, int n_kv, int head_dim) {
    auto ok_k = [](int K) { return K % 8 == 0 && K / 8 <= 3 * kMC && (size_t)K * 2 <= kSlotBytes; };
    const int G = n_kv > 0 ? n_q / n_kv : 0;
    return ok_k(d) && ok_k(ffn) && ok_k(n_q * head_dim) && (head_dim == 64 || head_dim == 128) && G >= 1 && G <= 4;
}

__device__ __forceinline__ int rows_per_stage(int K) {
    const int upt = (K / 8 + kMC - 1) / kMC;
    return upt == 1 ? 4 : (upt == 2 ? 2 : 1);
}

struct RowRange {
    int row0, nrows, units;
};
__device__ __forceinline__ RowRange cta_rows(int N, int gran, int G, int bid) {
    const int n_units = N / gran;
    const int base = n_units / G, rem = n_units % G;
    RowRange r;
    r.units = base + (bid < rem ? 1 : 0);
    r.row0 = (bid * base + (bid < rem ? bid : rem)) * gran;
    r.nrows = r.units * gran;
    return r;
}
__device__ __forceinline__ int job_gran(int epi) {
    return (epi == B200LLM_EPI_SWIGLU || epi == B200LLM_EPI_QKV_ROPE) ? 2 : 1;
}

// Grid barrier for the compute threads (the producer warp never takes part).
__device__ __forceinline__ void grid_sync(unsigned *ctr, unsigned target, int tid) {
    asm volatile("fence.proxy.async;" ::: "memory");  // KV rows written here are read by TMA elsewhere
    bar_sync(1, kMC);
    if (tid == 0) {
        __threadfence();
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
        unsigned v;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
        } while (v < target);
    }
    bar_sync(1, kMC);
}

// ---------------------------------------------------------------- producer: the weight stream
__device__ void mega_producer(const MegaParams &p, const MegaSmem &s, int G, int bid) {
    const uint64_t pol = l2_policy_evict_first();
    const int n_jobs = 4 * p.n_layers + 1;
    uint32_t gs = 0;
    for (int step = 0; step < p.n_steps; ++step) {
        for (int j = 0; j < n_jobs; ++j) {
            const MegaJob job = p.jobs[j];
            const RowRange rr = cta_rows(job.N, job_gran(job.epi), G, bid);
            const int rg = rows_per_stage(job.K);
            const size_t row_bytes = (size_t)job.K * 2;
            const uint8_t *src = reinterpret_cast<const uint8_t *>(job.w) + (size_t)rr.row0 * row_bytes;
            const int n_groups = (rr.nrows + rg - 1) / rg;
            for (int g = 0; g < n_groups; ++g, ++gs) {
                const uint32_t slot = gs % kSlots, use = gs / kSlots;
                if (use > 0) mbar_wait(&s.empty[slot], (use - 1) & 1);
                const int nr = min(rg, rr.nrows - g * rg);
                const uint32_t bytes = (uint32_t)(nr * row_bytes);
                mbar_expect_tx(&s.full[slot], bytes);
                tma_load_1d_hint(s.ring + (size_t)slot * kSlotBytes, src + (size_t)g * rg * row_bytes, bytes,
                                 &s.full[slot], pol);
            }
        }
    }
}

struct StepCtx;
__device__ __forceinline__ void mega_trace(const MegaParams &p, int &ti, int bid, int tid) {
    if (ti >= 0 && tid == 0 && ti < 1024) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.trace[(bid ? 1024 : 0) + ti++] = t;
    }
}

// ---------------------------------------------------------------- consumer: one GEMV phase
struct StepCtx {
    int tok, pos;
    int ti;               // trace cursor (-1: tracing off for this step)
    uint32_t gs;          // ring stage counter, in lock-step with the producer
    float best;           // LOGITS: running (max, argmax) of this thread
    int best_i;
};

__device__ void mega_gemv_phase(const MegaParams &p, const MegaJob &job, const MegaSmem &s, StepCtx &c, int G,
                                int bid, int tid) {
    const int warp = tid >> 5, lane = tid & 31;
    const int K = job.K;
    const int upt = (K / 8 + kMC - 1) / kMC;
    const int rg = upt == 1 ? 4 : (upt == 2 ? 2 : 1);
    const size_t row_bytes = (size_t)K * 2;

    // ---- activation slice of this thread -> registers (fp32) ----
    float xr[3][8];
    bool valid[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        const int k0 = (j * kMC + tid) * 8;
        valid[j] = j < upt && k0 < K;
        if (valid[j]) {
            if (job.x_src == MEGA_X_EMBED) {
                const uint16_t *e = reinterpret_cast<const uint16_t *>(p.embed) + (size_t)c.tok * p.d + k0;
                const uint4 v = *reinterpret_cast<const uint4 *>(e);
                xr[j][0] = bf16lo(v.x), xr[j][1] = bf16hi(v.x), xr[j][2] = bf16lo(v.y), xr[j][3] = bf16hi(v.y);
                xr[j][4] = bf16lo(v.z), xr[j][5] = bf16hi(v.z), xr[j][6] = bf16lo(v.w), xr[j][7] = bf16hi(v.w);
                if (((j * kMC + tid) % G) == bid) {  // residual stream h <- embedding row, each unit written once
                    float4 *hd4 = reinterpret_cast<float4 *>(p.h + k0);
                    hd4[0] = make_float4(xr[j][0], xr[j][1], xr[j][2], xr[j][3]);
                    hd4[1] = make_float4(xr[j][4], xr[j][5], xr[j][6], xr[j][7]);
                }
            } else {
                const float *x = job.x_src == MEGA_X_H ? p.h : (job.x_src == MEGA_X_ATT ? p.att : p.act);
                const float4 v0 = __ldcg(reinterpret_cast<const float4 *>(x + k0));
                const float4 v1 = __ldcg(reinterpret_cast<const float4 *>(x + k0 + 4));
                xr[j][0] = v0.x, xr[j][1] = v0.y, xr[j][2] = v0.z, xr[j][3] = v0.w;
                xr[j][4] = v1.x, xr[j][5] = v1.y, xr[j][6] = v1.z, xr[j][7] = v1.w;
            }
        } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) xr[j][e] = 0.f;
        }
    }
    if (job.gain != nullptr) {  // fused RMSNorm prologue
        float ss = 0.f;
#pragma unroll
        for (int j = 0; j < 3; ++j)
#pragma unroll
            for (int e = 0; e < 8; ++e) ss += xr[j][e] * xr[j][e];
        ss = warp_sum(ss);
        if (lane == 0) s.red[warp] = ss;
        bar_sync(1, kMC);
        float tot = 0.f;
#pragma unroll
        for (int w = 0; w < kMW; ++w) tot += s.red[w];
        const float sc = rsqrtf(tot / (float)K + p.rms_eps);
        const uint16_t *gain = reinterpret_cast<const uint16_t *>(job.gain);
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            if (valid[j]) {
                const uint4 gv = *reinterpret_cast<const uint4 *>(gain + (j * kMC + tid) * 8);
                const uint32_t gw[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    xr[j][2 * e] = bf16lo(gw[e]) * (xr[j][2 * e] * sc);
                    xr[j][2 * e + 1] = bf16hi(gw[e]) * (xr[j][2 * e + 1] * sc);
                }
            }
        }
    }

    const int gran = job_gran(job.epi);
    const RowRange rr = cta_rows(job.N, gran, G, bid);
    const int epi = job.epi;
    mega_trace(p, c.ti, bid, tid);  // x loaded (+ normalised)

    for (int blk0 = 0; blk0 < rr.nrows; blk0 += kRB) {
        const int blk_rows = min(kRB, rr.nrows - blk0);
        const int blk_groups = (blk_rows + rg - 1) / rg;
        for (int g = 0; g < blk_groups; ++g, ++c.gs) {
            const uint32_t slot = c.gs % kSlots, use = c.gs / kSlots;
            mbar_wait(&s.full[slot], use & 1);
            const uint8_t *st = s.ring + (size_t)slot * kSlotBytes;
            const int nr = min(rg, blk_rows - g * rg);
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                if (r < nr) {
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        if (valid[j]) {
                            const uint4 wv = *reinterpret_cast<const uint4 *>(st + (size_t)r * row_bytes +
                                                                              (size_t)(j * kMC + tid) * 16);
                            const uint32_t ww[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                acc[r] = fmaf(bf16lo(ww[e]), xr[j][2 * e], acc[r]);
                                acc[r] = fmaf(bf16hi(ww[e]), xr[j][2 * e + 1], acc[r]);
                            }
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&s.empty[slot]);
#pragma unroll
            for (int r = 0; r < 4; ++r) acc[r] = warp_sum(acc[r]);
            if (lane == 0) {
#pragma unroll
                for (int r = 0; r < 4; ++r)
                    if (r < nr) s.part[(g * rg + r) * kMW + warp] = acc[r];
            }
        }
        bar_sync(1, kMC);
        mega_trace(p, c.ti, bid, tid);  // ring consumed
        // ---- combine the 16 warp partials of every row of the block + fused epilogue ----
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
            if (epi == B200LLM_EPI_RESIDUAL) {
                p.h[n] = __ldcg(p.h + n) + v0;
            } else if (epi == B200LLM_EPI_SWIGLU) {
                p.act[n >> 1] = (v0 / (1.f + expf(-v0))) * v1;
            } else if (epi == B200LLM_EPI_LOGITS) {
                p.logits[n] = v0;
                if (v0 > c.best) c.best = v0, c.best_i = n;
            } else {  // QKV_ROPE
                const int hd = p.head_dim;
                const int qrows = p.n_q * hd, krows = p.n_kv * hd;
                if (n < qrows + krows) {
                    const float ang = (float)c.pos * p.inv_freq[(n % hd) >> 1];
                    float sn, cs;
                    sincosf(ang, &sn, &cs);
                    const float o0 = v0 * cs - v1 * sn, o1 = v1 * cs + v0 * sn;
                    v0 = o0, v1 = o1;
                }
                if (n < qrows) {
                    *reinterpret_cast<float2 *>(p.q + n) = make_float2(v0, v1);
                } else {
                    const int page = p.page_table[c.pos / kMPT];
                    const bool is_k = n < qrows + krows;
                    const int kn = is_k ? n - qrows : n - qrows - krows;
                    uint16_t *dst = reinterpret_cast<uint16_t *>(is_k ? job.k_pages : job.v_pages) +
                                    (((size_t)page * p.n_kv + kn / hd) * kMPT + (c.pos % kMPT)) * hd + kn % hd;
                    *reinterpret_cast<uint32_t *>(dst) =
                        (uint32_t)f32_to_bf16_bits(v0) | ((uint32_t)f32_to_bf16_bits(v1) << 16);
                }
            }
        }
        if (blk0 + kRB < rr.nrows) bar_sync(1, kMC);  // partials are overwritten by the next block
    }
}

#include "mega_attn.cuh"
#include "mega_kernel.cuh"

int mega_decode_launch(const MegaParams &p, cudaStream_t stream) {
    B200_REQUIRE(mega_supported(p.d, p.ffn, p.n_q, p.n_kv, p.head_dim), "mega: unsupported geometry");
    B200_REQUIRE(p.n_steps >= 1 && (p.sample || p.n_steps == 1), "mega: n_steps must be 1 when sampling is external");
    B200_REQUIRE(p.n_splits >= 1, "mega: n_splits < 1");
    const int GQ = p.n_q / p.n_kv;
#define B200_MEGA_CASE(HD_, G_) \
    if (p.head_dim == HD_ && GQ == G_) return mega_launch_t<HD_, G_>(p, stream);
    B200_MEGA_CASE(128, 1) B200_MEGA_CASE(128, 2) B200_MEGA_CASE(128, 3) B200_MEGA_CASE(128, 4)
    B200_MEGA_CASE(64, 1) B200_MEGA_CASE(64, 2) B200_MEGA_CASE(64, 4)
#undef B200_MEGA_CASE
    set_error("mega: unsupported head_dim %d / group %d", p.head_dim, GQ);
    return B200LLM_EINVAL;
}

}  // namespace b200
