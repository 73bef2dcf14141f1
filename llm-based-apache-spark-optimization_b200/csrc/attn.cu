// attn.cu -- paged-KV GQA decode attention (kernel K4 of SURVEY.md section 2.3).
//
// grid = (n_splits, n_kv, rows).  One CTA = one KV head of one query row over a slice
// of that row's KV pages (split-T / flash-decoding).  Each of the 4 warps streams its
// own pages: K and V tiles of one page ([16 tokens, head_dim] bf16, contiguous in the
// pool) are staged into a private 2-deep shared-memory ring by 1-D TMA bulk copies
// (cp.async.bulk + mbarrier).  Inside a warp, 8 lanes share a token (each lane owns
// head_dim/8 dims), so 4 tokens are scored per step; the q.k reduction is 3 warp
// shuffles, the softmax is an online (running max / sum) fp32 stream per 8-lane group.
// Streams are merged by shuffles (warp), shared memory (CTA) and -- across splits --
// through a global scratch buffer by the last CTA to arrive (atomic ticket), in fixed
// split order, so results are deterministic.
#include <mutex>

#include "common.cuh"
#include "kernels.h"

namespace b200 {

constexpr int kAttnWarps = 4;
constexpr int kAttnStages = 2;
constexpr int kPT = 16;  // tokens per page

struct AttnDev {
    b200llm_attn_args a;
    float *part_o;     // [rows, n_q, n_splits, HD]
    float *part_ml;    // [rows, n_q, n_splits, 2]
    unsigned *tickets; // [rows, n_kv]
    float scale_log2;  // 1/sqrt(hd) * log2(e)
    uint16_t *out_bf16; // when set, the output goes here as bf16 (the next GEMM's operand) instead of a.out
    int64_t out_lo_off; // != 0: the bf16 lo parts (value - hi) go to out_bf16 + out_lo_off
};

__device__ __forceinline__ void merge_state(float &m, float &l, float m2, float l2, float &c1, float &c2) {
    const float mn = fmaxf(m, m2);
    if (mn == -INFINITY) {
        c1 = 0.f, c2 = 0.f;
    } else {
        c1 = exp2f(m - mn), c2 = exp2f(m2 - mn);
    }
    l = l * c1 + l2 * c2;
    m = mn;
}

template <int HD, int G>
__global__ void __launch_bounds__(kAttnWarps * 32) attn_decode_kernel(const __grid_constant__ AttnDev p) {
    constexpr int EL = HD / 16;                  // elements per lane segment (two segments per lane)
    constexpr int TILE = kPT * HD * 2;           // bytes of one K (or V) page tile
    extern __shared__ __align__(128) uint8_t smem[];
    const b200llm_attn_args &a = p.a;
    const int split = blockIdx.x, hk = blockIdx.y, row = blockIdx.z;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int tg = lane >> 3, p8 = lane & 7;

    uint8_t *tiles = smem + (size_t)warp * kAttnStages * 2 * TILE;
    float *mrg = reinterpret_cast<float *>(smem + (size_t)kAttnWarps * kAttnStages * 2 * TILE);  // [warps][G][HD+2]
    uint64_t *bars = reinterpret_cast<uint64_t *>(mrg + kAttnWarps * G * (HD + 2));
    uint64_t *full = bars + warp * kAttnStages;
    __shared__ int s_is_last;

    if (lane == 0) {
        for (int s = 0; s < kAttnStages; ++s) mbar_init(&full[s], 1);
        mbar_fence_init();
    }
    __syncwarp();
    pdl_trigger();
    pdl_wait();  // q and the newly appended K/V come from the previous kernel

    const int pos = a.row_pos[row];
    const int ln = a.row_lane[row];
    const int n_pages = pos / kPT + 1;  // tokens 0..pos inclusive
    const int n_splits = gridDim.x;
    const int pps = (n_pages + n_splits - 1) / n_splits;
    const int pg_begin = split * pps;
    const int pg_end = min(n_pages, pg_begin + pps);
    const int32_t *ptab = a.page_table + (size_t)ln * a.max_pages_per_lane;
    const uint8_t *kbase = reinterpret_cast<const uint8_t *>(a.k_pages);
    const uint8_t *vbase = reinterpret_cast<const uint8_t *>(a.v_pages);

    // q for the G heads of this KV head, pre-scaled into the log2 domain
    float qa[G][EL], qb[G][EL];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const float *qp = a.q + (size_t)row * a.n_q * HD + (size_t)(hk * G + g) * HD;
#pragma unroll
        for (int e = 0; e < EL; ++e) {
            qa[g][e] = qp[p8 * EL + e] * p.scale_log2;
            qb[g][e] = qp[HD / 2 + p8 * EL + e] * p.scale_log2;
        }
    }
    float m[G], l[G], oa[G][EL], ob[G][EL];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        m[g] = -INFINITY, l[g] = 0.f;
#pragma unroll
        for (int e = 0; e < EL; ++e) oa[g][e] = 0.f, ob[g][e] = 0.f;
    }

    // ---- per-warp page stream ----
    const int my_first = pg_begin + warp;
    const int n_mine = my_first < pg_end ? (pg_end - my_first + kAttnWarps - 1) / kAttnWarps : 0;
    auto issue = [&](int i) {
        const int st = i % kAttnStages;
        const int page = ptab[my_first + i * kAttnWarps];
        const size_t off = ((size_t)page * (a.kv_page_heads > 0 ? a.kv_page_heads : a.n_kv) + hk) * (size_t)TILE;
        mbar_expect_tx(&full[st], 2 * TILE);
        tma_load_1d(tiles + (size_t)st * 2 * TILE, kbase + off, TILE, &full[st]);
        tma_load_1d(tiles + (size_t)st * 2 * TILE + TILE, vbase + off, TILE, &full[st]);
    };
    if (lane == 0) {
        for (int i = 0; i < min(n_mine, kAttnStages); ++i) issue(i);
    }
    for (int i = 0; i < n_mine; ++i) {
        const int st = i % kAttnStages;
        mbar_wait(&full[st], (i / kAttnStages) & 1);
        const uint8_t *ks = tiles + (size_t)st * 2 * TILE;
        const uint8_t *vs = ks + TILE;
        const int t0 = (my_first + i * kAttnWarps) * kPT;
        // Page-at-a-time softmax: all 16 scores first (independent work, the shuffles pipeline),
        // one running-max update and one rescale per page, then the P.V accumulation.  (A
        // token-at-a-time online softmax serialises ~50 dependent instructions per token.)
        float sc[kPT / 4][G];
#pragma unroll
        for (int it = 0; it < kPT / 4; ++it) {
            const int t = it * 4 + tg;
            float ka[EL], kb[EL];
            if constexpr (EL == 8) {
                const uint4 A = *reinterpret_cast<const uint4 *>(ks + (size_t)t * HD * 2 + p8 * 16);
                const uint4 B = *reinterpret_cast<const uint4 *>(ks + (size_t)t * HD * 2 + HD + p8 * 16);
                const uint32_t aw[4] = {A.x, A.y, A.z, A.w}, bw[4] = {B.x, B.y, B.z, B.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    ka[2 * e] = bf16lo(aw[e]), ka[2 * e + 1] = bf16hi(aw[e]);
                    kb[2 * e] = bf16lo(bw[e]), kb[2 * e + 1] = bf16hi(bw[e]);
                }
            } else {
                const uint2 A = *reinterpret_cast<const uint2 *>(ks + (size_t)t * HD * 2 + p8 * 8);
                const uint2 B = *reinterpret_cast<const uint2 *>(ks + (size_t)t * HD * 2 + HD + p8 * 8);
                ka[0] = bf16lo(A.x), ka[1] = bf16hi(A.x), ka[2] = bf16lo(A.y), ka[3] = bf16hi(A.y);
                kb[0] = bf16lo(B.x), kb[1] = bf16hi(B.x), kb[2] = bf16lo(B.y), kb[3] = bf16hi(B.y);
            }
            const bool ok = (t0 + t) <= pos;
#pragma unroll
            for (int g = 0; g < G; ++g) {
                float s = 0.f;
#pragma unroll
                for (int e = 0; e < EL; ++e) s = fmaf(qa[g][e], ka[e], fmaf(qb[g][e], kb[e], s));
                s += __shfl_xor_sync(0xffffffffu, s, 1);
                s += __shfl_xor_sync(0xffffffffu, s, 2);
                s += __shfl_xor_sync(0xffffffffu, s, 4);
                sc[it][g] = ok ? s : -INFINITY;
            }
        }
        float corr[G];
#pragma unroll
        for (int g = 0; g < G; ++g) {
            float mn = m[g];
#pragma unroll
            for (int it = 0; it < kPT / 4; ++it) mn = fmaxf(mn, sc[it][g]);
            corr[g] = mn == -INFINITY ? 1.f : exp2f(m[g] - mn);
            float ps = 0.f;
#pragma unroll
            for (int it = 0; it < kPT / 4; ++it) {
                sc[it][g] = mn == -INFINITY ? 0.f : exp2f(sc[it][g] - mn);  // exp2(-inf) = 0 masks the tail
                ps += sc[it][g];
            }
            l[g] = l[g] * corr[g] + ps;
            m[g] = mn;
#pragma unroll
            for (int e = 0; e < EL; ++e) oa[g][e] *= corr[g], ob[g][e] *= corr[g];
        }
#pragma unroll
        for (int it = 0; it < kPT / 4; ++it) {
            const int t = it * 4 + tg;
            float va[EL], vb[EL];
            if constexpr (EL == 8) {
                const uint4 C = *reinterpret_cast<const uint4 *>(vs + (size_t)t * HD * 2 + p8 * 16);
                const uint4 D = *reinterpret_cast<const uint4 *>(vs + (size_t)t * HD * 2 + HD + p8 * 16);
                const uint32_t cw[4] = {C.x, C.y, C.z, C.w}, dw[4] = {D.x, D.y, D.z, D.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    va[2 * e] = bf16lo(cw[e]), va[2 * e + 1] = bf16hi(cw[e]);
                    vb[2 * e] = bf16lo(dw[e]), vb[2 * e + 1] = bf16hi(dw[e]);
                }
            } else {
                const uint2 C = *reinterpret_cast<const uint2 *>(vs + (size_t)t * HD * 2 + p8 * 8);
                const uint2 D = *reinterpret_cast<const uint2 *>(vs + (size_t)t * HD * 2 + HD + p8 * 8);
                va[0] = bf16lo(C.x), va[1] = bf16hi(C.x), va[2] = bf16lo(C.y), va[3] = bf16hi(C.y);
                vb[0] = bf16lo(D.x), vb[1] = bf16hi(D.x), vb[2] = bf16lo(D.y), vb[3] = bf16hi(D.y);
            }
#pragma unroll
            for (int g = 0; g < G; ++g) {
#pragma unroll
                for (int e = 0; e < EL; ++e) {
                    oa[g][e] = fmaf(sc[it][g], va[e], oa[g][e]);
                    ob[g][e] = fmaf(sc[it][g], vb[e], ob[g][e]);
                }
            }
        }
        __syncwarp();
        if (lane == 0 && i + kAttnStages < n_mine) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(i + kAttnStages);
        }
    }

    // ---- merge the 4 token-group streams of the warp ----
#pragma unroll
    for (int off = 8; off <= 16; off <<= 1) {
#pragma unroll
        for (int g = 0; g < G; ++g) {
            const float m2 = __shfl_xor_sync(0xffffffffu, m[g], off);
            const float l2 = __shfl_xor_sync(0xffffffffu, l[g], off);
            float c1, c2;
            merge_state(m[g], l[g], m2, l2, c1, c2);
#pragma unroll
            for (int e = 0; e < EL; ++e) {
                const float o2a = __shfl_xor_sync(0xffffffffu, oa[g][e], off);
                const float o2b = __shfl_xor_sync(0xffffffffu, ob[g][e], off);
                oa[g][e] = oa[g][e] * c1 + o2a * c2;
                ob[g][e] = ob[g][e] * c1 + o2b * c2;
            }
        }
    }
    // ---- merge the warps through shared memory ----
    if (tg == 0) {
#pragma unroll
        for (int g = 0; g < G; ++g) {
            float *d = mrg + ((size_t)warp * G + g) * (HD + 2);
            if (p8 == 0) d[HD] = m[g], d[HD + 1] = l[g];
#pragma unroll
            for (int e = 0; e < EL; ++e) d[p8 * EL + e] = oa[g][e], d[HD / 2 + p8 * EL + e] = ob[g][e];
        }
    }
    __syncthreads();
    const bool single = n_splits == 1;
    for (int idx = tid; idx < G * HD; idx += kAttnWarps * 32) {
        const int g = idx / HD, dd = idx % HD;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < kAttnWarps; ++w) M = fmaxf(M, mrg[((size_t)w * G + g) * (HD + 2) + HD]);
        float L = 0.f, O = 0.f;
        if (M != -INFINITY) {
#pragma unroll
            for (int w = 0; w < kAttnWarps; ++w) {
                const float *s = mrg + ((size_t)w * G + g) * (HD + 2);
                const float c = exp2f(s[HD] - M);
                L += s[HD + 1] * c;
                O += s[dd] * c;
            }
        }
        const int hq = hk * G + g;
        if (single) {
            if (p.out_bf16 != nullptr) {
                const float y = O / L;
                const uint16_t yh = f32_to_bf16_bits(y);
                p.out_bf16[(size_t)row * a.n_q * HD + (size_t)hq * HD + dd] = yh;
                if (p.out_lo_off != 0)
                    p.out_bf16[p.out_lo_off + (size_t)row * a.n_q * HD + (size_t)hq * HD + dd] =
                        f32_to_bf16_bits(y - __uint_as_float((uint32_t)yh << 16));
            } else
                a.out[(size_t)row * a.n_q * HD + (size_t)hq * HD + dd] = O / L;
        } else {
            const size_t pi = ((size_t)row * a.n_q + hq) * n_splits + split;
            p.part_o[pi * HD + dd] = O;
            if (dd == 0) p.part_ml[pi * 2] = M, p.part_ml[pi * 2 + 1] = L;
        }
    }
    if (single) return;

    // ---- cross-split merge by the last CTA to arrive ----
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned prev = atomicAdd(&p.tickets[(size_t)row * a.n_kv + hk], 1u);
        s_is_last = (prev == (unsigned)(n_splits - 1));
    }
    __syncthreads();
    if (!s_is_last) return;
    __threadfence();
    for (int idx = tid; idx < G * HD; idx += kAttnWarps * 32) {
        const int g = idx / HD, dd = idx % HD;
        const int hq = hk * G + g;
        const size_t pb = ((size_t)row * a.n_q + hq) * n_splits;
        float M = -INFINITY;
        for (int s = 0; s < n_splits; ++s) M = fmaxf(M, __ldcg(&p.part_ml[(pb + s) * 2]));
        float L = 0.f, O = 0.f;
        for (int s = 0; s < n_splits; ++s) {
            const float ms = __ldcg(&p.part_ml[(pb + s) * 2]);
            if (ms == -INFINITY) continue;
            const float c = exp2f(ms - M);
            L += __ldcg(&p.part_ml[(pb + s) * 2 + 1]) * c;
            O += __ldcg(&p.part_o[(pb + s) * HD + dd]) * c;
        }
        if (p.out_bf16 != nullptr) {
            const float y = O / L;
            const uint16_t yh = f32_to_bf16_bits(y);
            p.out_bf16[(size_t)row * a.n_q * HD + (size_t)hq * HD + dd] = yh;
            if (p.out_lo_off != 0)
                p.out_bf16[p.out_lo_off + (size_t)row * a.n_q * HD + (size_t)hq * HD + dd] =
                    f32_to_bf16_bits(y - __uint_as_float((uint32_t)yh << 16));
        } else
            a.out[(size_t)row * a.n_q * HD + (size_t)hq * HD + dd] = O / L;
    }
    if (tid == 0) p.tickets[(size_t)row * a.n_kv + hk] = 0;  // leave the ticket at zero for the next launch
}

size_t attn_scratch_bytes(int bs, int n_q, int n_kv, int head_dim, int n_splits) {
    size_t po = (size_t)bs * n_q * n_splits * head_dim * 4;
    size_t pml = (size_t)bs * n_q * n_splits * 2 * 4;
    return kAttnTicketBytes + po + pml;
}

template <int HD, int G>
static int launch_attn(const AttnDev &d, cudaStream_t stream, bool pdl) {
    constexpr size_t smem = (size_t)kAttnWarps * kAttnStages * 2 * kPT * HD * 2 +
                            (size_t)kAttnWarps * G * (HD + 2) * 4 + kAttnWarps * kAttnStages * 8;
    // per launch: the attribute belongs to the (function, device) pair, and callers may use several devices and threads
    B200_CUDA_OK(cudaFuncSetAttribute(attn_decode_kernel<HD, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    LaunchCfg lc(dim3(d.a.n_splits, d.a.n_kv, d.a.bs), dim3(kAttnWarps * 32), smem, stream, pdl);
    B200_CUDA_OK(cudaLaunchKernelEx(&lc.cfg, attn_decode_kernel<HD, G>, d));
    return B200LLM_OK;
}

int attn_decode_launch(const b200llm_attn_args &a, cudaStream_t stream, bool pdl, void *out_bf16, int64_t out_lo_off) {
    B200_REQUIRE(a.page_tokens == kPT, "attn: page_tokens must be %d (got %d)", kPT, a.page_tokens);
    B200_REQUIRE(a.n_kv > 0 && a.n_q % a.n_kv == 0, "attn: n_q %d not a multiple of n_kv %d", a.n_q, a.n_kv);
    B200_REQUIRE(a.bs >= 1 && a.bs <= 65535 && a.n_splits >= 1 && a.n_splits <= 64, "attn: bad bs/n_splits");
    const size_t need = attn_scratch_bytes(a.bs, a.n_q, a.n_kv, a.head_dim, a.n_splits);
    if (a.n_splits > 1 && (a.scratch == nullptr || a.scratch_bytes < need)) {
        set_error("attn: scratch too small (%zu < %zu)", a.scratch_bytes, need);
        return B200LLM_ENOMEM;
    }
    AttnDev d{};
    d.a = a;
    d.out_bf16 = reinterpret_cast<uint16_t *>(out_bf16);
    d.out_lo_off = out_lo_off;
    d.scale_log2 = (float)(1.4426950408889634 / sqrt((double)a.head_dim));
    uint8_t *s = reinterpret_cast<uint8_t *>(a.scratch);
    B200_REQUIRE((size_t)a.bs * a.n_kv * 4 <= kAttnTicketBytes, "attn: bs * n_kv exceeds %zu tickets", kAttnTicketBytes / 4);
    d.tickets = reinterpret_cast<unsigned *>(s);
    d.part_o = reinterpret_cast<float *>(s + kAttnTicketBytes);
    d.part_ml = d.part_o + (size_t)a.bs * a.n_q * a.n_splits * a.head_dim;
    const int G = a.n_q / a.n_kv;
#define B200_ATTN_CASE(HD_, G_) \
    if (a.head_dim == HD_ && G == G_) return launch_attn<HD_, G_>(d, stream, pdl);
    B200_ATTN_CASE(128, 1) B200_ATTN_CASE(128, 2) B200_ATTN_CASE(128, 3) B200_ATTN_CASE(128, 4)
    B200_ATTN_CASE(64, 1) B200_ATTN_CASE(64, 2) B200_ATTN_CASE(64, 4)
#undef B200_ATTN_CASE
    set_error("attn: unsupported head_dim %d / group %d", a.head_dim, G);
    return B200LLM_EINVAL;
}


// ---------------------------------------------------------------- prefill chunks at short contexts
//
// attn_rows_kernel: the query rows of one prefill chunk (consecutive positions of ONE lane), four rows per CTA.
// The decode kernel above gives every row its own CTA and streams the row's whole prefix from L2 for it: a chunk of
// 160 rows re-reads the same K/V tiles 160 times (~200 MB of L2 reads per layer at 300 keys, which is what bounds it).
// Here one CTA = one KV head x 4 consecutive rows; every page's K and V tiles are fetched ONCE per CTA into a 3-deep
// shared ring (1-D TMA bulk copies) and all four warps -- one row each, with its own causal limit -- score it.  Same
// arithmetic per row as the decode kernel (page-at-a-time online softmax in the log2 domain, 8 lanes per token), no
// cross-warp merge (a warp owns its row), output as the bf16 hi + lo operand of the o_proj GEMM.
// Used below ~512 keys (and where a rank has few heads); beyond, the tensor-core kernel of attn_prefill.cu takes over.
constexpr int kRowsStages = 3;

template <int HD, int G>
__global__ void __launch_bounds__(kAttnWarps * 32) attn_rows_kernel(const __grid_constant__ AttnDev p) {
    constexpr int EL = HD / 16;
    constexpr int TILE = kPT * HD * 2;
    extern __shared__ __align__(128) uint8_t smem[];
    const b200llm_attn_args &a = p.a;
    const int hk = blockIdx.y;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int tg = lane >> 3, p8 = lane & 7;
    uint8_t *tiles = smem;                                                     // [stages][K tile | V tile]
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)kRowsStages * 2 * TILE);

    if (tid == 0) {
        for (int s = 0; s < kRowsStages; ++s) mbar_init(&full[s], 1);
        mbar_fence_init();
    }
    __syncthreads();
    pdl_trigger();
    pdl_wait();  // q and the chunk's K/V come from the previous kernels

    const int row_first = blockIdx.x * kAttnWarps;
    const int row = row_first + warp;
    const bool live = row < a.bs;
    const int row_last = min(row_first + kAttnWarps, a.bs) - 1;
    const int pos = live ? a.row_pos[row] : -1;          // a warp without a row masks everything
    const int n_pages = a.row_pos[row_last] / kPT + 1;   // positions ascend inside a chunk: the last row sees the most
    const int32_t *ptab = a.page_table + (size_t)a.row_lane[row_first] * a.max_pages_per_lane;
    const uint8_t *kbase = reinterpret_cast<const uint8_t *>(a.k_pages);
    const uint8_t *vbase = reinterpret_cast<const uint8_t *>(a.v_pages);

    float qa[G][EL], qb[G][EL];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const float *qp = a.q + (size_t)(live ? row : row_first) * a.n_q * HD + (size_t)(hk * G + g) * HD;
#pragma unroll
        for (int e = 0; e < EL; ++e) {
            qa[g][e] = qp[p8 * EL + e] * p.scale_log2;
            qb[g][e] = qp[HD / 2 + p8 * EL + e] * p.scale_log2;
        }
    }
    float m[G], l[G], oa[G][EL], ob[G][EL];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        m[g] = -INFINITY, l[g] = 0.f;
#pragma unroll
        for (int e = 0; e < EL; ++e) oa[g][e] = 0.f, ob[g][e] = 0.f;
    }
    auto issue = [&](int i) {
        const int st = i % kRowsStages;
        const size_t off = ((size_t)ptab[i] * (a.kv_page_heads > 0 ? a.kv_page_heads : a.n_kv) + hk) * (size_t)TILE;
        mbar_expect_tx(&full[st], 2 * TILE);
        tma_load_1d(tiles + (size_t)st * 2 * TILE, kbase + off, TILE, &full[st]);
        tma_load_1d(tiles + (size_t)st * 2 * TILE + TILE, vbase + off, TILE, &full[st]);
    };
    if (tid == 0)
        for (int i = 0; i < min(n_pages, kRowsStages); ++i) issue(i);
    for (int i = 0; i < n_pages; ++i) {
        const int st = i % kRowsStages;
        mbar_wait(&full[st], (i / kRowsStages) & 1);
        const int t0 = i * kPT;
        if (t0 <= pos) {   // (warp-uniform) the page holds keys this row may see
            const uint8_t *ks = tiles + (size_t)st * 2 * TILE;
            const uint8_t *vs = ks + TILE;
            float sc[kPT / 4][G];
#pragma unroll
            for (int it = 0; it < kPT / 4; ++it) {
                const int t = it * 4 + tg;
                float ka[EL], kb[EL];
                if constexpr (EL == 8) {
                    const uint4 A = *reinterpret_cast<const uint4 *>(ks + (size_t)t * HD * 2 + p8 * 16);
                    const uint4 B = *reinterpret_cast<const uint4 *>(ks + (size_t)t * HD * 2 + HD + p8 * 16);
                    const uint32_t aw[4] = {A.x, A.y, A.z, A.w}, bw[4] = {B.x, B.y, B.z, B.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        ka[2 * e] = bf16lo(aw[e]), ka[2 * e + 1] = bf16hi(aw[e]);
                        kb[2 * e] = bf16lo(bw[e]), kb[2 * e + 1] = bf16hi(bw[e]);
                    }
                } else {
                    const uint2 A = *reinterpret_cast<const uint2 *>(ks + (size_t)t * HD * 2 + p8 * 8);
                    const uint2 B = *reinterpret_cast<const uint2 *>(ks + (size_t)t * HD * 2 + HD + p8 * 8);
                    ka[0] = bf16lo(A.x), ka[1] = bf16hi(A.x), ka[2] = bf16lo(A.y), ka[3] = bf16hi(A.y);
                    kb[0] = bf16lo(B.x), kb[1] = bf16hi(B.x), kb[2] = bf16lo(B.y), kb[3] = bf16hi(B.y);
                }
                const bool ok = (t0 + t) <= pos;
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    float s = 0.f;
#pragma unroll
                    for (int e = 0; e < EL; ++e) s = fmaf(qa[g][e], ka[e], fmaf(qb[g][e], kb[e], s));
                    s += __shfl_xor_sync(0xffffffffu, s, 1);
                    s += __shfl_xor_sync(0xffffffffu, s, 2);
                    s += __shfl_xor_sync(0xffffffffu, s, 4);
                    sc[it][g] = ok ? s : -INFINITY;
                }
            }
#pragma unroll
            for (int g = 0; g < G; ++g) {
                float mn = m[g];
#pragma unroll
                for (int it = 0; it < kPT / 4; ++it) mn = fmaxf(mn, sc[it][g]);
                const float corr = mn == -INFINITY ? 1.f : exp2f(m[g] - mn);
                float ps = 0.f;
#pragma unroll
                for (int it = 0; it < kPT / 4; ++it) {
                    sc[it][g] = mn == -INFINITY ? 0.f : exp2f(sc[it][g] - mn);
                    ps += sc[it][g];
                }
                l[g] = l[g] * corr + ps;
                m[g] = mn;
#pragma unroll
                for (int e = 0; e < EL; ++e) oa[g][e] *= corr, ob[g][e] *= corr;
            }
#pragma unroll
            for (int it = 0; it < kPT / 4; ++it) {
                const int t = it * 4 + tg;
                float va[EL], vb[EL];
                if constexpr (EL == 8) {
                    const uint4 C = *reinterpret_cast<const uint4 *>(vs + (size_t)t * HD * 2 + p8 * 16);
                    const uint4 D = *reinterpret_cast<const uint4 *>(vs + (size_t)t * HD * 2 + HD + p8 * 16);
                    const uint32_t cw[4] = {C.x, C.y, C.z, C.w}, dw[4] = {D.x, D.y, D.z, D.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        va[2 * e] = bf16lo(cw[e]), va[2 * e + 1] = bf16hi(cw[e]);
                        vb[2 * e] = bf16lo(dw[e]), vb[2 * e + 1] = bf16hi(dw[e]);
                    }
                } else {
                    const uint2 C = *reinterpret_cast<const uint2 *>(vs + (size_t)t * HD * 2 + p8 * 8);
                    const uint2 D = *reinterpret_cast<const uint2 *>(vs + (size_t)t * HD * 2 + HD + p8 * 8);
                    va[0] = bf16lo(C.x), va[1] = bf16hi(C.x), va[2] = bf16lo(C.y), va[3] = bf16hi(C.y);
                    vb[0] = bf16lo(D.x), vb[1] = bf16hi(D.x), vb[2] = bf16lo(D.y), vb[3] = bf16hi(D.y);
                }
#pragma unroll
                for (int g = 0; g < G; ++g) {
#pragma unroll
                    for (int e = 0; e < EL; ++e) {
                        oa[g][e] = fmaf(sc[it][g], va[e], oa[g][e]);
                        ob[g][e] = fmaf(sc[it][g], vb[e], ob[g][e]);
                    }
                }
            }
        }
        __syncthreads();   // all four warps are done with the stage: its next page may land
        if (tid == 0 && i + kRowsStages < n_pages) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(i + kRowsStages);
        }
    }
    // merge the 4 token-group streams of the warp; every lane ends with the row's totals
#pragma unroll
    for (int off = 8; off <= 16; off <<= 1) {
#pragma unroll
        for (int g = 0; g < G; ++g) {
            const float m2 = __shfl_xor_sync(0xffffffffu, m[g], off);
            const float l2 = __shfl_xor_sync(0xffffffffu, l[g], off);
            float c1, c2;
            merge_state(m[g], l[g], m2, l2, c1, c2);
#pragma unroll
            for (int e = 0; e < EL; ++e) {
                const float o2a = __shfl_xor_sync(0xffffffffu, oa[g][e], off);
                const float o2b = __shfl_xor_sync(0xffffffffu, ob[g][e], off);
                oa[g][e] = oa[g][e] * c1 + o2a * c2;
                ob[g][e] = ob[g][e] * c1 + o2b * c2;
            }
        }
    }
    if (live && tg == 0) {
#pragma unroll
        for (int g = 0; g < G; ++g) {
            const float inv = 1.f / l[g];
            const size_t base = (size_t)row * a.n_q * HD + (size_t)(hk * G + g) * HD;
#pragma unroll
            for (int e = 0; e < EL; ++e) {
                const float ya = oa[g][e] * inv, yb = ob[g][e] * inv;
                const size_t ia = base + p8 * EL + e, ib = base + HD / 2 + p8 * EL + e;
                if (p.out_bf16 != nullptr) {
                    const uint16_t ha = f32_to_bf16_bits(ya), hb = f32_to_bf16_bits(yb);
                    p.out_bf16[ia] = ha, p.out_bf16[ib] = hb;
                    if (p.out_lo_off != 0) {
                        p.out_bf16[p.out_lo_off + ia] = f32_to_bf16_bits(ya - __uint_as_float((uint32_t)ha << 16));
                        p.out_bf16[p.out_lo_off + ib] = f32_to_bf16_bits(yb - __uint_as_float((uint32_t)hb << 16));
                    }
                } else {
                    a.out[ia] = ya, a.out[ib] = yb;
                }
            }
        }
    }
}

template <int HD, int G>
static int launch_attn_rows(const AttnDev &d, cudaStream_t stream, bool pdl) {
    constexpr size_t smem = (size_t)kRowsStages * 2 * kPT * HD * 2 + kRowsStages * 8;
    LaunchCfg lc(dim3((d.a.bs + kAttnWarps - 1) / kAttnWarps, d.a.n_kv), dim3(kAttnWarps * 32), smem, stream, pdl);
    B200_CUDA_OK(cudaLaunchKernelEx(&lc.cfg, attn_rows_kernel<HD, G>, d));
    return B200LLM_OK;
}

// rows = consecutive positions of one lane (a prefill chunk): a.row_pos ascending, a.row_lane all equal
int attn_rows_launch(const b200llm_attn_args &a, cudaStream_t stream, bool pdl, void *out_bf16, int64_t out_lo_off) {
    B200_REQUIRE(a.page_tokens == kPT, "attn: page_tokens must be %d (got %d)", kPT, a.page_tokens);
    B200_REQUIRE(a.n_kv > 0 && a.n_q % a.n_kv == 0 && a.bs >= 1, "attn: bad head counts / rows");
    AttnDev d{};
    d.a = a;
    d.out_bf16 = reinterpret_cast<uint16_t *>(out_bf16);
    d.out_lo_off = out_lo_off;
    d.scale_log2 = (float)(1.4426950408889634 / sqrt((double)a.head_dim));
    const int G = a.n_q / a.n_kv;
#define B200_ATTN_CASE(HD_, G_) \
    if (a.head_dim == HD_ && G == G_) return launch_attn_rows<HD_, G_>(d, stream, pdl);
    B200_ATTN_CASE(128, 1) B200_ATTN_CASE(128, 2) B200_ATTN_CASE(128, 3) B200_ATTN_CASE(128, 4)
    B200_ATTN_CASE(64, 1) B200_ATTN_CASE(64, 2) B200_ATTN_CASE(64, 4)
#undef B200_ATTN_CASE
    set_error("attn: unsupported head_dim %d / group %d", a.head_dim, G);
    return B200LLM_EINVAL;
}

// ---------------------------------------------------------------- no-allocation variant
//
// b200llm_op_attn_decode_static: the same split-T decode attention for callers that cannot
// hand us a scratch buffer because they cannot allocate one.  Two kinds of host program are in
// that position: one that gave every byte of HBM to its own pool at start-up (a serving process
// that reserves the whole card for weights and KV pages, so a late cudaMalloc fails), and one
// that calls from a thread where cudaMalloc / cudaFree are off limits because both take the
// allocator lock and synchronise the whole device (a latency-critical decode loop).
//
// The split partials and arrival tickets live in the array below instead.  It is part of the
// library image: every GPU that loads libb200llm.so carries these 4 MiB until the process
// exits, whether or not this entry point is ever called.
constexpr size_t kStaticScratchBytes = 4u << 20;   // the stated maximum: tickets + partials must fit
constexpr size_t kStaticTicketBytes = 4096;        // 1024 (row, kv head) tickets
static __device__ __align__(256) unsigned char g_attn_static_scratch[kStaticScratchBytes];

// In host code `g_attn_static_scratch` is only a handle for the runtime, not a device address.
// The address a kernel can use comes from cudaGetSymbolAddress, asked for once per device with
// that device current.
static std::mutex g_static_mu;
static void *g_static_ptr[64] = {nullptr};

static int static_scratch_ptr(void **out) {
    int dev = -1;
    B200_CUDA_OK(cudaGetDevice(&dev));
    B200_REQUIRE(dev >= 0 && dev < 64, "attn static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (g_static_ptr[dev] == nullptr) {
        void *ptr = nullptr;
        B200_CUDA_OK(cudaGetSymbolAddress(&ptr, g_attn_static_scratch));
        g_static_ptr[dev] = ptr;
    }
    *out = g_static_ptr[dev];
    return B200LLM_OK;
}

}  // namespace b200

extern "C" int b200llm_op_attn_decode_static(const b200llm_attn_args *a, void *stream) {
    using namespace b200;
    B200_REQUIRE(a && a->q && a->out && a->k_pages && a->v_pages && a->row_lane && a->row_pos && a->page_table,
                 "op_attn_decode_static: NULL argument");
    {
        int n_dev = 0, cur = 0, major = 0;
        if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) {
            set_error("no CUDA device: this library has no CPU fallback");
            return B200LLM_ENODEV;
        }
        B200_CUDA_OK(cudaGetDevice(&cur));
        B200_CUDA_OK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, cur));
        if (major != 10) {
            set_error("current device is sm_%dx; this library is built for sm_100a (B200) only", major);
            return B200LLM_ENODEV;
        }
    }
    {
        const int G = a->n_kv > 0 ? a->n_q / a->n_kv : 0;
        const bool known = (a->head_dim == 128 && G >= 1 && G <= 4) || (a->head_dim == 64 && (G == 1 || G == 2 || G == 4));
        B200_REQUIRE(known, "attn static: unsupported head_dim %d / group %d", a->head_dim, G);   // before anything is queued on the stream
    }
    B200_REQUIRE(a->page_tokens == kPT, "attn static: page_tokens must be %d (got %d)", kPT, a->page_tokens);
    B200_REQUIRE(a->n_kv > 0 && a->n_q % a->n_kv == 0, "attn static: n_q %d not a multiple of n_kv %d", a->n_q, a->n_kv);
    B200_REQUIRE(a->bs >= 1 && a->n_splits >= 1 && a->n_splits <= 64, "attn static: bad bs/n_splits");
    const size_t tickets = (size_t)a->bs * a->n_kv * 4;
    const size_t part_o = (size_t)a->bs * a->n_q * a->n_splits * a->head_dim * 4;
    const size_t part_ml = (size_t)a->bs * a->n_q * a->n_splits * 2 * 4;
    if (tickets > kStaticTicketBytes || kStaticTicketBytes + part_o + part_ml > kStaticScratchBytes) {
        set_error("attn static: bs %d x n_q %d x n_splits %d x head_dim %d needs %zu bytes of split scratch (and %zu of "
                  "tickets); the built-in buffer holds %zu (tickets %zu). Use b200llm_op_attn_decode with your own "
                  "scratch for larger launches.",
                  a->bs, a->n_q, a->n_splits, a->head_dim, kStaticTicketBytes + part_o + part_ml, tickets,
                  kStaticScratchBytes, kStaticTicketBytes);
        return B200LLM_ENOMEM;
    }
    void *base = nullptr;
    int rc = static_scratch_ptr(&base);   // first call on a device resolves the symbol address
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // the buffer is shared by every call and keeps the last call's contents: the tickets must
    // start at zero, whatever an earlier (possibly failed) launch left behind
    B200_CUDA_OK(cudaMemsetAsync(base, 0, kStaticTicketBytes, st));
    AttnDev d{};
    d.a = *a;
    d.a.scratch = base, d.a.scratch_bytes = kStaticScratchBytes;
    d.scale_log2 = (float)(1.4426950408889634 / sqrt((double)a->head_dim));
    uint8_t *s = reinterpret_cast<uint8_t *>(base);
    d.tickets = reinterpret_cast<unsigned *>(s);
    d.part_o = reinterpret_cast<float *>(s + kStaticTicketBytes);
    d.part_ml = d.part_o + (size_t)a->bs * a->n_q * a->n_splits * a->head_dim;
    const int G = a->n_q / a->n_kv;
#define B200_ATTN_CASE(HD_, G_) \
    if (a->head_dim == HD_ && G == G_) return launch_attn<HD_, G_>(d, st, false);
    B200_ATTN_CASE(128, 1) B200_ATTN_CASE(128, 2) B200_ATTN_CASE(128, 3) B200_ATTN_CASE(128, 4)
    B200_ATTN_CASE(64, 1) B200_ATTN_CASE(64, 2) B200_ATTN_CASE(64, 4)
#undef B200_ATTN_CASE
    set_error("attn static: unsupported head_dim %d / group %d", a->head_dim, G);
    return B200LLM_EINVAL;
}
