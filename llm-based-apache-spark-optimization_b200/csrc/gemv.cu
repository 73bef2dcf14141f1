// gemv.cu -- bf16 weight GEMV for decode at batch <= 4 (kernels K2/K5/K6/K7/K8 of
// SURVEY.md section 2.3), HBM-bound: every weight byte is read exactly once.
//
// One CTA owns a contiguous block of weight rows (= one contiguous byte range of W).
// A producer warp streams that range through a shared-memory ring with 1-D TMA bulk
// copies (cp.async.bulk + mbarrier complete_tx); the ring is filled BEFORE
// griddepcontrol.wait, so under programmatic dependent launch the weight stream of
// kernel i+1 is already in flight while kernel i drains.  Compute threads keep their
// K-slice of the activation vector(s) in registers (fp32), FMA bf16 weights from
// shared memory, warp-shuffle reduce, and combine per-warp partials once at the end,
// where the fused epilogue runs (residual add / SwiGLU / RoPE + paged-KV append /
// logits + per-CTA argmax).  Optional fused RMSNorm prologue.
#include "common.cuh"
#include "kernels.h"

namespace b200 {

constexpr int kGemvMaxComputeThreads = 512;
constexpr int kGemvRingBytes = 80 * 1024;  // ring + partials < 113 KB: two CTAs (this kernel's and the next one's, under PDL) fit one SM
constexpr int kGemvMaxStages = 16;
constexpr int kGemvPartBytes = 24 * 1024;

struct GemvPlan {
    int upt;       // 16-byte units of K per compute thread
    int nt;        // compute threads
    int rg;        // rows per stage
    int stages;
    int grid;
    int gran;      // row granularity of the CTA partition (2 when the epilogue pairs rows)
    size_t smem;
    int max_rows_cta;
};

static bool gemv_plan(int N, int K, int bs, int epi, int grid_hint, GemvPlan *pl) {
    if (K % 8 || K <= 0 || N <= 0 || bs < 1 || bs > 4) return false;
    const int units = K / 8;
    pl->upt = (units + kGemvMaxComputeThreads - 1) / kGemvMaxComputeThreads;
    if (pl->upt > 3) return false;
    pl->nt = ((units + pl->upt - 1) / pl->upt + 31) / 32 * 32;
    pl->rg = pl->upt == 1 ? 4 : (pl->upt == 2 ? 2 : 1);
    const size_t stage_bytes = (size_t)pl->rg * K * 2;
    int s = (int)(kGemvRingBytes / stage_bytes);
    pl->stages = s < 2 ? 2 : (s > kGemvMaxStages ? kGemvMaxStages : s);
    pl->gran = (epi == B200LLM_EPI_SWIGLU || epi == B200LLM_EPI_QKV_ROPE) ? 2 : 1;
    if (N % pl->gran) return false;
    const int nw = pl->nt / 32;
    const int n_units = N / pl->gran;
    int grid = grid_hint;
    if (grid <= 0) {
        for (int m = 1;; ++m) {
            grid = kNumSMs * m;
            int rows = ((n_units + grid - 1) / grid) * pl->gran;
            if ((size_t)rows * bs * nw * 4 <= (size_t)kGemvPartBytes) break;
        }
    }
    if (grid > n_units) grid = n_units;
    pl->grid = grid;
    pl->max_rows_cta = ((n_units + grid - 1) / grid) * pl->gran;
    size_t part = ((size_t)pl->max_rows_cta * bs * nw * 4 + 127) / 128 * 128;
    pl->smem = (size_t)pl->stages * stage_bytes + part + 2 * kGemvMaxStages * 8 + 64 * 4 + 128;
    return pl->smem <= 220 * 1024;
}

struct GemvDev {
    b200llm_gemv_args a;
    int gran, stages, nt, max_rows_cta;
};

template <int BS, int UPT>
__global__ void __launch_bounds__(kGemvMaxComputeThreads + 32, (BS * UPT <= 2) ? 2 : 1)
gemv_kernel(const __grid_constant__ GemvDev p) {
    constexpr int RG = UPT == 1 ? 4 : (UPT == 2 ? 2 : 1);
    extern __shared__ __align__(128) uint8_t smem[];
    const b200llm_gemv_args &a = p.a;
    const int K = a.K;
    const int nt = p.nt;
    const int nw = nt >> 5;
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const size_t row_bytes = (size_t)K * 2;
    const size_t stage_bytes = (size_t)RG * row_bytes;
    const int S = p.stages;

    uint8_t *ring = smem;
    float *part = reinterpret_cast<float *>(smem + (size_t)S * stage_bytes);
    const size_t part_bytes = ((size_t)p.max_rows_cta * BS * nw * 4 + 127) / 128 * 128;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)S * stage_bytes + part_bytes);
    uint64_t *empty = full + kGemvMaxStages;
    float *red = reinterpret_cast<float *>(empty + kGemvMaxStages);

    // --- CTA row range (in units of `gran` rows) ---
    const int n_units = a.N / p.gran;
    const int G = gridDim.x, bid = blockIdx.x;
    const int base = n_units / G, rem = n_units % G;
    const int my_units = base + (bid < rem ? 1 : 0);
    const int row0 = (bid * base + (bid < rem ? bid : rem)) * p.gran;
    const int nrows = my_units * p.gran;
    const int n_groups = (nrows + RG - 1) / RG;

    if (tid == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], nw);
        }
        mbar_fence_init();
    }
    __syncthreads();
    pdl_trigger();

    if (warp == nw) {
        // ===== producer warp: weights do not depend on the previous kernel -> no pdl_wait =====
        if (lane == 0) {
            const uint8_t *wsrc = reinterpret_cast<const uint8_t *>(a.w) + (size_t)row0 * row_bytes;
            const uint64_t pol = l2_policy_evict_first();
            for (int g = 0; g < n_groups; ++g) {
                const int slot = g % S, use = g / S;
                if (use > 0) mbar_wait(&empty[slot], (use - 1) & 1);
                const int nr = min(RG, nrows - g * RG);
                const uint32_t bytes = (uint32_t)(nr * row_bytes);
                mbar_expect_tx(&full[slot], bytes);
                tma_load_1d_hint(ring + (size_t)slot * stage_bytes, wsrc + (size_t)g * stage_bytes, bytes,
                                 &full[slot], pol);
            }
        }
        return;
    }

    // ===== compute warps =====
    pdl_wait();  // x (and h for the residual) come from the previous kernel

    float xr[BS][UPT][8];
    bool valid[UPT];
#pragma unroll
    for (int j = 0; j < UPT; ++j) {
        const int k0 = (j * nt + tid) * 8;
        valid[j] = k0 < K;
#pragma unroll
        for (int b = 0; b < BS; ++b) {
            if (valid[j]) {
                const float4 v0 = *reinterpret_cast<const float4 *>(a.x + (size_t)b * a.x_stride + k0);
                const float4 v1 = *reinterpret_cast<const float4 *>(a.x + (size_t)b * a.x_stride + k0 + 4);
                xr[b][j][0] = v0.x, xr[b][j][1] = v0.y, xr[b][j][2] = v0.z, xr[b][j][3] = v0.w;
                xr[b][j][4] = v1.x, xr[b][j][5] = v1.y, xr[b][j][6] = v1.z, xr[b][j][7] = v1.w;
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) xr[b][j][e] = 0.f;
            }
        }
    }

    if (a.norm_gain != nullptr) {
        // fused RMSNorm prologue: x <- gain * (x * rsqrt(mean(x^2) + eps)), every CTA redundantly
        float ss[BS];
#pragma unroll
        for (int b = 0; b < BS; ++b) {
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < UPT; ++j)
#pragma unroll
                for (int e = 0; e < 8; ++e) s += xr[b][j][e] * xr[b][j][e];
            ss[b] = warp_sum(s);
        }
        if (lane == 0) {
#pragma unroll
            for (int b = 0; b < BS; ++b) red[b * 16 + warp] = ss[b];
        }
        bar_sync(1, nt);
#pragma unroll
        for (int b = 0; b < BS; ++b) {
            float s = 0.f;
            for (int w = 0; w < nw; ++w) s += red[b * 16 + w];
            ss[b] = rsqrtf(s / (float)K + a.rms_eps);
        }
        const uint16_t *gain = reinterpret_cast<const uint16_t *>(a.norm_gain);
#pragma unroll
        for (int j = 0; j < UPT; ++j) {
            if (valid[j]) {
                const int k0 = (j * nt + tid) * 8;
                const uint4 gv = *reinterpret_cast<const uint4 *>(gain + k0);
                const uint32_t gw[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
                for (int b = 0; b < BS; ++b)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        xr[b][j][2 * e] = bf16lo(gw[e]) * (xr[b][j][2 * e] * ss[b]);
                        xr[b][j][2 * e + 1] = bf16hi(gw[e]) * (xr[b][j][2 * e + 1] * ss[b]);
                    }
            }
        }
    }

    // ----- main loop over row groups -----
    for (int g = 0; g < n_groups; ++g) {
        const int slot = g % S, use = g / S;
        mbar_wait(&full[slot], use & 1);
        const uint8_t *st = ring + (size_t)slot * stage_bytes;
        const int nr = min(RG, nrows - g * RG);
        float acc[RG][BS];
#pragma unroll
        for (int r = 0; r < RG; ++r)
#pragma unroll
            for (int b = 0; b < BS; ++b) acc[r][b] = 0.f;
#pragma unroll
        for (int r = 0; r < RG; ++r) {
            if (r < nr) {
#pragma unroll
                for (int j = 0; j < UPT; ++j) {
                    if (valid[j]) {
                        const uint4 wv =
                            *reinterpret_cast<const uint4 *>(st + (size_t)r * row_bytes + (size_t)(j * nt + tid) * 16);
                        const uint32_t ww[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float w0 = bf16lo(ww[e]), w1 = bf16hi(ww[e]);
#pragma unroll
                            for (int b = 0; b < BS; ++b) {
                                acc[r][b] = fmaf(w0, xr[b][j][2 * e], acc[r][b]);
                                acc[r][b] = fmaf(w1, xr[b][j][2 * e + 1], acc[r][b]);
                            }
                        }
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);  // ring slot may be refilled
#pragma unroll
        for (int r = 0; r < RG; ++r)
#pragma unroll
            for (int b = 0; b < BS; ++b) acc[r][b] = warp_sum(acc[r][b]);
        if (lane == 0) {
#pragma unroll
            for (int r = 0; r < RG; ++r)
                if (r < nr) {
#pragma unroll
                    for (int b = 0; b < BS; ++b) part[((size_t)(g * RG + r) * BS + b) * nw + warp] = acc[r][b];
                }
        }
    }
    bar_sync(1, nt);

    // ----- combine warp partials + fused epilogue -----
    const int epi = a.epi;
    float best[BS];
    int best_i[BS];
#pragma unroll
    for (int b = 0; b < BS; ++b) best[b] = -INFINITY, best_i[b] = 0x7fffffff;

    for (int u = tid; u < my_units; u += nt) {
        const int rl = u * p.gran;
        const int n = row0 + rl;
#pragma unroll
        for (int b = 0; b < BS; ++b) {
            float v0 = 0.f, v1 = 0.f;
            const float *p0 = part + ((size_t)rl * BS + b) * nw;
            for (int w = 0; w < nw; ++w) v0 += p0[w];
            if (p.gran == 2) {
                const float *p1 = part + ((size_t)(rl + 1) * BS + b) * nw;
                for (int w = 0; w < nw; ++w) v1 += p1[w];
            }
            if (epi == B200LLM_EPI_STORE) {
                a.out[(size_t)b * a.out_stride + n] = v0;
            } else if (epi == B200LLM_EPI_RESIDUAL) {
                a.out[(size_t)b * a.out_stride + n] += v0;
            } else if (epi == B200LLM_EPI_SWIGLU) {
                const float sg = v0 / (1.f + expf(-v0));
                a.out[(size_t)b * a.out_stride + (n >> 1)] = sg * v1;
            } else if (epi == B200LLM_EPI_LOGITS) {
                a.out[(size_t)b * a.out_stride + n] = v0;
                if (v0 > best[b]) best[b] = v0, best_i[b] = n;  // rows ascend per thread: first max kept
            } else {  // B200LLM_EPI_QKV_ROPE
                const int hd = a.head_dim;
                const int qrows = a.n_q * hd, krows = a.n_kv * hd;
                const int pos = a.row_pos[b];
                if (n < qrows + krows) {
                    const int i = (n % hd) >> 1;
                    const float ang = (float)pos * a.inv_freq[i];
                    float sn, cs;
                    sincosf(ang, &sn, &cs);
                    const float o0 = v0 * cs - v1 * sn;
                    const float o1 = v1 * cs + v0 * sn;
                    v0 = o0, v1 = o1;
                }
                if (n < qrows) {
                    a.out[(size_t)b * a.out_stride + n] = v0;
                    a.out[(size_t)b * a.out_stride + n + 1] = v1;
                } else {
                    const int ln = a.row_lane[b];
                    const int page = a.page_table[(size_t)ln * a.max_pages_per_lane + pos / a.page_tokens];
                    const int slot_t = pos % a.page_tokens;
                    const bool is_k = n < qrows + krows;
                    const int kn = is_k ? n - qrows : n - qrows - krows;
                    const int hk = kn / hd, dim = kn % hd;
                    uint16_t *dst = reinterpret_cast<uint16_t *>(is_k ? a.k_pages : a.v_pages) +
                                    (((size_t)page * (a.kv_page_heads > 0 ? a.kv_page_heads : a.n_kv) + hk) * a.page_tokens + slot_t) * hd + dim;
                    const uint32_t packed = (uint32_t)f32_to_bf16_bits(v0) | ((uint32_t)f32_to_bf16_bits(v1) << 16);
                    *reinterpret_cast<uint32_t *>(dst) = packed;
                }
            }
        }
    }

    if (epi == B200LLM_EPI_LOGITS) {
        // per-CTA (max, lowest argmax) -> partials, reduced by the sampler kernel
        float *rmax = red;  // [16], reused per batch row
        int *ridx = reinterpret_cast<int *>(red) + 16;
#pragma unroll
        for (int b = 0; b < BS; ++b) {
            float m = best[b];
            int mi = best_i[b];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float om = __shfl_xor_sync(0xffffffffu, m, o);
                const int oi = __shfl_xor_sync(0xffffffffu, mi, o);
                if (om > m || (om == m && oi < mi)) m = om, mi = oi;
            }
            bar_sync(1, nt);  // red reuse
            if (lane == 0) rmax[warp] = m, ridx[warp] = mi;
            bar_sync(1, nt);
            if (tid == 0) {
                float bm = rmax[0];
                int bi = ridx[0];
                for (int w = 1; w < nw; ++w)
                    if (rmax[w] > bm || (rmax[w] == bm && ridx[w] < bi)) bm = rmax[w], bi = ridx[w];
                a.part_max[(size_t)b * G + bid] = bm;
                a.part_idx[(size_t)b * G + bid] = bi + a.logits_idx_offset;
            }
        }
    }
}

template <int BS, int UPT>
static int launch_one(const GemvDev &dev, const GemvPlan &pl, cudaStream_t stream, bool pdl) {
    // per launch: the attribute belongs to the (function, device) pair, and callers may use several devices and threads
    B200_CUDA_OK(cudaFuncSetAttribute(gemv_kernel<BS, UPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    LaunchCfg lc(dim3(pl.grid), dim3(pl.nt + 32), pl.smem, stream, pdl);
    B200_CUDA_OK(cudaLaunchKernelEx(&lc.cfg, gemv_kernel<BS, UPT>, dev));
    return B200LLM_OK;
}

int gemv_launch(const b200llm_gemv_args &a, cudaStream_t stream, bool pdl) {
    GemvPlan pl;
    B200_REQUIRE(gemv_plan(a.N, a.K, a.bs, a.epi, a.grid_hint, &pl),
                 "gemv: unsupported shape N=%d K=%d bs=%d epi=%d (need K%%8==0, K<=12288, 1<=bs<=4)", a.N, a.K,
                 a.bs, a.epi);
    B200_REQUIRE(a.epi >= 0 && a.epi <= B200LLM_EPI_LOGITS, "gemv: bad epilogue %d", a.epi);
    B200_REQUIRE(a.x_stride % 4 == 0 && ((uintptr_t)a.x % 16) == 0 && ((uintptr_t)a.w % 16) == 0,
                 "gemv: x and w must be 16-byte aligned");
    if (a.epi == B200LLM_EPI_QKV_ROPE)
        B200_REQUIRE(a.N == (a.n_q + 2 * a.n_kv) * a.head_dim && a.head_dim % 2 == 0, "gemv: qkv rows mismatch");
    GemvDev dev;
    dev.a = a;
    dev.gran = pl.gran;
    dev.stages = pl.stages;
    dev.nt = pl.nt;
    dev.max_rows_cta = pl.max_rows_cta;
#define B200_GEMV_CASE(BS_, UPT_) \
    if (a.bs == BS_ && pl.upt == UPT_) return launch_one<BS_, UPT_>(dev, pl, stream, pdl);
    B200_GEMV_CASE(1, 1) B200_GEMV_CASE(1, 2) B200_GEMV_CASE(1, 3)
    B200_GEMV_CASE(2, 1) B200_GEMV_CASE(2, 2) B200_GEMV_CASE(2, 3)
    B200_GEMV_CASE(3, 1) B200_GEMV_CASE(3, 2) B200_GEMV_CASE(3, 3)
    B200_GEMV_CASE(4, 1) B200_GEMV_CASE(4, 2) B200_GEMV_CASE(4, 3)
#undef B200_GEMV_CASE
    set_error("gemv: no instantiation for bs=%d upt=%d", a.bs, pl.upt);
    return B200LLM_EINVAL;
}

int gemv_grid(int N, int K, int bs, int epi, int grid_hint) {
    GemvPlan pl;
    if (!gemv_plan(N, K, bs, epi, grid_hint, &pl)) return -1;
    return pl.grid;
}

}  // namespace b200
