// misc.cu -- the small kernels around the GEMVs: token-embedding gather (K1), the fused
// argmax / top-k / top-p / temperature sampler (K8 epilogue), the
// seeded weight stream (device twin of the spec in DESIGN.md "Weights"), and a residual add
// used only on the tensor-parallel path.
#include "common.cuh"
#include "kernels.h"

namespace b200 {

// ------------------------------------------------------------------ embed gather
__global__ void __launch_bounds__(256) embed_kernel(const __grid_constant__ EmbedArgs a, int32_t *lane_hist,
                                                    int32_t *step_ctr) {
    pdl_trigger();
    pdl_wait();
    const int row = blockIdx.x;
    int tok, pos;
    const int ln = a.row_lane[row];
    if (a.from_lanes) {
        tok = a.lane_token[ln];
        pos = a.lane_pos[ln];
    } else {
        tok = a.row_token[row];
        pos = a.row_pos[row];
    }
    if (threadIdx.x == 0) {
        if (a.from_lanes) {
            a.row_token[row] = tok;
            a.row_pos[row] = pos;
            if (row == 0 && step_ctr) *step_ctr += 1;
        }
        if (lane_hist) lane_hist[(size_t)ln * 64 + (pos & 63)] = tok;
    }
    tok = min(max(tok, 0), a.vocab - 1);
    const uint16_t *src = reinterpret_cast<const uint16_t *>(a.embed) + (size_t)tok * a.d;
    float *dst = a.h + (size_t)row * a.d;
    for (int k = threadIdx.x * 8; k < a.d; k += blockDim.x * 8) {
        const uint4 v = *reinterpret_cast<const uint4 *>(src + k);
        float4 o0, o1;
        o0.x = bf16lo(v.x), o0.y = bf16hi(v.x), o0.z = bf16lo(v.y), o0.w = bf16hi(v.y);
        o1.x = bf16lo(v.z), o1.y = bf16hi(v.z), o1.z = bf16lo(v.w), o1.w = bf16hi(v.w);
        *reinterpret_cast<float4 *>(dst + k) = o0;
        *reinterpret_cast<float4 *>(dst + k + 4) = o1;
    }
}

int embed_launch_ex(const EmbedArgs &a, int32_t *lane_hist, int32_t *step_ctr, cudaStream_t stream, bool pdl) {
    B200_REQUIRE(a.d % 8 == 0 && a.rows >= 1, "embed: d %% 8 != 0 or no rows");
    LaunchCfg lc(dim3(a.rows), dim3(256), 0, stream, pdl);
    B200_CUDA_OK(cudaLaunchKernelEx(&lc.cfg, embed_kernel, a, lane_hist, step_ctr));
    return B200LLM_OK;
}
int embed_launch(const EmbedArgs &a, cudaStream_t stream, bool pdl) {
    return embed_launch_ex(a, nullptr, nullptr, stream, pdl);
}

// ------------------------------------------------------------------ sampler
constexpr int kSampThreads = 1024;
constexpr int kSampCap = 1024;

__device__ __forceinline__ uint32_t fkey(float f) {  // order-preserving float -> uint
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

__global__ void __launch_bounds__(kSampThreads) sample_kernel(const __grid_constant__ SampleDev p) {
    __shared__ float s_val[kSampCap];
    __shared__ int s_idx[kSampCap];
    __shared__ unsigned s_hist[256];
    __shared__ float s_rmax[32];
    __shared__ int s_ridx[32];
    __shared__ unsigned s_prefix, s_want, s_count;
    __shared__ int s_token;
    pdl_trigger();
    pdl_wait();
    const b200llm_sample_args &a = p.a;
    const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ln = p.row_lane ? p.row_lane[row] : row;
    const b200llm_sampling sp = a.params ? a.params[ln] : b200llm_sampling{0.f, 0, 1.f, 1.f, 0};
    float *logits = const_cast<float *>(a.logits) + (size_t)row * a.vocab;
    const int V = a.vocab;
    const int32_t *hist = p.lane_hist ? p.lane_hist + (size_t)ln * 64 : (a.history ? a.history + (size_t)row * 64 : nullptr);

    const bool penal = sp.repeat_penalty != 1.f && hist != nullptr;
    if (penal) {
        // each distinct recent token once: logit > 0 ? logit / penalty : logit * penalty
        if (tid < 64) {
            const int t = hist[tid];
            bool dup = t < 0 || t >= V;
            for (int j = 0; j < tid && !dup; ++j) dup = hist[j] == t;
            if (!dup) {
                const float v = logits[t];
                logits[t] = v > 0.f ? v / sp.repeat_penalty : v * sp.repeat_penalty;
            }
        }
        __syncthreads();
    }

    if (sp.temperature <= 0.f) {
        // ---- greedy: argmax, ties -> lowest id ----
        float m = -INFINITY;
        int mi = 0x7fffffff;
        if (a.part_max != nullptr && !penal) {
            const int nr = p.n_ranks > 1 ? p.n_ranks : 1;
            for (int r = 0; r < nr; ++r)
                for (int i = tid; i < a.n_part; i += kSampThreads) {
                    const size_t at = (size_t)r * p.rank_stride + (size_t)row * a.n_part + i;
                    const float v = a.part_max[at];
                    const int vi = a.part_idx[at];
                    if (v > m || (v == m && vi < mi)) m = v, mi = vi;
                }
        } else {
            for (int i = tid; i < V; i += kSampThreads) {
                const float v = logits[i];
                if (v > m) m = v, mi = i;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float om = __shfl_xor_sync(0xffffffffu, m, o);
            const int oi = __shfl_xor_sync(0xffffffffu, mi, o);
            if (om > m || (om == m && oi < mi)) m = om, mi = oi;
        }
        if (lane == 0) s_rmax[warp] = m, s_ridx[warp] = mi;
        __syncthreads();
        if (tid == 0) {
            for (int w = 1; w < kSampThreads / 32; ++w)
                if (s_rmax[w] > m || (s_rmax[w] == m && s_ridx[w] < mi)) m = s_rmax[w], mi = s_ridx[w];
            s_token = mi;
        }
        __syncthreads();
    } else {
        // ---- top-k by radix select on order-preserving keys (4 x 8 bits) ----
        int k = sp.top_k > 0 ? min(sp.top_k, kSampCap) : kSampCap;
        k = min(k, V);
        if (tid == 0) s_prefix = 0, s_want = (unsigned)k;
        __syncthreads();
        for (int pass = 0; pass < 4; ++pass) {
            const int shift = 24 - 8 * pass;
            if (tid < 256) s_hist[tid] = 0;
            __syncthreads();
            const unsigned prefix = s_prefix;
            const unsigned pmask = pass == 0 ? 0u : (0xFFFFFFFFu << (shift + 8));
            for (int i = tid; i < V; i += kSampThreads) {
                const uint32_t key = fkey(logits[i]);
                if ((key & pmask) == prefix) atomicAdd(&s_hist[(key >> shift) & 255u], 1u);
            }
            __syncthreads();
            if (tid == 0) {
                unsigned want = s_want, acc = 0;
                int b = 255;
                for (; b > 0; --b) {
                    if (acc + s_hist[b] >= want) break;
                    acc += s_hist[b];
                }
                s_want = want - acc;
                s_prefix = prefix | ((unsigned)b << shift);
            }
            __syncthreads();
        }
        const uint32_t kth = s_prefix;  // key of the k-th largest logit
        if (tid == 0) s_count = 0;
        __syncthreads();
        for (int i = tid; i < V; i += kSampThreads) {
            const float v = logits[i];
            if (fkey(v) >= kth) {
                const unsigned slot = atomicAdd(&s_count, 1u);
                if (slot < (unsigned)kSampCap) s_val[slot] = v, s_idx[slot] = i;
            }
        }
        __syncthreads();
        const int n = min((int)s_count, kSampCap);
        for (int i = tid; i < kSampCap; i += kSampThreads)
            if (i >= n) s_val[i] = -INFINITY, s_idx[i] = 0x7fffffff;
        __syncthreads();
        // bitonic sort, descending by (value, then ascending id)
        for (int size = 2; size <= kSampCap; size <<= 1) {
            for (int stride = size >> 1; stride > 0; stride >>= 1) {
                const int i = tid, j = i ^ stride;
                if (j > i) {
                    const bool desc = (i & size) == 0;
                    const float vi = s_val[i], vj = s_val[j];
                    const int ii = s_idx[i], ij = s_idx[j];
                    const bool i_before_j = vi > vj || (vi == vj && ii < ij);
                    if (desc ? !i_before_j : i_before_j) {
                        s_val[i] = vj, s_val[j] = vi;
                        s_idx[i] = ij, s_idx[j] = ii;
                    }
                }
                __syncthreads();
            }
        }
        if (tid == 0) {
            const int kk = min(k, n);
            const float inv_t = 1.f / sp.temperature;
            const float v0 = s_val[0];
            // probabilities (unnormalised), sequential fp32 sums: same order as the oracle sampler
            float total = 0.f;
            for (int i = 0; i < kk; ++i) {
                const float e = expf((s_val[i] - v0) * inv_t);
                s_val[i] = e;
                total += e;
            }
            int keep = kk;
            if (sp.top_p < 1.f) {
                float cum = 0.f;
                for (int i = 0; i < kk; ++i) {
                    cum += s_val[i];
                    if (cum / total >= sp.top_p) {
                        keep = i + 1;
                        break;
                    }
                }
            }
            float mass = 0.f;
            for (int i = 0; i < keep; ++i) mass += s_val[i];
            const uint64_t ctr = p.lane_pos ? (uint64_t)p.lane_pos[ln] : a.step;
            const uint64_t z = mix64(sp.seed + (ctr + 1ULL) * 0x9E3779B97F4A7C15ULL);
            const float u = (float)(z >> 40) * (1.f / 16777216.f);
            const float r = u * mass;
            float cum = 0.f;
            int pick = keep - 1;
            for (int i = 0; i < keep; ++i) {
                cum += s_val[i];
                if (cum > r) {
                    pick = i;
                    break;
                }
            }
            s_token = s_idx[pick];
        }
        __syncthreads();
    }

    if (tid == 0) {
        const int tok = s_token;
        if (a.out_token) a.out_token[row] = tok;
        if (p.advance) {
            p.lane_token[ln] = tok;
            p.lane_pos[ln] += 1;
            if (p.gen_out && p.step_ctr) p.gen_out[(size_t)(*p.step_ctr - 1) * a.bs + row] = tok;
        }
    }
}

int sample_launch(const SampleDev &d, cudaStream_t stream, bool pdl) {
    B200_REQUIRE(d.a.bs >= 1 && d.a.vocab >= 1 && d.a.logits != nullptr, "sample: bad arguments");
    LaunchCfg lc(dim3(d.a.bs), dim3(kSampThreads), 0, stream, pdl);
    B200_CUDA_OK(cudaLaunchKernelEx(&lc.cfg, sample_kernel, d));
    return B200LLM_OK;
}

// ------------------------------------------------------------------ seeded weight stream
__global__ void fill_weights_kernel(uint16_t *out, uint64_t n, uint64_t seed, float scale, float mean) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t z = mix64(seed + (i + 1ULL) * 0x9E3779B97F4A7C15ULL);
        const int s = (int)(z & 0xFFFF) + (int)((z >> 16) & 0xFFFF) + (int)((z >> 32) & 0xFFFF) + (int)(z >> 48);
        const float v = __fmul_rn((float)(s - 131070), scale);
        out[i] = f32_to_bf16_bits(__fadd_rn(mean, v));
    }
}

int fill_weights_launch(void *dev_bf16, uint64_t n, uint64_t seed, float sigma, float mean, cudaStream_t s) {
    if (n == 0) return B200LLM_OK;
    const float scale = sigma / 37837.22765f;
    const unsigned blocks = (unsigned)((n + 1023) / 1024 > 148u * 16u ? 148u * 16u : (n + 1023) / 1024);
    fill_weights_kernel<<<blocks, 1024, 0, s>>>(reinterpret_cast<uint16_t *>(dev_bf16), n, seed, scale, mean);
    B200_CUDA_OK(cudaGetLastError());
    return B200LLM_OK;
}

// ------------------------------------------------------------------ dst += src (TP residual)
__global__ void add_inplace_kernel(float *dst, const float *src, size_t n) {
    pdl_trigger();
    pdl_wait();
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        dst[i] += src[i];
}
int add_inplace_launch(float *dst, const float *src, size_t n, cudaStream_t s, bool pdl) {
    const unsigned blocks = (unsigned)((n + 255) / 256 > 148 ? 148 : (n + 255) / 256);
    LaunchCfg lc(dim3(blocks), dim3(256), 0, s, pdl);
    B200_CUDA_OK(cudaLaunchKernelEx(&lc.cfg, add_inplace_kernel, dst, src, n));
    return B200LLM_OK;
}

}  // namespace b200
