// batch_ops.cu -- the small kernels around the tcgen05 GEMM on the batch / prefill path:
// RMSNorm -> bf16, fp32 -> bf16 cast, RoPE + paged-KV append, SwiGLU -> bf16.  All are
// one-pass, vectorised, and tiny next to the weight stream (activations are [rows, d]).
#include "common.cuh"
#include "kernels.h"

namespace b200 {

// bf16 hi part of v and the bf16 lo part (v - hi): together ~fp32 accurate as a GEMM operand pair
__device__ __forceinline__ void split_bf16(float v, uint16_t &hi, uint16_t &lo) {
    hi = f32_to_bf16_bits(v);
    lo = f32_to_bf16_bits(v - __uint_as_float((uint32_t)hi << 16));
}

// x fp32 [rows, d] -> bf16 [rows, d] = gain * (x * rsqrt(mean(x^2) + eps)); one CTA per row
// Also zeroes `zero_w` floats of row `row` of a second buffer (the fp32 output the next, k-split
// GEMM accumulates into), which saves a memset node between the two kernels.
__global__ void __launch_bounds__(256) rmsnorm_bf16_kernel(const float *x, const uint16_t *gain, uint16_t *out, int d,
                                                           float eps, float *zero, int zero_w, int64_t lo_off) {
    __shared__ float red[8];
    pdl_trigger();
    pdl_wait();
    const int row = blockIdx.x;
    if (zero != nullptr)
        for (int k = threadIdx.x * 4; k < zero_w; k += blockDim.x * 4)
            *reinterpret_cast<float4 *>(zero + (size_t)row * zero_w + k) = make_float4(0.f, 0.f, 0.f, 0.f);
    const float *xr = x + (size_t)row * d;
    float ss = 0.f;
    for (int k = threadIdx.x * 4; k < d; k += blockDim.x * 4) {
        const float4 v = *reinterpret_cast<const float4 *>(xr + k);
        ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) tot += red[w];
    const float sc = rsqrtf(tot / (float)d + eps);
    uint16_t *o = out + (size_t)row * d;
    for (int k = threadIdx.x * 4; k < d; k += blockDim.x * 4) {
        const float4 v = *reinterpret_cast<const float4 *>(xr + k);
        const uint2 g = *reinterpret_cast<const uint2 *>(gain + k);
        const float y[4] = {bf16lo(g.x) * (v.x * sc), bf16hi(g.x) * (v.y * sc), bf16lo(g.y) * (v.z * sc), bf16hi(g.y) * (v.w * sc)};
        uint16_t hi[4], lo[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) split_bf16(y[e], hi[e], lo[e]);
        *reinterpret_cast<uint2 *>(o + k) = make_uint2((uint32_t)hi[0] | ((uint32_t)hi[1] << 16), (uint32_t)hi[2] | ((uint32_t)hi[3] << 16));
        if (lo_off != 0)
            *reinterpret_cast<uint2 *>(o + lo_off + k) = make_uint2((uint32_t)lo[0] | ((uint32_t)lo[1] << 16), (uint32_t)lo[2] | ((uint32_t)lo[3] << 16));
    }
}

// fp32 [rows, n] (row stride in_stride) -> bf16 [rows, n]
__global__ void __launch_bounds__(256) cast_bf16_kernel(const float *x, int64_t in_stride, uint16_t *out, int n, int64_t lo_off) {
    pdl_trigger();
    pdl_wait();
    const int row = blockIdx.y;
    for (int k = (blockIdx.x * blockDim.x + threadIdx.x) * 4; k < n; k += gridDim.x * blockDim.x * 4) {
        const float4 v = *reinterpret_cast<const float4 *>(x + (size_t)row * in_stride + k);
        const float y[4] = {v.x, v.y, v.z, v.w};
        uint16_t hi[4], lo[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) split_bf16(y[e], hi[e], lo[e]);
        uint16_t *o = out + (size_t)row * n + k;
        *reinterpret_cast<uint2 *>(o) = make_uint2((uint32_t)hi[0] | ((uint32_t)hi[1] << 16), (uint32_t)hi[2] | ((uint32_t)hi[3] << 16));
        if (lo_off != 0)
            *reinterpret_cast<uint2 *>(o + lo_off) = make_uint2((uint32_t)lo[0] | ((uint32_t)lo[1] << 16), (uint32_t)lo[2] | ((uint32_t)lo[3] << 16));
    }
}

// gate/up fp32 [rows, 2 ffn] with interleaved rows (2i = gate_i, 2i+1 = up_i) -> bf16 [rows, ffn]
__global__ void __launch_bounds__(256) swiglu_bf16_kernel(const float *gu, uint16_t *out, int ffn, int64_t lo_off) {
    pdl_trigger();
    pdl_wait();
    const int row = blockIdx.y;
    const float *g = gu + (size_t)row * 2 * ffn;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) * 2; i < ffn; i += gridDim.x * blockDim.x * 2) {
        const float4 v = *reinterpret_cast<const float4 *>(g + 2 * i);  // gate_i, up_i, gate_i+1, up_i+1
        const float a = (v.x / (1.f + expf(-v.x))) * v.y;
        const float b = (v.z / (1.f + expf(-v.z))) * v.w;
        uint16_t ah, al, bh, bl;
        split_bf16(a, ah, al), split_bf16(b, bh, bl);
        *reinterpret_cast<uint32_t *>(out + (size_t)row * ffn + i) = (uint32_t)ah | ((uint32_t)bh << 16);
        if (lo_off != 0) *reinterpret_cast<uint32_t *>(out + lo_off + (size_t)row * ffn + i) = (uint32_t)al | ((uint32_t)bl << 16);
    }
}

// qkv fp32 [rows, (n_q + 2 n_kv) * hd] (q/k rows pair-interleaved) -> RoPE -> q fp32 [rows, n_q*hd],
// k, v bf16 into the lane's KV page of position row_pos[row]
struct RopeArgs {
    const float *qkv;
    float *q;
    void *k_pages, *v_pages;
    const float *inv_freq;
    const int32_t *row_lane, *row_pos, *page_table;
    int n_q, n_kv, hd, page_tokens, max_pages, kv_page_heads;
};
__global__ void __launch_bounds__(256) rope_append_kernel(const __grid_constant__ RopeArgs a) {
    pdl_trigger();
    pdl_wait();
    const int row = blockIdx.y;
    const int width = (a.n_q + 2 * a.n_kv) * a.hd;
    const int qrows = a.n_q * a.hd, krows = a.n_kv * a.hd;
    const int pos = a.row_pos[row];
    const int page = a.page_table[(size_t)a.row_lane[row] * a.max_pages + pos / a.page_tokens];
    const int slot = pos % a.page_tokens;
    const float *src = a.qkv + (size_t)row * width;
    for (int n = (blockIdx.x * blockDim.x + threadIdx.x) * 2; n < width; n += gridDim.x * blockDim.x * 2) {
        float v0 = src[n], v1 = src[n + 1];
        if (n < qrows + krows) {
            float sn, cs;
            sincosf((float)pos * a.inv_freq[(n % a.hd) >> 1], &sn, &cs);
            const float o0 = v0 * cs - v1 * sn, o1 = v1 * cs + v0 * sn;
            v0 = o0, v1 = o1;
        }
        if (n < qrows) {
            *reinterpret_cast<float2 *>(a.q + (size_t)row * qrows + n) = make_float2(v0, v1);
        } else {
            const bool is_k = n < qrows + krows;
            const int kn = is_k ? n - qrows : n - qrows - krows;
            uint16_t *dst = reinterpret_cast<uint16_t *>(is_k ? a.k_pages : a.v_pages) +
                            (((size_t)page * a.kv_page_heads + kn / a.hd) * a.page_tokens + slot) * a.hd + kn % a.hd;
            *reinterpret_cast<uint32_t *>(dst) = (uint32_t)f32_to_bf16_bits(v0) | ((uint32_t)f32_to_bf16_bits(v1) << 16);
        }
    }
}

int rmsnorm_bf16_launch(const float *x, const void *gain, void *out, int rows, int d, float eps, cudaStream_t s, bool pdl,
                        float *zero, int zero_w, int64_t lo_off) {
    B200_REQUIRE(d % 4 == 0 && rows >= 1 && zero_w % 4 == 0, "rmsnorm_bf16: bad shape");
    LaunchCfg lc(dim3(rows), dim3(256), 0, s, pdl);
    B200_CUDA_OK(cudaLaunchKernelEx(&lc.cfg, rmsnorm_bf16_kernel, x, reinterpret_cast<const uint16_t *>(gain),
                                    reinterpret_cast<uint16_t *>(out), d, eps, zero, zero_w, lo_off));
    return B200LLM_OK;
}
int cast_bf16_launch(const float *x, int64_t in_stride, void *out, int rows, int n, cudaStream_t s, bool pdl, int64_t lo_off) {
    B200_REQUIRE(n % 4 == 0 && rows >= 1, "cast_bf16: bad shape");
    LaunchCfg lc(dim3((n / 4 + 255) / 256, rows), dim3(256), 0, s, pdl);
    B200_CUDA_OK(cudaLaunchKernelEx(&lc.cfg, cast_bf16_kernel, x, in_stride, reinterpret_cast<uint16_t *>(out), n, lo_off));
    return B200LLM_OK;
}
int swiglu_bf16_launch(const float *gu, void *out, int rows, int ffn, cudaStream_t s, bool pdl, int64_t lo_off) {
    B200_REQUIRE(ffn % 2 == 0 && rows >= 1, "swiglu_bf16: bad shape");
    LaunchCfg lc(dim3((ffn / 2 + 255) / 256, rows), dim3(256), 0, s, pdl);
    B200_CUDA_OK(cudaLaunchKernelEx(&lc.cfg, swiglu_bf16_kernel, gu, reinterpret_cast<uint16_t *>(out), ffn, lo_off));
    return B200LLM_OK;
}
int rope_append_launch(const float *qkv, float *q, void *k_pages, void *v_pages, const float *inv_freq,
                       const int32_t *row_lane, const int32_t *row_pos, const int32_t *page_table, int rows, int n_q,
                       int n_kv, int hd, int page_tokens, int max_pages, cudaStream_t s, bool pdl, int kv_page_heads) {
    RopeArgs a{qkv, q, k_pages, v_pages, inv_freq, row_lane, row_pos, page_table, n_q, n_kv, hd, page_tokens, max_pages,
               kv_page_heads > 0 ? kv_page_heads : n_kv};
    const int width = (n_q + 2 * n_kv) * hd;
    LaunchCfg lc(dim3((width / 2 + 255) / 256, rows), dim3(256), 0, s, pdl);
    B200_CUDA_OK(cudaLaunchKernelEx(&lc.cfg, rope_append_kernel, a));
    return B200LLM_OK;
}

}  // namespace b200
