// dequant.cu -- GGUF block formats -> bf16 on the device (the loader of b200llm/gguf_loader.py).
//
// The reference's two models ship as quantised GGUF blobs (duckdb-nsql: Q4_0; llama3.2: Q4_K_M = Q4_K + Q6_K; both with
// F32 norms and sometimes F16 / Q8_0 / BF16 tensors).  The engine computes in bf16, so a checkpoint is expanded once at
// load time: the raw blocks are uploaded as they lie in the file and expanded here instead of in numpy on the host
// (a 7B Q4_0 file is 3.8 GB of blocks against 13.4 GB of bf16).  The arithmetic is ggml's reference de-quantiser, in the
// same order and without fused multiply-adds, so the bf16 result is bit-equal to rounding gguf-py's fp32 output:
//   Q8_0: d * q        Q4_0: d * (q - 8)        Q4_K: (d * sc_j) * q - (dmin * m_j)        Q6_K: ((q - 32) * sc_j) * d
#include <cuda_fp16.h>

#include "common.cuh"
#include "kernels.h"

namespace b200 {

__device__ __forceinline__ float dq_h2f(uint16_t h) { return __half2float(__ushort_as_half(h)); }
__device__ __forceinline__ uint32_t dq_pack(float a, float b) { return (uint32_t)f32_to_bf16_bits(a) | ((uint32_t)f32_to_bf16_bits(b) << 16); }

// one thread per element pair
__global__ void dequant_scalar_kernel(const uint8_t *__restrict__ src, uint32_t *__restrict__ out, int64_t n_pairs, int kind) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pairs; i += (int64_t)gridDim.x * blockDim.x) {
        float a, b;
        if (kind == 0) {          // F32
            a = reinterpret_cast<const float *>(src)[2 * i], b = reinterpret_cast<const float *>(src)[2 * i + 1];
        } else if (kind == 1) {   // F16
            a = dq_h2f(reinterpret_cast<const uint16_t *>(src)[2 * i]), b = dq_h2f(reinterpret_cast<const uint16_t *>(src)[2 * i + 1]);
        } else {                  // BF16
            out[i] = reinterpret_cast<const uint32_t *>(src)[i];
            continue;
        }
        out[i] = dq_pack(a, b);
    }
}

// Q8_0 (34 bytes: fp16 d, 32 int8) and Q4_0 (18 bytes: fp16 d, 16 bytes of nibbles; elements 0..15 = low nibbles, 16..31 = high):
// 16 threads per block of 32 elements, one output pair each
__global__ void dequant_q0_kernel(const uint8_t *__restrict__ src, uint32_t *__restrict__ out, int64_t n_blocks, int q4) {
    const int bb = q4 ? 18 : 34;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n_blocks * 16; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t blk = t >> 4;
        const int j = (int)(t & 15);          // elements 2j, 2j + 1
        const uint8_t *p = src + blk * bb;
        const float d = dq_h2f((uint16_t)(p[0] | (p[1] << 8)));
        float a, b;
        if (q4) {
            const int e0 = 2 * j, e1 = 2 * j + 1;
            const int q0 = e0 < 16 ? (p[2 + e0] & 0xF) : (p[2 + e0 - 16] >> 4);
            const int q1 = e1 < 16 ? (p[2 + e1] & 0xF) : (p[2 + e1 - 16] >> 4);
            a = __fmul_rn((float)(q0 - 8), d), b = __fmul_rn((float)(q1 - 8), d);
        } else {
            a = __fmul_rn((float)(int8_t)p[2 + 2 * j], d), b = __fmul_rn((float)(int8_t)p[2 + 2 * j + 1], d);
        }
        out[blk * 16 + j] = dq_pack(a, b);
    }
}

// Q4_K: super-block of 256 = 144 bytes (fp16 d, fp16 dmin, 12 bytes of 6-bit (scale, min) pairs, 128 bytes of nibbles).
// One warp per super-block, 8 consecutive elements per lane (one 16-byte store).
__global__ void dequant_q4k_kernel(const uint8_t *__restrict__ src, uint4 *__restrict__ out, int64_t n_blocks) {
    const int lane = threadIdx.x & 31;
    for (int64_t blk = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; blk < n_blocks; blk += ((int64_t)gridDim.x * blockDim.x) >> 5) {
        const uint8_t *p = src + blk * 144;
        const float d = dq_h2f((uint16_t)(p[0] | (p[1] << 8))), dmin = dq_h2f((uint16_t)(p[2] | (p[3] << 8)));
        const uint8_t *sc = p + 4, *qs = p + 16;
        const int j = lane >> 2, o = (lane & 3) * 8;   // sub-block of 32, offset inside it
        int s_, m_;
        if (j < 4) {
            s_ = sc[j] & 63, m_ = sc[j + 4] & 63;
        } else {
            s_ = (sc[j + 4] & 0x0F) | ((sc[j - 4] >> 6) << 4);
            m_ = (sc[j + 4] >> 4) | ((sc[j] >> 6) << 4);
        }
        const float scale = __fmul_rn((float)s_, d), mn = __fmul_rn((float)m_, dmin);
        const uint8_t *q = qs + 32 * (j >> 1) + o;
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int n = (j & 1) ? (q[e] >> 4) : (q[e] & 0x0F);
            v[e] = __fsub_rn(__fmul_rn((float)n, scale), mn);
        }
        out[blk * 32 + lane] = make_uint4(dq_pack(v[0], v[1]), dq_pack(v[2], v[3]), dq_pack(v[4], v[5]), dq_pack(v[6], v[7]));
    }
}

// Q6_K: super-block of 256 = 210 bytes (128 bytes ql, 64 bytes qh, 16 int8 scales, fp16 d).  Element e of half h (128
// elements): group g = e / 32, l = e % 32: q = (nibble of ql[h][(g & 1) * 32 + l], low for g < 2) | ((qh[h][l] >> 2 g) & 3) << 4,
// scale index 8 h + 2 g + l / 16.
__global__ void dequant_q6k_kernel(const uint8_t *__restrict__ src, uint4 *__restrict__ out, int64_t n_blocks) {
    const int lane = threadIdx.x & 31;
    for (int64_t blk = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; blk < n_blocks; blk += ((int64_t)gridDim.x * blockDim.x) >> 5) {
        const uint8_t *p = src + blk * 210;
        const uint8_t *ql = p, *qh = p + 128;
        const int8_t *sc = reinterpret_cast<const int8_t *>(p + 192);
        const float d = dq_h2f((uint16_t)(p[208] | (p[209] << 8)));
        const int h = lane >> 4, g = (lane & 15) >> 2, l0 = (lane & 3) * 8;
        const float s = (float)sc[8 * h + 2 * g + (l0 >> 4)];
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int l = l0 + e;
            const int L = ql[64 * h + (g & 1) * 32 + l];
            const int n = ((g < 2 ? (L & 0x0F) : (L >> 4)) | (((qh[32 * h + l] >> (2 * g)) & 3) << 4)) - 32;
            v[e] = __fmul_rn(__fmul_rn((float)n, s), d);
        }
        out[blk * 32 + lane] = make_uint4(dq_pack(v[0], v[1]), dq_pack(v[2], v[3]), dq_pack(v[4], v[5]), dq_pack(v[6], v[7]));
    }
}

// ggml type ids: 0 F32, 1 F16, 2 Q4_0, 8 Q8_0, 12 Q4_K, 14 Q6_K, 30 BF16
int dequant_bf16_launch(const void *blocks, int ggml_type, int64_t n, void *out, cudaStream_t st) {
    B200_REQUIRE(blocks && out && n > 0, "dequant: NULL argument or no elements");
    const int threads = 256;
    auto grid_for = [&](int64_t work) { return (int)((work + threads - 1) / threads > 148 * 16 ? 148 * 16 : (work + threads - 1) / threads); };
    const uint8_t *src = reinterpret_cast<const uint8_t *>(blocks);
    switch (ggml_type) {
    case 0: case 1: case 30:
        B200_REQUIRE(n % 2 == 0, "dequant: odd element count");
        dequant_scalar_kernel<<<grid_for(n / 2), threads, 0, st>>>(src, reinterpret_cast<uint32_t *>(out), n / 2, ggml_type == 0 ? 0 : (ggml_type == 1 ? 1 : 2));
        break;
    case 2: case 8:
        B200_REQUIRE(n % 32 == 0, "dequant: Q4_0 / Q8_0 tensors are whole blocks of 32");
        dequant_q0_kernel<<<grid_for(n / 2), threads, 0, st>>>(src, reinterpret_cast<uint32_t *>(out), n / 32, ggml_type == 2 ? 1 : 0);
        break;
    case 12:
        B200_REQUIRE(n % 256 == 0, "dequant: Q4_K tensors are whole super-blocks of 256");
        dequant_q4k_kernel<<<grid_for(n / 8), threads, 0, st>>>(src, reinterpret_cast<uint4 *>(out), n / 256);
        break;
    case 14:
        B200_REQUIRE(n % 256 == 0, "dequant: Q6_K tensors are whole super-blocks of 256");
        dequant_q6k_kernel<<<grid_for(n / 8), threads, 0, st>>>(src, reinterpret_cast<uint4 *>(out), n / 256);
        break;
    default:
        set_error("dequant: ggml type %d is not supported (F32, F16, BF16, Q4_0, Q8_0, Q4_K, Q6_K are)", ggml_type);
        return B200LLM_EINVAL;
    }
    B200_CUDA_OK(cudaGetLastError());
    return B200LLM_OK;
}

}  // namespace b200
