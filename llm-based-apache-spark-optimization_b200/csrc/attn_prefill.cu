// attn_prefill.cu -- causal attention of a prefill chunk on the tensor cores (kernel K10 of SURVEY.md section 2.3; the
// TTFT path of the reference's ollama.generate calls, whose prompts are whole Spark error texts and table schemas:
// Flask/app.py:98-107,154-166).
//
// One CTA = 128 consecutive query rows of one lane x one q head.  The rows' K/V rows (this chunk's and every earlier
// one's) are already in the paged cache; a tile of 128 keys is gathered page by page with 2-D TMA (a [16 tokens x 64
// dims] box per page and 64-dim block, 128-byte swizzle), so K lands as the K-major B operand of S = Q K^T and V, with
// the very same bytes, as the MN-major B operand of O = P V (no transposition anywhere).  Both products run on
// tcgen05.mma with the accumulators in TMEM (S: 128 columns, O: head_dim columns):
//   pass 1   S = Q K^T per key tile -> row maxima (causal mask on the diagonal tile)
//   pass 2   S again -> P = exp2(S - max) -> bf16 to shared memory -> O += P V, row sums; O / sum at the end
// Two passes instead of an online rescale of O: the second Q K^T costs 1/3 more tensor work in a kernel that is a few
// per cent of the prefill FLOPs, and O never has to leave TMEM between tiles.
// q and P go through the MMAs as bf16 hi + lo pairs (x = hi + lo to ~16 mantissa bits), like the activations of the
// prefill GEMMs (gemm_tc.cu): with single bf16 the logits of a 2k-token prompt leave the oracle's tolerance.
//   warps 0-3   one query row per thread: q -> operand tiles, softmax on its S row (tcgen05.ld), P tiles, output
//   warp 4      TMEM allocation, TMA gathers, MMA issue (converged warp, elected lane: operands in uniform registers)
#include <cuda.h>
#include <math.h>

#include "common.cuh"
#include "kernels.h"

namespace b200 {

constexpr int kApBM = 128;      // query rows per CTA
constexpr int kApBN = 128;      // keys per tile
constexpr int kApThreads = 160;
constexpr int kApPT = 16;       // tokens per KV page

struct ApParams {
    const float *q;            // fp32 [rows, n_q * hd], RoPE applied
    float *out;                // fp32 [rows, n_q * hd] or NULL
    uint16_t *out_bf16;        // bf16 hi parts [rows, n_q * hd] or NULL; lo parts at + out_lo_off elements (0: none)
    int64_t out_lo_off;
    const int32_t *page_table; // the lane's row
    int32_t n_ptab;            // its length
    int32_t rows, pos0;        // query row r sits at position pos0 + r and attends keys 0 .. pos0 + r
    int32_t n_q, gq;           // q heads, q heads per kv head
    int32_t kv_page_heads;     // (kv head, 16-token) tiles from one page to the next
    float scale_log2;          // log2(e) / sqrt(hd)
};

__device__ __forceinline__ void ap_tma_2d(void *dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
                     smem_u32(dst)),
                 "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
                 : "memory");
}
// 128-byte-swizzled operand tile: rows of 128 bytes, 8-row groups 1024 bytes apart (K-major: rows = M/N index;
// MN-major: rows = K index, the 64 MN elements of a row are the contiguous ones)
__device__ __forceinline__ uint64_t ap_desc(uint32_t addr) {
    uint64_t d = 0;
    d |= (uint64_t)((addr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__device__ __forceinline__ bool ap_elect() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.b32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void ap_mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t accum) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                 "l"(a), "l"(b), "r"(idesc), "r"(accum)
                 : "memory");
}
__device__ __forceinline__ void ap_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 32 consecutive fp32 columns of this thread's TMEM lane
__device__ __forceinline__ void ap_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, "
        "%18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ uint32_t ap_pack(float a, float b) {
    return (uint32_t)f32_to_bf16_bits(a) | ((uint32_t)f32_to_bf16_bits(b) << 16);
}
__device__ __forceinline__ float ap_round(float a) { return __uint_as_float((uint32_t)f32_to_bf16_bits(a) << 16); }

// shared memory: [Q hi][Q lo][K][V][P hi][P lo], each HD/64 (Q, K, V) or 2 (P) blocks of [128 rows x 64] = 16 KB
template <int HD>
__global__ void __launch_bounds__(kApThreads, 1)
attn_prefill_kernel(const __grid_constant__ CUtensorMap map_k, const __grid_constant__ CUtensorMap map_v,
                    const __grid_constant__ ApParams p) {
    constexpr int KB = HD / 64;                 // 64-wide blocks along head_dim
    constexpr int TILE = kApBM * 64 * 2;        // 16 KB
    extern __shared__ __align__(1024) uint8_t smem_dyn[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
    uint8_t *q_hi = smem, *q_lo = q_hi + KB * TILE, *k_sm = q_lo + KB * TILE, *v_sm = k_sm + KB * TILE;
    uint8_t *p_hi = v_sm + KB * TILE, *p_lo = p_hi + 2 * TILE;
    uint64_t *bars = reinterpret_cast<uint64_t *>(p_lo + 2 * TILE);   // [0] K/V landed, [1] S ready, [2] O += P V done
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 4);
    const int tid = threadIdx.x, warp = tid >> 5;
    const int row0 = blockIdx.x * kApBM, head = blockIdx.y, hk = head / p.gq;
    const int n_rows = min(kApBM, p.rows - row0);
    const int n_tiles = (p.pos0 + row0 + n_rows + kApBN - 1) / kApBN;   // key tiles up to the last row's position

    if (tid == 0) {
        mbar_init(&bars[0], 1), mbar_init(&bars[1], 1), mbar_init(&bars[2], 1);
        mbar_fence_init();
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
    }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    pdl_trigger();
    pdl_wait();   // q and this chunk's K/V rows come from the kernel before
    if (warp < 4) {   // q row -> bf16 hi / lo, scaled into log2 units, 128-byte-swizzled K-major tiles
        const int r = tid;
        const float *qp = p.q + (size_t)(row0 + r) * p.n_q * HD + (size_t)head * HD;
#pragma unroll 4
        for (int c = 0; c < HD / 8; ++c) {
            float v[8];
            if (r < n_rows) {
                const float4 a = *reinterpret_cast<const float4 *>(qp + c * 8), b = *reinterpret_cast<const float4 *>(qp + c * 8 + 4);
                v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w, v[4] = b.x, v[5] = b.y, v[6] = b.z, v[7] = b.w;
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = 0.f;
            }
            float hi[8], lo[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] *= p.scale_log2, hi[e] = ap_round(v[e]), lo[e] = v[e] - hi[e];
            const int off = (c >> 3) * TILE + r * 128 + (((c & 7) ^ (r & 7)) << 4);
            *reinterpret_cast<uint4 *>(q_hi + off) = make_uint4(ap_pack(hi[0], hi[1]), ap_pack(hi[2], hi[3]), ap_pack(hi[4], hi[5]), ap_pack(hi[6], hi[7]));
            *reinterpret_cast<uint4 *>(q_lo + off) = make_uint4(ap_pack(lo[0], lo[1]), ap_pack(lo[2], lo[3]), ap_pack(lo[4], lo[5]), ap_pack(lo[6], lo[7]));
        }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    const uint32_t t_s = tmem, t_o = tmem + 128;      // S: columns 0..127, O: 128..128 + HD
    // D fp32, A / B bf16, M = 128: S = Q K^T has both operands K-major (N = 128 keys);
    // O = P V takes V MN-major (bit 16), 64 head dims per instruction
    const uint32_t idesc_s = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kApBN >> 3) << 17) | ((uint32_t)(kApBM >> 4) << 24);
    const uint32_t idesc_o = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(kApBM >> 4) << 24);

    float row_max = -INFINITY, row_sum = 0.f;
    const int my_pos = p.pos0 + row0 + tid;           // (softmax threads) position of this thread's query row
    for (int it = 0; it < 2 * n_tiles; ++it) {
        const int pass = it >= n_tiles, j = pass ? it - n_tiles : it;
        if (warp == 4) {
            // ---- gather the key tile (and in pass 2 the value tile): one box per page and 64-dim block
            if (it > 0) mbar_wait(&bars[pass && j > 0 ? 2 : 1], (uint32_t)((pass && j > 0 ? j - 1 : it - 1) & 1));   // smem of the tile before is free
            if (ap_elect()) {
                mbar_expect_tx(&bars[0], (uint32_t)((pass ? 2 : 1) * KB * TILE));
                for (int pg = 0; pg < kApBN / kApPT; ++pg) {
                    const int pidx = j * (kApBN / kApPT) + pg;
                    const int page = pidx < p.n_ptab ? p.page_table[pidx] : 0;   // (keys past the sequence end are masked)
                    const int grow = (page * p.kv_page_heads + hk) * kApPT;
                    for (int kb = 0; kb < KB; ++kb) {
                        ap_tma_2d(k_sm + kb * TILE + pg * kApPT * 128, &map_k, kb * 64, grow, &bars[0]);
                        if (pass) ap_tma_2d(v_sm + kb * TILE + pg * kApPT * 128, &map_v, kb * 64, grow, &bars[0]);
                    }
                }
            }
            __syncwarp();
            mbar_wait(&bars[0], (uint32_t)(it & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (ap_elect()) {   // S = (Q hi + Q lo) K^T
#pragma unroll
                for (int half = 0; half < 2; ++half)
#pragma unroll
                    for (int kq = 0; kq < HD / 16; ++kq) {
                        const uint32_t koff = (uint32_t)(kq >> 2) * TILE + (uint32_t)(kq & 3) * 32;
                        ap_mma(t_s, ap_desc(smem_u32(half ? q_lo : q_hi) + koff), ap_desc(smem_u32(k_sm) + koff), idesc_s,
                               (half | kq) ? 1u : 0u);
                    }
                ap_commit(&bars[1]);
            }
            __syncwarp();
            if (pass) {   // wait for the P tiles, then O += (P hi + P lo) V
                asm volatile("bar.sync 1, 160;" ::: "memory");
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (ap_elect()) {
#pragma unroll
                    for (int half = 0; half < 2; ++half)
#pragma unroll
                        for (int kq = 0; kq < kApBN / 16; ++kq)     // 16 keys per instruction
#pragma unroll
                            for (int nb = 0; nb < KB; ++nb) {       // 64 head dims per instruction
                                const uint32_t a_off = (uint32_t)(kq >> 2) * TILE + (uint32_t)(kq & 3) * 32;
                                const uint32_t b_off = (uint32_t)nb * TILE + (uint32_t)kq * 16 * 128;
                                ap_mma(t_o + nb * 64, ap_desc(smem_u32(half ? p_lo : p_hi) + a_off), ap_desc(smem_u32(v_sm) + b_off),
                                       idesc_o, (j | half | kq) ? 1u : 0u);
                            }
                    ap_commit(&bars[2]);
                }
                __syncwarp();
            }
        } else {
            // ---- softmax threads: one query row each
            mbar_wait(&bars[1], (uint32_t)(it & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (pass && j > 0) {      // the P tiles of the tile before have been multiplied
                mbar_wait(&bars[2], (uint32_t)((j - 1) & 1));
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            }
            const uint32_t lane_base = t_s + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
            for (int cq = 0; cq < kApBN / 32; ++cq) {
                float sv[32];
                ap_ld32(lane_base + cq * 32, sv);
                const int key0 = j * kApBN + cq * 32;
                if (!pass) {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (key0 + i <= my_pos) row_max = fmaxf(row_max, sv[i]);
                } else {
                    float hi[32], lo[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const float pv = (key0 + i <= my_pos) ? exp2f(sv[i] - row_max) : 0.f;
                        row_sum += pv;
                        hi[i] = ap_round(pv), lo[i] = pv - hi[i];
                    }
#pragma unroll
                    for (int c = 0; c < 4; ++c) {   // 4 chunks of 8 keys: chunk index cq * 4 + c of 16 in the 128-key row
                        const int ch = cq * 4 + c;
                        const int off = (ch >> 3) * TILE + tid * 128 + (((ch & 7) ^ (tid & 7)) << 4);
                        *reinterpret_cast<uint4 *>(p_hi + off) = make_uint4(ap_pack(hi[8 * c], hi[8 * c + 1]), ap_pack(hi[8 * c + 2], hi[8 * c + 3]),
                                                                            ap_pack(hi[8 * c + 4], hi[8 * c + 5]), ap_pack(hi[8 * c + 6], hi[8 * c + 7]));
                        *reinterpret_cast<uint4 *>(p_lo + off) = make_uint4(ap_pack(lo[8 * c], lo[8 * c + 1]), ap_pack(lo[8 * c + 2], lo[8 * c + 3]),
                                                                            ap_pack(lo[8 * c + 4], lo[8 * c + 5]), ap_pack(lo[8 * c + 6], lo[8 * c + 7]));
                    }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            if (pass) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                asm volatile("bar.sync 1, 160;" ::: "memory");   // P tiles complete (and S consumed): the MMA warp goes on
            }
        }
        if (!pass) {
            // pass 1: the MMA warp must not overwrite S (next tile's MMAs) before every row has been read
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            asm volatile("bar.sync 2, 160;" ::: "memory");
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
    }
    // ---- O / row sum -> output
    if (warp < 4) {
        mbar_wait(&bars[2], (uint32_t)((n_tiles - 1) & 1));
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const float inv = 1.f / row_sum;
        const size_t o_at = (size_t)(row0 + tid) * p.n_q * HD + (size_t)head * HD;
#pragma unroll 1
        for (int cq = 0; cq < HD / 32; ++cq) {
            float ov[32];
            ap_ld32(t_o + ((uint32_t)(warp * 32) << 16) + cq * 32, ov);
            if (tid < n_rows) {
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    const float a = ov[i] * inv, b = ov[i + 1] * inv, c = ov[i + 2] * inv, d = ov[i + 3] * inv;
                    if (p.out) *reinterpret_cast<float4 *>(p.out + o_at + cq * 32 + i) = make_float4(a, b, c, d);
                    if (p.out_bf16) {
                        const float ha = ap_round(a), hb = ap_round(b), hc = ap_round(c), hd_ = ap_round(d);
                        *reinterpret_cast<uint2 *>(p.out_bf16 + o_at + cq * 32 + i) = make_uint2(ap_pack(ha, hb), ap_pack(hc, hd_));
                        if (p.out_lo_off)
                            *reinterpret_cast<uint2 *>(p.out_bf16 + p.out_lo_off + o_at + cq * 32 + i) =
                                make_uint2(ap_pack(a - ha, b - hb), ap_pack(c - hc, d - hd_));
                    }
                }
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    __syncthreads();
    if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem) : "memory");
}

int attn_prefill_launch(const void *map_k, const void *map_v, const float *q, float *out, void *out_bf16, int64_t out_lo_off,
                        const int32_t *page_table_row, int n_ptab, int rows, int pos0, int n_q, int n_kv, int head_dim, int kv_page_heads,
                        cudaStream_t stream, bool pdl) {
    B200_REQUIRE(rows >= 1 && pos0 >= 0 && n_kv >= 1 && n_q % n_kv == 0 && (head_dim == 64 || head_dim == 128),
                 "attn_prefill: bad shape (rows %d pos0 %d n_q %d n_kv %d head_dim %d)", rows, pos0, n_q, n_kv, head_dim);
    ApParams p{};
    p.q = q, p.out = out, p.out_bf16 = reinterpret_cast<uint16_t *>(out_bf16), p.out_lo_off = out_lo_off;
    p.page_table = page_table_row, p.n_ptab = n_ptab, p.rows = rows, p.pos0 = pos0, p.n_q = n_q, p.gq = n_q / n_kv;
    p.kv_page_heads = kv_page_heads > 0 ? kv_page_heads : n_kv;
    p.scale_log2 = (float)(1.4426950408889634 / sqrt((double)head_dim));
    const int kb = head_dim / 64;
    const size_t smem = (size_t)(4 * kb + 4) * kApBM * 64 * 2 + 64 + 1024;
    LaunchCfg lc(dim3((rows + kApBM - 1) / kApBM, n_q), dim3(kApThreads), smem, stream, pdl);
    if (head_dim == 128) {
        B200_CUDA_OK(cudaFuncSetAttribute(attn_prefill_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        B200_CUDA_OK(cudaLaunchKernelEx(&lc.cfg, attn_prefill_kernel<128>, *reinterpret_cast<const CUtensorMap *>(map_k),
                                        *reinterpret_cast<const CUtensorMap *>(map_v), p));
    } else {
        B200_CUDA_OK(cudaFuncSetAttribute(attn_prefill_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        B200_CUDA_OK(cudaLaunchKernelEx(&lc.cfg, attn_prefill_kernel<64>, *reinterpret_cast<const CUtensorMap *>(map_k),
                                        *reinterpret_cast<const CUtensorMap *>(map_v), p));
    }
    return B200LLM_OK;
}

}  // namespace b200
