// gemm_tc.cu -- tcgen05 / TMEM weight-streaming GEMM for batches > 4 rows and for prefill
// (kernel K9/K10 of SURVEY.md section 2.3):   out[t, n] (+)= sum_k X[t, k] * W[n, k]
//
// Swap-AB: the weight matrix feeds the M side of the MMA (128 weight rows per tile), the
// tokens are the N side (16..256), so one CTA streams a [128 x K/splits] slab of W exactly
// once while the small activation tile is re-read from L2.
//   warp 0   : TMA producer   (cp.async.bulk.tensor.2d, 128B-swizzled [128 x 64] W tiles and
//                              [BN x 64] X tiles into a multi-stage shared-memory ring)
//   warp 1   : MMA issuer     (one thread: tcgen05.mma.cta_group::1.kind::f16, M=128, N=BN,
//                              K=16, bf16 x bf16 -> fp32 in TMEM; tcgen05.commit frees stages)
//   warp 2   : TMEM allocator
//   warps 4-7: epilogue       (tcgen05.ld 32x32b of the warp's 32 accumulator lanes = weight
//                              rows; registers = tokens; coalesced fp32 store / red.add)
// grid = (N/128 row tiles, k_splits): small matrices are split along K so that ~all SMs stream
// (decode at batch 64 is still HBM-bound); k-split partial sums meet in the output with fp32
// red.global.add (the output is zeroed first, or already holds the residual stream).
#include <cuda.h>

#include "common.cuh"
#include "kernels.h"

namespace b200 {

constexpr int kTcBM = 128;   // weight rows per tile (UMMA M)
constexpr int kTcBK = 64;    // k per stage: 64 bf16 = one 128-byte swizzle row
constexpr int kTcThreads = 256;

struct TcParams {
    float *out;          // fp32 [T, out_stride]
    int64_t out_stride;
    int32_t T;           // valid tokens (rows of X)
    int32_t N, K;
    int32_t bn;          // token tile = UMMA N (multiple of 16, >= T)
    int32_t k_per_split; // multiple of 64
    int32_t stages;
    int32_t accumulate;  // 1: out += (red.add), 0: plain store (only valid with one k split)
    int32_t tmem_cols;   // power of two >= max(32, bn)
    int32_t has_lo;       // 1: a second X map holds the bf16 lo parts (x - hi); both are multiplied, ~fp32-accurate activations
    int64_t swiglu_lo_off; // != 0: lo parts of the fused SwiGLU output go to out_swiglu + swiglu_lo_off
    int32_t bk;           // k per stage: 64 (128-byte swizzle rows) or 32 (64-byte swizzle rows: half the stage, twice the stages)
    int32_t cs;           // CTAs per cluster (1, 2 or 4): neighbouring weight tiles that share every activation tile
    int32_t w_early;      // 1: the first ring of weight tiles is requested before griddepcontrol.wait
    const uint8_t *w_tm;  // != NULL (bk 64): the TILE-MAJOR copy of W ([N/128][K/64] tiles of 16 KB, stored as the swizzled operand,
                          // tcd.h): a weight tile is then ONE contiguous 16 KB bulk copy instead of a 2-D box of 128 rows of 128 bytes
    uint16_t *out_swiglu; // when set (needs one k split): rows (2i, 2i+1) = (gate_i, up_i) -> bf16 silu(gate) * up at
                          // out_swiglu[t, i], row stride N / 2, instead of the fp32 output
};

__device__ __forceinline__ void tma_load_2d(void *smem_dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(smem_dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}

// The same box, delivered to the same shared-memory offset of every CTA of the cluster in `mask` (and counted on each one's
// barrier at the same offset).  The source is always THIS GPU's memory: the activation buffers of the engine's own
// workspace (TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another
// GPU's memory hung the GPU).
__device__ __forceinline__ void tma_load_2d_mc(void *smem_dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar, uint16_t mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;" ::
            "r"(smem_u32(smem_dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask)
        : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// K-major operand tile in shared memory, 128-byte swizzle, rows of 128 bytes, 8-row groups 1024 B apart
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr, int bk = 64) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);          // start address
    d |= (uint64_t)1 << 16;                               // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)((bk == 64 ? 1024 : 512) >> 4) << 32;  // stride byte offset between 8-row groups (rows of 2 bk bytes)
    d |= (uint64_t)1 << 46;                               // descriptor version (sm_100)
    d |= (uint64_t)(bk == 64 ? 2 : 4) << 61;              // SWIZZLE_128B / SWIZZLE_64B
    return d;
}

__device__ __forceinline__ uint32_t umma_idesc_bf16(int m, int n) {
    return (1u << 4)      // D format f32
           | (1u << 7)    // A format bf16
           | (1u << 10)   // B format bf16
           | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);   // both operands K-major
}

__global__ void __launch_bounds__(kTcThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_x,
               const __grid_constant__ CUtensorMap map_xl, const __grid_constant__ TcParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int a_bytes = kTcBM * p.bk * 2;           // 16 KB (8 KB at bk 32)
    const int b_bytes = p.bn * p.bk * 2;            // multiple of 1 KB (bn % 16 == 0)
    const int b_pad = (b_bytes + 1023) & ~1023;
    const int stage_bytes = a_bytes + (p.has_lo ? 2 : 1) * b_pad;
    uint8_t *tiles = smem;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)p.stages * stage_bytes);
    uint64_t *empty = full + p.stages;
    uint64_t *tmem_full = empty + p.stages;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tmem_full + 1);

    const int row0 = blockIdx.x * kTcBM;
    const int k0 = blockIdx.y * p.k_per_split;
    const int n_k = min(p.k_per_split, p.K - k0) / p.bk;

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], (uint32_t)p.cs);   // every CTA of the cluster writes a slice of the stage's activation tile
        }
        mbar_init(tmem_full, 1);
        mbar_fence_init();
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
        if (p.has_lo) asm volatile("prefetch.tensormap [%0];" ::"l"(&map_xl) : "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"((uint32_t)p.tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (p.cs > 1) cluster_sync_all();   // the peers' barriers exist before anything is sent to them
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    const uint32_t crank = p.cs > 1 ? cluster_ctarank() : 0u;
    const uint16_t cmask = (uint16_t)((1u << p.cs) - 1u);
    pdl_trigger();
    // The weights do not depend on the previous kernel: the first ring of weight tiles is requested BEFORE the wait, so a
    // CTA that became resident while its predecessor still runs (programmatic dependent launch) has its HBM stream under
    // way when the activations arrive.  The stage barrier expects the whole stage; the activation tiles follow the wait.
    const int n_pre = p.w_early ? min(n_k, p.stages) : 0;
    // tile-major weights: this CTA's tiles (row tile blockIdx.x, k blocks k0 / 64 ...) are consecutive 16 KB pieces
    const uint8_t *w_tile0 = p.w_tm ? p.w_tm + ((size_t)blockIdx.x * (size_t)(p.K / kTcBK) + (size_t)(k0 / kTcBK)) * (size_t)a_bytes : nullptr;
    const uint64_t w_policy = l2_policy_evict_first();
    if (warp == 0 && lane == 0) {
        for (int i = 0; i < n_pre; ++i) {
            mbar_expect_tx(&full[i], (uint32_t)(a_bytes + (p.has_lo ? 2 : 1) * b_bytes));
            if (p.w_tm)
                tma_load_1d_hint(tiles + (size_t)i * stage_bytes, w_tile0 + (size_t)i * a_bytes, (uint32_t)a_bytes, &full[i], w_policy);
            else
                tma_load_2d(tiles + (size_t)i * stage_bytes, &map_w, k0 + i * p.bk, row0, &full[i]);
        }
    }
    pdl_wait();  // X (and the residual in `out`) come from the previous kernel

    if (warp == 0) {
        if (lane == 0) {
            for (int i = 0; i < n_k; ++i) {
                const int s = i % p.stages, use = i / p.stages;
                if (use > 0) mbar_wait(&empty[s], (use - 1) & 1);
                uint8_t *a_dst = tiles + (size_t)s * stage_bytes;
                if (i >= n_pre) {
                    mbar_expect_tx(&full[s], (uint32_t)(a_bytes + (p.has_lo ? 2 : 1) * b_bytes));
                    if (p.w_tm)
                        tma_load_1d_hint(a_dst, w_tile0 + (size_t)i * a_bytes, (uint32_t)a_bytes, &full[s], w_policy);
                    else
                        tma_load_2d(a_dst, &map_w, k0 + i * p.bk, row0, &full[s]);
                }
                if (p.cs == 1) {
                    tma_load_2d(a_dst + a_bytes, &map_x, k0 + i * p.bk, 0, &full[s]);
                    if (p.has_lo) tma_load_2d(a_dst + a_bytes + b_pad, &map_xl, k0 + i * p.bk, 0, &full[s]);
                } else {
                    // this CTA fetches rows [rank * bn / cs, +bn / cs) of the activation tile (the maps' box) for the whole cluster
                    const int r0 = (int)crank * (p.bn / p.cs);
                    const int soff = r0 * p.bk * 2;
                    tma_load_2d_mc(a_dst + a_bytes + soff, &map_x, k0 + i * p.bk, r0, &full[s], cmask);
                    if (p.has_lo) tma_load_2d_mc(a_dst + a_bytes + b_pad + soff, &map_xl, k0 + i * p.bk, r0, &full[s], cmask);
                }
            }
        }
    } else if (warp == 1) {
        // MMA issue: the whole warp walks the loop, converged, one elected lane issues -- descriptors and addresses are then
        // warp-uniform and stay in uniform registers (issued from lane 0 of a diverged warp, every tcgen05.mma sat in an
        // ELECT / R2UR.BROADCAST / BRA.U.ANY loop of ~145 clk, more than a 128 x 64..128 MMA takes to execute)
        {
            // hi and lo tiles of a stage are contiguous (bn * 2 bk bytes is a multiple of 1 KB), so up to bn = 128 they are ONE
            // B operand of N = 2 bn: columns [0, bn) = W x hi, [bn, 2 bn) = W x lo, added in the epilogue
            const bool stacked = p.has_lo && p.bn <= 128;
            const uint32_t idesc = umma_idesc_bf16(kTcBM, stacked ? 2 * p.bn : p.bn);
            const uint32_t tiles_a = smem_u32(tiles);
            for (int i = 0; i < n_k; ++i) {
                const int s = i % p.stages, use = i / p.stages;
                mbar_wait(&full[s], use & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t a_addr = tiles_a + (uint32_t)s * (uint32_t)stage_bytes;
                const uint64_t adesc = umma_desc_sw128(a_addr, p.bk);
                const uint64_t bdesc = umma_desc_sw128(a_addr + a_bytes, p.bk);
                const uint64_t bdesc_lo = umma_desc_sw128(a_addr + a_bytes + b_pad, p.bk);
                uint32_t el;
                asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.b32 %0, 1, 0, P;\n\t}" : "=r"(el));
                if (el) {
                    for (int k = 0; k < p.bk / 16; ++k) {
                        const uint32_t accum = (i > 0 || k > 0) ? 1u : 0u;
                        // advancing 16 k = 32 bytes inside the swizzle row: +2 in the address field
                        asm volatile(
                            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
                            "l"(adesc + (uint64_t)(2 * k)), "l"(bdesc + (uint64_t)(2 * k)), "r"(idesc), "r"(accum)
                            : "memory");
                        if (p.has_lo && !stacked)  // W x lo(X) into the same accumulator
                            asm volatile(
                                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
                                "l"(adesc + (uint64_t)(2 * k)), "l"(bdesc_lo + (uint64_t)(2 * k)), "r"(idesc), "r"(1u)
                                : "memory");
                    }
                    // frees the stage once the MMAs above have read it -- in every CTA of the cluster, whose producers write into it
                    if (p.cs == 1)
                        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                                         smem_u32(&empty[s]))
                                     : "memory");
                    else
                        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                                         smem_u32(&empty[s])),
                                     "h"(cmask)
                                     : "memory");
                }
                __syncwarp();
            }
            if (lane == 0)
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                                 smem_u32(tmem_full))
                             : "memory");
            __syncwarp();
        }
    } else if (warp >= 4) {
        // ===== epilogue: this warp owns accumulator lanes 32*(warp%4) .. +31 = 32 weight rows =====
        mbar_wait(tmem_full, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int q = warp & 3;
        const int n = row0 + q * 32 + lane;  // output column (weight row) of this thread
        for (int c0 = 0; c0 < p.bn; c0 += 16) {
            uint32_t v[16];
            const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)c0;
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, "
                "%14, %15}, [%16];"
                : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                  "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (p.has_lo && p.bn <= 128) {  // stacked hi | lo accumulator halves
                uint32_t w[16];
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, "
                    "%14, %15}, [%16];"
                    : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]),
                      "=r"(w[8]), "=r"(w[9]), "=r"(w[10]), "=r"(w[11]), "=r"(w[12]), "=r"(w[13]), "=r"(w[14]), "=r"(w[15])
                    : "r"(taddr + (uint32_t)p.bn));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(w[j]));
            }
            if (p.out_swiglu != nullptr) {
                // fused SwiGLU: gate and up of one unit sit in adjacent accumulator lanes = adjacent threads
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const int t = c0 + j;
                    const float mine = __uint_as_float(v[j]);
                    const float other = __shfl_xor_sync(0xffffffffu, mine, 1);
                    if ((lane & 1) == 0 && t < p.T && n < p.N) {
                        const float y = (mine / (1.f + expf(-mine))) * other;
                        const uint16_t yh = f32_to_bf16_bits(y);
                        const size_t at = (size_t)t * (p.N >> 1) + (n >> 1);
                        p.out_swiglu[at] = yh;
                        if (p.swiglu_lo_off != 0)
                            p.out_swiglu[p.swiglu_lo_off + at] = f32_to_bf16_bits(y - __uint_as_float((uint32_t)yh << 16));
                    }
                }
                continue;
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int t = c0 + j;
                if (t < p.T && n < p.N) {
                    float *dst = p.out + (size_t)t * p.out_stride + n;
                    if (p.accumulate)
                        atomicAdd(dst, __uint_as_float(v[j]));
                    else
                        *dst = __uint_as_float(v[j]);
                }
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    __syncthreads();
    if (p.cs > 1) cluster_sync_all();   // nobody leaves while a peer may still signal its barriers
    if (warp == 2) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)p.tmem_cols)
                     : "memory");
    }
}

// ---------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;

static int load_encode() {
    if (g_encode) return B200LLM_OK;
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    B200_CUDA_OK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (!fn || qres != cudaDriverEntryPointSuccess) {
        set_error("cuTensorMapEncodeTiled is not available from the driver");
        return B200LLM_ECUDA;
    }
    g_encode = (EncodeTiledFn)fn;
    return B200LLM_OK;
}

// bf16 row-major [rows, cols] -> 2-D map with a [box_rows x 64] box, 128-byte swizzle, zero fill out of bounds
int tc_make_map(void *map_out, const void *base, int64_t rows, int64_t cols, int box_rows, int bk) {
    int rc = load_encode();
    if (rc) return rc;
    B200_REQUIRE(cols % 8 == 0 && ((uintptr_t)base % 16) == 0 && box_rows >= 1 && box_rows <= 256,
                 "tensor map: bad shape/alignment");
    cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t gstr[1] = {(cuuint64_t)cols * 2};
    B200_REQUIRE(bk == 64 || bk == 32, "tensor map: k per box must be 64 or 32");
    cuuint32_t box[2] = {(cuuint32_t)bk, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(reinterpret_cast<CUtensorMap *>(map_out), CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                          const_cast<void *>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          bk == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with %d (rows %lld cols %lld box %d)", (int)r, (long long)rows,
                  (long long)cols, box_rows);
        return B200LLM_ECUDA;
    }
    return B200LLM_OK;
}

// CTAs per cluster for an [N, K] x [T, K] GEMM: neighbouring weight tiles share the activation tile, which each CTA fetches a
// slice of and multicasts (the activation tile is re-read per weight tile: at 256 rows with hi + lo operands it is 4x the
// weight bytes of a stage).  The activation maps' box must be tc_x_box(N, T) rows.  OFF by default (B200LLM_TC_CLUSTER=2|4
// turns it on): measured on the prefill of 296 / 2072 ids and on batch-64 decode, clusters of 2 are level and clusters of
// 4 are 10-15 % slower -- L2 reads of the activation tile are not what bounds this kernel (profiles/prefill_r2.txt).
int tc_cluster_size(int N, int T) {
    static const int forced = getenv("B200LLM_TC_CLUSTER") ? atoi(getenv("B200LLM_TC_CLUSTER")) : 1;
    const int tiles = N / kTcBM, bn = (T + 15) / 16 * 16;
    for (int cs = forced; cs > 1; cs >>= 1)
        if (tiles % cs == 0 && (bn / cs) % 8 == 0 && bn % cs == 0) return cs;
    return 1;
}
int tc_x_box(int N, int T) { return (T + 15) / 16 * 16 / tc_cluster_size(N, T); }
// k per stage: 64; B200LLM_TC_BK=32 halves the stage (64-byte swizzle rows) for twice the stages -- 256 rows with hi + lo
// operands are 80 KB per 64-wide stage, i.e. two stages.  Measured level to 8 % slower, so the pipeline depth is not
// the limit either: ncu shows the tensor pipe active throughout at ~36 % of the cuBLAS rate (profiles/prefill_r2.txt).
int tc_bk(int T, bool has_lo) {
    static const int forced = getenv("B200LLM_TC_BK") ? atoi(getenv("B200LLM_TC_BK")) : 64;
    (void)T, (void)has_lo;
    return forced == 32 ? 32 : 64;
}

int tc_gemm_launch(const void *map_w, const void *map_x, float *out, int64_t out_stride, int T, int N, int K,
                   int accumulate, cudaStream_t stream, bool pdl, void *out_swiglu, const void *map_x_lo,
                   int64_t swiglu_lo_off, int bk, const void *w_tile_major) {
    B200_REQUIRE(bk == 64 || bk == 32, "gemm_tc: bk must be 64 or 32");
    B200_REQUIRE(w_tile_major == nullptr || (bk == 64 && ((uintptr_t)w_tile_major % 16) == 0), "gemm_tc: tile-major weights need bk 64 and 16-byte alignment");
    B200_REQUIRE(T >= 1 && T <= 256 && N % kTcBM == 0 && K % kTcBK == 0, "gemm_tc: need 1 <= T <= 256, N %% 128 == 0, K %% 64 == 0 (T %d N %d K %d)", T, N, K);
    TcParams p{};
    p.out = out, p.out_stride = out_stride, p.T = T, p.N = N, p.K = K;
    p.bn = (T + 15) / 16 * 16;
    const int tiles = N / kTcBM;
    int ks = kNumSMs / tiles;
    const int k_iters = K / kTcBK;   // (k splits stay multiples of 64)
    p.bk = bk;
    if (ks < 1) ks = 1;
    if (ks > k_iters) ks = k_iters;
    if (!accumulate || out_swiglu != nullptr) ks = 1;  // a plain store / the fused SwiGLU needs the whole K in one CTA
    p.out_swiglu = reinterpret_cast<uint16_t *>(out_swiglu);
    p.has_lo = map_x_lo != nullptr ? 1 : 0;
    p.swiglu_lo_off = swiglu_lo_off;
    const int it_per = (k_iters + ks - 1) / ks;
    ks = (k_iters + it_per - 1) / it_per;
    p.k_per_split = it_per * kTcBK;
    p.accumulate = accumulate;
    int cols = 32;
    while (cols < ((p.has_lo && p.bn <= 128) ? 2 * p.bn : p.bn)) cols <<= 1;
    p.tmem_cols = cols;
    const int stage_bytes = kTcBM * bk * 2 + (p.has_lo ? 2 : 1) * ((p.bn * bk * 2 + 1023) & ~1023);
    int stages = (200 * 1024) / stage_bytes;
    p.stages = stages > 8 ? 8 : (stages < 2 ? 2 : stages);
    const size_t smem = (size_t)p.stages * stage_bytes + (2 * p.stages + 1) * 8 + 16 + 1024;
    // per launch: the attribute belongs to the (function, device) pair, and callers may use several devices and threads
    B200_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    p.cs = tc_cluster_size(N, T);
    static const int w_early = getenv("B200LLM_TC_W_EARLY") ? atoi(getenv("B200LLM_TC_W_EARLY")) : 1;
    p.w_early = w_early;
    p.w_tm = reinterpret_cast<const uint8_t *>(w_tile_major);
    LaunchCfg lc(dim3(tiles, ks), dim3(kTcThreads), smem, stream, pdl);
    cudaLaunchAttribute attrs[2];
    int n_attr = 0;
    if (pdl) attrs[n_attr++] = lc.attr[0];
    if (p.cs > 1) {
        attrs[n_attr].id = cudaLaunchAttributeClusterDimension;
        attrs[n_attr].val.clusterDim.x = (unsigned)p.cs, attrs[n_attr].val.clusterDim.y = 1, attrs[n_attr].val.clusterDim.z = 1;
        ++n_attr;
    }
    lc.cfg.attrs = n_attr ? attrs : nullptr, lc.cfg.numAttrs = n_attr;
    B200_CUDA_OK(cudaLaunchKernelEx(&lc.cfg, gemm_tc_kernel, *reinterpret_cast<const CUtensorMap *>(map_w),
                                    *reinterpret_cast<const CUtensorMap *>(map_x),
                                    *reinterpret_cast<const CUtensorMap *>(map_x_lo ? map_x_lo : map_x), p));
    return B200LLM_OK;
}

}  // namespace b200
