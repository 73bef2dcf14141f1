// ubench_issue.cu -- what does one [128 x 64] bf16 stage (4 x tcgen05.mma M128 x N8 x K16, A and B in shared memory) cost the
// ISSUING warp, by code shape?  (The decode kernel's consumer has to outrun this SM's share of the HBM stream, 16 KB / 0.36 us.)
//   mode 0: lane 0 of a diverged warp (operands in vector registers: ELECT / R2UR.BROADCAST / BRA.U.ANY loop per instruction)
//   mode 1: converged warp, elect.sync around every group of GRP stages (operands in uniform registers), __syncwarp after
//   mode 2: as 1 without the __syncwarp
//   mode 3: elect per stage + a tcgen05.commit per stage (onto a barrier nobody waits for)
//   mode 4: 3 + mbarrier.try_wait on an already complete barrier + tcgen05.fence::after_thread_sync per stage
//   mode 5: 3 + the try_wait only      mode 6: 3 + the fence only
//   mode 7: 4, but the NEXT stage's barrier is probed (test_wait) before this stage's MMAs are issued and the result used after
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_issue ubench_issue.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc_sw128(uint32_t a, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((a & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(sbo >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__device__ __forceinline__ bool elect() {
    uint32_t el;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.b32 %0, 1, 0, P;\n\t}" : "=r"(el));
    return el != 0;
}
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                 "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ uint32_t try_wait(uint64_t *bar, uint32_t par) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(par) : "memory");
    return ok;
}
__device__ __forceinline__ uint32_t test_wait(uint64_t *bar, uint32_t par) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(par) : "memory");
    return ok;
}

template <int MODE, int GRP>
__global__ void __launch_bounds__(128, 1) k(int n_stage, long long *clk, int n_warps) {
    extern __shared__ uint8_t dyn[];
    uint8_t *base = (uint8_t *)(((uintptr_t)dyn + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar[4];
    __shared__ uint32_t slot;
    const int tid = threadIdx.x, w = tid >> 5;
    for (int i = tid; i < (8 * 16384 + 1024) / 4; i += 128) ((uint32_t *)base)[i] = 0x3c003c00u;
    if (tid == 0) {
        for (int i = 0; i < 4; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[i])));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    if (tid < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;");
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = __shfl_sync(0xffffffffu, slot, 0);
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((8u >> 3) << 17) | ((128u >> 4) << 24);
    const uint32_t a0 = smem_u32(base), b0 = smem_u32(base + 8 * 16384);
    const uint32_t d = tmem + (uint32_t)__shfl_sync(0xffffffffu, w, 0) * 32;
    if (w < n_warps) {
        long long t0 = clock64();
        if (MODE == 0) {
            if ((tid & 31) == 0) {
                for (int st = 0; st < n_stage; ++st) {
                    const uint64_t ad = desc_sw128(a0 + (uint32_t)(st & 7) * 16384, 1024), bd = desc_sw128(b0, 0);
#pragma unroll
                    for (int kq = 0; kq < 4; ++kq) mma(d, ad + 2 * kq, bd + 2 * kq, idesc, (st > 0 || kq > 0) ? 1u : 0u);
                }
            }
        } else if (MODE <= 2) {
            for (int st = 0; st < n_stage; st += GRP) {
                if (elect()) {
#pragma unroll
                    for (int g = 0; g < GRP; ++g) {
                        const uint64_t ad = desc_sw128(a0 + (uint32_t)((st + g) & 7) * 16384, 1024), bd = desc_sw128(b0, 0);
#pragma unroll
                        for (int kq = 0; kq < 4; ++kq) mma(d, ad + 2 * kq, bd + 2 * kq, idesc, (st + g > 0 || kq > 0) ? 1u : 0u);
                    }
                }
                if (MODE == 1) __syncwarp();
            }
        } else {
            uint32_t ok_next = 1;
            for (int st = 0; st < n_stage; ++st) {
                if (MODE == 4 || MODE == 5)
                    while (!try_wait(&bar[2], 1)) {
                    }
                if (MODE == 7)
                    while (!ok_next) ok_next = try_wait(&bar[2], 1);
                if (MODE == 4 || MODE == 6 || MODE == 7) asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (MODE == 7) ok_next = test_wait(&bar[2], 1);
                if (elect()) {
                    const uint64_t ad = desc_sw128(a0 + (uint32_t)(st & 7) * 16384, 1024), bd = desc_sw128(b0, 0);
#pragma unroll
                    for (int kq = 0; kq < 4; ++kq) mma(d, ad + 2 * kq, bd + 2 * kq, idesc, (st > 0 || kq > 0) ? 1u : 0u);
                    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar[3])) : "memory");
                }
            }
        }
        long long t1 = clock64();
        if ((tid & 31) == 0) {
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar[w])) : "memory");
            while (!try_wait(&bar[w], 0)) {
            }
            long long t2 = clock64();
            clk[2 * w] = t1 - t0, clk[2 * w + 1] = t2 - t0;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

template <int MODE, int GRP>
void run(const char *name, long long *c, int n_warps) {
    long long h[8];
    const int smem = 9 * 16384 + 2048, n = 512;
    cudaFuncSetAttribute(k<MODE, GRP>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    k<MODE, GRP><<<1, 128, smem>>>(n, c, n_warps);
    k<MODE, GRP><<<1, 128, smem>>>(n, c, n_warps);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(h, c, 64, cudaMemcpyDeviceToHost);
    printf("%-62s warps=%d: %6.1f clk per 16 KB stage per warp, %6.1f aggregate %s\n", name, n_warps, (double)h[1] / n,
           (double)h[1] / n / n_warps, e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main() {
    long long *c;
    cudaMalloc(&c, 64);
    for (int nw : {1, 2}) {
        run<0, 1>("lane 0 of a diverged warp", c, nw);
        run<1, 1>("converged warp, elect per stage, syncwarp", c, nw);
        run<2, 1>("converged warp, elect per stage", c, nw);
        run<2, 4>("converged warp, elect per 4 stages", c, nw);
        run<3, 1>("elect per stage + tcgen05.commit per stage", c, nw);
        run<4, 1>("+ mbarrier.try_wait (complete) + tcgen05.fence per stage", c, nw);
        run<5, 1>("+ mbarrier.try_wait only", c, nw);
        run<6, 1>("+ tcgen05.fence only", c, nw);
        run<7, 1>("+ fence, next stage's barrier probed before the MMAs", c, nw);
    }
    return 0;
}
