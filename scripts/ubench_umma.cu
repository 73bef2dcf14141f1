// ubench_umma.cu -- how fast does ONE thread retire small-N tcgen05.mma instructions?
// (M = 128, K = 16, bf16, A and B from shared memory, 128-byte swizzle, D in TMEM.)
// Variants: N in {8, 16, 32, 64}; all MMAs into one accumulator vs rotating over 2 / 4
// accumulators; A operand walking over 1 / 8 different 16 KB stages.  Prints clk per MMA.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_umma ubench_umma.cu
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

__global__ void __launch_bounds__(128, 1) k(int n_mma, int N, int n_acc, int n_stage, long long *out, int M = 128, int a_sbo = 1024, int check = 0) {
    extern __shared__ uint8_t dyn[];
    uint8_t *base = (uint8_t *)(((uintptr_t)dyn + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    for (int i = threadIdx.x; i < (8 * 16384 + 16384) / 4; i += blockDim.x) ((uint32_t *)base)[i] = 0x3c003c00u + i;
    if (check) {
        // A block 0: row r (0..7) = r + 1 for every k (bf16), swizzle irrelevant for constant rows; rows 8.. of the
        // stage hold 100 (must NOT show up when a_sbo = 0 aliases every 8-row group onto group 0)
        for (int i = threadIdx.x; i < 16384 / 2; i += blockDim.x) {
            const int row = i / 64;
            const float v = row < 8 ? (float)(row + 1) : 100.f;
            ((uint16_t *)base)[i] = (uint16_t)(__float_as_uint(v) >> 16);
        }
        // B: all ones
        for (int i = threadIdx.x; i < 16384 / 2; i += blockDim.x) ((uint16_t *)(base + 8 * 16384))[i] = 0x3f80;
    }
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;");
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = slot;
    if (threadIdx.x == 0) {
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (((uint32_t)M >> 4) << 24);
        const uint32_t a0 = smem_u32(base), b0 = smem_u32(base + 8 * 16384);
        long long t0 = clock64();
        for (int i = 0; i < n_mma; ++i) {
            const uint32_t st = (uint32_t)(i >> 2) % (uint32_t)n_stage;
            const uint64_t ad = desc_sw128(a0 + st * 16384, a_sbo) + 2 * (i & 3);
            const uint64_t bd = desc_sw128(b0, 1024) + 2 * (i & 3);
            const uint32_t d = tmem + (uint32_t)(i % n_acc) * (N >= 128 ? 256 : 64);
            const uint32_t acc = i >= n_acc;
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                "l"(ad), "l"(bd), "r"(idesc), "r"(acc)
                : "memory");
        }
        long long t1 = clock64();
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        uint32_t ok = 0;
        while (!ok)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok)
                         : "r"(smem_u32(&bar))
                         : "memory");
        long long t2 = clock64();
        out[0] = t1 - t0;
        out[1] = t2 - t0;
    }
    __syncthreads();
    if (check) {
        asm volatile("tcgen05.fence::after_thread_sync;");
        uint32_t r0;
        const uint32_t taddr = tmem + ((uint32_t)((threadIdx.x >> 5) * 32) << 16);
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r0) : "r"(taddr) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        ((float *)(out + 2))[threadIdx.x] = __uint_as_float(r0);
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

// two issuer threads in different warps, each with its own accumulator and barrier
__global__ void __launch_bounds__(128, 1) k2(int n_mma, int N, long long *out, int n_issuers) {
    extern __shared__ uint8_t dyn[];
    uint8_t *base = (uint8_t *)(((uintptr_t)dyn + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar[4];
    __shared__ uint32_t slot;
    for (int i = threadIdx.x; i < (8 * 16384 + 16384) / 4; i += blockDim.x) ((uint32_t *)base)[i] = 0x3c003c00u + i;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[i])));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;");
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = slot;
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0 && w < n_issuers) {
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
        const uint32_t a0 = smem_u32(base), b0 = smem_u32(base + 8 * 16384);
        long long t0 = clock64();
        for (int i = 0; i < n_mma; ++i) {
            const uint32_t st = ((uint32_t)(i >> 2) * n_issuers + w) % 8u;
            const uint64_t ad = desc_sw128(a0 + st * 16384, 1024) + 2 * (i & 3);
            const uint64_t bd = desc_sw128(b0, 1024) + 2 * (i & 3);
            const uint32_t d = tmem + (uint32_t)w * 64;
            const uint32_t acc = i >= 1;
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                "l"(ad), "l"(bd), "r"(idesc), "r"(acc)
                : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar[w])) : "memory");
        uint32_t ok = 0;
        while (!ok)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok)
                         : "r"(smem_u32(&bar[w]))
                         : "memory");
        long long t2 = clock64();
        out[w] = t2 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

int main() {
    long long *d, h[2];
    cudaMalloc(&d, 16 + 512);
    const int smem = 10 * 16384 + 2048;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int n = 512;
    printf("%4s %5s %7s %12s %12s\n", "N", "accs", "stages", "issue clk/mma", "done clk/mma");
    for (int N : {8, 64, 128, 256})
        for (int na : {1, 2})
            for (int ns : {8}) {
                k<<<1, 128, smem>>>(n, N, na, ns, d);
                k<<<1, 128, smem>>>(n, N, na, ns, d);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) {
                    printf("N=%d: %s\n", N, cudaGetErrorString(e));
                    return 1;
                }
                cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
                printf("%4d %5d %7d %12.1f %12.1f\n", N, na, ns, (double)h[0] / n, (double)h[1] / n);
            }
    printf("-- M = 64 / A sbo = 0 variants (N, M, a_sbo): issue, done clk per mma\n");
    for (int N : {8, 128, 256})
        for (int M : {64, 128})
            for (int sbo : {1024, 0}) {
                k<<<1, 128, smem>>>(n, N, 1, 8, d, M, sbo, 0);
                k<<<1, 128, smem>>>(n, N, 1, 8, d, M, sbo, 0);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("N=%d M=%d sbo=%d: %s\n", N, M, sbo, cudaGetErrorString(e)); return 1; }
                cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
                printf("%4d %4d %5d %12.1f %12.1f\n", N, M, sbo, (double)h[0] / n, (double)h[1] / n);
            }
    cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    for (int ni : {1, 2, 4}) {
        long long hh[4];
        k2<<<1, 128, smem>>>(n, 8, d, ni);
        k2<<<1, 128, smem>>>(n, 8, d, ni);
        cudaDeviceSynchronize();
        cudaMemcpy(hh, d, 32, cudaMemcpyDeviceToHost);
        printf("issuers=%d (N=8, %d MMAs each): clk per MMA per issuer %.1f -> aggregate %.1f clk/MMA\n", ni, n, (double)hh[0] / n, (double)hh[0] / n / ni);
    }
    // correctness of the aliasing descriptor: one MMA (K = 16), A rows r = r + 1, B = ones -> D[m][0] = 16 (m % 8 + 1)
    for (int sbo : {1024, 0}) {
        float hf[128];
        k<<<1, 128, smem>>>(1, 8, 1, 1, d, 128, sbo, 1);
        cudaDeviceSynchronize();
        cudaMemcpy(hf, (char *)d + 16, 512, cudaMemcpyDeviceToHost);
        printf("a_sbo=%d D[0..19][0]:", sbo);
        for (int i = 0; i < 20; ++i) printf(" %g", hf[i]);
        printf(" ... D[127] %g\n", hf[127]);
    }
    return 0;
}
