// ubench_tmem_a.cu -- can the weight tile (the A operand, M = 128 x K = 64 bf16, 128-byte swizzled K-major in
// shared memory) be moved to TENSOR MEMORY with tcgen05.cp and multiplied from there (tcgen05.mma with A in TMEM)?
// Checks the result against the shared-memory-operand MMA the decode kernel uses today and times both.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_tmem_a ubench_tmem_a.cu
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
__device__ __forceinline__ uint16_t bf(float v) { return (uint16_t)(__float_as_uint(v) >> 16); }
__device__ __forceinline__ void commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void wait(uint64_t *bar, uint32_t par) {
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(par) : "memory");
}

// mode 0: A from shared memory (reference), 1: tcgen05.cp 128x256b then A from TMEM, 2: copies only, 3: MMAs from TMEM only (timing)
__global__ void __launch_bounds__(128, 1) k(int mode, int N, int n_rep, float *out, long long *clk, int n_acc = 1) {
    extern __shared__ uint8_t dyn[];
    uint8_t *base = (uint8_t *)(((uintptr_t)dyn + 1023) & ~(uintptr_t)1023);
    uint8_t *A = base, *B = base + 8 * 16384;
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int tid = threadIdx.x;
    for (int st = 0; st < 8; ++st)
        for (int i = tid; i < 128 * 8; i += 128) {   // (row, 16-byte chunk)
            const int r = i >> 3, c = i & 7;
            uint16_t v[8];
            for (int e = 0; e < 8; ++e) v[e] = bf((float)(((r * 7 + (c * 8 + e) * 3 + st) % 17) - 8));
            *(uint4 *)(A + st * 16384 + r * 128 + ((c ^ (r & 7)) * 16)) = *(uint4 *)v;
        }
    for (int i = tid; i < 16 * 8; i += 128) {        // operand block: 16 rows (rows >= 2 zero)
        const int r = i >> 3, c = i & 7;
        uint16_t v[8];
        for (int e = 0; e < 8; ++e) {
            const int kk = c * 8 + e;
            v[e] = r == 0 ? bf((float)((kk % 5) - 2)) : r == 1 ? bf((float)(kk % 3)) : 0;
        }
        *(uint4 *)(B + r * 128 + ((c ^ (r & 7)) * 16)) = *(uint4 *)v;
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
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
    const uint32_t tA = tmem + 128;   // columns 128.. hold A stages of 32 columns
    if (tid < 32) {   // whole warp, converged; one elected lane issues (operands stay in uniform registers)
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
        const uint32_t a0 = smem_u32(A), b0 = smem_u32(B);
        long long t0 = clock64();
        for (int rep = 0; rep < n_rep; ++rep) {
            const uint32_t st = (uint32_t)rep & 7u;
            const uint64_t ad = desc_sw128(a0 + st * 16384, 1024);
            const uint64_t bd = desc_sw128(b0, 1024);
            const uint32_t ta = tA + st * 32;
            uint32_t el;
            asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.b32 %0, 1, 0, P;\n\t}" : "=r"(el));
            if (el) {
            const uint32_t dacc = tmem + (uint32_t)(rep % n_acc) * 16;
            if (mode == 1 || mode == 2) {
#pragma unroll
                for (int kq = 0; kq < 4; ++kq)
                    asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(ta + 8 * kq), "l"(ad + (uint64_t)(2 * kq)) : "memory");
            }
#pragma unroll
            for (int kq = 0; kq < 4 && mode != 2; ++kq) {
                const uint32_t acc = (rep >= n_acc || kq > 0) ? 1u : 0u;
                if (mode == 0)
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(dacc),
                                 "l"(ad + (uint64_t)(2 * kq)), "l"(bd + (uint64_t)(2 * kq)), "r"(idesc), "r"(acc) : "memory");
                else
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(dacc),
                                 "r"(ta + 8 * kq), "l"(bd + (uint64_t)(2 * kq)), "r"(idesc), "r"(acc) : "memory");
            }
            }
            __syncwarp();
        }
        long long t1 = clock64();
        if (tid == 0) commit(&bar);
        __syncwarp();
        wait(&bar, 0);
        long long t2 = clock64();
        if (tid == 0) clk[0] = t1 - t0, clk[1] = t2 - t0;
    }
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    uint32_t r0, r1;
    const uint32_t taddr = tmem + ((uint32_t)((tid >> 5) * 32) << 16);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    out[2 * tid] = __uint_as_float(r0), out[2 * tid + 1] = __uint_as_float(r1);
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

int main() {
    float *d, h[2][256];
    long long *c, hc[2];
    cudaMalloc(&d, 1024);
    cudaMalloc(&c, 16);
    const int smem = 9 * 16384 + 2048;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    for (int N : {8, 16}) {
        for (int n_rep : {1, 8}) {
            for (int mode = 0; mode < 2; ++mode) {
                k<<<1, 128, smem>>>(mode, N, n_rep, d, c);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("N=%d mode=%d: %s\n", N, mode, cudaGetErrorString(e)); return 1; }
                cudaMemcpy(h[mode], d, 1024, cudaMemcpyDeviceToHost);
            }
            // host expectation
            double worst_ref = 0, worst_ts = 0;
            for (int r = 0; r < 128; ++r) {
                double e0 = 0, e1 = 0;
                for (int rep = 0; rep < n_rep; ++rep)
                    for (int kk = 0; kk < 64; ++kk) {
                        const double w = ((r * 7 + kk * 3 + (rep & 7)) % 17) - 8;
                        e0 += w * ((kk % 5) - 2), e1 += w * (kk % 3);
                    }
                worst_ref = fmax(worst_ref, fmax(fabs(h[0][2 * r] - e0), fabs(h[0][2 * r + 1] - e1)));
                worst_ts = fmax(worst_ts, fmax(fabs(h[1][2 * r] - e0), fabs(h[1][2 * r + 1] - e1)));
            }
            printf("N=%2d stages=%d: max |D - expected|: A from smem %g, A via tcgen05.cp from TMEM %g   (row 5: %g %g | %g %g)\n", N, n_rep,
                   worst_ref, worst_ts, h[0][10], h[0][11], h[1][10], h[1][11]);
        }
    }
    const char *names[4] = {"mma(A in smem)     ", "cp + mma(A in TMEM)", "cp only            ", "mma(A in TMEM) only"};
    for (int N : {8, 16})
        for (int mode = 0; mode < 4; ++mode) {
            const int n_rep = 256;
            for (int n_acc : {1, 4}) {
                k<<<1, 128, smem>>>(mode, N, n_rep, d, c, n_acc);
                k<<<1, 128, smem>>>(mode, N, n_rep, d, c, n_acc);
                cudaDeviceSynchronize();
                cudaMemcpy(hc, c, 16, cudaMemcpyDeviceToHost);
                printf("N=%2d %s accs=%d: issue %.1f clk per 16 KB stage, done %.1f clk per stage\n", N, names[mode], n_acc,
                       (double)hc[0] / n_rep, (double)hc[1] / n_rep);
            }
        }
    return 0;
}
