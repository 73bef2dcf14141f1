// ubench_sync.cu -- micro-benchmark of the persistent-kernel phase boundary on B200:
// 148 CTAs each write a slice of a vector, pass a grid barrier, then all read the whole vector.
// Prints the time (globaltimer ns) of: barrier alone, read of the vector after the barrier
// (ld.global.cg), with several vector sizes and read patterns.  Build: nvcc -arch=sm_100a -O3.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void bar512() { asm volatile("bar.sync 1, 512;" ::: "memory"); }
__device__ __forceinline__ void grid_sync(unsigned *ctr, unsigned target) {
    bar512();
    if (threadIdx.x == 0) {
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
        unsigned v;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
        } while (v < target);
    }
    bar512();
}

// mode 0: float4 per thread contiguous (coalesced 16 B/lane); mode 1: the megakernel's float2 pattern
// threads >= 512 are background streamers: they read `stream_bytes` of `big` over and over
// (each CTA its own contiguous slice) until the sync threads are done.
__global__ void __launch_bounds__(768, 1) k(float *vec, int n, unsigned *ctr, unsigned long long *out, int iters, int mode,
                                            float *sink, const uint4 *big, size_t stream_bytes, volatile int *stop) {
    const int G = gridDim.x, bid = blockIdx.x, tid = threadIdx.x;
    if (tid >= 512) {
        if (stream_bytes == 0) return;
        const size_t per_cta = stream_bytes / G / 16;  // uint4 elements
        const uint4 *base = big + (size_t)bid * per_cta;
        unsigned accu = 0;
        size_t i = tid - 512;
        while (!*stop) {
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                uint4 v;
                asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                             : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(base + i));
                accu ^= v.x ^ v.y ^ v.z ^ v.w;
                i += 256;
                if (i >= per_cta) i -= per_cta;
            }
        }
        if (accu == 0x12345u) sink[2] = 1.f;
        return;
    }
    unsigned bar = 0;
    float acc = 0.f;
    unsigned long long t_bar = 0, t_rd = 0, t_wr = 0;
    for (int it = 0; it < iters; ++it) {
        // each CTA writes its slice
        unsigned long long t0 = gtime();
        for (int i = bid * 512 + tid; i < n; i += G * 512) vec[i] = (float)(it + i);
        unsigned long long t1 = gtime();
        grid_sync(ctr, ++bar * G);
        unsigned long long t2 = gtime();
        if (mode == 0) {
            for (int i = tid * 4; i < n; i += 512 * 4) {
                float4 v = __ldcg(reinterpret_cast<const float4 *>(vec + i));
                acc += v.x + v.y + v.z + v.w;
            }
        } else {
            for (int item = tid; item < n / 4; item += 512) {
                const int kk = (item >> 2) * 16 + (item & 3) * 2;
                float2 a = __ldcg(reinterpret_cast<const float2 *>(vec + kk));
                float2 b = __ldcg(reinterpret_cast<const float2 *>(vec + kk + 8));
                acc += a.x + a.y + b.x + b.y;
            }
        }
        if (acc == 1.2345f) sink[0] = acc;  // keep the loads
        bar512();
        unsigned long long t3 = gtime();
        grid_sync(ctr, ++bar * G);  // nobody overwrites vec before all have read it
        if (it >= 4) t_wr += t1 - t0, t_bar += t2 - t1, t_rd += t3 - t2;
    }
    grid_sync(ctr, ++bar * G);
    if (bid == 0 && tid == 0) *stop = 1;
    if (tid == 0) out[bid * 3 + 0] = t_wr, out[bid * 3 + 1] = t_bar, out[bid * 3 + 2] = t_rd;
    if (acc == 1.2345f) sink[1] = acc;
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float *vec, *sink;
    unsigned *ctr;
    unsigned long long *out;
    cudaMalloc(&vec, 1 << 20);
    cudaMalloc(&sink, 64);
    cudaMalloc(&ctr, 4);
    cudaMalloc(&out, sms * 3 * 8);
    const int iters = 104;
    uint4 *big;
    const size_t big_bytes = (size_t)8 << 30;
    cudaMalloc(&big, big_bytes);
    cudaMemset(big, 1, big_bytes);
    int *stop;
    cudaMalloc(&stop, 4);
    const size_t sizes[] = {0, (size_t)32 << 20, (size_t)512 << 20, (size_t)8 << 30};
    for (size_t stream_bytes : sizes)
    for (int mode = 1; mode < 2; ++mode)
        for (int n : {3072, 8192}) {
            cudaMemset(ctr, 0, 4);
            cudaMemset(stop, 0, 4);
            printf("stream over %5zu MB: ", stream_bytes >> 20);
            void *args[] = {&vec, &n, &ctr, &out, (void *)&iters, &mode, &sink, &big, &stream_bytes, &stop};
            cudaError_t e = cudaLaunchCooperativeKernel((void *)k, dim3(sms), dim3(768), args, 0, 0);
            cudaDeviceSynchronize();
            if (e != cudaSuccess || cudaGetLastError() != cudaSuccess) {
                printf("launch failed: %s\n", cudaGetErrorString(e));
                return 1;
            }
            unsigned long long h[148 * 3];
            cudaMemcpy(h, out, sms * 3 * 8, cudaMemcpyDeviceToHost);
            double wr = 0, bar = 0, rd = 0, rdmax = 0;
            for (int b = 0; b < sms; ++b) {
                wr += h[b * 3], bar += h[b * 3 + 1], rd += h[b * 3 + 2];
                if ((double)h[b * 3 + 2] > rdmax) rdmax = (double)h[b * 3 + 2];
            }
            const double d = (double)sms * (iters - 4) * 1e3;
            printf("mode %d n=%6d floats: write %.2f us  barrier %.2f us  read-all %.2f us (max CTA %.2f us)\n", mode, n,
                   wr / d, bar / d, rd / d, rdmax / ((iters - 4) * 1e3));
        }
    return 0;
}
