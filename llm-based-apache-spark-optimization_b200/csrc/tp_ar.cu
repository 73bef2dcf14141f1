// tp_ar.cu -- one-shot all-reduce (+ residual add) of the tensor-parallel decode step over
// NVLink peer memory:   h[i] += sum over ranks of delta_rank[i]
//
// The 7B model at tensor parallel 2/4/8 needs two such reductions per layer of 4-64 KB each
// (o_proj and down_proj are row-parallel).  NCCL's all-reduce costs ~12 us at that size and
// needs a second kernel for the residual add; this kernel does both in one launch and one
// NVLink round trip:
//   * every rank owns an INBOX in a cudaMalloc'ed region shared with its peers through CUDA IPC:
//     inbox[phase][source rank][rows * d] floats and flag[phase][source rank][slice] sequence words;
//   * CTA j of every rank PUSHES slice j of its partial result into slot [phase][my rank] of every
//     peer's inbox (16-byte stores straight into peer memory), fences at system scope, and then
//     publishes the launch's sequence number in the peers' flag [phase][my rank][j];
//   * it then waits for the flags [phase][r][j] of all peers r in its OWN inbox (acquire loads on
//     local memory: no polling over NVLink), adds the pushed slices to its own partial and to the
//     residual stream.  Slices only depend on the same slice of the peers, so there is no grid-wide
//     or cross-GPU barrier.
// Two phases alternate: a rank can be at most one reduction ahead of a peer (it needs the peer's
// push of reduction k to finish k), so the slot it overwrites for k+1 was consumed in k-1.
// The sequence number lives in device memory (read at kernel start, bumped by the last CTA to
// finish) because the launch sits inside a replayed CUDA graph, where arguments are frozen.
#include "common.cuh"
#include "kernels.h"

namespace b200 {

struct TpArArgs {
    float *h;              // [n] residual stream, updated in place
    const float *delta;    // [n] this rank's partial result
    float *inbox[kTpMaxRanks];      // inbox base of every rank (own included), peer-mapped
    unsigned *flags[kTpMaxRanks];   // flag base of every rank
    unsigned *seq;         // local: [0] launch sequence number, [1] finished-CTA ticket, [2] token exchanges, [3] error word
    int32_t n;             // floats, multiple of 4
    int32_t rank, size;
    int32_t slot_floats;   // capacity of one inbox slot (floats)
    int32_t max_slices;    // flag words per (phase, source)
};

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Bounded wait for a peer's flag: a rank that has left the lock-step (a failed call, a crash, a different code path)
// must not leave its peers spinning inside a kernel that cannot be cancelled.  After 2 s the waiter records the
// failure in the local error word (seq[3]; the engine reports it as B200LLM_ENCCL on the next read / sync) and goes
// on, so the launch ends; its results are garbage and flagged as such.
__device__ __forceinline__ bool tp_wait_flag(const unsigned *f, unsigned want, unsigned *err) {
    if (ld_acquire_sys(f) == want) return true;
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (ld_acquire_sys(f) != want) {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 > 2000000000ull) {
            atomicExch(err, 1u);
            return false;
        }
    }
    return true;
}

constexpr int kArThreads = 256;
constexpr int kArSliceFloats = kArThreads * 4;   // one float4 per thread and slice

__global__ void __launch_bounds__(kArThreads) tp_allreduce_add_kernel(const __grid_constant__ TpArArgs a) {
    const unsigned s = *reinterpret_cast<volatile unsigned *>(a.seq) + 1;   // same value in every CTA of the launch
    const int phase = (int)(s & 1u);
    const int n_slices = (a.n + kArSliceFloats - 1) / kArSliceFloats;
    for (int j = blockIdx.x; j < n_slices; j += gridDim.x) {
        const int i = j * kArSliceFloats + threadIdx.x * 4;
        const bool live = i < a.n;
        float4 mine = make_float4(0.f, 0.f, 0.f, 0.f);
        if (live) mine = *reinterpret_cast<const float4 *>(a.delta + i);
        // push my slice into slot [phase][my rank] of every peer
        for (int r = 0; r < a.size; ++r) {
            if (r == a.rank || !live) continue;
            float *dst = a.inbox[r] + ((size_t)phase * a.size + a.rank) * a.slot_floats + i;
            *reinterpret_cast<float4 *>(dst) = mine;
        }
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x < a.size && threadIdx.x != a.rank)
            st_release_sys(a.flags[threadIdx.x] + ((size_t)phase * a.size + a.rank) * a.max_slices + j, s);
        // wait for the same slice of every peer in my own inbox
        if (threadIdx.x < a.size && threadIdx.x != a.rank) {
            const unsigned *f = a.flags[a.rank] + ((size_t)phase * a.size + threadIdx.x) * a.max_slices + j;
            tp_wait_flag(f, s, a.seq + 3);
        }
        __syncthreads();
        if (live) {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int r = 0; r < a.size; ++r) {   // rank order 0, 1, ..: every rank computes the same sum bit for bit
                float4 v = mine;
                if (r != a.rank)
                    v = __ldcv(reinterpret_cast<const float4 *>(a.inbox[a.rank] + ((size_t)phase * a.size + r) * a.slot_floats + i));
                acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
            }
            float4 hv = *reinterpret_cast<float4 *>(a.h + i);
            hv.x += acc.x, hv.y += acc.y, hv.z += acc.z, hv.w += acc.w;
            *reinterpret_cast<float4 *>(a.h + i) = hv;
        }
    }
    // the last CTA to finish publishes the sequence number for the next launch
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(a.seq + 1, 1u) == gridDim.x - 1) {
            a.seq[1] = 0;
            __threadfence();
            *reinterpret_cast<volatile unsigned *>(a.seq) = s;
        }
    }
}

int tp_allreduce_add_launch(float *h, const float *delta, int n, float *const *inbox, unsigned *const *flags,
                            unsigned *seq, int rank, int size, int slot_floats, int max_slices, cudaStream_t stream) {
    B200_REQUIRE(n % 4 == 0 && n <= slot_floats && size >= 2 && size <= kTpMaxRanks, "tp all-reduce: bad shape");
    const int n_slices = (n + kArSliceFloats - 1) / kArSliceFloats;
    B200_REQUIRE(n_slices <= max_slices, "tp all-reduce: %d slices exceed %d flags", n_slices, max_slices);
    TpArArgs a{};
    a.h = h, a.delta = delta, a.seq = seq, a.n = n, a.rank = rank, a.size = size;
    a.slot_floats = slot_floats, a.max_slices = max_slices;
    for (int r = 0; r < size; ++r) a.inbox[r] = inbox[r], a.flags[r] = flags[r];
    // every CTA of the launch reads the same sequence number: it is advanced only by the LAST CTA to finish (ticket
    // in seq[1]), and CTAs of the next launch on the stream start after this one has ended
    const int grid = n_slices < kNumSMs ? n_slices : kNumSMs;
    tp_allreduce_add_kernel<<<grid, kArThreads, 0, stream>>>(a);
    B200_CUDA_OK(cudaGetLastError());
    return B200LLM_OK;
}

}  // namespace b200
