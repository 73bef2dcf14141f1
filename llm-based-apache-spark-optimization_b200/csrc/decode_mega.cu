// decode_mega.cu -- the batch-1 decode step as ONE persistent kernel (one CTA per SM).
//
// At batch 1 a Llama-3.2-3B step is ~140 GEMV/attention phases of 3-15 us each; launched as
// separate kernels, ramp-up and drain cost more than the streaming itself (profiles/
// launches_r1a: 2.3 ms per step against a 0.98 ms HBM floor).  Here all phases of all steps of
// a decode call run inside one cooperative grid:
//   * a producer warp per CTA walks the whole weight schedule (layer after layer, step after
//     step) and streams this CTA's rows of every matrix through a 165 KB shared-memory ring
//     with 1-D TMA bulk copies (one per weight row, rows skewed by 16 B so that ldmatrix is
//     bank-conflict free); it never waits for anything but ring space, so HBM stays busy across
//     phase boundaries;
//   * 16 compute warps consume the ring on the tensor cores: the activation vector is split
//     into bf16 hi + lo parts (together ~fp32 accurate) that form rows 0/1 of the m16 operand
//     of mma.m16n8k16, 8 weight rows form the n8 operand (ldmatrix.x4 from the ring), fp32
//     accumulate.  A first version did this with CUDA-core FMAs and was issue-bound at 17 B/clk
//     per SM (profiles/mega_trace_r1a), below the 23 B/clk the HBM stream needs;
//   * per-warp partials are combined in shared memory, then the fused epilogues run (RMSNorm
//     prologue, RoPE + paged-KV append, SwiGLU, residual, logits + argmax);
//   * phases are separated by a grid barrier (one atomic + acquire poll per CTA) on the
//     consumer side only; attention (split-T over the lane's KV pages, K/V tiles staged by TMA)
//     runs between the qkv and o_proj phases while the ring fills with o_proj's weights;
//   * greedy sampling and token feedback happen in-kernel, so n_steps tokens need one launch.
#include "common.cuh"
#include "kernels.h"
#include "mega.h"

namespace b200 {

constexpr int kMC = 384;                       // compute threads (12 warps + the producer warp leave 152 registers
                                               // per thread; with 16 warps the 96-register cap spilled, and a spill
                                               // right after a grid barrier is an L2 round trip)
constexpr int kMW = kMC / 32;                  // compute warps
constexpr int kSPW = (64 + kMW - 1) / kMW;     // 32-wide k steps of a ring stage per warp (stage <= 2048 columns)
constexpr int kMaxKC = 2048;                   // columns of one ring stage (8 rows x kMaxKC)
constexpr int kSlotBytes = 33792;              // >= 8 * (kMaxKC * 2 + 16), multiple of 128
constexpr int kSlots = 5;
constexpr int kXtabBytes = 45056;              // activation operand table: K * 4 bytes (K <= 11264)
constexpr int kAttW = 4;                       // warps that stream KV pages in the attention phase
constexpr int kPartBytes = 8704;               // per-warp GEMV partials / attention merge buffer
constexpr int kRB = 128;                       // rows combined per partial block
constexpr int kMPT = 16;                       // tokens per KV page
constexpr int kMaxJobs = 4 * 32 + 1;           // phase descriptors kept in shared memory (<= 32 layers)
constexpr int kMaxPtab = 448;                  // page-table entries kept in shared memory

struct MegaSmem {
    uint8_t *ring, *xtab;  // xtab doubles as the K/V tile buffer of the attention phase
    float *part;
    uint64_t *full, *empty, *att_full;
    float *red;
    int *misc;
    float *rope;  // [head_dim/2][cos, sin] of the current position
    MegaJob *jobs;    // copy of the phase descriptors (global reads after a gpu-scope acquire miss in L1)
    int32_t *ptab;    // copy of the lane's page table (first kMaxPtab entries)
};

__device__ __forceinline__ MegaSmem carve(uint8_t *base) {
    MegaSmem s;
    s.ring = base;
    s.xtab = base + kSlots * kSlotBytes;
    s.part = reinterpret_cast<float *>(s.xtab + kXtabBytes);
    s.full = reinterpret_cast<uint64_t *>(reinterpret_cast<uint8_t *>(s.part) + kPartBytes);
    s.empty = s.full + kSlots;
    s.att_full = s.empty + kSlots;
    s.red = reinterpret_cast<float *>(s.att_full + kAttW);
    s.misc = reinterpret_cast<int *>(s.red + 64);
    s.rope = reinterpret_cast<float *>(s.misc + 64);
    s.jobs = reinterpret_cast<MegaJob *>(s.rope + 128);
    s.ptab = reinterpret_cast<int32_t *>(s.jobs + kMaxJobs);
    return s;
}

size_t mega_smem_bytes(int) {
    return (size_t)kSlots * kSlotBytes + kXtabBytes + kPartBytes + (2 * kSlots + kAttW) * 8 + 64 * 4 + 256 + 128 * 4 + kMaxJobs * sizeof(MegaJob) +
           kMaxPtab * 4;
}

__host__ __device__ __forceinline__ int k_parts(int K) {
    int p = (K + kMaxKC - 1) / kMaxKC;
    while ((K / 32) % p) ++p;
    return p;
}

bool mega_supported(int d, int ffn, int n_q, int n_kv, int head_dim) {
    auto ok_k = [](int K) { return K % 32 == 0 && (size_t)K * 4 <= (size_t)kXtabBytes; };
    const int G = n_kv > 0 ? n_q / n_kv : 0;
    const bool att_fits = (size_t)kAttW * 2 * kMPT * head_dim * 2 <= (size_t)kXtabBytes;
    return ok_k(d) && ok_k(ffn) && ok_k(n_q * head_dim) && (head_dim == 64 || head_dim == 128) && G >= 1 && G <= 4 &&
           att_fits && ffn % 4 == 0 && d % 8 == 0;
}

struct RowRange {
    int row0, nrows;
};
// rows of this CTA, in units of 8 rows (one n8 tensor-core operand)
__device__ __forceinline__ RowRange cta_rows(int N, int G, int bid) {
    const int n_units = N / 8;
    const int base = n_units / G, rem = n_units % G;
    RowRange r;
    r.row0 = (bid * base + (bid < rem ? bid : rem)) * 8;
    r.nrows = (base + (bid < rem ? 1 : 0)) * 8;
    return r;
}

// Grid barrier for the compute threads (the producer warp never takes part).
// (release/acquire at gpu scope; the CTA barrier in front makes the release cumulative over the
// writes of all compute threads.)
__device__ __forceinline__ void grid_sync(unsigned *ctr, unsigned target, int tid) {
    bar_sync(1, kMC);
    if (tid == 0) {
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
        // poll relaxed and fence once: an acquire load per poll costs an L1 invalidate (CCTL.IVALL)
        // per iteration, ~100 per barrier, which also throws out every warp's local-memory lines
        unsigned v;
        do {
            asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
        } while (v < target);
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
    }
    bar_sync(1, kMC);
}

// ---------------------------------------------------------------- producer: the weight stream
__device__ void mega_producer(const MegaParams &p, const MegaSmem &s, int G, int bid, int lane) {
    const uint64_t pol = l2_policy_evict_first();
    const int n_jobs = 4 * p.n_layers + 1;
    uint32_t gs = 0;
    int seq = 0;  // phases started so far (all steps)
    volatile int *gate = s.misc + 2;
    for (int step = 0; step < p.n_steps; ++step) {
        for (int j = 0; j < n_jobs; ++j, ++seq) {
            const MegaJob job = s.jobs[j];
            int ahead = 0;  // stages of this phase issued while its gate is still closed
            const RowRange rr = cta_rows(job.N, G, bid);
            const int parts = k_parts(job.K);
            const int kc = job.K / parts;
            const uint32_t row_smem = (uint32_t)kc * 2 + 16;
            const size_t row_bytes = (size_t)job.K * 2;
            const uint8_t *src = reinterpret_cast<const uint8_t *>(job.w) + (size_t)rr.row0 * row_bytes;
            for (int grp = 0; grp < rr.nrows / 8; ++grp) {
                for (int pt = 0; pt < parts; ++pt, ++gs) {
                    const uint32_t slot = gs % (uint32_t)p.ring_slots, use = gs / (uint32_t)p.ring_slots;
                    // Deep prefetch inflates the latency of everything else that goes to memory (a demand
                    // load queues behind ~120 KB per SM of bulk copies), and the consumers' barrier
                    // + activation loads between two phases are pure latency: only `gap_slots`
                    // stages of a phase may be in flight before its consumers have their activations.
                    if (ahead >= p.gap_slots) {
                        while (*gate <= seq) {
                        }
                    } else {
                        ++ahead;
                    }
                    if (use > 0) mbar_wait(&s.empty[slot], (use - 1) & 1);
                    if (lane == 0) mbar_expect_tx(&s.full[slot], ((job.x_src >> 8) & 1) ? 8u * row_smem : 8u * (uint32_t)kc * 2u);
                    __syncwarp();
                    uint8_t *dst = s.ring + (size_t)slot * kSlotBytes;
                    if (((job.x_src >> 8) & 1)) {
                        const size_t stage_bytes = 8 * (size_t)row_smem;
                        const uint8_t *g1 = reinterpret_cast<const uint8_t *>(job.w) +
                                            ((size_t)(rr.row0 / 8 + grp) * parts + pt) * stage_bytes;
                        if (lane == 0) tma_load_1d_hint(dst, g1, (uint32_t)stage_bytes, &s.full[slot], pol);
                        continue;
                    }
                    const uint8_t *g0 = src + (size_t)grp * 8 * row_bytes + (size_t)pt * kc * 2;
                    if (p.dbg & 2) {
                        if (lane == 0) tma_load_1d_hint(dst, g0, 8u * (uint32_t)kc * 2u, &s.full[slot], pol);
                        continue;
                    }
                    // lanes 0..7 issue one row each
                    if (lane < 8)
                        tma_load_1d_hint(dst + lane * row_smem, g0 + (size_t)lane * row_bytes, (uint32_t)kc * 2,
                                         &s.full[slot], pol);
                }
            }
        }
    }
}

struct StepCtx {
    int tok, pos;
    int ti;               // trace cursor (-1: tracing off for this step)
    int seq;              // GEMV phases started (gate value the producer waits for)
    uint32_t gs;          // ring stage counter, in lock-step with the producer
    float best;           // LOGITS: running (max, argmax) of this thread
    int best_i;
};

__device__ __forceinline__ void mega_trace(const MegaParams &p, int &ti, int bid, int tid) {
    if (ti >= 0 && tid == 0 && ti < 480) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.trace[(bid ? 1024 : 0) + ti++] = t;
    }
}

#include "mega_gemv.cuh"
#include "mega_attn.cuh"
#include "mega_kernel.cuh"

int mega_decode_launch(const MegaParams &p, cudaStream_t stream) {
    B200_REQUIRE(mega_supported(p.d, p.ffn, p.n_q, p.n_kv, p.head_dim), "mega: unsupported geometry");
    B200_REQUIRE(p.n_steps >= 1 && (p.sample || p.n_steps == 1), "mega: n_steps must be 1 when sampling is external");
    B200_REQUIRE(p.n_splits >= 1, "mega: n_splits < 1");
    B200_REQUIRE(4 * p.n_layers + 1 <= kMaxJobs, "mega: more than 32 layers");
    const int GQ = p.n_q / p.n_kv;
#define B200_MEGA_CASE(HD_, G_) \
    if (p.head_dim == HD_ && GQ == G_) return mega_launch_t<HD_, G_>(p, stream);
    B200_MEGA_CASE(128, 1) B200_MEGA_CASE(128, 2) B200_MEGA_CASE(128, 3) B200_MEGA_CASE(128, 4)
    B200_MEGA_CASE(64, 1) B200_MEGA_CASE(64, 2) B200_MEGA_CASE(64, 4)
#undef B200_MEGA_CASE
    set_error("mega: unsupported head_dim %d / group %d", p.head_dim, GQ);
    return B200LLM_EINVAL;
}

}  // namespace b200
