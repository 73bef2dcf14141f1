// decode_tc.cu -- the batch-1 decode step as ONE persistent kernel whose weight stream is
// consumed by the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM).
//
// decode_mega.cu consumes its ring with mma.sync from registers and is bound by instruction
// issue at about the HBM rate, so nothing the producer prefetches during a grid barrier is ever
// caught up on.  Here the consumer is one thread issuing tcgen05.mma straight from shared
// memory (M = 128 weight rows, N = 8, K = 16 per instruction, 4 KB of weights each; ~64 clk per
// instruction from one thread, scripts/ubench_umma.cu), i.e. ~120 GB/s per SM against the
// ~45 GB/s share of the HBM stream, so a ring that filled up during a barrier drains quickly:
//   warp 13      producer: walks the whole weight schedule (every phase of every step) and pulls
//                this CTA's share through a shared-memory ring of 16 KB stages (8-10 of them, what
//                the 227 KB leave beside the operand blocks: tcd_plan), one 1-D TMA bulk copy per
//                [128 rows x 64 k] tile.  The matrices are stored tile-major and pre-swizzled
//                (tcd.h), so a tile lands as a ready 128B-swizzled K-major operand and a CTA's
//                share of a matrix is one contiguous region of HBM.
//   warp 12      MMA issuer: A = weight tile from the ring, B = the activation vector as
//                [8 x 64 k] operand blocks (row 0 = bf16 hi part, row 1 = bf16 lo part, rows 2-7
//                zero; hi + lo is ~fp32 accurate), D[128, 8] fp32 in TMEM; tcgen05.commit hands
//                ring stages back to the producer and finished accumulators to the epilogue.
//                Its loop carries no integer division: one modulo per stage cost 5 % of the step.
//   warps 0-3    epilogue: tcgen05.ld of the warp's 32 accumulator lanes (= weight rows),
//                (col 0 + col 1) * norm scale, then red.global.add / store.
//   warps 0-11   phase prologues (activation vector -> operand blocks, fused RMSNorm / SwiGLU /
//                embedding / merge of split-attention partials), attention, sampling, the grid
//                barrier.
// Work split: the (row tile, k block) units of a matrix are numbered tile-major and cut into 148
// equal contiguous ranges (stream-K), so every CTA streams the same number of bytes (+-16 KB) in
// every phase; partial sums of a tile meet in global memory with fp32 red.add (the residual
// stream takes them directly, qkv and gate/up go to zeroed buffers whose consumers apply RoPE /
// SwiGLU on load).  The lm_head is cut at tile boundaries instead so that logits are plain stores.
// Tried and dropped: weights as the N = 256 operand, fully dynamic work
// distribution from an atomic counter, L2 prefetch beyond a full ring, two issuer threads.
#include "common.cuh"
#include "kernels.h"
#include "tcd.h"

namespace b200 {

constexpr int kTC = 384;              // compute threads
constexpr int kTW = kTC / 32;         // compute warps
constexpr int kTIssuers = 1;           // MMA issuer threads (one warp each): stage u of a segment goes to issuer u % kTIssuers.
                                       // Two issuers are ~3 % faster (a prefilled ring drains at 39 instead of 64 clk per MMA) but
                                       // rare launch failures / hangs appeared with them, so one is the default
constexpr int kTThreads = kTC + 32 * kTIssuers + 32;   // + MMA issuer warps + producer warp
constexpr int kTMaxStages = 10;        // ring stages: as many as fit beside the operand blocks (TcdParams::n_stages)
constexpr int kTStageBytes = 16384;   // [128 rows x 64 k] bf16
constexpr int kTBSlots = 64;          // most activation operand blocks ([8 x 64 k] bf16 = 1 KB each) a CTA may need (TcdParams::n_bslots)
constexpr int kTAcc = 8;              // accumulators in flight (16 TMEM columns apart)
constexpr int kTAccCols = 16;
constexpr int kTPT = 16;              // tokens per KV page
constexpr int kTMaxPtab = 448;        // page-table entries kept in shared memory
constexpr int kTMaxJobs = 4 * 32 + 1;

struct TcdSmem {
    uint8_t *ring, *bop;
    float *mrg;            // attention merge buffer [kTW][GQ][HD + 2]
    uint64_t *full, *empty, *acc_full, *acc_empty;
    uint32_t *tmem_slot;
    float *red;
    int *misc;
    float *rope;           // [head_dim/2][cos, sin] of the current position
    int32_t *ptab;
    TcdJob *jobs;
};

__host__ __device__ __forceinline__ size_t tcd_mrg_bytes(int hd, int gq) { return (size_t)kTW * gq * (hd + 2) * 4; }

// Layout: everything the compute warps address sits at COMPILE-TIME offsets from the base (merge buffer,
// barriers, scratch, page table, phase descriptors), so those pointers cost no registers; the operand blocks
// and the ring (1024-byte aligned, sizes known at launch) follow.
__device__ __forceinline__ TcdSmem tcd_carve(uint8_t *base, int hd, int gq, int n_jobs, int n_bslots) {
    TcdSmem s;
    s.mrg = reinterpret_cast<float *>(base);
    uint8_t *q = base + ((tcd_mrg_bytes(hd, gq) + 15) & ~(size_t)15);
    s.full = reinterpret_cast<uint64_t *>(q);
    s.empty = s.full + kTMaxStages;
    s.acc_full = s.empty + kTMaxStages;
    s.acc_empty = s.acc_full + kTAcc;
    s.tmem_slot = reinterpret_cast<uint32_t *>(s.acc_empty + kTAcc);
    s.red = reinterpret_cast<float *>(s.tmem_slot + 4);
    s.misc = reinterpret_cast<int *>(s.red + 64);
    s.rope = reinterpret_cast<float *>(s.misc + 64);
    s.ptab = reinterpret_cast<int32_t *>(s.rope + 128);
    s.jobs = reinterpret_cast<TcdJob *>(s.ptab + kTMaxPtab);
    const size_t head = (size_t)(reinterpret_cast<uint8_t *>(s.jobs) - base) + (size_t)n_jobs * sizeof(TcdJob);
    s.bop = base + ((head + 1023) & ~(size_t)1023);
    s.ring = s.bop + (size_t)n_bslots * 1024;
    return s;
}

static size_t tcd_fixed_smem_bytes(int hd, int gq, int n_jobs) {  // everything but the ring and the operand blocks
    return ((tcd_mrg_bytes(hd, gq) + 15) & ~(size_t)15) + (2 * kTMaxStages + 2 * kTAcc) * 8 + 16 + 64 * 4 + 64 * 4 + 128 * 4 +
           kTMaxPtab * 4 + (size_t)n_jobs * sizeof(TcdJob) + 2 * 1024;  // slack: 1024-byte alignment of the base and of the operand blocks
}
static size_t tcd_smem_bytes(int hd, int gq, int n_jobs, int n_stages, int n_bslots) {
    return (size_t)n_stages * kTStageBytes + (size_t)n_bslots * 1024 + tcd_fixed_smem_bytes(hd, gq, n_jobs);
}
constexpr size_t kTSmemLimit = 232448;  // 227 KB per CTA on sm_100

// This CTA's contiguous range [g0, g1) of the matrix's (row tile, k block) units, unit g = tile * nkb + kb.
struct TRange {
    int g0, g1, nkb;
};
__host__ __device__ __forceinline__ TRange tcd_range(int N, int K, bool whole_tiles, int G, int bid) {
    TRange r;
    r.nkb = K / 64;
    const int n_units = whole_tiles ? N / 128 : (N / 128) * r.nkb;
    const int base = n_units / G, rem = n_units % G;
    const int u0 = bid * base + (bid < rem ? bid : rem);
    const int u1 = u0 + base + (bid < rem ? 1 : 0);
    r.g0 = whole_tiles ? u0 * r.nkb : u0;
    r.g1 = whole_tiles ? u1 * r.nkb : u1;
    return r;
}

// operand blocks the CTA with the largest share needs for one matrix (0: the shape cannot be tiled)
int tcd_job_slots(int N, int K, bool whole_tiles) {
    if (N % 128 != 0 || K % 64 != 0 || N < 128 || K < 64) return 0;
    const TRange r = tcd_range(N, K, whole_tiles, kNumSMs, 0);  // CTA 0 has the largest share
    return (r.g1 - r.g0) < r.nkb ? (r.g1 - r.g0) : r.nkb;
}
bool tcd_job_fits(int N, int K, bool whole_tiles) {
    const int need = tcd_job_slots(N, K, whole_tiles);
    return need >= 1 && need <= kTBSlots;
}
int tcd_stages_for(int head_dim, int gq, int n_jobs, int n_bslots) {
    const size_t fixed = tcd_fixed_smem_bytes(head_dim, gq, n_jobs) + (size_t)n_bslots * 1024;
    if (fixed + 4 * (size_t)kTStageBytes > kTSmemLimit) return 0;
    const int st = (int)((kTSmemLimit - fixed) / kTStageBytes);
    return st > kTMaxStages ? kTMaxStages : st;
}

// ring stages and operand blocks of a model: the operand buffer is sized for the neediest matrix,
// the ring takes what is left of the 227 KB (false: not even 4 stages fit)
bool tcd_plan(int d, int ffn, int n_q, int n_kv, int head_dim, int vocab, int n_layers, int *n_stages, int *n_bslots) {
    const int qd = n_q * head_dim, qkv = (n_q + 2 * n_kv) * head_dim;
    const int shapes[5][2] = {{qkv, d}, {d, qd}, {2 * ffn, d}, {d, ffn}, {vocab, d}};
    int slots = 0;
    for (int i = 0; i < 5; ++i) {
        const int need = tcd_job_slots(shapes[i][0], shapes[i][1], i == 4);
        if (need < 1 || need > kTBSlots) return false;
        slots = need > slots ? need : slots;
    }
    const size_t fixed = tcd_fixed_smem_bytes(head_dim, n_kv > 0 ? n_q / n_kv : 1, 4 * n_layers + 1) + (size_t)slots * 1024;
    if (fixed + 4 * (size_t)kTStageBytes > kTSmemLimit) return false;
    int st = (int)((kTSmemLimit - fixed) / kTStageBytes);
    *n_stages = st > kTMaxStages ? kTMaxStages : st;
    *n_bslots = slots;
    return true;
}

bool tcd_supported(int d, int ffn, int n_q, int n_kv, int head_dim, int vocab, int n_layers) {
    const int G = n_kv > 0 ? n_q / n_kv : 0;
    if (!(head_dim == 64 || head_dim == 128) || G < 1 || G > 4 || (G == 3 && head_dim == 64)) return false;
    if (4 * n_layers + 1 > kTMaxJobs) return false;
    int st, sl;
    return tcd_plan(d, ffn, n_q, n_kv, head_dim, vocab, n_layers, &st, &sl);
}

// The in-kernel sampler keeps kTSampPerThread logit rows per compute thread and its scratch in the attention
// merge buffer; top-k beyond kTcdSampK (or "no top-k" with a temperature) stays on the stand-alone sample_kernel.
bool tcd_sampler_supported(int vocab_rows, int head_dim, int gq, float temperature, int top_k) {
    const int tiles = (vocab_rows + 127) / 128;
    const int rows_per_cta = (tiles + kNumSMs - 1) / kNumSMs * 128;
    if (rows_per_cta > 4 * kTC) return false;
    if (tcd_mrg_bytes(head_dim, gq) < 64 * 4 + 2 * kTW * 8 + kTcdSampK * 8 + kTcdSampK * 4) return false;
    return temperature <= 0.f || (top_k >= 1 && top_k <= kTcdSampK);
}

__device__ __forceinline__ bool tcd_whole(const TcdParams &p, int j, int n_jobs) {
    return j == n_jobs - 1 && !p.split_all;
}

// Grid barrier for the compute threads (producer and MMA warps never take part); see decode_mega.cu.
__device__ __forceinline__ unsigned tcd_poll(unsigned *ctr) {
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
    return v;
}
__device__ __forceinline__ void tcd_grid_sync(unsigned *ctr, unsigned target, int tid) {
    bar_sync(1, kTC);
    if (tid == 0) {
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
        // two polls in flight, half a round trip apart: the last arrival is seen ~0.2 us sooner than
        // with one dependent load after the other.  Relaxed polls and ONE fence at the end: an
        // acquire load per poll costs an L1 invalidate (CCTL.IVALL) per iteration
        unsigned a = tcd_poll(ctr);
        __nanosleep(300);   // (0 / 100 / 300 ns measure the same, 600 ns costs 1 %)
        for (;;) {
            const unsigned b = tcd_poll(ctr);
            if (a >= target) break;
            a = tcd_poll(ctr);
            if (b >= target) break;
        }
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
    }
    bar_sync(1, kTC);
}

// ---------------------------------------------------------------- producer: the weight stream
__device__ void tcd_producer(const TcdParams &p, const TcdSmem &s, int G, int bid) {
    const uint64_t pol = l2_policy_evict_first();
    const int n_jobs = 4 * p.n_layers + 1;
    // ring position: slot and the parity of its current use (no modulo of a run-time ring size in the loop)
    int slot = 0;
    uint32_t par = 0;
    bool first_round = true;
    for (int step = 0; step < p.n_steps; ++step) {
        for (int j = 0; j < n_jobs; ++j) {
            const TcdJob &job = s.jobs[j];
            const TRange r = tcd_range(job.N, job.K, tcd_whole(p, j, n_jobs), G, bid);
            const uint8_t *src = reinterpret_cast<const uint8_t *>(job.w) + (size_t)r.g0 * kTStageBytes;
            for (int g = r.g0; g < r.g1; ++g, src += kTStageBytes) {
                if (!first_round) mbar_wait(&s.empty[slot], par ^ 1);
                mbar_expect_tx(&s.full[slot], kTStageBytes);
                tma_load_1d_hint(s.ring + (size_t)slot * kTStageBytes, src, kTStageBytes, &s.full[slot], pol);
                if (++slot == p.n_stages) slot = 0, par ^= 1, first_round = false;
            }
        }
    }
}

// ---------------------------------------------------------------- MMA issuer
// K-major operand tile, 128-byte swizzle: rows of 128 bytes, 8-row groups `sbo` bytes apart
__device__ __forceinline__ uint64_t tcd_desc(uint32_t smem_addr, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(sbo >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__device__ __forceinline__ void tcd_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}

// Issuer `me` (0 or 1) takes every second stage of the CTA's range.  One thread sustains a
// tcgen05.mma every ~64 clk, two threads together one every ~39 clk (scripts/ubench_umma.cu), which
// is what drains a prefetched ring quickly after a barrier.  The issuers accumulate into separate
// column groups of the segment's accumulator (columns 0..7 and 8..15); the epilogue adds them.
__device__ void tcd_mma_warp(const TcdParams &p, const TcdSmem &s, int G, int bid, int lane, uint32_t tmem, int me) {
    const int n_jobs = 4 * p.n_layers + 1;
    // D fp32, A/B bf16, both K-major, M = 128, N = 8
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((8u >> 3) << 17) | ((128u >> 4) << 24);
    const uint32_t ring_a = smem_u32(s.ring), bop_a = smem_u32(s.bop);
    uint32_t as = 0;
    int slot = 0;      // ring position and the parity of its current use, in lock-step with the producer
    uint32_t par = 0;
    for (int step = 0; step < p.n_steps; ++step) {
        for (int j = 0; j < n_jobs; ++j) {
            const TcdJob &job = s.jobs[j];
            const TRange r = tcd_range(job.N, job.K, tcd_whole(p, j, n_jobs), G, bid);
            bar_sync(2, kTC + 32 * kTIssuers);  // the activation operand of this phase is in shared memory
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            // operand slot of the stage's k block: (kb - kb_start) mod nkb = stages since the range began, mod nkb
            // (kept incrementally: an integer division per stage would cost the issuer ~40 clk)
            int bslot = 0;
            for (int g = r.g0, t = r.g0 / r.nkb; lane == 0 && g < r.g1; ++t) {
                const int seg_end = min(r.g1, (t + 1) * r.nkb);
                const uint32_t aslot = as % kTAcc;
                if (as >= kTAcc) {
                    mbar_wait(&s.acc_empty[aslot], ((as / kTAcc) - 1) & 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                }
                const uint32_t d_addr = tmem + aslot * kTAccCols + me * 8;
                bool first = true;
                for (int u = 0; g < seg_end; ++g, ++u, bslot = (bslot + 1 == r.nkb ? 0 : bslot + 1),
                         par ^= (slot + 1 == p.n_stages), slot = (slot + 1 == p.n_stages ? 0 : slot + 1)) {
                    if ((u % kTIssuers) != me) continue;
                    mbar_wait(&s.full[slot], par);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint64_t adesc = tcd_desc(ring_a + (uint32_t)slot * kTStageBytes, 1024);
                    const uint64_t bdesc = tcd_desc(bop_a + (uint32_t)bslot * 1024, 0);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t accum = (first && k == 0) ? 0u : 1u;
                        asm volatile(
                            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_addr),
                            "l"(adesc + (uint64_t)(2 * k)), "l"(bdesc + (uint64_t)(2 * k)), "r"(idesc), "r"(accum)
                            : "memory");
                    }
                    first = false;
                    tcd_commit(&s.empty[slot]);  // frees the ring stage once the MMAs above have read it
                }
                if (first) {
                    mbar_arrive(&s.acc_full[aslot]);  // a one-stage segment: nothing for this issuer
                } else {
                    tcd_commit(&s.acc_full[aslot]);
                }
                ++as;
            }
            __syncwarp();
        }
    }
}

// Kept small: what lives across the whole step competes with the attention phase for the 128 registers, and a
// spilled value is reloaded behind a grid barrier at L2 latency.  (The tensor-parallel sequence numbers live in
// shared memory: s.misc[62] reductions, s.misc[63] token exchanges.)
struct TcdCtx {
    int tok, pos;
    int ti;               // trace cursor (-1: tracing off for this step)
    uint32_t as;          // accumulator counter, in lock-step with the MMA issuer
    float best;           // lm_head: running (max, argmax) of this thread
    int best_i;
};

// (debug) tagged globaltimer stamps of thread 0: tag in bits 48.., nanoseconds below
enum { TT_STEP = 1, TT_XCONV = 2, TT_XREADY = 3, TT_EPI = 4, TT_BAR = 5, TT_ATT_Q = 6, TT_ATT_KV = 7, TT_ATT_WARPS = 8,
       TT_ATT_DONE = 9 };
// Compiled in only with -DB200LLM_TCD_TRACE (make TRACE=1): the cursor and the stamps cost registers, and what the
// 128-register kernel spills it reloads right behind the grid barriers (an L2 round trip each, L1 was just invalidated).
__device__ __forceinline__ void tcd_trace(const TcdParams &p, int &ti, int bid, int tid, int tag) {
#ifndef B200LLM_TCD_TRACE
    (void)p, (void)ti, (void)bid, (void)tid, (void)tag;
    return;
#endif
    if (ti >= 0 && tid == 0 && ti < 1000) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.trace[(bid ? 1024 : 0) + ti++] = (t & 0xFFFFFFFFFFFFull) | ((unsigned long long)tag << 48);
    }
}

__device__ __forceinline__ uint32_t tcd_pack_bf16(float a, float b) {
    return (uint32_t)f32_to_bf16_bits(a) | ((uint32_t)f32_to_bf16_bits(b) << 16);
}
__device__ __forceinline__ float tcd_bf16_round(float a) { return __uint_as_float((uint32_t)f32_to_bf16_bits(a) << 16); }

// split-T factor of the attention phases at position `pos`: one half-page (one L2 round trip) per warp while splits last
__device__ __forceinline__ int tcd_att_ns(const TcdParams &p, int pos) {
    return min(p.n_splits, (2 * (pos / kTPT + 1) + kTW - 1) / kTW);
}

#include "decode_tc_phase.cuh"
#include "decode_tc_attn.cuh"
#include "decode_tc_sample.cuh"
#include "decode_tc_kernel.cuh"

}  // namespace b200
