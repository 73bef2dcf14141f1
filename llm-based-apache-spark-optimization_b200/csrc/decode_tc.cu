// decode_tc.cu -- a whole decode call (every phase of every step, 1..4 lanes) as ONE persistent kernel whose weight
// stream is consumed by the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM).
//
// decode_mega.cu consumes its ring with mma.sync from registers and is bound by instruction issue at about the HBM rate,
// so nothing the producer prefetches during a grid barrier is ever caught up on.  Here the consumer is a warp issuing
// tcgen05.mma straight from shared memory (M = 128 weight rows, N = 8, K = 16 per instruction, 4 KB of weights each;
// ~39 clk per instruction at best, scripts/ubench_umma.cu / ubench_issue.cu), several times this SM's share of the HBM
// stream, so a ring that filled up during a barrier drains quickly.  148 CTAs x 480 threads:
//   warp 14      producer: walks the whole weight schedule (every phase of every step) and pulls this CTA's share
//                through a shared-memory ring of 16 KB stages (an even number, 8 for the 3B: what the 227 KB leave
//                beside the operand blocks, tcd_plan), one 1-D TMA bulk copy per [128 rows x 64 k] tile.  The matrices
//                are stored tile-major and pre-swizzled (tcd.h), so a tile lands as a ready 128B-swizzled K-major
//                operand and a CTA's share of a matrix is one contiguous region of HBM.
//   warps 12-13  MMA issuers (kTIssuers): stage i of the CTA's stream belongs to issuer i % 2 and ring slot i % n_stages.
//                A = weight tile from the ring, B = the activations as [8 x 64 k] operand blocks (rows 2b / 2b + 1 =
//                bf16 hi / lo parts of lane b's activations, hi + lo ~ fp32 accurate; unused rows zero), D[128, 8] fp32
//                in TMEM (columns 2b / 2b + 1 = lane b); tcgen05.commit hands ring stages back to the producer and
//                finished accumulators to the epilogue.  The whole warp walks the loop converged and an elected lane
//                issues, with operands in uniform registers; no integer division in the loop.
//   warps 0-3    epilogue: tcgen05.ld of the warp's 32 accumulator lanes (= weight rows), (hi + lo column) * norm
//                scale per lane, then red.global.add (plain, or (sum, 1) pairs where the consumer polls) / store.
//   warps 0-11   phase prologues (activations -> operand blocks, fused RMSNorm / SwiGLU / embedding / merge of
//                split-attention partials), attention, argmax or the sampler chain, grid barriers.
// Work split: the (row tile, k block) units of a matrix are numbered tile-major and cut into 148 equal contiguous ranges
// (stream-K), so every CTA streams the same number of bytes (+-16 KB) in every phase; partial sums of a tile meet in
// global memory with fp32 red.add (the residual stream takes them directly, qkv and gate/up go to zeroed buffers whose
// consumers apply RoPE / SwiGLU on load).  The lm_head is cut at tile boundaries instead so that logits are plain stores.
// Synchronisation between phases is by data where a phase has one consumer (TcdParams::dataflow), else a grid barrier.
// Tried and dropped: weights as the N = 256 operand, fully dynamic work distribution from an atomic
// counter, L2 prefetch beyond a full ring, tensor memory as ring overflow.
#include "common.cuh"
#include "kernels.h"
#include "tcd.h"

// The file is compiled eight times (Makefile): part 0 holds the host-side planning functions and the dispatcher, parts
// 1..7 each instantiate the kernel for one (head_dim, q heads per kv head) pair -- the instantiations take minutes
// each, in parallel they take as long as one.
#ifndef B200_TCD_PART
#define B200_TCD_PART 0
#endif

namespace b200 {

constexpr int kTC = 384;              // compute threads
constexpr int kTW = kTC / 32;         // compute warps
constexpr int kTIssuers = 2;           // MMA issuer warps.  Stage i of the CTA's weight stream (counted over the whole launch) goes to
                                       // issuer i % 2 and lands in ring slot i % n_stages; n_stages is EVEN, so every ring slot belongs
                                       // to one issuer for good.  That matters: an mbarrier waiter tells phases apart by parity only, so
                                       // it has to see every phase of a barrier it waits on.  (Round 1 gave an issuer every second stage
                                       // of each SEGMENT; with an odd ring a slot changed hands, an issuer far ahead of the producer --
                                       // it is held back only by the 8 accumulators -- took a barrier still in its previous phase for
                                       // "landed", multiplied stale bytes and freed the stage early: rare faults in short phases, a
                                       // certain one in the 336-stage lm_head phase.)
constexpr int kTThreads = kTC + 32 * kTIssuers + 32;   // + MMA issuer warps + producer warp
// (the optional TMEM overflow queue is kept by ONE issuer: it is compiled in only when kTIssuers == 1)
constexpr int kTMaxStages = 10;        // ring stages: as many as fit beside the operand blocks (TcdParams::n_stages)
constexpr int kTStageBytes = 16384;   // [128 rows x 64 k] bf16
constexpr int kTBSlots = 64;          // most activation operand blocks ([8 x 64 k] bf16 = 1 KB each) a CTA may need (TcdParams::n_bslots)
constexpr int kTAcc = 8;              // accumulators in flight (16 TMEM columns apart)
constexpr int kTAccCols = 16;
// Overflow of the weight ring, in TENSOR MEMORY.  While the MMA issuer waits for a phase's activation operand (grid
// barrier + activation re-load, or the whole attention phase) the shared-memory ring is full and the HBM stream of
// this SM stalls.  Instead of idling, the issuer moves landed stages out of shared memory with tcgen05.cp
// (4 x 128x256b = the [128 x 64] bf16 tile as 128 lanes x 32 columns, the layout tcgen05.mma takes an A operand
// in from TMEM) and hands the shared-memory stage back to the producer, so the stream keeps running: 512 columns -
// 128 of accumulators = 12 more stages of 16 KB, ~4.3 us of this SM's share of the stream.  Moved stages are
// multiplied from TMEM, in order, before the ones still in shared memory (scripts/ubench_tmem_a.cu checks copy + MMA
// bit for bit; a copy costs as many tcgen05 issue slots as the MMAs of its stage, so stages are only moved while
// the issuer has nothing to multiply).
constexpr int kTSlots = 12;           // TMEM weight slots
constexpr int kTSlotCols = 32;
constexpr int kTSlotCol0 = kTAcc * kTAccCols;
constexpr int kTTmemCols = 512;
static_assert(kTSlotCol0 + kTSlots * kTSlotCols <= kTTmemCols, "tensor memory over-committed");
constexpr int kTPT = 16;              // tokens per KV page
constexpr int kTMaxPtab = 448;        // page-table entries kept in shared memory
constexpr int kTMaxJobs = 4 * 32 + 1;

struct TcdSmem {
    uint8_t *ring, *bop;
    float *mrg;            // attention merge buffer [kTW][GQ][HD + 2]
    float *qst;            // q | k | v sums of this CTA's attention item [(GQ + 2) * HD] (data-flow mode)
    uint64_t *full, *empty, *acc_full, *acc_empty, *tempty, *xrdy;
    uint32_t *tmem_slot;
    float *red;
    int *misc;
    float *rope;           // [head_dim/2][cos, sin] of the current position
    int32_t *ptab;
    TcdJob *jobs;
};

__host__ __device__ __forceinline__ size_t tcd_mrg_bytes(int hd, int gq) { return (size_t)kTW * gq * (hd + 2) * 4; }
__host__ __device__ __forceinline__ size_t tcd_qst_bytes(int hd, int gq) { return (size_t)(gq + 2) * hd * 4; }

// Layout: everything the compute warps address sits at COMPILE-TIME offsets from the base (merge buffer,
// barriers, scratch, page table, phase descriptors), so those pointers cost no registers; the operand blocks
// and the ring (1024-byte aligned, sizes known at launch) follow.
__device__ __forceinline__ TcdSmem tcd_carve(uint8_t *base, int hd, int gq, int n_jobs, int n_bslots, int lb) {   // lb: lanes the kernel instance can hold (1 or kTcdMaxLanes)
    TcdSmem s;
    s.mrg = reinterpret_cast<float *>(base);
    s.qst = reinterpret_cast<float *>(base + ((tcd_mrg_bytes(hd, gq) + 15) & ~(size_t)15));
    uint8_t *q = reinterpret_cast<uint8_t *>(s.qst) + tcd_qst_bytes(hd, gq);
    s.full = reinterpret_cast<uint64_t *>(q);
    s.empty = s.full + kTMaxStages;
    s.acc_full = s.empty + kTMaxStages;
    s.acc_empty = s.acc_full + kTAcc;
    s.tempty = s.acc_empty + kTAcc;
    s.xrdy = s.tempty + kTSlots;
    s.tmem_slot = reinterpret_cast<uint32_t *>(s.xrdy + 2);
    s.red = reinterpret_cast<float *>(s.tmem_slot + 4);
    s.misc = reinterpret_cast<int *>(s.red + 128);
    s.rope = reinterpret_cast<float *>(s.misc + 64);
    s.ptab = reinterpret_cast<int32_t *>(s.rope + 128 * lb);           // per lane: [head_dim/2][cos, sin]
    s.jobs = reinterpret_cast<TcdJob *>(s.ptab + kTMaxPtab * lb);      // per lane: the first kTMaxPtab page-table entries
    // (room for the most phase descriptors a launch may have, so that the operand blocks too sit at a compile-time offset:
    // with the offset depending on n_jobs, ptxas spilled the pointer and every compute thread reloaded it once per phase)
    (void)n_jobs;
    const size_t head = (size_t)(reinterpret_cast<uint8_t *>(s.jobs) - base) + (size_t)kTMaxJobs * sizeof(TcdJob);
    s.bop = base + ((head + 1023) & ~(size_t)1023);
    s.ring = s.bop + (size_t)n_bslots * 1024;
    return s;
}

static size_t tcd_fixed_smem_bytes(int hd, int gq, int n_jobs, int lb = 1) {  // everything but the ring and the operand blocks
    return ((tcd_mrg_bytes(hd, gq) + 15) & ~(size_t)15) + tcd_qst_bytes(hd, gq) + (2 * kTMaxStages + 2 * kTAcc + kTSlots + 2) * 8 + 16 + 128 * 4 + 64 * 4 + 128 * 4 * lb +
           kTMaxPtab * 4 * lb + (size_t)kTMaxJobs * sizeof(TcdJob) + 2 * 1024 + 0 * (size_t)n_jobs;  // slack: 1024-byte alignment of the base and of the operand blocks
}
static size_t tcd_smem_bytes(int hd, int gq, int n_jobs, int n_stages, int n_bslots, int lb = 1) {
    return (size_t)n_stages * kTStageBytes + (size_t)n_bslots * 1024 + tcd_fixed_smem_bytes(hd, gq, n_jobs, lb);
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

#if B200_TCD_PART == 0
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
    return (st > kTMaxStages ? kTMaxStages : st) & ~(kTIssuers - 1);   // a multiple of the issuer count (see kTIssuers)
}

// ring stages and operand blocks of a model: the operand buffer is sized for the neediest matrix,
// the ring takes what is left of the 227 KB (false: not even 4 stages fit)
bool tcd_plan(int d, int ffn, int n_q, int n_kv, int head_dim, int vocab, int n_layers, int *n_stages, int *n_bslots, int lanes) {
    const int qd = n_q * head_dim, qkv = (n_q + 2 * n_kv) * head_dim;
    const int shapes[5][2] = {{qkv, d}, {d, qd}, {2 * ffn, d}, {d, ffn}, {vocab, d}};
    int slots = 0;
    for (int i = 0; i < 5; ++i) {
        const int need = tcd_job_slots(shapes[i][0], shapes[i][1], i == 4);
        if (need < 1 || need > kTBSlots) return false;
        slots = need > slots ? need : slots;
    }
    const size_t fixed = tcd_fixed_smem_bytes(head_dim, n_kv > 0 ? n_q / n_kv : 1, 4 * n_layers + 1, lanes > 1 ? kTcdMaxLanes : 1) + (size_t)slots * 1024;
    if (fixed + 4 * (size_t)kTStageBytes > kTSmemLimit) return false;
    int st = (int)((kTSmemLimit - fixed) / kTStageBytes);
    *n_stages = (st > kTMaxStages ? kTMaxStages : st) & ~(kTIssuers - 1);
    *n_bslots = slots;
    return true;
}

bool tcd_supported(int d, int ffn, int n_q, int n_kv, int head_dim, int vocab, int n_layers) {
    const int G = n_kv > 0 ? n_q / n_kv : 0;
    if (!(head_dim == 64 || head_dim == 128) || G < 1 || G > 4 || (G == 3 && head_dim == 64)) return false;
    if (4 * n_layers + 1 > kTMaxJobs) return false;
    int st, sl;
    return tcd_plan(d, ffn, n_q, n_kv, head_dim, vocab, n_layers, &st, &sl, 1);
}

// The in-kernel sampler keeps kTSampPerThread logit rows per compute thread and its scratch in the attention
// merge buffer; top-k beyond kTcdSampK (or "no top-k" with a temperature) stays on the stand-alone sample_kernel.
bool tcd_sampler_supported(int vocab_rows, int head_dim, int gq, float temperature, int top_k) {
    const int tiles = (vocab_rows + 127) / 128;
    const int rows_per_cta = (tiles + kNumSMs - 1) / kNumSMs * 128;
    if (rows_per_cta > 4 * kTC) return false;
    if (temperature > 0.f && (top_k < 1 || top_k > kTcdSampK)) return false;
    // scratch in the attention merge buffer (s.mrg, s.qst): recent tokens, top-k keys and weights, one list of k keys per warp (decode_tc_sample.cuh)
    const int k = temperature > 0.f ? top_k : 1;
    // (the staging buffer of the attention items follows the merge buffer and is idle as well)
    const size_t room = ((tcd_mrg_bytes(head_dim, gq) + 15) & ~(size_t)15) + tcd_qst_bytes(head_dim, gq);
    return room >= 64 * 4 + (size_t)kTcdSampK * 8 + (size_t)kTcdSampK * 4 + (size_t)kTW * k * 8;
}

#endif  // part 0

__device__ __forceinline__ bool tcd_whole(const TcdParams &p, int j, int n_jobs) {
    return j == n_jobs - 1 && !p.split_all;
}

// Grid barrier for the compute threads (producer and MMA warps never take part); see decode_mega.cu.
__device__ __forceinline__ unsigned tcd_poll(unsigned *ctr) {
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
    return v;
}
// Two halves, so that the trace build can stamp between them: arrive = CTA barrier + thread 0's release (a fence that
// waits for the CTA's global writes to be performed, then the add); wait = poll, one fence, CTA barrier.
__device__ __forceinline__ void tcd_grid_arrive(unsigned *ctr, int tid) {
    bar_sync(1, kTC);
    if (tid == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
}
__device__ __forceinline__ void tcd_grid_wait(unsigned *ctr, unsigned target, int tid, volatile unsigned *done, int n_poll) {
    // Lane 0 of the first n_poll warps polls, each with one relaxed load in flight (measured: more than one poller per
    // CTA only adds contention on the counter's line, profiles/sweep_poll_r2.txt); the finder fences once (an acquire
    // load per poll would cost an L1 invalidate, CCTL.IVALL, per iteration) and tells the others through shared memory.
    if ((tid & 31) == 0 && (tid >> 5) < n_poll) {
        if (tid) __nanosleep((unsigned)(tid >> 5) * 1200u / (unsigned)n_poll);
        for (;;) {
            if (*done >= target) break;
            if (tcd_poll(ctr) >= target) {
                asm volatile("fence.acq_rel.gpu;" ::: "memory");
                *done = target;
                break;
            }
        }
    }
    bar_sync(1, kTC);
}
__device__ __forceinline__ void tcd_grid_sync(unsigned *ctr, unsigned target, int tid, volatile unsigned *done, int n_poll) {
    tcd_grid_arrive(ctr, tid);
    tcd_grid_wait(ctr, target, tid, done, n_poll);
}

// ---------------------------------------------------------------- producer: the weight stream
static __device__ void tcd_producer(const TcdParams &p, const TcdSmem &s, int G, int bid) {
    const uint64_t pol = l2_policy_evict_first();
    const int n_jobs = 4 * p.n_layers + 1;
    // ring position: slot and the parity of its current use (no modulo of a run-time ring size in the loop)
    int slot = 0;
    uint32_t par = 0;
    bool first_round = true;
    // in-flight limit: copy i is issued only once copy i - n_inflight has landed.  A full ring of copies in flight on
    // every SM is 3x what the HBM pipe needs (Little: ~6.4 MB at ~1 us) and queues in front of the latency-bound
    // traffic of the phase boundaries (barrier polls, activation re-loads)
    const int lim = (p.n_inflight > 0 && p.n_inflight < p.n_stages) ? p.n_inflight : 0;
    int wslot = 0, ahead = 0;
    uint32_t wpar = 0;
    for (int step = 0; step < p.n_steps; ++step) {
        for (int j = 0; j < n_jobs; ++j) {
            const TcdJob &job = s.jobs[j];
            const TRange r = tcd_range(job.N, job.K, tcd_whole(p, j, n_jobs), G, bid);
            const uint8_t *src = reinterpret_cast<const uint8_t *>(job.w) + (size_t)r.g0 * kTStageBytes;
            for (int g = r.g0; g < r.g1; ++g, src += kTStageBytes) {
                if (lim) {
                    if (ahead == lim) {
                        mbar_wait(&s.full[wslot], wpar);
                        if (++wslot == p.n_stages) wslot = 0, wpar ^= 1;
                    } else {
                        ++ahead;
                    }
                }
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
// one lane of a converged warp
__device__ __forceinline__ bool tcd_elect() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.b32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tcd_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}

// The MMA issuer warp.  A = a weight tile (in its ring stage, or in a TMEM slot if it was moved there while the
// issuer waited), B = the activation operand block in shared memory, D[128, 8] fp32 in TMEM.
// The WHOLE warp walks the schedule, converged, and one elected lane issues the tcgen05 instructions: every operand
// (descriptors, TMEM addresses, barrier addresses) is then the same in all lanes and derives from kernel parameters
// only, so ptxas keeps it in uniform registers.  Run by a single lane of a diverged warp, the same code cost
// ~145 clk per MMA (each UTCHMMA sat in an ELECT / 5 x R2UR.BROADCAST / BRA.U.ANY loop over "possibly different"
// lane values): 0.3 us per 16 KB stage, as slow as this SM's share of the HBM stream delivers them, so nothing the
// ring had prefetched during a barrier was ever caught up on (profiles/tcd_trace_r2a.txt).
// The loop carries no integer division.
struct TcdIssuer {
    int slot;          // ring stage that leaves shared memory next (moved or multiplied), and the parity of its use
    uint32_t par;
    int t_head, t_n;   // TMEM queue: oldest filled slot, number of filled slots
    int t_tail;        // next TMEM slot to fill (slots are filled and multiplied in ring order)
    uint32_t t_use;    // completed laps of t_tail = earlier fills of slot t_tail (phase of its tempty barrier)
};

// shape of phase j's matrix from the kernel parameters (the descriptors in shared memory hold the same numbers, but a
// value that went through a load is not provably warp-uniform)
__device__ __forceinline__ void tcd_job_shape(const TcdParams &p, int j, int n_jobs, int &N, int &K) {
    const int q = j & 3;
    if (j == n_jobs - 1) {
        N = (p.vocab + 127) / 128 * 128, K = p.d;
    } else if (q == 0) {
        N = (p.n_q + 2 * p.n_kv) * p.head_dim, K = p.d;
    } else if (q == 1) {
        N = p.d, K = p.n_q * p.head_dim;
    } else if (q == 2) {
        N = 2 * p.ffn, K = p.d;
    } else {
        N = p.d, K = p.ffn;
    }
}

// The four K = 16 MMAs of one [128 x 64] stage.  Descriptors come as (low, high) words: stepping 32 bytes along K only
// touches the 14-bit address field of the low word (tiles are 1024-byte aligned), so the step is a 32-bit immediate add.
__device__ __forceinline__ void tcd_mma4_ss(uint32_t d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi, uint32_t idesc,
                                            uint32_t first) {
#pragma unroll
    for (int k = 0; k < 4; ++k)
        asm volatile(
            "{\n\t.reg .pred p;\n\t.reg .b64 ra, rb;\n\tsetp.eq.b32 p, %6, 0;\n\t"
            "mov.b64 ra, {%1, %2};\n\tmov.b64 rb, {%3, %4};\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], ra, rb, %5, p;\n\t}" ::"r"(d),
            "r"(a_lo + 2 * k), "r"(a_hi), "r"(b_lo + 2 * k), "r"(b_hi), "r"(idesc), "r"(k == 0 ? first : 0u)
            : "memory");
}
__device__ __forceinline__ void tcd_mma4_ts(uint32_t d, uint32_t ta, uint32_t b_lo, uint32_t b_hi, uint32_t idesc, uint32_t first) {
#pragma unroll
    for (int k = 0; k < 4; ++k)
        asm volatile(
            "{\n\t.reg .pred p;\n\t.reg .b64 rb;\n\tsetp.eq.b32 p, %5, 0;\n\t"
            "mov.b64 rb, {%2, %3};\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], rb, %4, p;\n\t}" ::"r"(d),
            "r"(ta + 8 * k), "r"(b_lo + 2 * k), "r"(b_hi), "r"(idesc), "r"(k == 0 ? first : 0u)
            : "memory");
}

__device__ __forceinline__ void tcd_cp_stage(uint32_t ta, uint64_t adesc) {
#pragma unroll
    for (int k = 0; k < 4; ++k)
        asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(ta + 8 * k), "l"(adesc + (uint64_t)(2 * k)) : "memory");
}

// Move the oldest landed ring stage into a free TMEM slot, if there is one of each.  Never blocks.  (Converged warp.)
__device__ __forceinline__ bool tcd_try_move(const TcdParams &p, const TcdSmem &s, TcdIssuer &q, uint32_t tmem, uint32_t ring_a,
                                             int &left) {
    if (left == 0 || q.t_n >= p.tmem_slots) return false;
    if (!mbar_test_wait(&s.full[q.slot], q.par)) return false;
    // the MMAs that read this slot's previous tile were issued by this warp before (tcgen05 ops run in issue order);
    // the barrier check keeps the re-use safe whatever the pipeline does
    if (q.t_use > 0 && !mbar_test_wait(&s.tempty[q.t_tail], (q.t_use - 1) & 1)) return false;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (tcd_elect()) {
        tcd_cp_stage(tmem + kTSlotCol0 + (uint32_t)q.t_tail * kTSlotCols, tcd_desc(ring_a + (uint32_t)q.slot * kTStageBytes, 1024));
        tcd_commit(&s.empty[q.slot]);   // the ring stage goes back to the producer once the copies have read it
    }
    __syncwarp();
    if (++q.slot == p.n_stages) q.slot = 0, q.par ^= 1;
    if (++q.t_tail == kTSlots) q.t_tail = 0, ++q.t_use;
    ++q.t_n;
    --left;
    return true;
}

#ifdef B200LLM_TCD_TRACE
#define TCD_ISTAMP(tag_)                                                                                          \
    do {                                                                                                          \
        if (tr_on && lane == 0 && tr_n < 500) {                                                                   \
            unsigned long long t_;                                                                                \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                                \
            p.trace[1536 + tr_n++] = (t_ & 0xFFFFFFFFFFFFull) | ((unsigned long long)(tag_) << 48);               \
        }                                                                                                         \
    } while (0)
#ifdef B200LLM_TCD_TRACE_DEEP   // per-stage cycle accounting and stamps: slows the issuer loop itself
#define TCD_ICLK(var_) const long long var_ = clock64()
#define TCD_IACC(i_, v_) tr_acc[i_] += (v_)
#define TCD_ISTAMP_STAGE(tag_) TCD_ISTAMP(tag_)
#else
#define TCD_ICLK(var_) do { } while (0)
#define TCD_IACC(i_, v_) do { } while (0)
#define TCD_ISTAMP_STAGE(tag_) do { } while (0)
#endif
#else
#define TCD_ISTAMP(tag_) do { } while (0)
#define TCD_ISTAMP_STAGE(tag_) do { } while (0)
#define TCD_ICLK(var_) do { } while (0)
#define TCD_IACC(i_, v_) do { } while (0)
#endif

static __device__ void tcd_mma_warp(const TcdParams &p, const TcdSmem &s, int G, int bid, int lane, uint32_t tmem_in, int me) {
    (void)lane;
#ifdef B200LLM_TCD_TRACE
    bool tr_on = false;
    int tr_n = 0;
    long long tr_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const long long tr_t0 = clock64();
#endif
    const int n_jobs = 4 * p.n_layers + 1;
    // D fp32, A/B bf16, both K-major, M = 128, N = 8
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((8u >> 3) << 17) | ((128u >> 4) << 24);
    const uint32_t tmem = __shfl_sync(0xffffffffu, tmem_in, 0);   // (read from shared memory: make it warp-uniform)
    const uint32_t ring_a = smem_u32(s.ring), bop_a = smem_u32(s.bop);
    // descriptor words (tcd_desc): address >> 4 | 1 << 16 below; 8-row group stride >> 4, version 1, 128-byte swizzle above
    const uint32_t a_lo0 = ((ring_a & 0x3FFFFu) >> 4) | (1u << 16), a_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
    const uint32_t b_lo0 = ((bop_a & 0x3FFFFu) >> 4) | (1u << 16), b_hi = (1u << 14) | (2u << 29);
    uint32_t as = 0;
    TcdIssuer q;
    q.slot = 0, q.par = 0, q.t_head = 0, q.t_n = 0, q.t_use = 0, q.t_tail = 0;
    // stages the producer will still deliver (the mover must not wait for a copy that is never issued)
    int left = 0;
    for (int j = 0; j < n_jobs; ++j) {
        int N, K;
        tcd_job_shape(p, j, n_jobs, N, K);
        const TRange r = tcd_range(N, K, tcd_whole(p, j, n_jobs), G, bid);
        left += r.g1 - r.g0;
    }
    left *= p.n_steps;
    const bool use_tmem = kTIssuers == 1 && p.tmem_slots > 0;
    uint32_t xph = 0;
    uint32_t turn = 0;   // stream index of the next stage, mod kTIssuers: whose stage it is
    for (int step = 0; step < p.n_steps; ++step) {
        for (int j = 0; j < n_jobs; ++j) {
            int N, K;
            tcd_job_shape(p, j, n_jobs, N, K);
            const TRange r = tcd_range(N, K, tcd_whole(p, j, n_jobs), G, bid);
#ifdef B200LLM_TCD_TRACE
            tr_on = p.trace != nullptr && bid == 0 && me == 0 && step == p.trace_step && j >= 4 && j < 8;
#endif
            TCD_ICLK(c0);
            // wait for the activation operand of this phase; meanwhile (one issuer only) keep this SM's weight stream moving
            if (use_tmem) {
                while (!mbar_test_wait(s.xrdy, xph & 1))
                    if (!tcd_try_move(p, s, q, tmem, ring_a, left) && p.poll_ns > 0) __nanosleep((unsigned)p.poll_ns);
            } else {
                mbar_wait(s.xrdy, xph & 1);
            }
            ++xph;
            __syncwarp();
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            TCD_ICLK(c1);
            TCD_IACC(1, c1 - c0);
            TCD_ISTAMP(10);
            // operand slot of the stage's k block: (kb - kb_start) mod nkb = stages since the range began, mod nkb
            // (kept incrementally: an integer division per stage would cost the issuer ~40 clk)
            int bslot = 0;
            for (int g = r.g0, t = r.g0 / r.nkb; g < r.g1; ++t) {
                const int seg_end = min(r.g1, (t + 1) * r.nkb);
                const uint32_t aslot = as % kTAcc;
                TCD_ICLK(c2);
                if (as >= kTAcc) {
                    mbar_wait(&s.acc_empty[aslot], ((as / kTAcc) - 1) & 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                }
                TCD_ICLK(c3);
                TCD_IACC(3, c3 - c2);
                // each issuer accumulates its stages of the segment into its own column group (columns 0..7 / 8..15 of the
                // segment's accumulator) and the epilogue adds the groups; a one-stage segment goes to group 0 whoever has it
                const uint32_t d_addr = tmem + aslot * kTAccCols + (seg_end - g > 1 ? (uint32_t)me * 8u : 0u);
                uint32_t first = 1u;
                for (; g < seg_end; ++g, turn = (turn + 1 == kTIssuers ? 0 : turn + 1), bslot = (bslot + 1 == r.nkb ? 0 : bslot + 1)) {
                    const uint32_t b_lo = b_lo0 + (uint32_t)bslot * 64u;
                    if (kTIssuers == 1 && q.t_n > 0) {   // the oldest stage was moved to TMEM (its copies precede these MMAs in issue order)
                        if (tcd_elect()) {
                            tcd_mma4_ts(d_addr, tmem + kTSlotCol0 + (uint32_t)q.t_head * kTSlotCols, b_lo, b_hi, idesc, first);
                            tcd_commit(&s.tempty[q.t_head]);   // the slot may be refilled once these MMAs have read it
                        }
                        if (++q.t_head == kTSlots) q.t_head = 0;
                        --q.t_n;
                        first = 0u;
                        continue;
                    }
                    if (kTIssuers == 1 || turn == (uint32_t)me) {
                        TCD_ICLK(c4);
                        mbar_wait(&s.full[q.slot], q.par);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        TCD_ICLK(c5);
                        TCD_IACC(0, c5 - c4);
                        TCD_IACC(4, 1);
                        TCD_ISTAMP_STAGE(11);
                        if (tcd_elect()) {
                            tcd_mma4_ss(d_addr, a_lo0 + (uint32_t)q.slot * 1024u, a_hi, b_lo, b_hi, idesc, first);
                            tcd_commit(&s.empty[q.slot]);  // frees the ring stage once the MMAs above have read it
                        }
                        first = 0u;
                    }
                    if (++q.slot == p.n_stages) q.slot = 0, q.par ^= 1;
                    --left;
                }
                if (tcd_elect()) {
                    if (first) {
                        mbar_arrive(&s.acc_full[aslot]);   // a one-stage segment: nothing for this issuer
                    } else {
                        tcd_commit(&s.acc_full[aslot]);
                    }
                }
                __syncwarp();
                TCD_ISTAMP_STAGE(13);
                ++as;
            }
            TCD_ISTAMP(12);
        }
    }
#ifdef B200LLM_TCD_TRACE
    if (p.trace != nullptr && bid == 0 && lane == 0 && me == 0) {
#ifdef B200LLM_TCD_TRACE_DEEP
        tr_acc[6] = clock64() - tr_t0;
        tr_acc[2] = tr_acc[6] - tr_acc[0] - tr_acc[1] - tr_acc[3];
        for (int i = 0; i < 7; ++i) p.trace[4000 + i] = (unsigned long long)tr_acc[i];
#endif
    }
#endif
}

// Kept small: what lives across the whole step competes with the attention phase for the 128 registers, and a
// spilled value is reloaded behind a grid barrier at L2 latency.  (The tensor-parallel sequence numbers live in
// shared memory: s.misc[62] reductions, s.misc[63] token exchanges.)
// NB = lanes the kernel instance decodes together (1, or kTcdMaxLanes for 2..4).  With one lane the token and position
// live in registers; with several they live in shared memory (misc[16 + b], misc[20 + b]): tcd_pos / tcd_tok.
template <int NB>
struct TcdCtxT {
    int tok, pos;         // (NB == 1)
    int ti;               // trace cursor (-1: tracing off for this step)
    uint32_t as;          // accumulator counter, in lock-step with the MMA issuer
    float best[NB];       // lm_head: running (max, argmax) of this thread, per lane
    int best_i[NB];
};
template <int NB>
__device__ __forceinline__ int tcd_pos(const TcdCtxT<NB> &c, const TcdSmem &s, int b) {
    if constexpr (NB == 1) return c.pos;
    else return reinterpret_cast<volatile int *>(s.misc)[16 + b];
}
template <int NB>
__device__ __forceinline__ int tcd_tok(const TcdCtxT<NB> &c, const TcdSmem &s, int b) {
    if constexpr (NB == 1) return c.tok;
    else return reinterpret_cast<volatile int *>(s.misc)[20 + b];
}
// lane b's first page-table entries (shared memory) and its whole table (global)
template <int NB>
__device__ __forceinline__ int tcd_page(const TcdParams &p, const TcdSmem &s, int b, int pg) {
    if constexpr (NB == 1) return pg < kTMaxPtab ? s.ptab[pg] : p.page_table[pg];
    else return pg < kTMaxPtab ? s.ptab[b * kTMaxPtab + pg] : p.page_tables[b][pg];
}

// (debug) tagged globaltimer stamps of thread 0: tag in bits 48.., nanoseconds below
enum { TT_STEP = 1, TT_XCONV = 2, TT_XREADY = 3, TT_EPI = 4, TT_BAR = 5, TT_ATT_Q = 6, TT_ATT_KV = 7, TT_ATT_WARPS = 8,
       TT_ATT_DONE = 9, TT_BAR_ARRIVED = 14 };
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
