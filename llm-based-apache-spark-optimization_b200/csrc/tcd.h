// tcd.h -- parameters of the tcgen05 persistent decode kernel (decode_tc.cu): one launch = n_steps decode steps of 1..4 lanes.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

// One weight-streaming phase: y (+)= W x over a whole matrix.  W is the TILE-MAJOR copy of the
// matrix: [N/128 row tiles][K/64 k blocks][128 rows][64 k] bf16, every 16 KB tile stored the way
// a 128-byte-swizzled K-major tcgen05 operand lies in shared memory, so that one 1-D bulk copy
// drops a ready-to-use MMA operand into the ring (weights.py: tile_major()).
struct TcdJob {
    const void *w;        // tile-major bf16
    const void *gain;     // bf16 [K] RMSNorm gain fused in front, or NULL
    void *k_pages;        // qkv job: this layer's K / V page regions (used by the attention phase)
    void *v_pages;
    float *out;           // fp32 [N]: partial sums are added (TCD_EPI_ADD) or rows stored (TCD_EPI_LOGITS)
    int32_t N, K;         // N % 128 == 0, K % 64 == 0
    int32_t epi;          // TCD_EPI_*
    int32_t x_src;        // TCD_X_* in the low byte; bit 8: this phase also clears the qkv and gate/up sum buffers
};
enum { TCD_EPI_ADD = 0, TCD_EPI_LOGITS = 1, TCD_EPI_ADD_CNT = 2 };   // ADD_CNT: out is float2 [N]: (partial sum, 1) is added, see TcdParams::dataflow
enum { TCD_X_H = 0, TCD_X_ATT = 1, TCD_X_SWIGLU = 2, TCD_X_EMBED = 3 };

struct TcdParams {
    const TcdJob *jobs;   // device array: 4 per layer (qkv, o, gate_up, down) + lm_head
    int32_t n_layers;
    int32_t d, n_q, n_kv, head_dim, ffn, vocab;
    int32_t kv_page_heads;   // (kv head, 16-token) tiles from one KV page to the next: n_kv in a pool of this engine's own, more in a shared pool
    float rms_eps;
    float scale_log2;     // log2(e) / sqrt(head_dim)
    // activations (fp32), one row per lane (lane b at base + b * row size, see `nb` below)
    float *h;             // residual stream [d]
    float *qkv;           // raw q|k|v sums of the current layer [(n_q + 2 n_kv) * head_dim], q/k rows pair-interleaved
    float *att;           // attention output [n_q * head_dim]
    float *gu;            // raw gate/up sums [2 ffn], interleaved
    float *logits;        // [vocab]
    float *part_max;      // [grid]
    int32_t *part_idx;    // [grid]
    const void *embed;    // bf16 [vocab, d]
    const float *inv_freq;
    // lane state
    int32_t lane;
    int32_t *lane_token, *lane_pos, *lane_hist;
    const int32_t *page_table;  // this lane's row
    int32_t n_ptab;
    int32_t *gen_out;     // [n_steps]
    int32_t n_steps;
    // attention split partials
    float *part_o;        // [n_q, n_splits, head_dim]
    float *part_ml;       // [n_q, n_splits, 2]
    unsigned *tickets;    // [n_kv]
    int32_t n_splits;
    unsigned *bar_ctr;    // grid barrier counter (zeroed before every launch)
    int32_t sample;       // 1: greedy argmax + feedback inside the kernel; 0: stop after the logits;
                          // 2: the sampler chain inside the kernel (decode_tc_sample.cuh): repeat penalty over the lane's
                          //    last 64 tokens -> top-k (1..64) -> top-p -> temperature -> draw; temperature 0 = argmax of
                          //    the penalised logits
    float temperature, top_p, repeat_penalty;
    int32_t top_k;
    uint64_t seed;
    unsigned long long *cand;   // [grid][64] per-CTA candidate keys of the sampler (sample == 2)
    int32_t n_stages;     // ring stages of 16 KB (<= 10) and operand blocks of 1 KB (<= 64) in shared memory: tcd_plan()
    int32_t n_bslots;
    int32_t n_poll;       // warps whose lane 0 polls the grid-barrier counter (1..12)
    // Data-flow synchronisation (tp_size 1): a phase whose bit is set adds (partial sum, 1.0f) pairs with ONE vector
    // red per row instead of a bare fp32 red, and its consumer polls the pairs of the rows it needs until each count
    // equals the number of CTAs that contribute to the row's tile: the data says when it is complete, so the grid
    // barrier between producer and consumer phase (release fence + counter + poll, then the re-load: three dependent
    // L2 round trips) becomes one polling load.  bit 0: gate/up sums -> down_proj.  bit 1: attention -> o_proj: the
    // attention items (one writer per value) store (value, tag) pairs, tag = step * n_layers + layer + 1 of this launch
    // (the host clears the buffer before a launch), and o_proj's prologue polls the pairs of its k-slices for the tag.
    // bit 2: q|k|v sums -> attention: a CTA with an attention item polls the counts of its head's rows, then runs the item.
    int32_t dataflow;
    float2 *gu2;          // [2 ffn] (sum, count) pairs of the gate/up phase
    float2 *qkv2;         // [(n_q + 2 n_kv) head_dim] (sum, count) pairs of the qkv phase
    float2 *att2;         // [n_q head_dim] (value, tag) | part_o2 [n_q, n_splits, head_dim] (value, tag) | part_ml2 [n_q, n_splits] float4 (max, sum, tag, -)
    int32_t tmem_slots;   // weight stages (0..12) the issuer may park in tensor memory while it waits (decode_tc.cu)
    int32_t poll_ns;      // pause of the issuer between two probes while it waits for a phase's operand
    int32_t n_inflight;   // bulk copies a CTA keeps in flight (0: the whole ring)
    int32_t split_all;    // (test hook) k-split the last job too instead of giving whole tiles
    // ---- several lanes in one launch (nb = 2..4, tp_size 1): the N = 8 operand block has room for four (hi, lo) row pairs, so
    // the weights stream ONCE for up to four sequences.  Lane b's activations, sums, partials, logits and candidates live at
    // base + b * (the buffer's size for one lane); `lane`, `page_table` and the sampler settings above describe lane 0 and the
    // arrays below all lanes.  gen_out is [n_steps][nb].
    int32_t nb;
    int32_t lanes[4];
    const int32_t *page_tables[4];
    float temperature_l[4], top_p_l[4], repeat_penalty_l[4];
    int32_t top_k_l[4];
    uint64_t seed_l[4];
    int32_t part_stride;   // floats from one lane's (part_o | part_ml) block to the next
    int32_t att2_stride;   // float2 from one lane's att2 block to the next
    // ---- tensor parallel (tp_size > 1): row-parallel phases (bit 10 of x_src) add into `delta`, then an
    // all-reduce phase pushes delta to the peers' inboxes and adds all ranks' partials to h (see tp_ar.cu)
    int32_t tp_rank, tp_size;
    int32_t vocab_global;  // token ids are clamped against this; `vocab` is this rank's share
    float *delta;          // [d] this rank's partial sums of a row-parallel phase
    float *ar_inbox[8];    // inbox base of every rank: [2 phases][tp_size][ar_slot_floats]
    unsigned *ar_flags[8]; // flag base of every rank: [2 phases][tp_size][ar_max_slices]
    unsigned *ar_seq;      // local: [0] reduction sequence number (shared with tp_ar.cu), [2] token-exchange sequence
    unsigned long long *tok_inbox[8];  // per rank: [2 phases][tp_size] {max as bits | token id << 32}, and
    unsigned *tok_flags[8];            // [2 phases][tp_size] sequence words
    int32_t ar_slot_floats, ar_max_slices;
    unsigned long long *trace;  // optional [2][1024] globaltimer stamps of CTA 0 and CTA G-1 (debug)
    int32_t trace_step;
};

int tcd_decode_launch(const TcdParams &p, cudaStream_t stream);
bool tcd_supported(int d, int ffn, int n_q, int n_kv, int head_dim, int vocab, int n_layers);
bool tcd_plan(int d, int ffn, int n_q, int n_kv, int head_dim, int vocab, int n_layers, int *n_stages, int *n_bslots, int lanes = 1);
int tcd_job_slots(int N, int K, bool whole_tiles);
int tcd_stages_for(int head_dim, int gq, int n_jobs, int n_bslots);
// per-phase check that the activation operand of one CTA fits its 64 k-block buffer
bool tcd_job_fits(int N, int K, bool whole_tiles);
// can the in-kernel sampler (sample == 2) take these settings for a lm_head of `vocab_rows` (padded) rows?
bool tcd_sampler_supported(int vocab_rows, int head_dim, int gq, float temperature, int top_k);
constexpr int kTcdMaxLanes = 4; // lanes one launch of the persistent kernel can decode together
constexpr int kTcdSampK = 64;   // largest top-k of the in-kernel sampler = candidate keys per CTA in TcdParams::cand

}  // namespace b200
