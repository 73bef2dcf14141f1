// mega.h -- parameters of the persistent batch-1 decode kernel (decode_mega.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

// One weight-streaming phase (a GEMV over a whole matrix, rows split over the CTAs).
struct MegaJob {
    const void *w;        // bf16 [N, K]
    const void *gain;     // bf16 [K] RMSNorm gain fused in front, or NULL
    void *k_pages;        // QKV_ROPE: this layer's K / V page regions
    void *v_pages;
    int32_t N, K;
    int32_t epi;          // B200LLM_EPI_*
    int32_t x_src;        // MEGA_X_* in the low byte; bit 8: w is the stage-major copy (one bulk copy per stage)
};
enum { MEGA_X_H = 0, MEGA_X_ATT = 1, MEGA_X_ACT = 2, MEGA_X_EMBED = 3 };

struct MegaParams {
    const MegaJob *jobs;  // device array: 4 per layer (qkv, o, gate_up, down) + lm_head
    int32_t n_layers;
    int32_t d, n_q, n_kv, head_dim, ffn, vocab;
    int32_t kv_page_heads;   // (kv head, 16-token) tiles from one KV page to the next (n_kv, or more in a shared pool)
    float rms_eps;
    float scale_log2;     // log2(e) / sqrt(head_dim)
    // activations (fp32), all one row (batch 1)
    float *h, *q, *att, *act, *logits;
    float *part_max;      // [grid]
    int32_t *part_idx;    // [grid]
    const void *embed;    // bf16 [vocab, d]
    const float *inv_freq;
    // lane state
    int32_t lane;
    int32_t *lane_token, *lane_pos, *lane_hist;
    const int32_t *page_table;  // this lane's row
    int32_t n_ptab;       // entries of it that are valid for this call
    int32_t *gen_out;     // [n_steps]
    int32_t n_steps;
    // attention split partials
    float *part_o;        // [n_q, n_splits, head_dim]
    float *part_ml;       // [n_q, n_splits, 2]
    unsigned *tickets;    // [n_kv]
    int32_t n_splits;
    // grid barrier counter (zeroed before every launch)
    unsigned *bar_ctr;
    int32_t sample;       // 1: greedy argmax + feedback inside the kernel; 0: stop after logits
    unsigned long long *trace;  // optional [2][1024] globaltimer stamps of CTA 0 and CTA G-1 (debug)
    int32_t trace_step;
    int32_t dbg;          // debug bits: 1 = consumer skips the MMAs, 2 = one bulk copy per stage (results invalid)
    int32_t gap_slots;    // ring stages the producer may prefetch for a phase before its consumers hold their activations
    int32_t ring_slots;   // ring stages in use (<= 5); fewer = shallower prefetch (tuning knob)
};

int mega_decode_launch(const MegaParams &p, cudaStream_t stream);
size_t mega_smem_bytes(int head_dim);
bool mega_supported(int d, int ffn, int n_q, int n_kv, int head_dim);

}  // namespace b200
