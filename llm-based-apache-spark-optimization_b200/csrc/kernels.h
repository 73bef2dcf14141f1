// kernels.h -- internal launch functions of libb200llm (host side, C++).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200llm.h"

namespace b200 {

// gemv.cu
int gemv_launch(const b200llm_gemv_args &a, cudaStream_t stream, bool pdl);
int gemv_grid(int N, int K, int bs, int epi, int grid_hint);

// attn.cu
int attn_decode_launch(const b200llm_attn_args &a, cudaStream_t stream, bool pdl, void *out_bf16 = nullptr,
                       int64_t out_lo_off = 0);
// prefill chunk (consecutive rows of one lane) at short contexts: 4 rows per CTA share every K/V tile (attn.cu)
int attn_rows_launch(const b200llm_attn_args &a, cudaStream_t stream, bool pdl, void *out_bf16, int64_t out_lo_off);
size_t attn_scratch_bytes(int bs, int n_q, int n_kv, int head_dim, int n_splits);
// The arrival tickets live in a fixed-size region at the start of the scratch buffer, so that
// launches with different batch sizes sharing one buffer never put partials where another
// launch keeps tickets (tickets must stay zero between launches).
constexpr size_t kAttnTicketBytes = 256 * 1024;  // 65536 (row, kv head) tickets

// attn_prefill.cu: causal attention of `rows` consecutive positions of one lane (prefill chunk) on tcgen05; map_k / map_v are
// tc_make_map(base of the layer's K / V region, rows of 16-token page tiles, head_dim, box 16)
int attn_prefill_launch(const void *map_k, const void *map_v, const float *q, float *out, void *out_bf16, int64_t out_lo_off,
                        const int32_t *page_table_row, int n_ptab, int rows, int pos0, int n_q, int n_kv, int head_dim,
                        int kv_page_heads, cudaStream_t stream, bool pdl);

// misc.cu
struct EmbedArgs {
    const void *embed;        // bf16 [vocab, d]
    float *h;                 // fp32 [rows, d]
    int32_t *row_token;       // [rows] (decode: filled from lane state)
    int32_t *row_pos;         // [rows]
    const int32_t *row_lane;  // [rows]
    const int32_t *lane_token;  // [lanes]
    const int32_t *lane_pos;    // [lanes]
    int rows, d, vocab;
    int from_lanes;           // 1: decode (gather token/pos from lane state); 0: prefill (row_* already set)
};
int embed_launch(const EmbedArgs &a, cudaStream_t stream, bool pdl);
// engine variant: also records the token in the lane's 64-entry history ring and bumps the step counter
int embed_launch_ex(const EmbedArgs &a, int32_t *lane_hist, int32_t *step_ctr, cudaStream_t stream, bool pdl);

struct SampleDev {
    b200llm_sample_args a;
    // engine-side feedback (all may be NULL for the bare op)
    const int32_t *row_lane;  // [bs]
    int32_t *lane_token;      // [lanes]  <- sampled token
    int32_t *lane_pos;        // [lanes]  += 1
    int32_t *lane_hist;       // [lanes, 64] ring of recent tokens
    int32_t *gen_out;         // [steps, bs] ring of generated tokens (row-major by step)
    const int32_t *step_ctr;  // device step counter (index into gen_out), or NULL
    int32_t *step_ctr_w;      // incremented by row 0's CTA
    const uint64_t *rng_ctr;  // device RNG counter or NULL (then a.step is used)
    int advance;              // 1: write feedback
    int n_ranks;              // > 1: partials are [rank][row][n_part] (vocab-parallel all-gather)
    size_t rank_stride;
};
int sample_launch(const SampleDev &a, cudaStream_t stream, bool pdl);

// gemm_tc.cu: tcgen05 GEMM  out[t, n] (+)= sum_k X[t, k] W[n, k]  (bf16 in, fp32 out), T <= 256
int tc_make_map(void *map_out_128B, const void *base, int64_t rows, int64_t cols, int box_rows, int bk = 64);
int tc_bk(int T, bool has_lo);       // k per stage (64 or 32) tc_gemm_launch should be given for T rows: the maps' box is [box_rows x bk]
int tc_cluster_size(int N, int T);   // CTAs per cluster tc_gemm_launch will use
int tc_x_box(int N, int T);          // rows of the activation maps' box for that launch: ceil16(T) / cluster size
int tc_gemm_launch(const void *map_w, const void *map_x, float *out, int64_t out_stride, int T, int N, int K,
                   int accumulate, cudaStream_t stream, bool pdl, void *out_swiglu = nullptr,
                   const void *map_x_lo = nullptr, int64_t swiglu_lo_off = 0, int bk = 64,
                   const void *w_tile_major = nullptr);   // tile-major copy of W (tcd.h): weight tiles as 16 KB bulk copies; map_w unused
// dequant.cu: GGUF blocks (ggml type id) -> bf16, bit-equal to rounding ggml's reference de-quantiser
int dequant_bf16_launch(const void *blocks, int ggml_type, int64_t n_elements, void *out_bf16, cudaStream_t stream);
// tp_ar.cu: one-shot all-reduce + residual add over NVLink peer memory (h += sum over ranks of delta)
constexpr int kTpMaxRanks = 8;
int tp_allreduce_add_launch(float *h, const float *delta, int n, float *const *inbox, unsigned *const *flags,
                            unsigned *seq, int rank, int size, int slot_floats, int max_slices, cudaStream_t stream);
// batch_ops.cu
int rmsnorm_bf16_launch(const float *x, const void *gain, void *out, int rows, int d, float eps, cudaStream_t s, bool pdl,
                        float *zero = nullptr, int zero_w = 0, int64_t lo_off = 0);
int cast_bf16_launch(const float *x, int64_t in_stride, void *out, int rows, int n, cudaStream_t s, bool pdl, int64_t lo_off = 0);
int swiglu_bf16_launch(const float *gu, void *out, int rows, int ffn, cudaStream_t s, bool pdl, int64_t lo_off = 0);
int rope_append_launch(const float *qkv, float *q, void *k_pages, void *v_pages, const float *inv_freq,
                       const int32_t *row_lane, const int32_t *row_pos, const int32_t *page_table, int rows, int n_q,
                       int n_kv, int hd, int page_tokens, int max_pages, cudaStream_t s, bool pdl, int kv_page_heads = 0);

int fill_weights_launch(void *dev_bf16, uint64_t n, uint64_t seed, float sigma, float mean, cudaStream_t s);
int add_inplace_launch(float *dst, const float *src, size_t n, cudaStream_t s, bool pdl);

}  // namespace b200
