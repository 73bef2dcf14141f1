/*
 * b200llm.h -- C ABI of the B200-native Llama decode engine (libb200llm.so).
 *
 * Drop-in boundary.  The reference (Rajwardhan0511/LLM-Based-Apache-Spark-Optimization)
 * reaches the decode path only through
 *     res = ollama.generate(model=..., system=..., prompt=...);  res.response
 * at six call sites: Flask/app.py:102-107 and :160-166, FastAPI/app.py:85-90 and
 * :105-111, Model_Evaluation_&_Comparision.py:23-45 and :114-120.  Everything below
 * that call (Ollama server + llama.cpp: tokenise -> prefill -> decode loop -> sampler)
 * is replaced by this library; the Python package `ollama/` shipped in this repo binds
 * these entry points with ctypes.
 *
 * Conventions
 *   - plain C, no C++/torch types; every function returns 0 on success or a negative
 *     B200LLM_E* code; b200llm_last_error() gives the message of the calling thread's
 *     last failure (valid until that thread's next failing call).
 *   - all device buffers (weights, KV pool, workspace) are allocated by the host
 *     (PyTorch) and only BORROWED: the engine never frees them and the caller keeps
 *     them alive until b200llm_destroy().  The engine owns its stream, events and
 *     CUDA graphs only.
 *   - one engine = one GPU (one process per GPU; tensor-parallel groups pass
 *     tp_rank/tp_size and an NCCL unique id).
 *   - there is no CPU fallback: every entry point fails with B200LLM_ENODEV when no
 *     sm_100 device is present.
 */
#ifndef B200LLM_H_
#define B200LLM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200LLM_ABI_VERSION 2

enum {
    B200LLM_OK = 0,
    B200LLM_EINVAL = -1,   /* bad argument / unsupported shape            */
    B200LLM_ENODEV = -2,   /* no sm_100 device, or wrong device current   */
    B200LLM_ECUDA = -3,    /* a CUDA runtime call or kernel failed        */
    B200LLM_ESTATE = -4,   /* call order violated (weights not bound ...) */
    B200LLM_ENCCL = -5,    /* NCCL missing or an NCCL call failed         */
    B200LLM_ENOMEM = -6    /* caller-provided buffer too small            */
};

typedef struct b200llm_engine b200llm_engine;

/* Geometry of the whole (un-sharded) model + this rank's engine limits.
 * Replaces: the model Ollama resolves from the `model=` string of
 * ollama.generate (Flask/app.py:103,161; FastAPI/app.py:86,106). */
typedef struct b200llm_config {
    int32_t abi_version;        /* B200LLM_ABI_VERSION */
    int32_t d;                  /* hidden size */
    int32_t n_layers;
    int32_t n_q;                /* query heads (global) */
    int32_t n_kv;               /* key/value heads (global) */
    int32_t head_dim;           /* 64 or 128 */
    int32_t ffn;                /* MLP inner size (global) */
    int32_t vocab;              /* global vocabulary */
    float rms_eps;
    int32_t page_tokens;        /* tokens per KV page (16) */
    int32_t max_lanes;          /* concurrent sequences the engine tracks */
    int32_t max_pages_per_lane; /* page-table width: max context = this * page_tokens */
    int32_t max_rows;           /* max rows (tokens) of one forward pass: >= max decode batch and prefill chunk */
    int32_t tp_rank, tp_size;   /* tensor-parallel placement (1 = replica) */
    int32_t device;             /* CUDA ordinal; must be the current device of the calling thread */
    int32_t use_graphs;         /* capture the decode step into a CUDA graph */
    int32_t use_pdl;            /* programmatic dependent launch between kernels */
    int32_t attn_splits;        /* split-T factor of decode attention, 0 = auto */
    int32_t use_megakernel;     /* batch-1 greedy decode as one persistent kernel */
    int32_t use_tc;             /* passes of more than 4 rows (batches, prefill chunks) on the tcgen05 GEMM:
                                   activations rounded to bf16 at every matmul input */
    void *stream;               /* cudaStream_t to run on; NULL = engine creates its own */
    const void *nccl_unique_id; /* 128-byte ncclUniqueId when tp_size > 1, else NULL */
} b200llm_config;

/* Packed weight tensors (bf16 unless noted; row-major [rows, cols]).  Layout is the
 * engine's own (DESIGN.md "Data layout in HBM"); b200llm.weights.pack() builds it
 * from HF-named tensors. */
enum {
    B200LLM_W_EMBED = 0,      /* [vocab, d]                       global            */
    B200LLM_W_LM_HEAD = 1,    /* [vocab/tp, d]   rows of this rank (may alias EMBED) */
    B200LLM_W_FINAL_NORM = 2, /* [d]                                                 */
    B200LLM_W_INV_FREQ = 3,   /* fp32 [head_dim/2]                global             */
    B200LLM_W_ATTN_NORM = 4,  /* [d]                              per layer          */
    B200LLM_W_QKV = 5,        /* [(n_q+2 n_kv)/tp * head_dim, d]  per layer; q,k rows pair-interleaved */
    B200LLM_W_O = 6,          /* [d, n_q/tp * head_dim]           per layer          */
    B200LLM_W_MLP_NORM = 7,   /* [d]                              per layer          */
    B200LLM_W_GATE_UP = 8,    /* [2 ffn/tp, d]  row 2i = gate_i, row 2i+1 = up_i     */
    B200LLM_W_DOWN = 9,       /* [d, ffn/tp]                      per layer          */
    /* Optional second copies for the persistent batch-1 kernel, "stage-major": the matrix cut into
     * stages of 8 rows x kc columns (kc = K / parts, parts = smallest p >= ceil(K/2048) dividing
     * K/32), stored [row group][part][8][kc + 8] so that one stage -- rows already skewed by 16
     * bytes for conflict-free ldmatrix -- is one contiguous bulk copy.  rows/cols in the
     * descriptor are the logical [N, K]. */
    B200LLM_W_QKV_ST = 10,
    B200LLM_W_O_ST = 11,
    B200LLM_W_GATE_UP_ST = 12,
    B200LLM_W_DOWN_ST = 13,
    B200LLM_W_LM_HEAD_ST = 14,
    /* Optional copies for the tcgen05 persistent batch-1 kernel, "tile-major": the matrix cut into
     * [128 rows x 64 k] tiles stored [N/128][K/64][128][64], each 16 KB tile laid out as a
     * 128-byte-swizzled K-major tensor-core operand (16-byte chunk c of row r at chunk c ^ (r & 7)),
     * so one bulk copy lands a ready MMA operand.  Needs N % 128 == 0 and K % 64 == 0; rows/cols
     * in the descriptor are the logical [N, K].  The lm_head alone may have a row count that is
     * no multiple of 128 (vocab 32016; 32000 over 4 ranks): its copy is padded with zero rows to
     * whole tiles and its descriptor carries the PADDED row count; the kernel neither stores nor
     * ranks the padding rows.  When all five are bound they take precedence over the stage-major
     * copies. */
    B200LLM_W_QKV_TM = 15,
    B200LLM_W_O_TM = 16,
    B200LLM_W_GATE_UP_TM = 17,
    B200LLM_W_DOWN_TM = 18,
    B200LLM_W_LM_HEAD_TM = 19
};

typedef struct b200llm_tensor_desc {
    int32_t kind;   /* B200LLM_W_* */
    int32_t layer;  /* -1 for global tensors */
    const void *dev;
    int64_t rows, cols;
} b200llm_tensor_desc;

/* Sampler chain of one lane.  Replaces: Ollama's sampler (Modelfile PARAMETERs /
 * `options=`); temperature 0 = greedy argmax with lowest-id tie-break (parity mode). */
typedef struct b200llm_sampling {
    float temperature;    /* 0 = greedy */
    int32_t top_k;        /* <= 0: 1024 candidates */
    float top_p;          /* >= 1: off */
    float repeat_penalty; /* 1 = off; applied to the last 64 tokens of the lane */
    uint64_t seed;
} b200llm_sampling;

typedef struct b200llm_timing {
    float prefill_ms;      /* device time of the last b200llm_prefill  */
    float decode_ms;       /* device time of the last b200llm_decode   */
    int32_t prefill_tokens;
    int32_t decode_steps;
    int64_t kernel_launches; /* kernels of this library launched by the last decode call */
} b200llm_timing;

const char *b200llm_last_error(void);
int b200llm_abi_version(void);

/* ---- engine life cycle ---------------------------------------------------------- */
int b200llm_create(const b200llm_config *cfg, b200llm_engine **out);
int b200llm_destroy(b200llm_engine *e);
/* Rank 0 of a tensor-parallel group calls this and hands the 128 bytes to the other
 * ranks (any host channel); every rank passes them in b200llm_config.nccl_unique_id. */
int b200llm_nccl_unique_id(void *out128);

/* One-shot all-reduce over NVLink peer memory for tensor parallel (optional; without it the
 * engine uses ncclAllReduce).  Every rank calls b200llm_tp_export, which allocates the rank's
 * inbox region and returns its 64-byte CUDA IPC handle; the host gathers the handles of all
 * ranks (any channel) and passes them, in rank order, to b200llm_tp_import on every rank.  From
 * then on the two reductions per layer (after o_proj and down_proj) are one kernel each that
 * pushes the rank's partial sums into the peers' inboxes and adds what the peers pushed.
 * All ranks must be processes on one node whose GPUs have peer access (NVLink / NVSwitch). */
int b200llm_tp_export(b200llm_engine *e, void *handle64);
int b200llm_tp_import(b200llm_engine *e, const void *handles, int32_t n_ranks);
/* Bytes of scratch (activations, logits, attention split partials) the host must bind. */
int b200llm_workspace_bytes(const b200llm_config *cfg, size_t *bytes);
/* Bytes of one KV page across all layers of this rank: n_layers * 2 * (n_kv/tp) * page_tokens * head_dim * 2. */
int b200llm_kv_page_bytes(const b200llm_config *cfg, size_t *bytes);
int b200llm_bind_weights(b200llm_engine *e, const b200llm_tensor_desc *descs, int32_t n);
int b200llm_bind_kv_pool(b200llm_engine *e, void *dev, size_t bytes);
/* A KV pool that several engines share (the reference serves two models from one process: duckdb-nsql at
 * Flask/app.py:102-107 and llama3.2 at :160-166; BASELINE.json config 5 co-locates both on every GPU).  The pool is
 * cut into slots of `slot_bytes`, a page of this engine is ONE slot (page id = slot index, page i at dev + i *
 * slot_bytes), laid out page-major: [layer][K|V][kv head][page_tokens][head_dim].  slot_bytes must be a multiple of
 * page_tokens * head_dim * 2 and at least b200llm_kv_page_bytes.  Engines with different page sizes share one pool by
 * agreeing on a block size and taking slot_bytes = block / k each (host side: b200llm/kv_pool.py); the library never
 * allocates or frees pages, the host's allocator decides which slots a lane's page table names. */
int b200llm_bind_kv_pool_shared(b200llm_engine *e, void *dev, size_t bytes, size_t slot_bytes);
int b200llm_bind_workspace(b200llm_engine *e, void *dev, size_t bytes);

/* ---- sequence ("lane") state ------------------------------------------------------ */
/* Host-side page allocator decides the pages; this copies lane's table (n entries). */
int b200llm_set_page_table(b200llm_engine *e, int32_t lane, const int32_t *pages, int32_t n);
int b200llm_set_sampling(b200llm_engine *e, int32_t lane, const b200llm_sampling *s);
/* Overwrite the lane's next input token / position (teacher forcing, resume). */
int b200llm_set_lane(b200llm_engine *e, int32_t lane, int32_t next_token, int32_t pos);
/* Reset the lane's 64-token repeat-penalty window and seed it with ids[0..n) as the tokens at positions
 * pos0 .. pos0+n-1 (n <= 64).  Used when a lane starts on KV pages computed by another request (prefix
 * cache): the window must hold what a full prefill of the same prompt would have left in it. */
int b200llm_set_history(b200llm_engine *e, int32_t lane, const int32_t *ids, int32_t n, int32_t pos0);

/* Tensor parallel, measurement only (no counterpart in the reference): time `iters` back-to-back all-reduces of a
 * [rows, d] fp32 block through the engine's own reduction path (the one-shot kernel over NVLink peer memory once
 * b200llm_tp_import has run, else ncclAllReduce + add).  Collective: every rank of the group calls it with the same
 * arguments.  The residual stream is used as scratch.  bench.py reports the all-reduce share of a decode step from it. */
int b200llm_debug_tp_allreduce_us(b200llm_engine *e, int32_t rows, int32_t iters, float *us_per_call);

/* ---- the hot path ---------------------------------------------------------------- */
/* Prefill: run `n` prompt ids (host buffer) of `lane` starting at position pos0, append
 * their K/V to the lane's pages, sample the first new token into the lane. Asynchronous. */
int b200llm_prefill(b200llm_engine *e, int32_t lane, const int32_t *ids, int32_t n, int32_t pos0);
/* Decode: n_steps autoregressive steps for the given lanes as one batch; sampled tokens
 * feed back on the device (no host round trip between steps). Asynchronous. */
int b200llm_decode(b200llm_engine *e, const int32_t *lanes, int32_t n_lanes, int32_t n_steps);
/* Wait and copy out the tokens produced by the last decode call: out[step * n_lanes + i]. */
int b200llm_read_tokens(b200llm_engine *e, int32_t *out, int32_t n_steps, int32_t n_lanes);
/* Wait and read the first token sampled by the last prefill of `lane`. */
int b200llm_read_lane(b200llm_engine *e, int32_t lane, int32_t *next_token, int32_t *pos);
/* Test hook for logits parity: fp32 logits (this rank's vocab rows) of batch row `row`
 * of the last forward pass. */
int b200llm_step_logits(b200llm_engine *e, int32_t row, float *out, int32_t n);
int b200llm_sync(b200llm_engine *e);
int b200llm_last_timing(b200llm_engine *e, b200llm_timing *t);
/* The cudaStream_t the engine launches on (for CUDA-event timing by the caller). */
void *b200llm_stream(b200llm_engine *e);
/* Debug: phase timestamps (ns, %globaltimer) the persistent decode kernel records when the
 * environment variable B200LLM_MEGA_TRACE is set: [2][1024] for the first and last CTA. */
int b200llm_debug_read_trace(b200llm_engine *e, uint64_t *out, int32_t n);

/* ---- single kernels (unit-test and profiling entry points; same kernels the engine runs) */

/* bf16 weight GEMV  y[b, n] = sum_k W[n, k] * x[b, k],  fp32 accumulate, bs <= 4.
 * Optional fused RMSNorm prologue (norm_gain != NULL) and epilogue `epi`. */
enum {
    B200LLM_EPI_STORE = 0,    /* y -> out fp32 [bs, N]                                   */
    B200LLM_EPI_RESIDUAL = 1, /* out[b, n] += y                                          */
    B200LLM_EPI_SWIGLU = 2,   /* rows (2i, 2i+1) = (gate, up): out[b, i] = silu(g) * u   */
    B200LLM_EPI_QKV_ROPE = 3, /* RoPE on q/k pairs; q -> out fp32; k, v -> bf16 KV pages */
    B200LLM_EPI_LOGITS = 4    /* y -> out fp32 and per-CTA (max, argmax) partials        */
};
typedef struct b200llm_gemv_args {
    const void *w;           /* bf16 [N, K] */
    const float *x;          /* fp32 [bs, x_stride] */
    int64_t x_stride;
    const void *norm_gain;   /* bf16 [K] or NULL */
    float rms_eps;
    float *out;              /* fp32 [bs, out_stride] */
    int64_t out_stride;
    int32_t N, K, bs, epi;
    /* QKV_ROPE only */
    int32_t n_q, n_kv, head_dim, page_tokens, max_pages_per_lane;
    const float *inv_freq;   /* fp32 [head_dim/2] */
    const int32_t *row_lane; /* [bs] */
    const int32_t *row_pos;  /* [bs] */
    const int32_t *page_table; /* [lanes, max_pages_per_lane] */
    void *k_pages;           /* bf16 [pages, n_kv, page_tokens, head_dim] of this layer */
    void *v_pages;
    /* LOGITS only */
    float *part_max;         /* fp32 [bs, grid] */
    int32_t *part_idx;       /* int32 [bs, grid] */
    int32_t logits_idx_offset; /* added to the argmax ids (vocab-parallel shards) */
    int32_t grid_hint;       /* 0 = auto */
    int32_t kv_page_heads;   /* QKV_ROPE: (kv head, 16-token) tiles from one page to the next; 0 = n_kv (pages are
                              * [n_kv, page_tokens, head_dim] and lie back to back); larger in a shared pool, see
                              * b200llm_bind_kv_pool_shared */
    int32_t reserved0;
} b200llm_gemv_args;
int b200llm_op_gemv(const b200llm_gemv_args *a, void *stream);
/* Number of CTAs b200llm_op_gemv will launch for (N, K, bs, epi): size of the LOGITS partials. */
int b200llm_op_gemv_grid(int32_t N, int32_t K, int32_t bs, int32_t epi, int32_t *grid);

/* One weight-streaming phase of the tcgen05 persistent decode kernel on its own (test hook):
 * out[N] = W x with W given tile-major (B200LLM_W_*_TM layout), x and out fp32 on the device.
 * k_split = 0 cuts the matrix at row-tile boundaries (the lm_head mode, plain stores),
 * k_split = 1 into 148 equal (tile, k block) ranges whose partial sums meet with fp32 atomics
 * (the mode of every other phase).  Synchronous. */
int b200llm_op_gemv_tm(const void *w_tm, const float *x, float *out, int32_t N, int32_t K, int32_t k_split,
                       void *stream);

/* Paged GQA decode attention with split-T partials merged by the last CTA to finish.
 * `scratch` holds the split partials and arrival counters; it must be zeroed once
 * before first use (the kernel leaves the counters at zero). */
typedef struct b200llm_attn_args {
    const float *q;          /* fp32 [bs, n_q * head_dim] (RoPE applied) */
    float *out;              /* fp32 [bs, n_q * head_dim] */
    const void *k_pages;     /* bf16 [pages, n_kv, page_tokens, head_dim] */
    const void *v_pages;
    const int32_t *row_lane; /* [bs] */
    const int32_t *row_pos;  /* [bs]: attends to tokens 0..pos inclusive */
    const int32_t *page_table;
    int32_t bs, n_q, n_kv, head_dim, page_tokens, max_pages_per_lane, n_splits;
    int32_t kv_page_heads;   /* (kv head, 16-token) tiles from one page to the next; 0 = n_kv (see b200llm_gemv_args) */
    void *scratch;
    size_t scratch_bytes;
} b200llm_attn_args;
int b200llm_op_attn_decode(const b200llm_attn_args *a, void *stream);
int b200llm_op_attn_scratch_bytes(int32_t bs, int32_t n_q, int32_t n_kv, int32_t head_dim,
                                  int32_t n_splits, size_t *bytes);

/* Causal attention of a prefill chunk on the tensor cores (kernel K10; csrc/attn_prefill.cu): `rows` consecutive query
 * rows of ONE lane, row r at position pos0 + r attending keys 0 .. pos0 + r, whose K/V rows are in the paged cache
 * already (this chunk's included).  Replaces what llama.cpp's prompt processing does below the reference's
 * ollama.generate calls (Flask/app.py:102-107,160-166) -- the TTFT path.  q is fp32 [rows, n_q * head_dim] with RoPE
 * applied; out fp32 of the same shape.  k_pages / v_pages: bf16 tiles [page][kv_page_heads (>= n_kv)][16][head_dim] as
 * for b200llm_op_attn_decode; n_pages bounds the page ids; page_table_row is the lane's row (n_ptab entries). */
typedef struct b200llm_attn_prefill_args {
    const float *q;
    float *out;
    const void *k_pages;
    const void *v_pages;
    const int32_t *page_table_row;
    int32_t n_ptab, rows, pos0, n_q, n_kv, head_dim, n_pages, kv_page_heads;   /* kv_page_heads 0 = n_kv */
} b200llm_attn_prefill_args;
int b200llm_op_attn_prefill(const b200llm_attn_prefill_args *a, void *stream);

/* The same attention launch for callers that cannot allocate device memory.  (This entry
 * point has no counterpart in the reference; it exists for host programs of two kinds: those
 * that handed the whole card to their own memory pool at start-up, where a later cudaMalloc
 * simply fails, and those that call from a latency-critical thread, where the allocator lock
 * and the device-wide synchronisation inside cudaMalloc/cudaFree are not acceptable.)
 *
 * `scratch` and `scratch_bytes` of the argument block are ignored: split partials and arrival
 * tickets go to a 4 MiB buffer built into the library (every GPU that loads libb200llm.so
 * holds it until the process exits).  Limits: bs * n_kv <= 1024 and
 * bs * n_q * n_splits * (head_dim + 2) * 4 + 4096 <= 4 MiB (for example 8 rows x 32 heads x
 * 16 splits x 128); anything larger is refused with B200LLM_ENOMEM and a message that names
 * the sizes -- use b200llm_op_attn_decode with your own scratch then.
 *
 * No cudaMalloc, cudaFree or other allocating call is made.  The first call on a device looks
 * up the buffer's device address (cudaGetSymbolAddress, with that device current), and if it is
 * also the first use of this library on that device, the CUDA runtime loads the library's module
 * there at that moment, which does reserve device memory for the code and for the 4 MiB buffer:
 * make that first call early, outside any stream capture, while memory can still be had.  It
 * returns B200LLM_ENODEV, before touching `stream`, when the current device is not an sm_100
 * part.  Every call shares the one buffer and clears its
 * ticket area on `stream` before launching, so calls that may overlap on the GPU (different
 * streams, or different threads) race with each other: serialise them. */
int b200llm_op_attn_decode_static(const b200llm_attn_args *a, void *stream);

/* tcgen05 tensor-core GEMM of the batch (> 4 rows) and prefill path:
 *   out[t, n] (+)= sum_k x[t, k] * w[n, k],  w bf16 [N, K], x bf16 [T, K], out fp32 [T, out_stride].
 * 1 <= T <= 256, N % 128 == 0, K % 64 == 0.  accumulate = 1 adds onto `out` (small matrices are
 * split along K and meet in `out` with fp32 atomics); accumulate = 0 stores (whole K per CTA). */
int b200llm_op_gemm_tc(const void *w, const void *x, float *out, int64_t out_stride, int32_t T, int32_t N, int32_t K,
                       int32_t accumulate, void *stream);

/* Sampler over fp32 logits [bs, vocab] (+ the LOGITS partials for greedy rows). */
typedef struct b200llm_sample_args {
    const float *logits;     /* fp32 [bs, vocab]; may be modified (repeat penalty) */
    int32_t bs, vocab;
    const float *part_max;   /* [bs, n_part] or NULL */
    const int32_t *part_idx;
    int32_t n_part;
    const b200llm_sampling *params; /* DEVICE pointer, [bs] */
    const int32_t *history;  /* DEVICE int32 [bs, 64] recent tokens (-1 = empty) or NULL */
    uint64_t step;           /* RNG counter */
    int32_t *out_token;      /* DEVICE int32 [bs] */
} b200llm_sample_args;
int b200llm_op_sample(const b200llm_sample_args *a, void *stream);

/* Seeded random-init weights, generated on the device (no checkpoints exist offline):
 * element i of stream `seed` (see DESIGN.md "Weights"); bf16 out. */
int b200llm_op_fill_weights(void *dev_bf16, uint64_t n, uint64_t seed, float sigma, float mean,
                            void *stream);

/* A GGUF tensor as it lies in the file -> bf16, on the device (the checkpoints the reference's models ship as:
 * `ollama.generate(model="duckdb-nsql" | "llama3.2")`, Flask/app.py:102-107,160-166, pulls Q4_0 / Q4_K_M blobs).
 * `blocks`: DEVICE copy of the raw tensor bytes; ggml_type: 0 F32, 1 F16, 2 Q4_0, 8 Q8_0, 12 Q4_K, 14 Q6_K, 30 BF16;
 * n_elements: whole blocks (32 for Q4_0 / Q8_0, 256 for Q4_K / Q6_K, even otherwise); out_bf16: DEVICE, n_elements
 * bf16.  ggml's reference arithmetic in ggml's order, so the result is bit-equal to rounding the fp32 values
 * gguf-py's de-quantiser produces. */
int b200llm_op_dequant_bf16(const void *blocks, int32_t ggml_type, int64_t n_elements, void *out_bf16, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* B200LLM_H_ */
