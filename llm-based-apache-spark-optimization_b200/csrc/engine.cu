// engine.cu -- host side of the C ABI (include/b200llm.h): the decode engine that replaces
// what sits below `ollama.generate` in the reference (Ollama server + llama.cpp runner;
// call sites Flask/app.py:102-107,160-166, FastAPI/app.py:85-90,105-111).
//
// The engine owns a stream, the CUDA graphs of the decode step (one per batch size) and,
// for tensor-parallel groups, an NCCL communicator.  All device memory is borrowed from
// the host (PyTorch).  One forward pass over `rows` token rows is
//   embed -> L x [ rmsnorm+qkv GEMV (+RoPE, +paged-KV append) -> split-T paged attention
//                  -> o_proj GEMV (+residual | all-reduce) -> rmsnorm+gate/up GEMV (+SwiGLU)
//                  -> down GEMV (+residual | all-reduce) ] -> rmsnorm+lm_head GEMV (+argmax
//   partials) -> sampler (feeds the token and position back on the device).
#include <dlfcn.h>
#include <stdarg.h>
#include <math.h>
#include <string.h>

#include <map>
#include <tuple>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "kernels.h"
#include "mega.h"
#include "tcd.h"

namespace b200 {

static thread_local char g_err[1024] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// ---------------------------------------------------------------- NCCL (dlopen'ed, optional)
typedef struct ncclComm *ncclComm_t;
struct Id128 {
    char b[128];
};
struct NcclApi {
    void *lib = nullptr;
    int (*GetUniqueId)(void *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, /* ncclUniqueId by value: 128 bytes */ Id128, int) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};
static NcclApi g_nccl;
static std::mutex g_nccl_mu;
constexpr int kNcclFloat32 = 7, kNcclInt8 = 0, kNcclSum = 0;

static int nccl_load() {
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (g_nccl.lib) return B200LLM_OK;
    void *lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!lib) {
        const char *env = getenv("B200LLM_NCCL_LIB");
        lib = dlopen(env ? env : "libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    }
    if (!lib) {
        set_error("NCCL: cannot load libnccl.so.2 (set B200LLM_NCCL_LIB): %s", dlerror());
        return B200LLM_ENCCL;
    }
    g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))dlsym(lib, "ncclGetUniqueId");
    g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))dlsym(lib, "ncclCommInitRank");
    g_nccl.AllReduce = (decltype(g_nccl.AllReduce))dlsym(lib, "ncclAllReduce");
    g_nccl.AllGather = (decltype(g_nccl.AllGather))dlsym(lib, "ncclAllGather");
    g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))dlsym(lib, "ncclCommDestroy");
    g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))dlsym(lib, "ncclGetErrorString");
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.AllGather) {
        set_error("NCCL: symbols missing in libnccl.so.2");
        return B200LLM_ENCCL;
    }
    g_nccl.lib = lib;
    return B200LLM_OK;
}
#define B200_NCCL_OK(expr)                                                                         \
    do {                                                                                           \
        int _r = (expr);                                                                           \
        if (_r != 0) {                                                                             \
            set_error("%s failed: %s", #expr, g_nccl.GetErrorString ? g_nccl.GetErrorString(_r) : "?"); \
            return B200LLM_ENCCL;                                                                  \
        }                                                                                          \
    } while (0)

}  // namespace b200

using namespace b200;

constexpr int kMaxDecodeSteps = 4096;

struct LayerW {
    const void *attn_norm = nullptr, *wqkv = nullptr, *wo = nullptr, *mlp_norm = nullptr, *wgu = nullptr,
               *wdown = nullptr;
    const void *wqkv_st = nullptr, *wo_st = nullptr, *wgu_st = nullptr, *wdown_st = nullptr;  // stage-major copies
    const void *wqkv_tm = nullptr, *wo_tm = nullptr, *wgu_tm = nullptr, *wdown_tm = nullptr;  // tile-major copies
};

struct b200llm_engine {
    b200llm_config cfg;
    int nq_l, nkv_l, ffn_l, vocab_l, max_ctx;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    bool pdl = false;
    // weights
    const void *embed = nullptr, *lm_head = nullptr, *final_norm = nullptr, *lm_head_st = nullptr, *lm_head_tm = nullptr;
    const float *inv_freq = nullptr;
    std::vector<LayerW> layers;
    bool weights_bound = false;
    // kv pool: [layer][K|V][page][kv head][token][dim] bf16
    uint8_t *kv = nullptr;
    int n_pages = 0;
    size_t kv_half_stride = 0;  // bytes from one layer's K region to its V region (and on to the next layer's K)
    int kv_page_heads = 0;      // (kv head, 16-token) tiles from one page to the next: nkv_l, or a shared pool's slot size
    bool kv_shared = false;     // pages are page-major slots of a pool shared with other engines
    bool lean = false;          // only the tile-major matrices are bound: every pass runs the persistent kernel or the tcgen05 GEMM path
    // workspace
    uint8_t *ws = nullptr;
    float *h, *q, *att, *act, *delta, *logits, *part_max;
    int32_t *part_idx, *row_lane, *row_pos, *row_token, *pf_lane, *pf_pos, *pf_token, *lane_token, *lane_pos,
        *lane_hist, *page_table, *gen_out, *step_ctr;
    float *gath_max;
    int32_t *gath_idx;
    // persistent batch-1 decode kernel
    MegaJob *mega_jobs = nullptr;
    unsigned *bar_ctr = nullptr;
    unsigned long long *trace = nullptr;
    unsigned long long *samp_cand = nullptr;
    float2 *tcd_gu2 = nullptr, *tcd_att2 = nullptr, *tcd_qkv2 = nullptr;
    size_t tcd_att2_bytes = 0, tcd_att2_lane_bytes = 0;
    // batch / prefill path
    uint16_t *xb = nullptr, *xb2 = nullptr;
    int64_t xb_lo = 0, xb2_lo = 0;   // element offsets of the lo parts inside xb / xb2 (0: single-bf16 activations)
    bool tc_pdl = true;
    float *qkv_out = nullptr, *gu_out = nullptr;
    struct MapBox { alignas(128) unsigned char b[128]; };
    std::map<std::tuple<const void *, int64_t, int64_t, int>, MapBox *> maps;
    bool mega_ready = false;
    bool tcd_ready = false;
    int tcd_dataflow = 0;   // TcdParams::dataflow bits the phase descriptors were built for
    // one-shot all-reduce over peer memory (tp_ar.cu): own region + the peers' regions opened through CUDA IPC
    uint8_t *ar_region = nullptr;
    void *ar_peer_base[kTpMaxRanks] = {};
    float *ar_inbox[kTpMaxRanks] = {};
    unsigned *ar_flags[kTpMaxRanks] = {};
    unsigned *ar_seq = nullptr;
    unsigned long long *tok_inbox[kTpMaxRanks] = {};   // greedy-token exchange of the tensor-parallel persistent kernel
    unsigned *tok_flags[kTpMaxRanks] = {};
    int ar_slot_floats = 0, ar_max_slices = 0;
    bool ar_ready = false;
    std::vector<b200llm_sampling> samp_host;
    std::vector<int> lane_pos_host, lane_pages_host;   // host mirror: next position and page-table length of every lane
    b200llm_sampling *lane_samp;
    void *attn_scratch;
    size_t attn_scratch_bytes;
    int lm_grid = 0;
    int attn_splits = 1;
    // graphs
    std::map<int, cudaGraphExec_t> graphs;
    std::map<int, int64_t> graph_launches;  // kernels per captured step
    // nccl
    ncclComm_t comm = nullptr;
    // timing
    cudaEvent_t ev[4];
    b200llm_timing timing{};
    int64_t launches = 0;  // counted while building a pass
    int last_rows = 0;
    int last_logit_rows = 0;
};

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static int check_cfg(const b200llm_config *c) {
    B200_REQUIRE(c != nullptr, "config is NULL");
    B200_REQUIRE(c->abi_version == B200LLM_ABI_VERSION, "ABI version mismatch: caller %d, library %d", c->abi_version,
                 B200LLM_ABI_VERSION);
    B200_REQUIRE(c->tp_size >= 1 && c->tp_rank >= 0 && c->tp_rank < c->tp_size, "bad tp_rank/tp_size");
    B200_REQUIRE(c->head_dim == 64 || c->head_dim == 128, "head_dim must be 64 or 128");
    B200_REQUIRE(c->page_tokens == 16, "page_tokens must be 16");
    B200_REQUIRE(c->n_q % c->tp_size == 0 && c->n_kv % c->tp_size == 0 && c->ffn % c->tp_size == 0 &&
                     c->vocab % c->tp_size == 0,
                 "heads/ffn/vocab must divide by tp_size");
    B200_REQUIRE(c->n_q % c->n_kv == 0, "n_q must be a multiple of n_kv");
    B200_REQUIRE(c->d % 8 == 0 && (c->ffn / c->tp_size) % 8 == 0, "d and ffn/tp must be multiples of 8");
    B200_REQUIRE(c->max_rows >= 1 && c->max_lanes >= 1 && c->max_pages_per_lane >= 1, "bad engine limits");
    return B200LLM_OK;
}

struct WsLayout {
    size_t h, q, att, act, delta, logits, part_max, part_idx, row_lane, row_pos, row_token, pf_lane, pf_pos,
        pf_token, lane_token, lane_pos, lane_hist, page_table, gen_out, step_ctr, lane_samp, gath_max, gath_idx,
        attn_scratch, mega_jobs, bar_ctr, trace, samp_cand, tcd_gu2, tcd_att2, tcd_qkv2, xb, xb2, qkv_out, gu_out, total;
    size_t attn_scratch_bytes, tcd_att2_bytes, tcd_att2_lane_bytes;
    size_t xb_lo_off, xb2_lo_off;
    int lm_grid, attn_splits;
};

static int auto_splits(const b200llm_config *c) {
    if (c->attn_splits > 0) return c->attn_splits > 64 ? 64 : c->attn_splits;
    int nkv_l = c->n_kv / c->tp_size;
    int s = (2 * kNumSMs + nkv_l - 1) / nkv_l;
    return s < 1 ? 1 : (s > 32 ? 32 : s);
}

static void ws_layout(const b200llm_config *c, WsLayout *w) {
    const int tp = c->tp_size;
    const size_t R = c->max_rows, Lm = c->max_lanes;
    const size_t nq_l = c->n_q / tp, nkv_l = c->n_kv / tp, ffn_l = c->ffn / tp, vocab_l = c->vocab / tp;
    const size_t max_ctx = (size_t)c->max_pages_per_lane * c->page_tokens;
    w->lm_grid = gemv_grid((int)vocab_l, c->d, 1, B200LLM_EPI_LOGITS, 0);
    // batched lm_head (bs up to 4) may pick a larger grid; use the maximum
    for (int bs = 2; bs <= 4; ++bs) {
        int g = gemv_grid((int)vocab_l, c->d, bs, B200LLM_EPI_LOGITS, 0);
        if (g > w->lm_grid) w->lm_grid = g;
    }
    w->attn_splits = auto_splits(c);
    w->attn_scratch_bytes = attn_scratch_bytes((int)R, (int)nq_l, (int)nkv_l, c->head_dim, w->attn_splits);
    size_t o = 0;
    auto take = [&](size_t bytes) {
        size_t at = o;
        o = align_up(o + bytes, 256);
        return at;
    };
    w->h = take(R * c->d * 4);
    w->q = take(R * nq_l * c->head_dim * 4);
    w->att = take(R * nq_l * c->head_dim * 4);
    w->act = take(R * ffn_l * 4);
    w->delta = take(R * c->d * 4);
    w->logits = take(R * vocab_l * 4);
    const size_t part_n = R * w->lm_grid > (size_t)kTcdMaxLanes * kNumSMs ? R * w->lm_grid : (size_t)kTcdMaxLanes * kNumSMs;
    w->part_max = take(part_n * 4);   // (the persistent kernel keeps [lanes][148] partials here)
    w->part_idx = take(part_n * 4);
    w->row_lane = take(R * 4);
    w->row_pos = take(R * 4);
    w->row_token = take(R * 4);
    w->pf_lane = take(max_ctx * 4);
    w->pf_pos = take(max_ctx * 4);
    w->pf_token = take(max_ctx * 4);
    w->lane_token = take(Lm * 4);
    w->lane_pos = take(Lm * 4);
    w->lane_hist = take(Lm * 64 * 4);
    w->page_table = take(Lm * c->max_pages_per_lane * 4);
    w->gen_out = take((size_t)kMaxDecodeSteps * R * 4);
    w->step_ctr = take(256);
    w->lane_samp = take(Lm * sizeof(b200llm_sampling));
    w->gath_max = take(R * tp * w->lm_grid * 4);
    w->gath_idx = take(R * tp * w->lm_grid * 4);
    w->attn_scratch = take(w->attn_scratch_bytes);
    // (two sets of phase descriptors: the persistent kernel's second set is for 3-4 lanes, see tcd_prepare)
    w->mega_jobs = take(2 * (size_t)(4 * c->n_layers + 1) * (sizeof(MegaJob) > sizeof(TcdJob) ? sizeof(MegaJob) : sizeof(TcdJob)));
    w->bar_ctr = take(256);
    w->trace = take(4096 * 8);
    // the persistent kernel decodes up to kTcdMaxLanes lanes per launch: one block of each of these per lane
    const size_t TL = R < (size_t)kTcdMaxLanes ? R : (size_t)kTcdMaxLanes;
    w->samp_cand = take(TL * kNumSMs * kTcdSampK * 8);        // per-CTA candidate keys of the in-kernel sampler
    w->tcd_gu2 = take(TL * 2 * ffn_l * 8);                    // (sum, count) pairs of the persistent kernel's gate/up phase
    w->tcd_qkv2 = take(TL * (nq_l + 2 * nkv_l) * c->head_dim * 8);   // ... and of its qkv phase
    {   // (value, tag) pairs of its attention phase: outputs, split-T partials, their (max, sum, tag, -)
        const size_t ns = (size_t)(kNumSMs / (nkv_l > 0 ? nkv_l : 1) < 1 ? 1 : kNumSMs / (nkv_l > 0 ? nkv_l : 1));
        w->tcd_att2_lane_bytes = nq_l * c->head_dim * 8 + nq_l * ns * c->head_dim * 8 + nq_l * ns * 16;
        w->tcd_att2_bytes = TL * w->tcd_att2_lane_bytes;
        w->tcd_att2 = take(w->tcd_att2_bytes);
    }
    {   // batch / prefill path (tcgen05 GEMM): bf16 activations and fp32 GEMM outputs
        size_t widest = c->d > (int)(nq_l * c->head_dim) ? c->d : nq_l * c->head_dim;
        if (ffn_l > widest) widest = ffn_l;
        // bf16 GEMM operands: hi parts, then (at a fixed element offset) the lo parts x - hi
        w->xb_lo_off = R * widest;
        w->xb2_lo_off = R * ffn_l;
        w->xb = take(2 * R * widest * 2);
        w->xb2 = take(2 * R * ffn_l * 2);   // SwiGLU output of the fused gate/up GEMM epilogue (xb is that GEMM's input)
        w->qkv_out = take(R * (nq_l + 2 * nkv_l) * c->head_dim * 4);
        w->gu_out = take(R * 2 * ffn_l * 4);
    }
    w->total = o;
}

extern "C" {

const char *b200llm_last_error(void) { return g_err; }
int b200llm_abi_version(void) { return B200LLM_ABI_VERSION; }

int b200llm_workspace_bytes(const b200llm_config *cfg, size_t *bytes) {
    int rc = check_cfg(cfg);
    if (rc) return rc;
    B200_REQUIRE(bytes != nullptr, "bytes is NULL");
    WsLayout w;
    ws_layout(cfg, &w);
    *bytes = w.total;
    return B200LLM_OK;
}

int b200llm_kv_page_bytes(const b200llm_config *cfg, size_t *bytes) {
    int rc = check_cfg(cfg);
    if (rc) return rc;
    B200_REQUIRE(bytes != nullptr, "bytes is NULL");
    *bytes = (size_t)cfg->n_layers * 2 * (cfg->n_kv / cfg->tp_size) * cfg->page_tokens * cfg->head_dim * 2;
    return B200LLM_OK;
}

static int require_device_any() {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        set_error("no CUDA device: this library has no CPU fallback (%s)", cudaGetErrorString(e));
        return B200LLM_ENODEV;
    }
    int cur = 0, major = 0;
    B200_CUDA_OK(cudaGetDevice(&cur));
    B200_CUDA_OK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, cur));
    if (major != 10) {
        set_error("current device is sm_%dx; this library is built for sm_100a (B200) only", major);
        return B200LLM_ENODEV;
    }
    return B200LLM_OK;
}

static int require_device(int device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        set_error("no CUDA device: this library has no CPU fallback (%s)", cudaGetErrorString(e));
        return B200LLM_ENODEV;
    }
    int cur = -1;
    B200_CUDA_OK(cudaGetDevice(&cur));
    if (cur != device) {
        set_error("device %d requested but device %d is current on this thread", device, cur);
        return B200LLM_ENODEV;
    }
    cudaDeviceProp p;
    B200_CUDA_OK(cudaGetDeviceProperties(&p, device));
    if (p.major != 10) {
        set_error("device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, p.major, p.minor);
        return B200LLM_ENODEV;
    }
    return B200LLM_OK;
}

int b200llm_create(const b200llm_config *cfg, b200llm_engine **out) {
    int rc = check_cfg(cfg);
    if (rc) return rc;
    B200_REQUIRE(out != nullptr, "out is NULL");
    rc = require_device(cfg->device);
    if (rc) return rc;
    b200llm_engine *e = new b200llm_engine();
    e->cfg = *cfg;
    e->nq_l = cfg->n_q / cfg->tp_size;
    e->nkv_l = cfg->n_kv / cfg->tp_size;
    e->ffn_l = cfg->ffn / cfg->tp_size;
    e->vocab_l = cfg->vocab / cfg->tp_size;
    e->max_ctx = cfg->max_pages_per_lane * cfg->page_tokens;
    e->layers.resize(cfg->n_layers);
    e->pdl = cfg->use_pdl != 0;
    e->tc_pdl = getenv("B200LLM_TC_NO_PDL") == nullptr;
    if (cfg->stream) {
        e->stream = (cudaStream_t)cfg->stream;
    } else {
        if (cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) != cudaSuccess) {
            set_error("cudaStreamCreate failed");
            delete e;
            return B200LLM_ECUDA;
        }
        e->own_stream = true;
    }
    for (int i = 0; i < 4; ++i) cudaEventCreate(&e->ev[i]);
    if (cfg->tp_size > 1) {
        rc = nccl_load();
        if (rc) {
            delete e;
            return rc;
        }
        if (!cfg->nccl_unique_id) {
            set_error("tp_size > 1 needs nccl_unique_id");
            delete e;
            return B200LLM_EINVAL;
        }
        Id128 id;
        memcpy(id.b, cfg->nccl_unique_id, 128);
        int r = g_nccl.CommInitRank(&e->comm, cfg->tp_size, id, cfg->tp_rank);
        if (r != 0) {
            set_error("ncclCommInitRank failed: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
            delete e;
            return B200LLM_ENCCL;
        }
    }
    *out = e;
    return B200LLM_OK;
}

// ---------------------------------------------------------------- peer-memory all-reduce set-up
static void ar_layout(const b200llm_engine *e, size_t *flag_bytes, size_t *total) {
    const int tp = e->cfg.tp_size;
    const size_t slot = (size_t)e->cfg.max_rows * e->cfg.d;
    const size_t slices = (slot + 1023) / 1024;
    *flag_bytes = align_up(2 * (size_t)tp * slices * 4 + 256, 256);   // flags, then the local sequence words
    *total = *flag_bytes + 2 * (size_t)tp * slot * 4;
}

static void ar_pointers(b200llm_engine *e, int r, uint8_t *base) {
    size_t flag_bytes, total;
    ar_layout(e, &flag_bytes, &total);
    e->ar_flags[r] = reinterpret_cast<unsigned *>(base);
    e->ar_inbox[r] = reinterpret_cast<float *>(base + flag_bytes);
    // the last 256 bytes of the flag area: sequence words (16 B), token-exchange flags (+64), token inbox (+128)
    e->tok_flags[r] = reinterpret_cast<unsigned *>(base + flag_bytes - 256 + 64);
    e->tok_inbox[r] = reinterpret_cast<unsigned long long *>(base + flag_bytes - 256 + 128);
}

int b200llm_tp_export(b200llm_engine *e, void *handle64) {
    B200_REQUIRE(e && handle64, "tp_export: NULL argument");
    B200_REQUIRE(e->cfg.tp_size >= 2 && e->cfg.tp_size <= kTpMaxRanks, "tp_export: tp_size %d not in [2, %d]", e->cfg.tp_size,
                 kTpMaxRanks);
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
    int rc = require_device(e->cfg.device);
    if (rc) return rc;
    size_t flag_bytes, total;
    ar_layout(e, &flag_bytes, &total);
    if (!e->ar_region) {
        B200_CUDA_OK(cudaMalloc(&e->ar_region, total));
        B200_CUDA_OK(cudaMemset(e->ar_region, 0, total));
        const size_t slot = (size_t)e->cfg.max_rows * e->cfg.d;
        e->ar_slot_floats = (int)slot, e->ar_max_slices = (int)((slot + 1023) / 1024);
        ar_pointers(e, e->cfg.tp_rank, e->ar_region);
        e->ar_seq = reinterpret_cast<unsigned *>(e->ar_region + flag_bytes - 256);
    }
    cudaIpcMemHandle_t h;
    B200_CUDA_OK(cudaIpcGetMemHandle(&h, e->ar_region));
    memcpy(handle64, &h, 64);
    return B200LLM_OK;
}

int b200llm_tp_import(b200llm_engine *e, const void *handles, int32_t n) {
    B200_REQUIRE(e && handles && e->ar_region, "tp_import: call tp_export first");
    B200_REQUIRE(n == e->cfg.tp_size, "tp_import: %d handles for tp_size %d", n, e->cfg.tp_size);
    int rc = require_device(e->cfg.device);
    if (rc) return rc;
    for (int r = 0; r < n; ++r) {
        if (r == e->cfg.tp_rank || e->ar_peer_base[r]) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, reinterpret_cast<const uint8_t *>(handles) + 64 * (size_t)r, 64);
        void *base = nullptr;
        B200_CUDA_OK(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
        e->ar_peer_base[r] = base;
        ar_pointers(e, r, reinterpret_cast<uint8_t *>(base));
    }
    e->ar_ready = true;
    for (auto &kv : e->graphs) cudaGraphExecDestroy(kv.second);   // steps captured with the NCCL all-reduce are stale
    e->graphs.clear();
    return B200LLM_OK;
}

int b200llm_nccl_unique_id(void *out128) {
    B200_REQUIRE(out128 != nullptr, "out is NULL");
    int rc = nccl_load();
    if (rc) return rc;
    B200_NCCL_OK(g_nccl.GetUniqueId(out128));
    return B200LLM_OK;
}

int b200llm_destroy(b200llm_engine *e) {
    if (!e) return B200LLM_OK;
    cudaStreamSynchronize(e->stream);
    for (auto &kv : e->graphs) cudaGraphExecDestroy(kv.second);
    for (auto &m : e->maps) delete m.second;
    if (e->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(e->comm);
    for (int r = 0; r < kTpMaxRanks; ++r)
        if (e->ar_peer_base[r]) cudaIpcCloseMemHandle(e->ar_peer_base[r]);
    if (e->ar_region) cudaFree(e->ar_region);
    for (int i = 0; i < 4; ++i) cudaEventDestroy(e->ev[i]);
    if (e->own_stream) cudaStreamDestroy(e->stream);
    delete e;
    return B200LLM_OK;
}

#define B200_TRY(expr)        \
    do {                      \
        int _rc = (expr);     \
        if (_rc) return _rc;  \
    } while (0)

// A bound buffer must be device memory of the engine's GPU (a tensor made on another GPU of the process would fault in the
// first kernel that touches it, far from the cause)
static int check_on_device(const b200llm_engine *e, const void *ptr, const char *what, int index) {
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, ptr) != cudaSuccess) {
        cudaGetLastError();
        set_error("%s %d: not a CUDA pointer", what, index);
        return B200LLM_EINVAL;
    }
    if (at.type != cudaMemoryTypeDevice && at.type != cudaMemoryTypeManaged) {
        set_error("%s %d: not device memory", what, index);
        return B200LLM_EINVAL;
    }
    if (at.type == cudaMemoryTypeDevice && at.device != e->cfg.device) {
        set_error("%s %d lives on GPU %d, the engine on GPU %d", what, index, at.device, e->cfg.device);
        return B200LLM_EINVAL;
    }
    return B200LLM_OK;
}

int b200llm_bind_weights(b200llm_engine *e, const b200llm_tensor_desc *descs, int32_t n) {
    B200_REQUIRE(e && descs && n > 0, "bind_weights: bad arguments");
    const b200llm_config &c = e->cfg;
    for (int i = 0; i < n; ++i) {
        const b200llm_tensor_desc &t = descs[i];
        B200_REQUIRE(t.dev != nullptr && ((uintptr_t)t.dev % 16) == 0, "tensor %d: NULL or unaligned device pointer", i);
        B200_TRY(check_on_device(e, t.dev, "tensor", i));
        auto shape = [&](int64_t r, int64_t cc) {
            if (t.rows != r || t.cols != cc) {
                set_error("tensor kind %d layer %d: shape [%lld, %lld], expected [%lld, %lld]", t.kind, t.layer,
                          (long long)t.rows, (long long)t.cols, (long long)r, (long long)cc);
                return false;
            }
            return true;
        };
        const int64_t qkv_rows = (int64_t)(e->nq_l + 2 * e->nkv_l) * c.head_dim;
        if (t.kind >= B200LLM_W_ATTN_NORM && t.kind != B200LLM_W_LM_HEAD_ST && t.kind != B200LLM_W_LM_HEAD_TM) B200_REQUIRE(t.layer >= 0 && t.layer < c.n_layers, "tensor %d: bad layer", i);
        switch (t.kind) {
            case B200LLM_W_EMBED: if (!shape(c.vocab, c.d)) return B200LLM_EINVAL; e->embed = t.dev; break;
            case B200LLM_W_LM_HEAD: if (!shape(e->vocab_l, c.d)) return B200LLM_EINVAL; e->lm_head = t.dev; break;
            case B200LLM_W_FINAL_NORM: if (!shape(1, c.d)) return B200LLM_EINVAL; e->final_norm = t.dev; break;
            case B200LLM_W_INV_FREQ: if (!shape(1, c.head_dim / 2)) return B200LLM_EINVAL; e->inv_freq = (const float *)t.dev; break;
            case B200LLM_W_ATTN_NORM: if (!shape(1, c.d)) return B200LLM_EINVAL; e->layers[t.layer].attn_norm = t.dev; break;
            case B200LLM_W_QKV: if (!shape(qkv_rows, c.d)) return B200LLM_EINVAL; e->layers[t.layer].wqkv = t.dev; break;
            case B200LLM_W_O: if (!shape(c.d, (int64_t)e->nq_l * c.head_dim)) return B200LLM_EINVAL; e->layers[t.layer].wo = t.dev; break;
            case B200LLM_W_MLP_NORM: if (!shape(1, c.d)) return B200LLM_EINVAL; e->layers[t.layer].mlp_norm = t.dev; break;
            case B200LLM_W_GATE_UP: if (!shape(2 * (int64_t)e->ffn_l, c.d)) return B200LLM_EINVAL; e->layers[t.layer].wgu = t.dev; break;
            case B200LLM_W_DOWN: if (!shape(c.d, e->ffn_l)) return B200LLM_EINVAL; e->layers[t.layer].wdown = t.dev; break;
            case B200LLM_W_QKV_ST: if (!shape(qkv_rows, c.d)) return B200LLM_EINVAL; e->layers[t.layer].wqkv_st = t.dev; break;
            case B200LLM_W_O_ST: if (!shape(c.d, (int64_t)e->nq_l * c.head_dim)) return B200LLM_EINVAL; e->layers[t.layer].wo_st = t.dev; break;
            case B200LLM_W_GATE_UP_ST: if (!shape(2 * (int64_t)e->ffn_l, c.d)) return B200LLM_EINVAL; e->layers[t.layer].wgu_st = t.dev; break;
            case B200LLM_W_DOWN_ST: if (!shape(c.d, e->ffn_l)) return B200LLM_EINVAL; e->layers[t.layer].wdown_st = t.dev; break;
            case B200LLM_W_LM_HEAD_ST: if (!shape(e->vocab_l, c.d)) return B200LLM_EINVAL; e->lm_head_st = t.dev; break;
            case B200LLM_W_QKV_TM: if (!shape(qkv_rows, c.d)) return B200LLM_EINVAL; e->layers[t.layer].wqkv_tm = t.dev; break;
            case B200LLM_W_O_TM: if (!shape(c.d, (int64_t)e->nq_l * c.head_dim)) return B200LLM_EINVAL; e->layers[t.layer].wo_tm = t.dev; break;
            case B200LLM_W_GATE_UP_TM: if (!shape(2 * (int64_t)e->ffn_l, c.d)) return B200LLM_EINVAL; e->layers[t.layer].wgu_tm = t.dev; break;
            case B200LLM_W_DOWN_TM: if (!shape(c.d, e->ffn_l)) return B200LLM_EINVAL; e->layers[t.layer].wdown_tm = t.dev; break;
            case B200LLM_W_LM_HEAD_TM:   // rows padded with zeros to whole 128-row tiles
                if (!shape((e->vocab_l + 127) / 128 * 128, c.d)) return B200LLM_EINVAL;
                e->lm_head_tm = t.dev;
                break;
            default: set_error("tensor %d: unknown kind %d", i, t.kind); return B200LLM_EINVAL;
        }
    }
    // Either the row-major matrices (every path available) or, without them, all tile-major ones ("lean": half the weight
    // footprint; the GEMV kernels, which read row-major rows, are then not available -- one GPU, vocabulary a multiple of 128)
    bool norms = e->embed && e->final_norm && e->inv_freq;
    for (auto &l : e->layers) norms = norms && l.attn_norm && l.mlp_norm;
    bool rows_all = e->lm_head != nullptr, tiles_all = e->lm_head_tm != nullptr;
    for (auto &l : e->layers) {
        rows_all = rows_all && l.wqkv && l.wo && l.wgu && l.wdown;
        tiles_all = tiles_all && l.wqkv_tm && l.wo_tm && l.wgu_tm && l.wdown_tm;
    }
    e->lean = norms && !rows_all && tiles_all;
    if (e->lean)
        B200_REQUIRE(c.tp_size == 1 && c.vocab % 128 == 0 && c.use_tc && c.use_megakernel,
                     "bind_weights: tile-major matrices alone need tp_size 1, a vocabulary that is a multiple of 128, use_tc and use_megakernel");
    e->weights_bound = norms && (rows_all || tiles_all);
    return B200LLM_OK;
}

int b200llm_bind_kv_pool(b200llm_engine *e, void *dev, size_t bytes) {
    B200_REQUIRE(e && dev && ((uintptr_t)dev % 256) == 0, "bind_kv_pool: NULL or unaligned pool");
    B200_TRY(check_on_device(e, dev, "KV pool", 0));
    size_t pb;
    b200llm_kv_page_bytes(&e->cfg, &pb);
    e->n_pages = (int)(bytes / pb);
    B200_REQUIRE(e->n_pages >= 1, "bind_kv_pool: pool smaller than one page (%zu < %zu)", bytes, pb);
    e->kv = (uint8_t *)dev;
    e->kv_half_stride = (size_t)e->n_pages * e->nkv_l * e->cfg.page_tokens * e->cfg.head_dim * 2;
    e->kv_page_heads = e->nkv_l, e->kv_shared = false;
    for (auto &kv : e->graphs) cudaGraphExecDestroy(kv.second);   // captured steps carry the old addresses
    e->graphs.clear();
    e->mega_ready = e->tcd_ready = false;
    return B200LLM_OK;
}

int b200llm_bind_kv_pool_shared(b200llm_engine *e, void *dev, size_t bytes, size_t slot_bytes) {
    B200_REQUIRE(e && dev && ((uintptr_t)dev % 256) == 0, "bind_kv_pool_shared: NULL or unaligned pool");
    B200_TRY(check_on_device(e, dev, "KV pool", 0));
    size_t pb;
    b200llm_kv_page_bytes(&e->cfg, &pb);
    const size_t tile = (size_t)e->cfg.page_tokens * e->cfg.head_dim * 2;
    B200_REQUIRE(slot_bytes >= pb && slot_bytes % tile == 0 && slot_bytes % 256 == 0,
                 "bind_kv_pool_shared: slot of %zu bytes must hold a %zu-byte page and be a multiple of %zu", slot_bytes, pb, tile);
    e->n_pages = (int)(bytes / slot_bytes);
    B200_REQUIRE(e->n_pages >= 1, "bind_kv_pool_shared: pool smaller than one slot (%zu < %zu)", bytes, slot_bytes);
    e->kv = (uint8_t *)dev;
    e->kv_half_stride = (size_t)e->nkv_l * tile;        // page-major: layer l's K tiles, then its V tiles, inside every page
    e->kv_page_heads = (int)(slot_bytes / tile), e->kv_shared = true;
    for (auto &kv : e->graphs) cudaGraphExecDestroy(kv.second);
    e->graphs.clear();
    e->mega_ready = e->tcd_ready = false;
    return B200LLM_OK;
}

int b200llm_bind_workspace(b200llm_engine *e, void *dev, size_t bytes) {
    B200_REQUIRE(e && dev && ((uintptr_t)dev % 256) == 0, "bind_workspace: NULL or unaligned workspace");
    B200_TRY(check_on_device(e, dev, "workspace", 0));
    WsLayout w;
    ws_layout(&e->cfg, &w);
    if (bytes < w.total) {
        set_error("bind_workspace: %zu bytes given, %zu needed", bytes, w.total);
        return B200LLM_ENOMEM;
    }
    uint8_t *b = (uint8_t *)dev;
    e->ws = b;
    e->h = (float *)(b + w.h), e->q = (float *)(b + w.q), e->att = (float *)(b + w.att);
    e->act = (float *)(b + w.act), e->delta = (float *)(b + w.delta), e->logits = (float *)(b + w.logits);
    e->part_max = (float *)(b + w.part_max), e->part_idx = (int32_t *)(b + w.part_idx);
    e->row_lane = (int32_t *)(b + w.row_lane), e->row_pos = (int32_t *)(b + w.row_pos);
    e->row_token = (int32_t *)(b + w.row_token);
    e->pf_lane = (int32_t *)(b + w.pf_lane), e->pf_pos = (int32_t *)(b + w.pf_pos);
    e->pf_token = (int32_t *)(b + w.pf_token);
    e->lane_token = (int32_t *)(b + w.lane_token), e->lane_pos = (int32_t *)(b + w.lane_pos);
    e->lane_hist = (int32_t *)(b + w.lane_hist), e->page_table = (int32_t *)(b + w.page_table);
    e->gen_out = (int32_t *)(b + w.gen_out), e->step_ctr = (int32_t *)(b + w.step_ctr);
    e->lane_samp = (b200llm_sampling *)(b + w.lane_samp);
    e->gath_max = (float *)(b + w.gath_max), e->gath_idx = (int32_t *)(b + w.gath_idx);
    e->attn_scratch = b + w.attn_scratch;
    e->attn_scratch_bytes = w.attn_scratch_bytes;
    e->lm_grid = w.lm_grid;
    e->attn_splits = w.attn_splits;
    e->mega_jobs = (MegaJob *)(b + w.mega_jobs);
    e->bar_ctr = (unsigned *)(b + w.bar_ctr);
    e->trace = (unsigned long long *)(b + w.trace);
    e->samp_cand = (unsigned long long *)(b + w.samp_cand);
    e->tcd_gu2 = (float2 *)(b + w.tcd_gu2);
    e->tcd_qkv2 = (float2 *)(b + w.tcd_qkv2);
    e->tcd_att2 = (float2 *)(b + w.tcd_att2), e->tcd_att2_bytes = w.tcd_att2_bytes, e->tcd_att2_lane_bytes = w.tcd_att2_lane_bytes;
    const bool hilo = getenv("B200LLM_TC_BF16ACT") == nullptr;  // default: hi + lo operand pairs, ~fp32-accurate activations
    e->xb_lo = hilo ? (int64_t)w.xb_lo_off : 0, e->xb2_lo = hilo ? (int64_t)w.xb2_lo_off : 0;
    e->xb = (uint16_t *)(b + w.xb), e->xb2 = (uint16_t *)(b + w.xb2), e->qkv_out = (float *)(b + w.qkv_out), e->gu_out = (float *)(b + w.gu_out);
    e->mega_ready = false;
    e->tcd_ready = false;
    e->samp_host.assign(e->cfg.max_lanes, b200llm_sampling{0.f, 0, 1.f, 1.f, 0});
    e->lane_pos_host.assign(e->cfg.max_lanes, 0);
    e->lane_pages_host.assign(e->cfg.max_lanes, 0);
    // zero everything once: attention tickets must start at zero, lanes at token 0 / pos 0,
    // sampling = greedy, page tables -> page 0
    B200_CUDA_OK(cudaMemsetAsync(b, 0, w.total, e->stream));
    B200_CUDA_OK(cudaMemsetAsync(e->lane_hist, 0xFF, (size_t)e->cfg.max_lanes * 64 * 4, e->stream));
    std::vector<b200llm_sampling> sp(e->cfg.max_lanes, b200llm_sampling{0.f, 0, 1.f, 1.f, 0});
    B200_CUDA_OK(cudaMemcpyAsync(e->lane_samp, sp.data(), sp.size() * sizeof(b200llm_sampling),
                                 cudaMemcpyHostToDevice, e->stream));
    B200_CUDA_OK(cudaStreamSynchronize(e->stream));
    return B200LLM_OK;
}

static int require_ready(b200llm_engine *e) {
    if (!e) {
        set_error("engine is NULL");
        return B200LLM_EINVAL;
    }
    if (!e->weights_bound || !e->kv || !e->ws) {
        set_error("engine not ready: bind weights (%d), KV pool (%d) and workspace (%d) first", (int)e->weights_bound,
                  e->kv != nullptr, e->ws != nullptr);
        return B200LLM_ESTATE;
    }
    return B200LLM_OK;
}

int b200llm_set_page_table(b200llm_engine *e, int32_t lane, const int32_t *pages, int32_t n) {
    B200_REQUIRE(e && e->ws && pages, "set_page_table: bad arguments");
    B200_REQUIRE(lane >= 0 && lane < e->cfg.max_lanes && n >= 0 && n <= e->cfg.max_pages_per_lane,
                 "set_page_table: lane %d / n %d out of range", lane, n);
    for (int i = 0; i < n; ++i)
        B200_REQUIRE(pages[i] >= 0 && pages[i] < e->n_pages, "set_page_table: page %d out of range [0, %d)", pages[i],
                     e->n_pages);
    B200_CUDA_OK(cudaMemcpyAsync(e->page_table + (size_t)lane * e->cfg.max_pages_per_lane, pages, (size_t)n * 4,
                                 cudaMemcpyHostToDevice, e->stream));
    e->lane_pages_host[lane] = n;
    return B200LLM_OK;
}

int b200llm_set_sampling(b200llm_engine *e, int32_t lane, const b200llm_sampling *s) {
    B200_REQUIRE(e && e->ws && s && lane >= 0 && lane < e->cfg.max_lanes, "set_sampling: bad arguments");
    B200_REQUIRE(s->temperature >= 0.f && s->repeat_penalty > 0.f, "set_sampling: bad temperature/repeat_penalty");
    // vocab-parallel lm_head: only the ranks' (max, argmax) partials are exchanged, so a temperature or a repeat
    // penalty would sample from this rank's vocabulary shard alone
    B200_REQUIRE(e->cfg.tp_size == 1 || (s->temperature <= 0.f && s->repeat_penalty == 1.f),
                 "set_sampling: tensor parallel decodes greedily only (temperature 0, repeat_penalty 1); got %g / %g",
                 (double)s->temperature, (double)s->repeat_penalty);
    B200_CUDA_OK(cudaMemcpyAsync(e->lane_samp + lane, s, sizeof(*s), cudaMemcpyHostToDevice, e->stream));
    e->samp_host[lane] = *s;
    return B200LLM_OK;
}

int b200llm_set_lane(b200llm_engine *e, int32_t lane, int32_t next_token, int32_t pos) {
    B200_REQUIRE(e && e->ws && lane >= 0 && lane < e->cfg.max_lanes, "set_lane: bad arguments");
    B200_REQUIRE(pos >= 0 && pos < e->max_ctx, "set_lane: pos %d out of range", pos);
    B200_CUDA_OK(cudaMemcpyAsync(e->lane_token + lane, &next_token, 4, cudaMemcpyHostToDevice, e->stream));
    B200_CUDA_OK(cudaMemcpyAsync(e->lane_pos + lane, &pos, 4, cudaMemcpyHostToDevice, e->stream));
    e->lane_pos_host[lane] = pos;
    if (pos == 0)
        B200_CUDA_OK(cudaMemsetAsync(e->lane_hist + (size_t)lane * 64, 0xFF, 64 * 4, e->stream));
    return B200LLM_OK;
}

int b200llm_set_history(b200llm_engine *e, int32_t lane, const int32_t *ids, int32_t n, int32_t pos0) {
    B200_REQUIRE(e && e->ws && lane >= 0 && lane < e->cfg.max_lanes, "set_history: bad arguments");
    B200_REQUIRE(n >= 0 && n <= 64 && (n == 0 || ids) && pos0 >= 0, "set_history: %d ids at pos %d", n, pos0);
    int32_t ring[64];
    for (int i = 0; i < 64; ++i) ring[i] = -1;
    for (int i = 0; i < n; ++i) ring[(pos0 + i) & 63] = ids[i];   // same slot the embedding kernel uses: pos & 63
    B200_CUDA_OK(cudaMemcpyAsync(e->lane_hist + (size_t)lane * 64, ring, sizeof(ring), cudaMemcpyHostToDevice, e->stream));
    return B200LLM_OK;   // the pageable source is staged before the call returns
}

}  // extern "C"

// ---------------------------------------------------------------- one forward pass
struct PassRows {
    int rows;
    int32_t *row_lane, *row_pos, *row_token;
    bool from_lanes;
    int logit_row0;   // first row that needs logits
    int logit_rows;   // 0 = skip lm_head + sampler
    int lane = -1;    // prefill chunk: its rows are consecutive positions pos0 .. of this lane
    int pos0 = 0;
};

// GEMV over `rows` rows in groups of <= 4 (the batch > 4 tensor-core path replaces this loop)
static int gemv_rows(b200llm_engine *e, b200llm_gemv_args a, int rows, const PassRows &pr, int row0) {
    const float *x = a.x;
    float *out = a.out;
    for (int r = 0; r < rows; r += 4) {
        a.bs = rows - r < 4 ? rows - r : 4;
        a.x = x + (size_t)r * a.x_stride;
        a.out = out + (size_t)r * a.out_stride;
        a.row_lane = pr.row_lane + row0 + r;
        a.row_pos = pr.row_pos + row0 + r;
        if (a.epi == B200LLM_EPI_LOGITS) {
            a.part_max = e->part_max + (size_t)r * e->lm_grid;
            a.part_idx = e->part_idx + (size_t)r * e->lm_grid;
        }
        B200_TRY(gemv_launch(a, e->stream, e->pdl));
        e->launches++;
    }
    return B200LLM_OK;
}

static int tp_allreduce_add(b200llm_engine *e, int rows) {
    const size_t n = (size_t)rows * e->cfg.d;
    if (e->ar_ready) {   // peers connected (b200llm_tp_import): one kernel, one NVLink round trip
        B200_TRY(tp_allreduce_add_launch(e->h, e->delta, (int)n, e->ar_inbox, e->ar_flags, e->ar_seq, e->cfg.tp_rank,
                                         e->cfg.tp_size, e->ar_slot_floats, e->ar_max_slices, e->stream));
        e->launches++;
        return B200LLM_OK;
    }
    B200_NCCL_OK(g_nccl.AllReduce(e->delta, e->delta, n, kNcclFloat32, kNcclSum, e->comm, e->stream));
    B200_TRY(add_inplace_launch(e->h, e->delta, n, e->stream, false));
    e->launches++;
    return B200LLM_OK;
}

// Tensor parallel: did a wait for a peer time out in one of the kernels queued so far?  (Call with the stream drained.)
static int tp_check_peers(b200llm_engine *e) {
    if (e->cfg.tp_size <= 1 || !e->ar_seq) return B200LLM_OK;
    unsigned err = 0;
    B200_CUDA_OK(cudaMemcpy(&err, e->ar_seq + 3, 4, cudaMemcpyDeviceToHost));
    if (err) {
        set_error("tensor parallel: rank %d waited 2 s for a peer inside a kernel (a rank left the lock-step); results of "
                  "this engine are invalid from that call on", e->cfg.tp_rank);
        return B200LLM_ENCCL;
    }
    return B200LLM_OK;
}

// final norm + lm_head GEMV + sampler (vocab-parallel under tensor parallel: the ranks' (max, argmax) partials are gathered)
static int lm_head_and_sample(b200llm_engine *e, const PassRows &pr, bool advance) {
    const b200llm_config &c = e->cfg;
    const bool tp = c.tp_size > 1;
    b200llm_gemv_args lm{};
    lm.w = e->lm_head, lm.x = e->h + (size_t)pr.logit_row0 * c.d, lm.x_stride = c.d, lm.norm_gain = e->final_norm;
    lm.rms_eps = c.rms_eps, lm.out = e->logits, lm.out_stride = e->vocab_l, lm.N = e->vocab_l, lm.K = c.d;
    lm.epi = B200LLM_EPI_LOGITS, lm.grid_hint = e->lm_grid, lm.logits_idx_offset = c.tp_rank * e->vocab_l;
    B200_TRY(gemv_rows(e, lm, pr.logit_rows, pr, pr.logit_row0));

    SampleDev sd{};
    sd.a.logits = e->logits, sd.a.bs = pr.logit_rows, sd.a.vocab = e->vocab_l;
    sd.a.part_max = e->part_max, sd.a.part_idx = e->part_idx, sd.a.n_part = e->lm_grid;
    sd.a.params = e->lane_samp, sd.a.out_token = nullptr;
    if (tp) {
        // vocab-parallel greedy: gather every rank's per-row (max, argmax) and reduce those.
        // TODO(tp sampling): temperature > 0 needs the full logits gathered; greedy only here.
        B200_NCCL_OK(g_nccl.AllGather(e->part_max, e->gath_max, (size_t)pr.logit_rows * e->lm_grid * 4, kNcclInt8,
                                      e->comm, e->stream));
        B200_NCCL_OK(g_nccl.AllGather(e->part_idx, e->gath_idx, (size_t)pr.logit_rows * e->lm_grid * 4, kNcclInt8,
                                      e->comm, e->stream));
        // gathered layout is [rank][row][grid]
        sd.a.part_max = e->gath_max, sd.a.part_idx = e->gath_idx;
        sd.n_ranks = c.tp_size, sd.rank_stride = (size_t)pr.logit_rows * e->lm_grid;
    }
    sd.row_lane = pr.row_lane + pr.logit_row0;
    sd.lane_token = e->lane_token, sd.lane_pos = e->lane_pos, sd.lane_hist = e->lane_hist;
    sd.gen_out = pr.from_lanes ? e->gen_out : nullptr, sd.step_ctr = pr.from_lanes ? e->step_ctr : nullptr;
    sd.advance = advance ? 1 : 0;
    B200_TRY(sample_launch(sd, e->stream, e->pdl && !tp));
    e->launches++;
    return B200LLM_OK;
}


static int forward_pass_gemv(b200llm_engine *e, const PassRows &pr, bool advance) {
    const b200llm_config &c = e->cfg;
    const int rows = pr.rows;
    const bool tp = c.tp_size > 1;
    EmbedArgs ea{};
    ea.embed = e->embed, ea.h = e->h, ea.row_token = pr.row_token, ea.row_pos = pr.row_pos, ea.row_lane = pr.row_lane;
    ea.lane_token = e->lane_token, ea.lane_pos = e->lane_pos, ea.rows = rows, ea.d = c.d, ea.vocab = c.vocab;
    ea.from_lanes = pr.from_lanes ? 1 : 0;
    B200_TRY(embed_launch_ex(ea, e->lane_hist, pr.from_lanes ? e->step_ctr : nullptr, e->stream, e->pdl));
    e->launches++;

    const size_t qd = (size_t)e->nq_l * c.head_dim;
    for (int l = 0; l < c.n_layers; ++l) {
        const LayerW &w = e->layers[l];
        uint8_t *kl = e->kv + (size_t)l * 2 * e->kv_half_stride;
        uint8_t *vl = kl + e->kv_half_stride;
        b200llm_gemv_args g{};
        // --- rmsnorm + qkv + rope + kv append ---
        g.w = w.wqkv, g.x = e->h, g.x_stride = c.d, g.norm_gain = w.attn_norm, g.rms_eps = c.rms_eps;
        g.out = e->q, g.out_stride = (int64_t)qd, g.N = (e->nq_l + 2 * e->nkv_l) * c.head_dim, g.K = c.d;
        g.epi = B200LLM_EPI_QKV_ROPE, g.n_q = e->nq_l, g.n_kv = e->nkv_l, g.head_dim = c.head_dim;
        g.page_tokens = c.page_tokens, g.max_pages_per_lane = c.max_pages_per_lane, g.inv_freq = e->inv_freq;
        g.page_table = e->page_table, g.k_pages = kl, g.v_pages = vl, g.kv_page_heads = e->kv_page_heads;
        B200_TRY(gemv_rows(e, g, rows, pr, 0));
        // --- attention ---
        b200llm_attn_args at{};
        at.q = e->q, at.out = e->att, at.k_pages = kl, at.v_pages = vl, at.row_lane = pr.row_lane, at.row_pos = pr.row_pos;
        at.page_table = e->page_table, at.bs = rows, at.n_q = e->nq_l, at.n_kv = e->nkv_l, at.head_dim = c.head_dim;
        at.page_tokens = c.page_tokens, at.max_pages_per_lane = c.max_pages_per_lane, at.kv_page_heads = e->kv_page_heads;
        // enough CTAs to fill the GPU, no more splits than that needs
        int want = (kNumSMs + rows * e->nkv_l - 1) / (rows * e->nkv_l);
        if (want > 6) want = 6;  // the last-CTA merge is serial in the number of splits (profiles/launches_r1a)
        at.n_splits = want < 1 ? 1 : (want > e->attn_splits ? e->attn_splits : want);
        at.scratch = e->attn_scratch, at.scratch_bytes = e->attn_scratch_bytes;
        B200_TRY(attn_decode_launch(at, e->stream, e->pdl));
        e->launches++;
        // --- o_proj ---
        b200llm_gemv_args o{};
        o.w = w.wo, o.x = e->att, o.x_stride = (int64_t)qd, o.N = c.d, o.K = (int)qd;
        o.out = tp ? e->delta : e->h, o.out_stride = c.d, o.epi = tp ? B200LLM_EPI_STORE : B200LLM_EPI_RESIDUAL;
        B200_TRY(gemv_rows(e, o, rows, pr, 0));
        if (tp) B200_TRY(tp_allreduce_add(e, rows));
        // --- rmsnorm + gate/up + swiglu ---
        b200llm_gemv_args u{};
        u.w = w.wgu, u.x = e->h, u.x_stride = c.d, u.norm_gain = w.mlp_norm, u.rms_eps = c.rms_eps;
        u.out = e->act, u.out_stride = e->ffn_l, u.N = 2 * e->ffn_l, u.K = c.d, u.epi = B200LLM_EPI_SWIGLU;
        B200_TRY(gemv_rows(e, u, rows, pr, 0));
        // --- down ---
        b200llm_gemv_args dn{};
        dn.w = w.wdown, dn.x = e->act, dn.x_stride = e->ffn_l, dn.N = c.d, dn.K = e->ffn_l;
        dn.out = tp ? e->delta : e->h, dn.out_stride = c.d, dn.epi = tp ? B200LLM_EPI_STORE : B200LLM_EPI_RESIDUAL;
        B200_TRY(gemv_rows(e, dn, rows, pr, 0));
        if (tp) B200_TRY(tp_allreduce_add(e, rows));
    }
    e->last_rows = rows;
    e->last_logit_rows = pr.logit_rows;
    if (pr.logit_rows <= 0) return B200LLM_OK;
    return lm_head_and_sample(e, pr, advance);
}

// ---------------------------------------------------------------- persistent batch-1 decode
static int mega_prepare(b200llm_engine *e) {
    const b200llm_config &c = e->cfg;
    std::vector<MegaJob> jobs(4 * c.n_layers + 1);
    const int qd = c.n_q * c.head_dim;
    bool staged = e->lm_head_st != nullptr;
    for (auto &w : e->layers) staged = staged && w.wqkv_st && w.wo_st && w.wgu_st && w.wdown_st;
    const int st = staged ? 1 : 0;
    for (int l = 0; l < c.n_layers; ++l) {
        const LayerW &w = e->layers[l];
        uint8_t *kl = e->kv + (size_t)l * 2 * e->kv_half_stride;
        jobs[4 * l + 0] = MegaJob{st ? w.wqkv_st : w.wqkv, w.attn_norm, kl, kl + e->kv_half_stride,
                                  (c.n_q + 2 * c.n_kv) * c.head_dim, c.d, B200LLM_EPI_QKV_ROPE,
                                  (l == 0 ? MEGA_X_EMBED : MEGA_X_H) | (st << 8)};
        jobs[4 * l + 1] = MegaJob{st ? w.wo_st : w.wo, nullptr, nullptr, nullptr, c.d, qd, B200LLM_EPI_RESIDUAL, MEGA_X_ATT | (st << 8)};
        jobs[4 * l + 2] = MegaJob{st ? w.wgu_st : w.wgu, w.mlp_norm, nullptr, nullptr, 2 * c.ffn, c.d, B200LLM_EPI_SWIGLU, MEGA_X_H | (st << 8)};
        jobs[4 * l + 3] = MegaJob{st ? w.wdown_st : w.wdown, nullptr, nullptr, nullptr, c.d, c.ffn, B200LLM_EPI_RESIDUAL, MEGA_X_ACT | (st << 8)};
    }
    jobs[4 * c.n_layers] = MegaJob{st ? e->lm_head_st : e->lm_head, e->final_norm, nullptr, nullptr, c.vocab, c.d,
                                   B200LLM_EPI_LOGITS, MEGA_X_H | (st << 8)};
    B200_CUDA_OK(cudaMemcpyAsync(e->mega_jobs, jobs.data(), jobs.size() * sizeof(MegaJob), cudaMemcpyHostToDevice,
                                 e->stream));
    B200_CUDA_OK(cudaStreamSynchronize(e->stream));  // `jobs` is a pageable temporary
    e->mega_ready = true;
    e->tcd_ready = false;
    return B200LLM_OK;
}

static int mega_decode(b200llm_engine *e, int lane, int n_steps) {
    const b200llm_config &c = e->cfg;
    if (!e->mega_ready) B200_TRY(mega_prepare(e));
    MegaParams p{};
    p.jobs = e->mega_jobs, p.n_layers = c.n_layers;
    p.d = c.d, p.n_q = c.n_q, p.n_kv = c.n_kv, p.head_dim = c.head_dim, p.ffn = c.ffn, p.vocab = c.vocab;
    p.kv_page_heads = e->kv_page_heads;
    p.rms_eps = c.rms_eps;
    p.scale_log2 = (float)(1.4426950408889634 / sqrt((double)c.head_dim));
    p.h = e->h, p.q = e->q, p.att = e->att, p.act = e->act, p.logits = e->logits;
    p.part_max = e->part_max, p.part_idx = e->part_idx;
    p.embed = e->embed, p.inv_freq = e->inv_freq;
    p.lane = lane, p.lane_token = e->lane_token, p.lane_pos = e->lane_pos, p.lane_hist = e->lane_hist;
    p.page_table = e->page_table + (size_t)lane * c.max_pages_per_lane;
    p.n_ptab = c.max_pages_per_lane;
    p.gen_out = e->gen_out, p.n_steps = n_steps;
    int ns = kNumSMs / c.n_kv;
    ns = ns < 1 ? 1 : (ns > e->attn_splits ? e->attn_splits : ns);
    p.n_splits = ns;
    uint8_t *sc = reinterpret_cast<uint8_t *>(e->attn_scratch);
    p.tickets = reinterpret_cast<unsigned *>(sc);
    p.part_o = reinterpret_cast<float *>(sc + kAttnTicketBytes);
    p.part_ml = p.part_o + (size_t)c.n_q * ns * c.head_dim;
    p.bar_ctr = e->bar_ctr;
    p.sample = 1;
    p.ring_slots = 5;
    p.gap_slots = 1 << 30;  // gate off: measured best (profiles/); B200LLM_MEGA_GAP=n closes it after n stages
    if (const char *gp = getenv("B200LLM_MEGA_GAP")) p.gap_slots = atoi(gp);
    if (const char *dbg = getenv("B200LLM_MEGA_DBG")) p.dbg = atoi(dbg);
    if (const char *rs = getenv("B200LLM_MEGA_SLOTS")) {
        int v = atoi(rs);
        if (v >= 1 && v <= 5) p.ring_slots = v;
    }
    if (getenv("B200LLM_MEGA_TRACE")) p.trace = e->trace, p.trace_step = n_steps > 1 ? 1 : 0;
    e->last_rows = 1, e->last_logit_rows = 1;
    return mega_decode_launch(p, e->stream);
}

// ---------------------------------------------------------------- persistent batch-1 decode, tcgen05 consumer
static bool tcd_usable(const b200llm_engine *e) {
    const b200llm_config &c = e->cfg;
    if (getenv("B200LLM_NO_TCD")) return false;
    bool tiled = e->lm_head_tm != nullptr;
    for (auto &w : e->layers) tiled = tiled && w.wqkv_tm && w.wo_tm && w.wgu_tm && w.wdown_tm;
    if (c.tp_size > 1 && !e->ar_ready) return false;   // tensor parallel needs the peer inboxes (b200llm_tp_import)
    return tiled && tcd_supported(c.d, e->ffn_l, e->nq_l, e->nkv_l, c.head_dim, (e->vocab_l + 127) / 128 * 128, c.n_layers);
}

static int tcd_dataflow_bits(const b200llm_engine *e) {
    // (tensor parallel: the three edges below are local to a rank; the all-reduce phases keep their grid barriers)
    // bit 3 (tensor parallel only): the row-parallel sums are (sum, count) pairs in `delta` and the all-reduce phase's slice
    // owners poll them -- no grid barrier between o_proj / down_proj and their reduction (needs room for d pairs in `delta`)
    const char *df = getenv("B200LLM_TCD_DATAFLOW");
    int bits = df ? atoi(df) : 15;
    if (e->cfg.tp_size == 1 || e->cfg.max_rows < 2) bits &= ~8;
    return bits;
}

// data-flow bits of a launch with `nb` lanes: with 3-4 lanes the gate/up -> down edge goes back to a grid barrier -- the
// (sum, count) pairs are twice the bytes of the sums, every CTA polls its slice of them for every lane, and at 3+
// lanes that costs more than the barrier it saves (3B, 4 lanes: 1.716 against 1.757 ms per step).
// (B200LLM_TCD_EDGE0_LANES: the most lanes that use the edge.)
static int tcd_dataflow_for(const b200llm_engine *e, int nb) {
    static const int edge0_lanes = getenv("B200LLM_TCD_EDGE0_LANES") ? atoi(getenv("B200LLM_TCD_EDGE0_LANES")) : 2;
    return nb > edge0_lanes ? (e->tcd_dataflow & ~1) : e->tcd_dataflow;
}

static int tcd_prepare(b200llm_engine *e) {
    const b200llm_config &c = e->cfg;
    e->tcd_dataflow = tcd_dataflow_bits(e);
    const size_t n_jobs = 4 * (size_t)c.n_layers + 1;
    std::vector<TcdJob> jobs(2 * n_jobs);   // set 0: all edges, set 1: without the gate/up -> down edge (3-4 lanes)
    const int nq = e->nq_l, nkv = e->nkv_l, ffn = e->ffn_l;   // this rank's share (= the model's at tp 1)
    const int qd = nq * c.head_dim;
    const bool tp = c.tp_size > 1;
    // tensor parallel: o_proj and down_proj are row-parallel, their partial sums go to `delta` (cleared by the
    // phase two before, bit 10) and an all-reduce phase adds every rank's delta to h
    float *row_out = tp ? e->delta : e->h;
    const int row_epi = (tp && (e->tcd_dataflow & 8)) ? TCD_EPI_ADD_CNT : TCD_EPI_ADD;
    for (int set = 0; set < 2; ++set) {
        const int df = set == 0 ? e->tcd_dataflow : (e->tcd_dataflow & ~1);   // set 1: without the gate/up -> down edge
        TcdJob *jb = jobs.data() + set * n_jobs;
        for (int l = 0; l < c.n_layers; ++l) {
            const LayerW &w = e->layers[l];
            uint8_t *kl = e->kv + (size_t)l * 2 * e->kv_half_stride;
            jb[4 * l + 0] = TcdJob{w.wqkv_tm, w.attn_norm, kl, kl + e->kv_half_stride,
                                   (df & 4) ? reinterpret_cast<float *>(e->tcd_qkv2) : e->qkv_out, (nq + 2 * nkv) * c.head_dim, c.d,
                                   (df & 4) ? TCD_EPI_ADD_CNT : TCD_EPI_ADD, (l == 0 ? TCD_X_EMBED : TCD_X_H) | 0x400};
            jb[4 * l + 1] = TcdJob{w.wo_tm, nullptr, nullptr, nullptr, row_out, c.d, qd, row_epi, TCD_X_ATT | 0x100};   // 0x100: clears the gate/up sums
            jb[4 * l + 2] = TcdJob{w.wgu_tm, w.mlp_norm, nullptr, nullptr, (df & 1) ? reinterpret_cast<float *>(e->tcd_gu2) : e->gu_out,
                                   2 * ffn, c.d, (df & 1) ? TCD_EPI_ADD_CNT : TCD_EPI_ADD, TCD_X_H | 0x400 | 0x200};   // 0x200: clears the qkv sums
            jb[4 * l + 3] = TcdJob{w.wdown_tm, nullptr, nullptr, nullptr, row_out, c.d, ffn, row_epi, TCD_X_SWIGLU};
        }
        // the lm_head streams whole tiles: N counts the zero rows that pad the last one, p.vocab bounds what is stored
        jb[4 * c.n_layers] = TcdJob{e->lm_head_tm, e->final_norm, nullptr, nullptr, e->logits, (e->vocab_l + 127) / 128 * 128, c.d, TCD_EPI_LOGITS, TCD_X_H};
    }
    B200_CUDA_OK(cudaMemcpyAsync(e->mega_jobs, jobs.data(), jobs.size() * sizeof(TcdJob), cudaMemcpyHostToDevice, e->stream));
    B200_CUDA_OK(cudaStreamSynchronize(e->stream));  // `jobs` is a pageable temporary
    e->tcd_ready = true;
    e->mega_ready = false;  // the two kernels share the descriptor array
    return B200LLM_OK;
}

// One launch of the persistent kernel: n_steps decode steps of `nb` lanes (1..kTcdMaxLanes; several lanes only at tp_size 1).
// in_kernel_sampler: the sampler chain runs as the kernel's last phase with every lane's own settings (a greedy lane is
// temperature 0 there); otherwise all lanes are greedy.
static int tcd_decode(b200llm_engine *e, const int32_t *lanes, int nb, int n_steps, bool in_kernel_sampler) {
    const b200llm_config &c = e->cfg;
    const int lane = lanes[0];
    const b200llm_sampling &sp = e->samp_host[lane];
    if (!e->tcd_ready || e->tcd_dataflow != tcd_dataflow_bits(e)) B200_TRY(tcd_prepare(e));
    TcdParams p{};
    p.dataflow = tcd_dataflow_for(e, nb), p.gu2 = e->tcd_gu2, p.att2 = e->tcd_att2, p.qkv2 = e->tcd_qkv2;
    p.jobs = reinterpret_cast<const TcdJob *>(e->mega_jobs) + ((p.dataflow & 1) == (e->tcd_dataflow & 1) ? 0 : 4 * (size_t)c.n_layers + 1), p.n_layers = c.n_layers;
    p.d = c.d, p.n_q = e->nq_l, p.n_kv = e->nkv_l, p.head_dim = c.head_dim, p.ffn = e->ffn_l, p.vocab = e->vocab_l;
    p.kv_page_heads = e->kv_page_heads;
    p.vocab_global = c.vocab, p.tp_rank = c.tp_rank, p.tp_size = c.tp_size;
    if (c.tp_size > 1) {
        p.delta = e->delta, p.ar_seq = e->ar_seq, p.ar_slot_floats = e->ar_slot_floats, p.ar_max_slices = e->ar_max_slices;
        for (int r = 0; r < c.tp_size; ++r)
            p.ar_inbox[r] = e->ar_inbox[r], p.ar_flags[r] = e->ar_flags[r], p.tok_inbox[r] = e->tok_inbox[r], p.tok_flags[r] = e->tok_flags[r];
    }
    p.rms_eps = c.rms_eps;
    p.scale_log2 = (float)(1.4426950408889634 / sqrt((double)c.head_dim));
    p.h = e->h, p.qkv = e->qkv_out, p.att = e->att, p.gu = e->gu_out, p.logits = e->logits;
    p.part_max = e->part_max, p.part_idx = e->part_idx;
    p.embed = e->embed, p.inv_freq = e->inv_freq;
    p.lane = lane, p.lane_token = e->lane_token, p.lane_pos = e->lane_pos, p.lane_hist = e->lane_hist;
    p.page_table = e->page_table + (size_t)lane * c.max_pages_per_lane;
    p.n_ptab = c.max_pages_per_lane;
    p.gen_out = e->gen_out, p.n_steps = n_steps;
    // attention items = lanes x kv heads x splits, at most one per CTA
    int ns = kNumSMs / (e->nkv_l * nb);
    ns = ns < 1 ? 1 : (ns > e->attn_splits ? e->attn_splits : ns);
    {   // (the att2 block of a lane is sized for the single-lane split count)
        const int ns1 = kNumSMs / e->nkv_l < 1 ? 1 : kNumSMs / e->nkv_l;
        ns = ns > ns1 ? ns1 : ns;
    }
    p.n_splits = ns;
    uint8_t *sc = reinterpret_cast<uint8_t *>(e->attn_scratch);
    p.tickets = reinterpret_cast<unsigned *>(sc);
    p.part_o = reinterpret_cast<float *>(sc + kAttnTicketBytes);
    p.part_ml = p.part_o + (size_t)e->nq_l * ns * c.head_dim;
    p.part_stride = e->nq_l * ns * (c.head_dim + 2);
    p.att2_stride = (int32_t)(e->tcd_att2_lane_bytes / 8);
    p.nb = nb;
    for (int b = 0; b < nb; ++b) {
        const b200llm_sampling &sb = e->samp_host[lanes[b]];
        p.lanes[b] = lanes[b];
        p.page_tables[b] = e->page_table + (size_t)lanes[b] * c.max_pages_per_lane;
        p.temperature_l[b] = sb.temperature > 0.f ? sb.temperature : 0.f, p.top_k_l[b] = sb.top_k, p.top_p_l[b] = sb.top_p;
        p.repeat_penalty_l[b] = sb.repeat_penalty, p.seed_l[b] = sb.seed;
    }
    p.bar_ctr = e->bar_ctr;
    p.sample = 1;
    if (in_kernel_sampler) {   // penalty -> top-k -> top-p -> temperature -> draw as the kernel's last phase
        p.sample = 2, p.cand = e->samp_cand;
        p.temperature = sp.temperature, p.top_k = sp.top_k, p.top_p = sp.top_p, p.repeat_penalty = sp.repeat_penalty;
        p.seed = sp.seed;
    }
    B200_REQUIRE(nb == 1 || (size_t)nb * p.part_stride * 4 + kAttnTicketBytes <= e->attn_scratch_bytes, "tcd: attention scratch too small for %d lanes", nb);
    B200_REQUIRE(tcd_plan(c.d, e->ffn_l, e->nq_l, e->nkv_l, c.head_dim, (e->vocab_l + 127) / 128 * 128, c.n_layers, &p.n_stages, &p.n_bslots, nb),
                 "tcd: geometry does not fit");
    if (const char *ns = getenv("B200LLM_TCD_STAGES")) p.n_stages = (atoi(ns) < p.n_stages ? atoi(ns) : p.n_stages) & ~1;
    if (const char *nf = getenv("B200LLM_TCD_INFLIGHT")) p.n_inflight = atoi(nf);
    p.tmem_slots = 0;   // measured: parking stages in tensor memory during the gaps loses 1-3 % (profiles/tcd_tmem_overflow_r2.txt)
    if (const char *ts = getenv("B200LLM_TCD_TMEM")) p.tmem_slots = atoi(ts) < 0 ? 0 : (atoi(ts) > 12 ? 12 : atoi(ts));
    p.poll_ns = 100;
    if (const char *pn = getenv("B200LLM_TCD_POLL_NS")) p.poll_ns = atoi(pn);
    p.n_poll = 1;
    if (const char *np = getenv("B200LLM_TCD_POLL")) p.n_poll = atoi(np) < 1 ? 1 : (atoi(np) > 12 ? 12 : atoi(np));
    if (getenv("B200LLM_MEGA_TRACE")) p.trace = e->trace, p.trace_step = n_steps > 1 ? 1 : 0;
    // the qkv and gate/up phases ADD into these rows; later layers find them cleared by the o_proj phase
    B200_CUDA_OK(cudaMemsetAsync(e->qkv_out, 0, (size_t)nb * (e->nq_l + 2 * e->nkv_l) * c.head_dim * 4, e->stream));
    B200_CUDA_OK(cudaMemsetAsync(e->gu_out, 0, (size_t)nb * 2 * e->ffn_l * 4, e->stream));
    if (p.dataflow & 1) B200_CUDA_OK(cudaMemsetAsync(e->tcd_gu2, 0, (size_t)nb * 2 * e->ffn_l * 8, e->stream));
    if (p.dataflow & 2) B200_CUDA_OK(cudaMemsetAsync(e->tcd_att2, 0, (size_t)nb * e->tcd_att2_lane_bytes, e->stream));
    if (p.dataflow & 4) B200_CUDA_OK(cudaMemsetAsync(e->tcd_qkv2, 0, (size_t)nb * (e->nq_l + 2 * e->nkv_l) * c.head_dim * 8, e->stream));   // no tag of an earlier launch survives
    e->last_rows = nb, e->last_logit_rows = nb;
    return tcd_decode_launch(p, e->stream);
}

// ---------------------------------------------------------------- batch / prefill pass (tcgen05 GEMM)
static int get_map(b200llm_engine *e, const void *base, int64_t rows, int64_t cols, int box, const void **out, int bk = 64) {
    auto key = std::make_tuple(base, rows, cols, box + 1000 * bk);
    auto it = e->maps.find(key);
    if (it == e->maps.end()) {
        auto *m = new b200llm_engine::MapBox();
        int rc = tc_make_map(m->b, base, rows, cols, box, bk);
        if (rc) {
            delete m;
            return rc;
        }
        it = e->maps.emplace(key, m).first;
    }
    *out = it->second->b;
    return B200LLM_OK;
}

static bool tc_usable(const b200llm_engine *e, int rows) {
    const b200llm_config &c = e->cfg;
    auto okn = [](int n) { return n % 128 == 0; };
    auto okk = [](int k) { return k % 64 == 0; };
    const bool lm_ok = c.tp_size > 1 || okn(c.vocab);   // tensor parallel keeps the GEMV lm_head (partials are gathered)
    // from 3 rows on the GEMM path is faster per step than the per-phase GEMV kernels (3B: 2.4 ms at any batch <= 16 against
    // 2.3 / 3.3 ms at 2 / 4 rows; profiles/harness_r2.txt)
    static const int min_rows = getenv("B200LLM_TC_MIN_ROWS") ? atoi(getenv("B200LLM_TC_MIN_ROWS")) : 3;
    return c.use_tc && rows >= (e->lean ? 1 : min_rows) && rows <= 256 && okn((e->nq_l + 2 * e->nkv_l) * c.head_dim) && okn(c.d) &&
           okn(2 * e->ffn_l) && lm_ok && okk(c.d) && okk(e->nq_l * c.head_dim) && okk(e->ffn_l) &&
           getenv("B200LLM_NO_TC") == nullptr;
}

// out (+)= X W^T through the tcgen05 kernel; X = e->xb viewed as [max_rows, K]
static int tc_gemm(b200llm_engine *e, const void *w, int N, int K, float *out, int64_t out_stride, int T,
                   bool add_to_existing, bool prezeroed = false, const void *x_buf = nullptr, void *out_swiglu = nullptr,
                   const void *w_tm = nullptr) {
    const void *mw, *mx, *mxl = nullptr;
    const uint16_t *xb = x_buf ? reinterpret_cast<const uint16_t *>(x_buf) : e->xb;
    const int64_t lo = xb == e->xb2 ? e->xb2_lo : e->xb_lo;
    const int bk = tc_bk(T, lo != 0);
    B200_TRY(get_map(e, xb, e->cfg.max_rows, K, tc_x_box(N, T), &mx, bk));
    if (w != nullptr) {
        B200_TRY(get_map(e, w, N, K, 128, &mw, bk));
    } else {   // lean engine: the weight tiles come from the tile-major copy, the kernel never touches its weight map
        B200_REQUIRE(w_tm != nullptr, "gemm: neither a row-major nor a tile-major matrix");
        mw = mx;
    }
    if (lo != 0) B200_TRY(get_map(e, xb + lo, e->cfg.max_rows, K, tc_x_box(N, T), &mxl, bk));
    int accumulate = 1;
    if (out_swiglu != nullptr) {
        accumulate = 0;
    } else if (!add_to_existing) {
        if (N / 128 >= kNumSMs) {
            accumulate = 0;  // enough row tiles: no k split, plain stores
        } else if (!prezeroed) {
            B200_CUDA_OK(cudaMemsetAsync(out, 0, (size_t)T * out_stride * 4, e->stream));
        }
    }
    // the tile-major copy (bound for the persistent decode kernel) serves this kernel too: one 16 KB bulk copy per weight tile
    static const bool use_tm = !(getenv("B200LLM_TC_TM") && atoi(getenv("B200LLM_TC_TM")) == 0);
    if (!use_tm || bk != 64 || tc_cluster_size(N, T) != 1) {
        B200_REQUIRE(w != nullptr, "gemm: this engine holds tile-major matrices only (B200LLM_TC_TM=0, B200LLM_TC_BK=32 and B200LLM_TC_CLUSTER need the row-major ones)");
        w_tm = nullptr;
    }
    B200_TRY(tc_gemm_launch(mw, mx, out, out_stride, T, N, K, accumulate, e->stream, e->tc_pdl, out_swiglu, mxl,
                            out_swiglu != nullptr ? e->xb2_lo : 0, bk, w_tm));
    e->launches++;
    return B200LLM_OK;
}

static int forward_pass_tc(b200llm_engine *e, const PassRows &pr, bool advance) {
    const b200llm_config &c = e->cfg;
    const int rows = pr.rows;
    cudaStream_t st = e->stream;
    const bool pdl = e->tc_pdl;
    EmbedArgs ea{};
    ea.embed = e->embed, ea.h = e->h, ea.row_token = pr.row_token, ea.row_pos = pr.row_pos, ea.row_lane = pr.row_lane;
    ea.lane_token = e->lane_token, ea.lane_pos = e->lane_pos, ea.rows = rows, ea.d = c.d, ea.vocab = c.vocab;
    ea.from_lanes = pr.from_lanes ? 1 : 0;
    B200_TRY(embed_launch_ex(ea, e->lane_hist, pr.from_lanes ? e->step_ctr : nullptr, st, pdl));
    e->launches++;
    // this rank's share of heads and MLP units (the whole model at tp 1); o_proj and down_proj are row-parallel
    // under tensor parallel: their GEMMs fill `delta`, the all-reduce kernel adds every rank's delta to h
    const bool tp = c.tp_size > 1;
    const int n_q = e->nq_l, n_kv = e->nkv_l, ffn = e->ffn_l;
    const int qd = n_q * c.head_dim, qkv_w = (n_q + 2 * n_kv) * c.head_dim;
    for (int l = 0; l < c.n_layers; ++l) {
        const LayerW &w = e->layers[l];
        uint8_t *kl = e->kv + (size_t)l * 2 * e->kv_half_stride;
        uint8_t *vl = kl + e->kv_half_stride;
        B200_TRY(rmsnorm_bf16_launch(e->h, w.attn_norm, e->xb, rows, c.d, c.rms_eps, st, pdl, e->qkv_out, qkv_w, e->xb_lo));
        B200_TRY(tc_gemm(e, w.wqkv, qkv_w, c.d, e->qkv_out, qkv_w, rows, false, true, nullptr, nullptr, w.wqkv_tm));
        B200_TRY(rope_append_launch(e->qkv_out, e->q, kl, vl, e->inv_freq, pr.row_lane, pr.row_pos, e->page_table, rows,
                                    n_q, n_kv, c.head_dim, c.page_tokens, c.max_pages_per_lane, st, pdl, e->kv_page_heads));
        b200llm_attn_args at{};
        at.q = e->q, at.out = e->att, at.k_pages = kl, at.v_pages = vl, at.row_lane = pr.row_lane, at.row_pos = pr.row_pos;
        at.page_table = e->page_table, at.bs = rows, at.n_q = n_q, at.n_kv = n_kv, at.head_dim = c.head_dim;
        at.page_tokens = c.page_tokens, at.max_pages_per_lane = c.max_pages_per_lane, at.kv_page_heads = e->kv_page_heads;
        int want = (kNumSMs + rows * n_kv - 1) / (rows * n_kv);
        if (want > 6) want = 6;
        static const int forced_splits = getenv("B200LLM_BATCH_SPLITS") ? atoi(getenv("B200LLM_BATCH_SPLITS")) : 0;
        if (forced_splits > 0 && pr.lane < 0) want = forced_splits;
        at.n_splits = want < 1 ? 1 : (want > e->attn_splits ? e->attn_splits : want);
        at.scratch = e->attn_scratch, at.scratch_bytes = e->attn_scratch_bytes;
        // (short contexts: the decode-style kernel, one CTA group per row, still wins -- 296 ids: 8.9 vs 9.7 ms; from ~512 keys on
        // the tiled kernel does -- 1024 ids: 27.8 vs 30.2 ms, 7B at 2072 ids: 132 vs 182 ms)
        // (it runs one CTA per (128-query tile, q head): with few heads per rank -- tensor parallel 4 / 8 of the 7B: 16 / 8 CTAs per
        // 256-row chunk -- the decode-style kernel's rows x kv heads x splits CTAs are faster: TTFT at TP 8 134 against 118 ms)
        if (pr.lane >= 0 && pr.pos0 + rows >= 512 && n_q * ((rows + 127) / 128) >= 32 && getenv("B200LLM_NO_PREFILL_ATTN") == nullptr) {
            // prefill chunk: query-tiled causal attention on the tensor cores (attn_prefill.cu) instead of one decode-style
            // pass over the whole prefix per query row; output as bf16 hi + lo straight into the o_proj operand
            // rows of [head_dim] bf16 from the layer's K (V) base to the end of what page ids can reach: the layer's own region in
            // a pool of this engine's own, the rest of the whole pool when pages are page-major slots of a shared one
            const int64_t all_rows = (int64_t)e->n_pages * e->kv_page_heads * c.page_tokens;
            const int64_t k_rows = e->kv_shared ? all_rows - (int64_t)((kl - e->kv) / ((size_t)c.head_dim * 2)) : all_rows;
            const int64_t v_rows = e->kv_shared ? all_rows - (int64_t)((vl - e->kv) / ((size_t)c.head_dim * 2)) : all_rows;
            const void *mk, *mv;
            B200_TRY(get_map(e, kl, k_rows, c.head_dim, c.page_tokens, &mk));
            B200_TRY(get_map(e, vl, v_rows, c.head_dim, c.page_tokens, &mv));
            B200_TRY(attn_prefill_launch(mk, mv, e->q, nullptr, e->xb, e->xb_lo, e->page_table + (size_t)pr.lane * c.max_pages_per_lane,
                                         c.max_pages_per_lane, rows, pr.pos0, n_q, n_kv, c.head_dim, e->kv_page_heads, st, pdl));
        } else if (pr.lane >= 0 && rows >= 4 && (n_kv * ((rows + 3) / 4) >= 1000 || getenv("B200LLM_ROWS_ATTN") != nullptr) &&
                   getenv("B200LLM_NO_ROWS_ATTN") == nullptr) {
            // prefill chunk at a short context, many kv heads (the 7B: 32): four rows per CTA share every K/V tile
            // (attn_rows_kernel) instead of one CTA per row streaming the row's whole prefix from L2.  A warp there walks all
            // pages of its row, so it needs the CTAs to fill the GPU twice over: 7B at 296 ids 18.3 -> 17.4 ms, but the 3B
            // (8 kv heads: 320 CTAs) 8.42 -> 8.64 ms, which therefore keeps the kernel below
            B200_TRY(attn_rows_launch(at, st, pdl, e->xb, e->xb_lo));
        } else {
            B200_TRY(attn_decode_launch(at, st, pdl, e->xb, e->xb_lo));  // output rounded to bf16 straight into the o_proj operand
        }
        if (tp) {
            B200_TRY(tc_gemm(e, w.wo, c.d, qd, e->delta, c.d, rows, false, false, nullptr, nullptr, w.wo_tm));
            B200_TRY(tp_allreduce_add(e, rows));
        } else {
            B200_TRY(tc_gemm(e, w.wo, c.d, qd, e->h, c.d, rows, true, false, nullptr, nullptr, w.wo_tm));
        }
        // gate/up: with >= 74 row tiles the GEMM runs without a k split and applies SwiGLU in its epilogue
        const bool fuse_swiglu = 2 * ffn / 128 > kNumSMs / 2;
        if (fuse_swiglu) {
            B200_TRY(rmsnorm_bf16_launch(e->h, w.mlp_norm, e->xb, rows, c.d, c.rms_eps, st, pdl, nullptr, 0, e->xb_lo));
            B200_TRY(tc_gemm(e, w.wgu, 2 * ffn, c.d, e->gu_out, 2 * ffn, rows, false, true, nullptr, e->xb2, w.wgu_tm));
            if (tp) {
                B200_TRY(tc_gemm(e, w.wdown, c.d, ffn, e->delta, c.d, rows, false, false, e->xb2, nullptr, w.wdown_tm));
                B200_TRY(tp_allreduce_add(e, rows));
            } else {
                B200_TRY(tc_gemm(e, w.wdown, c.d, ffn, e->h, c.d, rows, true, false, e->xb2, nullptr, w.wdown_tm));
            }
            e->launches += 4;
        } else {
            B200_TRY(rmsnorm_bf16_launch(e->h, w.mlp_norm, e->xb, rows, c.d, c.rms_eps, st, pdl, e->gu_out, 2 * ffn, e->xb_lo));
            B200_TRY(tc_gemm(e, w.wgu, 2 * ffn, c.d, e->gu_out, 2 * ffn, rows, false, true, nullptr, nullptr, w.wgu_tm));
            B200_TRY(swiglu_bf16_launch(e->gu_out, e->xb, rows, ffn, st, pdl, e->xb_lo));
            if (tp) {
                B200_TRY(tc_gemm(e, w.wdown, c.d, ffn, e->delta, c.d, rows, false, false, nullptr, nullptr, w.wdown_tm));
                B200_TRY(tp_allreduce_add(e, rows));
            } else {
                B200_TRY(tc_gemm(e, w.wdown, c.d, ffn, e->h, c.d, rows, true, false, nullptr, nullptr, w.wdown_tm));
            }
            e->launches += 5;
        }
    }
    e->last_rows = rows;
    e->last_logit_rows = pr.logit_rows;
    if (pr.logit_rows <= 0) return B200LLM_OK;
    if (tp) return lm_head_and_sample(e, pr, advance);
    B200_TRY(rmsnorm_bf16_launch(e->h + (size_t)pr.logit_row0 * c.d, e->final_norm, e->xb, pr.logit_rows, c.d, c.rms_eps,
                                 st, pdl, nullptr, 0, e->xb_lo));
    B200_TRY(tc_gemm(e, e->lm_head, c.vocab, c.d, e->logits, c.vocab, pr.logit_rows, false, false, nullptr, nullptr, e->lm_head_tm));
    SampleDev sd{};
    sd.a.logits = e->logits, sd.a.bs = pr.logit_rows, sd.a.vocab = c.vocab;
    sd.a.part_max = nullptr, sd.a.part_idx = nullptr, sd.a.n_part = 0;  // full-vocabulary argmax over the logits
    sd.a.params = e->lane_samp, sd.a.out_token = nullptr;
    sd.row_lane = pr.row_lane + pr.logit_row0;
    sd.lane_token = e->lane_token, sd.lane_pos = e->lane_pos, sd.lane_hist = e->lane_hist;
    sd.gen_out = pr.from_lanes ? e->gen_out : nullptr, sd.step_ctr = pr.from_lanes ? e->step_ctr : nullptr;
    sd.advance = advance ? 1 : 0;
    B200_TRY(sample_launch(sd, st, pdl));
    e->launches += 2;
    return B200LLM_OK;
}

static int forward_pass(b200llm_engine *e, const PassRows &pr, bool advance) {
    if (tc_usable(e, pr.rows)) return forward_pass_tc(e, pr, advance);
    B200_REQUIRE(!e->lean, "this pass (%d rows) needs the GEMV kernels, which read the row-major matrices this engine was not given", pr.rows);
    return forward_pass_gemv(e, pr, advance);
}

extern "C" {

int b200llm_prefill(b200llm_engine *e, int32_t lane, const int32_t *ids, int32_t n, int32_t pos0) {
    B200_TRY(require_ready(e));
    const b200llm_config &c = e->cfg;
    B200_REQUIRE(ids && n >= 1 && lane >= 0 && lane < c.max_lanes, "prefill: bad arguments");
    B200_REQUIRE(pos0 >= 0 && pos0 + n <= e->max_ctx, "prefill: %d tokens at pos %d exceed the lane's %d-token context",
                 n, pos0, e->max_ctx);
    B200_REQUIRE(pos0 + n <= e->lane_pages_host[lane] * c.page_tokens,
                 "prefill: %d tokens at pos %d need %d pages, lane %d has %d (b200llm_set_page_table)", n, pos0,
                 (pos0 + n + c.page_tokens - 1) / c.page_tokens, lane, e->lane_pages_host[lane]);
    e->lane_pos_host[lane] = pos0 + n;
    std::vector<int32_t> pos(n), lanes(n, lane);
    for (int i = 0; i < n; ++i) pos[i] = pos0 + i;
    B200_CUDA_OK(cudaEventRecord(e->ev[0], e->stream));
    B200_CUDA_OK(cudaMemcpyAsync(e->pf_token, ids, (size_t)n * 4, cudaMemcpyHostToDevice, e->stream));
    B200_CUDA_OK(cudaMemcpyAsync(e->pf_pos, pos.data(), (size_t)n * 4, cudaMemcpyHostToDevice, e->stream));
    B200_CUDA_OK(cudaMemcpyAsync(e->pf_lane, lanes.data(), (size_t)n * 4, cudaMemcpyHostToDevice, e->stream));
    const int last = pos0 + n - 1;
    B200_CUDA_OK(cudaMemcpyAsync(e->lane_pos + lane, &last, 4, cudaMemcpyHostToDevice, e->stream));
    if (pos0 == 0) B200_CUDA_OK(cudaMemsetAsync(e->lane_hist + (size_t)lane * 64, 0xFF, 64 * 4, e->stream));
    // pageable sources above are staged by the runtime before the call returns
    // chunks of equal size (a multiple of 16 rows) rather than full ones and a remainder: on the GEMM path a pass costs its
    // weight stream plus an activation re-read that grows with the rows, so 296 ids as 160 + 136 beat 256 + 40
    int chunk = c.max_rows;
    static const bool balanced = !(getenv("B200LLM_PREFILL_BALANCED") && atoi(getenv("B200LLM_PREFILL_BALANCED")) == 0);
    if (balanced && n > c.max_rows) {
        const int n_chunks = (n + c.max_rows - 1) / c.max_rows;
        chunk = ((n + n_chunks - 1) / n_chunks + 15) / 16 * 16;
        if (chunk > c.max_rows) chunk = c.max_rows;
    }
    for (int c0 = 0; c0 < n; c0 += chunk) {
        PassRows pr{};
        pr.rows = n - c0 < chunk ? n - c0 : chunk;
        pr.row_lane = e->pf_lane + c0, pr.row_pos = e->pf_pos + c0, pr.row_token = e->pf_token + c0;
        pr.from_lanes = false;
        pr.lane = lane, pr.pos0 = pos0 + c0;
        const bool is_last = c0 + pr.rows == n;
        pr.logit_row0 = pr.rows - 1;
        pr.logit_rows = is_last ? 1 : 0;
        B200_TRY(forward_pass(e, pr, true));
    }
    B200_CUDA_OK(cudaEventRecord(e->ev[1], e->stream));
    e->timing.prefill_tokens = n;
    return B200LLM_OK;
}

int b200llm_decode(b200llm_engine *e, const int32_t *lanes, int32_t n_lanes, int32_t n_steps) {
    B200_TRY(require_ready(e));
    const b200llm_config &c = e->cfg;
    B200_REQUIRE(lanes && n_lanes >= 1 && n_lanes <= c.max_rows, "decode: n_lanes %d out of range [1, %d]", n_lanes,
                 c.max_rows);
    B200_REQUIRE(n_steps >= 1 && n_steps <= kMaxDecodeSteps, "decode: n_steps %d out of range [1, %d]", n_steps,
                 kMaxDecodeSteps);
    for (int i = 0; i < n_lanes; ++i)
        B200_REQUIRE(lanes[i] >= 0 && lanes[i] < c.max_lanes, "decode: lane %d out of range", lanes[i]);
    // every K/V row the call appends must land in a page the caller has put in the lane's table
    for (int i = 0; i < n_lanes; ++i)
        B200_REQUIRE(e->lane_pos_host[lanes[i]] + n_steps <= e->lane_pages_host[lanes[i]] * c.page_tokens,
                     "decode: lane %d at pos %d + %d steps exceeds its %d pages (b200llm_set_page_table)", lanes[i],
                     e->lane_pos_host[lanes[i]], n_steps, e->lane_pages_host[lanes[i]]);
    for (int i = 0; i < n_lanes; ++i) e->lane_pos_host[lanes[i]] += n_steps;
    B200_CUDA_OK(cudaMemcpyAsync(e->row_lane, lanes, (size_t)n_lanes * 4, cudaMemcpyHostToDevice, e->stream));
    B200_CUDA_OK(cudaMemsetAsync(e->step_ctr, 0, 4, e->stream));
    PassRows pr{};
    pr.rows = n_lanes, pr.row_lane = e->row_lane, pr.row_pos = e->row_pos, pr.row_token = e->row_token;
    pr.from_lanes = true, pr.logit_row0 = 0, pr.logit_rows = n_lanes;

    B200_CUDA_OK(cudaEventRecord(e->ev[2], e->stream));
    const b200llm_sampling &sp0 = e->samp_host[lanes[0]];
    const bool greedy = sp0.temperature <= 0.f && sp0.repeat_penalty == 1.f;
    // the reference's calls pass no options (Ollama's defaults: temperature 0.8, top-k 40, top-p 0.9, penalty 1.1):
    // the persistent kernel samples in its last phase, so they are one launch per call like greedy ones
    const bool in_kernel = !greedy && c.tp_size == 1 && getenv("B200LLM_NO_TCD_SAMPLER") == nullptr &&
                           tcd_sampler_supported((e->vocab_l + 127) / 128 * 128, c.head_dim, e->nq_l / e->nkv_l,
                                                 sp0.temperature, sp0.top_k);
    // 2..4 lanes ride the same kernel (the operand block has room for four lanes' hi / lo rows): every lane greedy, or
    // every non-greedy lane within what the in-kernel sampler takes
    static const int tcd_max_lanes = getenv("B200LLM_TCD_LANES") ? atoi(getenv("B200LLM_TCD_LANES")) : kTcdMaxLanes;
    bool multi = n_lanes >= 2 && n_lanes <= kTcdMaxLanes && n_lanes <= tcd_max_lanes && c.tp_size == 1, multi_sampled = false;
    for (int i = 0; multi && i < n_lanes; ++i) {
        const b200llm_sampling &si = e->samp_host[lanes[i]];
        const bool g = si.temperature <= 0.f && si.repeat_penalty == 1.f;
        if (!g) {
            multi_sampled = true;
            multi = getenv("B200LLM_NO_TCD_SAMPLER") == nullptr && si.repeat_penalty > 0.f &&
                    tcd_sampler_supported((e->vocab_l + 127) / 128 * 128, c.head_dim, e->nq_l / e->nkv_l, si.temperature, si.top_k);
        }
        for (int j = 0; j < i; ++j) multi = multi && lanes[j] != lanes[i];
    }
    if (c.use_megakernel && multi && tcd_usable(e)) {
        B200_TRY(tcd_decode(e, lanes, n_lanes, n_steps, multi_sampled));
        e->timing.kernel_launches = 1;
    } else if (c.use_megakernel && n_lanes == 1 && (greedy || in_kernel) && tcd_usable(e)) {
        B200_TRY(tcd_decode(e, lanes, 1, n_steps, in_kernel));
        e->timing.kernel_launches = 1;
    } else if (c.use_megakernel && n_lanes == 1 && c.tp_size == 1 && greedy && !e->lean &&
               mega_supported(c.d, c.ffn, c.n_q, c.n_kv, c.head_dim)) {
        B200_TRY(mega_decode(e, lanes[0], n_steps));
        e->timing.kernel_launches = 1;
    } else if (c.use_graphs) {
        auto it = e->graphs.find(n_lanes);
        if (it == e->graphs.end()) {
            cudaGraph_t graph = nullptr;
            e->launches = 0;
            B200_CUDA_OK(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
            int rc = forward_pass(e, pr, true);
            cudaError_t ce = cudaStreamEndCapture(e->stream, &graph);
            if (rc) {
                if (graph) cudaGraphDestroy(graph);
                return rc;
            }
            B200_CUDA_OK(ce);
            cudaGraphExec_t exec = nullptr;
            B200_CUDA_OK(cudaGraphInstantiate(&exec, graph, 0));
            cudaGraphDestroy(graph);
            it = e->graphs.emplace(n_lanes, exec).first;
            e->graph_launches[n_lanes] = e->launches;
        }
        for (int s = 0; s < n_steps; ++s) B200_CUDA_OK(cudaGraphLaunch(it->second, e->stream));
        e->timing.kernel_launches = e->graph_launches[n_lanes] * n_steps;
    } else {
        e->launches = 0;
        for (int s = 0; s < n_steps; ++s) B200_TRY(forward_pass(e, pr, true));
        e->timing.kernel_launches = e->launches;
    }
    B200_CUDA_OK(cudaEventRecord(e->ev[3], e->stream));
    e->timing.decode_steps = n_steps;
    return B200LLM_OK;
}

int b200llm_read_tokens(b200llm_engine *e, int32_t *out, int32_t n_steps, int32_t n_lanes) {
    B200_TRY(require_ready(e));
    B200_REQUIRE(out && n_steps >= 1 && n_steps <= kMaxDecodeSteps && n_lanes >= 1 && n_lanes <= e->cfg.max_rows,
                 "read_tokens: bad arguments");
    B200_CUDA_OK(cudaMemcpyAsync(out, e->gen_out, (size_t)n_steps * n_lanes * 4, cudaMemcpyDeviceToHost, e->stream));
    B200_CUDA_OK(cudaStreamSynchronize(e->stream));
    return tp_check_peers(e);
}

int b200llm_read_lane(b200llm_engine *e, int32_t lane, int32_t *next_token, int32_t *pos) {
    B200_TRY(require_ready(e));
    B200_REQUIRE(lane >= 0 && lane < e->cfg.max_lanes, "read_lane: lane out of range");
    int32_t t = 0, p = 0;
    B200_CUDA_OK(cudaMemcpyAsync(&t, e->lane_token + lane, 4, cudaMemcpyDeviceToHost, e->stream));
    B200_CUDA_OK(cudaMemcpyAsync(&p, e->lane_pos + lane, 4, cudaMemcpyDeviceToHost, e->stream));
    B200_CUDA_OK(cudaStreamSynchronize(e->stream));
    if (next_token) *next_token = t;
    if (pos) *pos = p;
    return B200LLM_OK;
}

int b200llm_step_logits(b200llm_engine *e, int32_t row, float *out, int32_t n) {
    B200_TRY(require_ready(e));
    B200_REQUIRE(out && row >= 0 && row < e->last_logit_rows && n >= 1 && n <= e->vocab_l,
                 "step_logits: row %d / n %d out of range (last pass had %d logit rows, %d vocab rows)", row, n,
                 e->last_logit_rows, e->vocab_l);
    B200_CUDA_OK(cudaMemcpyAsync(out, e->logits + (size_t)row * e->vocab_l, (size_t)n * 4, cudaMemcpyDeviceToHost,
                                 e->stream));
    B200_CUDA_OK(cudaStreamSynchronize(e->stream));
    return B200LLM_OK;
}

int b200llm_sync(b200llm_engine *e) {
    B200_REQUIRE(e, "engine is NULL");
    B200_CUDA_OK(cudaStreamSynchronize(e->stream));
    return tp_check_peers(e);
}

int b200llm_last_timing(b200llm_engine *e, b200llm_timing *t) {
    B200_REQUIRE(e && t, "last_timing: bad arguments");
    B200_CUDA_OK(cudaStreamSynchronize(e->stream));
    if (e->timing.prefill_tokens > 0) cudaEventElapsedTime(&e->timing.prefill_ms, e->ev[0], e->ev[1]);
    if (e->timing.decode_steps > 0) cudaEventElapsedTime(&e->timing.decode_ms, e->ev[2], e->ev[3]);
    cudaGetLastError();
    *t = e->timing;
    return B200LLM_OK;
}

void *b200llm_stream(b200llm_engine *e) { return e ? (void *)e->stream : nullptr; }

int b200llm_debug_tp_allreduce_us(b200llm_engine *e, int32_t rows, int32_t iters, float *us_per_call) {
    B200_TRY(require_ready(e));
    B200_REQUIRE(us_per_call && e->cfg.tp_size > 1 && rows >= 1 && rows <= e->cfg.max_rows && iters >= 1 && iters <= 100000,
                 "debug_tp_allreduce_us: needs a tensor-parallel engine, 1 <= rows <= max_rows, 1 <= iters <= 100000");
    // `iters` reductions of a [rows, d] fp32 block back to back, as the per-phase path issues them (collective: every
    // rank calls this with the same arguments); the residual stream is scratch here and is left changed
    cudaEvent_t a, b;
    B200_CUDA_OK(cudaEventCreate(&a));
    B200_CUDA_OK(cudaEventCreate(&b));
    B200_CUDA_OK(cudaMemsetAsync(e->delta, 0, (size_t)rows * e->cfg.d * 4, e->stream));
    B200_CUDA_OK(cudaEventRecord(a, e->stream));
    for (int i = 0; i < iters; ++i) B200_TRY(tp_allreduce_add(e, rows));
    B200_CUDA_OK(cudaEventRecord(b, e->stream));
    B200_CUDA_OK(cudaStreamSynchronize(e->stream));
    float ms = 0.f;
    B200_CUDA_OK(cudaEventElapsedTime(&ms, a, b));
    cudaEventDestroy(a), cudaEventDestroy(b);
    *us_per_call = ms * 1000.f / (float)iters;
    return tp_check_peers(e);
}

int b200llm_debug_read_trace(b200llm_engine *e, uint64_t *out, int32_t n) {
    B200_REQUIRE(e && e->ws && out && n >= 1 && n <= 4096, "debug_read_trace: bad arguments");
    B200_CUDA_OK(cudaMemcpyAsync(out, e->trace, (size_t)n * 8, cudaMemcpyDeviceToHost, e->stream));
    B200_CUDA_OK(cudaStreamSynchronize(e->stream));
    return B200LLM_OK;
}

// ---------------------------------------------------------------- single-kernel entry points
int b200llm_op_gemv(const b200llm_gemv_args *a, void *stream) {
    B200_REQUIRE(a && a->w && a->x && a->out, "op_gemv: NULL argument");
    int rc = require_device_any();
    if (rc) return rc;
    return gemv_launch(*a, (cudaStream_t)stream, false);
}

int b200llm_op_gemv_grid(int32_t N, int32_t K, int32_t bs, int32_t epi, int32_t *grid) {
    B200_REQUIRE(grid, "op_gemv_grid: NULL argument");
    int g = gemv_grid(N, K, bs, epi, 0);
    B200_REQUIRE(g > 0, "op_gemv_grid: unsupported shape");
    *grid = g;
    return B200LLM_OK;
}

int b200llm_op_gemv_tm(const void *w_tm, const float *x, float *out, int32_t N, int32_t K, int32_t k_split,
                       void *stream) {
    B200_REQUIRE(w_tm && x && out, "op_gemv_tm: NULL argument");
    B200_REQUIRE(tcd_job_fits(N, K, !k_split), "op_gemv_tm: N %% 128, K %% 64, or the per-CTA share does not fit");
    int rc = require_device_any();
    if (rc) return rc;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    // scratch of this test entry point: one descriptor, per-CTA argmax partials, lane state, barrier counter
    uint8_t *scr = nullptr;
    B200_CUDA_OK(cudaMalloc(&scr, 8192));
    B200_CUDA_OK(cudaMemsetAsync(scr, 0, 8192, st));
    TcdJob job{w_tm, nullptr, nullptr, nullptr, out, N, K, k_split ? TCD_EPI_ADD : TCD_EPI_LOGITS, TCD_X_H};
    B200_CUDA_OK(cudaMemcpyAsync(scr, &job, sizeof(job), cudaMemcpyHostToDevice, st));
    if (k_split) B200_CUDA_OK(cudaMemsetAsync(out, 0, (size_t)N * 4, st));
    TcdParams p{};
    p.jobs = reinterpret_cast<const TcdJob *>(scr), p.n_layers = 0;
    p.d = K, p.n_q = 1, p.n_kv = 1, p.head_dim = 128, p.ffn = 0, p.vocab = N;
    p.kv_page_heads = 1;
    p.vocab_global = N, p.tp_rank = 0, p.tp_size = 1;
    p.h = const_cast<float *>(x), p.logits = out;
    p.part_max = reinterpret_cast<float *>(scr + 1024), p.part_idx = reinterpret_cast<int32_t *>(scr + 2048);
    p.lane = 0, p.lane_token = reinterpret_cast<int32_t *>(scr + 3072), p.lane_pos = reinterpret_cast<int32_t *>(scr + 3200);
    p.lane_hist = reinterpret_cast<int32_t *>(scr + 3328);
    p.page_table = reinterpret_cast<const int32_t *>(scr + 3840), p.n_ptab = 0;
    p.inv_freq = reinterpret_cast<const float *>(scr + 4096);
    p.embed = x;
    p.gen_out = reinterpret_cast<int32_t *>(scr + 3712), p.n_steps = 1, p.n_splits = 1;
    p.bar_ctr = reinterpret_cast<unsigned *>(scr + 3776);
    p.sample = 0, p.split_all = k_split ? 1 : 0;
    p.nb = 1, p.lanes[0] = 0, p.page_tables[0] = p.page_table, p.repeat_penalty_l[0] = 1.f;
    p.n_poll = 1, p.tmem_slots = 0;
    p.n_bslots = tcd_job_slots(N, K, !k_split);
    p.n_stages = tcd_stages_for(128, 1, 1, p.n_bslots);
    rc = tcd_decode_launch(p, st);
    cudaError_t ce = cudaStreamSynchronize(st);
    cudaFree(scr);
    if (rc) return rc;
    B200_CUDA_OK(ce);
    return B200LLM_OK;
}

int b200llm_op_attn_decode(const b200llm_attn_args *a, void *stream) {
    B200_REQUIRE(a && a->q && a->out && a->k_pages && a->v_pages && a->row_lane && a->row_pos && a->page_table,
                 "op_attn_decode: NULL argument");
    int rc = require_device_any();
    if (rc) return rc;
    return attn_decode_launch(*a, (cudaStream_t)stream, false);
}

int b200llm_op_attn_prefill(const b200llm_attn_prefill_args *a, void *stream) {
    B200_REQUIRE(a && a->q && a->out && a->k_pages && a->v_pages && a->page_table_row, "op_attn_prefill: NULL argument");
    int rc = require_device_any();
    if (rc) return rc;
    const int kvph = a->kv_page_heads > 0 ? a->kv_page_heads : a->n_kv;
    alignas(128) unsigned char mk[128], mv[128];
    B200_TRY(tc_make_map(mk, a->k_pages, (int64_t)a->n_pages * kvph * 16, a->head_dim, 16));
    B200_TRY(tc_make_map(mv, a->v_pages, (int64_t)a->n_pages * kvph * 16, a->head_dim, 16));
    return attn_prefill_launch(mk, mv, a->q, a->out, nullptr, 0, a->page_table_row, a->n_ptab, a->rows, a->pos0, a->n_q, a->n_kv,
                               a->head_dim, kvph, (cudaStream_t)stream, false);
}

int b200llm_op_attn_scratch_bytes(int32_t bs, int32_t n_q, int32_t n_kv, int32_t head_dim, int32_t n_splits,
                                  size_t *bytes) {
    B200_REQUIRE(bytes && bs >= 1 && n_q >= 1 && n_kv >= 1 && n_splits >= 1, "op_attn_scratch_bytes: bad arguments");
    *bytes = attn_scratch_bytes(bs, n_q, n_kv, head_dim, n_splits);
    return B200LLM_OK;
}

int b200llm_op_gemm_tc(const void *w, const void *x, float *out, int64_t out_stride, int32_t T, int32_t N, int32_t K,
                       int32_t accumulate, void *stream) {
    B200_REQUIRE(w && x && out, "op_gemm_tc: NULL argument");
    int rc = require_device_any();
    if (rc) return rc;
    alignas(128) unsigned char mw[128], mx[128];
    const int bk = tc_bk(T, false);
    B200_TRY(tc_make_map(mw, w, N, K, 128, bk));
    B200_TRY(tc_make_map(mx, x, T, K, tc_x_box(N, T), bk));
    return tc_gemm_launch(mw, mx, out, out_stride, T, N, K, accumulate, (cudaStream_t)stream, false, nullptr, nullptr, 0, bk);
}

int b200llm_op_sample(const b200llm_sample_args *a, void *stream) {
    B200_REQUIRE(a && a->logits && a->out_token, "op_sample: NULL argument");
    int rc = require_device_any();
    if (rc) return rc;
    SampleDev d{};
    d.a = *a;
    return sample_launch(d, (cudaStream_t)stream, false);
}

int b200llm_op_fill_weights(void *dev_bf16, uint64_t n, uint64_t seed, float sigma, float mean, void *stream) {
    B200_REQUIRE(dev_bf16 != nullptr || n == 0, "op_fill_weights: NULL argument");
    int rc = require_device_any();
    if (rc) return rc;
    return fill_weights_launch(dev_bf16, n, seed, sigma, mean, (cudaStream_t)stream);
}

int b200llm_op_dequant_bf16(const void *blocks, int32_t ggml_type, int64_t n_elements, void *out_bf16, void *stream) {
    int rc = require_device_any();
    if (rc) return rc;
    return dequant_bf16_launch(blocks, ggml_type, n_elements, out_bf16, (cudaStream_t)stream);
}

}  // extern "C"
