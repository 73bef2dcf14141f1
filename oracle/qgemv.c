/* oracle/qgemv.c -- TEST / BASELINE INFRASTRUCTURE, not product code.
 *
 * A restatement of the arithmetic llama.cpp runs on the CPU for one decode step of a quantised GGUF model -- the path
 * the reference really uses (Ollama -> llama.cpp; `ollama.generate`, /root/reference/Model_Evaluation_&_Comparision.py:113-119
 * times it).  llama.cpp is not in /root/reference and cannot be built here, so this file restates the published block
 * formats and dot products of ggml (ggml-quants.c / ggml-common.h, as documented by gguf-py 0.19 `quants.py`, which the
 * tests use as the checker):
 *
 *   Q4_K  super-block of 256 weights, 144 bytes: fp16 d, fp16 dmin, 12 bytes of 6-bit (scale, min) pairs for 8 sub-blocks
 *         of 32, 128 bytes of 4-bit quants.  w = d * sc_j * q - dmin * m_j.           (llama3.2's Q4_K_M matrices)
 *   Q8_K  activation block of 256: fp32 d, 256 int8, 16 int16 sums of 16.
 *   Q4_0  block of 32 weights, 18 bytes: fp16 d, 16 bytes of 4-bit quants.  w = d * (q - 8).   (duckdb-nsql's Q4_0)
 *   Q8_0  activation block of 32: fp16 d, 32 int8.
 *
 * y = W x is computed as llama.cpp does: x is quantised once per matrix input (Q8_K / Q8_0), every row is an integer
 * dot product per block, scaled in fp32.  Rows are spread over OpenMP threads.  The AVX-512 VNNI path (vpdpbusd) is chosen
 * at run time; the scalar path is the definition and the checker of the SIMD path.
 *
 * "restatement of llama.cpp math, not llama.cpp": no repacking, no llamafile sgemm, no NUMA placement; attention and the
 * norm / rope / softmax element-wise work (a few % of a step at this context) are left out of the timed step; all matrices
 * are taken in the one named format (Q4_K_M files keep some tensors in Q6_K).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#if defined(__x86_64__)
#include <immintrin.h>
#endif
#ifdef _OPENMP
#include <omp.h>
#endif

#define QK_K 256
#define QK0 32

typedef struct { uint16_t d, dmin; uint8_t scales[12]; uint8_t qs[QK_K / 2]; } block_q4_K;   /* 144 bytes */
typedef struct { float d; int8_t qs[QK_K]; int16_t bsums[QK_K / 16]; } block_q8_K;          /* 292 bytes */
typedef struct { uint16_t d; uint8_t qs[QK0 / 2]; } block_q4_0;                             /* 18 bytes */
typedef struct { uint16_t d; int8_t qs[QK0]; } block_q8_0;                                  /* 34 bytes */

int qgemv_sizeof_q4_K(void) { return (int)sizeof(block_q4_K); }
int qgemv_sizeof_q8_K(void) { return (int)sizeof(block_q8_K); }
int qgemv_sizeof_q4_0(void) { return (int)sizeof(block_q4_0); }
int qgemv_sizeof_q8_0(void) { return (int)sizeof(block_q8_0); }

/* ---- fp16 <-> fp32 (IEEE half, as ggml's GGML_FP16_TO_FP32) ---- */
static inline float h2f(uint16_t h) {
    const uint32_t s = (uint32_t)(h & 0x8000) << 16, e = (h >> 10) & 0x1f, m = h & 0x3ff;
    uint32_t u;
    if (e == 0) {
        if (m == 0) u = s;
        else {
            int k = 0;
            uint32_t mm = m;
            while (!(mm & 0x400)) mm <<= 1, ++k;
            u = s | ((uint32_t)(113 - k) << 23) | ((mm & 0x3ff) << 13);
        }
    } else if (e == 31) u = s | 0x7f800000u | (m << 13);
    else u = s | ((e + 112) << 23) | (m << 13);
    float f;
    memcpy(&f, &u, 4);
    return f;
}
static inline uint16_t f2h(float f) {   /* round to nearest even */
    uint32_t u;
    memcpy(&u, &f, 4);
    const uint32_t s = (u >> 16) & 0x8000;
    const int32_t e = (int32_t)((u >> 23) & 0xff) - 127 + 15;
    uint32_t m = u & 0x7fffff;
    if (((u >> 23) & 0xff) == 0xff) return (uint16_t)(s | 0x7c00 | (m ? 0x200 : 0));
    if (e >= 31) return (uint16_t)(s | 0x7c00);
    if (e <= 0) {
        if (e < -10) return (uint16_t)s;
        m |= 0x800000;
        const int sh = 14 - e;
        uint32_t r = m >> sh;
        const uint32_t rem = m & ((1u << sh) - 1), half = 1u << (sh - 1);
        if (rem > half || (rem == half && (r & 1))) ++r;
        return (uint16_t)(s | r);
    }
    uint32_t r = ((uint32_t)e << 10) | (m >> 13);
    const uint32_t rem = m & 0x1fff;
    if (rem > 0x1000 || (rem == 0x1000 && (r & 1))) ++r;
    return (uint16_t)(s | r);
}

static inline int nearest_int(float f) { return (int)lrintf(f); }

/* ---- activation quantisation (ggml quantize_row_q8_K_ref / quantize_row_q8_0_ref) ---- */
void qgemv_quantize_q8_K(const float *x, void *vy, int64_t k) {
    block_q8_K *y = (block_q8_K *)vy;
    for (int64_t i = 0; i < k / QK_K; ++i) {
        float max = 0.f, amax = 0.f;
        for (int j = 0; j < QK_K; ++j) {
            const float ax = fabsf(x[j]);
            if (ax > amax) amax = ax, max = x[j];
        }
        if (amax == 0.f) {
            memset(&y[i], 0, sizeof(block_q8_K));
            x += QK_K;
            continue;
        }
        const float iscale = -127.f / max;
        for (int j = 0; j < QK_K; ++j) {
            int v = nearest_int(iscale * x[j]);
            y[i].qs[j] = (int8_t)(v > 127 ? 127 : v);
        }
        for (int j = 0; j < QK_K / 16; ++j) {
            int sum = 0;
            for (int t = 0; t < 16; ++t) sum += y[i].qs[16 * j + t];
            y[i].bsums[j] = (int16_t)sum;
        }
        y[i].d = 1.f / iscale;
        x += QK_K;
    }
}

void qgemv_quantize_q8_0(const float *x, void *vy, int64_t k) {
    block_q8_0 *y = (block_q8_0 *)vy;
    for (int64_t i = 0; i < k / QK0; ++i) {
        float amax = 0.f;
        for (int j = 0; j < QK0; ++j) amax = fmaxf(amax, fabsf(x[i * QK0 + j]));
        const float d = amax / 127.f, id = d ? 1.f / d : 0.f;
        y[i].d = f2h(d);
        for (int j = 0; j < QK0; ++j) y[i].qs[j] = (int8_t)nearest_int(x[i * QK0 + j] * id);
    }
}

/* 6-bit (scale, min) of sub-block j (ggml get_scale_min_k4) */
static inline void scale_min_k4(int j, const uint8_t *q, uint8_t *d, uint8_t *m) {
    if (j < 4) {
        *d = q[j] & 63, *m = q[j + 4] & 63;
    } else {
        *d = (uint8_t)((q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4));
        *m = (uint8_t)((q[j + 4] >> 4) | ((q[j] >> 6) << 4));
    }
}

/* ---- scalar dot products: the definition (compiled twice: baseline x86-64 and auto-vectorised for AVX2, picked at load time,
 * so a host without AVX-512 VNNI still gets SIMD code) ---- */
#if defined(__x86_64__) && defined(__GNUC__)
#define QG_CLONES __attribute__((target_clones("avx2", "default")))
#else
#define QG_CLONES
#endif
QG_CLONES static float dot_q4_K_q8_K_scalar(int64_t n, const block_q4_K *x, const block_q8_K *y) {
    float sumf = 0.f;
    for (int64_t i = 0; i < n / QK_K; ++i) {
        const float d = h2f(x[i].d) * y[i].d, dmin = h2f(x[i].dmin) * y[i].d;
        int32_t sumi = 0, summ = 0;
        for (int j = 0; j < 8; ++j) {
            uint8_t sc, m;
            scale_min_k4(j, x[i].scales, &sc, &m);
            const uint8_t *q = x[i].qs + 32 * (j / 2);
            const int8_t *q8 = y[i].qs + 32 * j;
            int32_t s = 0;
            if (j & 1) for (int t = 0; t < 32; ++t) s += (q[t] >> 4) * q8[t];
            else       for (int t = 0; t < 32; ++t) s += (q[t] & 0xF) * q8[t];
            sumi += s * sc;
            summ += m * (y[i].bsums[2 * j] + y[i].bsums[2 * j + 1]);
        }
        sumf += d * (float)sumi - dmin * (float)summ;
    }
    return sumf;
}

QG_CLONES static float dot_q4_0_q8_0_scalar(int64_t n, const block_q4_0 *x, const block_q8_0 *y) {
    float sumf = 0.f;
    for (int64_t i = 0; i < n / QK0; ++i) {
        int32_t s = 0;
        for (int t = 0; t < 16; ++t) {
            s += ((x[i].qs[t] & 0xF) - 8) * y[i].qs[t];
            s += ((x[i].qs[t] >> 4) - 8) * y[i].qs[t + 16];
        }
        sumf += (float)s * h2f(x[i].d) * h2f(y[i].d);
    }
    return sumf;
}

/* ---- AVX-512 VNNI dot products (checked against the scalar ones in tests/test_qgemv.py) ---- */
#if defined(__x86_64__)
#define QG_TARGET __attribute__((target("avx512f,avx512bw,avx512vl,avx512dq,avx512vnni,avx2,fma")))

QG_TARGET static float dot_q4_K_q8_K_vnni(int64_t n, const block_q4_K *x, const block_q8_K *y) {
    static const uint32_t kmask1 = 0x3f3f3f3f, kmask2 = 0x0f0f0f0f, kmask3 = 0x03030303;
    const __m512i m4 = _mm512_set1_epi8(0xF);
    __m512 acc = _mm512_setzero_ps();
    float acc_m = 0.f;
    for (int64_t i = 0; i < n / QK_K; ++i) {
        const float d = h2f(x[i].d) * y[i].d, dmin = h2f(x[i].dmin) * y[i].d;
        uint32_t utmp[4];
        memcpy(utmp, x[i].scales, 12);   /* -> utmp[0..1] = 8 scales, utmp[2..3] = 8 mins, one byte each (ggml's unpacking) */
        utmp[3] = ((utmp[2] >> 4) & kmask2) | (((utmp[1] >> 6) & kmask3) << 4);
        const uint32_t uaux = utmp[1] & kmask1;
        utmp[1] = (utmp[2] & kmask2) | (((utmp[0] >> 6) & kmask3) << 4);
        utmp[2] = uaux;
        utmp[0] &= kmask1;
        const uint8_t *sc = (const uint8_t *)utmp, *mn = sc + 8;
        int32_t summ = 0;
        for (int j = 0; j < 8; ++j) summ += mn[j] * (y[i].bsums[2 * j] + y[i].bsums[2 * j + 1]);
        __m512i sumi = _mm512_setzero_si512();
        for (int g = 0; g < 4; ++g) {   /* 64 weights: low nibbles = sub-block 2g, high nibbles = sub-block 2g + 1 */
            const __m256i q = _mm256_loadu_si256((const __m256i *)(x[i].qs + 32 * g));
            __m512i q4 = _mm512_inserti64x4(_mm512_castsi256_si512(q), _mm256_srli_epi16(q, 4), 1);
            q4 = _mm512_and_si512(q4, m4);
            const __m512i q8 = _mm512_loadu_si512((const void *)(y[i].qs + 64 * g));
            const __m512i p = _mm512_dpbusd_epi32(_mm512_setzero_si512(), q4, q8);
            const __m512i scv = _mm512_inserti64x4(_mm512_set1_epi32(sc[2 * g]), _mm256_set1_epi32(sc[2 * g + 1]), 1);
            sumi = _mm512_add_epi32(sumi, _mm512_mullo_epi32(p, scv));
        }
        acc = _mm512_fmadd_ps(_mm512_set1_ps(d), _mm512_cvtepi32_ps(sumi), acc);
        acc_m += dmin * (float)summ;
    }
    return _mm512_reduce_add_ps(acc) - acc_m;
}

QG_TARGET static float dot_q4_0_q8_0_vnni(int64_t n, const block_q4_0 *x, const block_q8_0 *y) {
    const __m256i m4 = _mm256_set1_epi8(0xF), eight = _mm256_set1_epi8(8);
    __m256 acc = _mm256_setzero_ps();
    for (int64_t i = 0; i < n / QK0; ++i) {
        const __m128i q = _mm_loadu_si128((const __m128i *)x[i].qs);
        __m256i q4 = _mm256_inserti128_si256(_mm256_castsi128_si256(q), _mm_srli_epi16(q, 4), 1);   /* low nibbles | high nibbles */
        q4 = _mm256_and_si256(q4, m4);
        const __m256i q8 = _mm256_loadu_si256((const __m256i *)y[i].qs);
        __m256i p = _mm256_dpbusd_epi32(_mm256_setzero_si256(), q4, q8);
        p = _mm256_sub_epi32(p, _mm256_dpbusd_epi32(_mm256_setzero_si256(), eight, q8));            /* sum (q - 8) q8 */
        acc = _mm256_fmadd_ps(_mm256_set1_ps(h2f(x[i].d) * h2f(y[i].d)), _mm256_cvtepi32_ps(p), acc);
    }
    __m128 s = _mm_add_ps(_mm256_castps256_ps128(acc), _mm256_extractf128_ps(acc, 1));
    s = _mm_add_ps(s, _mm_movehl_ps(s, s));
    s = _mm_add_ss(s, _mm_movehdup_ps(s));
    return _mm_cvtss_f32(s);
}
#endif

void qgemv_set_threads(int n) {
#ifdef _OPENMP
    if (n >= 1) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
static int g_force_scalar = 0;
void qgemv_force_scalar(int on) { g_force_scalar = on; }
int qgemv_has_vnni(void) {
#if defined(__x86_64__)
    return __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512vl") &&
           __builtin_cpu_supports("avx512vnni");
#else
    return 0;
#endif
}

/* ---- y = W x: rows over threads, one quantised activation vector ---- */
void qgemv_q4_K(const void *w, const void *xq, float *y, int64_t n_rows, int64_t k) {
    const block_q4_K *W = (const block_q4_K *)w;
    const int64_t nb = k / QK_K;
    const int simd = qgemv_has_vnni() && !g_force_scalar;
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < n_rows; ++r) {
#if defined(__x86_64__)
        if (simd) { y[r] = dot_q4_K_q8_K_vnni(k, W + r * nb, (const block_q8_K *)xq); continue; }
#endif
        y[r] = dot_q4_K_q8_K_scalar(k, W + r * nb, (const block_q8_K *)xq);
    }
}
void qgemv_q4_0(const void *w, const void *xq, float *y, int64_t n_rows, int64_t k) {
    const block_q4_0 *W = (const block_q4_0 *)w;
    const int64_t nb = k / QK0;
    const int simd = qgemv_has_vnni() && !g_force_scalar;
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < n_rows; ++r) {
#if defined(__x86_64__)
        if (simd) { y[r] = dot_q4_0_q8_0_vnni(k, W + r * nb, (const block_q8_0 *)xq); continue; }
#endif
        y[r] = dot_q4_0_q8_0_scalar(k, W + r * nb, (const block_q8_0 *)xq);
    }
}

/* ---- one decode step of a Llama geometry: every weight matrix once, as a GEMV on quantised blocks ---- */
static void fill_blocks(uint8_t *p, int64_t n_blocks, int block_bytes, int n_half, uint64_t seed) {
    /* pseudo-random quants; the fp16 scale fields get small finite values so that sums stay finite */
    const uint16_t d = f2h(0.002f), dmin = f2h(0.001f);
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < n_blocks; ++b) {
        uint64_t s = seed + (uint64_t)b * 0x9E3779B97F4A7C15ull;
        uint8_t *q = p + b * block_bytes;
        for (int i = 0; i < block_bytes; i += 8) {
            s ^= s >> 12, s ^= s << 25, s ^= s >> 27;
            const uint64_t v = s * 0x2545F4914F6CDD1Dull;
            memcpy(q + i, &v, (size_t)(block_bytes - i < 8 ? block_bytes - i : 8));
        }
        memcpy(q, &d, 2);
        if (n_half == 2) memcpy(q + 2, &dmin, 2);
    }
}

typedef struct { uint8_t *w; int64_t n, k; } qmat;

/* fmt 0: Q4_K x Q8_K, fmt 1: Q4_0 x Q8_0.  Runs n_warm + n_steps decode steps (all GEMVs of the model: per layer qkv, o,
 * gate|up, down; then the lm_head; activations re-quantised before every matrix as llama.cpp does; SwiGLU between gate|up and
 * down) and returns the seconds the n_steps took, or a negative number when memory cannot be had.  *bytes_per_step gets
 * the quantised weight bytes one step reads. */
double qgemv_decode_bench(int fmt, int d, int n_layers, int n_q, int n_kv, int hd, int ffn, int vocab, int n_steps, int n_warm,
                          double *bytes_per_step, int *n_threads) {
    const int bb = fmt == 0 ? (int)sizeof(block_q4_K) : (int)sizeof(block_q4_0), bk = fmt == 0 ? QK_K : QK0;
    const int64_t shapes[5][2] = {{(int64_t)(n_q + 2 * n_kv) * hd, d}, {d, (int64_t)n_q * hd}, {2 * (int64_t)ffn, d}, {d, ffn}, {vocab, d}};
    for (int m = 0; m < 5; ++m)
        if (shapes[m][1] % bk) return -2.0;
    const int n_mats = 4 * n_layers + 1;
    qmat *mats = (qmat *)calloc((size_t)n_mats, sizeof(qmat));
    double total = 0;
    for (int i = 0; i < n_mats; ++i) {
        const int kind = i == n_mats - 1 ? 4 : i % 4;
        mats[i].n = shapes[kind][0], mats[i].k = shapes[kind][1];
        const int64_t nb = mats[i].n * (mats[i].k / bk);
        mats[i].w = (uint8_t *)malloc((size_t)nb * bb);
        if (!mats[i].w) return -1.0;
        fill_blocks(mats[i].w, nb, bb, fmt == 0 ? 2 : 1, 0xB200ull + (uint64_t)i);
        total += (double)nb * bb;
    }
    if (bytes_per_step) *bytes_per_step = total;
    int64_t maxdim = vocab;
    if (2 * (int64_t)ffn > maxdim) maxdim = 2 * (int64_t)ffn;
    float *x = (float *)malloc((size_t)maxdim * 4), *yv = (float *)malloc((size_t)maxdim * 4), *h = (float *)malloc((size_t)d * 4);
    void *xq = malloc((size_t)(maxdim / 32 + 8) * 40);
    for (int i = 0; i < d; ++i) h[i] = 0.01f * (float)((i * 37) % 101 - 50);
#ifdef _OPENMP
    if (n_threads) *n_threads = omp_get_max_threads();
#else
    if (n_threads) *n_threads = 1;
#endif
    struct timespec t0, t1;
    double secs = 0;
    for (int step = 0; step < n_warm + n_steps; ++step) {
        if (step == n_warm) clock_gettime(CLOCK_MONOTONIC, &t0);
        for (int i = 0; i < n_mats; ++i) {
            const qmat *m = &mats[i];
            const int kind = i == n_mats - 1 ? 4 : i % 4;
            const float *src = h;
            if (kind == 1) src = yv;                       /* o_proj reads the (stand-in) attention output: the q rows of qkv */
            if (kind == 3) {                               /* SwiGLU of the gate|up rows */
                for (int64_t t = 0; t < ffn; ++t) x[t] = (yv[t] / (1.f + expf(-yv[t]))) * yv[ffn + t];
                src = x;
            }
            if (fmt == 0) qgemv_quantize_q8_K(src, xq, m->k);
            else qgemv_quantize_q8_0(src, xq, m->k);
            if (fmt == 0) qgemv_q4_K(m->w, xq, yv, m->n, m->k);
            else qgemv_q4_0(m->w, xq, yv, m->n, m->k);
            if (kind == 1 || kind == 3)                    /* residual add, kept bounded */
                for (int t = 0; t < d; ++t) h[t] = 0.5f * h[t] + 1e-3f * yv[t];
        }
    }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    secs = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
    for (int i = 0; i < n_mats; ++i) free(mats[i].w);
    free(mats), free(x), free(yv), free(h), free(xq);
    return secs;
}
