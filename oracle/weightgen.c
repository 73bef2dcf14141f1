/*
 * oracle/weightgen.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the seeded random-init weight stream that both the oracle
 * and the B200 engine consume ("same random-init weights", BASELINE.json
 * north_star; SURVEY.md section 8(d) "Weights").  The reference repo has no
 * weights of its own (they live in Ollama's GGUF blobs),
 * so the stream is defined here, integer-only up to one fp32 multiply and one
 * fp32 add, so that it is bit-identical on any CPU and on the GPU
 * (product twin: csrc/weightgen.cu, checked bit-for-bit in tests/).
 *
 *   z   = splitmix64 finaliser of (seed + (idx+1) * 0x9E3779B97F4A7C15)
 *   S   = sum of the four 16-bit lanes of z          (Irwin-Hall, n = 4)
 *   v   = (float)(S - 131070) * scale                (scale = sigma / 37837.2277)
 *   out = bf16_rne(mean + v)
 *
 * Build: see oracle/Makefile (gcc -O3 -fopenmp -ffp-contract=off).
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>

static inline uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

static inline uint16_t f32_to_bf16_rne(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    u += 0x7FFFu + ((u >> 16) & 1u); /* finite inputs only */
    return (uint16_t)(u >> 16);
}

/* Fills out[0..n) with elements [first, first+n) of stream `seed`. */
void oracle_weightgen_bf16(uint16_t *out, uint64_t n, uint64_t first, uint64_t seed, float scale,
                           float mean) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        uint64_t z = mix64(seed + ((uint64_t)i + first + 1ULL) * 0x9E3779B97F4A7C15ULL);
        int32_t s = (int32_t)(z & 0xFFFF) + (int32_t)((z >> 16) & 0xFFFF) +
                    (int32_t)((z >> 32) & 0xFFFF) + (int32_t)(z >> 48);
        float v = (float)(s - 131070) * scale;
        float w = mean + v;
        out[i] = f32_to_bf16_rne(w);
    }
}

/* Uniform synthetic token ids in [0, modulo): SURVEY.md 8(d) prompts. */
void oracle_tokengen_i32(int32_t *out, uint64_t n, uint64_t seed, uint32_t modulo) {
    for (uint64_t i = 0; i < n; ++i) {
        uint64_t z = mix64(seed + (i + 1ULL) * 0x9E3779B97F4A7C15ULL);
        out[i] = (int32_t)((z >> 11) % modulo);
    }
}
