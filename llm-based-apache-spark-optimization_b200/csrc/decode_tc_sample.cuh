// decode_tc_sample.cuh -- the sampler as the last phase of the tcgen05 persistent decode kernel (decode_tc.cu).
//
// The reference's six ollama.generate call sites pass no `options` (Flask/app.py:102-107,160-166;
// FastAPI/app.py:85-90,105-111; Model_Evaluation_&_Comparision.py:23-45,114-120), so Ollama's defaults
// apply: repeat penalty 1.1 over the last 64 tokens -> top-k 40 -> top-p 0.9 -> temperature 0.8 -> draw.
// Same chain and the same fp32 arithmetic order as sample_kernel (misc.cu) / oracle/sampler.py, so that
// a sampled decode call is ONE launch like a greedy one:
//   1. lm_head epilogue has stored this CTA's logit rows (whole 128-row tiles);
//   2. the CTA that owns the row of a recent token applies the repeat penalty to it in place;
//   3. every CTA extracts its k best rows (k <= kTSampK) as packed keys, value-descending then
//      id-ascending, and publishes them; the grid barrier that ends the phase publishes them;
//   4. every CTA merges the G sorted lists redundantly (k rounds of "largest list head"), so all CTAs
//      hold the same global top-k without a second barrier, and thread 0 of each runs the short serial
//      tail (temperature softmax, top-p prefix, draw) bit for bit like the stand-alone sampler.
// Steps 3 and 4 are two-level: every warp first takes the k best of ITS keys / lists with warp-wide maxima only (no block
// barrier inside the k rounds), then one warp merges the 12 warp lists the same way -- the top k of a union is the top
// k of the union of the parts' top k.  (One level with a block-wide maximum per round cost 34 us per sampled token, this
// costs about a quarter of it.)
// Scratch: the attention merge buffer and the staging buffer behind it (s.mrg, s.qst: contiguous), idle outside the attention phases.
#pragma once

constexpr int kTSampK = kTcdSampK; // largest top-k the in-kernel sampler takes (Ollama's default is 40)
constexpr int kTSampPerThread = 4; // logit rows of the CTA's lm_head share per compute thread (vocab <= 148 * 1536)

__device__ __forceinline__ uint32_t tcd_fkey(float f) {  // order-preserving float -> uint (as misc.cu)
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float tcd_unfkey(uint32_t k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
// larger key = larger logit, then lower token id; 0 = empty
__device__ __forceinline__ unsigned long long tcd_cand_key(float v, int id) {
    return ((unsigned long long)tcd_fkey(v) << 32) | (unsigned long long)(0x7fffffffu - (uint32_t)id);
}
__device__ __forceinline__ unsigned long long tcd_warp_max64(unsigned long long k) {
    const uint32_t hi = (uint32_t)(k >> 32);
    const uint32_t mh = __reduce_max_sync(0xffffffffu, hi);
    const uint32_t ml = __reduce_max_sync(0xffffffffu, hi == mh ? (uint32_t)k : 0u);
    return ((unsigned long long)mh << 32) | ml;
}
__device__ __forceinline__ uint64_t tcd_mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

struct TcdSampScratch {
    int *hist;                    // [64] the lane's recent tokens
    unsigned long long *top;      // [kTSampK] the global top-k keys, best first
    float *ev;                    // [kTSampK] exp((v_i - v_0) / T)
    unsigned long long *wl;       // [kTW][k] the k best keys of every warp's share, best first
};
__device__ __forceinline__ TcdSampScratch tcd_samp_scratch(const TcdSmem &s) {
    TcdSampScratch t;
    t.hist = reinterpret_cast<int *>(s.mrg);
    t.top = reinterpret_cast<unsigned long long *>(s.mrg + 64);
    t.ev = reinterpret_cast<float *>(t.top + kTSampK);
    t.wl = reinterpret_cast<unsigned long long *>(t.ev + kTSampK);
    return t;
}
// bytes of scratch for top-k `k`: hist + top + ev + kTW lists of k keys
__host__ __device__ __forceinline__ size_t tcd_samp_scratch_bytes(int k) { return 64 * 4 + kTSampK * 8 + kTSampK * 4 + (size_t)kTW * k * 8; }

// Warp 0: merge the kTW sorted warp lists in t.wl (k entries each, 0 = none) into out[0..k), best first.
__device__ __forceinline__ void tcd_merge_warp_lists(const TcdSampScratch &t, int k, int lane, unsigned long long *out) {
    int at = 0;
    unsigned long long head = lane < kTW ? t.wl[lane * k] : 0ull;
    for (int round = 0; round < k; ++round) {
        const unsigned long long best = tcd_warp_max64(head);
        if (best != 0ull && head == best) {   // keys are unique: exactly one lane advances
            ++at;
            head = at < k ? t.wl[lane * k + at] : 0ull;
        }
        if (lane == 0) out[round] = best;
    }
}

// Steps 2 and 3: repeat penalty on the rows this CTA owns, then its k best rows -> p.cand[bid][0..k).
// (b: the lane of a multi-lane launch; its logits, candidate lists and sampler settings are lane b's)
static __device__ void tcd_sample_local(const TcdParams &p, const TcdJob &job, const TcdSmem &s, int k, int G, int bid, int tid, int b) {
    const int warp = tid >> 5, lane = tid & 31;
    const TcdSampScratch t = tcd_samp_scratch(s);
    float *logits = p.logits + (size_t)b * p.vocab;
    const float repeat_penalty = p.repeat_penalty_l[b];
    const TRange r = tcd_range(job.N, job.K, true, G, bid);
    const int row0 = (r.g0 / r.nkb) * 128;
    const int row1 = min((r.g1 / r.nkb) * 128, p.vocab);
    const int id0 = p.tp_rank * p.vocab;     // global id of this rank's row 0 (vocab-parallel lm_head)
    bar_sync(1, kTC);                        // the epilogue warps have stored this CTA's logit rows
    if (repeat_penalty != 1.f) {
        if (tid < 64) t.hist[tid] = __ldcg(p.lane_hist + (size_t)p.lanes[b] * 64 + tid);
        bar_sync(1, kTC);
        if (tid < 64) {   // each distinct recent token once: logit > 0 ? logit / penalty : logit * penalty
            const int tk = t.hist[tid];
            const int row = tk - id0;
            bool skip = tk < 0 || tk >= p.vocab_global || row < row0 || row >= row1;
            for (int j = 0; j < tid && !skip; ++j) skip = t.hist[j] == tk;
            if (!skip) {
                const float v = __ldcg(logits + row);
                logits[row] = v > 0.f ? v / repeat_penalty : v * repeat_penalty;
            }
        }
        bar_sync(1, kTC);
    }
    unsigned long long key[kTSampPerThread];
#pragma unroll
    for (int m = 0; m < kTSampPerThread; ++m) {
        const int n = row0 + tid + m * kTC;
        key[m] = n < row1 ? tcd_cand_key(__ldcg(logits + n), id0 + n) : 0ull;
    }
    // level 1: every warp's k best of its own 128 keys, no block barrier inside the rounds
    for (int round = 0; round < k; ++round) {
        unsigned long long mx = key[0];
#pragma unroll
        for (int m = 1; m < kTSampPerThread; ++m) mx = key[m] > mx ? key[m] : mx;
        const unsigned long long best = tcd_warp_max64(mx);
#pragma unroll
        for (int m = 0; m < kTSampPerThread; ++m) key[m] = key[m] == best ? 0ull : key[m];   // keys are unique (ids are); 0 stays 0
        if (lane == 0) t.wl[warp * k + round] = best;
    }
    bar_sync(1, kTC);
    // level 2: warp 0 merges the 12 lists into this CTA's list (published by thread 0's release in the grid barrier that follows)
    if (warp == 0) tcd_merge_warp_lists(t, k, lane, p.cand + ((size_t)b * G + bid) * kTSampK);
    bar_sync(1, kTC);   // the lists are rewritten by the next lane / the merge phase
}

// Step 4, after the grid barrier: merge the G lists, then the serial tail.  Returns the token (the same in every CTA).
static __device__ int tcd_sample_merge(const TcdParams &p, const TcdSmem &s, int pos, int k, int G, int tid, int b) {
    const int warp = tid >> 5, lane = tid & 31;
    const TcdSampScratch t = tcd_samp_scratch(s);
    const float temperature = p.temperature_l[b], top_p = p.top_p_l[b];
    // level 1: warp w merges the CTA lists w, w + 12, ... (lane l holds the head of list w + 12 l) into its warp list
    const int li = warp + kTW * lane;
    const unsigned long long *list = p.cand + ((size_t)b * G + (li < G ? li : 0)) * kTSampK;
    unsigned long long h0 = 0, h1 = 0, h2 = 0, h3 = 0;
    int nxt = 0, used = 0;
    auto refill = [&]() {   // the next four entries of this lane's list (entries past k are stale)
        const ulonglong2 a = __ldcg(reinterpret_cast<const ulonglong2 *>(list + nxt));
        const ulonglong2 c2 = __ldcg(reinterpret_cast<const ulonglong2 *>(list + nxt + 2));
        h0 = nxt + 0 < k ? a.x : 0ull, h1 = nxt + 1 < k ? a.y : 0ull;
        h2 = nxt + 2 < k ? c2.x : 0ull, h3 = nxt + 3 < k ? c2.y : 0ull;
        nxt += 4, used = 0;
    };
    if (li < G) refill();
    for (int round = 0; round < k; ++round) {
        const unsigned long long best = tcd_warp_max64(h0);
        if (best != 0ull && h0 == best) {   // this lane's list supplied the winner: advance it
            h0 = h1, h1 = h2, h2 = h3, h3 = 0ull;
            if (++used == 4 && nxt < k) refill();
        }
        if (lane == 0) t.wl[warp * k + round] = best;
    }
    bar_sync(1, kTC);
    // level 2: warp 0 merges the 12 warp lists into the global top-k
    if (warp == 0) tcd_merge_warp_lists(t, k, lane, t.top);
    bar_sync(1, kTC);
    int kk = 0;     // candidates that exist (the vocabulary may hold fewer than k rows)
    for (int i = 0; i < k; ++i) kk += t.top[i] != 0ull ? 1 : 0;
    const float v0 = tcd_unfkey((uint32_t)(t.top[0] >> 32));
    int tok = (int)(0x7fffffffu - (uint32_t)t.top[0]);
    if (temperature > 0.f) {
        const float inv_t = 1.f / temperature;
        if (tid < kk) t.ev[tid] = expf((tcd_unfkey((uint32_t)(t.top[tid] >> 32)) - v0) * inv_t);
        bar_sync(1, kTC);
        if (tid == 0) {   // sequential fp32 sums, the order of sample_kernel / oracle/sampler.py
            float total = 0.f;
            for (int i = 0; i < kk; ++i) total += t.ev[i];
            int keep = kk;
            if (top_p < 1.f) {
                float cum = 0.f;
                for (int i = 0; i < kk; ++i) {
                    cum += t.ev[i];
                    if (cum / total >= top_p) {
                        keep = i + 1;
                        break;
                    }
                }
            }
            float mass = 0.f;
            for (int i = 0; i < keep; ++i) mass += t.ev[i];
            const uint64_t z = tcd_mix64(p.seed_l[b] + ((uint64_t)pos + 1ULL) * 0x9E3779B97F4A7C15ULL);
            const float u = (float)(z >> 40) * (1.f / 16777216.f);
            const float rr = u * mass;
            float cum = 0.f;
            int pick = keep - 1;
            for (int i = 0; i < keep; ++i) {
                cum += t.ev[i];
                if (cum > rr) {
                    pick = i;
                    break;
                }
            }
            s.misc[56] = (int)(0x7fffffffu - (uint32_t)t.top[pick]);
        }
        bar_sync(1, kTC);
        tok = s.misc[56];
    }
    bar_sync(1, kTC);   // the scratch (attention merge buffer) and misc[56] are free again
    return tok;
}
