// mega_kernel.cuh -- the persistent decode kernel itself and its launcher (included by
// decode_mega.cu after the phase functions).
#pragma once

template <int HD, int GQ>
__global__ void __launch_bounds__(kMC + 32, 1) mega_decode_kernel(const __grid_constant__ MegaParams p) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const MegaSmem s = carve(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int G = gridDim.x, bid = blockIdx.x;

    if (tid == 0) {
        for (int i = 0; i < kSlots; ++i) {
            mbar_init(&s.full[i], 1);
            mbar_init(&s.empty[i], kMW);
        }
        for (int i = 0; i < kAttW; ++i) mbar_init(&s.att_full[i], 1);
        s.misc[2] = 0;
        mbar_fence_init();
    }
    {   // phase descriptors and page table -> shared memory (read once)
        const int nj = 4 * p.n_layers + 1;
        const uint32_t *src = reinterpret_cast<const uint32_t *>(p.jobs);
        uint32_t *dst = reinterpret_cast<uint32_t *>(s.jobs);
        for (int i = tid; i < nj * (int)(sizeof(MegaJob) / 4); i += kMC + 32) dst[i] = src[i];
        for (int i = tid; i < min(p.n_ptab, kMaxPtab); i += kMC + 32) s.ptab[i] = p.page_table[i];
    }
    __syncthreads();

    if (warp == kMW) {  // producer warp: streams weights for every phase of every step
        mega_producer(p, s, G, bid, lane);  // whole warp: lanes 0..7 each issue one row copy per stage
        return;
    }

    StepCtx c;
    c.gs = 0;
    c.seq = 0;
    c.tok = __ldcg(p.lane_token + p.lane);
    c.pos = __ldcg(p.lane_pos + p.lane);
    unsigned bar_no = 0;
    const int n_jobs = 4 * p.n_layers + 1;

    for (int step = 0; step < p.n_steps; ++step) {
        c.ti = (p.trace != nullptr && step == p.trace_step && (bid == 0 || bid == G - 1)) ? 0 : -1;
        mega_trace(p, c.ti, bid, tid);
        c.tok = min(max(c.tok, 0), p.vocab - 1);
        if (bid == 0 && tid == 0) p.lane_hist[(size_t)p.lane * 64 + (c.pos & 63)] = c.tok;
        if (tid < HD / 2) {  // RoPE angles of this position, used by the qkv epilogues of every layer
            float sn, cs;
            sincosf((float)c.pos * p.inv_freq[tid], &sn, &cs);
            s.rope[2 * tid] = cs, s.rope[2 * tid + 1] = sn;
        }
        for (int l = 0; l < p.n_layers; ++l) {
            const MegaJob *jb = s.jobs + 4 * l;
            mega_attn_prefetch<HD>(p, jb[0], s, c, G, bid, tid);
            mega_gemv_phase(p, jb[0], s, c, G, bid, tid);  // rmsnorm + qkv + rope + kv append
            mega_trace(p, c.ti, bid, tid);
            grid_sync(p.bar_ctr, ++bar_no * G, tid);
            mega_trace(p, c.ti, bid, tid);
            mega_attn_phase<HD, GQ>(p, jb[0], s, c, G, bid, tid);
            mega_trace(p, c.ti, bid, tid);
            grid_sync(p.bar_ctr, ++bar_no * G, tid);
            mega_trace(p, c.ti, bid, tid);
            mega_gemv_phase(p, jb[1], s, c, G, bid, tid);  // o_proj + residual
            mega_trace(p, c.ti, bid, tid);
            grid_sync(p.bar_ctr, ++bar_no * G, tid);
            mega_trace(p, c.ti, bid, tid);
            mega_gemv_phase(p, jb[2], s, c, G, bid, tid);  // rmsnorm + gate/up + swiglu
            mega_trace(p, c.ti, bid, tid);
            grid_sync(p.bar_ctr, ++bar_no * G, tid);
            mega_trace(p, c.ti, bid, tid);
            mega_gemv_phase(p, jb[3], s, c, G, bid, tid);  // down + residual
            mega_trace(p, c.ti, bid, tid);
            grid_sync(p.bar_ctr, ++bar_no * G, tid);
            mega_trace(p, c.ti, bid, tid);
        }
        // ---- final norm + lm_head: logits, per-CTA (max, lowest argmax) ----
        c.best = -INFINITY, c.best_i = 0x7fffffff;
        mega_gemv_phase(p, s.jobs[n_jobs - 1], s, c, G, bid, tid);
        {
            float m = c.best;
            int mi = c.best_i;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float om = __shfl_xor_sync(0xffffffffu, m, o);
                const int oi = __shfl_xor_sync(0xffffffffu, mi, o);
                if (om > m || (om == m && oi < mi)) m = om, mi = oi;
            }
            bar_sync(1, kMC);
            if (lane == 0) s.red[warp] = m, s.misc[4 + warp] = mi;
            bar_sync(1, kMC);
            if (tid == 0) {
                for (int w = 1; w < kMW; ++w)
                    if (s.red[w] > m || (s.red[w] == m && s.misc[4 + w] < mi)) m = s.red[w], mi = s.misc[4 + w];
                p.part_max[bid] = m;
                p.part_idx[bid] = mi;
            }
        }
        grid_sync(p.bar_ctr, ++bar_no * G, tid);
        if (!p.sample) break;  // caller samples from the logits (temperature > 0)
        // ---- greedy sampling: every CTA reduces the G partials redundantly (no extra barrier) ----
        {
            float m = -INFINITY;
            int mi = 0x7fffffff;
            for (int i = tid; i < G; i += kMC) {
                const float v = __ldcg(p.part_max + i);
                const int vi = __ldcg(p.part_idx + i);
                if (v > m || (v == m && vi < mi)) m = v, mi = vi;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float om = __shfl_xor_sync(0xffffffffu, m, o);
                const int oi = __shfl_xor_sync(0xffffffffu, mi, o);
                if (om > m || (om == m && oi < mi)) m = om, mi = oi;
            }
            if (lane == 0) s.red[32 + warp] = m, s.misc[24 + warp] = mi;
            bar_sync(1, kMC);
            for (int w = 0; w < kMW; ++w) {
                const float om = s.red[32 + w];
                const int oi = s.misc[24 + w];
                if (om > m || (om == m && oi < mi)) m = om, mi = oi;
            }
            c.tok = mi;
            c.pos += 1;
            if (bid == 0 && tid == 0) {
                p.lane_token[p.lane] = c.tok;
                p.lane_pos[p.lane] = c.pos;
                p.gen_out[step] = c.tok;
            }
            bar_sync(1, kMC);  // red/misc are reused by the next step
        }
    }
}

template <int HD, int GQ>
static int mega_launch_t(const MegaParams &p, cudaStream_t stream) {
    const size_t smem = mega_smem_bytes(HD);
    static bool attr_set = false;
    if (!attr_set) {
        B200_CUDA_OK(cudaFuncSetAttribute(mega_decode_kernel<HD, GQ>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)smem));
        attr_set = true;
    }
    int dev = 0, sms = 0;
    B200_CUDA_OK(cudaGetDevice(&dev));
    B200_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    B200_CUDA_OK(cudaMemsetAsync(p.bar_ctr, 0, sizeof(unsigned), stream));
    // cooperative launch: all CTAs must be co-resident, the grid barrier depends on it
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(sms);
    cfg.blockDim = dim3(kMC + 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    B200_CUDA_OK(cudaLaunchKernelEx(&cfg, mega_decode_kernel<HD, GQ>, p));
    return B200LLM_OK;
}
