// decode_tc_kernel.cuh -- the tcgen05 persistent decode kernel and its launcher (included by decode_tc.cu).
#pragma once

// NB: lanes the instance decodes together -- 1 (every pointer of TcdParams is lane 0's) or kTcdMaxLanes (p.nb = 2..4 lanes:
// operand rows 2b / 2b + 1 and accumulator columns 2b / 2b + 1 belong to lane b, so the weights stream once for all of them)
template <int HD, int GQ, int NB>
__global__ void __launch_bounds__(kTThreads, 1) tcd_decode_kernel(const __grid_constant__ TcdParams p) {
    // the swizzled operand tiles need 1024-byte alignment; declared, not computed: with the base a compile-time shared-memory
    // address the regions below cost no register at all (a base aligned at run time was spilled and reloaded once per phase)
    extern __shared__ __align__(1024) uint8_t smem_dyn[];
    uint8_t *smem_raw = smem_dyn;
    const int n_jobs = 4 * p.n_layers + 1;
    const TcdSmem s = tcd_carve(smem_raw, HD, GQ, n_jobs, p.n_bslots, NB);
    const int nb = NB == 1 ? 1 : p.nb;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int G = gridDim.x, bid = blockIdx.x;

    if (tid == 0) {
        for (int i = 0; i < kTMaxStages; ++i) {
            mbar_init(&s.full[i], 1);
            mbar_init(&s.empty[i], 1);
        }
        for (int i = 0; i < kTAcc; ++i) {
            mbar_init(&s.acc_full[i], kTIssuers);
            mbar_init(&s.acc_empty[i], 4);
        }
        for (int i = 0; i < kTSlots; ++i) mbar_init(&s.tempty[i], 1);
        mbar_init(s.xrdy, kTW);
        mbar_fence_init();
    }
    if (warp == kTW) {  // TMEM, all of it: kTAcc accumulators of kTAccCols columns, then kTSlots weight slots of kTSlotCols
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s.tmem_slot)),
                     "r"((uint32_t)kTTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    {   // phase descriptors and page table -> shared memory (read once); operand blocks start as zero
        const uint32_t *src = reinterpret_cast<const uint32_t *>(p.jobs);
        uint32_t *dst = reinterpret_cast<uint32_t *>(s.jobs);
        for (int i = tid; i < n_jobs * (int)(sizeof(TcdJob) / 4); i += kTThreads) dst[i] = src[i];
        if constexpr (NB == 1) {
            for (int i = tid; i < min(p.n_ptab, kTMaxPtab); i += kTThreads) s.ptab[i] = p.page_table[i];
        } else {
            for (int b = 0; b < nb; ++b)
                for (int i = tid; i < min(p.n_ptab, kTMaxPtab); i += kTThreads) s.ptab[b * kTMaxPtab + i] = p.page_tables[b][i];
        }
        for (int i = tid; i < p.n_bslots * 1024 / 16; i += kTThreads) reinterpret_cast<uint4 *>(s.bop)[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *s.tmem_slot;

    if (warp == kTW + kTIssuers) {
        if (lane == 0) tcd_producer(p, s, G, bid);
    } else if (warp >= kTW) {
        tcd_mma_warp(p, s, G, bid, lane, tmem, warp - kTW);
    } else {
        TcdCtxT<NB> c;
        c.as = 0;
        if (p.tp_size > 1 && tid == 0) s.misc[62] = (int)__ldcg(p.ar_seq), s.misc[63] = (int)__ldcg(p.ar_seq + 2);
        if (p.tp_size > 1) bar_sync(1, kTC);
        if constexpr (NB == 1) {
            c.tok = __ldcg(p.lane_token + p.lane);
            c.pos = __ldcg(p.lane_pos + p.lane);
        } else {
            c.tok = c.pos = 0;
            if (tid < nb) s.misc[16 + tid] = __ldcg(p.lane_pos + p.lanes[tid]), s.misc[20 + tid] = __ldcg(p.lane_token + p.lanes[tid]);
            bar_sync(1, kTC);
        }
        unsigned bar_no = 0;
        volatile unsigned *bar_done = reinterpret_cast<volatile unsigned *>(s.misc + 60);   // last barrier seen complete
        if (tid == 0) *bar_done = 0u;
        // one weight-streaming phase, compute-warp side
        auto phase = [&](int j) {
            const TcdJob &job = s.jobs[j];
            const TRange r = tcd_range(job.N, job.K, tcd_whole(p, j, n_jobs), G, bid);
            const TcdScale<NB> sc = tcd_prologue<NB>(p, job, s, c, r, G, bid, tid);
            tcd_trace(p, c.ti, bid, tid, TT_XREADY);
            if (warp < 4) tcd_epilogue<NB>(p, job, s, c, r, sc, warp, lane, tmem);
            tcd_trace(p, c.ti, bid, tid, TT_EPI);
        };
        for (int step = 0; step < p.n_steps; ++step) {
#ifdef B200LLM_TCD_TRACE
            c.ti = (p.trace != nullptr && step == p.trace_step && (bid == 0 || bid == G - 1)) ? 0 : -1;
#else
            c.ti = -1;
#endif
            tcd_trace(p, c.ti, bid, tid, TT_STEP);
            if constexpr (NB == 1) {
                c.tok = min(max(c.tok, 0), p.vocab_global - 1);
                if (bid == 0 && tid == 0) p.lane_hist[(size_t)p.lane * 64 + (c.pos & 63)] = c.tok;
                if (tid < HD / 2) {  // RoPE angles of this position
                    float sn, cs;
                    sincosf((float)c.pos * p.inv_freq[tid], &sn, &cs);
                    s.rope[2 * tid] = cs, s.rope[2 * tid + 1] = sn;
                }
            } else {
                if (tid < nb) {
                    const int tk = min(max(s.misc[20 + tid], 0), p.vocab_global - 1);
                    s.misc[20 + tid] = tk;
                    if (bid == 0) p.lane_hist[(size_t)p.lanes[tid] * 64 + (s.misc[16 + tid] & 63)] = tk;
                }
                for (int i = tid; i < nb * (HD / 2); i += kTC) {   // RoPE angles of every lane's position
                    const int b = i / (HD / 2), j = i % (HD / 2);
                    float sn, cs;
                    sincosf((float)s.misc[16 + b] * p.inv_freq[j], &sn, &cs);
                    s.rope[b * 128 + 2 * j] = cs, s.rope[b * 128 + 2 * j + 1] = sn;
                }
                bar_sync(1, kTC);   // (the first prologue reads the tokens)
            }
            for (int l = 0; l < p.n_layers; ++l) {
                // CTA-uniform loop state is parked in shared memory across the attention phase (the register peak of the
                // kernel): left to ptxas it goes to local memory and comes back at L2 latency behind the grid barriers
                // (7 % of all stall samples sat on ONE such reload, profiles/tcd_full_r1e_summary.txt)
                if (tid == 0) s.misc[57] = step, s.misc[58] = l, s.misc[59] = (int)bar_no + ((p.dataflow & 4) ? 0 : 1);
                const TcdJob &qkv_job = s.jobs[4 * l];
#ifdef B200LLM_TCD_TRACE
                int arr_i = (p.trace != nullptr && step == p.trace_step && l == 1 && tid == 0) ? 0 : -1;
#else
                constexpr int arr_i = -1;
#endif
                auto arrive_stamp = [&]() {  // (debug) when does each CTA reach the barriers of layer 1?
#ifdef B200LLM_TCD_TRACE
                    if (arr_i >= 0) {
                        unsigned long long t;
                        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                        p.trace[2048 + arr_i * G + bid] = t;
                        ++arr_i;
                    }
#endif
                };
                tcd_attn_prefetch<HD, NB>(p, qkv_job, s, c, G, bid, tid);
                phase(4 * l);      // rmsnorm + qkv (raw sums)
                arrive_stamp();
                if (!(p.dataflow & 4)) {
                    tcd_grid_arrive(p.bar_ctr, tid);
                    tcd_trace(p, c.ti, bid, tid, TT_BAR_ARRIVED);
                    tcd_grid_wait(p.bar_ctr, ++bar_no * G, tid, bar_done, p.n_poll);
                } else {   // data-flow mode: only the CTAs with an attention item wait, for their head's rows
                    tcd_attn_wait_qkv<HD, GQ, NB>(p, qkv_job, s, c, G, bid, tid);
                }
                tcd_trace(p, c.ti, bid, tid, TT_BAR);
                tcd_attn_phase<HD, GQ, NB>(p, qkv_job, s, c, G, bid, tid);  // rope + kv append + attention
                step = reinterpret_cast<volatile int *>(s.misc)[57];   // written before the barrier above
                l = reinterpret_cast<volatile int *>(s.misc)[58];
                bar_no = (unsigned)reinterpret_cast<volatile int *>(s.misc)[59];
                tcd_trace(p, c.ti, bid, tid, TT_ATT_DONE);
                arrive_stamp();
                if (!(p.dataflow & 2)) {   // (data-flow mode: the o_proj prologue polls the (value, tag) pairs it needs)
                    tcd_grid_arrive(p.bar_ctr, tid);
                    tcd_trace(p, c.ti, bid, tid, TT_BAR_ARRIVED);
                    tcd_grid_wait(p.bar_ctr, ++bar_no * G, tid, bar_done, p.n_poll);
                    tcd_trace(p, c.ti, bid, tid, TT_BAR);
                }
                phase(4 * l + 1);  // o_proj + residual
                arrive_stamp();
                if (!(p.tp_size > 1 && (p.dataflow & 8))) {   // (tensor parallel, data-flow mode: the slice owners poll the sums' counts)
                    tcd_grid_arrive(p.bar_ctr, tid);
                    tcd_trace(p, c.ti, bid, tid, TT_BAR_ARRIVED);
                    tcd_grid_wait(p.bar_ctr, ++bar_no * G, tid, bar_done, p.n_poll);
                    tcd_trace(p, c.ti, bid, tid, TT_BAR);
                }
                if (p.tp_size > 1) {   // o_proj is row-parallel: h += sum over ranks of the partial sums
                    tcd_allreduce_phase(p, s, G, bid, tid, s.jobs[4 * l + 1].N, s.jobs[4 * l + 1].K);
                    tcd_grid_sync(p.bar_ctr, ++bar_no * G, tid, bar_done, p.n_poll);
                }
                phase(4 * l + 2);  // rmsnorm + gate/up (raw sums)
                arrive_stamp();
                if (!(p.dataflow & 1)) {   // (data-flow mode: the down_proj prologue polls the (sum, count) pairs it needs)
                    tcd_grid_arrive(p.bar_ctr, tid);
                    tcd_trace(p, c.ti, bid, tid, TT_BAR_ARRIVED);
                    tcd_grid_wait(p.bar_ctr, ++bar_no * G, tid, bar_done, p.n_poll);
                    tcd_trace(p, c.ti, bid, tid, TT_BAR);
                }
                phase(4 * l + 3);  // swiglu + down + residual
                arrive_stamp();
                if (!(p.tp_size > 1 && (p.dataflow & 8))) {
                    tcd_grid_arrive(p.bar_ctr, tid);
                    tcd_trace(p, c.ti, bid, tid, TT_BAR_ARRIVED);
                    tcd_grid_wait(p.bar_ctr, ++bar_no * G, tid, bar_done, p.n_poll);
                    tcd_trace(p, c.ti, bid, tid, TT_BAR);
                }
                if (p.tp_size > 1) {   // down_proj is row-parallel too
                    tcd_allreduce_phase(p, s, G, bid, tid, s.jobs[4 * l + 3].N, s.jobs[4 * l + 3].K);
                    tcd_grid_sync(p.bar_ctr, ++bar_no * G, tid, bar_done, p.n_poll);
                }
            }
            // ---- final norm + lm_head: logits, per-CTA (max, lowest argmax) ----
#pragma unroll
            for (int b = 0; b < NB; ++b) c.best[b] = -INFINITY, c.best_i[b] = 0x7fffffff;
            phase(n_jobs - 1);
            if (p.sample == 2) {
                // ---- sampler chain in the kernel: penalty + per-CTA top-k, barrier, redundant merge + draw (lane after lane) ----
                for (int b = 0; b < nb; ++b) {
                    const int k = p.temperature_l[b] > 0.f ? p.top_k_l[b] : 1;
                    tcd_sample_local(p, s.jobs[n_jobs - 1], s, k, G, bid, tid, b);
                }
                tcd_grid_arrive(p.bar_ctr, tid);
                tcd_trace(p, c.ti, bid, tid, TT_BAR_ARRIVED);
                tcd_grid_wait(p.bar_ctr, ++bar_no * G, tid, bar_done, p.n_poll);
                tcd_trace(p, c.ti, bid, tid, TT_BAR);
                if constexpr (NB == 1) {
                    c.tok = tcd_sample_merge(p, s, c.pos, p.temperature_l[0] > 0.f ? p.top_k_l[0] : 1, G, tid, 0);
                    c.pos += 1;
                    if (bid == 0 && tid == 0) {
                        p.lane_token[p.lane] = c.tok;
                        p.lane_pos[p.lane] = c.pos;
                        p.gen_out[step] = c.tok;
                    }
                } else {
                    for (int b = 0; b < nb; ++b) {
                        const int pos = s.misc[16 + b];
                        const int tok = tcd_sample_merge(p, s, pos, p.temperature_l[b] > 0.f ? p.top_k_l[b] : 1, G, tid, b);
                        if (tid == 0) {
                            s.misc[16 + b] = pos + 1, s.misc[20 + b] = tok;
                            if (bid == 0) p.lane_token[p.lanes[b]] = tok, p.lane_pos[p.lanes[b]] = pos + 1, p.gen_out[step * nb + b] = tok;
                        }
                    }
                    bar_sync(1, kTC);
                }
                continue;
            }
            for (int b = 0; b < nb; ++b) {
                // (static indices only: an array indexed by b would live in local memory and every epilogue would load and store it)
                float m = c.best[0];
                int mi = c.best_i[0];
#pragma unroll
                for (int l = 1; l < NB; ++l)
                    if (l == b) m = c.best[l], mi = c.best_i[l];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const float om = __shfl_xor_sync(0xffffffffu, m, o);
                    const int oi = __shfl_xor_sync(0xffffffffu, mi, o);
                    if (om > m || (om == m && oi < mi)) m = om, mi = oi;
                }
                bar_sync(1, kTC);
                if (lane == 0) s.red[warp] = m, s.misc[4 + warp] = mi;
                bar_sync(1, kTC);
                if (tid == 0) {
                    for (int w = 1; w < kTW; ++w)
                        if (s.red[w] > m || (s.red[w] == m && s.misc[4 + w] < mi)) m = s.red[w], mi = s.misc[4 + w];
                    p.part_max[b * G + bid] = m;
                    p.part_idx[b * G + bid] = mi;
                }
            }
            tcd_grid_sync(p.bar_ctr, ++bar_no * G, tid, bar_done, p.n_poll);
            tcd_trace(p, c.ti, bid, tid, TT_BAR);
            if (!p.sample) break;  // caller samples from the logits (temperature > 0)
            // ---- greedy sampling: every CTA reduces the G partials redundantly (no extra barrier) ----
            for (int b = 0; b < nb; ++b) {
                float m = -INFINITY;
                int mi = 0x7fffffff;
                for (int i = tid; i < G; i += kTC) {
                    const float v = __ldcg(p.part_max + b * G + i);
                    const int vi = __ldcg(p.part_idx + b * G + i);
                    if (v > m || (v == m && vi < mi)) m = v, mi = vi;
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const float om = __shfl_xor_sync(0xffffffffu, m, o);
                    const int oi = __shfl_xor_sync(0xffffffffu, mi, o);
                    if (om > m || (om == m && oi < mi)) m = om, mi = oi;
                }
                if (lane == 0) s.red[32 + warp] = m, s.misc[24 + warp] = mi;
                bar_sync(1, kTC);
                for (int w = 0; w < kTW; ++w) {
                    const float om = s.red[32 + w];
                    const int oi = s.misc[24 + w];
                    if (om > m || (om == m && oi < mi)) m = om, mi = oi;
                }
                if (p.tp_size > 1) {   // vocab-parallel lm_head: this rank's winner against the peers'
                    bar_sync(1, kTC);  // all threads have read red[32..] / misc[24..]
                    mi = tcd_token_exchange(p, s, m, mi + p.tp_rank * p.vocab, bid, tid);
                }
                if constexpr (NB == 1) {
                    c.tok = mi;
                    c.pos += 1;
                    if (bid == 0 && tid == 0) {
                        p.lane_token[p.lane] = c.tok;
                        p.lane_pos[p.lane] = c.pos;
                        p.gen_out[step] = c.tok;
                    }
                } else if (tid == 0) {
                    const int pos = s.misc[16 + b] + 1;
                    s.misc[16 + b] = pos, s.misc[20 + b] = mi;
                    if (bid == 0) p.lane_token[p.lanes[b]] = mi, p.lane_pos[p.lanes[b]] = pos, p.gen_out[step * nb + b] = mi;
                }
                bar_sync(1, kTC);  // red/misc are reused by the next lane / step
            }
        }
        if (p.tp_size > 1 && bid == 0 && tid == 0) {   // sequence numbers for the next launch (and for tp_ar.cu's kernel)
            *reinterpret_cast<volatile unsigned *>(p.ar_seq) = (unsigned)s.misc[62];
            *reinterpret_cast<volatile unsigned *>(p.ar_seq + 2) = (unsigned)s.misc[63];
        }
    }
    // every role has finished its schedule: all MMAs are committed and read back
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == kTW)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)kTTmemCols)
                     : "memory");
}

template <int HD, int GQ, int NB>
int tcd_launch_t(const TcdParams &p, cudaStream_t stream) {
    B200_REQUIRE(p.n_stages >= 4 && p.n_stages <= kTMaxStages && p.n_stages % kTIssuers == 0 && p.n_bslots >= 1 && p.n_bslots <= kTBSlots,
                 "tcd: bad ring plan (%d stages, %d operand blocks)", p.n_stages, p.n_bslots);
    const size_t smem = tcd_smem_bytes(HD, GQ, 4 * p.n_layers + 1, p.n_stages, p.n_bslots, NB);
    B200_REQUIRE(smem <= kTSmemLimit, "tcd: %zu bytes of shared memory needed", smem);
    B200_CUDA_OK(cudaFuncSetAttribute(tcd_decode_kernel<HD, GQ, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int dev = 0, sms = 0;
    B200_CUDA_OK(cudaGetDevice(&dev));
    B200_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    B200_REQUIRE(sms == kNumSMs, "tcd: built for %d SMs, device has %d", kNumSMs, sms);
    B200_CUDA_OK(cudaMemsetAsync(p.bar_ctr, 0, sizeof(unsigned), stream));
    // cooperative launch: all CTAs must be co-resident, the grid barrier depends on it
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(sms);
    cfg.blockDim = dim3(kTThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    B200_CUDA_OK(cudaLaunchKernelEx(&cfg, tcd_decode_kernel<HD, GQ, NB>, p));
    return B200LLM_OK;
}

// one (head_dim, q heads per kv head) pair per translation unit, see B200_TCD_PART in decode_tc.cu
#if B200_TCD_PART == 0
extern template int tcd_launch_t<128, 1, 1>(const TcdParams &, cudaStream_t);
extern template int tcd_launch_t<128, 1, kTcdMaxLanes>(const TcdParams &, cudaStream_t);
extern template int tcd_launch_t<128, 2, 1>(const TcdParams &, cudaStream_t);
extern template int tcd_launch_t<128, 2, kTcdMaxLanes>(const TcdParams &, cudaStream_t);
extern template int tcd_launch_t<128, 3, 1>(const TcdParams &, cudaStream_t);
extern template int tcd_launch_t<128, 3, kTcdMaxLanes>(const TcdParams &, cudaStream_t);
extern template int tcd_launch_t<128, 4, 1>(const TcdParams &, cudaStream_t);
extern template int tcd_launch_t<128, 4, kTcdMaxLanes>(const TcdParams &, cudaStream_t);
extern template int tcd_launch_t<64, 1, 1>(const TcdParams &, cudaStream_t);
extern template int tcd_launch_t<64, 1, kTcdMaxLanes>(const TcdParams &, cudaStream_t);
extern template int tcd_launch_t<64, 2, 1>(const TcdParams &, cudaStream_t);
extern template int tcd_launch_t<64, 2, kTcdMaxLanes>(const TcdParams &, cudaStream_t);
extern template int tcd_launch_t<64, 4, 1>(const TcdParams &, cudaStream_t);
extern template int tcd_launch_t<64, 4, kTcdMaxLanes>(const TcdParams &, cudaStream_t);

int tcd_decode_launch(const TcdParams &p, cudaStream_t stream) {
    B200_REQUIRE(p.n_steps >= 1 && (p.sample || p.n_steps == 1), "tcd: n_steps must be 1 when sampling is external");
    B200_REQUIRE(p.n_splits >= 1, "tcd: n_splits < 1");
    B200_REQUIRE(p.nb >= 1 && p.nb <= kTcdMaxLanes, "tcd: %d lanes (1..%d)", p.nb, kTcdMaxLanes);
    B200_REQUIRE(p.nb == 1 || (p.tp_size == 1 && p.sample != 0 && p.trace == nullptr), "tcd: several lanes need tp_size 1 and in-kernel sampling");
    B200_REQUIRE(p.nb * p.n_kv * p.n_splits <= kNumSMs || !(p.dataflow & 4), "tcd: more attention items than CTAs");
    if (p.sample == 2) {
        B200_REQUIRE(p.cand != nullptr && !p.split_all && p.tp_size == 1, "tcd: in-kernel sampler needs a candidate buffer, whole lm_head tiles and tp_size 1");
        for (int b = 0; b < p.nb; ++b) {
            B200_REQUIRE(tcd_sampler_supported(p.vocab, p.head_dim, p.n_kv > 0 ? p.n_q / p.n_kv : 1, p.temperature_l[b], p.top_k_l[b]),
                         "tcd: in-kernel sampler does not take top_k %d at temperature %g (1..%d) or this vocabulary", p.top_k_l[b],
                         (double)p.temperature_l[b], kTcdSampK);
            B200_REQUIRE(p.repeat_penalty_l[b] > 0.f, "tcd: repeat_penalty must be positive");
        }
    }
    B200_REQUIRE(4 * p.n_layers + 1 <= kTMaxJobs, "tcd: more than 32 layers");
    const int GQ = p.n_kv > 0 ? p.n_q / p.n_kv : 1;
    static const bool force_multi = getenv("B200LLM_TCD_FORCE_MULTI") != nullptr;   // (measurement) one lane on the multi-lane instance
#define B200_TCD_CASE(HD_, G_)                                                  \
    if (p.head_dim == HD_ && GQ == G_)                                          \
        return (p.nb == 1 && !force_multi) ? tcd_launch_t<HD_, G_, 1>(p, stream) : tcd_launch_t<HD_, G_, kTcdMaxLanes>(p, stream);
    B200_TCD_CASE(128, 1) B200_TCD_CASE(128, 2) B200_TCD_CASE(128, 3) B200_TCD_CASE(128, 4)
    B200_TCD_CASE(64, 1) B200_TCD_CASE(64, 2) B200_TCD_CASE(64, 4)
#undef B200_TCD_CASE
    set_error("tcd: unsupported head_dim %d / group %d", p.head_dim, GQ);
    return B200LLM_EINVAL;
}
#elif B200_TCD_PART == 1
template int tcd_launch_t<128, 1, 1>(const TcdParams &, cudaStream_t);
template int tcd_launch_t<128, 1, kTcdMaxLanes>(const TcdParams &, cudaStream_t);
#elif B200_TCD_PART == 2
template int tcd_launch_t<128, 2, 1>(const TcdParams &, cudaStream_t);
template int tcd_launch_t<128, 2, kTcdMaxLanes>(const TcdParams &, cudaStream_t);
#elif B200_TCD_PART == 3
template int tcd_launch_t<128, 3, 1>(const TcdParams &, cudaStream_t);
template int tcd_launch_t<128, 3, kTcdMaxLanes>(const TcdParams &, cudaStream_t);
#elif B200_TCD_PART == 4
template int tcd_launch_t<128, 4, 1>(const TcdParams &, cudaStream_t);
template int tcd_launch_t<128, 4, kTcdMaxLanes>(const TcdParams &, cudaStream_t);
#elif B200_TCD_PART == 5
template int tcd_launch_t<64, 1, 1>(const TcdParams &, cudaStream_t);
template int tcd_launch_t<64, 1, kTcdMaxLanes>(const TcdParams &, cudaStream_t);
#elif B200_TCD_PART == 6
template int tcd_launch_t<64, 2, 1>(const TcdParams &, cudaStream_t);
template int tcd_launch_t<64, 2, kTcdMaxLanes>(const TcdParams &, cudaStream_t);
#elif B200_TCD_PART == 7
template int tcd_launch_t<64, 4, 1>(const TcdParams &, cudaStream_t);
template int tcd_launch_t<64, 4, kTcdMaxLanes>(const TcdParams &, cudaStream_t);
#endif
