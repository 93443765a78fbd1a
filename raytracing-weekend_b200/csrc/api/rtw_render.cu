// rtw_render.cu — C-ABI render entry points.
// Replaces Director::renderFrame's single optixLaunch (Director.cpp:971-986):
// the host drives wavefront iterations until every sample has been flushed.
#include "rtw_ctx.h"

#include <algorithm>
#include <cstring>

using namespace rtw;

static int ensure_wave(rtw_ctx* c, int j, int P) {
    rtw_ctx::Wave& wv = c->waves[j];
    const bool stream = c->cfg.stream;
    if (!wv.stream) {
        if (j == 0) wv.stream = c->stream;
        else RTW_CUDA(c, cudaStreamCreateWithFlags(&wv.stream, cudaStreamNonBlocking));
        for (auto& e : wv.ring_events) RTW_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        RTW_CUDA(c, cudaEventCreateWithFlags(&wv.ev_done, cudaEventDisableTiming));
        for (auto& e : wv.ev_iter) RTW_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        RTW_CUDA(c, cudaMallocHost(reinterpret_cast<void**>(&wv.h_counters), sizeof(uint32_t) * RTW_C_WORDS * 8));
    }
    const int sort_rays = c->opt_sort_rays >= 0 ? c->opt_sort_rays : 1;   // -1 = auto (rtw_ctx.h): every walker scene
    const bool want_sort = sort_rays > 0 && !stream && c->cfg.walker && !c->cfg.small && !(c->scene.root & RTW_REF_LEAF);
    const int key_ob = c->opt_sort_obits, key_db = std::max(1, std::min(c->opt_sort_dbits, 6 - key_ob));   // at most 2^18 bins
    const int key_bits = 3 * (key_ob + key_db);
    const int n_bins = (1 << key_bits) + (c->opt_sort_camera ? RTW_CAMERA_BINS : 0);   // + the camera rays' direction bins (kernels/wavefront.cu)
    auto set_key_cells = [&](DWave& w) {   // the ray keys' origin cells span the bounds of the primitive centroids
        w.key_obits = key_ob;
        w.key_dbits = key_db;
        w.key_camera = c->opt_sort_camera;
        w.n_bins = n_bins;
        for (int k = 0; k < 3; k++) {
            const float ext = c->key_hi[k] - c->key_lo[k];
            (&w.key_lo.x)[k] = c->key_lo[k];
            (&w.key_scale.x)[k] = ext > 0.f ? (float)(1 << c->opt_sort_obits) / ext : 0.f;
        }
    };
    if (wv.capacity >= P && wv.for_stream == stream && (wv.w.bins != nullptr) == want_sort && (!want_sort || wv.key_bits == n_bins)) {
        wv.w.P = P;
        set_key_cells(wv.w);
        return RTW_OK;
    }
    RTW_CUDA(c, cudaStreamSynchronize(wv.stream));
    dfree_all(wv.allocs);
    wv.capacity = 0;
    auto& pool = wv.allocs;
    DWave& w = wv.w;
    w = DWave{};
    if (stream) {   // 2 buffers x 4 classes x 5 float4 per path: any class may hold every path in flight
        RTW_CUDA(c, dalloc(pool, &w.sq, (size_t)2 * 4 * 5 * P));
    } else {
        RTW_CUDA(c, dalloc(pool, &w.rec, P));
        RTW_CUDA(c, dalloc(pool, &w.rad, P));
        RTW_CUDA(c, dalloc(pool, &w.queue, (size_t)RTW_Q_SLOTS * P));
        RTW_CUDA(c, dalloc(pool, &w.sh_o, P));
        RTW_CUDA(c, dalloc(pool, &w.sh_d, P));
        RTW_CUDA(c, dalloc(pool, &w.sh_c, P));
    }
    RTW_CUDA(c, dalloc(pool, &w.counters, RTW_C_WORDS));
    if (want_sort) {
        RTW_CUDA(c, dalloc(pool, &w.bins, (size_t)n_bins + 1 + (size_t)(n_bins + 1023) / 1024));   // counts / first positions, the total, the chunk sums
        RTW_CUDA(c, dalloc(pool, &w.bin_rank, P));
        RTW_CUDA(c, dalloc(pool, &w.order, P));
        wv.key_bits = n_bins;
    }
    set_key_cells(w);
    w.P = P;
    wv.capacity = P;
    wv.for_stream = stream;
    return RTW_OK;
}

static inline uint64_t u64_at(const uint32_t* words, int lo) {
    return (uint64_t)words[lo] | ((uint64_t)words[lo + 1] << 32);
}

extern "C" int rtw_render_accum(rtw_ctx* c, int Nx, int Ny, int sample_first, int sample_stride, int sample_count,
                                int max_depth, float* accum_dev) {
    return rtw_render_accum_rows(c, Nx, Ny, 0, Ny, sample_first, sample_stride, sample_count, max_depth, accum_dev);
}

extern "C" int rtw_render_accum_rows(rtw_ctx* c, int Nx, int Ny, int row_first, int row_count, int sample_first, int sample_stride,
                                     int sample_count, int max_depth, float* accum_dev) {
    if (!c) return RTW_ERR_INVALID;
    if (!c->accel_built) return fail(c, RTW_ERR_STATE, "render before rtw_build_accel");
    if (Nx <= 0 || Ny <= 0 || sample_count < 0 || sample_first < 0 || sample_stride < 0 || max_depth < 0 ||
        max_depth > 255 || !accum_dev || row_first < 0 || row_count < 0 || row_first + row_count > Ny)
        return fail(c, RTW_ERR_INVALID, "bad render arguments");
    RTW_CUDA(c, cudaSetDevice(c->device));
    if ((unsigned long long)Nx * (unsigned long long)Ny > 0x7fffffffull) return fail(c, RTW_ERR_INVALID, "image too large");
    const unsigned long long npix = (unsigned long long)Nx * (unsigned long long)row_count;   // pixels of this call's tile
    const uint32_t pixel_first = (uint32_t)row_first * (uint32_t)Nx;
    const unsigned long long total = npix * (unsigned long long)sample_count;

    uint64_t launches = 0;
    rtw_stats& st = c->stats;
    const double build_ms = st.build_ms;
    st = rtw_stats{};
    st.build_ms = build_ms;
    if (max_depth == 0 && total > 0) {
        RTW_CUDA(c, cudaEventRecord(c->ev_begin, c->stream));   // no ray is ever traced: black samples, counted
        RTW_CUDA(c, launch_add_count(c->stream, (int)npix, sample_count, reinterpret_cast<float4*>(accum_dev) + pixel_first, &launches));
        RTW_CUDA(c, cudaEventRecord(c->ev_end, c->stream));
        RTW_CUDA(c, cudaStreamSynchronize(c->stream));
        float ms0 = 0.f;
        cudaEventElapsedTime(&ms0, c->ev_begin, c->ev_end);
        st.render_ms = ms0;
        st.paths = total;
        st.kernel_launches = launches;
        return RTW_OK;
    }

    // Waves: the frame's samples are dealt round-robin to n_waves independent pools, each on its own stream.
    // Their launches overlap on the device — while one wave's traversal launch saturates instruction issue,
    // another's shading launch streams path state through HBM, and the tail of every launch (CTAs running out)
    // is filled by the other wave's CTAs.  Measured on the B200, C2: 2666 -> 3042 Mpaths/s with two waves of 2^21
    // slots against one of 2^22 (tools/dual_ctx.py).  A sample is the same random walk in whichever wave it runs
    // (seed = tea<64>(pixel, k)), and every wave adds into the same accumulator with atomics.
    int n_waves = c->opt_waves > 0 ? c->opt_waves : (total >= (1ull << 23) ? 2 : 1);
    if (sample_count < n_waves) n_waves = sample_count > 0 ? sample_count : 1;
    // Pool size (measured on the B200, tools/diag_tail.py, C2 at 32-1024 spp): 2^22 slots in total is the fastest
    // at every frame length from 64 spp (23 M paths) up; 2^24 slots lose 28 % (the gathers through the queues then
    // span 2.7 GB).  Option pool_paths overrides (slots per wave).
    // Round 2 (tools/sweep_c2_pool.sh, after the hit record moved into the slot record): on the by-value small scenes a
    // frame of >= 2^28 paths (C2: 3.7e8) is 3 % faster with 2.75 M slots per wave (3003 / 3024 against 2926 / 2900
    // Mpaths/s, two runs each; 2.25 M: 2824, 3 M: 2938, 4 M: 3015); the 128 spp share of one GPU of eight is not
    // (2737 against 2770 with 2.5 M), so shorter frames keep 2^22 in total.
    long long P_all = total >= (1ull << 23) ? (1ll << 22) : (1ll << 21);
    if (c->cfg.small && !c->cfg.stream && total >= (1ull << 28)) P_all = 2ll * 2883584;
    struct Run {
        unsigned long long total = 0;
        uint64_t iter = 0;
        bool done = false, sparse = false, filled_prev = false, tail = false;
    } run[rtw_ctx::kMaxWaves];
    const bool stream = c->cfg.stream;
    for (int j = 0; j < n_waves; j++) {
        // wave j renders samples j, j + n_waves, ... of this call's samples
        const int count_j = (sample_count - j + n_waves - 1) / n_waves;
        run[j].total = npix * (unsigned long long)count_j;
        long long P = c->opt_pool_paths > 0 ? c->opt_pool_paths : P_all / n_waves;
        if ((unsigned long long)P > run[j].total) P = (long long)std::max<unsigned long long>(run[j].total, 1024ull);
        P = (P + 255) / 256 * 256;
        int rc = ensure_wave(c, j, (int)P);   // (re)allocation, if any, happens before the frame's clock starts
        if (rc != RTW_OK) return rc;
    }
    RTW_CUDA(c, cudaEventRecord(c->ev_begin, c->stream));
    for (int j = 0; j < n_waves; j++) {
        DWave& w = c->waves[j].w;
        w.accum = reinterpret_cast<float4*>(accum_dev);
        w.Nx = Nx;
        w.Ny = Ny;
        w.pixel_first = pixel_first;
        w.pixel_count = (uint32_t)npix;
        w.max_depth = max_depth;
        w.sample_first = sample_first + j * sample_stride;
        w.sample_stride = sample_stride * n_waves;
        w.total = run[j].total;
        w.trace_fill = 0;
        w.trace_use = 0;
        run[j].done = (run[j].total == 0);
        cudaStream_t sj = c->waves[j].stream;
        if (j > 0) RTW_CUDA(c, cudaStreamWaitEvent(sj, c->ev_begin, 0));   // the caller's work on `stream` (e.g. the clear) comes first
        if (stream) RTW_CUDA(c, stream_init(sj, w, &launches));
        else RTW_CUDA(c, wavefront_init(sj, w, &launches));
    }
    const int lag = c->opt_lag;
    // Per-stage timing (option "profile"): CUDA events around the stages of SAMPLED iterations of wave 0, read
    // back after the frame — no host sync inside the frame.  With several waves the stages of wave 0 share the
    // device with the other waves' launches, so their sum exceeds wave 0's share of the frame.
    const int kMaxSampled = 64;    // (reading 5 events per sampled iteration back costs ~1.5 us each after the frame)
    uint64_t prof_stride = 1;
    int sampled = 0;
    if (c->opt_profile) {
        const uint64_t est = run[0].total / (uint64_t)c->waves[0].w.P * 5 + 24;   // ~5 bounces per path
        prof_stride = est / kMaxSampled + 1;
        while ((int)c->prof_events.size() < 5 * kMaxSampled) {
            cudaEvent_t e = nullptr;
            RTW_CUDA(c, cudaEventCreate(&e));
            c->prof_events.push_back(e);
        }
    }
    // Sparse mode: once the counters show fewer than half of the slots with a ray in flight (the drain tail of
    // the frame, or a frame smaller than the pool), the small-scene traversal walks a list of those slots —
    // kept by the shading and regeneration kernels from one iteration earlier — instead of scanning the pool.
    bool all_done = false;
    uint64_t iterations = 0;
    while (!all_done) {
        // Phase experiments (options, both OFF by default; tools/gpu_variants.sh, three runs each, C2): the waves
        // left to run freely settle half-way between "same stage side by side" and "opposite stages" and render
        // 3040-3053 Mpaths/s; forced into LOCKSTEP (iteration i of every wave starts when all waves have finished
        // iteration i - 1) 2939-2942; started half an iteration apart (STAGGER) 2927-2976.
        if (c->opt_lockstep && n_waves > 1) {
            int live = 0;
            for (int j = 0; j < n_waves; j++) live += run[j].done ? 0 : 1;
            if (live > 1) {
                for (int j = 0; j < n_waves; j++)
                    if (!run[j].done) RTW_CUDA(c, cudaEventRecord(c->waves[j].ev_iter[run[j].iter & 1], c->waves[j].stream));
                for (int j = 0; j < n_waves; j++)
                    for (int k = 0; k < n_waves; k++)
                        if (k != j && !run[j].done && !run[k].done)
                            RTW_CUDA(c, cudaStreamWaitEvent(c->waves[j].stream, c->waves[k].ev_iter[run[k].iter & 1], 0));
            }
        }
        for (int j = 0; j < n_waves; j++) {
            Run& r = run[j];
            if (r.done) continue;
            rtw_ctx::Wave& wv = c->waves[j];
            DWave& w = wv.w;
            w.parity = (int)(r.iter & 1);
            w.first_iter = r.iter == 0 ? 1 : 0;
            if (r.tail) {
                // Drain tail (kernels/stream.cu k_tail_*): no sample left to start and few paths alive — instead of
                // ~30 more iterations of four nearly empty launches, one launch runs every surviving path to its end.
                w.trace_use = r.filled_prev ? 1 : 0;   // the last shading launch listed the survivors (sparse mode)
                RTW_CUDA(c, launch_tail(wv.stream, c->scene, w, c->cfg, &launches));
                r.done = true;
                continue;
            }
            w.trace_fill = (r.sparse && c->cfg.small && !stream) ? 1 : 0;
            w.trace_use = (w.trace_fill && r.filled_prev) ? 1 : 0;
            r.filled_prev = w.trace_fill != 0;
            cudaEvent_t* sev = nullptr;
            if (j == 0 && c->opt_profile && sampled < kMaxSampled && r.iter % prof_stride == 0) sev = c->prof_events.data() + 5 * sampled++;
            LaunchCfg cfg = c->cfg;
            cfg.sort_shadow = c->opt_sort_rays >= 2;   // (auto bins the closest-hit rays only)
            cfg.mid_event = nullptr;
            if (c->opt_stagger && r.iter == 0 && n_waves > 1) {
                if ((j & 1) == 0 && j + 1 < n_waves) cfg.mid_event = wv.ev_done;
                if (j & 1) RTW_CUDA(c, cudaStreamWaitEvent(wv.stream, c->waves[j - 1].ev_done, 0));
            }
            if (stream) RTW_CUDA(c, stream_iteration(wv.stream, c->scene, w, cfg, &launches, sev));
            else RTW_CUDA(c, wavefront_iteration(wv.stream, c->scene, w, cfg, &launches, sev));
            uint32_t* snap = wv.h_counters + (r.iter % 8) * RTW_C_WORDS;
            RTW_CUDA(c, cudaMemcpyAsync(snap, w.counters, sizeof(uint32_t) * RTW_C_WORDS, cudaMemcpyDeviceToHost, wv.stream));
            RTW_CUDA(c, cudaEventRecord(wv.ring_events[r.iter % 8], wv.stream));
            r.iter++;
            if (r.iter > (1ull << 40)) return fail(c, RTW_ERR_STATE, "wavefront did not terminate");
        }
        all_done = true;
        for (int j = 0; j < n_waves; j++) {
            Run& r = run[j];
            if (r.done) continue;
            if (r.iter >= (uint64_t)lag) {
                rtw_ctx::Wave& wv = c->waves[j];
                const uint64_t k = r.iter - lag;
                RTW_CUDA(c, cudaEventSynchronize(wv.ring_events[k % 8]));
                const uint32_t* s = wv.h_counters + (k % 8) * RTW_C_WORDS;
                if (s[RTW_C_ALIVE] == 0) r.done = true;
                if (s[RTW_C_ALIVE] < (uint32_t)(wv.w.P / 2)) r.sparse = true;
                if (!stream && c->opt_tail_paths > 0 && s[RTW_C_ALIVE] < (uint32_t)c->opt_tail_paths &&
                    u64_at(s, RTW_C_NEXT_LO) >= r.total)
                    r.tail = true;
            }
            if (!r.done) all_done = false;
        }
    }
    for (int j = 1; j < n_waves; j++) {   // the frame ends on `stream` when every wave has ended
        RTW_CUDA(c, cudaEventRecord(c->waves[j].ev_done, c->waves[j].stream));
        RTW_CUDA(c, cudaStreamWaitEvent(c->stream, c->waves[j].ev_done, 0));
    }
    RTW_CUDA(c, cudaEventRecord(c->ev_end, c->stream));
    for (int j = 0; j < n_waves; j++)
        RTW_CUDA(c, cudaMemcpyAsync(c->waves[j].h_counters, c->waves[j].w.counters, sizeof(uint32_t) * RTW_C_WORDS,
                                    cudaMemcpyDeviceToHost, c->waves[j].stream));
    for (int j = 0; j < n_waves; j++) RTW_CUDA(c, cudaStreamSynchronize(c->waves[j].stream));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, c->ev_begin, c->ev_end);
    st.render_ms = ms;
    st.paths = total;
    for (int j = 0; j < n_waves; j++) {
        st.rays_closest += u64_at(c->waves[j].h_counters, RTW_C_RAYS_LO);
        st.rays_shadow += u64_at(c->waves[j].h_counters, RTW_C_SHRAYS_LO);
        st.nee_unoccluded += u64_at(c->waves[j].h_counters, RTW_C_UNOCC_LO);
        iterations += run[j].iter;
    }
    st.kernel_launches = launches;
    st.iterations = iterations;
    st.waves = (uint64_t)n_waves;
    st.pool_paths = (uint64_t)c->waves[0].w.P;
    if (sampled > 0) {
        double a = 0, b = 0, d = 0, e = 0;
        for (int i = 0; i < sampled; i++) {
            cudaEvent_t* ev = c->prof_events.data() + 5 * i;
            float x = 0;
            cudaEventElapsedTime(&x, ev[0], ev[1]); a += x;
            cudaEventElapsedTime(&x, ev[1], ev[2]); b += x;
            cudaEventElapsedTime(&x, ev[2], ev[3]); d += x;
            cudaEventElapsedTime(&x, ev[3], ev[4]); e += x;
        }
        const double scale = (double)iterations / (double)sampled;   // extrapolate to all iterations of all waves
        st.logic_ms = a * scale; st.traverse_ms = b * scale; st.shade_ms = d * scale; st.shadow_ms = e * scale;
    }
    return check_overflow(c, &st.stack_overflows);
}

extern "C" int rtw_tonemap(rtw_ctx* c, int Nx, int Ny, int spp_total, const float* accum_dev, float* rgba_out_dev) {
    if (!c || Nx <= 0 || Ny <= 0 || spp_total <= 0 || !accum_dev || !rgba_out_dev) return RTW_ERR_INVALID;
    if ((unsigned long long)Nx * (unsigned long long)Ny > 0x7fffffffull) return fail(c, RTW_ERR_INVALID, "image size out of range");
    RTW_CUDA(c, cudaSetDevice(c->device));
    uint64_t l = 0;
    RTW_CUDA(c, launch_tonemap(c->stream, Nx * Ny, spp_total, reinterpret_cast<const float4*>(accum_dev),
                               reinterpret_cast<float4*>(rgba_out_dev), &l));
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    c->stats.kernel_launches += l;
    return RTW_OK;
}

extern "C" int rtw_render(rtw_ctx* c, int Nx, int Ny, int spp, int max_depth, float* rgba_out_host) {
    if (!c) return RTW_ERR_INVALID;
    if (!rgba_out_host || spp <= 0 || Nx <= 0 || Ny <= 0) return fail(c, RTW_ERR_INVALID, "bad render arguments");
    RTW_CUDA(c, cudaSetDevice(c->device));
    if (Nx <= 0 || Ny <= 0 || (unsigned long long)Nx * (unsigned long long)Ny > 0x7fffffffull) return fail(c, RTW_ERR_INVALID, "image size out of range");
    const int npix = Nx * Ny;
    if (c->frame_pixels != npix) {
        if (c->d_accum) cudaFree(c->d_accum);
        if (c->d_out) cudaFree(c->d_out);
        c->d_accum = nullptr;
        c->d_out = nullptr;
        c->frame_pixels = 0;
        RTW_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->d_accum), sizeof(float4) * npix));
        RTW_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->d_out), sizeof(float4) * npix));
        c->frame_pixels = npix;
    }
    RTW_CUDA(c, cudaMemsetAsync(c->d_accum, 0, sizeof(float4) * npix, c->stream));
    // option seed_base (SURVEY.md §8b `seed_base`): the frame's samples are k = seed_base .. seed_base + spp - 1, each
    // seeded tea<64>(pixel, k) — two frames with disjoint ranges are independent estimates of the same image
    int rc = rtw_render_accum(c, Nx, Ny, c->opt_seed_base, 1, spp, max_depth, reinterpret_cast<float*>(c->d_accum));
    if (rc != RTW_OK) return rc;
    c->frame_spp = spp;
    uint64_t l = 0;
    RTW_CUDA(c, launch_tonemap(c->stream, npix, spp, c->d_accum, c->d_out, &l));
    c->stats.kernel_launches += l;
    RTW_CUDA(c, cudaMemcpyAsync(rgba_out_host, c->d_out, sizeof(float4) * npix, cudaMemcpyDeviceToHost, c->stream));
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    return RTW_OK;
}

extern "C" int rtw_get_linear(rtw_ctx* c, float* out, int n_floats) {
    if (!c || !out) return RTW_ERR_INVALID;
    if (!c->d_accum || c->frame_spp <= 0) return fail(c, RTW_ERR_STATE, "no frame rendered");
    if (n_floats != c->frame_pixels * 4) return fail(c, RTW_ERR_INVALID, "n_floats != Nx*Ny*4");
    RTW_CUDA(c, cudaSetDevice(c->device));
    RTW_CUDA(c, cudaMemcpy(out, c->d_accum, sizeof(float) * n_floats, cudaMemcpyDeviceToHost));
    const float inv = 1.0f / (float)c->frame_spp;
    for (int i = 0; i < n_floats; i += 4) {
        out[i] *= inv; out[i + 1] *= inv; out[i + 2] *= inv;  // w keeps the sample count
    }
    return RTW_OK;
}

extern "C" int rtw_denoise(rtw_ctx* c, int Nx, int Ny, int iterations, float sigma, const float* in_host, float* out_host) {
    if (!c) return RTW_ERR_INVALID;
    if (!in_host || !out_host || Nx <= 0 || Ny <= 0 || iterations < 1 || iterations > 8 || !(sigma > 0.f) ||
        (unsigned long long)Nx * (unsigned long long)Ny > 0x7fffffffull)
        return fail(c, RTW_ERR_INVALID, "bad denoise arguments");
    RTW_CUDA(c, cudaSetDevice(c->device));
    const size_t bytes = sizeof(float4) * (size_t)Nx * Ny;
    float4 *a = nullptr, *b = nullptr, *res = nullptr;
    RTW_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&a), bytes));
    if (cudaMalloc(reinterpret_cast<void**>(&b), bytes) != cudaSuccess) { cudaFree(a); return fail(c, RTW_ERR_CUDA, "cudaMalloc"); }
    uint64_t l = 0;
    cudaError_t e = cudaMemcpyAsync(a, in_host, bytes, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = launch_denoise(c->stream, Nx, Ny, iterations, sigma, a, b, &res, &l);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_host, res, bytes, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(a);
    cudaFree(b);
    if (e != cudaSuccess) return cuda_fail(c, e, "rtw_denoise");
    c->stats.kernel_launches += l;
    return RTW_OK;
}
