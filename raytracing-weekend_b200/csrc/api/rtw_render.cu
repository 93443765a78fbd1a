// rtw_render.cu — C-ABI render entry points.
// Replaces Director::renderFrame's single optixLaunch (Director.cpp:971-986):
// the host drives wavefront iterations until every sample has been flushed.
#include "rtw_ctx.h"

#include <algorithm>
#include <cstring>

using namespace rtw;

static int ensure_wave(rtw_ctx* c, int P) {
    const bool stream = c->cfg.stream;
    if (c->wave_capacity >= P && c->wave_stream == stream) {
        c->wave.P = P;
        return RTW_OK;
    }
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    dfree_all(c->wave_allocs);
    c->wave_capacity = 0;
    auto& pool = c->wave_allocs;
    DWave& w = c->wave;
    w = DWave{};
    if (stream) {   // 2 buffers x 4 classes x 5 float4 per path: any class may hold every path in flight
        RTW_CUDA(c, dalloc(pool, &w.sq, (size_t)2 * 4 * 5 * P));
    } else {
        RTW_CUDA(c, dalloc(pool, &w.rec, P));
        RTW_CUDA(c, dalloc(pool, &w.hit, P));
        RTW_CUDA(c, dalloc(pool, &w.rad, P));
        RTW_CUDA(c, dalloc(pool, &w.queue, (size_t)RTW_Q_SLOTS * P));
        RTW_CUDA(c, dalloc(pool, &w.sh_o, P));
        RTW_CUDA(c, dalloc(pool, &w.sh_d, P));
        RTW_CUDA(c, dalloc(pool, &w.sh_c, P));
    }
    RTW_CUDA(c, dalloc(pool, &w.counters, RTW_C_WORDS));
    w.P = P;
    c->wave_capacity = P;
    c->wave_stream = stream;
    return RTW_OK;
}

static inline uint64_t u64_at(const uint32_t* words, int lo) {
    return (uint64_t)words[lo] | ((uint64_t)words[lo + 1] << 32);
}

extern "C" int rtw_render_accum(rtw_ctx* c, int Nx, int Ny, int sample_first, int sample_stride, int sample_count,
                                int max_depth, float* accum_dev) {
    if (!c) return RTW_ERR_INVALID;
    if (!c->accel_built) return fail(c, RTW_ERR_STATE, "render before rtw_build_accel");
    if (Nx <= 0 || Ny <= 0 || sample_count < 0 || sample_first < 0 || sample_stride < 0 || max_depth < 0 ||
        max_depth > 255 || !accum_dev)
        return fail(c, RTW_ERR_INVALID, "bad render arguments");
    RTW_CUDA(c, cudaSetDevice(c->device));
    const unsigned long long npix = (unsigned long long)Nx * (unsigned long long)Ny;
    if (npix > 0x7fffffffull) return fail(c, RTW_ERR_INVALID, "image too large");
    const unsigned long long total = npix * (unsigned long long)sample_count;
    // Pool size (measured on the B200, tools/diag_tail.py, C2 at 32-1024 spp): 2^22 slots is the fastest at every
    // frame length from 64 spp (23 M paths) up — 18.9 ms against 19.8 ms with 2^21 at 128 spp, the share one GPU
    // of eight renders — while 2^24 slots lose 28 % (the gathers through the queues then span 2.7 GB).
    // Option pool_paths overrides.
    long long P = c->opt_pool_paths > 0 ? c->opt_pool_paths : (total >= (1ull << 23) ? (1ll << 22) : (1ll << 21));
    if ((unsigned long long)P > total) P = (long long)std::max<unsigned long long>(total, 1024ull);
    P = (P + 255) / 256 * 256;
    int rc = ensure_wave(c, (int)P);
    if (rc != RTW_OK) return rc;
    DWave& w = c->wave;
    w.accum = reinterpret_cast<float4*>(accum_dev);
    w.Nx = Nx;
    w.Ny = Ny;
    w.max_depth = max_depth;
    w.sample_first = sample_first;
    w.sample_stride = sample_stride;
    w.total = total;

    uint64_t launches = 0;
    rtw_stats& st = c->stats;
    const double build_ms = st.build_ms;
    st = rtw_stats{};
    st.build_ms = build_ms;
    RTW_CUDA(c, cudaEventRecord(c->ev_begin, c->stream));
    if (max_depth == 0 && total > 0) {   // no ray is ever traced: black samples, counted
        RTW_CUDA(c, launch_add_count(c->stream, (int)npix, sample_count, w.accum, &launches));
        RTW_CUDA(c, cudaEventRecord(c->ev_end, c->stream));
        RTW_CUDA(c, cudaStreamSynchronize(c->stream));
        float ms0 = 0.f;
        cudaEventElapsedTime(&ms0, c->ev_begin, c->ev_end);
        st.render_ms = ms0;
        st.paths = total;
        st.kernel_launches = launches;
        return RTW_OK;
    }
    const bool stream = c->cfg.stream;
    if (stream) RTW_CUDA(c, stream_init(c->stream, w, &launches));
    else RTW_CUDA(c, wavefront_init(c->stream, w, &launches));
    const int lag = c->opt_lag;
    uint64_t iter = 0;
    bool done = (total == 0);
    // Per-stage timing (option "profile"): CUDA events around the stages of SAMPLED
    // iterations, read back after the frame — no host sync inside the frame.
    const int kMaxSampled = 256;
    uint64_t prof_stride = 1;
    int sampled = 0;
    if (c->opt_profile) {
        const uint64_t est = total / (uint64_t)P * 5 + 24;   // ~5 bounces per path
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
    bool sparse = false, filled_prev = false;
    w.trace_fill = 0;
    w.trace_use = 0;
    while (!done) {
        w.parity = (int)(iter & 1);
        w.trace_fill = (sparse && c->cfg.small && !stream) ? 1 : 0;
        w.trace_use = (w.trace_fill && filled_prev) ? 1 : 0;
        filled_prev = w.trace_fill != 0;
        cudaEvent_t* sev = nullptr;
        if (c->opt_profile && sampled < kMaxSampled && iter % prof_stride == 0) sev = c->prof_events.data() + 5 * sampled++;
        if (stream) RTW_CUDA(c, stream_iteration(c->stream, c->scene, w, c->cfg, &launches, sev));
        else RTW_CUDA(c, wavefront_iteration(c->stream, c->scene, w, c->cfg, &launches, sev));
        uint32_t* snap = c->h_counters + (iter % 8) * RTW_C_WORDS;
        RTW_CUDA(c, cudaMemcpyAsync(snap, w.counters, sizeof(uint32_t) * RTW_C_WORDS, cudaMemcpyDeviceToHost, c->stream));
        RTW_CUDA(c, cudaEventRecord(c->ring_events[iter % 8], c->stream));
        iter++;
        if (iter >= (uint64_t)lag) {
            const uint64_t j = iter - lag;
            RTW_CUDA(c, cudaEventSynchronize(c->ring_events[j % 8]));
            const uint32_t* s = c->h_counters + (j % 8) * RTW_C_WORDS;
            if (s[RTW_C_ALIVE] == 0) done = true;
            if (s[RTW_C_ALIVE] < (uint32_t)(P / 2)) sparse = true;
        }
        if (iter > (1ull << 40)) return fail(c, RTW_ERR_STATE, "wavefront did not terminate");
    }
    RTW_CUDA(c, cudaEventRecord(c->ev_end, c->stream));
    uint32_t* fin = c->h_counters;
    RTW_CUDA(c, cudaMemcpyAsync(fin, w.counters, sizeof(uint32_t) * RTW_C_WORDS, cudaMemcpyDeviceToHost, c->stream));
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, c->ev_begin, c->ev_end);
    st.render_ms = ms;
    st.paths = total;
    st.rays_closest = u64_at(fin, RTW_C_RAYS_LO);
    st.rays_shadow = u64_at(fin, RTW_C_SHRAYS_LO);
    st.kernel_launches = launches;
    st.iterations = iter;
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
        const double scale = (double)iter / (double)sampled;   // extrapolate to all iterations
        st.logic_ms = a * scale; st.traverse_ms = b * scale; st.shade_ms = d * scale; st.shadow_ms = e * scale;
    }
    return RTW_OK;
}

extern "C" int rtw_tonemap(rtw_ctx* c, int Nx, int Ny, int spp_total, const float* accum_dev, float* rgba_out_dev) {
    if (!c || Nx <= 0 || Ny <= 0 || spp_total <= 0 || !accum_dev || !rgba_out_dev) return RTW_ERR_INVALID;
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
    int rc = rtw_render_accum(c, Nx, Ny, 0, 1, spp, max_depth, reinterpret_cast<float*>(c->d_accum));
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
