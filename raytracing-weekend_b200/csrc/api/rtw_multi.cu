// rtw_multi.cu — several GPUs of one box behind the C-ABI (include/rtw.h "multi-GPU").
//
// One context, one worker thread and one stream set per GPU; the scene and its LBVH are replicated (every GPU
// builds its own).  A frame is sharded either by SAMPLE (GPU g of G renders samples k = g (mod G) of every
// pixel; one ncclReduce(sum, float, root 0) of the fp32 accumulators per frame) or by image TILE (GPU g renders
// all samples of the rows [g*Ny/G, (g+1)*Ny/G); the row blocks are sent to GPU 0 with ncclSend/ncclRecv — the
// low-sample-count variant, whose result is bit-identical to the 1-GPU frame at 1 spp).  GPU 0 tonemaps and
// the image goes back to the host.  The reference is single-GPU (Director.cpp:971-986); nothing there maps to this.
//
// NCCL is looked up at run time (dlopen of libnccl.so.2 — the copy a host process such as PyTorch has already
// loaded, if any), so librtw.so itself loads on a box without NCCL; if it cannot be found, or if one GPU is
// listed twice (tests on a one-GPU box), the exchange falls back to peer copies + an add kernel.
#include <dlfcn.h>

#include <atomic>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "rtw_ctx.h"

using namespace rtw;

namespace {

// ---- the few NCCL entry points used, resolved with dlsym -----------------------------------------------
typedef struct ncclComm* ncclComm_t;
typedef int ncclResult_t;                 // ncclSuccess == 0
enum { kNcclFloat = 7, kNcclSum = 0 };    // ncclDataType_t::ncclFloat32, ncclRedOp_t::ncclSum (nccl.h)
struct Nccl {
    void* lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, int, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool load() {
        // First whatever NCCL the process already has (a host such as PyTorch brings its own copy, and a second
        // library with the same soname cannot be loaded beside it); else the system's, kept out of the global
        // symbol scope.  A host that imports PyTorch AFTER this call will find this copy under torch's soname:
        // import it first.
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            lib = dlopen(name, RTLD_NOW | RTLD_NOLOAD);
            if (lib) break;
        }
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            if (lib) break;
            lib = dlopen(name, RTLD_NOW | RTLD_LOCAL);
        }
        if (!lib) return false;
#define RTW_NCCL_SYM(field, sym) field = reinterpret_cast<decltype(field)>(dlsym(lib, sym)); if (!field) return false;
        RTW_NCCL_SYM(CommInitAll, "ncclCommInitAll")
        RTW_NCCL_SYM(CommDestroy, "ncclCommDestroy")
        RTW_NCCL_SYM(Reduce, "ncclReduce")
        RTW_NCCL_SYM(Send, "ncclSend")
        RTW_NCCL_SYM(Recv, "ncclRecv")
        RTW_NCCL_SYM(GroupStart, "ncclGroupStart")
        RTW_NCCL_SYM(GroupEnd, "ncclGroupEnd")
        RTW_NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef RTW_NCCL_SYM
        return true;
    }
};

__global__ void k_accum_add(int n, const float4* __restrict__ src, float4* __restrict__ dst) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = src[i];
    float4 b = dst[i];
    b.x += a.x; b.y += a.y; b.z += a.z; b.w += a.w;
    dst[i] = b;
}

}  // namespace

struct rtw_multi {
    std::vector<int> devices;
    std::vector<rtw_ctx*> ctx;
    std::vector<float4*> accum;        // per GPU: Nx*Ny linear sums
    std::vector<float4*> staging;      // on GPU 0, one per peer (fallback exchange only)
    float4* out0 = nullptr;            // on GPU 0: tonemapped frame
    int frame_pixels = 0;
    Nccl nccl;
    std::vector<ncclComm_t> comms;
    bool use_nccl = false;
    int opt_shard = 0;                 // 0 = auto (tiles below 2^23 paths per frame), 1 = samples, 2 = tiles
    int last_shard = 1;
    std::string error;
    rtw_stats total{};
    std::vector<rtw_stats> per_device;
    double exchange_ms = 0.0;
};

static int mfail(rtw_multi* m, int code, const std::string& msg) {
    if (m) m->error = msg;
    return code;
}

// run fn(g) on one host thread per GPU; returns the first non-OK status
template <class F>
static int for_each_gpu(rtw_multi* m, F fn) {
    const int G = (int)m->ctx.size();
    std::vector<int> rc(G, RTW_OK);
    if (G == 1) {
        rc[0] = fn(0);
    } else {
        std::vector<std::thread> th;
        for (int g = 0; g < G; g++) th.emplace_back([&, g]() { rc[g] = fn(g); });
        for (auto& t : th) t.join();
    }
    for (int g = 0; g < G; g++)
        if (rc[g] != RTW_OK) {
            m->error = "GPU " + std::to_string(m->devices[g]) + ": " + (m->ctx[g] ? m->ctx[g]->error : std::string("failed"));
            return rc[g];
        }
    return RTW_OK;
}

// The multi-GPU entry points hop between devices on the CALLER's thread (allocation, peer set-up, tear-down); the
// caller's current device is put back when they return (a host that also uses the CUDA runtime itself — PyTorch in the
// tests — would otherwise find its allocations landing on the last GPU of the list).
struct DeviceGuard {
    int dev = -1;
    DeviceGuard() { if (cudaGetDevice(&dev) != cudaSuccess) { dev = -1; cudaGetLastError(); } }
    ~DeviceGuard() { if (dev >= 0) cudaSetDevice(dev); }
};

extern "C" int rtw_create_multi(const int* devices, int n, rtw_multi** out) {
    DeviceGuard restore_callers_device;
    if (!devices || n < 1 || n > 64 || !out) return RTW_ERR_INVALID;
    rtw_multi* m = new rtw_multi();
    m->devices.assign(devices, devices + n);
    m->ctx.assign(n, nullptr);
    m->accum.assign(n, nullptr);
    m->per_device.assign(n, rtw_stats{});
    for (int g = 0; g < n; g++) {
        int rc = rtw_create(devices[g], &m->ctx[g]);
        if (rc != RTW_OK) {
            for (int k = 0; k < g; k++) rtw_destroy(m->ctx[k]);
            delete m;
            return rc;
        }
    }
    bool distinct = true;
    for (int a = 0; a < n; a++)
        for (int b = a + 1; b < n; b++)
            if (devices[a] == devices[b]) distinct = false;
    if (n > 1 && distinct && m->nccl.load()) {
        m->comms.assign(n, nullptr);
        if (m->nccl.CommInitAll(m->comms.data(), n, devices) == 0) m->use_nccl = true;
        else m->comms.clear();
    }
    if (n > 1 && !m->use_nccl) {      // peer copies need peer access where the GPUs differ
        for (int g = 1; g < n; g++) {
            if (devices[g] == devices[0]) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[0], devices[g]);
            if (can) {
                cudaSetDevice(devices[0]);
                cudaError_t e = cudaDeviceEnablePeerAccess(devices[g], 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            }
        }
    }
    *out = m;
    return RTW_OK;
}

extern "C" int rtw_multi_destroy(rtw_multi* m) {
    DeviceGuard restore_callers_device;
    if (!m) return RTW_ERR_INVALID;
    for (size_t g = 0; g < m->ctx.size(); g++) {
        cudaSetDevice(m->devices[g]);
        if (m->accum[g]) cudaFree(m->accum[g]);
        if (m->use_nccl && m->comms[g]) m->nccl.CommDestroy(m->comms[g]);
    }
    cudaSetDevice(m->devices[0]);
    for (float4* s : m->staging) if (s) cudaFree(s);
    if (m->out0) cudaFree(m->out0);
    for (rtw_ctx* c : m->ctx) rtw_destroy(c);
    delete m;
    return RTW_OK;
}

extern "C" const char* rtw_multi_last_error(const rtw_multi* m) { return m ? m->error.c_str() : "null context"; }
extern "C" int rtw_multi_device_count(const rtw_multi* m) { return m ? (int)m->ctx.size() : 0; }
extern "C" int rtw_multi_uses_nccl(const rtw_multi* m) { return (m && m->use_nccl) ? 1 : 0; }

extern "C" int rtw_multi_set_option(rtw_multi* m, const char* name, int64_t value) {
    DeviceGuard restore_callers_device;
    if (!m || !name) return RTW_ERR_INVALID;
    if (!strcmp(name, "shard")) {
        if (value < 0 || value > 2) return mfail(m, RTW_ERR_INVALID, "shard out of range [0, 2] (0 = auto, 1 = samples, 2 = tiles)");
        m->opt_shard = (int)value;
        return RTW_OK;
    }
    for (size_t g = 0; g < m->ctx.size(); g++) {
        int rc = rtw_set_option(m->ctx[g], name, value);
        if (rc != RTW_OK) return mfail(m, rc, m->ctx[g]->error);
    }
    return RTW_OK;
}

extern "C" int rtw_multi_upload_scene(rtw_multi* m, const rtw_scene_desc* scene) {
    DeviceGuard restore_callers_device;
    if (!m || !scene) return RTW_ERR_INVALID;
    return for_each_gpu(m, [&](int g) { return rtw_upload_scene(m->ctx[g], scene); });
}
extern "C" int rtw_multi_build_accel(rtw_multi* m) {
    DeviceGuard restore_callers_device;
    if (!m) return RTW_ERR_INVALID;
    return for_each_gpu(m, [&](int g) { return rtw_build_accel(m->ctx[g]); });
}

static int ensure_frame(rtw_multi* m, int npix) {
    if (m->frame_pixels == npix) return RTW_OK;
    const int G = (int)m->ctx.size();
    for (int g = 0; g < G; g++) {
        if (cudaSetDevice(m->devices[g]) != cudaSuccess) return mfail(m, RTW_ERR_CUDA, "cudaSetDevice");
        if (m->accum[g]) cudaFree(m->accum[g]);
        m->accum[g] = nullptr;
        if (cudaMalloc(reinterpret_cast<void**>(&m->accum[g]), sizeof(float4) * (size_t)npix) != cudaSuccess)
            return mfail(m, RTW_ERR_CUDA, "cudaMalloc(accumulator)");
    }
    cudaSetDevice(m->devices[0]);
    for (float4* s : m->staging) if (s) cudaFree(s);
    m->staging.assign(G, nullptr);
    if (!m->use_nccl)
        for (int g = 1; g < G; g++)
            if (cudaMalloc(reinterpret_cast<void**>(&m->staging[g]), sizeof(float4) * (size_t)npix) != cudaSuccess)
                return mfail(m, RTW_ERR_CUDA, "cudaMalloc(staging)");
    if (m->out0) cudaFree(m->out0);
    if (cudaMalloc(reinterpret_cast<void**>(&m->out0), sizeof(float4) * (size_t)npix) != cudaSuccess) return mfail(m, RTW_ERR_CUDA, "cudaMalloc(frame)");
    m->frame_pixels = npix;
    return RTW_OK;
}

// rows of GPU g in tile mode
static inline void tile_rows(int Ny, int g, int G, int& first, int& count) {
    first = (int)((long long)g * Ny / G);
    count = (int)((long long)(g + 1) * Ny / G) - first;
}

extern "C" int rtw_multi_render(rtw_multi* m, int Nx, int Ny, int spp, int max_depth, float* rgba_out_host) {
    DeviceGuard restore_callers_device;
    if (!m) return RTW_ERR_INVALID;
    if (!rgba_out_host || spp <= 0 || Nx <= 0 || Ny <= 0 || (unsigned long long)Nx * (unsigned long long)Ny > 0x7fffffffull)
        return mfail(m, RTW_ERR_INVALID, "bad render arguments");
    const int G = (int)m->ctx.size();
    const int npix = Nx * Ny;
    int rc = ensure_frame(m, npix);
    if (rc != RTW_OK) return rc;
    // tiles for short frames (the per-frame fixed cost of a pool dominates when every GPU holds only a few samples
    // of every pixel), samples otherwise (perfect load balance whatever the image shows)
    int shard = m->opt_shard;
    if (shard == 0) shard = ((unsigned long long)npix * (unsigned long long)spp < (1ull << 23) || spp < G) ? 2 : 1;
    if (G == 1) shard = 1;
    m->last_shard = shard;

    // ---- every GPU renders its share into its own accumulator ----
    rc = for_each_gpu(m, [&](int g) -> int {
        rtw_ctx* c = m->ctx[g];
        if (cudaSetDevice(m->devices[g]) != cudaSuccess) return fail(c, RTW_ERR_CUDA, "cudaSetDevice");
        if (cudaMemsetAsync(m->accum[g], 0, sizeof(float4) * (size_t)npix, c->stream) != cudaSuccess) return fail(c, RTW_ERR_CUDA, "cudaMemsetAsync");
        int r;
        if (shard == 1) {
            const int count = spp / G + (g < spp % G ? 1 : 0);
            r = rtw_render_accum(c, Nx, Ny, g, G, count, max_depth, reinterpret_cast<float*>(m->accum[g]));
        } else {
            int first, count;
            tile_rows(Ny, g, G, first, count);
            r = rtw_render_accum_rows(c, Nx, Ny, first, count, 0, 1, spp, max_depth, reinterpret_cast<float*>(m->accum[g]));
        }
        m->per_device[g] = c->stats;
        return r;
    });
    if (rc != RTW_OK) return rc;

    // ---- one exchange per frame ----
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaSetDevice(m->devices[0]);
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, m->ctx[0]->stream);
    if (G > 1 && m->use_nccl) {
        std::atomic<int> bad(0);
        rc = for_each_gpu(m, [&](int g) -> int {
            rtw_ctx* c = m->ctx[g];
            cudaSetDevice(m->devices[g]);
            ncclResult_t r = 0;
            if (shard == 1) {
                r = m->nccl.Reduce(m->accum[g], m->accum[g], (size_t)npix * 4, kNcclFloat, kNcclSum, 0, m->comms[g], c->stream);
            } else {
                m->nccl.GroupStart();
                if (g == 0) {
                    for (int p = 1; p < G && r == 0; p++) {
                        int first, count;
                        tile_rows(Ny, p, G, first, count);
                        if (count > 0) r = m->nccl.Recv(m->accum[0] + (size_t)first * Nx, (size_t)count * Nx * 4, kNcclFloat, p, m->comms[0], c->stream);
                    }
                } else {
                    int first, count;
                    tile_rows(Ny, g, G, first, count);
                    if (count > 0) r = m->nccl.Send(m->accum[g] + (size_t)first * Nx, (size_t)count * Nx * 4, kNcclFloat, 0, m->comms[g], c->stream);
                }
                const ncclResult_t r2 = m->nccl.GroupEnd();
                if (r == 0) r = r2;
            }
            if (r != 0) { bad = 1; return fail(c, RTW_ERR_CUDA, std::string("NCCL: ") + m->nccl.GetErrorString(r)); }
            if (cudaStreamSynchronize(c->stream) != cudaSuccess) return fail(c, RTW_ERR_CUDA, "cudaStreamSynchronize after the exchange");
            return RTW_OK;
        });
        if (rc != RTW_OK) return rc;
    } else if (G > 1) {
        // fallback: peer copies into GPU 0 (+ an add kernel for sample shards)
        cudaStream_t s0 = m->ctx[0]->stream;
        for (int g = 1; g < G; g++) {
            if (shard == 1) {
                if (cudaMemcpyPeerAsync(m->staging[g], m->devices[0], m->accum[g], m->devices[g], sizeof(float4) * (size_t)npix, s0) != cudaSuccess)
                    return mfail(m, RTW_ERR_CUDA, "cudaMemcpyPeerAsync");
                k_accum_add<<<(npix + 255) / 256, 256, 0, s0>>>(npix, m->staging[g], m->accum[0]);
            } else {
                int first, count;
                tile_rows(Ny, g, G, first, count);
                if (count > 0 && cudaMemcpyPeerAsync(m->accum[0] + (size_t)first * Nx, m->devices[0], m->accum[g] + (size_t)first * Nx, m->devices[g],
                                                     sizeof(float4) * (size_t)count * Nx, s0) != cudaSuccess)
                    return mfail(m, RTW_ERR_CUDA, "cudaMemcpyPeerAsync");
            }
        }
        if (cudaGetLastError() != cudaSuccess) return mfail(m, RTW_ERR_CUDA, "exchange");
    }
    cudaSetDevice(m->devices[0]);
    cudaEventRecord(e1, m->ctx[0]->stream);

    // ---- GPU 0: tonemap, image back to the host ----
    rc = rtw_tonemap(m->ctx[0], Nx, Ny, spp, reinterpret_cast<const float*>(m->accum[0]), reinterpret_cast<float*>(m->out0));
    if (rc != RTW_OK) return mfail(m, rc, m->ctx[0]->error);
    if (cudaMemcpy(rgba_out_host, m->out0, sizeof(float4) * (size_t)npix, cudaMemcpyDeviceToHost) != cudaSuccess) return mfail(m, RTW_ERR_CUDA, "cudaMemcpy(frame)");
    float ex = 0.f;
    cudaEventElapsedTime(&ex, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    m->exchange_ms = ex;

    rtw_stats t{};
    for (int g = 0; g < G; g++) {
        const rtw_stats& s = m->per_device[g];
        t.paths += s.paths; t.rays_closest += s.rays_closest; t.rays_shadow += s.rays_shadow;
        t.kernel_launches += s.kernel_launches; t.iterations += s.iterations;
        if (s.render_ms > t.render_ms) t.render_ms = s.render_ms;      // the GPUs run side by side: the slowest one
        if (s.build_ms > t.build_ms) t.build_ms = s.build_ms;
        t.waves = s.waves; t.pool_paths = s.pool_paths;
    }
    t.kernel_launches += 1 + (uint64_t)((G > 1 && !m->use_nccl && shard == 1) ? G - 1 : 0);
    m->total = t;
    return RTW_OK;
}

extern "C" int rtw_multi_get_linear(rtw_multi* m, float* rgba_linear_mean_host, int n_floats, int spp) {
    DeviceGuard restore_callers_device;
    if (!m || !rgba_linear_mean_host || spp <= 0) return RTW_ERR_INVALID;
    if (n_floats != m->frame_pixels * 4) return mfail(m, RTW_ERR_INVALID, "n_floats != Nx*Ny*4");
    cudaSetDevice(m->devices[0]);
    if (cudaMemcpy(rgba_linear_mean_host, m->accum[0], sizeof(float) * (size_t)n_floats, cudaMemcpyDeviceToHost) != cudaSuccess)
        return mfail(m, RTW_ERR_CUDA, "cudaMemcpy(accumulator)");
    const float inv = 1.0f / (float)spp;
    for (int i = 0; i < n_floats; i += 4) {
        rgba_linear_mean_host[i] *= inv; rgba_linear_mean_host[i + 1] *= inv; rgba_linear_mean_host[i + 2] *= inv;
    }
    return RTW_OK;
}

extern "C" int rtw_multi_get_stats(rtw_multi* m, rtw_stats* total, rtw_stats* per_device, double* exchange_ms, int* shard) {
    if (!m) return RTW_ERR_INVALID;
    if (total) *total = m->total;
    if (per_device) memcpy(per_device, m->per_device.data(), sizeof(rtw_stats) * m->per_device.size());
    if (exchange_ms) *exchange_ms = m->exchange_ms;
    if (shard) *shard = m->last_shard;
    return RTW_OK;
}
