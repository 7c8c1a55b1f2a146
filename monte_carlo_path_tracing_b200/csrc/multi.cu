// csrc/multi.cu -- multi-GPU behind the C ABI: mcpt_multi_* / mcpt_render_multi.
//
// The path shards by samples per pixel: every (pixel, sample) is independent in the reference (main.cpp:567-577, samples
// only meet in `sum = sum + I`), and the renderer's uniforms are keyed on the GLOBAL sample index, so device r of n
// renders all pixels for its contiguous range of sample indices, normalised by the global spp.  The scene is built
// once (one SAH BVH build on the host) and its packed blob is replicated to every device; each device gets one host
// thread, one stream and one workspace.  The films meet in ONE collective: ncclReduce(float32, sum, W*H*3) onto
// devices[0] over NVLink, then one device-to-host copy -- there is no exchange inside the render, so no fused
// compute + collective kernel is warranted.
//
// NCCL is bound at run time (dlopen of libnccl.so.2, the same library a host program or PyTorch already carries):
// libmcpt.so keeps no link-time dependency on it, and a process that never renders on more than one device never
// loads it.  There is no fallback: without NCCL a multi-device render fails with MCPT_ERR_NCCL.
#include <dlfcn.h>
#include <nccl.h>
#include <algorithm>
#include <cstring>
#include <memory>
#include <thread>
#include "mcpt_internal.h"

namespace mcpt {
namespace {

struct NcclApi {
    void* lib = nullptr;
    decltype(&ncclCommInitAll) CommInitAll = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclReduce) Reduce = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclGetVersion) GetVersion = nullptr;
};

const NcclApi& nccl()
{
    static NcclApi api;
    static bool tried = false;
    if (!tried) {
        tried = true;
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            api.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (api.lib) break;
        }
        if (api.lib) {
            api.CommInitAll = (decltype(api.CommInitAll))dlsym(api.lib, "ncclCommInitAll");
            api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.lib, "ncclCommDestroy");
            api.Reduce = (decltype(api.Reduce))dlsym(api.lib, "ncclReduce");
            api.GroupStart = (decltype(api.GroupStart))dlsym(api.lib, "ncclGroupStart");
            api.GroupEnd = (decltype(api.GroupEnd))dlsym(api.lib, "ncclGroupEnd");
            api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.lib, "ncclGetErrorString");
            api.GetVersion = (decltype(api.GetVersion))dlsym(api.lib, "ncclGetVersion");
        }
    }
    if (!api.lib || !api.CommInitAll || !api.CommDestroy || !api.Reduce || !api.GroupStart || !api.GroupEnd)
        throw Error(MCPT_ERR_NCCL, "libnccl.so.2 not found (or incomplete): a multi-device render needs NCCL for its one reduce");
    return api;
}

#define MCPT_NCCL(expr)                                                                                      \
    do {                                                                                                     \
        ncclResult_t r__ = (expr);                                                                           \
        if (r__ != ncclSuccess)                                                                              \
            throw ::mcpt::Error(MCPT_ERR_NCCL, std::string(#expr) + ": " +                                   \
                                                   (nccl().GetErrorString ? nccl().GetErrorString(r__) : "NCCL error"));  \
    } while (0)

struct Slot {  // what a device's worker thread leaves for the caller
    int code = MCPT_OK;
    std::string msg;
    McptStats st;
};

// Runs body(r) for r in [0, n) on one host thread each; rethrows the first failure on the calling thread
// (mcpt_last_error is thread-local, so worker threads must not be the ones reporting).
template <class F> void per_device(int n, std::vector<Slot>& slots, F&& body)
{
    std::vector<std::thread> th;
    for (int r = 0; r < n; r++)
        th.emplace_back([&, r] {
            try {
                body(r);
            } catch (const Error& e) {
                slots[r].code = e.code; slots[r].msg = e.what();
            } catch (const std::exception& e) {
                slots[r].code = MCPT_ERR_INVALID; slots[r].msg = e.what();
            }
        });
    for (auto& t : th) t.join();
    for (int r = 0; r < n; r++)
        if (slots[r].code != MCPT_OK) throw Error(slots[r].code, "device " + std::to_string(r) + ": " + slots[r].msg);
}

}  // namespace
}  // namespace mcpt

struct McptMulti {
    std::vector<int> devices;
    std::vector<McptScene*> scenes;
    std::vector<cudaStream_t> streams;
    std::vector<float*> films;  // W*H*3 floats per device
    size_t film_bytes = 0;
    std::vector<ncclComm_t> comms;
    cudaEvent_t ev[2] = {nullptr, nullptr};  // on devices[0]: around the reduce
};

using namespace mcpt;

extern "C" {

void mcpt_multi_destroy(McptMulti* m)
{
    if (!m) return;
    for (size_t r = 0; r < m->devices.size(); r++) {
        cudaSetDevice(m->devices[r]);
        if (r < m->comms.size() && m->comms[r]) { try { nccl().CommDestroy(m->comms[r]); } catch (...) {} }
        if (r < m->films.size() && m->films[r]) cudaFree(m->films[r]);
        if (r < m->streams.size() && m->streams[r]) cudaStreamDestroy(m->streams[r]);
        if (r == 0) for (auto& e : m->ev) if (e) cudaEventDestroy(e);
        if (r < m->scenes.size() && m->scenes[r]) mcpt_scene_destroy(m->scenes[r]);
    }
    delete m;
}

int mcpt_multi_create(const McptSceneDesc* desc, const int32_t* devices, int32_t n, uint32_t build_flags, McptMulti** out)
{
    if (out) *out = nullptr;
    McptMulti* m = nullptr;
    int rc = guarded([&] {
        if (!out || !devices || n < 1) throw Error(MCPT_ERR_INVALID, "mcpt_multi_create: null argument or no devices");
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); throw Error(MCPT_ERR_NO_DEVICE, "no CUDA device available: libmcpt has no CPU fallback"); }
        for (int r = 0; r < n; r++) {
            if (devices[r] < 0 || devices[r] >= ndev) throw Error(MCPT_ERR_INVALID, "device index out of range");
            for (int q = 0; q < r; q++) if (devices[q] == devices[r]) throw Error(MCPT_ERR_INVALID, "a device is listed twice");
        }
        m = new McptMulti();
        m->devices.assign(devices, devices + n);
        m->scenes.assign(n, nullptr); m->streams.assign(n, nullptr); m->films.assign(n, nullptr);
        // one BVH build; the other devices receive a copy of the packed blob
        McptScene* first = nullptr;
        int rc0 = mcpt_scene_create_ex(desc, devices[0], build_flags, &first);
        if (rc0) throw Error(rc0, mcpt_last_error());
        m->scenes[0] = first;
        std::vector<Slot> slots(n);
        per_device(n, slots, [&](int r) {
            MCPT_CUDA(cudaSetDevice(m->devices[r]));
            if (r > 0) m->scenes[r] = scene_clone(first, m->devices[r]);
            MCPT_CUDA(cudaStreamCreateWithFlags(&m->streams[r], cudaStreamNonBlocking));
        });
        MCPT_CUDA(cudaSetDevice(m->devices[0]));
        for (auto& e : m->ev) MCPT_CUDA(cudaEventCreate(&e));
        if (n > 1) {
            m->comms.assign(n, nullptr);
            MCPT_NCCL(nccl().CommInitAll(m->comms.data(), n, m->devices.data()));
        }
        *out = m;
    });
    if (rc != MCPT_OK && m) mcpt_multi_destroy(m);
    return rc;
}

int mcpt_multi_device_count(const McptMulti* m) { return m ? (int)m->devices.size() : 0; }
McptScene* mcpt_multi_scene(McptMulti* m, int32_t r) { return (m && r >= 0 && (size_t)r < m->scenes.size()) ? m->scenes[r] : nullptr; }

int mcpt_multi_upload(McptMulti* m)
{
    return guarded([&] {
        if (!m) throw Error(MCPT_ERR_INVALID, "null handle");
        std::vector<Slot> slots(m->devices.size());
        per_device((int)m->devices.size(), slots, [&](int r) {
            int rc = mcpt_scene_upload(m->scenes[r]);
            if (rc) throw Error(rc, mcpt_last_error());
        });
    });
}

int mcpt_multi_render(McptMulti* m, const McptCamera* cam, const McptRenderParams* params, float* out_rgb, McptStats* stats)
{
    return guarded([&] {
        if (!m || !cam || !params || !out_rgb) throw Error(MCPT_ERR_INVALID, "null argument");
        if (params->width <= 0 || params->height <= 0 || params->spp <= 0) throw Error(MCPT_ERR_INVALID, "width, height and spp must be positive");
        const int n = (int)m->devices.size();
        const size_t count = (size_t)params->width * params->height * 3, bytes = count * sizeof(float);
        const int spp_total = params->spp_total > 0 ? params->spp_total : params->spp;
        std::vector<Slot> slots(n);
        per_device(n, slots, [&](int r) {
            MCPT_CUDA(cudaSetDevice(m->devices[r]));
            if (m->film_bytes < bytes || !m->films[r]) {
                if (m->films[r]) MCPT_CUDA(cudaFree(m->films[r]));
                m->films[r] = nullptr;
                MCPT_CUDA(cudaMalloc((void**)&m->films[r], bytes));
            }
            // contiguous share of the sample range, remainder to the low ranks
            const int base = params->spp / n, rem = params->spp % n;
            const int cnt = base + (r < rem ? 1 : 0), off = r * base + std::min(r, rem);
            std::memset(&slots[r].st, 0, sizeof(McptStats));
            if (cnt == 0) { MCPT_CUDA(cudaMemsetAsync(m->films[r], 0, bytes, m->streams[r])); MCPT_CUDA(cudaStreamSynchronize(m->streams[r])); return; }
            McptRenderParams p = *params;
            p.spp = cnt; p.spp_offset = params->spp_offset + off; p.spp_total = spp_total;
            p.flags &= ~(uint32_t)MCPT_ACCUMULATE;
            render_device(m->scenes[r], cam, &p, m->films[r], m->streams[r], &slots[r].st);
        });
        m->film_bytes = std::max(m->film_bytes, bytes);
        MCPT_CUDA(cudaSetDevice(m->devices[0]));
        float ms_reduce = 0.f;
        if (n > 1) {
            // the path's one collective: sum of the partial films onto devices[0]
            MCPT_CUDA(cudaEventRecord(m->ev[0], m->streams[0]));
            MCPT_NCCL(nccl().GroupStart());
            for (int r = 0; r < n; r++) MCPT_NCCL(nccl().Reduce(m->films[r], m->films[r], count, ncclFloat32, ncclSum, 0, m->comms[r], m->streams[r]));
            MCPT_NCCL(nccl().GroupEnd());
            MCPT_CUDA(cudaSetDevice(m->devices[0]));
            MCPT_CUDA(cudaEventRecord(m->ev[1], m->streams[0]));
            for (int r = 0; r < n; r++) { MCPT_CUDA(cudaSetDevice(m->devices[r])); MCPT_CUDA(cudaStreamSynchronize(m->streams[r])); }
            MCPT_CUDA(cudaSetDevice(m->devices[0]));
            cudaEventElapsedTime(&ms_reduce, m->ev[0], m->ev[1]);
        }
        MCPT_CUDA(cudaMemcpyAsync(out_rgb, m->films[0], bytes, cudaMemcpyDeviceToHost, m->streams[0]));
        MCPT_CUDA(cudaStreamSynchronize(m->streams[0]));
        if (stats) {
            McptStats S;
            std::memset(&S, 0, sizeof S);
            for (int r = 0; r < n; r++) {
                const McptStats& s = slots[r].st;
                S.paths += s.paths; S.primary_hits += s.primary_hits; S.vertices += s.vertices; S.rays_closest += s.rays_closest;
                S.rays_shadow += s.rays_shadow; S.rays_light += s.rays_light; S.elements += s.elements; S.light_sweeps += s.light_sweeps; S.light_evals += s.light_evals;
                S.passes += s.passes; S.bounces += s.bounces; S.kernel_launches += s.kernel_launches;
                S.queue_peak = std::max(S.queue_peak, s.queue_peak);
                // device time of the call = the slowest device (they run side by side) + the reduce
                S.ms_total = std::max(S.ms_total, s.ms_total); S.ms_bounce = std::max(S.ms_bounce, s.ms_bounce);
                S.ms_shade = std::max(S.ms_shade, s.ms_shade); S.ms_trace = std::max(S.ms_trace, s.ms_trace);
                S.ms_sweep = std::max(S.ms_sweep, s.ms_sweep); S.ms_prep = std::max(S.ms_prep, s.ms_prep);
            }
            S.ms_total += ms_reduce;
            S.ms_other = S.ms_total - S.ms_bounce;
            *stats = S;
        }
    });
}

int mcpt_render_multi(const McptSceneDesc* desc, const int32_t* devices, int32_t n, const McptCamera* cam,
                      const McptRenderParams* params, float* out_rgb, McptStats* stats)
{
    McptMulti* m = nullptr;
    int rc = mcpt_multi_create(desc, devices, n, MCPT_BUILD_HOST_SAH, &m);
    if (rc) return rc;
    rc = mcpt_multi_render(m, cam, params, out_rgb, stats);
    std::string msg = rc ? mcpt_last_error() : "";
    mcpt_multi_destroy(m);
    if (rc) set_last_error(msg);
    return rc;
}

}  // extern "C"
