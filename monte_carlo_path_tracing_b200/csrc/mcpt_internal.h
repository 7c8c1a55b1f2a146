// csrc/mcpt_internal.h -- McptScene, error plumbing, shared between api.cu / batch.cu / render.cu.
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <vector>
#include "../../include/mcpt.h"
#include "bvh.h"
#include "host_scene.h"
#include "light_tree.h"
#include "scene_dev.cuh"

namespace mcpt {

void set_last_error(const std::string& msg);

#define MCPT_CUDA(expr)                                                                                      \
    do {                                                                                                     \
        cudaError_t e__ = (expr);                                                                            \
        if (e__ != cudaSuccess)                                                                              \
            throw ::mcpt::Error(MCPT_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__) + " (" +  \
                                                   __FILE__ + ":" + std::to_string(__LINE__) + ")");         \
    } while (0)

// Runs `body`, mapping exceptions to McptError codes (the C ABI never throws and never exits).
template <class F> int guarded(F&& body)
{
    try {
        body();
        return MCPT_OK;
    } catch (const Error& e) {
        set_last_error(e.what());
        return e.code;
    } catch (const std::bad_alloc&) {
        set_last_error("out of host memory");
        return MCPT_ERR_OOM;
    } catch (const std::exception& e) {
        set_last_error(e.what());
        return MCPT_ERR_INVALID;
    }
}

// Device buffers of the wavefront renderer, grown on demand and reused across calls.
struct Workspace {
    void* base = nullptr;
    size_t bytes = 0;
    void* pinned = nullptr;  // small pinned staging area for counters
    float* d_image = nullptr;  // W*H*3 output staging for mcpt_render
    size_t image_bytes = 0;
    size_t mem_budget = 0;  // free device memory + our own workspace at the last cudaMemGetInfo (render.cu)
};

}  // namespace mcpt

struct McptScene {
    int device = 0;
    uint32_t P = 0, S = 0, Mn = 0, L = 0, C = 0;
    // host copies of the description (the caller keeps ownership of its own arrays)
    std::vector<float> positions, normals;
    std::vector<int32_t> material_id, light_prim, light_class;
    std::vector<double> class_radiance;
    std::vector<double> geom_normals;  // [P][3]  get_unique_normal_of_facet
    std::vector<double> light_area;    // [L]
    mcpt::Bvh bvh, lbvh;
    mcpt::LightTree ltree;  // light hierarchy of MCPT_LIGHT_TREE
    uint32_t list_max = 0;  // longest padded list
    uint32_t nlists = 0;    // per-primitive survivor lists of the light sweeps (light_lists.h); 0 = not in use
    size_t list_entries = 0;
    double list_fraction = 0, list_build_ms = 0;
    double build_tree_ms = 0, build_total_ms = 0;  // scene BVH: tree construction / whole scene_create
    int32_t builder = 0;                           // MCPT_BUILD_*
    // one packed, pinned host image of everything the kernels read + its device twin
    void* h_blob = nullptr;
    void* d_blob = nullptr;
    size_t blob_bytes = 0;
    mcpt::DScene d{};      // device pointers into d_blob
    float extent = 1.f;    // largest scene dimension
    float calib_ms[2] = {0.f, 0.f};  // time of the layout calibration batch through the 4-wide / compressed 8-wide BVH (0: not run)
    float render_ms[2] = {0.f, 0.f}; // the same two kernels timed on the first large bounce of the first render (0: not yet)
    bool layout_free = false;        // the layout is the library's to choose (not forced by MCPT_BVH8, 8-wide tree usable)
    bool layout_final = false;       // settled on real rays
    mcpt::Workspace ws;
    cudaEvent_t ev[10] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
};

namespace mcpt {
// render.cu
void render_device(McptScene* scene, const McptCamera* cam, const McptRenderParams* params, float* d_out_rgb,
                   cudaStream_t stream, McptStats* stats);
// api.cu: replica of a scene on another device (no second BVH build)
McptScene* scene_clone(const McptScene* src, int device);
// render_parity.cu: MCPT_PARITY_FP64
void render_parity_device(McptScene* scene, const McptCamera* cam, const McptRenderParams* params, float* d_out_rgb,
                          cudaStream_t stream, McptStats* stats);
// render.cu: times a batch of rays through both wide layouts and sets d.use8
void calibrate_trace_layout(McptScene* scene);
// render.cu test hooks: the production trace kernels (k_extend / k_extend_exact, k_shadow / k_shadow_exact) on explicit rays
void trace_rays_device(McptScene* scene, int n, const float4* d_org, const float4* d_dir, int leaf_mode, float4* d_hit);
void visibility_device(McptScene* scene, int n, const float4* d_org, const float4* d_dir, const int2* d_ids, int leaf_mode, int* d_visible);
// render.cu: the same visibility test with its scratch in a fixed file-scope device array (no allocation; launches on `stream`, no sync)
void visibility_noalloc_device(McptScene* scene, size_t n, const float4* d_org, const float4* d_dir, const int2* d_ids, int leaf_mode,
                               int* d_visible, cudaStream_t stream);
void tonemap_device(McptScene* scene, const float* rgb, size_t n, double max_radiance, double gamma, uint8_t* out);
void workspace_free(Workspace& ws);
}  // namespace mcpt
