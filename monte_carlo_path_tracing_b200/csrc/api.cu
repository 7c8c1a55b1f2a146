// csrc/api.cu -- C ABI: scene lifetime, upload, introspection, host loaders, render entry points.
#include <chrono>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <memory>
#include <type_traits>
#include <numeric>
#include "bvh_tree.h"
#include "mcpt_internal.h"
#include "light_tree.h"
#include "light_lists.h"

namespace mcpt {
static thread_local std::string g_last_error;
void set_last_error(const std::string& msg) { g_last_error = msg; }

namespace {
struct BlobLayout {
    size_t off_nodes, off_nodes4, off_tris, off_nodes8, off_tris8, off_lnodes, off_ltris, off_pverts, off_pnorms, off_mats, off_lights, off_lsoa, off_lng64, off_larea64,
        off_crad64, off_class_begin, off_class_cdf, off_tri_cdf, off_ltree, off_ltree_leaf, off_llist, off_lrange, off_prim_list, total;
};
size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

void validate(const McptSceneDesc* d)
{
    if (!d) throw Error(MCPT_ERR_INVALID, "null scene description");
    if (d->num_prims == 0) throw Error(MCPT_ERR_INVALID, "scene has no triangles");
    if (!d->positions || !d->normals || !d->material_id || !d->mat_kd || !d->mat_ks || !d->mat_ns)
        throw Error(MCPT_ERR_INVALID, "scene description has null geometry/material arrays");
    if (d->num_lights && (!d->light_prim || !d->light_class || !d->class_radiance))
        throw Error(MCPT_ERR_INVALID, "scene description has null light arrays");
    for (uint32_t p = 0; p < d->num_prims; p++)
        if (d->material_id[p] < 0 || (uint32_t)d->material_id[p] >= d->num_materials)
            throw Error(MCPT_ERR_INVALID, "material_id out of range (the reference's materials.at() would throw, main.cpp:292)");
    int prev = 0;
    std::vector<int> per_class(d->num_classes, 0);
    for (uint32_t l = 0; l < d->num_lights; l++) {
        if (d->light_prim[l] < 0 || (uint32_t)d->light_prim[l] >= d->num_prims) throw Error(MCPT_ERR_INVALID, "light_prim out of range");
        int c = d->light_class[l];
        if (c < 0 || (uint32_t)c >= d->num_classes || c < prev) throw Error(MCPT_ERR_INVALID, "light_class must be non-decreasing and < num_classes");
        prev = c;
        per_class[c]++;
    }
    for (uint32_t c = 0; c < d->num_classes; c++)
        if (per_class[c] == 0)
            throw Error(MCPT_ERR_INVALID, "a <light> class has no triangles (lightsTriangles.at() throws in the reference, Mylight.cpp:126)");
}
}  // namespace
}  // namespace mcpt

using namespace mcpt;

extern "C" {

const char* mcpt_last_error(void) { return g_last_error.c_str(); }
int mcpt_version(void) { return MCPT_VERSION; }
int mcpt_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int mcpt_scene_upload(McptScene* s)
{
    return guarded([&] {
        if (!s) throw Error(MCPT_ERR_INVALID, "null scene");
        MCPT_CUDA(cudaSetDevice(s->device));
        MCPT_CUDA(cudaMemcpy(s->d_blob, s->h_blob, s->blob_bytes, cudaMemcpyHostToDevice));
    });
}

size_t mcpt_scene_device_bytes(const McptScene* s) { return s ? s->blob_bytes : 0; }

int mcpt_scene_create(const McptSceneDesc* desc, int device, McptScene** out)
{
    return mcpt_scene_create_ex(desc, device, MCPT_BUILD_HOST_SAH, out);
}

int mcpt_scene_trace_layout(const McptScene* s, int32_t* width, float calib_ms[2], float render_ms[2])
{
    return guarded([&] {
        if (!s) throw Error(MCPT_ERR_INVALID, "null scene");
        if (width) *width = s->d.use8 ? 8 : 4;
        if (calib_ms) { calib_ms[0] = s->calib_ms[0]; calib_ms[1] = s->calib_ms[1]; }
        if (render_ms) { render_ms[0] = s->render_ms[0]; render_ms[1] = s->render_ms[1]; }
    });
}

int mcpt_scene_light_lists(const McptScene* s, uint32_t* lists, uint64_t* entries, double* mean_fraction, double* build_ms)
{
    return guarded([&] {
        if (!s) throw Error(MCPT_ERR_INVALID, "null scene");
        if (lists) *lists = s->nlists;
        if (entries) *entries = s->list_entries;
        if (mean_fraction) *mean_fraction = s->list_fraction;
        if (build_ms) *build_ms = s->list_build_ms;
    });
}

int mcpt_scene_light_list_of_prim(const McptScene* s, uint32_t prim, int32_t* out_lights, int32_t* count)
{
    return guarded([&] {
        if (!s || !count) throw Error(MCPT_ERR_INVALID, "null argument");
        if (prim >= s->P) throw Error(MCPT_ERR_INVALID, "prim out of range");
        *count = -1;
        if (!s->nlists) return;
        const char* hb = (const char*)s->h_blob;
        const char* db = (const char*)s->d_blob;
        const int32_t id = ((const int32_t*)(hb + ((const char*)s->d.prim_list - db)))[prim];
        if (id < 0) return;
        const int32_t* rg = (const int32_t*)(hb + ((const char*)s->d.lrange - db)) + 2 * (size_t)id;
        const uint32_t* e = (const uint32_t*)(hb + ((const char*)s->d.llist - db)) + rg[0];
        int32_t n = 0;
        for (int32_t k = 0; k < rg[1]; k++) {
            const int32_t l = (int32_t)((e[k] / (uint32_t)kLsoaBlock) * 32 + (e[k] % (uint32_t)kLsoaBlock));  // lights.cuh lsoa_light
            if (l < (int32_t)s->L) { if (out_lights) out_lights[n] = l; n++; }
        }
        *count = n;
    });
}

int mcpt_scene_build_info(const McptScene* s, double* tree_ms, double* total_ms, int32_t* builder)
{
    return guarded([&] {
        if (!s) throw Error(MCPT_ERR_INVALID, "null scene");
        if (tree_ms) *tree_ms = s->build_tree_ms;
        if (total_ms) *total_ms = s->build_total_ms;
        if (builder) *builder = s->builder;
    });
}

int mcpt_scene_create_ex(const McptSceneDesc* desc, int device, uint32_t build_flags, McptScene** out)
{
    if (out) *out = nullptr;
    const auto t_begin = std::chrono::steady_clock::now();
    McptScene* s = nullptr;
    int rc = guarded([&] {
        if (!out) throw Error(MCPT_ERR_INVALID, "null output pointer");
        validate(desc);
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
            cudaGetLastError();
            throw Error(MCPT_ERR_NO_DEVICE, "no CUDA device available: libmcpt has no CPU fallback");
        }
        if (device < 0 || device >= ndev) throw Error(MCPT_ERR_INVALID, "device index out of range");
        MCPT_CUDA(cudaSetDevice(device));
        s = new McptScene();
        s->device = device;
        const uint32_t P = desc->num_prims, L = desc->num_lights, C = desc->num_classes, Mn = desc->num_materials;
        s->P = P; s->S = desc->num_shapes; s->Mn = Mn; s->L = L; s->C = C;
        s->positions.assign(desc->positions, desc->positions + (size_t)P * 9);
        s->normals.assign(desc->normals, desc->normals + (size_t)P * 9);
        s->material_id.assign(desc->material_id, desc->material_id + P);
        if (L) {
            s->light_prim.assign(desc->light_prim, desc->light_prim + L);
            s->light_class.assign(desc->light_class, desc->light_class + L);
            s->class_radiance.assign(desc->class_radiance, desc->class_radiance + (size_t)C * 3);
        }
        // geometric normals (Myobj.cpp:680-709) and light areas (Mylight.cpp:66-69), FP64
        s->geom_normals.resize((size_t)P * 3);
        for (uint32_t p = 0; p < P; p++) {
            dvec g = unique_normal(&s->positions[(size_t)p * 9], &s->normals[(size_t)p * 9]);
            for (int c = 0; c < 3; c++) s->geom_normals[(size_t)p * 3 + c] = g[c];
        }
        s->light_area.resize(L);
        for (uint32_t l = 0; l < L; l++) {
            const float* v = &s->positions[(size_t)s->light_prim[l] * 9];
            double e1[3], e2[3], n[3];
            for (int c = 0; c < 3; c++) { e1[c] = (double)v[3 + c] - (double)v[c]; e2[c] = (double)v[6 + c] - (double)v[c]; }
            n[0] = e1[1] * e2[2] - e1[2] * e2[1]; n[1] = e1[2] * e2[0] - e1[0] * e2[2]; n[2] = e1[0] * e2[1] - e1[1] * e2[0];
            double len = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
            double inv = 1.0 / len;
            double u[3] = {n[0] * inv, n[1] * inv, n[2] * inv};
            double ans = 0;
            for (int c = 0; c < 3; c++) ans += n[c] * u[c];
            s->light_area[l] = 0.5 * ans;
        }
        // BVHs
        std::vector<int32_t> all(P);
        std::iota(all.begin(), all.end(), 0);
        {
            const auto t0 = std::chrono::steady_clock::now();
            BvhTree T;
            bool gpu = false;
            if (build_flags & MCPT_BUILD_GPU_LBVH) {
                gpu = build_tree_gpu(s->positions.data(), P, all.data(), P, device, T, nullptr);
                s->build_tree_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
                if (gpu) {
                    emit_bvh(s->positions.data(), T, s->bvh);
                    if (s->bvh.max_depth4 * 3 + 1 > (uint32_t)kBvhStackDepth || s->bvh.max_depth + 1 > (uint32_t)kBvhStackDepth) {
                        gpu = false;  // too deep for the traversal stacks (long runs of equal Morton codes): SAH instead
                        s->bvh = Bvh();
                        T = BvhTree();
                    }
                }
            }
            if (!gpu) {
                build_tree_host(s->positions.data(), all.data(), P, T);
                s->build_tree_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
                emit_bvh(s->positions.data(), T, s->bvh);
            }
            s->builder = gpu ? MCPT_BUILD_GPU_LBVH : MCPT_BUILD_HOST_SAH;
        }
        std::vector<int32_t> lp(s->light_prim);
        std::sort(lp.begin(), lp.end());
        s->lbvh = build_bvh(s->positions.data(), lp.data(), lp.size());
        {   // light hierarchy for MCPT_LIGHT_TREE (light_tree.h)
            std::vector<double> lsum(L);
            for (uint32_t l = 0; l < L; l++) {
                const double* r = &s->class_radiance[(size_t)s->light_class[l] * 3];
                lsum[l] = r[0] + r[1] + r[2];
            }
            s->ltree = build_light_tree(s->positions.data(), s->light_prim.data(), s->geom_normals.data(), s->light_area.data(), lsum.data(), L);
        }
        if (s->bvh.max_depth4 * 3 + 1 > (uint32_t)kBvhStackDepth || s->bvh.max_depth + 1 > (uint32_t)kBvhStackDepth)
            throw Error(MCPT_ERR_INVALID, "scene BVH is too deep for the traversal stack (the builders fall back to median splits well before: corrupt input?)");
        if (s->lbvh.max_depth + 1 > (uint32_t)kBvhStackDepth)
            throw Error(MCPT_ERR_INVALID, "light BVH is too deep for the traversal stack");
        s->extent = std::max({s->bvh.hi[0] - s->bvh.lo[0], s->bvh.hi[1] - s->bvh.lo[1], s->bvh.hi[2] - s->bvh.lo[2], 1e-30f});

        // pack the blob
        BlobLayout B;
        size_t o = 0;
        auto place = [&](size_t bytes) { size_t at = o; o = align_up(o + std::max<size_t>(bytes, 16)); return at; };
        B.off_nodes = place(s->bvh.nodes.size() * sizeof(BvhNode));
        B.off_nodes4 = place(s->bvh.nodes4.size() * sizeof(f4));
        B.off_tris = place(s->bvh.tris.size() * sizeof(f4));
        B.off_nodes8 = place(s->bvh.nodes8.size() * sizeof(f4));
        B.off_tris8 = place(s->bvh.tris8.size() * sizeof(f4));
        B.off_lnodes = place(s->lbvh.nodes.size() * sizeof(BvhNode));
        B.off_ltris = place(s->lbvh.tris.size() * sizeof(f4));
        B.off_pverts = place((size_t)P * 3 * 16);
        B.off_pnorms = place((size_t)P * 3 * 16);
        B.off_mats = place((size_t)Mn * 2 * 16);
        B.off_lights = place((size_t)L * 5 * 16);
        std::vector<int32_t> light_of_prim(P, -1);
        for (uint32_t l = 0; l < L; l++) light_of_prim[s->light_prim[l]] = (int32_t)l;
        // per-primitive survivor lists of the light sweeps (light_lists.h); their padding entries point at a padding light of
        // the table, so a table that is full to the last block gets one more block when (and only when) lists are in use
        uint32_t Lpad = (L + 63) / 64 * 64;
        LightLists lists = build_light_lists(s->positions.data(), s->normals.data(), light_of_prim.data(), P, s->light_prim.data(),
                                             s->geom_normals.data(), s->light_area.data(), L, Lpad > L ? Lpad : Lpad + 64);
        if (lists.nlists && Lpad == L) Lpad += 64;
        s->nlists = lists.nlists; s->list_max = lists.max_len; s->list_entries = lists.entries.size(); s->list_fraction = lists.mean_fraction; s->list_build_ms = lists.build_ms;
        B.off_lsoa = place((size_t)Lpad * kLsoaPlanes * 16);
        B.off_lng64 = place((size_t)L * 3 * 8);
        B.off_larea64 = place((size_t)L * 8);
        B.off_crad64 = place((size_t)C * 3 * 8);
        B.off_class_begin = place((size_t)(C + 1) * 4);
        B.off_class_cdf = place((size_t)C * 4);
        B.off_tri_cdf = place((size_t)L * 4);
        B.off_ltree = place(s->ltree.nodes.size() * sizeof(f4));
        B.off_ltree_leaf = place((size_t)L * 4);
        B.off_llist = place(lists.entries.size() * 4);
        B.off_lrange = place(lists.range.size() * 4);
        B.off_prim_list = place(lists.nlists ? (size_t)P * 4 : 0);
        B.total = o;
        s->blob_bytes = B.total;
        MCPT_CUDA(cudaMallocHost(&s->h_blob, B.total));
        std::memset(s->h_blob, 0, B.total);
        char* hb = (char*)s->h_blob;
        std::memcpy(hb + B.off_nodes, s->bvh.nodes.data(), s->bvh.nodes.size() * sizeof(BvhNode));
        std::memcpy(hb + B.off_nodes4, s->bvh.nodes4.data(), s->bvh.nodes4.size() * sizeof(f4));
        std::memcpy(hb + B.off_tris, s->bvh.tris.data(), s->bvh.tris.size() * sizeof(f4));
        std::memcpy(hb + B.off_nodes8, s->bvh.nodes8.data(), s->bvh.nodes8.size() * sizeof(f4));
        std::memcpy(hb + B.off_tris8, s->bvh.tris8.data(), s->bvh.tris8.size() * sizeof(f4));
        std::memcpy(hb + B.off_lnodes, s->lbvh.nodes.data(), s->lbvh.nodes.size() * sizeof(BvhNode));
        std::memcpy(hb + B.off_ltris, s->lbvh.tris.data(), s->lbvh.tris.size() * sizeof(f4));
        if (L) {
            std::memcpy(hb + B.off_ltree, s->ltree.nodes.data(), s->ltree.nodes.size() * sizeof(f4));
            std::memcpy(hb + B.off_ltree_leaf, s->ltree.leaf_of_light.data(), (size_t)L * 4);
        }
        if (lists.nlists) {
            std::memcpy(hb + B.off_llist, lists.entries.data(), lists.entries.size() * 4);
            std::memcpy(hb + B.off_lrange, lists.range.data(), lists.range.size() * 4);
            std::memcpy(hb + B.off_prim_list, lists.prim_list.data(), (size_t)P * 4);
        }
        f4* pv = (f4*)(hb + B.off_pverts);
        f4* pn = (f4*)(hb + B.off_pnorms);
        for (uint32_t p = 0; p < P; p++) {
            const float* v = &s->positions[(size_t)p * 9];
            const float* n = &s->normals[(size_t)p * 9];
            for (int k = 0; k < 3; k++) {
                pv[3 * p + k] = f4{v[3 * k], v[3 * k + 1], v[3 * k + 2], 0.f};
                pn[3 * p + k] = f4{n[3 * k], n[3 * k + 1], n[3 * k + 2], 0.f};
            }
            std::memcpy(&pv[3 * p + 0].w, &s->material_id[p], 4);
            std::memcpy(&pv[3 * p + 1].w, &light_of_prim[p], 4);
        }
        f4* mt = (f4*)(hb + B.off_mats);
        for (uint32_t m = 0; m < Mn; m++) {
            mt[2 * m + 0] = f4{desc->mat_kd[3 * m], desc->mat_kd[3 * m + 1], desc->mat_kd[3 * m + 2], desc->mat_ns[m]};
            mt[2 * m + 1] = f4{desc->mat_ks[3 * m], desc->mat_ks[3 * m + 1], desc->mat_ks[3 * m + 2], 0.f};
        }
        f4* lt = (f4*)(hb + B.off_lights);
        double* lng = (double*)(hb + B.off_lng64);
        double* lar = (double*)(hb + B.off_larea64);
        double* crd = (double*)(hb + B.off_crad64);
        int32_t* cb = (int32_t*)(hb + B.off_class_begin);
        float* ccdf = (float*)(hb + B.off_class_cdf);
        float* tcdf = (float*)(hb + B.off_tri_cdf);
        for (uint32_t c = 0; c < C * 3; c++) crd[c] = s->class_radiance[c];
        f4* so = (f4*)(hb + B.off_lsoa);
        auto sidx = [](uint32_t l, int plane) { return (size_t)(l >> 5) * kLsoaBlock + (size_t)plane * 32 + (l & 31); };  // lights.cuh lsoa_index
        for (uint32_t l = L; l < Lpad; l++) so[sidx(l, 4)].w = std::numeric_limits<float>::infinity();  // padding: always culled
        for (uint32_t l = 0; l < L; l++) {
            int32_t prim = s->light_prim[l], cls = s->light_class[l];
            const float* v = &s->positions[(size_t)prim * 9];
            const double* g = &s->geom_normals[(size_t)prim * 3];
            const double* r = &s->class_radiance[(size_t)cls * 3];
            lt[5 * l + 0] = f4{v[0], v[1], v[2], 0.f};
            lt[5 * l + 1] = f4{v[3], v[4], v[5], 0.f};
            lt[5 * l + 2] = f4{v[6], v[7], v[8], (float)s->light_area[l]};
            lt[5 * l + 3] = f4{(float)g[0], (float)g[1], (float)g[2], (float)(r[0] + r[1] + r[2])};
            lt[5 * l + 4] = f4{(float)r[0], (float)r[1], (float)r[2], 0.f};
            std::memcpy(&lt[5 * l + 0].w, &prim, 4);
            std::memcpy(&lt[5 * l + 1].w, &cls, 4);
            const double a2 = 2.0 * s->light_area[l];
            so[sidx(l, 0)] = f4{v[0], v[1], v[2], 0.f};
            // edges from the first vertex and their three dot products, rounded once from FP64 (lights.cuh: light_weight_pair)
            double e1[3], e2[3], d11 = 0, d22 = 0, d12 = 0;
            for (int c = 0; c < 3; c++) {
                e1[c] = (double)(float)((double)v[3 + c] - (double)v[c]); e2[c] = (double)(float)((double)v[6 + c] - (double)v[c]);
                d11 += e1[c] * e1[c]; d22 += e2[c] * e2[c]; d12 += e1[c] * e2[c];
            }
            so[sidx(l, 1)] = f4{(float)e1[0], (float)e1[1], (float)e1[2], 0.f};
            so[sidx(l, 2)] = f4{(float)e2[0], (float)e2[1], (float)e2[2], 0.f};
            so[sidx(l, 4)] = f4{(float)d11, (float)d22, (float)d12, (float)(1e-8 * a2)};
            so[sidx(l, 3)] = f4{-(float)(g[0] * a2), -(float)(g[1] * a2), -(float)(g[2] * a2), 2.0f * (float)(r[0] + r[1] + r[2])};
            for (int c = 0; c < 3; c++) lng[3 * l + c] = g[c];
            lar[l] = s->light_area[l];
        }
        // class ranges + the two discrete distributions of Mylight.cpp:112-135 as cumulative tables
        for (uint32_t c = 0; c <= C; c++) cb[c] = 0;
        for (uint32_t l = 0; l < L; l++) cb[s->light_class[l] + 1] = (int32_t)l + 1;
        for (uint32_t c = 1; c <= C; c++) cb[c] = std::max(cb[c], cb[c - 1]);
        double csum = 0;
        for (uint32_t c = 0; c < C; c++) csum += crd[3 * c] + crd[3 * c + 1] + crd[3 * c + 2];
        double acc = 0;
        for (uint32_t c = 0; c < C; c++) {
            acc += (crd[3 * c] + crd[3 * c + 1] + crd[3 * c + 2]) / csum;
            ccdf[c] = (c + 1 == C) ? 1.0f : (float)acc;
            double asum = 0, a2 = 0;
            for (int32_t l = cb[c]; l < cb[c + 1]; l++) asum += s->light_area[l];
            for (int32_t l = cb[c]; l < cb[c + 1]; l++) {
                a2 += s->light_area[l] / asum;
                tcdf[l] = (l + 1 == cb[c + 1]) ? 1.0f : (float)a2;
            }
        }
        MCPT_CUDA(cudaMalloc(&s->d_blob, B.total));
        char* db = (char*)s->d_blob;
        DScene& d = s->d;
        d.nodes = (const float4*)(db + B.off_nodes); d.tris = (const float4*)(db + B.off_tris);
        d.nodes4 = (const float4*)(db + B.off_nodes4);
        d.nodes8 = (const float4*)(db + B.off_nodes8); d.tris8 = (const float4*)(db + B.off_tris8);
        d.use8 = 0;
        d.refill4 = kRefill4; d.refill8 = kRefill8;
        d.lnodes = (const float4*)(db + B.off_lnodes); d.ltris = (const float4*)(db + B.off_ltris);
        d.pverts = (const float4*)(db + B.off_pverts); d.pnorms = (const float4*)(db + B.off_pnorms);
        d.mats = (const float4*)(db + B.off_mats); d.lights = (const float4*)(db + B.off_lights);
        d.lsoa = (const float4*)(db + B.off_lsoa); d.Lpad = (int32_t)Lpad;
        d.lng64 = (const double*)(db + B.off_lng64); d.larea64 = (const double*)(db + B.off_larea64);
        d.crad64 = (const double*)(db + B.off_crad64); d.class_begin = (const int32_t*)(db + B.off_class_begin);
        d.class_cdf = (const float*)(db + B.off_class_cdf); d.tri_cdf = (const float*)(db + B.off_tri_cdf);
        d.ltree = (const float4*)(db + B.off_ltree); d.ltree_leaf = (const int32_t*)(db + B.off_ltree_leaf);
        d.llist = (const uint32_t*)(db + B.off_llist); d.lrange = (const int2*)(db + B.off_lrange);
        d.prim_list = (const int32_t*)(db + B.off_prim_list); d.nlists = (int32_t)lists.nlists; d.list_max = (int32_t)lists.max_len;
        d.P = (int32_t)P; d.L = (int32_t)L; d.C = (int32_t)C; d.M = (int32_t)Mn;
        d.tmin = s->extent * 1e-6f;
        MCPT_CUDA(cudaMemcpy(s->d_blob, s->h_blob, B.total, cudaMemcpyHostToDevice));
        for (int i = 0; i < 10; i++) MCPT_CUDA(cudaEventCreate(&s->ev[i]));
        // Which wide layout the trace kernels walk: measured on this scene (render.cu: calibrate_trace_layout) unless the
        // 8-wide tree is too deep for its stack or MCPT_BVH8=0/1 forces the choice (A/B measurements, tests).
        if (s->bvh.max_depth8 + 2 <= (uint32_t)kBvh8StackDepth && !s->bvh.nodes8.empty()) {
            if (const char* e = std::getenv("MCPT_BVH8")) d.use8 = std::atoi(e) != 0;
            else { calibrate_trace_layout(s); s->layout_free = true; }
        }
        s->build_total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
        *out = s;
    });
    if (rc != MCPT_OK && s) { mcpt_scene_destroy(s); }
    return rc;
}

}  // extern "C"

namespace mcpt {
// A replica of `src` on another device: the host-side description and the packed blob are copied, the device pointers
// rebased -- no second BVH build (multi.cu).
McptScene* scene_clone(const McptScene* src, int device)
{
    MCPT_CUDA(cudaSetDevice(device));
    std::unique_ptr<McptScene> s(new McptScene(*src));
    s->device = device;
    s->h_blob = nullptr; s->d_blob = nullptr;
    s->ws = Workspace();
    for (auto& e : s->ev) e = nullptr;
    McptScene* raw = s.release();
    try {
        MCPT_CUDA(cudaMallocHost(&raw->h_blob, src->blob_bytes));
        std::memcpy(raw->h_blob, src->h_blob, src->blob_bytes);
        MCPT_CUDA(cudaMalloc(&raw->d_blob, src->blob_bytes));
        MCPT_CUDA(cudaMemcpy(raw->d_blob, raw->h_blob, src->blob_bytes, cudaMemcpyHostToDevice));
        const char* ob = (const char*)src->d_blob;
        char* nb = (char*)raw->d_blob;
        auto rebase = [&](auto& ptr) { using T = std::remove_reference_t<decltype(ptr)>; ptr = (T)(nb + ((const char*)ptr - ob)); };
        DScene& d = raw->d;
        rebase(d.nodes); rebase(d.nodes4); rebase(d.tris); rebase(d.nodes8); rebase(d.tris8); rebase(d.lnodes); rebase(d.ltris); rebase(d.pverts); rebase(d.pnorms);
        rebase(d.mats); rebase(d.lights); rebase(d.lsoa); rebase(d.lng64); rebase(d.larea64); rebase(d.crad64);
        rebase(d.class_begin); rebase(d.class_cdf); rebase(d.tri_cdf); rebase(d.ltree); rebase(d.ltree_leaf);
        rebase(d.llist); rebase(d.lrange); rebase(d.prim_list);
        for (int i = 0; i < 10; i++) MCPT_CUDA(cudaEventCreate(&raw->ev[i]));
    } catch (...) {
        mcpt_scene_destroy(raw);
        throw;
    }
    return raw;
}
}  // namespace mcpt

extern "C" {

void mcpt_scene_destroy(McptScene* s)
{
    if (!s) return;
    cudaSetDevice(s->device);
    workspace_free(s->ws);
    if (s->d_blob) cudaFree(s->d_blob);
    if (s->h_blob) cudaFreeHost(s->h_blob);
    for (int i = 0; i < 10; i++)
        if (s->ev[i]) cudaEventDestroy(s->ev[i]);
    delete s;
}

int mcpt_scene_geom_normals(const McptScene* s, double* out)
{
    return guarded([&] {
        if (!s || !out) throw Error(MCPT_ERR_INVALID, "null argument");
        std::memcpy(out, s->geom_normals.data(), s->geom_normals.size() * sizeof(double));
    });
}
int mcpt_scene_light_areas(const McptScene* s, double* out)
{
    return guarded([&] {
        if (!s || !out) throw Error(MCPT_ERR_INVALID, "null argument");
        std::memcpy(out, s->light_area.data(), s->light_area.size() * sizeof(double));
    });
}
int mcpt_scene_bvh_info(const McptScene* s, int which, uint32_t* nodes, uint32_t* leaves, uint32_t* max_depth, double* sah)
{
    return guarded([&] {
        if (!s) throw Error(MCPT_ERR_INVALID, "null scene");
        const Bvh& b = which ? s->lbvh : s->bvh;
        if (nodes) *nodes = (uint32_t)b.nodes.size();
        if (leaves) *leaves = b.num_leaves;
        if (max_depth) *max_depth = b.max_depth;
        if (sah) *sah = b.sah_cost;
    });
}

// Structural check of a built BVH (both layouts): boxes nest, every listed prim sits in exactly one leaf with its own
// vertices inside the leaf box, depths fit the traversal stacks.  Returns the number of violations.
static uint32_t check_bvh(const mcpt::Bvh& b, const float* positions, uint32_t num_prims, const std::vector<int32_t>& ids)
{
    using namespace mcpt;
    uint32_t errors = 0;
    std::vector<uint32_t> seen(num_prims, 0);
    struct Item { int32_t code; float lo[3], hi[3]; };
    std::vector<Item> stack;
    const float inf = std::numeric_limits<float>::infinity();
    stack.push_back(Item{0, {-inf, -inf, -inf}, {inf, inf, inf}});
    while (!stack.empty()) {
        Item it = stack.back();
        stack.pop_back();
        if (it.code >= 0) {
            if ((size_t)it.code >= b.nodes.size()) { errors++; continue; }
            const BvhNode& n = b.nodes[it.code];
            Item c0{n.c0, {n.n0.x, n.n0.z, n.n2.x}, {n.n0.y, n.n0.w, n.n2.y}};
            Item c1{n.c1, {n.n1.x, n.n1.z, n.n2.z}, {n.n1.y, n.n1.w, n.n2.w}};
            for (Item* c : {&c0, &c1}) {
                if (c->lo[0] > c->hi[0]) continue;  // empty child
                for (int k = 0; k < 3; k++)
                    if (c->lo[k] < it.lo[k] - 1e-3f * std::fabs(it.lo[k]) - 1e-6f || c->hi[k] > it.hi[k] + 1e-3f * std::fabs(it.hi[k]) + 1e-6f) errors++;
                stack.push_back(*c);
            }
        } else {
            int code = ~it.code, first = code >> 3, cnt = (code & 7) + 1;
            for (int i = 0; i < cnt; i++) {
                if ((size_t)(first + i) * 3 + 2 >= b.tris.size()) { errors++; continue; }
                int32_t prim;
                std::memcpy(&prim, &b.tris[3 * (first + i)].w, 4);
                if (prim < 0 || (uint32_t)prim >= num_prims) { errors++; continue; }
                seen[prim]++;
                for (int v = 0; v < 3; v++) {
                    const f4& p = b.tris[3 * (first + i) + v];
                    const float* q = positions + (size_t)prim * 9 + 3 * v;
                    if (p.x != q[0] || p.y != q[1] || p.z != q[2]) errors++;
                    float c[3] = {p.x, p.y, p.z};
                    for (int k = 0; k < 3; k++)
                        if (c[k] < it.lo[k] || c[k] > it.hi[k]) errors++;
                }
            }
        }
    }
    for (int32_t id : ids) if (seen[id] != 1) errors++;
    // the 4-wide collapse of the same tree (bvh.h, nodes4): same leaves, same containment
    {
        std::vector<uint32_t> seen4(num_prims, 0);
        stack.push_back(Item{0, {-inf, -inf, -inf}, {inf, inf, inf}});
        while (!stack.empty()) {
            Item it = stack.back();
            stack.pop_back();
            if (it.code >= 0) {
                if ((size_t)it.code * 7 + 6 >= b.nodes4.size()) { errors++; continue; }
                const f4* q = &b.nodes4[(size_t)it.code * 7];
                for (int k = 0; k < 4; k++) {
                    auto comp = [k](const f4& v) { return k == 0 ? v.x : k == 1 ? v.y : k == 2 ? v.z : v.w; };
                    Item c;
                    float cf = comp(q[6]);
                    std::memcpy(&c.code, &cf, 4);
                    for (int a = 0; a < 3; a++) { c.lo[a] = comp(q[2 * a]); c.hi[a] = comp(q[2 * a + 1]); }
                    if (c.lo[0] > c.hi[0]) continue;  // unused slot
                    for (int a = 0; a < 3; a++)
                        if (c.lo[a] < it.lo[a] - 1e-3f * std::fabs(it.lo[a]) - 1e-6f || c.hi[a] > it.hi[a] + 1e-3f * std::fabs(it.hi[a]) + 1e-6f) errors++;
                    stack.push_back(c);
                }
            } else {
                int code = ~it.code, first = code >> 3, cnt = (code & 7) + 1;
                for (int i = 0; i < cnt; i++) {
                    if ((size_t)(first + i) * 3 + 2 >= b.tris.size()) { errors++; continue; }
                    int32_t prim;
                    std::memcpy(&prim, &b.tris[3 * (first + i)].w, 4);
                    if (prim < 0 || (uint32_t)prim >= num_prims) { errors++; continue; }
                    seen4[prim]++;
                    for (int v = 0; v < 3; v++) {
                        const f4& p = b.tris[3 * (first + i) + v];
                        float c[3] = {p.x, p.y, p.z};
                        for (int a = 0; a < 3; a++)
                            if (c[a] < it.lo[a] || c[a] > it.hi[a]) errors++;
                    }
                }
            }
        }
        for (int32_t id : ids) if (seen4[id] != 1) errors++;
        if (b.max_depth4 > b.max_depth + 1 || b.max_depth4 * 3 + 1 > (uint32_t)kBvhStackDepth) errors++;
    }
    // the compressed 8-wide collapse (bvh.h, nodes8 / tris8): every prim in exactly one slot, every triangle inside its slot's
    // DEQUANTISED box (p + q * 2^e evaluated in float, as the traversal does), every inner child's node box inside its slot's box
    if (!ids.empty()) {
        uint32_t cat[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        const uint32_t e_before = errors;
        std::vector<uint32_t> seen8(num_prims, 0);
        struct Item8 { uint32_t node; float lo[3], hi[3]; uint32_t depth; };
        std::vector<Item8> st8;
        st8.push_back(Item8{0, {-inf, -inf, -inf}, {inf, inf, inf}, 1});
        uint32_t depth8 = 0;
        while (!st8.empty()) {
            const Item8 it = st8.back();
            st8.pop_back();
            depth8 = std::max(depth8, it.depth);
            if ((size_t)it.node * 5 + 4 >= b.nodes8.size()) { errors++; continue; }
            const f4* N = &b.nodes8[(size_t)it.node * 5];
            uint32_t w[20];
            std::memcpy(w, N, 80);
            const float p[3] = {N[0].x, N[0].y, N[0].z};
            float scale[3];
            for (int a = 0; a < 3; a++) { const uint32_t e = ((w[3] >> (8 * a)) & 255u) << 23; std::memcpy(&scale[a], &e, 4); }
            const uint32_t imask = w[3] >> 24, child_base = w[4], tri_base = w[5];
            const uint8_t* meta = (const uint8_t*)&w[6];
            const uint8_t* q = (const uint8_t*)&w[8];  // qlo.x[8] qlo.y[8] qlo.z[8] qhi.x[8] qhi.y[8] qhi.z[8]
            for (int a = 0; a < 3; a++)
                if (p[a] < it.lo[a] - 1e-3f * std::fabs(it.lo[a]) - 1e-6f) { errors++; cat[0]++; }
            uint32_t inner_seen = 0;
            for (int sl = 0; sl < 8; sl++) {
                const uint32_t m = meta[sl];
                if (m == 0) { if ((imask >> sl) & 1u) errors++; continue; }
                float lo[3], hi[3];
                for (int a = 0; a < 3; a++) { lo[a] = p[a] + (float)q[8 * a + sl] * scale[a]; hi[a] = p[a] + (float)q[24 + 8 * a + sl] * scale[a]; }
                for (int a = 0; a < 3; a++)
                    // a slot box may overhang the box its node got in the parent by one quantisation step (outward rounding on
                    // a finer grid); what the traversal relies on is checked below: triangles lie inside EVERY ancestor's box
                    if (lo[a] > hi[a] || hi[a] > it.hi[a] + 1.01f * scale[a] + 1e-3f * std::fabs(it.hi[a]) + 1e-6f) { errors++; cat[1]++; }
                const bool inner = (m & 0x18u) == 0x18u;
                if (inner != (((imask >> sl) & 1u) != 0)) errors++;
                if (inner) {
                    if ((m >> 5) != 1u || (m & 7u) != (uint32_t)sl) errors++;
                    Item8 c;
                    c.node = child_base + inner_seen++; c.depth = it.depth + 1;
                    for (int a = 0; a < 3; a++) { c.lo[a] = std::max(lo[a], it.lo[a]); c.hi[a] = std::min(hi[a], it.hi[a]); }  // intersection of the ancestors' boxes
                    st8.push_back(c);
                } else {
                    const uint32_t bits = m >> 5, off = m & 31u;
                    const uint32_t cnt = bits == 1 ? 1 : bits == 3 ? 2 : bits == 7 ? 3 : 0;
                    if (cnt == 0 || off + cnt > 24) { errors++; continue; }
                    for (uint32_t i = 0; i < cnt; i++) {
                        const size_t t = (size_t)tri_base + off + i;
                        if (t * 3 + 2 >= b.tris8.size()) { errors++; continue; }
                        int32_t prim;
                        std::memcpy(&prim, &b.tris8[3 * t].w, 4);
                        if (prim < 0 || (uint32_t)prim >= num_prims) { errors++; continue; }
                        seen8[prim]++;
                        for (int v = 0; v < 3; v++) {
                            const f4& pt = b.tris8[3 * t + v];
                            const float* src = positions + (size_t)prim * 9 + 3 * v;
                            if (pt.x != src[0] || pt.y != src[1] || pt.z != src[2]) errors++;
                            const float c[3] = {pt.x, pt.y, pt.z};
                            for (int a = 0; a < 3; a++)
                                if (c[a] < std::max(lo[a], it.lo[a]) || c[a] > std::min(hi[a], it.hi[a])) { errors++; cat[2]++; }
                        }
                    }
                }
            }
        }
        for (int32_t id : ids) if (seen8[id] != 1) { errors++; cat[3]++; }
        if (depth8 != b.max_depth8 || b.max_depth8 + 2 > (uint32_t)kBvh8StackDepth) { errors++; cat[4]++; }
        if (b.tris8.size() != b.tris.size()) { errors++; cat[5]++; }
        if (errors != e_before && std::getenv("MCPT_DEBUG_BVH8"))
            std::fprintf(stderr, "nodes8 check: origin %u, slot box %u, triangle outside %u, seen %u, depth %u (%u vs %u), size %u\n", cat[0], cat[1], cat[2], cat[3], cat[4], depth8, b.max_depth8, cat[5]);
    }
    uint32_t total = 0;
    for (uint32_t v : seen) total += v;
    if (total != ids.size()) errors++;
    return errors;
}

int mcpt_host_bvh_check(const McptSceneDesc* desc, int lights_only, uint32_t out4[4], double* sah_cost)
{
    return guarded([&] {
        validate(desc);
        std::vector<int32_t> ids;
        if (lights_only) { ids.assign(desc->light_prim, desc->light_prim + desc->num_lights); std::sort(ids.begin(), ids.end()); }
        else { ids.resize(desc->num_prims); std::iota(ids.begin(), ids.end(), 0); }
        Bvh b = build_bvh(desc->positions, ids.data(), ids.size());
        const uint32_t errors = check_bvh(b, desc->positions, desc->num_prims, ids);
        if (out4) { out4[0] = (uint32_t)b.nodes.size(); out4[1] = b.num_leaves; out4[2] = b.max_depth; out4[3] = errors; }
        if (sah_cost) *sah_cost = b.sah_cost;
    });
}

// Host-only: build the light hierarchy of MCPT_LIGHT_TREE for `desc` and validate it (light_tree.h).
// out4 = {nodes, max_depth, errors, lights}.
int mcpt_host_light_tree_check(const McptSceneDesc* desc, uint32_t out4[4])
{
    return guarded([&] {
        validate(desc);
        if (!out4) throw Error(MCPT_ERR_INVALID, "null argument");
        const uint32_t L = desc->num_lights;
        std::vector<double> gn((size_t)desc->num_prims * 3, 0.0), area(L), lsum(L);
        for (uint32_t l = 0; l < L; l++) {
            const int32_t p = desc->light_prim[l];
            const dvec g = unique_normal(desc->positions + (size_t)p * 9, desc->normals + (size_t)p * 9);
            for (int c = 0; c < 3; c++) gn[(size_t)p * 3 + c] = g[c];
            const float* v = desc->positions + (size_t)p * 9;
            double e1[3], e2[3];
            for (int c = 0; c < 3; c++) { e1[c] = (double)v[3 + c] - v[c]; e2[c] = (double)v[6 + c] - v[c]; }
            const double n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
            area[l] = 0.5 * std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
            const double* r = desc->class_radiance + (size_t)desc->light_class[l] * 3;
            lsum[l] = r[0] + r[1] + r[2];
        }
        const LightTree T = build_light_tree(desc->positions, desc->light_prim, gn.data(), area.data(), lsum.data(), L);
        out4[0] = T.num_nodes; out4[1] = T.max_depth; out4[2] = check_light_tree(T, desc->positions, desc->light_prim, gn.data(), L); out4[3] = L;
    });
}

int mcpt_host_light_lists(const McptSceneDesc* desc, uint8_t* mask, uint32_t out3[3])
{
    return guarded([&] {
        validate(desc);
        if (!mask || !out3) throw Error(MCPT_ERR_INVALID, "null argument");
        const uint32_t P = desc->num_prims, L = desc->num_lights;
        std::vector<double> gn((size_t)P * 3, 0.0), area(L);
        std::vector<int32_t> light_of_prim(P, -1);
        for (uint32_t l = 0; l < L; l++) {
            const int32_t p = desc->light_prim[l];
            light_of_prim[p] = (int32_t)l;
            const dvec g = unique_normal(desc->positions + (size_t)p * 9, desc->normals + (size_t)p * 9);
            for (int c = 0; c < 3; c++) gn[(size_t)p * 3 + c] = g[c];
            const float* v = desc->positions + (size_t)p * 9;
            double e1[3], e2[3];
            for (int c = 0; c < 3; c++) { e1[c] = (double)v[3 + c] - v[c]; e2[c] = (double)v[6 + c] - v[c]; }
            const double n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
            area[l] = 0.5 * std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
        }
        const uint32_t Lpad = (L + 63) / 64 * 64;
        const LightLists R = build_light_lists(desc->positions, desc->normals, light_of_prim.data(), P, desc->light_prim, gn.data(), area.data(), L,
                                               Lpad > L ? Lpad : Lpad + 64);
        out3[0] = R.nlists; out3[1] = R.max_len; out3[2] = (uint32_t)(1000.0 * R.mean_fraction + 0.5);
        if (!R.nlists) return;
        std::memset(mask, 0, (size_t)P * L);
        for (uint32_t p = 0; p < P; p++) {
            const int32_t id = R.prim_list[p];
            if (id < 0) continue;
            for (int32_t k = 0; k < R.range[2 * id + 1]; k++) {
                const uint32_t e = R.entries[(size_t)R.range[2 * id] + k], l = ((e >> 7) << 5) + (e & 31);
                if (l < L) mask[(size_t)p * L + l] = 1;
            }
        }
    });
}

int mcpt_scene_light_tree_info(const McptScene* s, uint32_t* nodes, uint32_t* max_depth)
{
    return guarded([&] {
        if (!s) throw Error(MCPT_ERR_INVALID, "null scene");
        if (nodes) *nodes = s->ltree.num_nodes;
        if (max_depth) *max_depth = s->ltree.max_depth;
    });
}
int mcpt_scene_light_tree_nodes(const McptScene* s, float* out_nodes12, int32_t* out_leaf)
{
    return guarded([&] {
        if (!s || !out_nodes12 || !out_leaf) throw Error(MCPT_ERR_INVALID, "null argument");
        std::memcpy(out_nodes12, s->ltree.nodes.data(), s->ltree.nodes.size() * sizeof(f4));
        std::memcpy(out_leaf, s->ltree.leaf_of_light.data(), s->ltree.leaf_of_light.size() * sizeof(int32_t));
    });
}

// the same check on the scene BVH a McptScene actually holds (whichever builder made it)
int mcpt_scene_bvh_check(const McptScene* s, uint32_t* errors)
{
    return guarded([&] {
        if (!s || !errors) throw Error(MCPT_ERR_INVALID, "null argument");
        std::vector<int32_t> ids(s->P);
        std::iota(ids.begin(), ids.end(), 0);
        *errors = check_bvh(s->bvh, s->positions.data(), s->P, ids);
    });
}

// ---- render ------------------------------------------------------------------------------------------
int mcpt_render_device(McptScene* scene, const McptCamera* cam, const McptRenderParams* params, float* d_out, void* stream,
                       McptStats* stats)
{
    return guarded([&] {
        if (!scene || !cam || !params || !d_out) throw Error(MCPT_ERR_INVALID, "null argument");
        MCPT_CUDA(cudaSetDevice(scene->device));
        render_device(scene, cam, params, d_out, (cudaStream_t)stream, stats);
    });
}

int mcpt_render(McptScene* scene, const McptCamera* cam, const McptRenderParams* params, float* out_rgb, McptStats* stats)
{
    return guarded([&] {
        if (!scene || !cam || !params || !out_rgb) throw Error(MCPT_ERR_INVALID, "null argument");
        if (params->width <= 0 || params->height <= 0) throw Error(MCPT_ERR_INVALID, "bad resolution");
        MCPT_CUDA(cudaSetDevice(scene->device));
        size_t bytes = (size_t)params->width * params->height * 3 * sizeof(float);
        Workspace& ws = scene->ws;
        if (ws.image_bytes < bytes) {
            if (ws.d_image) MCPT_CUDA(cudaFree(ws.d_image));
            ws.d_image = nullptr; ws.image_bytes = 0;
            MCPT_CUDA(cudaMalloc((void**)&ws.d_image, bytes));
            ws.image_bytes = bytes;
        }
        McptRenderParams p = *params;
        p.flags &= ~(uint32_t)MCPT_ACCUMULATE;
        render_device(scene, cam, &p, ws.d_image, nullptr, stats);
        MCPT_CUDA(cudaMemcpy(out_rgb, ws.d_image, bytes, cudaMemcpyDeviceToHost));
    });
}

int mcpt_tonemap(McptScene* scene, const float* rgb, size_t n, double max_radiance, double gamma, uint8_t* out)
{
    return guarded([&] {
        if (!scene || !rgb || !out) throw Error(MCPT_ERR_INVALID, "null argument");
        MCPT_CUDA(cudaSetDevice(scene->device));
        tonemap_device(scene, rgb, n, max_radiance, gamma, out);
    });
}

// ---- host loaders --------------------------------------------------------------------------------------
struct McptHostScene {
    FlatScene flat;
    McptSceneDesc desc;
    CameraDesc cam;
};

int mcpt_host_scene_load(const char* obj_path, const char* xml_path, McptHostScene** out)
{
    if (out) *out = nullptr;
    return guarded([&] {
        if (!obj_path || !xml_path || !out) throw Error(MCPT_ERR_INVALID, "null argument");
        Myobj obj(obj_path);
        Mylight lights(xml_path);
        obj.read();                           // main.cpp:500
        lights.read();                        // main.cpp:503
        lights.gather_light_triangles(obj);   // main.cpp:504
        McptHostScene* h = new McptHostScene();
        h->flat = flatten(obj, lights);
        h->desc = h->flat.desc();
        h->cam = lights.camera;
        *out = h;
    });
}
const McptSceneDesc* mcpt_host_scene_desc(const McptHostScene* h) { return h ? &h->desc : nullptr; }
int mcpt_host_scene_camera(const McptHostScene* h, McptCamera* cam, int32_t* width, int32_t* height, double* fovy)
{
    if (!h || !h->cam.present) return 0;
    const CameraDesc& c = h->cam;
    if (cam) {
        double wl = 0;
        for (int k = 0; k < 3; k++) {
            cam->eye[k] = c.eye[k]; cam->w[k] = c.lookat[k] - c.eye[k]; cam->up[k] = c.up[k];
            wl += cam->w[k] * cam->w[k];
        }
        cam->pixellen = std::tan(c.fovy * (M_PI / 180.0) * 0.5) * std::sqrt(wl) / (c.height * 0.5);
    }
    if (width) *width = c.width;
    if (height) *height = c.height;
    if (fovy) *fovy = c.fovy;
    return 1;
}
void mcpt_host_scene_free(McptHostScene* h) { delete h; }

}  // extern "C"
