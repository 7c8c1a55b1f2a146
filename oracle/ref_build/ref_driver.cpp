// oracle/ref_build/ref_driver.cpp -- TEST INFRASTRUCTURE.
//
// Headless C-ABI driver around the UNMODIFIED reference (compiled from /root/reference by
// oracle/Makefile into oracle/_ref/libmcpt_ref.so).  It replaces only main() (main.cpp:497-599):
// scene path, camera, resolution, spp, estimator and row range become arguments, the EasyX
// window is gone, and the clock-seeded engines are replaced by oracle/stream.h (rng_shim.h).
// Everything else -- Myobj, Mylight, BRDF, shade / shade_with_brdf / shade_with_mis and the
// globals `veach` / `lights` they read -- is the reference's own object code.
//
// Used by tests/ (to pin the C restatement and to generate tests/golden/) and by bench.py's
// reference arm / cpu_baseline.  Never linked into the product.
#include "Myobj.h"
#include "Mylight.h"
#include "BRDF.h"
#include "matrix3d.h"
#include <array>
#include <cstdint>
#include <iostream>
#include <sstream>
#include <vector>

// main.cpp:23-24, 269, 348, 402
extern Myobj veach;
extern Mylight lights;
RadianceRGB shade(intersec_result point, vec wo);
RadianceRGB shade_with_brdf(intersec_result point, vec wo);
RadianceRGB shade_with_mis(intersec_result point, vec wo);

oracle_stream g_oracle_stream;
int g_oracle_depth = 0;
int g_oracle_max_depth = 0;
uint64_t g_oracle_vertices = 0;

namespace {
std::vector<size_t> g_shape_offset;  // prim = shape_offset[s] + f
std::vector<std::string> g_class_names;

struct Quiet {  // the loaders print banners to std::cout (Myobj.cpp:12,26-27, Mylight.cpp:13,27-29)
    std::streambuf* old;
    std::ostringstream sink;
    Quiet() : old(std::cout.rdbuf(sink.rdbuf())) {}
    ~Quiet() { std::cout.rdbuf(old); }
};

inline triangle prim_to_tri(long long prim)
{
    if (prim < 0) return triangle((size_t)-1, (size_t)-1);  // main.cpp:572 triangle(-1,-1)
    size_t s = 0;
    while (s + 1 < g_shape_offset.size() - 1 && (size_t)prim >= g_shape_offset[s + 1]) s++;
    return triangle(s, (size_t)prim - g_shape_offset[s]);
}
inline long long tri_to_prim(size_t s, size_t f) { return (long long)(g_shape_offset[s] + f); }
inline vec V3(const double* p) { return vec(p[0], p[1], p[2]); }
}  // namespace

extern "C" {

// Myobj::read + Mylight::read + gather_light_triangles (main.cpp:500-504).  One scene per process:
// the reference's caches and grid are never cleared (Myobj.h:76-90).
int ref_load(const char* obj_path, const char* xml_path)
{
    Quiet q;
    veach.inputfile = obj_path;
    lights.inputfile = xml_path;
    veach.read();
    lights.read();
    lights.gather_light_triangles(veach.reader);
    const auto& shapes = veach.reader.GetShapes();
    g_shape_offset.assign(1, 0);
    for (size_t s = 0; s < shapes.size(); s++)
        g_shape_offset.push_back(g_shape_offset.back() + shapes[s].mesh.num_face_vertices.size());
    g_class_names.clear();
    for (auto& kv : lights.lightsRadiance) g_class_names.push_back(kv.first);
    return 0;
}

// cal_scene_boundingbox(start) + meshing(n0) (main.cpp:519-520)
void ref_build_grid(const double eye[3], int n0)
{
    veach.cal_scene_boundingbox(V3(eye));
    veach.meshing(n0);
}

void ref_counts(int* P, int* S, int* M, int* L, int* C)
{
    *S = (int)g_shape_offset.size() - 1;
    *P = (int)g_shape_offset.back();
    *M = (int)veach.reader.GetMaterials().size();
    int l = 0;
    for (auto& kv : lights.lightsTriangles) l += (int)kv.second.size();
    *L = l;
    *C = (int)lights.lightsRadiance.size();
}

// Flat view of what the reference loaded, in the layout of McptSceneDesc (include/mcpt.h):
// light triangles in the reference's own iteration order (std::map by mtlname, then gather order).
void ref_flatten(float* pos, float* nor, int* matid, unsigned* shape_offset, float* kd, float* ks, float* ns,
                 int* light_prim, int* light_class, double* class_rad, double* geom_normal, double* light_area)
{
    const auto& shapes = veach.reader.GetShapes();
    const auto& mats = veach.reader.GetMaterials();
    size_t p = 0;
    for (size_t s = 0; s < shapes.size(); s++) {
        shape_offset[s] = (unsigned)g_shape_offset[s];
        for (size_t f = 0; f < shapes[s].mesh.num_face_vertices.size(); f++, p++) {
            auto v = veach.get_vertexes_of_facet(s, f);
            auto n = veach.get_normals_of_facet(s, f);
            for (int k = 0; k < 3; k++)
                for (int c = 0; c < 3; c++) {
                    pos[p * 9 + k * 3 + c] = (float)v[k].xyz[c];
                    nor[p * 9 + k * 3 + c] = (float)n[k].xyz[c];
                }
            matid[p] = shapes[s].mesh.material_ids[f];
            if (geom_normal) {
                vec g = veach.get_unique_normal_of_facet(s, f);
                for (int c = 0; c < 3; c++) geom_normal[p * 3 + c] = g.xyz[c];
            }
        }
    }
    shape_offset[shapes.size()] = (unsigned)g_shape_offset.back();
    for (size_t m = 0; m < mats.size(); m++) {
        for (int c = 0; c < 3; c++) {
            kd[m * 3 + c] = mats[m].diffuse[c];
            ks[m * 3 + c] = mats[m].specular[c];
        }
        ns[m] = mats[m].shininess;
    }
    int ci = 0, li = 0;
    for (auto& kv : lights.lightsRadiance) {
        for (int c = 0; c < 3; c++) class_rad[ci * 3 + c] = kv.second.RGB[c];
        auto it = lights.lightsTriangles.find(kv.first);
        if (it != lights.lightsTriangles.end())
            for (auto& lt : it->second) {
                light_prim[li] = (int)tri_to_prim(lt.s, lt.f);
                light_class[li] = ci;
                if (light_area) light_area[li] = lt.area;
                li++;
            }
        ci++;
    }
}

void ref_set_max_depth(int d) { g_oracle_max_depth = d; }
uint64_t ref_vertices(void) { return g_oracle_vertices; }
uint64_t ref_draws(void) { return g_oracle_stream.draws; }

// ---- deterministic probes ------------------------------------------------------------------
// Myobj::intersect_with_triangle (Myobj.cpp:165-192)
void ref_intersect_tri(int n, const double* ro, const double* rd, const int* prim, int* hit, double* bgt)
{
    for (int i = 0; i < n; i++) {
        triangle t = prim_to_tri(prim[i]);
        intersec_result r = veach.intersect_with_triangle(V3(ro + 3 * i), V3(rd + 3 * i), t.s, t.f);
        hit[i] = r.isIntersec;
        bgt[3 * i] = r.beta; bgt[3 * i + 1] = r.gamma; bgt[3 * i + 2] = r.t;
    }
}

// mode 0: Myobj::closet_ray_intersect (grid DDA, Myobj.cpp:334-474)
// mode 1: closet_ray_intersect_light_triangle (Myobj.cpp:476-622)
// mode 2: brute force over intersect_with_triangle == the commented-out definition Myobj.cpp:194-213
//         (strict <, (s,f) ascending) with the exclude-id rule of Myobj.cpp:417
// mode 3: brute force over light triangles only
void ref_closest(int n, const double* ro, const double* rd, const int* exclude, int mode, int* prim, double* bgt)
{
    const auto& shapes = veach.reader.GetShapes();
    for (int i = 0; i < n; i++) {
        triangle ex = prim_to_tri(exclude ? exclude[i] : -1);
        vec o = V3(ro + 3 * i), d = V3(rd + 3 * i);
        intersec_result r;
        if (mode == 0) r = veach.closet_ray_intersect(o, d, ex);
        else if (mode == 1) r = veach.closet_ray_intersect_light_triangle(o, d, ex, lights.islight);
        else {
            intersec_result ans(false, 0, 0, 0, 0, DBL_MAX);
            for (size_t s = 0; s < shapes.size(); s++)
                for (size_t f = 0; f < shapes[s].mesh.num_face_vertices.size(); f++) {
                    triangle t(s, f);
                    if (t == ex) continue;
                    if (mode == 3 && lights.islight.find(t) == lights.islight.end()) continue;
                    intersec_result rs = veach.intersect_with_triangle(o, d, s, f);
                    if (rs.isIntersec && rs.t < ans.t) ans = rs;
                }
            r = ans;
        }
        prim[i] = r.isIntersec ? (int)tri_to_prim(r.s, r.f) : -1;
        // misses report zeros (the reference leaves t = DBL_MAX in `ans`, Myobj.cpp:397)
        bgt[3 * i] = r.isIntersec ? r.beta : 0; bgt[3 * i + 1] = r.isIntersec ? r.gamma : 0; bgt[3 * i + 2] = r.isIntersec ? r.t : 0;
    }
}

// BRDF::get_brdf_phong (BRDF.cpp:17-25)
void ref_brdf_eval(int n, const double* N, const double* wi, const double* wr, const double* kd, const double* ks,
                   const double* ns, double* rgb)
{
    for (int i = 0; i < n; i++) {
        BRDF b = BRDF::get_brdf_phong(V3(N + 3 * i), V3(wi + 3 * i), V3(wr + 3 * i), V3(kd + 3 * i), V3(ks + 3 * i), ns[i]);
        for (int c = 0; c < 3; c++) rgb[3 * i + c] = b.RGB[c];
    }
}

// BRDF::sample_from_phong (BRDF.cpp:28-104); u = (u_lobe, xi1, xi2) per call
void ref_brdf_sample(int n, const double* N, const double* wr, const double* kd, const double* ks, const double* ns,
                     const double* u, double* dir, double* pdf)
{
    for (int i = 0; i < n; i++) {
        oracle_stream_list(&g_oracle_stream, u + 3 * i, 3);
        sampledRay r = BRDF::sample_from_phong(V3(N + 3 * i), V3(wr + 3 * i), V3(kd + 3 * i), V3(ks + 3 * i), ns[i]);
        for (int c = 0; c < 3; c++) dir[3 * i + c] = r.dir.xyz[c];
        pdf[i] = r.pdf;
    }
}

// BRDF::eval_sample_from_phong_pdf (BRDF.cpp:107-132)
void ref_brdf_pdf(int n, const double* N, const double* wi, const double* wr, const double* kd, const double* ks,
                  const double* ns, double* pdf)
{
    for (int i = 0; i < n; i++)
        pdf[i] = BRDF::eval_sample_from_phong_pdf(V3(N + 3 * i), V3(wi + 3 * i), V3(wr + 3 * i), V3(kd + 3 * i),
                                                  V3(ks + 3 * i), ns[i]);
}

// Mylight::select_a_point_from_lights (Mylight.cpp:102-160); up to 4 uniforms per call
void ref_light_sample_uniform(int n, const double* u, int* prim, double* point, double* Le, double* pdf)
{
    for (int i = 0; i < n; i++) {
        oracle_stream_list(&g_oracle_stream, u + 4 * i, 4);
        sampledLightPoint s = lights.select_a_point_from_lights(veach);
        prim[i] = (int)tri_to_prim(s.s, s.f);
        for (int c = 0; c < 3; c++) { point[3 * i + c] = s.coord.xyz[c]; Le[3 * i + c] = s.I.RGB[c]; }
        pdf[i] = s.prob;
    }
}

// staged == 0: select_a_point_from_lights_spherical_triangle (Mylight.cpp:163-318)
// staged == 1: prepared_for_... + lights_spherical_triangle_sampling (+ eval_..._pdf of query prim)
//              (Mylight.cpp:322-493).  wsum / nsurv only in staged mode (member state).
void ref_light_sample_spherical(int n, int staged, const double* x1, const double* N, const double* u, const int* query,
                                int* prim, double* point, double* Le, double* pdf, double* wsum, int* nsurv,
                                double* query_pdf)
{
    for (int i = 0; i < n; i++) {
        oracle_stream_list(&g_oracle_stream, u + 3 * i, 3);
        vec x = V3(x1 + 3 * i), nn = V3(N + 3 * i);
        if (staged) lights.prepared_for_lights_spherical_triangle_sampling(x, nn, veach);
        sampledLightPoint s = staged ? lights.lights_spherical_triangle_sampling(x, nn, veach)
                                     : lights.select_a_point_from_lights_spherical_triangle(x, nn, veach);
        // the "no surviving light" dummy returns (s,f)=(0,0), I=0, prob=1 (Mylight.cpp:263-266)
        prim[i] = (int)tri_to_prim(s.s, s.f);
        for (int c = 0; c < 3; c++) { point[3 * i + c] = s.coord.xyz[c]; Le[3 * i + c] = s.I.RGB[c]; }
        pdf[i] = s.prob;
        if (staged) {
            if (wsum) wsum[i] = lights.weights_sum;
            if (nsurv) nsurv[i] = (int)lights.sphericalTriangles.size();
            if (query && query_pdf)
                query_pdf[i] = lights.eval_spherical_triangle_sampling_pdf(prim_to_tri(query[i]), veach);
        }
    }
}

// RadianceRGB::tone_mapping (RadianceRGB.cpp:51-67)
void ref_tonemap(int n, const double* rgb, double max_radiance, double gamma, int* out)
{
    for (int i = 0; i < n; i++) {
        RadianceRGB r(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
        std::array<int, 3> o = r.tone_mapping(max_radiance, gamma);
        for (int c = 0; c < 3; c++) out[3 * i + c] = o[c];
    }
}

// ---- the render loop, main.cpp:546-588 with literals turned into arguments --------------------
// strategy: 0 = shade (main.cpp:269, as checked in: spherical-triangle NEE, F9)
//           1 = shade_with_brdf (main.cpp:348)       2 = shade_with_mis (main.cpp:402, incl. F10)
// rows [i0,i1), cols [j0,j1); out = double[(i1-i0)*(j1-j0)*3], linear radiance, like buffer[i][j].
// The uniform stream restarts at every (pixel, global sample index) -> row bands can be rendered by
// independent processes (the reference is not thread-safe).
void ref_render(const double eye[3], const double w_[3], const double up[3], double pixellen, int W, int H, int i0,
                int i1, int j0, int j1, int spp, int spp_offset, int spp_total, int strategy, int max_depth,
                uint64_t seed, double* out, uint64_t* counters)
{
    g_oracle_max_depth = max_depth;
    uint64_t v0 = g_oracle_vertices, primary_hits = 0;
    vec eyev = V3(eye), w = V3(w_);
    // main.cpp:548-553
    vec N = w.normalized();
    vec UP = V3(up);
    vec Vv = N.cross_product(UP).normalized();
    vec U = Vv.cross_product(N).normalized();
    matrix3d T = matrix3d(U, Vv, N);
    const double ci = (H - 1) * 0.5, cj = (W - 1) * 0.5;  // 359.5 / 639.5 at 720x1280
    const double inv = 1.0 / spp_total;                   // the literal 0.1 at 10 spp (main.cpp:575)
    for (int i = i0; i < i1; i++)
        for (int j = j0; j < j1; j++) {
            // main.cpp:563-564
            vec delta(-pixellen * (i - ci), pixellen * (j - cj), 0);
            vec dir = (T * (delta + vec(0, 0, w.norm2()))).normalized();
            RadianceRGB sum(0, 0, 0);
            for (int k = 0; k < spp; k++) {
                oracle_stream_seed(&g_oracle_stream, seed, (uint64_t)i * (uint64_t)W + (uint64_t)j,
                                   (uint64_t)(spp_offset + k));
                // main.cpp:572-577
                intersec_result rs = veach.closet_ray_intersect(eyev, dir, triangle(-1, -1));
                if (rs.isIntersec) {
                    primary_hits++;
                    RadianceRGB I = (strategy == 0   ? shade(rs, dir * -1)
                                     : strategy == 1 ? shade_with_brdf(rs, dir * -1)
                                                     : shade_with_mis(rs, dir * -1)) * inv;
                    sum = sum + I;
                }
            }
            double* o = out + ((size_t)(i - i0) * (size_t)(j1 - j0) + (size_t)(j - j0)) * 3;
            o[0] = sum.RGB[0]; o[1] = sum.RGB[1]; o[2] = sum.RGB[2];
        }
    if (counters) {
        counters[0] = g_oracle_vertices - v0;  // non-emissive vertices shaded (RR engines built)
        counters[1] = primary_hits;
    }
}

}  // extern "C"
