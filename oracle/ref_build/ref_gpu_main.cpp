// oracle/ref_build/ref_gpu_main.cpp -- the REFERENCE-SIDE BINDING, compiled: the reference's main() (main.cpp:497-599)
// with its pixel loop (main.cpp:557-588) handed to libmcpt.so.
//
// Everything up to the loop is the reference's own code, unmodified, linked from oracle/_ref/libmcpt_ref.so: the
// globals `Myobj veach` / `Mylight lights` (main.cpp:23-24) and veach.read(); lights.read();
// lights.gather_light_triangles(veach.reader) (main.cpp:500-504) -- i.e. tiny_obj_loader and pugixml load the scene,
// not the repo's own host loaders.  upload_scene() flattens what those two objects hold through the reference's own
// accessors into McptSceneDesc, mcpt_render replaces the loop, and RadianceRGB::tone_mapping (main.cpp:583) produces
// the bytes.  The EasyX window / saveimage (graphics API) is replaced by a PFM (linear radiance) + PPM writer.
// INTEGRATION.md quotes this file; tests/test_gpu_render.py::test_reference_side_binding_is_a_drop_in checks that the
// film equals, bit for bit, what the repo's own driver (csrc/mcpt_main.cpp) renders from the same files.
//
//   ref_gpu_main scene.obj scene.xml out.pfm [--spp 10] [--strategy mis|nee|nee_uniform|brdf|mis_single] [--seed 1]
//                [--max-depth 0] [--ppm out.ppm] [--leaf exact] [--parity] [--gpus N]
// Camera: the <camera> node of the XML (README.md:339-343) read with pugixml; the reference hard-codes its literals
// (main.cpp:507-517, F7).
#include "Myobj.h"
#include "Mylight.h"
#include "RadianceRGB.h"
#include "pugixml.hpp"
#include "mcpt.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <execinfo.h>
#include <iostream>
#include <signal.h>
#include <unistd.h>
#include <string>
#include <vector>

extern Myobj veach;     // main.cpp:23
extern Mylight lights;  // main.cpp:24

// (s,f) -> flat prim = shape_offset[s] + f               (triangle key, Myobj.h:44-66)
// Myobj& is non-const on purpose: get_vertexes_of_facet / get_normals_of_facet fill caches (Myobj.h:107-113).
static int upload_scene(Myobj& veach, Mylight& lights, int device, McptSceneDesc* keep, std::vector<std::vector<char>>& store,
                        McptScene** scene)
{
    const auto& shapes = veach.reader.GetShapes();
    const auto& mats = veach.reader.GetMaterials();
    std::vector<uint32_t> shape_offset{0};
    for (auto& sh : shapes) shape_offset.push_back(shape_offset.back() + (uint32_t)sh.mesh.num_face_vertices.size());
    const uint32_t P = shape_offset.back();
    std::vector<float> pos(9 * (size_t)P), nor(9 * (size_t)P);
    std::vector<int32_t> mat(P);
    for (size_t s = 0; s < shapes.size(); s++)
        for (size_t f = 0; f < shapes[s].mesh.num_face_vertices.size(); f++) {
            const uint32_t p = shape_offset[s] + (uint32_t)f;
            auto v = veach.get_vertexes_of_facet(s, f);         // Myobj.cpp:626-650
            auto n = veach.get_normals_of_facet(s, f);          // Myobj.cpp:653-678
            for (int k = 0; k < 3; k++)
                for (int c = 0; c < 3; c++) { pos[9 * p + 3 * k + c] = (float)v[k].xyz[c]; nor[9 * p + 3 * k + c] = (float)n[k].xyz[c]; }
            mat[p] = shapes[s].mesh.material_ids[f];            // main.cpp:292
        }
    std::vector<float> kd, ks, ns;
    for (auto& m : mats) {                                      // tiny_obj_loader.h:185-193
        kd.insert(kd.end(), m.diffuse, m.diffuse + 3); ks.insert(ks.end(), m.specular, m.specular + 3); ns.push_back(m.shininess);
    }
    // lights in the reference's iteration order: std::map by mtlname, then gather order (Mylight.cpp:41-86,169-173);
    // a <light> whose material no face uses still counts as a class (lightsRadiance, Mylight.h:106)
    std::vector<int32_t> light_prim, light_class;
    std::vector<double> class_radiance;
    int cls = 0;
    for (auto& kv : lights.lightsRadiance) {
        class_radiance.insert(class_radiance.end(), kv.second.RGB, kv.second.RGB + 3);
        auto it = lights.lightsTriangles.find(kv.first);        // Mylight.h:108
        if (it != lights.lightsTriangles.end())
            for (auto& t : it->second) { light_prim.push_back((int32_t)(shape_offset[t.s] + t.f)); light_class.push_back(cls); }
        cls++;
    }
    McptSceneDesc d{P, (uint32_t)shapes.size(), (uint32_t)mats.size(), (uint32_t)light_prim.size(), (uint32_t)cls,
                    pos.data(), nor.data(), mat.data(), shape_offset.data(), kd.data(), ks.data(), ns.data(),
                    light_prim.data(), light_class.data(), class_radiance.data()};
    if (keep) {  // mcpt_render_multi wants the description again: keep the arrays alive
        auto hold = [&](const void* p, size_t bytes) { store.emplace_back((const char*)p, (const char*)p + bytes); return (const void*)store.back().data(); };
        *keep = d;
        keep->positions = (const float*)hold(pos.data(), pos.size() * 4); keep->normals = (const float*)hold(nor.data(), nor.size() * 4);
        keep->material_id = (const int32_t*)hold(mat.data(), mat.size() * 4); keep->shape_offset = (const uint32_t*)hold(shape_offset.data(), shape_offset.size() * 4);
        keep->mat_kd = (const float*)hold(kd.data(), kd.size() * 4); keep->mat_ks = (const float*)hold(ks.data(), ks.size() * 4);
        keep->mat_ns = (const float*)hold(ns.data(), ns.size() * 4);
        keep->light_prim = (const int32_t*)hold(light_prim.data(), light_prim.size() * 4);
        keep->light_class = (const int32_t*)hold(light_class.data(), light_class.size() * 4);
        keep->class_radiance = (const double*)hold(class_radiance.data(), class_radiance.size() * 8);
    }
    if (!scene) return 0;
    return mcpt_scene_create(&d, device, scene);  // the library copies everything; the vectors may die here
}

static void on_crash(int sig)
{
    void* frames[64];
    const int n = backtrace(frames, 64);
    const char msg[] = "ref_gpu_main: fatal signal, backtrace:\n";
    if (write(2, msg, sizeof msg - 1) < 0) {}
    backtrace_symbols_fd(frames, n, 2);
    _exit(128 + sig);
}

int main(int argc, char** argv)
{
    signal(SIGSEGV, on_crash);
    signal(SIGABRT, on_crash);
    if (argc < 4) { std::fprintf(stderr, "usage: %s scene.obj scene.xml out.pfm [options]\n", argv[0]); return 2; }
    int spp = 10, strategy = MCPT_MIS, max_depth = 0, leaf = MCPT_LEAF_FAST, gpus = 1;  // main.cpp:567,575
    unsigned flags = 0;
    unsigned long long seed = 1;
    std::string ppm;
    for (int i = 4; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() -> const char* { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "--spp") spp = std::atoi(next());
        else if (a == "--seed") seed = std::strtoull(next(), nullptr, 10);
        else if (a == "--max-depth") max_depth = std::atoi(next());
        else if (a == "--ppm") ppm = next();
        else if (a == "--parity") flags |= MCPT_PARITY_FP64;
        else if (a == "--gpus") gpus = std::atoi(next());
        else if (a == "--leaf") leaf = std::string(next()) == "exact" ? MCPT_LEAF_EXACT : MCPT_LEAF_FAST;
        else if (a == "--strategy") {
            std::string s = next();
            strategy = s == "mis" ? MCPT_MIS : s == "mis_single" ? MCPT_MIS_SINGLE : s == "brdf" ? MCPT_BRDF_ONLY : s == "nee_uniform" ? MCPT_NEE_UNIFORM : MCPT_NEE_SPHERICAL;
        } else { std::fprintf(stderr, "unknown option %s\n", a.c_str()); return 2; }
    }
    // main.cpp:500-504, the reference's own loaders (tiny_obj_loader, pugixml)
    veach.inputfile = argv[1];
    lights.inputfile = argv[2];
    veach.read();
    lights.read();
    lights.gather_light_triangles(veach.reader);

    // camera: main.cpp:507-517 hard-codes these; here they come from the <camera> node the README documents
    pugi::xml_document doc;
    if (!doc.load_file(argv[2])) { std::cerr << "cannot parse " << argv[2] << "\n"; return 1; }
    pugi::xml_node c = doc.child("camera");
    if (!c) { std::cerr << argv[2] << " has no <camera> node\n"; return 1; }
    const int W = c.attribute("width").as_int(), H = c.attribute("height").as_int();
    const double fovy = c.attribute("fovy").as_double();
    auto v3 = [&](const char* name) { pugi::xml_node n = c.child(name); return vec(n.attribute("x").as_double(), n.attribute("y").as_double(), n.attribute("z").as_double()); };
    vec start = v3("eye"), lookat = v3("lookat"), UP = v3("up");
    vec w = lookat - start;                                         // main.cpp:546

    McptCamera cam;
    for (int k = 0; k < 3; k++) { cam.eye[k] = start.xyz[k]; cam.w[k] = w.xyz[k]; cam.up[k] = UP.xyz[k]; }
    cam.pixellen = std::tan(fovy * M_PI / 360.0) * w.norm2() / (H * 0.5);  // main.cpp:547 with the radians done right (F7)
    McptRenderParams prm;
    std::memset(&prm, 0, sizeof prm);
    prm.width = W; prm.height = H;                                  // main.cpp:496,539
    prm.spp = spp; prm.spp_total = spp;                             // main.cpp:567,575
    prm.strategy = strategy;                                        // which shade* (main.cpp:269 / 348 / 402)
    prm.max_depth = max_depth; prm.flags = flags; prm.seed = seed; prm.leaf_mode = leaf;
    std::vector<float> film((size_t)W * H * 3);
    McptStats st;
    if (gpus > 1) {
        // spp sharded over `gpus` devices inside the library, one NCCL reduce of the films (include/mcpt.h)
        McptSceneDesc desc;
        std::vector<std::vector<char>> store;
        store.reserve(16);
        upload_scene(veach, lights, 0, &desc, store, nullptr);
        std::vector<int32_t> devices(gpus);
        for (int g = 0; g < gpus; g++) devices[g] = g;
        if (mcpt_render_multi(&desc, devices.data(), gpus, &cam, &prm, film.data(), &st)) { std::cerr << mcpt_last_error() << "\n"; return 1; }
    } else {
        McptScene* scene = nullptr;
        std::vector<std::vector<char>> store;
        if (upload_scene(veach, lights, 0, nullptr, store, &scene)) { std::cerr << mcpt_last_error() << "\n"; return 1; }
        if (mcpt_render(scene, &cam, &prm, film.data(), &st)) { std::cerr << mcpt_last_error() << "\n"; return 1; }
        mcpt_scene_destroy(scene);
    }
    std::printf("%dx%d, %d spp: %.2f ms on the device, %.1f Msamples/s\n", W, H, spp, st.ms_total, st.paths / st.ms_total / 1e3);

    if (FILE* f = std::fopen(argv[3], "wb")) {  // linear radiance, bottom row first
        std::fprintf(f, "PF\n%d %d\n-1.0\n", W, H);
        for (int i = H - 1; i >= 0; i--) std::fwrite(&film[(size_t)i * W * 3], sizeof(float), (size_t)W * 3, f);
        std::fclose(f);
    } else { std::perror(argv[3]); return 1; }
    if (!ppm.empty()) {
        if (FILE* f = std::fopen(ppm.c_str(), "wb")) {
            std::fprintf(f, "P6\n%d %d\n255\n", W, H);
            for (int i = 0; i < H; i++)
                for (int j = 0; j < W; j++) {
                    const float* p = &film[((size_t)i * W + j) * 3];
                    RadianceRGB px(p[0], p[1], p[2]);                                  // main.cpp:582 buffer[i][j]
                    std::array<int, 3> rgb = px.tone_mapping(380, 0.25);               // main.cpp:583
                    unsigned char b[3] = {(unsigned char)rgb[0], (unsigned char)rgb[1], (unsigned char)rgb[2]};
                    std::fwrite(b, 1, 3, f);
                }
            std::fclose(f);
        } else { std::perror(ppm.c_str()); return 1; }
    }
    return 0;
}
