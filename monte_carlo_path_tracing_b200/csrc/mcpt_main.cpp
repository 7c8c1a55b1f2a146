// csrc/mcpt_main.cpp -- the reference's main() (main.cpp:497-599) with the pixel loop handed to libmcpt.so.
//
// Same sequence as the reference: read the OBJ (veach.read(), main.cpp:500), read the XML lights and gather
// the emissive triangles (lights.read(); lights.gather_light_triangles(), :503-504), set up the camera
// (:507-553, here from the XML <camera> node of README.md:339-343 or from --eye/--lookat/--up/--fovy), render
// (:557-588 -> mcpt_render), tone-map (RadianceRGB::tone_mapping(380, 0.25), :583 -> mcpt_tonemap) and save
// the image (saveimage, :597 -> binary PPM; the EasyX window is a graphics API and is not reproduced).
// Every parameter the reference hard-codes in source is a command-line option.  Uses only include/mcpt.h.
//
//   mcpt_render scene.obj scene.xml out.ppm [--spp 10] [--strategy mis|nee|nee_uniform|brdf] [--max-depth 0]
//               [--width W --height H] [--seed 1] [--device 0] [--fix-pdf] [--fix-measure] [--jitter]
//               [--lmax 380 --gamma 0.25] [--pfm out.pfm] [--ref-fov]
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "../../include/mcpt.h"

static int fail(const char* what)
{
    std::fprintf(stderr, "mcpt_render: %s: %s\n", what, mcpt_last_error());
    return 1;
}

int main(int argc, char** argv)
{
    if (argc < 4) {
        std::fprintf(stderr, "usage: %s scene.obj scene.xml out.ppm [options]  (see the header of mcpt_main.cpp)\n", argv[0]);
        return 2;
    }
    const char *obj = argv[1], *xml = argv[2], *out = argv[3];
    int spp = 10;                    // main.cpp:567
    int strategy = MCPT_MIS;         // main.cpp:575 calls shade_with_mis as checked in
    int max_depth = 0, width = 0, height = 0, device = 0;
    unsigned flags = 0;
    unsigned long long seed = 1;
    double lmax = 380.0, gamma = 0.25;  // main.cpp:583
    bool ref_fov = false, have_cam_args = false;
    double eye[3] = {0, 0, 0}, lookat[3] = {0, 0, 1}, up[3] = {0, 1, 0}, fovy = 0;
    std::string pfm;
    auto vec3 = [&](int& i, double* v) { for (int k = 0; k < 3 && i + 1 < argc; k++) v[k] = std::atof(argv[++i]); };
    for (int i = 4; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() -> const char* { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "--spp") spp = std::atoi(next());
        else if (a == "--strategy") {
            std::string s = next();
            strategy = s == "mis" ? MCPT_MIS : s == "brdf" ? MCPT_BRDF_ONLY : s == "nee_uniform" ? MCPT_NEE_UNIFORM : MCPT_NEE_SPHERICAL;
        } else if (a == "--max-depth") max_depth = std::atoi(next());
        else if (a == "--width") width = std::atoi(next());
        else if (a == "--height") height = std::atoi(next());
        else if (a == "--seed") seed = std::strtoull(next(), nullptr, 10);
        else if (a == "--device") device = std::atoi(next());
        else if (a == "--fix-pdf") flags |= MCPT_FIX_MIXTURE_PDF;
        else if (a == "--fix-measure") flags |= MCPT_FIX_NEE_MEASURE;
        else if (a == "--jitter") flags |= MCPT_JITTER;
        else if (a == "--lmax") lmax = std::atof(next());
        else if (a == "--gamma") gamma = std::atof(next());
        else if (a == "--pfm") pfm = next();
        else if (a == "--ref-fov") ref_fov = true;  // F7: the reference's tan(fovy/360) instead of tan(fovy/2 in radians)
        else if (a == "--eye") { vec3(i, eye); have_cam_args = true; }
        else if (a == "--lookat") vec3(i, lookat);
        else if (a == "--up") vec3(i, up);
        else if (a == "--fovy") fovy = std::atof(next());
        else { std::fprintf(stderr, "unknown option %s\n", a.c_str()); return 2; }
    }

    McptHostScene* hs = nullptr;
    if (mcpt_host_scene_load(obj, xml, &hs)) return fail("loading the scene");  // the reference exit(1)s here
    const McptSceneDesc* desc = mcpt_host_scene_desc(hs);
    std::printf("scene: %u triangles, %u shapes, %u materials, %u light triangles in %u classes\n", desc->num_prims,
                desc->num_shapes, desc->num_materials, desc->num_lights, desc->num_classes);

    McptCamera cam;
    int32_t cw = 0, chh = 0;
    double cfov = 0;
    const int has_cam = mcpt_host_scene_camera(hs, &cam, &cw, &chh, &cfov);
    if (have_cam_args || !has_cam) {
        if (!have_cam_args) { std::fprintf(stderr, "mcpt_render: the XML has no <camera>; pass --eye/--lookat/--up/--fovy\n"); return 2; }
        for (int k = 0; k < 3; k++) { cam.eye[k] = eye[k]; cam.w[k] = lookat[k] - eye[k]; cam.up[k] = up[k]; }
        cfov = fovy;
    } else if (fovy > 0) cfov = fovy;
    if (width <= 0) width = cw > 0 ? cw : 1280;    // main.cpp:539
    if (height <= 0) height = chh > 0 ? chh : 720;
    const double wl = std::sqrt(cam.w[0] * cam.w[0] + cam.w[1] * cam.w[1] + cam.w[2] * cam.w[2]);
    // main.cpp:547: pixellen = tan(fovy/360) * |w| / 360 at H = 720, i.e. tan(.) * |w| / (H/2)
    cam.pixellen = (ref_fov ? std::tan(cfov / 360.0) : std::tan(cfov * M_PI / 360.0)) * wl / (height * 0.5);

    McptScene* scene = nullptr;
    if (mcpt_scene_create(desc, device, &scene)) return fail("creating the device scene");
    uint32_t nodes = 0, leaves = 0, depth = 0;
    double sah = 0;
    mcpt_scene_bvh_info(scene, 0, &nodes, &leaves, &depth, &sah);
    std::printf("bvh: %u nodes, %u leaves, depth %u, SAH cost %.2f; %.1f MB on device %d\n", nodes, leaves, depth, sah,
                mcpt_scene_device_bytes(scene) / 1e6, device);

    McptRenderParams prm;
    std::memset(&prm, 0, sizeof prm);
    prm.width = width; prm.height = height; prm.spp = spp; prm.spp_total = spp; prm.max_depth = max_depth;
    prm.strategy = strategy; prm.flags = flags; prm.seed = seed;
    std::vector<float> film((size_t)width * height * 3);
    McptStats st;
    if (mcpt_render(scene, &cam, &prm, film.data(), &st)) return fail("rendering");
    std::printf("%dx%d, %d spp: %.2f ms on the device, %.1f Msamples/s, %.1f Msample-bounces/s, %.1f Mrays/s, %u kernel launches\n",
                width, height, spp, st.ms_total, st.paths / st.ms_total / 1e3, st.vertices / st.ms_total / 1e3,
                (st.rays_closest + st.rays_shadow + st.rays_light) / st.ms_total / 1e3, st.kernel_launches);

    std::vector<uint8_t> rgb8(film.size());
    if (mcpt_tonemap(scene, film.data(), (size_t)width * height, lmax, gamma, rgb8.data())) return fail("tone mapping");
    if (FILE* f = std::fopen(out, "wb")) {
        std::fprintf(f, "P6\n%d %d\n255\n", width, height);
        std::fwrite(rgb8.data(), 1, rgb8.size(), f);
        std::fclose(f);
    } else { std::perror(out); return 1; }
    if (!pfm.empty()) {
        if (FILE* f = std::fopen(pfm.c_str(), "wb")) {  // linear radiance, bottom row first
            std::fprintf(f, "PF\n%d %d\n-1.0\n", width, height);
            for (int i = height - 1; i >= 0; i--) std::fwrite(&film[(size_t)i * width * 3], sizeof(float), (size_t)width * 3, f);
            std::fclose(f);
        } else { std::perror(pfm.c_str()); return 1; }
    }
    mcpt_scene_destroy(scene);
    mcpt_host_scene_free(hs);
    return 0;
}
