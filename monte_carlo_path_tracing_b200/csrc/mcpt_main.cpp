// csrc/mcpt_main.cpp -- the reference's main() (main.cpp:497-599) with the pixel loop handed to libmcpt.so.
//
// Same sequence as the reference: read the OBJ (veach.read(), main.cpp:500), read the XML lights and gather
// the emissive triangles (lights.read(); lights.gather_light_triangles(), :503-504), set up the camera
// (:507-553, here from the XML <camera> node of README.md:339-343 or from --eye/--lookat/--up/--fovy), render
// (:557-588 -> mcpt_render), tone-map (RadianceRGB::tone_mapping(380, 0.25), :583 -> mcpt_tonemap) and save
// the image (saveimage, :597 -> a 24-bit BMP when the output name ends in .bmp, else binary PPM; the EasyX window is a
// graphics API and is not reproduced).
// Every parameter the reference hard-codes in source is a command-line option.  Uses only include/mcpt.h.
//
//   mcpt_render scene.obj scene.xml out.ppm|out.bmp [--spp 10] [--strategy mis|mis_single|nee|nee_uniform|brdf] [--max-depth 0]
//               [--width W --height H] [--seed 1] [--device 0] [--gpus N]   (N > 1: spp sharded over devices 0..N-1 inside
//               the library, one NCCL reduce of the films -- mcpt_multi_*) [--fix-pdf] [--fix-measure] [--jitter] [--light-tree]
//               [--leaf fast|exact] [--parity]   (MCPT_LEAF_EXACT / MCPT_PARITY_FP64: the reference's FP64 arithmetic)
//               [--lmax 380 --gamma 0.25] [--pfm out.pfm] [--ref-fov]
//               [--gpu-bvh] [--every N] [--checkpoint file [--resume]] [--target-rel-error e] [--time-budget seconds]
//
// Progressive rendering (the reference keeps one all-or-nothing `buffer`, main.cpp:496,582, and lost day-long renders
// to any crash): the spp are rendered in batches of N (default 64); the driver keeps the running sum of sample
// radiances in FP64 and the sum of squared batch means of the luminance, writes both to the checkpoint file after
// every batch (tmp + rename) and, with --resume, continues from it.  Samples are keyed by their global index
// (spp_offset), so a resumed render adds exactly the samples an uninterrupted one would have.  --target-rel-error
// stops when the mean over pixels of stderr(luminance) / (luminance + 1 % of the image mean) falls below e;
// --time-budget stops after that many seconds of device time.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>
#include "../../include/mcpt.h"

static int fail(const char* what)
{
    std::fprintf(stderr, "mcpt_render: %s: %s\n", what, mcpt_last_error());
    return 1;
}

// checkpoint file: header, then sum[W*H*3] and sumsq_lum[W*H] as doubles
struct CkptHeader {
    char magic[8];  // "MCPTCKP2"
    int32_t width, height, strategy, max_depth, batch, leaf_mode;
    uint32_t flags, num_prims, num_lights, batches;
    uint64_t seed, spp_done;
    uint64_t scene_hash;  // FNV-1a over every array of the McptSceneDesc: an edited OBJ/MTL/XML with the same counts is a different scene
    McptCamera cam;       // a resumed render must look through the same camera
};

static uint64_t fnv1a(uint64_t h, const void* data, size_t bytes)
{
    const unsigned char* p = (const unsigned char*)data;
    for (size_t i = 0; i < bytes; i++) { h ^= p[i]; h *= 1099511628211ull; }
    return h;
}
static uint64_t scene_hash(const McptSceneDesc* d)
{
    uint64_t h = 14695981039346656037ull;
    h = fnv1a(h, d->positions, (size_t)d->num_prims * 36); h = fnv1a(h, d->normals, (size_t)d->num_prims * 36);
    h = fnv1a(h, d->material_id, (size_t)d->num_prims * 4); h = fnv1a(h, d->shape_offset, (size_t)(d->num_shapes + 1) * 4);
    h = fnv1a(h, d->mat_kd, (size_t)d->num_materials * 12); h = fnv1a(h, d->mat_ks, (size_t)d->num_materials * 12);
    h = fnv1a(h, d->mat_ns, (size_t)d->num_materials * 4);
    h = fnv1a(h, d->light_prim, (size_t)d->num_lights * 4); h = fnv1a(h, d->light_class, (size_t)d->num_lights * 4);
    h = fnv1a(h, d->class_radiance, (size_t)d->num_classes * 24);
    return h;
}

static bool ckpt_write(const std::string& path, const CkptHeader& h, const std::vector<double>& sum, const std::vector<double>& sq)
{
    const std::string tmp = path + ".tmp";
    FILE* f = std::fopen(tmp.c_str(), "wb");
    if (!f) return false;
    bool ok = std::fwrite(&h, sizeof h, 1, f) == 1 && std::fwrite(sum.data(), 8, sum.size(), f) == sum.size() &&
              std::fwrite(sq.data(), 8, sq.size(), f) == sq.size();
    ok = (std::fclose(f) == 0) && ok;
    return ok && std::rename(tmp.c_str(), path.c_str()) == 0;
}

static bool ckpt_read(const std::string& path, CkptHeader* h, std::vector<double>& sum, std::vector<double>& sq)
{
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    bool ok = std::fread(h, sizeof *h, 1, f) == 1 && std::memcmp(h->magic, "MCPTCKP2", 8) == 0 && h->width > 0 && h->height > 0;
    if (ok) {
        sum.resize((size_t)h->width * h->height * 3);
        sq.resize((size_t)h->width * h->height);
        ok = std::fread(sum.data(), 8, sum.size(), f) == sum.size() && std::fread(sq.data(), 8, sq.size(), f) == sq.size();
    }
    std::fclose(f);
    return ok;
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
    int max_depth = 0, width = 0, height = 0, device = 0, leaf_mode = MCPT_LEAF_FAST, gpus = 1;
    unsigned flags = 0;
    unsigned long long seed = 1;
    double lmax = 380.0, gamma = 0.25;  // main.cpp:583
    bool ref_fov = false, have_cam_args = false, have_lookat = false, have_up = false;
    double eye[3] = {0, 0, 0}, lookat[3] = {0, 0, 1}, up[3] = {0, 1, 0}, fovy = 0;
    std::string pfm, ckpt;
    int every = 64;
    bool resume = false, gpu_bvh = false;
    double target_err = 0, time_budget = 0;
    auto vec3 = [&](int& i, double* v) { for (int k = 0; k < 3 && i + 1 < argc; k++) v[k] = std::atof(argv[++i]); };
    for (int i = 4; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() -> const char* { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "--spp") spp = std::atoi(next());
        else if (a == "--strategy") {
            std::string s = next();
            if (s == "mis") strategy = MCPT_MIS;
            else if (s == "mis_single") strategy = MCPT_MIS_SINGLE;
            else if (s == "brdf") strategy = MCPT_BRDF_ONLY;
            else if (s == "nee_uniform") strategy = MCPT_NEE_UNIFORM;
            else if (s == "nee") strategy = MCPT_NEE_SPHERICAL;
            else { std::fprintf(stderr, "unknown --strategy %s (mis, mis_single, nee, nee_uniform, brdf)\n", s.c_str()); return 2; }
        } else if (a == "--leaf") {
            std::string s = next();
            if (s == "exact") leaf_mode = MCPT_LEAF_EXACT;
            else if (s == "fast") leaf_mode = MCPT_LEAF_FAST;
            else { std::fprintf(stderr, "unknown --leaf %s (fast, exact)\n", s.c_str()); return 2; }
        } else if (a == "--parity") flags |= MCPT_PARITY_FP64;
        else if (a == "--max-depth") max_depth = std::atoi(next());
        else if (a == "--width") width = std::atoi(next());
        else if (a == "--height") height = std::atoi(next());
        else if (a == "--seed") seed = std::strtoull(next(), nullptr, 10);
        else if (a == "--device") device = std::atoi(next());
        else if (a == "--gpus") gpus = std::atoi(next());
        else if (a == "--fix-pdf") flags |= MCPT_FIX_MIXTURE_PDF;
        else if (a == "--fix-measure") flags |= MCPT_FIX_NEE_MEASURE;
        else if (a == "--jitter") flags |= MCPT_JITTER;
        else if (a == "--light-tree") flags |= MCPT_LIGHT_TREE;  // O(log L) light selection (not a reference estimator)
        else if (a == "--lmax") lmax = std::atof(next());
        else if (a == "--gamma") gamma = std::atof(next());
        else if (a == "--pfm") pfm = next();
        else if (a == "--gpu-bvh") gpu_bvh = true;  // Morton LBVH built on the device instead of the host SAH build
        else if (a == "--every") every = std::atoi(next());
        else if (a == "--checkpoint") ckpt = next();
        else if (a == "--resume") resume = true;
        else if (a == "--target-rel-error") target_err = std::atof(next());
        else if (a == "--time-budget") time_budget = std::atof(next());
        else if (a == "--ref-fov") ref_fov = true;  // F7: the reference's tan(fovy/360) instead of tan(fovy/2 in radians)
        else if (a == "--eye") { vec3(i, eye); have_cam_args = true; }
        else if (a == "--lookat") { vec3(i, lookat); have_lookat = true; }
        else if (a == "--up") { vec3(i, up); have_up = true; }
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
    } else {
        // the XML camera, with whichever of --lookat / --up / --fovy was given laid over it
        if (have_lookat) for (int k = 0; k < 3; k++) cam.w[k] = lookat[k] - cam.eye[k];
        if (have_up) for (int k = 0; k < 3; k++) cam.up[k] = up[k];
        if (fovy > 0) cfov = fovy;
    }
    if (width <= 0) width = cw > 0 ? cw : 1280;    // main.cpp:539
    if (height <= 0) height = chh > 0 ? chh : 720;
    const double wl = std::sqrt(cam.w[0] * cam.w[0] + cam.w[1] * cam.w[1] + cam.w[2] * cam.w[2]);
    // main.cpp:547: pixellen = tan(fovy/360) * |w| / 360 at H = 720, i.e. tan(.) * |w| / (H/2)
    cam.pixellen = (ref_fov ? std::tan(cfov / 360.0) : std::tan(cfov * M_PI / 360.0)) * wl / (height * 0.5);

    McptScene* scene = nullptr;
    McptMulti* multi = nullptr;
    if (gpus > 1) {
        std::vector<int32_t> devs(gpus);
        for (int g = 0; g < gpus; g++) devs[g] = device + g;
        if (mcpt_multi_create(desc, devs.data(), gpus, gpu_bvh ? MCPT_BUILD_GPU_LBVH : MCPT_BUILD_HOST_SAH, &multi)) return fail("creating the device scenes");
        scene = mcpt_multi_scene(multi, 0);
        std::printf("%d devices: spp sharded inside the library, one NCCL reduce of the films\n", gpus);
    } else if (mcpt_scene_create_ex(desc, device, gpu_bvh ? MCPT_BUILD_GPU_LBVH : MCPT_BUILD_HOST_SAH, &scene)) return fail("creating the device scene");
    uint32_t nodes = 0, leaves = 0, depth = 0;
    double sah = 0, tree_ms = 0, total_ms = 0;
    int32_t builder = 0;
    mcpt_scene_bvh_info(scene, 0, &nodes, &leaves, &depth, &sah);
    mcpt_scene_build_info(scene, &tree_ms, &total_ms, &builder);
    std::printf("bvh (%s, tree %.1f ms, scene ready in %.1f ms): %u nodes, %u leaves, depth %u, SAH cost %.2f; %.1f MB on device %d\n",
                builder == MCPT_BUILD_GPU_LBVH ? "device LBVH" : "host SAH", tree_ms, total_ms, nodes, leaves, depth, sah,
                mcpt_scene_device_bytes(scene) / 1e6, device);

    if (every <= 0) every = spp;
    const size_t npix = (size_t)width * height;
    std::vector<double> sum(npix * 3, 0.0), sq(npix, 0.0);
    CkptHeader hd;
    std::memset(&hd, 0, sizeof hd);
    std::memcpy(hd.magic, "MCPTCKP2", 8);
    hd.width = width; hd.height = height; hd.strategy = strategy; hd.max_depth = max_depth; hd.batch = every; hd.flags = flags;
    hd.num_prims = desc->num_prims; hd.num_lights = desc->num_lights; hd.seed = seed;
    hd.leaf_mode = leaf_mode; hd.scene_hash = scene_hash(desc); hd.cam = cam;
    if (resume && !ckpt.empty()) {
        CkptHeader old;
        std::vector<double> s0, q0;
        if (ckpt_read(ckpt, &old, s0, q0)) {
            const bool same = old.width == width && old.height == height && old.strategy == strategy && old.max_depth == max_depth &&
                              old.batch == every && old.flags == flags && old.num_prims == hd.num_prims &&
                              old.num_lights == hd.num_lights && old.seed == seed && old.leaf_mode == leaf_mode &&
                              old.scene_hash == hd.scene_hash && std::memcmp(&old.cam, &cam, sizeof cam) == 0;
            if (!same) { std::fprintf(stderr, "mcpt_render: %s was written for a different scene, camera or parameters\n", ckpt.c_str()); return 2; }
            hd.spp_done = old.spp_done; hd.batches = old.batches;
            sum.swap(s0); sq.swap(q0);
            std::printf("resumed %s at %llu spp\n", ckpt.c_str(), (unsigned long long)hd.spp_done);
        } else std::printf("no usable checkpoint at %s: starting from 0 spp\n", ckpt.c_str());
    }

    McptRenderParams prm;
    std::memset(&prm, 0, sizeof prm);
    prm.width = width; prm.height = height; prm.max_depth = max_depth;
    prm.strategy = strategy; prm.flags = flags; prm.seed = seed; prm.leaf_mode = leaf_mode;
    std::vector<float> film(npix * 3);
    McptStats tot;
    std::memset(&tot, 0, sizeof tot);
    double rel_err = 0;
    while (hd.spp_done < (uint64_t)spp) {
        const int n = (int)std::min<uint64_t>((uint64_t)every, (uint64_t)spp - hd.spp_done);
        prm.spp = n; prm.spp_total = n; prm.spp_offset = (int32_t)hd.spp_done;  // film = mean of this batch
        McptStats st;
        if (multi ? mcpt_multi_render(multi, &cam, &prm, film.data(), &st) : mcpt_render(scene, &cam, &prm, film.data(), &st)) return fail("rendering");
        double mean_lum = 0, err_acc = 0;
        for (size_t p = 0; p < npix; p++) {
            const double r = film[3 * p], g = film[3 * p + 1], b = film[3 * p + 2];
            sum[3 * p] += n * r; sum[3 * p + 1] += n * g; sum[3 * p + 2] += n * b;
            if (n == every) { const double l = (r + g + b) / 3; sq[p] += l * l; }  // full batches only: equal weights
        }
        hd.spp_done += (uint64_t)n;
        if (n == every) hd.batches++;
        tot.ms_total += st.ms_total; tot.paths += st.paths; tot.vertices += st.vertices; tot.kernel_launches += st.kernel_launches;
        tot.rays_closest += st.rays_closest; tot.rays_shadow += st.rays_shadow; tot.rays_light += st.rays_light;
        // standard error of the mean luminance from the spread of the batch means
        const uint32_t B = hd.batches;
        if (B >= 2) {
            for (size_t p = 0; p < npix; p++) mean_lum += (sum[3 * p] + sum[3 * p + 1] + sum[3 * p + 2]) / 3 / (double)hd.spp_done;
            mean_lum /= (double)npix;
            for (size_t p = 0; p < npix; p++) {
                const double m = (sum[3 * p] + sum[3 * p + 1] + sum[3 * p + 2]) / 3 / (double)hd.spp_done;
                const double var = std::max(0.0, sq[p] / B - m * m) / (B - 1);
                err_acc += std::sqrt(var) / (m + 0.01 * mean_lum + 1e-300);
            }
            rel_err = err_acc / (double)npix;
        }
        if (!ckpt.empty() && !ckpt_write(ckpt, hd, sum, sq)) { std::perror(ckpt.c_str()); return 1; }
        if (spp > every)
            std::printf("  %llu / %d spp, %.1f ms on the device%s\n", (unsigned long long)hd.spp_done, spp, tot.ms_total,
                        B >= 2 ? (", mean relative error " + std::to_string(rel_err)).c_str() : "");
        if (target_err > 0 && B >= 2 && rel_err < target_err) break;
        if (time_budget > 0 && tot.ms_total >= time_budget * 1e3) break;
    }
    const double inv = hd.spp_done ? 1.0 / (double)hd.spp_done : 0.0;
    for (size_t k = 0; k < film.size(); k++) film[k] = (float)(sum[k] * inv);
    const McptStats& st = tot;
    std::printf("%dx%d, %llu spp: %.2f ms on the device, %.1f Msamples/s, %.1f Msample-bounces/s, %.1f Mrays/s, %u kernel launches\n",
                width, height, (unsigned long long)hd.spp_done, st.ms_total, st.paths / st.ms_total / 1e3, st.vertices / st.ms_total / 1e3,
                (st.rays_closest + st.rays_shadow + st.rays_light) / st.ms_total / 1e3, st.kernel_launches);

    std::vector<uint8_t> rgb8(film.size());
    if (mcpt_tonemap(scene, film.data(), (size_t)width * height, lmax, gamma, rgb8.data())) return fail("tone mapping");
    const std::string outs = out;
    const bool bmp = outs.size() > 4 && (outs.substr(outs.size() - 4) == ".bmp" || outs.substr(outs.size() - 4) == ".BMP");
    if (FILE* f = std::fopen(out, "wb")) {
        if (bmp) {
            // what the reference's saveimage() leaves behind (main.cpp:597): 24-bit uncompressed BMP, bottom row first, BGR
            const uint32_t row = ((uint32_t)width * 3 + 3) & ~3u, data = row * (uint32_t)height, off = 54;
            uint8_t h[54] = {'B', 'M'};
            auto put32 = [&](int at, uint32_t v) { h[at] = v & 255; h[at + 1] = (v >> 8) & 255; h[at + 2] = (v >> 16) & 255; h[at + 3] = v >> 24; };
            put32(2, off + data); put32(10, off); put32(14, 40); put32(18, (uint32_t)width); put32(22, (uint32_t)height);
            h[26] = 1; h[28] = 24; put32(34, data); put32(38, 2835); put32(42, 2835);
            std::fwrite(h, 1, 54, f);
            std::vector<uint8_t> line(row, 0);
            for (int i = height - 1; i >= 0; i--) {
                for (int j = 0; j < width; j++) {
                    const uint8_t* p = &rgb8[((size_t)i * width + j) * 3];
                    line[3 * j] = p[2]; line[3 * j + 1] = p[1]; line[3 * j + 2] = p[0];
                }
                std::fwrite(line.data(), 1, row, f);
            }
        } else {
            std::fprintf(f, "P6\n%d %d\n255\n", width, height);
            std::fwrite(rgb8.data(), 1, rgb8.size(), f);
        }
        std::fclose(f);
    } else { std::perror(out); return 1; }
    if (!pfm.empty()) {
        if (FILE* f = std::fopen(pfm.c_str(), "wb")) {  // linear radiance, bottom row first
            std::fprintf(f, "PF\n%d %d\n-1.0\n", width, height);
            for (int i = height - 1; i >= 0; i--) std::fwrite(&film[(size_t)i * width * 3], sizeof(float), (size_t)width * 3, f);
            std::fclose(f);
        } else { std::perror(pfm.c_str()); return 1; }
    }
    if (multi) mcpt_multi_destroy(multi);
    else mcpt_scene_destroy(scene);
    mcpt_host_scene_free(hs);
    return 0;
}
