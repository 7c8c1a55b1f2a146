/* include/mcpt.h -- C ABI of libmcpt.so, the B200 (sm_100a) drop-in for the render loop of
 * luotong96/Monte_Carlo_Path_Tracing.
 *
 * The reference has no plugin/FFI interface: its "API" is the pixel/sample loop of main()
 * (main.cpp:557-588) calling RadianceRGB shade*(intersec_result, vec) (main.cpp:269,348,402)
 * on two file-scope globals, `Myobj veach` and `Mylight lights` (main.cpp:23-24).  The seam cut
 * here is that whole loop: inputs = what Myobj / Mylight hold after read() + gather_light_triangles()
 * (main.cpp:500-504) flattened to plain arrays + the camera literals (main.cpp:507-553) + the
 * loop constants (resolution, spp, which shade*), output = the film `buffer[i][j]` (main.cpp:496,582)
 * as linear radiance.  Plain pointers and sizes only; no C++ or torch types.  Every entry point
 * returns 0 on success or a negative McptError and never calls exit() (the reference exit(1)s on
 * loader failure, Myobj.cpp:15-20, Mylight.cpp:16-20); mcpt_last_error() gives the message.
 *
 * One host thread per McptScene.  All pointers are HOST pointers unless named d_*.
 */
#ifndef MCPT_H
#define MCPT_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define MCPT_VERSION 100

typedef enum McptError {
    MCPT_OK = 0,
    MCPT_ERR_INVALID = -1,        /* bad argument / malformed scene */
    MCPT_ERR_CUDA = -2,           /* CUDA runtime error (message in mcpt_last_error) */
    MCPT_ERR_NO_DEVICE = -3,      /* no CUDA device: there is NO CPU fallback */
    MCPT_ERR_IO = -4,             /* file not found / parse error (host loaders) */
    MCPT_ERR_QUEUE_OVERFLOW = -5, /* MIS path-tree outgrew the ray queue; retry with smaller batch_spp */
    MCPT_ERR_OOM = -6,
    MCPT_ERR_NCCL = -7            /* multi-device render: libnccl.so.2 missing, or an NCCL call failed */
} McptError;

/* Which shade* body runs (main.cpp:269 / 348 / 402).  Same numbering as oracle/mcpt_oracle.h. */
typedef enum McptStrategy {
    MCPT_NEE_SPHERICAL = 0, /* shade(): 1 spherical-triangle light sample + 1 BRDF sample (main.cpp:298) */
    MCPT_BRDF_ONLY = 1,     /* shade_with_brdf() */
    MCPT_MIS = 2,           /* shade_with_mis(): both samples recursed, balance heuristic (path tree, F6) */
    MCPT_NEE_UNIFORM = 3,   /* shade() with the uniform-area sampler of the commented line main.cpp:297 */
    MCPT_MIS_SINGLE = 4     /* NOT a reference estimator: textbook single-path MIS -- the spherical-
                             * triangle light sample contributes emitted light only (shadow ray, balance weight), only the
                             * BRDF sample continues, and emission found by it is weighted pdf_b / (pdf_b + pdf_light).
                             * Same expectation as MCPT_MIS with MCPT_FIX_MIXTURE_PDF | MCPT_FIX_NEE_MEASURE, no path
                             * tree, one visibility ray per vertex: the fast mode. */
} McptStrategy;

/* flags: 0 == the reference's estimators as written (F8, F9 reproduced).  The stale-MIS-state bug
 * F10 (Mylight scratch overwritten by recursion #1, main.cpp:443/464/487) is never reproduced: the
 * light pdf is always evaluated for the current vertex (== oracle flag ORACLE_FIX_STALE_MIS). */
enum {
    MCPT_FIX_NEE_MEASURE = 2,  /* F9: NEE_SPHERICAL uses the solid-angle estimator Le f cos / pdf */
    MCPT_FIX_MIXTURE_PDF = 4,  /* F8: BRDF-sampled directions are weighted with the mixture pdf */
    MCPT_JITTER = 16,          /* uniform sub-pixel jitter (the reference has none, F4) */
    MCPT_ACCUMULATE = 64,      /* mcpt_render_device: add into d_out instead of overwriting */
    MCPT_PARITY_FP64 = 256,    /* PARITY render mode: the whole estimator -- camera ray, closest hits, vertex, Phong eval /
                                * sample / pdf, both light samplers, roulette, accumulation -- runs in FP64 in the reference's
                                * expression order without FMA contraction (csrc/render_parity.cu), so a film is comparable
                                * PIXEL BY PIXEL with the oracle fed the same keyed uniforms (oracle flag ORACLE_RNG_KEYED),
                                * not just statistically.  Implies MCPT_LEAF_EXACT.  Orders of magnitude slower than the
                                * FP32 wavefront; it exists so that the renderer itself, not a test kernel, meets the
                                * bit-exact-ids / 1e-5 bar.  Ignores MCPT_JITTER-less shortcuts and MCPT_LIGHT_TREE. */
    MCPT_LIGHT_TREE = 128      /* NOT a reference estimator: the spherical strategies pick their
                                * light triangle by descending a light hierarchy (csrc/light_tree.h, O(log L) per vertex)
                                * instead of the O(L) loop of Mylight.cpp:169-276; the chosen triangle is still sampled by
                                * Arvo's spherical-triangle sampler and the pdf used is P(triangle | vertex) / sA.  Same
                                * expectation, different variance. */
};

/* Flattened Myobj + Mylight.  prim = shape_offset[s] + f for the reference's (s,f) (Myobj.h:44-66). */
typedef struct McptSceneDesc {
    uint32_t num_prims;       /* P  triangles (tinyobj triangulates, tiny_obj_loader.h:509-531) */
    uint32_t num_shapes;      /* S */
    uint32_t num_materials;   /* M */
    uint32_t num_lights;      /* L  emissive triangles (Mylight::lightsTriangles, Mylight.h:108) */
    uint32_t num_classes;     /* C  <light> nodes (Mylight::lightsRadiance, Mylight.h:106) */
    const float* positions;   /* [P][3][3] get_vertexes_of_facet (Myobj.cpp:626-650); tinyobj real_t = float */
    const float* normals;     /* [P][3][3] get_normals_of_facet (Myobj.cpp:653-678) */
    const int32_t* material_id;   /* [P] mesh.material_ids[f] (main.cpp:292) */
    const uint32_t* shape_offset; /* [S+1] */
    const float* mat_kd;      /* [M][3] material_t::diffuse */
    const float* mat_ks;      /* [M][3] material_t::specular */
    const float* mat_ns;      /* [M]    material_t::shininess */
    const int32_t* light_prim;    /* [L] in the reference's iteration order: class (mtlname ascending,
                                     std::map), then gather order (s,f ascending) (Mylight.cpp:41-86,169-173) */
    const int32_t* light_class;   /* [L] non-decreasing */
    const double* class_radiance; /* [C][3] RadianceRGB(std::string) = stod (RadianceRGB.cpp:17-27) */
} McptSceneDesc;

/* Camera exactly as main.cpp:546-553 parametrises it: dir(i,j) = normalize(T * (-pl*(i-(H-1)/2),
 * pl*(j-(W-1)/2), |w|)), T = [U V N], N = w/|w|, V = normalize(N x up), U = normalize(V x N). */
typedef struct McptCamera {
    double eye[3];
    double w[3];      /* lookat - eye, NOT normalised (|w| enters the formula) */
    double up[3];
    double pixellen;  /* tan(fovy/2) * |w| / (H/2); the reference's literal is F7 */
} McptCamera;

/* Ray/triangle test used by mcpt_render.  MCPT_LEAF_FAST: FP32 Moeller-Trumbore over the 4-wide
 * BVH, any-hit visibility rays -- the throughput path.  MCPT_LEAF_EXACT: every ray of the render (camera, extension,
 * visibility, light-only) is resolved by the reference's own test, the FP64 Cramer expression tree of
 * Myobj::intersect_with_triangle (Myobj.cpp:165-192) evaluated without FMA contraction inside FP64 slab tests, and
 * visibility is the reference's closest-hit-and-compare-ids (main.cpp:309-311) instead of an any-hit search: the ids
 * the wavefront sees are the ones the reference's exhaustive search returns for the same ray.  Shading stays FP32
 * unless MCPT_PARITY_FP64 is also set. */
typedef enum McptLeafMode { MCPT_LEAF_FAST = 0, MCPT_LEAF_EXACT = 1 } McptLeafMode;

typedef struct McptRenderParams {
    int32_t width, height;
    int32_t spp;         /* samples per pixel rendered by THIS call */
    int32_t spp_offset;  /* global index of the first sample (spp sharding across GPUs) */
    int32_t spp_total;   /* normalisation: out = sum / spp_total (main.cpp:575 `* 0.1` at 10 spp); 0 -> spp */
    int32_t max_depth;   /* RR forced to fail at vertices of depth >= max_depth; <=0 unlimited (F5) */
    int32_t strategy;    /* McptStrategy */
    uint32_t flags;
    uint64_t seed;
    float p_rr;          /* 0 -> 0.6 (main.cpp:321,375,429) */
    int32_t batch_spp;   /* samples per pixel in flight per wavefront pass; 0 -> auto */
    int32_t leaf_mode;   /* McptLeafMode: which ray/triangle test every ray of the render uses */
    int32_t reserved;    /* must be 0 */
} McptRenderParams;

typedef struct McptStats {
    uint64_t paths;          /* primary samples = W*H*spp */
    uint64_t primary_hits;
    uint64_t vertices;       /* sample-bounces: non-emissive vertices shaded (RR draws) */
    uint64_t rays_closest;   /* closest-hit rays traced (primary rays counted once per pixel if no jitter) */
    uint64_t rays_shadow;    /* NEE visibility rays */
    uint64_t rays_light;     /* light-only rays (MIS light pdf, main.cpp:484) */
    uint32_t passes;         /* wavefront passes (batches) */
    uint32_t bounces;        /* kernel bounce iterations summed over passes */
    uint32_t kernel_launches;
    uint32_t queue_peak;     /* largest ray-queue length seen */
    float ms_total;          /* device time of the call, CUDA events on the launch stream */
    float ms_bounce;         /* time in the per-bounce pipeline: k_shade, k_gather, k_extend, k_shadow, k_runsum */
    float ms_other;
    float ms_shade;          /* part of ms_bounce spent in k_shade (CUDA events on the launch stream) */
    float ms_trace;          /* part of ms_bounce spent in k_gather + k_extend + k_shadow (+ k_runsum) */
    uint64_t elements;       /* ray-queue elements processed by k_shade, summed over bounces (>= vertices) */
    uint64_t light_sweeps;   /* vertices for which the whole light table was evaluated (the O(L) loop of Mylight.cpp:329-421):
                              * once per shadable PIXEL per pass at the first vertex (F4), once per sampling vertex after it */
    uint64_t light_evals;    /* (light, vertex) spherical-triangle weights those sweeps evaluated: light_sweeps x the padded table
                              * length, or less when the sweeps run over per-primitive survivor lists (csrc/light_lists.h) */
    float ms_sweep;          /* part of ms_shade spent in the light-sweep kernel alone (k_lightsel / k_lightsel_sorted / k_lighttree) */
    float ms_prep;           /* k_lightprep: the first-vertex sweeps, once per pixel and pass (part of ms_other) */
} McptStats;

typedef struct McptScene McptScene;

/* ---- scene --------------------------------------------------------------------------------- */
/* Copies the arrays, computes geometric normals (get_unique_normal_of_facet, Myobj.cpp:680-709) and
 * light areas/CDFs (Mylight.cpp:66-69,112-135), builds the SAH BVHs (replacing cal_scene_boundingbox +
 * meshing, Myobj.cpp:78-162) and uploads everything to `device`.  Caller keeps its arrays. */
int mcpt_scene_create(const McptSceneDesc* desc, int device, McptScene** out);

/* Same, choosing the builder of the scene BVH.  MCPT_BUILD_GPU_LBVH: Morton-code LBVH built on the device
 * (csrc/bvh_gpu.cu; 45 ms instead of 670 ms at 10^6 triangles, SAH cost 8 % higher, rendering 9 % slower).  Results of
 * every query and of the renderer are identical for both builders (closest hit, ties by prim id, do not depend on the
 * tree).  A tree too deep for the traversal stack falls back to the host SAH builder.
 * mcpt_scene_build_info: milliseconds spent building the scene BVH (tree construction, and layout emission + the rest
 * of scene_create separately) and the builder actually used. */
enum { MCPT_BUILD_HOST_SAH = 0, MCPT_BUILD_GPU_LBVH = 1 };
int mcpt_scene_create_ex(const McptSceneDesc* desc, int device, uint32_t build_flags, McptScene** out);
int mcpt_scene_build_info(const McptScene*, double* tree_ms, double* total_ms, int32_t* builder);
/* Which wide BVH the trace kernels walk for this scene -- 4 (112 B nodes, distance-sorted children) or 8 (compressed 80 B
 * nodes, octant order; csrc/bvh.h).  Hits and films are identical for both, so the library measures: calib_ms = time of
 * a 512 Ki-ray batch of random surface rays through each layout at scene creation (the first guess), render_ms = time of
 * both kernels on the first bounce of >= 256 Ki real rays of the first render, which settles it (0, 0: not measured --
 * MCPT_BVH8=0/1 forced the choice, or no such bounce yet). */
int mcpt_scene_trace_layout(const McptScene*, int32_t* width, float calib_ms[2], float render_ms[2]);
/* Per-primitive survivor lists of the O(L) spherical-triangle light selection (csrc/light_lists.h): for every non-emissive
 * primitive the lights that are NOT culled (Mylight.cpp:178-197: vertex behind the light / light below the horizon) at
 * every point of it.  The sweeps evaluate the list instead of the table: identical weights and probabilities, about half
 * the evaluations on veach-mis.  lists = 0: not in use for this scene (few lights, too many primitives, MCPT_LIGHT_LISTS=0);
 * mean_fraction = mean list length / lights. */
int mcpt_scene_light_lists(const McptScene*, uint32_t* lists, uint64_t* entries, double* mean_fraction, double* build_ms);
/* The list of primitive `prim` as light indices (ascending; padding dropped): *count entries written to out_lights
 * (capacity num_lights).  *count = -1 when the primitive has no list (emissive, or lists not in use). */
int mcpt_scene_light_list_of_prim(const McptScene*, uint32_t prim, int32_t* out_lights, int32_t* count);
/* Host-only (no GPU): build the lists for `desc` and write them as a membership matrix, mask[prim * num_lights + light] =
 * 1 when the light is on the primitive's list (all 0 for emissive primitives).  out3 = {lists, longest padded list,
 * 1000 x mean fraction}; lists = 0: the builder declined (see above), mask untouched. */
int mcpt_host_light_lists(const McptSceneDesc* desc, uint8_t* mask, uint32_t out3[3]);
void mcpt_scene_destroy(McptScene* scene);
/* bytes of scene data resident on the device / re-upload them (used by the e2e timing) */
size_t mcpt_scene_device_bytes(const McptScene* scene);
int mcpt_scene_upload(McptScene* scene);

/* ---- render: the replacement for main.cpp:557-588 ----------------------------------------------
 * out_rgb: W*H*3 floats, row-major like buffer[i][j] (i = row), linear radiance (no tone mapping). */
int mcpt_render(McptScene* scene, const McptCamera* cam, const McptRenderParams* params, float* out_rgb,
                McptStats* stats);
/* Same, output left on the device (d_out_rgb: W*H*3 floats) and launched on `stream` (a cudaStream_t,
 * NULL = default) so callers can chain an NCCL reduce; returns after the work is enqueued AND complete. */
int mcpt_render_device(McptScene* scene, const McptCamera* cam, const McptRenderParams* params, float* d_out_rgb,
                       void* stream, McptStats* stats);
/* ---- multi-GPU behind the same ABI -------------------------------------------------
 * The loop shards by samples per pixel: device r of n renders ALL pixels for a contiguous share of the sample indices
 * [spp_offset, spp_offset + spp) (uniforms are keyed on the global index, so the film does not depend on n beyond the
 * FP32 order of one sum), normalised by spp_total.  One SAH BVH build; the packed scene is replicated to every device;
 * one host thread + stream + workspace per device; ONE collective -- ncclReduce(float32, sum, W*H*3) onto devices[0]
 * over NVLink -- then one device-to-host copy into out_rgb.  NCCL is bound at run time (dlopen libnccl.so.2); without
 * it these calls return MCPT_ERR_NCCL.  mcpt_multi_upload re-sends the scene blob to every device (e2e timing).
 * mcpt_render_multi = create + render + destroy, the one-call form of main.cpp:557-588 on n GPUs.
 * stats: counters summed over devices; ms_* = the slowest device (+ the reduce in ms_total), device-timed. */
typedef struct McptMulti McptMulti;
int mcpt_multi_create(const McptSceneDesc* desc, const int32_t* devices, int32_t n, uint32_t build_flags, McptMulti** out);
int mcpt_multi_device_count(const McptMulti*);
McptScene* mcpt_multi_scene(McptMulti*, int32_t r); /* replica r (owned by the McptMulti): introspection, mcpt_tonemap */
int mcpt_multi_upload(McptMulti*);
int mcpt_multi_render(McptMulti*, const McptCamera* cam, const McptRenderParams* params, float* out_rgb, McptStats* stats);
void mcpt_multi_destroy(McptMulti*);
int mcpt_render_multi(const McptSceneDesc* desc, const int32_t* devices, int32_t n, const McptCamera* cam,
                      const McptRenderParams* params, float* out_rgb, McptStats* stats);
/* RadianceRGB::tone_mapping (RadianceRGB.cpp:51-67) on the device: rgb8 = clamp(floor(255*(L/Lmax)^g+.5)) */
int mcpt_tonemap(McptScene* scene, const float* rgb, size_t num_pixels, double max_radiance, double gamma,
                 uint8_t* out_rgb8);

/* ---- deterministic parity entry points (explicit uniforms in, values out) ------------------------
 * precision: 0 = FP64 device code in the reference's expression order, 1 = the FP32 code the renderer runs. */
/* Myobj::intersect_with_triangle (Myobj.cpp:165-192) */
int mcpt_intersect_tri_batch(McptScene*, size_t n, const double* ro, const double* rd, const int32_t* prim,
                             int precision, int32_t* hit, double* bgt);
/* closest hit over all prims except `exclude` (Myobj::closet_ray_intersect, Myobj.cpp:334-474) or, with
 * light_only, over emissive prims (closet_ray_intersect_light_triangle, :476-622); semantics = exhaustive
 * search with strict < on t and ascending prim order (Myobj.cpp:194-213).  prim = -1 on miss.
 * precision 1 runs the renderer's own FP32 code (k_extend's stream traversal of the 4-wide BVH; with light_only the
 * trace_fast over the light BVH that k_shade calls for the MIS light pdf); precision 2 runs the wavefront's
 * MCPT_LEAF_EXACT kernel (k_extend_exact) on the FP32 roundings of the rays, as the render queues hold them. */
int mcpt_intersect_batch(McptScene*, size_t n, const double* ro, const double* rd, const int32_t* exclude,
                         int light_only, int precision, int32_t* prim, double* bgt);
/* The NEE visibility test of main.cpp:309-311 -- "the closest hit along wl IS the sampled light triangle" -- on explicit
 * rays (origin ro on prim `exclude`, unit direction rd aimed at light triangle `light_prim`), through the renderer's own
 * kernels: leaf_mode MCPT_LEAF_FAST = k_shadow (any-hit search for something that beats the light point under the
 * closest-hit order), MCPT_LEAF_EXACT = k_shadow_exact (closest hit, compare ids).  visible[i] = 1 / 0, or -1 when rd
 * misses the light triangle itself (the renderer queues no visibility ray then). */
int mcpt_visibility_batch(McptScene*, size_t n, const double* ro, const double* rd, const int32_t* exclude,
                          const int32_t* light_prim, int leaf_mode, int32_t* visible);
/* The same test for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  The call allocates nothing
 * and frees nothing -- no cudaMalloc / cudaFree, no allocating library routine: the scratch mcpt_visibility_batch allocates
 * per call (visibility-ray queue, per-ray accumulators, counters: 72 bytes per ray) lives in one fixed device array inside
 * the library, sized for MCPT_NOALLOC_MAX_RAYS rays (4.5 MiB on every GPU that loads libmcpt.so, until the process exits).
 * Inputs and outputs are the CALLER'S DEVICE buffers on the scene's device, in the layout the render queues use:
 *   d_org [n] float4 = origin xyz | unused,   d_dir [n] float4 = unit direction xyz | unused,
 *   d_ids [n] int2 = prim the origin lies on (excluded, -1 none) | light prim aimed at (must be < num_prims),
 *   d_visible [n] int32 out = 1 / 0 / -1 as above.
 * n > MCPT_NOALLOC_MAX_RAYS returns MCPT_ERR_INVALID (mcpt_last_error names the limit); n = 0 is a no-op.  The kernels are
 * launched on `stream` (a cudaStream_t; NULL = the default stream) and the call returns WITHOUT synchronising: the results
 * are there once the stream reaches that point.  All calls on one GPU share the scratch array, which keeps the previous
 * call's contents (the call clears what must start at zero itself): calls in one stream are ordered and safe, calls that
 * overlap on the GPU (two streams, two host threads) race and must be serialised by the caller.  The array's address is
 * resolved once per device (cudaGetSymbolAddress) on the first call for that device. */
#define MCPT_NOALLOC_MAX_RAYS 65536
int mcpt_visibility_noalloc(McptScene*, size_t n, const void* d_org, const void* d_dir, const void* d_ids, int leaf_mode,
                            void* d_visible, void* stream);
/* BRDF::get_brdf_phong / sample_from_phong / eval_sample_from_phong_pdf (BRDF.cpp:17-132) */
int mcpt_phong_eval_batch(McptScene*, size_t n, const double* N, const double* wi, const double* wr,
                          const double* kd, const double* ks, const double* ns, int precision, double* rgb);
int mcpt_phong_sample_batch(McptScene*, size_t n, const double* N, const double* wr, const double* kd,
                            const double* ks, const double* ns, const double* u3, int precision, double* dir,
                            double* pdf);
int mcpt_phong_pdf_batch(McptScene*, size_t n, const double* N, const double* wi, const double* wr,
                         const double* kd, const double* ks, const double* ns, int precision, double* pdf);
/* Mylight::select_a_point_from_lights (Mylight.cpp:102-160); u4 = (u_class, u_tri, xi1, xi2) consumed in the
 * reference's order (u_class / u_tri skipped when there is one class / one triangle) */
int mcpt_light_sample_uniform_batch(McptScene*, size_t n, const double* u4, int precision, int32_t* prim,
                                    double* point, double* Le, double* pdf);
/* prepared_for_lights_spherical_triangle_sampling + lights_spherical_triangle_sampling +
 * eval_spherical_triangle_sampling_pdf(query) (Mylight.cpp:322-493; one-shot variant :163-318 is the same
 * arithmetic).  u3 = (u_tri, xi1, xi2); out: chosen prim, point on the light, Le, pdf (solid angle),
 * weights_sum, number of surviving spherical triangles, pdf of `query` prim at this vertex. */
int mcpt_light_sample_spherical_batch(McptScene*, size_t n, const double* x1, const double* N, const double* u3,
                                      const int32_t* query, int precision, int32_t* prim, double* point, double* Le,
                                      double* pdf, double* wsum, int32_t* nsurv, double* query_pdf);

/* ---- introspection ---------------------------------------------------------------------------- */
/* geometric normals [P][3] and light areas [L] as computed at scene_create (FP64) */
int mcpt_scene_geom_normals(const McptScene*, double* out_P3);
int mcpt_scene_light_areas(const McptScene*, double* out_L);
/* BVH summary: nodes, leaves, max depth, SAH cost (for tests and DESIGN.md) */
int mcpt_scene_bvh_info(const McptScene*, int which /*0 scene, 1 lights*/, uint32_t* nodes, uint32_t* leaves,
                        uint32_t* max_depth, double* sah_cost);
/* structural check of the scene BVH the scene holds, both layouts (number of violations; 0 = sound) */
int mcpt_scene_bvh_check(const McptScene*, uint32_t* errors);
/* Host-only (no GPU needed): build the SAH BVH for `desc` (all prims, or the emissive ones) and validate it:
 * every prim appears in exactly one leaf, every triangle lies inside its leaf box, every child box inside its
 * parent's.  out4 = {nodes, leaves, max_depth, errors}. */
int mcpt_host_bvh_check(const McptSceneDesc* desc, int lights_only, uint32_t out4[4], double* sah_cost);
/* Light hierarchy of MCPT_LIGHT_TREE.  Host-only structural check (every light in exactly one leaf, every ancestor's
 * sphere / normal cone holds the light): out4 = {nodes, max_depth, errors, lights}. */
int mcpt_host_light_tree_check(const McptSceneDesc* desc, uint32_t out4[4]);
/* The hierarchy a McptScene holds: node count and depth; and the nodes themselves in the device layout of
 * csrc/light_tree.h (12 floats per node: centre | radius, cone axis | cos theta_o, power | left | right | parent as int
 * bits) with the leaf node of every light -- what a host-side restatement of the walk needs (tests/test_gpu_parity.py). */
int mcpt_scene_light_tree_info(const McptScene*, uint32_t* nodes, uint32_t* max_depth);
int mcpt_scene_light_tree_nodes(const McptScene*, float* out_nodes12, int32_t* out_leaf_of_light);
/* Parity / test entry point: at vertex (x1, N) draw a light with uniform u (prim, selection probability) and report
 * the selection probability of every light in `query_light` [n] (index into the light list, -1 = none) together
 * with that light's FP32 spherical-triangle weight sA * (R+G+B) (0 = culled at this vertex). */
int mcpt_light_tree_batch(McptScene*, size_t n, const double* x1, const double* N, const double* u, const int32_t* query_light,
                          int32_t* light, double* prob, double* query_prob, double* query_weight);
/* FP32 multiply-add ceiling of `device`, measured with register-resident FMA chains (no memory traffic), in TFLOP/s:
 * three-register scalar FFMA and the packed FFMA2 form the light sweeps use.  bench.py's fp32 roofline denominator. */
int mcpt_measure_fp32(int device, double* tflops_scalar, double* tflops_packed);
const char* mcpt_last_error(void);
int mcpt_version(void);
int mcpt_device_count(void);

/* ---- host loaders: the host-side mirror of Myobj / Mylight (csrc/host_scene.h) -------------------
 * Parses <stem>.obj (+ its mtllib) and <stem>.xml like Myobj::read / Mylight::read +
 * gather_light_triangles (Myobj.cpp:10-28, Mylight.cpp:11-100) and hands back a flat desc whose arrays
 * live until mcpt_host_scene_free.  has_camera = 0 if the XML has no <camera> node (README.md:339-343). */
typedef struct McptHostScene McptHostScene;
int mcpt_host_scene_load(const char* obj_path, const char* xml_path, McptHostScene** out);
const McptSceneDesc* mcpt_host_scene_desc(const McptHostScene*);
int mcpt_host_scene_camera(const McptHostScene*, McptCamera* cam, int32_t* width, int32_t* height, double* fovy_deg);
void mcpt_host_scene_free(McptHostScene*);

#ifdef __cplusplus
}
#endif
#endif /* MCPT_H */
