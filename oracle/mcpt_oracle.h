/* oracle/mcpt_oracle.h -- TEST INFRASTRUCTURE, not product code.
 *
 * CPU restatement, in plain C and FP64, of the reference's per-pixel path-tracing loop
 * (luotong96/Monte_Carlo_Path_Tracing: main.cpp:269-494,546-588; Myobj.cpp:78-192,334-709;
 * Mylight.cpp:32-493; BRDF.cpp:17-132; RadianceRGB.cpp:51-67).  Every function cites the
 * reference lines it follows and keeps the reference's expression order so that, fed the same
 * uniform stream (oracle/stream.h), it reproduces the compiled reference (oracle/_ref, built from
 * the unmodified sources) value for value.  tests/test_oracle_pin.py checks exactly that, and
 * tests/golden/ holds vectors produced by oracle/_ref for boxes where /root/reference is absent.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm may load this.
 * The product (monte_carlo_path_tracing_b200/libmcpt.so) never links or calls it.
 */
#ifndef MCPT_ORACLE_H
#define MCPT_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct OracleScene OracleScene;

/* estimator selection (which shade* body, main.cpp:269/348/402) */
enum {
    ORACLE_NEE_SPHERICAL = 0, /* shade() as checked in: select_a_point_from_lights_spherical_triangle (main.cpp:298) */
    ORACLE_BRDF_ONLY = 1,     /* shade_with_brdf() */
    ORACLE_MIS = 2,           /* shade_with_mis() */
    ORACLE_NEE_UNIFORM = 3    /* shade() with the commented-out line main.cpp:297 instead of :298 */
};

/* behaviour switches; 0 == the reference exactly as written */
enum {
    ORACLE_FIX_STALE_MIS = 1,      /* F10: save/restore the Mylight scratch around recursion #1 (main.cpp:464) */
    ORACLE_FIX_NEE_MEASURE = 2,    /* F9: NEE_SPHERICAL uses the solid-angle estimator Le f cos / pdf */
    ORACLE_FIX_MIXTURE_PDF = 4,    /* F8: BRDF-sampled directions use the mixture pdf (BRDF.cpp:107-132) */
    ORACLE_BRUTE_FORCE = 8,        /* closest hit by exhaustive search (Myobj.cpp:194-213) instead of the grid */
    ORACLE_RNG_KEYED = 16          /* uniforms from the keyed Philox table of libmcpt's renderer (stream.h) instead of
                                      the sequential stream: the oracle then walks the same paths as an
                                      MCPT_PARITY_FP64 render with the same seed */
};

/* Flat scene, same layout as McptSceneDesc (include/mcpt.h).  Arrays are copied. */
OracleScene* oracle_scene_create(int P, int S, int M, int L, int C, const float* positions, const float* normals,
                                 const int* material_id, const unsigned* shape_offset, const float* kd,
                                 const float* ks, const float* ns, const int* light_prim, const int* light_class,
                                 const double* class_radiance);
void oracle_scene_destroy(OracleScene*);

/* cal_scene_boundingbox + meshing (Myobj.cpp:78-162).  Returns cells per axis used (max 103, Myobj.h:76). */
int oracle_build_grid(OracleScene*, const double eye[3], int n0);

/* get_unique_normal_of_facet for every prim (Myobj.cpp:680-709), light areas (Mylight.cpp:66-69) */
void oracle_geom_normals(OracleScene*, double* out_P3);
void oracle_light_areas(OracleScene*, double* out_L);

void oracle_intersect_tri(OracleScene*, int n, const double* ro, const double* rd, const int* prim, int* hit,
                          double* bgt);
/* mode 0 grid (Myobj.cpp:334-474), 1 grid light-only (:476-622), 2 brute force, 3 brute force light-only */
void oracle_closest(OracleScene*, int n, const double* ro, const double* rd, const int* exclude, int mode, int* prim,
                    double* bgt);
void oracle_brdf_eval(int n, const double* N, const double* wi, const double* wr, const double* kd, const double* ks,
                      const double* ns, double* rgb);
void oracle_brdf_sample(int n, const double* N, const double* wr, const double* kd, const double* ks, const double* ns,
                        const double* u3, double* dir, double* pdf);
void oracle_brdf_pdf(int n, const double* N, const double* wi, const double* wr, const double* kd, const double* ks,
                     const double* ns, double* pdf);
void oracle_light_sample_uniform(OracleScene*, int n, const double* u4, int* prim, double* point, double* Le,
                                 double* pdf);
void oracle_light_sample_spherical(OracleScene*, int n, int staged, const double* x1, const double* N, const double* u3,
                                   const int* query, int* prim, double* point, double* Le, double* pdf, double* wsum,
                                   int* nsurv, double* query_pdf);
void oracle_tonemap(int n, const double* rgb, double max_radiance, double gamma, int* out);

/* row (node, call) of the keyed uniform table of ORACLE_RNG_KEYED (stream.h) */
void oracle_keyed_uniforms(uint64_t seed, uint32_t pixel, uint32_t sample, uint64_t node, uint32_t call, double out[4]);

/* main.cpp:546-588 with literals as arguments; same signature as ref_render + flags.
 * counters[0] non-emissive vertices shaded, [1] primary hits, [2] closest-hit rays, [3] light-only rays */
void oracle_render(OracleScene*, const double eye[3], const double w[3], const double up[3], double pixellen, int W,
                   int H, int i0, int i1, int j0, int j1, int spp, int spp_offset, int spp_total, int strategy,
                   int max_depth, int flags, uint64_t seed, double* out, uint64_t* counters);

#ifdef __cplusplus
}
#endif
#endif
