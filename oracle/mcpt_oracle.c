/* oracle/mcpt_oracle.c -- TEST INFRASTRUCTURE, not product code.  See mcpt_oracle.h.
 *
 * Plain-C FP64 restatement of the reference algorithm.  Expression order follows the reference
 * so results match the compiled reference bit for bit on x86-64 without FMA contraction
 * (-ffp-contract=off); citations are file:line into /root/reference.
 */
#include "mcpt_oracle.h"
#include "stream.h"
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define EPS 1e-8                    /* vec.h:7, Myobj.h:69, Mylight.h:101, BRDF.cpp:8 */
#define PI 3.14159265358979323846   /* std::numbers::pi */
#define P_RR 0.6                    /* main.cpp:321,375,429 */

/* ---- vec (vec.cpp) -------------------------------------------------------------------------- */
typedef struct { double x[3]; } v3;
static inline v3 V(double a, double b, double c) { v3 r = {{a, b, c}}; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.x[0] + b.x[0], a.x[1] + b.x[1], a.x[2] + b.x[2]); }  /* vec.cpp:38 */
static inline v3 vsub(v3 a, v3 b) { return V(a.x[0] - b.x[0], a.x[1] - b.x[1], a.x[2] - b.x[2]); }  /* vec.cpp:43 */
static inline v3 vmul(v3 a, double c) { return V(a.x[0] * c, a.x[1] * c, a.x[2] * c); }              /* vec.cpp:49 */
static inline double vdot(v3 a, v3 b)                                                                /* vec.cpp:73-81 */
{
    double ans = 0;
    for (int i = 0; i < 3; i++) ans += a.x[i] * b.x[i];
    return ans;
}
static inline v3 vcross(v3 a, v3 b)                                                                  /* vec.cpp:67-70 */
{
    return V(a.x[1] * b.x[2] - a.x[2] * b.x[1], a.x[2] * b.x[0] - a.x[0] * b.x[2], a.x[0] * b.x[1] - a.x[1] * b.x[0]);
}
static inline double vnorm(v3 a) { return sqrt(a.x[0] * a.x[0] + a.x[1] * a.x[1] + a.x[2] * a.x[2]); } /* vec.cpp:55 */
static inline v3 vnormalized(v3 a)                                                                   /* vec.cpp:99-103 */
{
    double len = vnorm(a);
    return V(a.x[0] / len, a.x[1] / len, a.x[2] / len);
}
static inline double vdet(v3 a, v3 b, v3 c) { return vdot(vcross(a, b), c); }                        /* vec.cpp:84-87 */
/* matrix3d(v1,v2,v3) * x with columns v1,v2,v3 (matrix3d.cpp:8-22,29-40) */
static inline v3 mat_cols_mul(v3 c0, v3 c1, v3 c2, v3 x)
{
    v3 b = V(0, 0, 0);
    for (int i = 0; i < 3; i++) {
        b.x[i] += c0.x[i] * x.x[0];
        b.x[i] += c1.x[i] * x.x[1];
        b.x[i] += c2.x[i] * x.x[2];
    }
    return b;
}
/* linear_interpolation (main.cpp:26-29) */
static inline v3 lerp3(const v3* v, double a, double b, double c) { return vadd(vadd(vmul(v[0], a), vmul(v[1], b)), vmul(v[2], c)); }

/* ---- scene ---------------------------------------------------------------------------------- */
typedef struct { double beta, gamma, t; int hit; int prim; } hit_t; /* intersec_result, Myobj.h:10-41 */

#define GRID_CAP 103 /* Myobj.h:76 */
typedef struct { int* tris; int n, cap; } cell_t;

struct OracleScene {
    int P, S, M, L, C;
    v3* pos;      /* [P][3] double copies of the float vertices (tinyobj real_t = float, F2) */
    v3* nor;      /* [P][3] */
    int* matid;
    unsigned* shape_offset;
    v3 *kd, *ks;
    double* ns;
    int* light_prim;   /* [L] reference order: class (mtlname asc), then (s,f) asc */
    int* light_class;  /* [L] */
    double* light_area;
    v3* class_rad;     /* [C] */
    int* class_begin;  /* [C+1] ranges into light_prim */
    int* light_of_prim; /* [P] index into light arrays or -1 (Mylight::islight) */
    v3* gnormal;       /* [P] lazily = get_unique_normal_of_facet */
    unsigned char* gnormal_ok;
    /* grid (Myobj.h:76-81) */
    cell_t* cells;
    double xyzmm[3][2];
    double cellw;
    /* per-ray mailbox replacing std::map<triangle,intersec_result> intersecMap (Myobj.cpp:395) */
    unsigned* mail_stamp;
    hit_t* mail_hit;
    unsigned stamp;
    /* Mylight scratch (Mylight.h:115-120): survivors of the last prepare, in light order */
    int* st_light;     /* light index of each surviving spherical triangle */
    v3 *st_A, *st_B, *st_C;
    double *st_alpha, *st_c, *st_sA, *st_w;
    int st_n;
    double st_wsum;
    unsigned char* st_in; /* [L] indiceMap membership */
    /* run state */
    oracle_stream stream;
    int max_depth, flags;
    uint64_t n_vertices, n_rays, n_light_rays;
    /* ORACLE_RNG_KEYED: the (seed, pixel, sample) of the path being traced and the table row handed to the stream */
    uint64_t k_seed;
    uint32_t k_pixel, k_sample;
    double k_buf[4];
};

/* ORACLE_RNG_KEYED: point the stream at positions [first, first + n) of table row (node, call) */
static void keyed_list(OracleScene* s, uint64_t node, uint32_t call, int first, int n)
{
    if (!(s->flags & ORACLE_RNG_KEYED)) return;
    oracle_keyed_draw4(s->k_seed, s->k_pixel, s->k_sample, node, call, s->k_buf);
    oracle_stream_list(&s->stream, s->k_buf + first, (size_t)n);
    s->stream.keyed = 1;
}

OracleScene* oracle_scene_create(int P, int S, int M, int L, int C, const float* positions, const float* normals,
                                 const int* material_id, const unsigned* shape_offset, const float* kd,
                                 const float* ks, const float* ns, const int* light_prim, const int* light_class,
                                 const double* class_radiance)
{
    OracleScene* s = (OracleScene*)calloc(1, sizeof(*s));
    s->P = P; s->S = S; s->M = M; s->L = L; s->C = C;
    s->pos = (v3*)malloc(sizeof(v3) * 3 * (size_t)P);
    s->nor = (v3*)malloc(sizeof(v3) * 3 * (size_t)P);
    for (size_t i = 0; i < (size_t)P * 3; i++) {
        s->pos[i] = V(positions[i * 3], positions[i * 3 + 1], positions[i * 3 + 2]);
        s->nor[i] = V(normals[i * 3], normals[i * 3 + 1], normals[i * 3 + 2]);
    }
    s->matid = (int*)malloc(sizeof(int) * (size_t)P);
    memcpy(s->matid, material_id, sizeof(int) * (size_t)P);
    s->shape_offset = (unsigned*)malloc(sizeof(unsigned) * (size_t)(S + 1));
    memcpy(s->shape_offset, shape_offset, sizeof(unsigned) * (size_t)(S + 1));
    s->kd = (v3*)malloc(sizeof(v3) * (size_t)M);
    s->ks = (v3*)malloc(sizeof(v3) * (size_t)M);
    s->ns = (double*)malloc(sizeof(double) * (size_t)M);
    for (int m = 0; m < M; m++) {
        s->kd[m] = V(kd[m * 3], kd[m * 3 + 1], kd[m * 3 + 2]);
        s->ks[m] = V(ks[m * 3], ks[m * 3 + 1], ks[m * 3 + 2]);
        s->ns[m] = ns[m];
    }
    s->light_prim = (int*)malloc(sizeof(int) * (size_t)(L + 1));
    s->light_class = (int*)malloc(sizeof(int) * (size_t)(L + 1));
    s->light_area = (double*)malloc(sizeof(double) * (size_t)(L + 1));
    s->light_of_prim = (int*)malloc(sizeof(int) * (size_t)P);
    for (int p = 0; p < P; p++) s->light_of_prim[p] = -1;
    s->class_rad = (v3*)malloc(sizeof(v3) * (size_t)(C + 1));
    s->class_begin = (int*)calloc((size_t)C + 2, sizeof(int));
    for (int c = 0; c < C; c++) s->class_rad[c] = V(class_radiance[c * 3], class_radiance[c * 3 + 1], class_radiance[c * 3 + 2]);
    for (int l = 0; l < L; l++) {
        s->light_prim[l] = light_prim[l];
        s->light_class[l] = light_class[l];
        s->light_of_prim[light_prim[l]] = l;
        s->class_begin[light_class[l] + 1] = l + 1;
        /* area, Mylight.cpp:66-69 */
        const v3* abc = s->pos + 3 * (size_t)light_prim[l];
        v3 n = vcross(vsub(abc[1], abc[0]), vsub(abc[2], abc[0]));
        n = vmul(n, 1.0 / vnorm(n));
        s->light_area[l] = 0.5 * vdet(vsub(abc[1], abc[0]), vsub(abc[2], abc[0]), n);
    }
    for (int c = 1; c <= C; c++)
        if (s->class_begin[c] < s->class_begin[c - 1]) s->class_begin[c] = s->class_begin[c - 1];
    s->gnormal = (v3*)malloc(sizeof(v3) * (size_t)P);
    s->gnormal_ok = (unsigned char*)calloc((size_t)P, 1);
    s->mail_stamp = (unsigned*)calloc((size_t)P, sizeof(unsigned));
    s->mail_hit = (hit_t*)malloc(sizeof(hit_t) * (size_t)P);
    size_t Ln = (size_t)L + 1;
    s->st_light = (int*)malloc(sizeof(int) * Ln);
    s->st_A = (v3*)malloc(sizeof(v3) * Ln); s->st_B = (v3*)malloc(sizeof(v3) * Ln); s->st_C = (v3*)malloc(sizeof(v3) * Ln);
    s->st_alpha = (double*)malloc(sizeof(double) * Ln); s->st_c = (double*)malloc(sizeof(double) * Ln);
    s->st_sA = (double*)malloc(sizeof(double) * Ln); s->st_w = (double*)malloc(sizeof(double) * Ln);
    s->st_in = (unsigned char*)calloc(Ln, 1);
    return s;
}

void oracle_scene_destroy(OracleScene* s)
{
    if (!s) return;
    if (s->cells) {
        for (size_t i = 0; i < (size_t)GRID_CAP * GRID_CAP * GRID_CAP; i++) free(s->cells[i].tris);
        free(s->cells);
    }
    free(s->pos); free(s->nor); free(s->matid); free(s->shape_offset); free(s->kd); free(s->ks); free(s->ns);
    free(s->light_prim); free(s->light_class); free(s->light_area); free(s->light_of_prim); free(s->class_rad);
    free(s->class_begin); free(s->gnormal); free(s->gnormal_ok); free(s->mail_stamp); free(s->mail_hit);
    free(s->st_light); free(s->st_A); free(s->st_B); free(s->st_C); free(s->st_alpha); free(s->st_c); free(s->st_sA);
    free(s->st_w); free(s->st_in);
    free(s);
}

/* get_unique_normal_of_facet, Myobj.cpp:680-709 (ties choose -n, :702-708) */
static v3 geom_normal(OracleScene* s, int prim)
{
    if (s->gnormal_ok[prim]) return s->gnormal[prim];
    const v3* abc = s->pos + 3 * (size_t)prim;
    const v3* nabc = s->nor + 3 * (size_t)prim;
    v3 na = vnormalized(nabc[0]), nb = vnormalized(nabc[1]), nc = vnormalized(nabc[2]);
    v3 n = vnormalized(vcross(vsub(abc[1], abc[0]), vsub(abc[2], abc[0])));
    v3 nr = vmul(n, -1);
    double w = vdot(n, na) + vdot(n, nb) + vdot(n, nc);
    double wr = vdot(nr, na) + vdot(nr, nb) + vdot(nr, nc);
    v3 r = (w > wr) ? n : nr;
    s->gnormal[prim] = r;
    s->gnormal_ok[prim] = 1;
    return r;
}

void oracle_geom_normals(OracleScene* s, double* out)
{
    for (int p = 0; p < s->P; p++) {
        v3 g = geom_normal(s, p);
        out[3 * p] = g.x[0]; out[3 * p + 1] = g.x[1]; out[3 * p + 2] = g.x[2];
    }
}
void oracle_light_areas(OracleScene* s, double* out) { memcpy(out, s->light_area, sizeof(double) * (size_t)s->L); }

/* ---- grid build: cal_scene_boundingbox (Myobj.cpp:78-107) + meshing (:110-162) --------------- */
static inline cell_t* cell_at(OracleScene* s, int i, int j, int k)
{
    return &s->cells[((size_t)i * GRID_CAP + (size_t)j) * GRID_CAP + (size_t)k];
}

int oracle_build_grid(OracleScene* s, const double eye[3], int n0)
{
    if (!s->cells) s->cells = (cell_t*)calloc((size_t)GRID_CAP * GRID_CAP * GRID_CAP, sizeof(cell_t));
    for (int i = 0; i < 3; i++) s->xyzmm[i][0] = s->xyzmm[i][1] = eye[i];
    for (size_t q = 0; q < (size_t)s->P * 3; q++)
        for (int i = 0; i < 3; i++) {
            s->xyzmm[i][0] = fmin(s->xyzmm[i][0], s->pos[q].x[i]);
            s->xyzmm[i][1] = fmax(s->xyzmm[i][1], s->pos[q].x[i]);
        }
    double len[3];
    for (int i = 0; i < 3; i++) len[i] = s->xyzmm[i][1] - s->xyzmm[i][0];
    double d = fmax(fmax(len[0], len[1]), len[2]) / pow(n0, 1.0 / 3);
    s->cellw = d;
    int maxidx = 0;
    for (int p = 0; p < s->P; p++) {
        double xyz[3][2];
        for (int k = 0; k < 3; k++) { xyz[k][0] = DBL_MAX; xyz[k][1] = -DBL_MAX; }
        for (int v = 0; v < 3; v++)
            for (int i = 0; i < 3; i++) {
                xyz[i][0] = fmin(xyz[i][0], s->pos[3 * (size_t)p + v].x[i]);
                xyz[i][1] = fmax(xyz[i][1], s->pos[3 * (size_t)p + v].x[i]);
            }
        int xyzi[3][2];
        for (int k = 0; k < 3; k++) {
            xyzi[k][0] = (int)floor((xyz[k][0] - s->xyzmm[k][0]) / d);
            xyzi[k][1] = (int)floor((xyz[k][1] - s->xyzmm[k][0]) / d);
            if (xyzi[k][1] > maxidx) maxidx = xyzi[k][1];
        }
        for (int i = xyzi[0][0]; i <= xyzi[0][1]; i++)
            for (int j = xyzi[1][0]; j <= xyzi[1][1]; j++)
                for (int k = xyzi[2][0]; k <= xyzi[2][1]; k++) {
                    if (i < 0 || j < 0 || k < 0 || i >= GRID_CAP || j >= GRID_CAP || k >= GRID_CAP) continue;
                    cell_t* c = cell_at(s, i, j, k);
                    if (c->n == c->cap) {
                        c->cap = c->cap ? c->cap * 2 : 4;
                        c->tris = (int*)realloc(c->tris, sizeof(int) * (size_t)c->cap);
                    }
                    c->tris[c->n++] = p;
                }
    }
    return maxidx + 1;
}

/* ---- Myobj::intersect_with_triangle, Myobj.cpp:165-192 ---------------------------------------- */
static hit_t intersect_tri(const OracleScene* s, v3 ro, v3 rd, int prim)
{
    hit_t r = {0, 0, 0, 0, prim};
    const v3* abc = s->pos + 3 * (size_t)prim;
    v3 a = abc[0], b = abc[1], c = abc[2];
    double detA = vdet(vsub(a, b), vsub(a, c), rd);
    if (fabs(detA) < EPS) return r;
    double beta = vdet(vsub(a, ro), vsub(a, c), rd) / detA;
    double gamma = vdet(vsub(a, b), vsub(a, ro), rd) / detA;
    double t = vdet(vsub(a, b), vsub(a, c), vsub(a, ro)) / detA;
    if (beta < 0 || gamma < 0 || beta + gamma > 1 || t < 0 || fabs(t) < EPS) return r;
    r.hit = 1; r.beta = beta; r.gamma = gamma; r.t = t;
    return r;
}

/* exhaustive search = the commented-out definition Myobj.cpp:194-213 (strict <, (s,f) ascending),
 * with the exclude-id rule of Myobj.cpp:417 and, for light_only, the skip of Myobj.cpp:559 */
static hit_t closest_brute(OracleScene* s, v3 ro, v3 rd, int exclude, int light_only)
{
    hit_t ans = {0, 0, DBL_MAX, 0, 0};
    if (light_only) {
        /* light prims ascending == (s,f) ascending */
        for (int p = 0; p < s->P; p++) {
            if (s->light_of_prim[p] < 0 || p == exclude) continue;
            hit_t rs = intersect_tri(s, ro, rd, p);
            if (rs.hit && rs.t < ans.t) ans = rs;
        }
        return ans;
    }
    for (int p = 0; p < s->P; p++) {
        if (p == exclude) continue;
        hit_t rs = intersect_tri(s, ro, rd, p);
        if (rs.hit && rs.t < ans.t) ans = rs;
    }
    return ans;
}

/* ---- grid traversal: closet_ray_intersect (Myobj.cpp:334-474) and
 *      closet_ray_intersect_light_triangle (Myobj.cpp:476-622) ------------------------------------
 * The two differ in (a) the light-only skip (:559) and (b) how the DDA advances: the first steps
 * the single axis with the smallest ts (:452-470), the second steps every axis whose ts is within
 * eps of the minimum (:597-619). */
static hit_t closest_grid(OracleScene* s, v3 ro, v3 rd, int exclude, int light_only)
{
    v3 org = V(s->xyzmm[0][0], s->xyzmm[1][0], s->xyzmm[2][0]);
    v3 xyz0 = vmul(vsub(ro, org), 1.0 / s->cellw);
    v3 abc = rd;
    int xyz[3], sign[3], nxyz[3];
    double ts[3], len[3];
    for (int i = 0; i < 3; i++) xyz[i] = (int)floor(xyz0.x[i]);
    for (int i = 0; i < 3; i++) {
        sign[i] = abc.x[i] < 0 ? -1 : 1;
        if (fabs(abc.x[i]) < EPS) sign[i] = 0;
    }
    for (int i = 0; i < 3; i++) {
        if (sign[i]) {
            if (fabs(floor(xyz0.x[i]) - xyz0.x[i]) < EPS) nxyz[i] = xyz[i] + sign[i];
            else nxyz[i] = sign[i] > 0 ? (int)ceil(xyz0.x[i]) : (int)floor(xyz0.x[i]);
            ts[i] = (nxyz[i] - xyz0.x[i]) / abc.x[i];
        } else {
            nxyz[i] = -1;
            ts[i] = DBL_MAX;
        }
    }
    for (int i = 0; i < 3; i++) len[i] = s->xyzmm[i][1] - s->xyzmm[i][0];

    if (++s->stamp == 0) { memset(s->mail_stamp, 0, sizeof(unsigned) * (size_t)s->P); s->stamp = 1; }
    hit_t ans = {0, 0, DBL_MAX, 0, 0};
    for (;;) {
        int reachbound = 0;
        for (int i = 0; i < 3; i++)
            if (xyz[i] < 0 || xyz[i] > (int)floor(len[i] / s->cellw) + 2) { reachbound = 1; break; }
        if (reachbound) break;
        if (xyz[0] < GRID_CAP && xyz[1] < GRID_CAP && xyz[2] < GRID_CAP) {
            cell_t* c = cell_at(s, xyz[0], xyz[1], xyz[2]);
            for (int q = 0; q < c->n; q++) {
                int tri = c->tris[q];
                if (light_only && s->light_of_prim[tri] < 0) continue; /* :559 */
                if (tri == exclude) continue;                          /* :417 / :562 */
                if (s->mail_stamp[tri] != s->stamp) {                  /* :420-423 */
                    s->mail_stamp[tri] = s->stamp;
                    s->mail_hit[tri] = intersect_tri(s, ro, rd, tri);
                }
                hit_t rs = s->mail_hit[tri];
                if (rs.hit) {
                    /* :429-441 accept only hits whose point lies in the current cell */
                    v3 cp = vmul(vsub(vadd(ro, vmul(rd, rs.t)), org), 1.0 / s->cellw);
                    int out = 0;
                    for (int i = 0; i < 3; i++)
                        if ((int)floor(cp.x[i]) != xyz[i]) { out = 1; break; }
                    if (out) continue;
                    if (rs.t < ans.t) ans = rs;
                }
            }
        }
        if (ans.hit) return ans;
        double t = DBL_MAX;
        int ind = -1;
        if (!light_only) {
            for (int i = 0; i < 3; i++) {
                if (sign[i] == 0) continue;
                if (ts[i] < t) { ind = i; t = ts[i]; }
            }
            if (ind < 0) break; /* the reference prints "464!!!!" and indexes xyz[-1] (UB); we stop */
            xyz[ind] += sign[ind];
            nxyz[ind] += sign[ind];
            ts[ind] = (nxyz[ind] - xyz0.x[ind]) / abc.x[ind];
        } else {
            for (int i = 0; i < 3; i++)
                if (ts[i] < t) { ind = i; t = ts[i]; }
            if (t == DBL_MAX) break; /* "612!!!!" -- would loop forever in the reference */
            for (int i = 0; i < 3; i++)
                if (fabs(t - ts[i]) < EPS) {
                    xyz[i] += sign[i];
                    nxyz[i] += sign[i];
                    ts[i] = (nxyz[i] - xyz0.x[i]) / abc.x[i];
                }
        }
    }
    return ans;
}

static hit_t closest(OracleScene* s, v3 ro, v3 rd, int exclude)
{
    s->n_rays++;
    return (s->flags & ORACLE_BRUTE_FORCE) ? closest_brute(s, ro, rd, exclude, 0) : closest_grid(s, ro, rd, exclude, 0);
}
static hit_t closest_light(OracleScene* s, v3 ro, v3 rd, int exclude)
{
    s->n_light_rays++;
    return (s->flags & ORACLE_BRUTE_FORCE) ? closest_brute(s, ro, rd, exclude, 1) : closest_grid(s, ro, rd, exclude, 1);
}

void oracle_intersect_tri(OracleScene* s, int n, const double* ro, const double* rd, const int* prim, int* hit, double* bgt)
{
    for (int i = 0; i < n; i++) {
        hit_t r = intersect_tri(s, V(ro[3 * i], ro[3 * i + 1], ro[3 * i + 2]), V(rd[3 * i], rd[3 * i + 1], rd[3 * i + 2]), prim[i]);
        hit[i] = r.hit;
        bgt[3 * i] = r.beta; bgt[3 * i + 1] = r.gamma; bgt[3 * i + 2] = r.t;
    }
}

void oracle_closest(OracleScene* s, int n, const double* ro, const double* rd, const int* exclude, int mode, int* prim, double* bgt)
{
    for (int i = 0; i < n; i++) {
        v3 o = V(ro[3 * i], ro[3 * i + 1], ro[3 * i + 2]), d = V(rd[3 * i], rd[3 * i + 1], rd[3 * i + 2]);
        int ex = exclude ? exclude[i] : -1;
        hit_t r = mode == 0 ? closest_grid(s, o, d, ex, 0) : mode == 1 ? closest_grid(s, o, d, ex, 1)
                : mode == 2 ? closest_brute(s, o, d, ex, 0) : closest_brute(s, o, d, ex, 1);
        prim[i] = r.hit ? r.prim : -1;
        bgt[3 * i] = r.hit ? r.beta : 0; bgt[3 * i + 1] = r.hit ? r.gamma : 0; bgt[3 * i + 2] = r.hit ? r.t : 0;
    }
}

/* ---- libstdc++ std::discrete_distribution (bits/random.tcc:2657-2714) --------------------------
 * <2 weights: no draw, index 0, probability 1.  Otherwise p_i = w_i / sum, cumulative partial sums
 * with the last forced to 1.0, ONE canonical draw u, result = first index with cp >= u. */
static int discrete_pick(oracle_stream* st, const double* w, int n, double* prob_out)
{
    if (n < 2) { if (st->keyed) st->list_i++; if (prob_out) *prob_out = 1.0; return 0; }
    double sum = 0.0;
    for (int i = 0; i < n; i++) sum += w[i];
    double u = oracle_stream_next(st);
    double cp = 0.0;
    int idx = n - 1;
    for (int i = 0; i < n; i++) {
        double p = w[i] / sum;
        cp = (i == 0) ? p : cp + p;
        double c = (i == n - 1) ? 1.0 : cp;
        if (c >= u) { idx = i; break; }
    }
    if (prob_out) *prob_out = w[idx] / sum;
    return idx;
}

/* ---- BRDF (BRDF.cpp) -------------------------------------------------------------------------- */
/* get_brdf_phong, BRDF.cpp:17-25 */
static v3 brdf_phong(v3 n, v3 wi, v3 wr, v3 kd, v3 ks, double ns)
{
    v3 R = vadd(vmul(wi, -1), vmul(n, 2 * vdot(wi, n)));
    v3 ans = vmul(kd, 1.0 / PI);
    if (vdot(wr, R) > 0) ans = vadd(ans, vmul(ks, (ns + 1) * pow(vdot(wr, R), ns) / (2 * PI)));
    return ans;
}

/* tangent frame used by both lobes, BRDF.cpp:58-71 / :86-99 (F11: NaN for axis == (-1,0,0)) */
static v3 frame_dir(v3 axis, double theta, double phi)
{
    v3 nx;
    if (fabs(vdot(axis, V(1, 0, 0)) - 1) > EPS) nx = vnormalized(vcross(axis, V(1, 0, 0)));
    else nx = vnormalized(vcross(axis, V(0, 1, 0)));
    v3 ny = vnormalized(vcross(axis, nx));
    return vnormalized(mat_cols_mul(nx, ny, axis, V(sin(theta) * cos(phi), sin(theta) * sin(phi), cos(theta))));
}

/* sample_from_phong, BRDF.cpp:28-104: draws u_lobe, xi1, xi2; returns the SINGLE-LOBE pdf (F8) */
static v3 brdf_sample(oracle_stream* st, v3 n, v3 wr, v3 kd, v3 ks, double ns, double* pdf_out)
{
    double d = vdot(kd, V(1, 1, 1)) / 3;
    double s = vdot(ks, V(1, 1, 1)) / 3;
    double pdf = 1;
    double w[2] = {d, s}, pl;
    int ind = discrete_pick(st, w, 2, &pl);
    pdf *= pl;
    double ksi1 = oracle_stream_next(st);
    double ksi2 = oracle_stream_next(st);
    if (ind == 0) {
        double theta = 0.5 * acos(fmax(-1, fmin(1, 1 - 2 * ksi1)));
        double phi = 2 * PI * ksi2;
        pdf *= cos(theta) / PI;
        *pdf_out = pdf;
        return frame_dir(n, theta, phi);
    }
    double theta = acos(fmax(-1, fmin(1, pow(ksi1, 1 / (ns + 1)))));
    double phi = 2 * PI * ksi2;
    pdf *= (ns + 1) / (2 * PI) * pow(ksi1, ns / (ns + 1));
    v3 R = vnormalized(vadd(vmul(wr, -1), vmul(n, 2 * vdot(wr, n))));
    *pdf_out = pdf;
    return frame_dir(R, theta, phi);
}

/* eval_sample_from_phong_pdf, BRDF.cpp:107-132 (mixture pdf) */
static double brdf_pdf(v3 n, v3 wi, v3 wr, v3 kd, v3 ks, double ns)
{
    double d = vdot(kd, V(1, 1, 1)) / 3;
    double s = vdot(ks, V(1, 1, 1)) / 3;
    double sum = d + s;
    double prob_d = d / sum, prob_s = s / sum;
    double cosTheta = vdot(wi, n);
    if (cosTheta < 0) prob_d *= 0; else prob_d *= cosTheta / PI;
    v3 R = vnormalized(vadd(vmul(wr, -1), vmul(n, 2 * vdot(wr, n))));
    double cosThetaS = vdot(wi, R);
    if (cosThetaS < 0) prob_s *= 0; else prob_s *= (ns + 1) / (2 * PI) * pow(cosThetaS, ns);
    return prob_d + prob_s;
}

static inline v3 P3(const double* p) { return V(p[0], p[1], p[2]); }
static inline void S3(double* o, v3 v) { o[0] = v.x[0]; o[1] = v.x[1]; o[2] = v.x[2]; }

void oracle_brdf_eval(int n, const double* N, const double* wi, const double* wr, const double* kd, const double* ks, const double* ns, double* rgb)
{
    for (int i = 0; i < n; i++) S3(rgb + 3 * i, brdf_phong(P3(N + 3 * i), P3(wi + 3 * i), P3(wr + 3 * i), P3(kd + 3 * i), P3(ks + 3 * i), ns[i]));
}
void oracle_brdf_sample(int n, const double* N, const double* wr, const double* kd, const double* ks, const double* ns, const double* u3, double* dir, double* pdf)
{
    oracle_stream st; memset(&st, 0, sizeof st);
    for (int i = 0; i < n; i++) {
        oracle_stream_list(&st, u3 + 3 * i, 3);
        S3(dir + 3 * i, brdf_sample(&st, P3(N + 3 * i), P3(wr + 3 * i), P3(kd + 3 * i), P3(ks + 3 * i), ns[i], pdf + i));
    }
}
void oracle_brdf_pdf(int n, const double* N, const double* wi, const double* wr, const double* kd, const double* ks, const double* ns, double* pdf)
{
    for (int i = 0; i < n; i++) pdf[i] = brdf_pdf(P3(N + 3 * i), P3(wi + 3 * i), P3(wr + 3 * i), P3(kd + 3 * i), P3(ks + 3 * i), ns[i]);
}

/* RadianceRGB::tone_mapping, RadianceRGB.cpp:51-67 */
void oracle_tonemap(int n, const double* rgb, double max_radiance, double gamma, int* out)
{
    double A = pow(max_radiance, -gamma);
    for (int i = 0; i < 3 * n; i++) {
        double r = A * pow(rgb[i], gamma);
        int v = (int)floor(r * 255 + 0.5);
        out[i] = v > 255 ? 255 : (v < 0 ? 0 : v); /* std::max(std::min(v,255),0) */
    }
}

/* ---- lights (Mylight.cpp) --------------------------------------------------------------------- */
typedef struct { int prim; v3 coord; v3 I; double prob; } light_pt; /* sampledLightPoint, Mylight.h:67-97 */

static inline double rgb_sum(v3 a) { return a.x[0] + a.x[1] + a.x[2]; } /* RadianceRGB::sum, RadianceRGB.cpp:70-73 */

/* select_a_point_from_lights, Mylight.cpp:102-160.  Draws: u_class (if >1 class), u_tri (if the class
 * has >1 triangle), xi1, xi2.  Class weight is R+G+B (not x area); pdf is in AREA measure. */
static light_pt light_sample_uniform(OracleScene* s, oracle_stream* st)
{
    double proba = 1.0, p;
    double* w = (double*)malloc(sizeof(double) * (size_t)(s->C > s->L ? s->C : s->L) + 8);
    for (int c = 0; c < s->C; c++) w[c] = rgb_sum(s->class_rad[c]);
    int rind = discrete_pick(st, w, s->C, &p);
    proba *= p;
    int b = s->class_begin[rind], e = s->class_begin[rind + 1];
    for (int l = b; l < e; l++) w[l - b] = s->light_area[l];
    int aind = discrete_pick(st, w, e - b, &p);
    free(w);
    proba *= p;
    int l = b + aind;
    double ksi1 = oracle_stream_next(st);
    double beta = 1 - sqrt(1 - ksi1);
    double ksi2 = oracle_stream_next(st);
    double gamma = (1 - beta) * ksi2;
    double alpha = 1 - beta - gamma;
    const v3* vs = s->pos + 3 * (size_t)s->light_prim[l];
    light_pt r;
    r.coord = vadd(vadd(vmul(vs[0], alpha), vmul(vs[1], beta)), vmul(vs[2], gamma));
    r.I = s->class_rad[rind];
    proba *= 1.0 / s->light_area[l];
    r.prim = s->light_prim[l];
    r.prob = proba;
    return r;
}

/* The per-vertex loop over all light triangles shared by select_a_point_from_lights_spherical_triangle
 * (Mylight.cpp:169-259) and prepared_for_lights_spherical_triangle_sampling (:329-421): culls,
 * projection to the unit sphere, winding fix, arcs a,b,c, dihedral angles, sA = alpha+beta+gamma-pi,
 * weight = sA * (R+G+B).  Fills the scene scratch (== the Mylight members in the staged variant). */
static void light_prepare(OracleScene* s, v3 x1, v3 n)
{
    s->st_n = 0;
    s->st_wsum = 0;
    memset(s->st_in, 0, (size_t)s->L);
    for (int l = 0; l < s->L; l++) {
        const v3* points = s->pos + 3 * (size_t)s->light_prim[l];
        v3 nl = geom_normal(s, s->light_prim[l]);
        double tmp = vdot(nl, vsub(x1, points[0]));
        if (tmp < 0 || fabs(tmp) < EPS) continue; /* :180-185 x1 behind the light */
        double tmps[3];
        for (int i = 0; i < 3; i++) tmps[i] = vdot(n, vsub(points[i], x1));
        if ((tmps[0] < 0 || fabs(tmps[0]) < EPS) && (tmps[1] < 0 || fabs(tmps[1]) < EPS) && (tmps[2] < 0 || fabs(tmps[2]) < EPS))
            continue; /* :194 light entirely below the horizon */
        v3 A = vnormalized(vsub(points[0], x1));
        v3 B = vnormalized(vsub(points[1], x1));
        v3 C = vnormalized(vsub(points[2], x1));
        if (vdot(vcross(vnormalized(vsub(C, A)), vnormalized(vsub(B, A))), n) < 0) { v3 t = B; B = C; C = t; } /* :206-211 */
        double a = acos(fmax(-1, fmin(1, vdot(B, C))));
        double b = acos(fmax(-1, fmin(1, vdot(A, C))));
        double c = acos(fmax(-1, fmin(1, vdot(A, B))));
        if (a < EPS || b < EPS || c < EPS) continue; /* :219 */
        double alpha = acos(fmax(-1, fmin(1, vdot(vmul(vnormalized(vcross(B, A)), -1), vnormalized(vcross(A, C))))));
        double beta = acos(fmax(-1, fmin(1, vdot(vmul(vnormalized(vcross(C, B)), -1), vnormalized(vcross(B, A))))));
        double gamma = acos(fmax(-1, fmin(1, vdot(vmul(vnormalized(vcross(A, C)), -1), vnormalized(vcross(C, B))))));
        if (alpha < EPS || beta < EPS || gamma < EPS) continue; /* :230 */
        double sA = alpha + beta + gamma - PI;
        if (sA < 0) continue; /* :238 */
        double weight = sA * rgb_sum(s->class_rad[s->light_class[l]]);
        if (weight < 0) continue;
        if (isinf(weight) || isnan(weight)) continue;
        int k = s->st_n++;
        s->st_light[k] = l; s->st_A[k] = A; s->st_B[k] = B; s->st_C[k] = C;
        s->st_alpha[k] = alpha; s->st_c[k] = c; s->st_sA[k] = sA; s->st_w[k] = weight;
        s->st_in[l] = 1;
        s->st_wsum += weight;
    }
}

/* Selection + Arvo SampleTriangle + re-intersection, Mylight.cpp:262-317 / :424-482.
 * Draws: u_tri (if >1 survivor), xi1, xi2.  prob = (R+G+B) / weights_sum, SOLID-ANGLE measure. */
static light_pt light_sample_prepared(OracleScene* s, oracle_stream* st, v3 x1, v3 n)
{
    light_pt r;
    if (s->st_n == 0 || fabs(s->st_wsum) < EPS) { /* :263-266 dummy */
        r.prim = 0; r.coord = vadd(vmul(n, -1), x1); r.I = V(0, 0, 0); r.prob = 1;
        return r;
    }
    int rind = discrete_pick(st, s->st_w, s->st_n, 0);
    v3 A = s->st_A[rind], B = s->st_B[rind], C = s->st_C[rind];
    double alpha = s->st_alpha[rind], cc = s->st_c[rind], area = s->st_sA[rind];
    double ksi1 = oracle_stream_next(st);
    double ksi2 = oracle_stream_next(st);
    /* SampleTriangle(ksi1,ksi2), :289-297 */
    double sA1 = ksi1 * area;
    double ss = sin(sA1 - alpha);
    double t = cos(sA1 - alpha);
    double u = t - cos(alpha);
    double v = ss + sin(alpha) * cos(cc);
    double q = ((v * t - u * ss) * cos(alpha) - v) / ((v * ss + u * t) * sin(alpha));
    v3 C1 = vnormalized(vadd(vmul(A, q), vmul(vnormalized(vsub(C, vmul(A, vdot(C, A)))), sqrt(1 - q * q))));
    double z = 1 - ksi2 * (1 - vdot(C1, B));
    v3 Pd = vnormalized(vadd(vmul(B, z), vmul(vnormalized(vsub(C1, vmul(B, vdot(C1, B)))), sqrt(1 - z * z))));
    int l = s->st_light[rind];
    hit_t rs = intersect_tri(s, x1, Pd, s->light_prim[l]); /* :311; a miss leaves t = 0 -> coord = x1 */
    r.prim = s->light_prim[l];
    r.coord = vadd(x1, vmul(Pd, rs.t));
    r.I = s->class_rad[s->light_class[l]];
    r.prob = rgb_sum(r.I) / s->st_wsum;
    return r;
}

/* eval_spherical_triangle_sampling_pdf, Mylight.cpp:484-493 (reads the scratch of the LAST prepare) */
static double light_pdf_prepared(const OracleScene* s, int prim)
{
    int l = s->light_of_prim[prim];
    if (l < 0 || !s->st_in[l]) return 0;
    if (fabs(s->st_wsum) < EPS) return 0;
    return rgb_sum(s->class_rad[s->light_class[l]]) / s->st_wsum;
}

void oracle_light_sample_uniform(OracleScene* s, int n, const double* u4, int* prim, double* point, double* Le, double* pdf)
{
    oracle_stream st; memset(&st, 0, sizeof st);
    for (int i = 0; i < n; i++) {
        oracle_stream_list(&st, u4 + 4 * i, 4);
        light_pt r = light_sample_uniform(s, &st);
        prim[i] = r.prim; S3(point + 3 * i, r.coord); S3(Le + 3 * i, r.I); pdf[i] = r.prob;
    }
}

void oracle_light_sample_spherical(OracleScene* s, int n, int staged, const double* x1, const double* N, const double* u3,
                                   const int* query, int* prim, double* point, double* Le, double* pdf, double* wsum,
                                   int* nsurv, double* query_pdf)
{
    (void)staged; /* one-shot and staged variants are the same arithmetic (Mylight.cpp:163-318 vs :322-482) */
    oracle_stream st; memset(&st, 0, sizeof st);
    for (int i = 0; i < n; i++) {
        oracle_stream_list(&st, u3 + 3 * i, 3);
        v3 x = P3(x1 + 3 * i), nn = P3(N + 3 * i);
        light_prepare(s, x, nn);
        light_pt r = light_sample_prepared(s, &st, x, nn);
        prim[i] = r.prim; S3(point + 3 * i, r.coord); S3(Le + 3 * i, r.I); pdf[i] = r.prob;
        if (wsum) wsum[i] = s->st_wsum;
        if (nsurv) nsurv[i] = s->st_n;
        if (query && query_pdf) query_pdf[i] = light_pdf_prepared(s, query[i]);
    }
}

/* the keyed table itself, for the tests (known-answer vectors of Philox4x32-10; equality with csrc/rng.cuh) */
void oracle_keyed_uniforms(uint64_t seed, uint32_t pixel, uint32_t sample, uint64_t node, uint32_t call, double out[4])
{
    oracle_keyed_draw4(seed, pixel, sample, node, call, out);
}

/* ---- integrators (main.cpp) ------------------------------------------------------------------- */
typedef struct { v3 p, N; int prim, mat, is_light; } vertex_t;

/* common prologue of all three shade* bodies: main.cpp:273-288 / :351-366 / :406-421.
 * returns 0 = continue, 1 = return *out */
static int shade_prologue(OracleScene* s, hit_t point, v3 wo, vertex_t* vx, v3* out)
{
    const v3* vs = s->pos + 3 * (size_t)point.prim;
    const v3* ns = s->nor + 3 * (size_t)point.prim;
    vx->p = lerp3(vs, 1.0 - point.beta - point.gamma, point.beta, point.gamma);
    vx->N = vnormalized(lerp3(ns, 1.0 - point.beta - point.gamma, point.beta, point.gamma));
    vx->prim = point.prim;
    vx->mat = s->matid[point.prim];
    if (vdot(vx->N, wo) < 0) { *out = V(0, 0, 0); return 1; }
    int l = s->light_of_prim[point.prim];
    if (l >= 0) { *out = s->class_rad[s->light_class[l]]; return 1; }
    return 0;
}

/* Russian roulette draw with the depth rule of rng_shim.h (oracle_rr_engine): at depth >= max_depth
 * the draw is forced to fail and consumes nothing. */
static int rr_survives(OracleScene* s, int depth, uint64_t node)
{
    s->n_vertices++;
    if (s->max_depth > 0 && depth >= s->max_depth) return 0;
    keyed_list(s, node, 0, 0, 1);
    double ksi = oracle_stream_next(&s->stream);
    return !(ksi > P_RR);
}

static inline v3 rgb_mul(v3 a, v3 b) { return V(a.x[0] * b.x[0], a.x[1] * b.x[1], a.x[2] * b.x[2]); } /* RadianceRGB.cpp:44-47 */

/* shade(), main.cpp:269-344; uniform != 0 swaps in the commented-out sampler of main.cpp:297 */
static v3 shade_nee(OracleScene* s, hit_t point, v3 wo, int depth, int uniform, uint64_t node)
{
    vertex_t vx; v3 out;
    if (shade_prologue(s, point, wo, &vx, &out)) return out;
    v3 kd = s->kd[vx.mat], ks = s->ks[vx.mat]; double ns = s->ns[vx.mat];
    v3 L_dir = V(0, 0, 0);
    light_pt lp;
    if (uniform) { keyed_list(s, node, 1, 0, 4); lp = light_sample_uniform(s, &s->stream); }
    else { light_prepare(s, vx.p, vx.N); keyed_list(s, node, 1, 0, 3); lp = light_sample_prepared(s, &s->stream, vx.p, vx.N); }
    v3 x1 = lp.coord;
    v3 n1 = geom_normal(s, lp.prim);
    v3 wl = vnormalized(vsub(x1, vx.p));
    if (vdot(wl, vx.N) > 0 && vdot(vmul(wl, -1), n1) > 0) {
        hit_t rstmp = closest(s, vx.p, wl, vx.prim);
        if (rstmp.hit && rstmp.prim == lp.prim) {
            v3 brdf = brdf_phong(vx.N, wl, wo, kd, ks, ns);
            double g;
            if (!uniform && (s->flags & ORACLE_FIX_NEE_MEASURE)) g = vdot(wl, vx.N) / lp.prob;
            else g = vdot(wl, vx.N) * vdot(vmul(wl, -1), n1) / vdot(vsub(x1, vx.p), vsub(x1, vx.p)) / lp.prob; /* :314 (F9) */
            L_dir = vmul(rgb_mul(lp.I, brdf), g);
        }
    }
    v3 L_indir = V(0, 0, 0);
    depth++; /* the RR engine of this frame is now alive (rng_shim.h) */
    if (!rr_survives(s, depth, node)) return vadd(L_dir, L_indir);
    double pdf;
    keyed_list(s, node, 0, 1, 3);
    v3 wi = brdf_sample(&s->stream, vx.N, wo, kd, ks, ns, &pdf);
    if (vdot(wi, vx.N) < 0) return vadd(L_dir, L_indir);
    if (s->flags & ORACLE_FIX_MIXTURE_PDF) pdf = brdf_pdf(vx.N, wi, wo, kd, ks, ns);
    hit_t rsq = closest(s, vx.p, wi, vx.prim);
    if (rsq.hit && s->light_of_prim[rsq.prim] < 0) { /* :338 indirect only off non-lights */
        v3 brdf = brdf_phong(vx.N, wi, wo, kd, ks, ns);
        L_indir = vmul(rgb_mul(shade_nee(s, rsq, vmul(wi, -1), depth, uniform, (uint64_t)depth + 1), brdf), vdot(wi, vx.N) / pdf / P_RR);
    }
    return vadd(L_dir, L_indir);
}

/* shade_with_brdf(), main.cpp:348-399 */
static v3 shade_brdf(OracleScene* s, hit_t point, v3 wo, int depth, uint64_t node)
{
    vertex_t vx; v3 out;
    if (shade_prologue(s, point, wo, &vx, &out)) return out;
    v3 kd = s->kd[vx.mat], ks = s->ks[vx.mat]; double ns = s->ns[vx.mat];
    v3 L = V(0, 0, 0);
    depth++;
    if (!rr_survives(s, depth, node)) return L;
    double pdf;
    keyed_list(s, node, 0, 1, 3);
    v3 wi = brdf_sample(&s->stream, vx.N, wo, kd, ks, ns, &pdf);
    if (vdot(wi, vx.N) < 0) return L;
    if (s->flags & ORACLE_FIX_MIXTURE_PDF) pdf = brdf_pdf(vx.N, wi, wo, kd, ks, ns);
    hit_t rsq = closest(s, vx.p, wi, vx.prim);
    if (rsq.hit) {
        v3 brdf = brdf_phong(vx.N, wi, wo, kd, ks, ns);
        L = vmul(rgb_mul(shade_brdf(s, rsq, vmul(wi, -1), depth, (uint64_t)depth + 1), brdf), vdot(wi, vx.N) / pdf / P_RR);
    }
    return L;
}

/* shade_with_mis(), main.cpp:402-494.  Without ORACLE_FIX_STALE_MIS the light pdf at :487 reads the
 * scratch left by the deepest prepare of recursion #1 (F10), exactly like the reference. */
static v3 shade_mis(OracleScene* s, hit_t point, v3 wo, int depth, uint64_t node)
{
    vertex_t vx; v3 out;
    if (shade_prologue(s, point, wo, &vx, &out)) return out;
    v3 kd = s->kd[vx.mat], ks = s->ks[vx.mat]; double ns = s->ns[vx.mat];
    depth++;
    if (!rr_survives(s, depth, node)) return V(0, 0, 0);
    v3 L_light = V(0, 0, 0);
    light_prepare(s, vx.p, vx.N);
    keyed_list(s, node, 1, 0, 3);
    light_pt lp = light_sample_prepared(s, &s->stream, vx.p, vx.N);
    v3 x1 = lp.coord;
    v3 wl = vnormalized(vsub(x1, vx.p));
    /* own-vertex scratch, for ORACLE_FIX_STALE_MIS */
    double own_wsum = s->st_wsum;
    unsigned char* own_in = 0;
    if (s->flags & ORACLE_FIX_STALE_MIS) {
        own_in = (unsigned char*)malloc((size_t)s->L + 1);
        memcpy(own_in, s->st_in, (size_t)s->L);
    }
    if (vdot(wl, vx.N) > 0) {
        hit_t rstmp = closest(s, vx.p, wl, vx.prim);
        if (rstmp.hit) {
            v3 brdf = brdf_phong(vx.N, wl, wo, kd, ks, ns);
            double phongPdf = brdf_pdf(vx.N, wl, wo, kd, ks, ns);
            L_light = vmul(rgb_mul(shade_mis(s, rstmp, vmul(wl, -1), depth, node * 2), brdf), vdot(wl, vx.N) / (lp.prob + phongPdf) / P_RR);
        }
    }
    if (own_in) {
        memcpy(s->st_in, own_in, (size_t)s->L);
        s->st_wsum = own_wsum;
        free(own_in);
    }
    v3 L_brdf = V(0, 0, 0);
    double pdf;
    keyed_list(s, node, 0, 1, 3);
    v3 wi = brdf_sample(&s->stream, vx.N, wo, kd, ks, ns, &pdf);
    if (vdot(wi, vx.N) < 0) return vadd(L_light, L_brdf);
    if (s->flags & ORACLE_FIX_MIXTURE_PDF) pdf = brdf_pdf(vx.N, wi, wo, kd, ks, ns);
    hit_t rsq = closest(s, vx.p, wi, vx.prim);
    if (rsq.hit) {
        v3 brdf = brdf_phong(vx.N, wi, wo, kd, ks, ns);
        double light_pdf;
        hit_t rs = closest_light(s, vx.p, wi, vx.prim);
        if (rs.hit) light_pdf = light_pdf_prepared(s, rs.prim);
        else light_pdf = 0;
        L_brdf = vmul(rgb_mul(shade_mis(s, rsq, vmul(wi, -1), depth, node * 2 + 1), brdf), vdot(wi, vx.N) / (pdf + light_pdf) / P_RR);
    }
    return vadd(L_light, L_brdf);
}

/* ---- render loop, main.cpp:546-588 ------------------------------------------------------------ */
void oracle_render(OracleScene* s, const double eye[3], const double w_[3], const double up[3], double pixellen, int W,
                   int H, int i0, int i1, int j0, int j1, int spp, int spp_offset, int spp_total, int strategy,
                   int max_depth, int flags, uint64_t seed, double* out, uint64_t* counters)
{
    s->max_depth = max_depth;
    s->flags = flags;
    s->n_vertices = s->n_rays = s->n_light_rays = 0;
    uint64_t primary_hits = 0;
    v3 eyev = P3(eye), w = P3(w_);
    v3 N = vnormalized(w);                       /* main.cpp:549 */
    v3 UP = P3(up);
    v3 Vv = vnormalized(vcross(N, UP));          /* :551 */
    v3 U = vnormalized(vcross(Vv, N));           /* :552 */
    const double ci = (H - 1) * 0.5, cj = (W - 1) * 0.5;
    const double inv = 1.0 / spp_total;
    for (int i = i0; i < i1; i++)
        for (int j = j0; j < j1; j++) {
            v3 delta = V(-pixellen * (i - ci), pixellen * (j - cj), 0);                         /* :563 */
            v3 dir = vnormalized(mat_cols_mul(U, Vv, N, vadd(delta, V(0, 0, vnorm(w)))));       /* :564 */
            v3 sum = V(0, 0, 0);
            for (int k = 0; k < spp; k++) {
                oracle_stream_seed(&s->stream, seed, (uint64_t)i * (uint64_t)W + (uint64_t)j, (uint64_t)(spp_offset + k));
                s->k_seed = seed; s->k_pixel = (uint32_t)(i * W + j); s->k_sample = (uint32_t)(spp_offset + k);
                hit_t rs = closest(s, eyev, dir, -1);                                           /* :572 */
                if (rs.hit) {
                    primary_hits++;
                    v3 wo = vmul(dir, -1);
                    v3 I = strategy == ORACLE_BRDF_ONLY ? shade_brdf(s, rs, wo, 0, 1)
                         : strategy == ORACLE_MIS       ? shade_mis(s, rs, wo, 0, 1)
                         : shade_nee(s, rs, wo, 0, strategy == ORACLE_NEE_UNIFORM, 1);
                    sum = vadd(sum, vmul(I, inv));                                              /* :575-576 */
                }
            }
            double* o = out + ((size_t)(i - i0) * (size_t)(j1 - j0) + (size_t)(j - j0)) * 3;
            o[0] = sum.x[0]; o[1] = sum.x[1]; o[2] = sum.x[2];
        }
    if (counters) {
        counters[0] = s->n_vertices; counters[1] = primary_hits; counters[2] = s->n_rays; counters[3] = s->n_light_rays;
    }
}
