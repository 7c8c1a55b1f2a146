// csrc/light_tree.cpp -- host builder of the light hierarchy (light_tree.h).  Top-down, binned splits scored with the
// surface-area-orientation heuristic: cost(side) = power * box area * measure of the side's normal cone, scaled by
// (largest centroid extent / extent of the split axis) so that long thin groups are cut across their length.
#include "light_tree.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>

#include "bvh_tree.h"

namespace mcpt {
namespace {

constexpr double kPi = 3.14159265358979323846;
constexpr int kBins = 12;

struct Item {
    BvhBox box;      // of the triangle's three vertices
    float c[3];      // centroid
    double n[3];     // emitting-side unit normal
    double power;    // area * (R+G+B)
    int32_t light;
};

struct Cone { double axis[3]; double cos_o; };

// cone around the normalised sum of the normals that holds every normal of items[b, e)
Cone cone_of(const std::vector<Item>& it, size_t b, size_t e)
{
    Cone c{{0, 0, 1}, -1.0};
    double s[3] = {0, 0, 0};
    for (size_t i = b; i < e; i++) for (int a = 0; a < 3; a++) s[a] += it[i].n[a];
    const double len = std::sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2]);
    if (!(len > 1e-3 * (double)(e - b))) return c;  // normals cancel: the whole sphere
    for (int a = 0; a < 3; a++) c.axis[a] = s[a] / len;
    double m = 1.0;
    for (size_t i = b; i < e; i++) m = std::min(m, c.axis[0] * it[i].n[0] + c.axis[1] * it[i].n[1] + c.axis[2] * it[i].n[2]);
    c.cos_o = m;
    return c;
}

// solid-angle-like measure of the directions a cone of one-sided emitters (theta_e = pi/2) can light
double cone_measure(double cos_o)
{
    const double to = std::acos(std::min(1.0, std::max(-1.0, cos_o)));
    const double tw = std::min(to + 0.5 * kPi, kPi);
    return 2 * kPi * (1 - std::cos(to)) + 0.5 * kPi * (2 * tw * std::sin(to) - std::cos(to - 2 * tw) - 2 * to * std::sin(to) + std::cos(to));
}

double box_area(const BvhBox& b)
{
    const double d[3] = {(double)b.hi[0] - b.lo[0], (double)b.hi[1] - b.lo[1], (double)b.hi[2] - b.lo[2]};
    if (d[0] < 0 || d[1] < 0 || d[2] < 0) return 0;
    const double a = 2 * (d[0] * d[1] + d[1] * d[2] + d[2] * d[0]);
    return a > 0 ? a : d[0] + d[1] + d[2];  // a segment or a point still has to cost something
}

struct Builder {
    std::vector<Item> it;
    LightTree* T;
    uint32_t depth_max = 0;

    int32_t new_node()
    {
        const int32_t id = (int32_t)(T->nodes.size() / 3);
        T->nodes.resize(T->nodes.size() + 3);
        return id;
    }
    void fill(int32_t id, size_t b, size_t e, int32_t left, int32_t right, int32_t parent)
    {
        BvhBox box;
        box.reset();
        double power = 0;
        for (size_t i = b; i < e; i++) { box.grow(it[i].box); power += it[i].power; }
        Cone c;
        if (e - b == 1) { c.axis[0] = it[b].n[0]; c.axis[1] = it[b].n[1]; c.axis[2] = it[b].n[2]; c.cos_o = 1.0; }
        else { c = cone_of(it, b, e); c.cos_o = std::max(-1.0, c.cos_o - 1e-5); }
        const double cx = 0.5 * ((double)box.lo[0] + box.hi[0]), cy = 0.5 * ((double)box.lo[1] + box.hi[1]), cz = 0.5 * ((double)box.lo[2] + box.hi[2]);
        const double dx = (double)box.hi[0] - box.lo[0], dy = (double)box.hi[1] - box.lo[1], dz = (double)box.hi[2] - box.lo[2];
        const double r = 0.5 * std::sqrt(dx * dx + dy * dy + dz * dz) * (1.0 + 1e-5);
        f4 t0{(float)cx, (float)cy, (float)cz, (float)r}, t1{(float)c.axis[0], (float)c.axis[1], (float)c.axis[2], (float)c.cos_o}, t2{(float)power, 0, 0, 0};
        // the float centre is not the exact centre: grow the radius by the rounding of the centre
        t0.w = std::nextafter(t0.w + (float)(std::fabs(cx - (double)t0.x) + std::fabs(cy - (double)t0.y) + std::fabs(cz - (double)t0.z)), INFINITY);
        std::memcpy(&t2.y, &left, 4); std::memcpy(&t2.z, &right, 4); std::memcpy(&t2.w, &parent, 4);
        T->nodes[3 * (size_t)id] = t0; T->nodes[3 * (size_t)id + 1] = t1; T->nodes[3 * (size_t)id + 2] = t2;
    }

    // where to cut items[b, e) in two: returns the first index of the right half after partitioning in place
    size_t split(size_t b, size_t e)
    {
        const size_t n = e - b;
        if (n == 2) return b + 1;
        BvhBox cb;
        cb.reset();
        for (size_t i = b; i < e; i++) cb.grow(it[i].c);
        const double ext[3] = {(double)cb.hi[0] - cb.lo[0], (double)cb.hi[1] - cb.lo[1], (double)cb.hi[2] - cb.lo[2]};
        const double ext_max = std::max({ext[0], ext[1], ext[2]});
        double best = INFINITY;
        int best_axis = -1, best_cut = 0;
        for (int ax = 0; ax < 3 && ext_max > 0; ax++) {
            if (!(ext[ax] > 0)) continue;
            struct Bin { BvhBox box; double nsum[3]; double power; size_t count; Cone cone; } bin[kBins];
            for (auto& k : bin) { k.box.reset(); k.nsum[0] = k.nsum[1] = k.nsum[2] = 0; k.power = 0; k.count = 0; k.cone = Cone{{0, 0, 1}, 1.0}; }
            auto bin_of = [&](const Item& x) { return std::min(kBins - 1, (int)(kBins * ((double)x.c[ax] - cb.lo[ax]) / ext[ax])); };
            for (size_t i = b; i < e; i++) {
                Bin& k = bin[bin_of(it[i])];
                k.box.grow(it[i].box); k.power += it[i].power; k.count++;
                for (int a = 0; a < 3; a++) k.nsum[a] += it[i].n[a];
            }
            for (auto& k : bin) {  // per-bin cone: axis from the normal sum, spread by a pass over the members below
                const double len = std::sqrt(k.nsum[0] * k.nsum[0] + k.nsum[1] * k.nsum[1] + k.nsum[2] * k.nsum[2]);
                if (k.count && len > 1e-3 * (double)k.count) { for (int a = 0; a < 3; a++) k.cone.axis[a] = k.nsum[a] / len; }
                else k.cone.cos_o = -1.0;
            }
            for (size_t i = b; i < e; i++) {
                Bin& k = bin[bin_of(it[i])];
                if (k.cone.cos_o > -1.0) k.cone.cos_o = std::min(k.cone.cos_o, k.cone.axis[0] * it[i].n[0] + k.cone.axis[1] * it[i].n[1] + k.cone.axis[2] * it[i].n[2]);
            }
            auto side_cost = [&](int from, int to) -> double {  // bins [from, to)
                BvhBox box;
                box.reset();
                double power = 0, s[3] = {0, 0, 0};
                size_t cnt = 0;
                for (int k = from; k < to; k++) {
                    if (!bin[k].count) continue;
                    box.grow(bin[k].box); power += bin[k].power; cnt += bin[k].count;
                    for (int a = 0; a < 3; a++) s[a] += bin[k].nsum[a];
                }
                if (!cnt) return INFINITY;
                const double len = std::sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2]);
                double cos_o = -1.0;
                if (len > 1e-3 * (double)cnt) {  // widest (bin axis angle + bin spread) around the side's mean normal
                    double widest = 0;
                    for (int k = from; k < to; k++) {
                        if (!bin[k].count) continue;
                        const double d = (s[0] * bin[k].cone.axis[0] + s[1] * bin[k].cone.axis[1] + s[2] * bin[k].cone.axis[2]) / len;
                        widest = std::max(widest, std::acos(std::min(1.0, std::max(-1.0, d))) + std::acos(std::min(1.0, std::max(-1.0, bin[k].cone.cos_o))));
                    }
                    cos_o = std::cos(std::min(widest, kPi));
                }
                return power * box_area(box) * cone_measure(cos_o);
            };
            for (int cut = 1; cut < kBins; cut++) {
                const double c = (side_cost(0, cut) + side_cost(cut, kBins)) * (ext_max / ext[ax]);
                if (c < best) { best = c; best_axis = ax; best_cut = cut; }
            }
        }
        if (best_axis >= 0 && std::isfinite(best)) {
            const int ax = best_axis;
            auto mid = std::partition(it.begin() + b, it.begin() + e, [&](const Item& x) {
                return std::min(kBins - 1, (int)(kBins * ((double)x.c[ax] - cb.lo[ax]) / ext[ax])) < best_cut;
            });
            const size_t m = (size_t)(mid - it.begin());
            if (m > b && m < e) return m;
        }
        return b + n / 2;  // coincident centroids: halve by order
    }

    // builds the subtree of items[b, e) into node `id`
    void build(int32_t id, size_t b, size_t e, int32_t parent, uint32_t depth)
    {
        depth_max = std::max(depth_max, depth);
        if (e - b == 1) {
            fill(id, b, e, ~it[b].light, -1, parent);
            T->leaf_of_light[it[b].light] = id;
            return;
        }
        const size_t m = split(b, e);
        const int32_t l = new_node(), r = new_node();
        fill(id, b, e, l, r, parent);
        build(l, b, m, id, depth + 1);
        build(r, m, e, id, depth + 1);
    }
};

}  // namespace

LightTree build_light_tree(const float* positions, const int32_t* light_prim, const double* geom_normals, const double* area,
                           const double* lsum, uint32_t L)
{
    LightTree T;
    T.leaf_of_light.assign(L, -1);
    if (!L) return T;
    Builder B;
    B.T = &T;
    B.it.resize(L);
    for (uint32_t l = 0; l < L; l++) {
        Item& x = B.it[l];
        const float* v = positions + (size_t)light_prim[l] * 9;
        x.box.reset();
        for (int k = 0; k < 3; k++) x.box.grow(v + 3 * k);
        for (int a = 0; a < 3; a++) {
            x.c[a] = (v[a] + v[3 + a] + v[6 + a]) * (1.0f / 3.0f);
            x.n[a] = geom_normals[(size_t)light_prim[l] * 3 + a];
        }
        x.power = area[l] * lsum[l];
        x.light = (int32_t)l;
    }
    T.nodes.reserve((size_t)(2 * L - 1) * 3);
    const int32_t root = B.new_node();
    B.build(root, 0, L, -1, 0);
    T.num_nodes = (uint32_t)(T.nodes.size() / 3);
    T.max_depth = B.depth_max;
    return T;
}

uint32_t check_light_tree(const LightTree& T, const float* positions, const int32_t* light_prim, const double* geom_normals, uint32_t L)
{
    uint32_t bad = 0;
    if (!L) return T.nodes.empty() ? 0 : 1;
    if (T.num_nodes != 2 * L - 1 || T.nodes.size() != (size_t)T.num_nodes * 3 || T.leaf_of_light.size() != L) return 1;
    auto geti = [](float f) { int32_t i; std::memcpy(&i, &f, 4); return i; };
    std::vector<int> seen(L, 0);
    for (uint32_t id = 0; id < T.num_nodes; id++) {
        const f4 t2 = T.nodes[3 * (size_t)id + 2];
        const int32_t left = geti(t2.y), right = geti(t2.z), parent = geti(t2.w);
        if ((id == 0) != (parent < 0)) bad++;
        if (left < 0) {
            const int32_t l = ~left;
            if (l < 0 || (uint32_t)l >= L) { bad++; continue; }
            seen[l]++;
            if (T.leaf_of_light[l] != (int32_t)id) bad++;
            // every ancestor's sphere holds the light's vertices and every ancestor's cone its normal
            const float* v = positions + (size_t)light_prim[l] * 9;
            const double* g = geom_normals + (size_t)light_prim[l] * 3;
            uint32_t steps = 0;
            for (int32_t a = (int32_t)id; a >= 0 && steps <= T.num_nodes; a = geti(T.nodes[3 * (size_t)a + 2].w), steps++) {
                const f4 a0 = T.nodes[3 * (size_t)a], a1 = T.nodes[3 * (size_t)a + 1];
                for (int k = 0; k < 3; k++) {
                    const double dx = v[3 * k] - (double)a0.x, dy = v[3 * k + 1] - (double)a0.y, dz = v[3 * k + 2] - (double)a0.z;
                    if (std::sqrt(dx * dx + dy * dy + dz * dz) > (double)a0.w * (1 + 1e-6) + 1e-30) bad++;
                }
                if (a1.w > -1.f && a1.x * g[0] + a1.y * g[1] + a1.z * g[2] < (double)a1.w - 1e-6) bad++;
            }
            if (steps > T.num_nodes) bad++;  // parent links loop
            continue;
        }
        if (left <= (int32_t)id || right <= (int32_t)id || (uint32_t)left >= T.num_nodes || (uint32_t)right >= T.num_nodes) { bad++; continue; }
        for (int32_t ch : {left, right})
            if (geti(T.nodes[3 * (size_t)ch + 2].w) != (int32_t)id) bad++;
        const double pl = T.nodes[3 * (size_t)left + 2].x, pr = T.nodes[3 * (size_t)right + 2].x;
        if (std::fabs(pl + pr - (double)t2.x) > 1e-4 * std::fabs((double)t2.x) + 1e-30) bad++;
    }
    for (uint32_t l = 0; l < L; l++) if (seen[l] != 1) bad++;
    return bad;
}

}  // namespace mcpt
