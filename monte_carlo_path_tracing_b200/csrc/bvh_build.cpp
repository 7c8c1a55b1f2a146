// csrc/bvh_build.cpp -- binned-SAH top-down BVH builder (host), emitting the layout of bvh.h.
#include "bvh.h"
#include "bvh_tree.h"
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <limits>

namespace mcpt {
namespace {

using Box = BvhBox;
using TmpNode = BvhTmpNode;
struct Ref { Box box; float c[3]; int32_t prim; };

constexpr int kBins = 16;
#ifndef MCPT_COST_TRAV
#define MCPT_COST_TRAV 1.0f
#endif
constexpr float kCostTrav = MCPT_COST_TRAV, kCostTri = 1.0f;

struct Builder {
    std::vector<Ref> refs;
    std::vector<TmpNode> tmp;
    uint32_t max_depth = 0;

    int32_t build(uint32_t lo, uint32_t hi, uint32_t depth)
    {
        int32_t idx = (int32_t)tmp.size();
        tmp.emplace_back();
        Box box, cbox;
        box.reset(); cbox.reset();
        for (uint32_t i = lo; i < hi; i++) { box.grow(refs[i].box); cbox.grow(refs[i].c); }
        tmp[idx].box = box;
        max_depth = std::max(max_depth, depth);
        uint32_t n = hi - lo;
        auto make_leaf = [&]() { tmp[idx].first = lo; tmp[idx].count = n; return idx; };
        if (n <= 1) return make_leaf();

        uint32_t mid = 0;
        bool split_found = false;
        if (depth < (uint32_t)kBvhStackDepth - 24) {
            // binned SAH over the three axes
            float best = std::numeric_limits<float>::infinity();
            int best_axis = -1, best_bin = -1;
            for (int a = 0; a < 3; a++) {
                float ext = cbox.hi[a] - cbox.lo[a];
                if (!(ext > 0)) continue;
                Box bb[kBins]; uint32_t bc[kBins];
                for (int b = 0; b < kBins; b++) { bb[b].reset(); bc[b] = 0; }
                float scale = kBins / ext;
                for (uint32_t i = lo; i < hi; i++) {
                    int b = std::min(kBins - 1, std::max(0, (int)((refs[i].c[a] - cbox.lo[a]) * scale)));
                    bb[b].grow(refs[i].box); bc[b]++;
                }
                float la[kBins]; uint32_t lc[kBins];
                Box acc; acc.reset(); uint32_t cnt = 0;
                for (int b = 0; b < kBins - 1; b++) { acc.grow(bb[b]); cnt += bc[b]; la[b] = acc.half_area(); lc[b] = cnt; }
                acc.reset(); cnt = 0;
                for (int b = kBins - 1; b > 0; b--) {
                    acc.grow(bb[b]); cnt += bc[b];
                    if (lc[b - 1] == 0 || cnt == 0) continue;
                    float cost = la[b - 1] * lc[b - 1] + acc.half_area() * cnt;
                    if (cost < best) { best = cost; best_axis = a; best_bin = b; }
                }
            }
            if (best_axis >= 0) {
                float pa = box.half_area();
                float split_cost = kCostTrav + (pa > 0 ? best / pa : 0.f) * kCostTri;
                if (n <= (uint32_t)kBvhMaxLeaf && split_cost >= n * kCostTri) return make_leaf();
                float ext = cbox.hi[best_axis] - cbox.lo[best_axis];
                float scale = kBins / ext, base = cbox.lo[best_axis];
                int a = best_axis, bsplit = best_bin;
                auto it = std::partition(refs.begin() + lo, refs.begin() + hi, [&](const Ref& r) {
                    int b = std::min(kBins - 1, std::max(0, (int)((r.c[a] - base) * scale)));
                    return b < bsplit;
                });
                mid = (uint32_t)(it - refs.begin());
                split_found = mid > lo && mid < hi;
            }
        }
        if (!split_found) {
            if (n <= (uint32_t)kBvhMaxLeaf) return make_leaf();
            // object median on the widest centroid axis (also the depth guard)
            int a = 0;
            float e[3] = {cbox.hi[0] - cbox.lo[0], cbox.hi[1] - cbox.lo[1], cbox.hi[2] - cbox.lo[2]};
            if (e[1] > e[a]) a = 1;
            if (e[2] > e[a]) a = 2;
            mid = lo + n / 2;
            std::nth_element(refs.begin() + lo, refs.begin() + mid, refs.begin() + hi,
                             [a](const Ref& x, const Ref& y) { return x.c[a] < y.c[a]; });
        }
        int32_t l = build(lo, mid, depth + 1);
        int32_t r = build(mid, hi, depth + 1);
        tmp[idx].left = l; tmp[idx].right = r;
        return idx;
    }
};

inline void pad_box(const Box& b, float tiny, float lo[3], float hi[3])
{
    for (int a = 0; a < 3; a++) {
        lo[a] = b.lo[a] - (std::fabs(b.lo[a]) * 2e-7f + tiny);
        hi[a] = b.hi[a] + (std::fabs(b.hi[a]) * 2e-7f + tiny);
    }
}


// Collapse the binary tree into 4-wide nodes (bvh.h, nodes4), depth-first so a subtree stays contiguous.
void emit_wide(const std::vector<TmpNode>& tmp, int32_t root, float tiny, Bvh& out)
{
    const float inf = std::numeric_limits<float>::infinity();
    auto leaf_code = [](const TmpNode& t) { return (int32_t)~((t.first << 3) | (t.count - 1)); };
    struct Item { int32_t tmp_idx, out_idx; uint32_t depth; };
    std::vector<Item> todo;
    auto alloc = [&]() { int32_t i = (int32_t)(out.nodes4.size() / 7); out.nodes4.resize(out.nodes4.size() + 7, f4{0, 0, 0, 0}); return i; };
    todo.push_back({root, alloc(), 1});
    while (!todo.empty()) {
        const Item it = todo.back();
        todo.pop_back();
        out.max_depth4 = std::max(out.max_depth4, it.depth);
        int32_t kids[4];
        int nk = 0;
        if (tmp[it.tmp_idx].left < 0) kids[nk++] = it.tmp_idx;  // lone leaf root
        else { kids[nk++] = tmp[it.tmp_idx].left; kids[nk++] = tmp[it.tmp_idx].right; }
        while (nk < 4) {
            int best = -1;
            float best_area = -1.f;
            for (int k = 0; k < nk; k++)
                if (tmp[kids[k]].left >= 0 && tmp[kids[k]].box.half_area() > best_area) { best = k; best_area = tmp[kids[k]].box.half_area(); }
            if (best < 0) break;
            const int32_t open = kids[best];
            kids[best] = tmp[open].left;
            kids[nk++] = tmp[open].right;
        }
        float q[7][4];
        for (int k = 0; k < 4; k++) {
            for (int a = 0; a < 3; a++) { q[2 * a][k] = inf; q[2 * a + 1][k] = -inf; }
            int32_t code = kBvhNoChild;
            if (k < nk) {
                const TmpNode& c = tmp[kids[k]];
                float lo[3], hi[3];
                pad_box(c.box, tiny, lo, hi);
                for (int a = 0; a < 3; a++) { q[2 * a][k] = lo[a]; q[2 * a + 1][k] = hi[a]; }
                if (c.left < 0) code = leaf_code(c);
                else { code = alloc(); todo.push_back({kids[k], code, it.depth + 1}); }
            }
            std::memcpy(&q[6][k], &code, 4);
        }
        for (int r = 0; r < 7; r++) out.nodes4[(size_t)it.out_idx * 7 + r] = f4{q[r][0], q[r][1], q[r][2], q[r][3]};
    }
}

// Collapse the binary tree into compressed 8-wide nodes (bvh.h, nodes8 / tris8).
void emit_wide8(const float* positions, const std::vector<TmpNode>& tmp, const std::vector<int32_t>& order, int32_t root, float tiny, Bvh& out)
{
    struct Kid { Box box; int32_t tmp_idx; uint32_t first, count; bool leaf; };  // leaf: triangles order[first .. first + count), count <= 3
    auto tri_box = [&](uint32_t first, uint32_t count) {
        Box b; b.reset();
        for (uint32_t i = first; i < first + count; i++) for (int k = 0; k < 3; k++) b.grow(positions + (size_t)order[i] * 9 + 3 * k);
        return b;
    };
    auto make_kid = [&](int32_t t) {
        Kid k; k.box = tmp[t].box; k.tmp_idx = t; k.first = tmp[t].first; k.count = tmp[t].count; k.leaf = tmp[t].left < 0;
        return k;
    };
    struct Item { int32_t tmp_idx, out_idx; uint32_t depth; };
    std::vector<Item> todo;
    auto alloc = [&](int n) { int32_t i = (int32_t)(out.nodes8.size() / 5); out.nodes8.resize(out.nodes8.size() + 5 * (size_t)n, f4{0, 0, 0, 0}); return i; };
    todo.push_back({root, alloc(1), 1});
    while (!todo.empty()) {
        const Item it = todo.back();
        todo.pop_back();
        out.max_depth8 = std::max(out.max_depth8, it.depth);
        Kid kids[8];
        int nk = 0;
        if (tmp[it.tmp_idx].left < 0) kids[nk++] = make_kid(it.tmp_idx);  // lone leaf root
        else { kids[nk++] = make_kid(tmp[it.tmp_idx].left); kids[nk++] = make_kid(tmp[it.tmp_idx].right); }
        // open the inner child with the largest surface area while eight slots still hold everything; a leaf of more than 3
        // triangles needs two slots (it is split in two below: a slot addresses at most 3 triangles)
        auto need = [&](int32_t t) { return (tmp[t].left < 0 && tmp[t].count > 3) ? 2 : 1; };
        int total = 0;
        for (int k = 0; k < nk; k++) total += need(kids[k].tmp_idx);
        for (;;) {
            int best = -1;
            float best_area = -1.f;
            for (int k = 0; k < nk; k++) {
                if (kids[k].leaf) continue;
                const int after = total - 1 + need(tmp[kids[k].tmp_idx].left) + need(tmp[kids[k].tmp_idx].right);
                if (after <= 8 && kids[k].box.half_area() > best_area) { best = k; best_area = kids[k].box.half_area(); }
            }
            if (best < 0) break;
            const int32_t open = kids[best].tmp_idx;
            total += need(tmp[open].left) + need(tmp[open].right) - 1;
            kids[best] = make_kid(tmp[open].left);
            kids[nk++] = make_kid(tmp[open].right);
        }
        for (int k = 0, n0 = nk; k < n0; k++) {
            if (!kids[k].leaf || kids[k].count <= 3) continue;
            const Kid whole = kids[k];
            const uint32_t h = whole.count / 2;
            Kid a = whole, b = whole;
            a.count = h; a.box = tri_box(a.first, a.count);
            b.first = whole.first + h; b.count = whole.count - h; b.box = tri_box(b.first, b.count);
            kids[k] = a; kids[nk++] = b;
        }
        // node box, origin, per-axis exponent
        Box nb; nb.reset();
        Box padded[8];
        for (int k = 0; k < nk; k++) {
            float lo[3], hi[3];
            pad_box(kids[k].box, tiny, lo, hi);
            for (int a = 0; a < 3; a++) { padded[k].lo[a] = lo[a]; padded[k].hi[a] = hi[a]; }
            nb.grow(padded[k]);
        }
        int ex[3];
        float scale[3];
        for (int a = 0; a < 3; a++) {
            const float ext = std::max(nb.hi[a] - nb.lo[a], 1e-30f);
            int e = (int)std::ceil(std::log2(ext / 255.0f));
            // q = ceil((hi - p) / 2^e) must fit 255 with the rounding of the division
            while (std::ceil((nb.hi[a] - nb.lo[a]) / std::ldexp(1.0f, e)) > 255.0f) e++;
            e = std::max(-126, std::min(127, e));
            ex[a] = e + 127;
            scale[a] = std::ldexp(1.0f, e);
        }
        // slots: greedy assignment of children to the sign octants of their offset from the node centre
        int slot_of[8], kid_in[8];
        for (int k = 0; k < 8; k++) { slot_of[k] = -1; kid_in[k] = -1; }
        {
            float c0[3];
            for (int a = 0; a < 3; a++) c0[a] = 0.5f * (nb.lo[a] + nb.hi[a]);
            float cost[8][8];
            for (int k = 0; k < nk; k++)
                for (int sl = 0; sl < 8; sl++) {
                    float d = 0.f;
                    for (int a = 0; a < 3; a++) {
                        const float off = 0.5f * (kids[k].box.lo[a] + kids[k].box.hi[a]) - c0[a];
                        d += ((sl >> a) & 1) ? -off : off;
                    }
                    cost[k][sl] = d;
                }
            for (int n = 0; n < nk; n++) {
                int bk = -1, bs = -1;
                float bc = -std::numeric_limits<float>::infinity();
                for (int k = 0; k < nk; k++) {
                    if (slot_of[k] >= 0) continue;
                    for (int sl = 0; sl < 8; sl++)
                        if (kid_in[sl] < 0 && cost[k][sl] > bc) { bc = cost[k][sl]; bk = k; bs = sl; }
                }
                slot_of[bk] = bs; kid_in[bs] = bk;
            }
        }
        // children: inner ones get consecutive nodes in slot order; leaf triangles are appended to tris8
        uint32_t imask = 0;
        int n_inner = 0;
        for (int sl = 0; sl < 8; sl++) if (kid_in[sl] >= 0 && !kids[kid_in[sl]].leaf) { imask |= 1u << sl; n_inner++; }
        const int32_t child_base = n_inner ? alloc(n_inner) : 0;
        const uint32_t tri_base = (uint32_t)(out.tris8.size() / 3);
        uint8_t meta[8], q[6][8];
        int inner_seen = 0;
        uint32_t tri_off = 0;
        for (int sl = 0; sl < 8; sl++) {
            meta[sl] = 0;
            for (int r = 0; r < 6; r++) q[r][sl] = r < 3 ? 255 : 0;  // inverted box: never entered
            const int k = kid_in[sl];
            if (k < 0) continue;
            for (int a = 0; a < 3; a++) {
                float ql = std::floor((padded[k].lo[a] - nb.lo[a]) / scale[a]);
                float qh = std::ceil((padded[k].hi[a] - nb.lo[a]) / scale[a]);
                ql = std::max(0.f, std::min(255.f, ql)); qh = std::max(0.f, std::min(255.f, qh));
                // outward in the arithmetic the traversal uses (p + q * 2^e in float)
                while (ql > 0.f && nb.lo[a] + ql * scale[a] > padded[k].lo[a]) ql -= 1.f;
                while (qh < 255.f && nb.lo[a] + qh * scale[a] < padded[k].hi[a]) qh += 1.f;
                q[a][sl] = (uint8_t)ql; q[3 + a][sl] = (uint8_t)qh;
            }
            if (kids[k].leaf) {
                meta[sl] = (uint8_t)((((1u << kids[k].count) - 1u) << 5) | tri_off);
                for (uint32_t i = kids[k].first; i < kids[k].first + kids[k].count; i++) {
                    const int32_t prim = order[i];
                    const float* v = positions + (size_t)prim * 9;
                    f4 a = {v[0], v[1], v[2], 0.f};
                    std::memcpy(&a.w, &prim, 4);
                    out.tris8.push_back(a);
                    out.tris8.push_back(f4{v[3], v[4], v[5], 0.f});
                    out.tris8.push_back(f4{v[6], v[7], v[8], 0.f});
                }
                tri_off += kids[k].count;
            } else {
                meta[sl] = (uint8_t)((1u << 5) | (24u + (uint32_t)sl));
                todo.push_back({kids[k].tmp_idx, child_base + inner_seen, it.depth + 1});
                inner_seen++;
            }
        }
        auto pack4 = [](const uint8_t* b) { uint32_t u = (uint32_t)b[0] | ((uint32_t)b[1] << 8) | ((uint32_t)b[2] << 16) | ((uint32_t)b[3] << 24); float f; std::memcpy(&f, &u, 4); return f; };
        const uint8_t eb[4] = {(uint8_t)ex[0], (uint8_t)ex[1], (uint8_t)ex[2], (uint8_t)imask};
        float cb, tb;
        std::memcpy(&cb, &child_base, 4); std::memcpy(&tb, &tri_base, 4);
        f4* N = &out.nodes8[(size_t)it.out_idx * 5];
        N[0] = f4{nb.lo[0], nb.lo[1], nb.lo[2], pack4(eb)};
        N[1] = f4{cb, tb, pack4(meta), pack4(meta + 4)};
        N[2] = f4{pack4(q[0]), pack4(q[0] + 4), pack4(q[1]), pack4(q[1] + 4)};
        N[3] = f4{pack4(q[2]), pack4(q[2] + 4), pack4(q[3]), pack4(q[3] + 4)};
        N[4] = f4{pack4(q[4]), pack4(q[4] + 4), pack4(q[5]), pack4(q[5] + 4)};
    }
}

}  // namespace

// Tree -> the two traversal layouts of bvh.h.  Walks the tree from the root, so nodes a builder left unreachable
// (the GPU builder's subtrees below a collapsed leaf) cost nothing; inner nodes are numbered depth-first.
void emit_bvh(const float* positions, const BvhTree& T, Bvh& out)
{
    const float inf = std::numeric_limits<float>::infinity();
    const size_t count = T.order.size();
    if (count == 0) {
        BvhNode n; std::memset(&n, 0, sizeof n);
        n.n0 = {inf, -inf, inf, -inf}; n.n1 = n.n0; n.n2 = {inf, -inf, inf, -inf};
        n.c0 = n.c1 = ~0;
        out.nodes.push_back(n);
        for (int k = 0; k < 3; k++) out.tris.push_back(f4{0, 0, 0, 0});
        for (int k = 0; k < 6; k++) out.nodes4.push_back(k % 2 ? f4{-inf, -inf, -inf, -inf} : f4{inf, inf, inf, inf});
        { float nc; int32_t code = kBvhNoChild; std::memcpy(&nc, &code, 4); out.nodes4.push_back(f4{nc, nc, nc, nc}); }
        return;
    }
    const std::vector<TmpNode>& tmp = T.tmp;
    const int32_t root = T.root;
    const Box& all = tmp[root].box;
    for (int a = 0; a < 3; a++) { out.lo[a] = all.lo[a]; out.hi[a] = all.hi[a]; }
    float ext = std::max(std::max(all.hi[0] - all.lo[0], all.hi[1] - all.lo[1]), all.hi[2] - all.lo[2]);
    float tiny = std::max(ext, 1e-30f) * 1e-7f;

    // triangles in leaf order == T.order (leaves own contiguous ranges)
    out.tris.resize(count * 3);
    for (size_t i = 0; i < count; i++) {
        const int32_t prim = T.order[i];
        const float* v = positions + (size_t)prim * 9;
        f4 a = {v[0], v[1], v[2], 0.f};
        std::memcpy(&a.w, &prim, 4);
        out.tris[3 * i + 0] = a;
        out.tris[3 * i + 1] = f4{v[3], v[4], v[5], 0.f};
        out.tris[3 * i + 2] = f4{v[6], v[7], v[8], 0.f};
    }

    auto leaf_code = [](const TmpNode& t) { return (int32_t)~((t.first << 3) | (t.count - 1)); };
    auto set_child = [&](BvhNode& n, int which, const TmpNode& c, int32_t code) {
        float lo[3], hi[3];
        pad_box(c.box, tiny, lo, hi);
        if (which == 0) { n.n0 = {lo[0], hi[0], lo[1], hi[1]}; n.n2.x = lo[2]; n.n2.y = hi[2]; n.c0 = code; }
        else { n.n1 = {lo[0], hi[0], lo[1], hi[1]}; n.n2.z = lo[2]; n.n2.w = hi[2]; n.c1 = code; }
    };
    if (tmp[root].left < 0) {  // a lone root leaf gets a wrapper node
        BvhNode n; std::memset(&n, 0, sizeof n);
        set_child(n, 0, tmp[root], leaf_code(tmp[root]));
        n.n1 = {inf, -inf, inf, -inf}; n.n2.z = inf; n.n2.w = -inf; n.c1 = ~0;
        out.nodes.push_back(n);
        out.num_leaves = 1;
        out.sah_cost = kCostTrav + tmp[root].count * kCostTri;
        emit_wide(tmp, root, tiny, out);
        emit_wide8(positions, tmp, T.order, root, tiny, out);
        return;
    }
    const double root_area = std::max(tmp[root].box.half_area(), 1e-30f);
    double sah = 0;
    struct Item { int32_t tmp_idx, out_idx; uint32_t depth; };
    std::vector<Item> todo;
    out.nodes.reserve(count);
    out.nodes.emplace_back();
    todo.push_back({root, 0, 0});
    while (!todo.empty()) {
        const Item it = todo.back();
        todo.pop_back();
        const TmpNode& t = tmp[it.tmp_idx];
        sah += t.box.half_area() / root_area * kCostTrav;
        BvhNode n; std::memset(&n, 0, sizeof n);
        const int32_t kid[2] = {t.left, t.right};
        for (int k = 0; k < 2; k++) {
            const TmpNode& c = tmp[kid[k]];
            int32_t code;
            if (c.left < 0) {
                code = leaf_code(c);
                out.num_leaves++;
                sah += c.box.half_area() / root_area * c.count * kCostTri;
                out.max_depth = std::max(out.max_depth, it.depth + 1);
            } else {
                code = (int32_t)out.nodes.size();
                out.nodes.emplace_back();
                todo.push_back({kid[k], code, it.depth + 1});
            }
            set_child(n, k, c, code);
        }
        out.nodes[it.out_idx] = n;
    }
    out.sah_cost = sah;
    emit_wide(tmp, root, tiny, out);
    emit_wide8(positions, tmp, T.order, root, tiny, out);
}

bool build_tree_host(const float* positions, const int32_t* prims, size_t count, BvhTree& T)
{
    Builder B;
    B.refs.resize(count);
    for (size_t i = 0; i < count; i++) {
        Ref& r = B.refs[i];
        r.prim = prims[i];
        const float* v = positions + (size_t)prims[i] * 9;
        r.box.reset();
        for (int k = 0; k < 3; k++) r.box.grow(v + 3 * k);
        for (int a = 0; a < 3; a++) r.c[a] = 0.5f * (r.box.lo[a] + r.box.hi[a]);
    }
    T.order.resize(count);
    if (count == 0) return true;
    B.tmp.reserve(count * 2);
    T.root = B.build(0, (uint32_t)count, 0);
    for (size_t i = 0; i < count; i++) T.order[i] = B.refs[i].prim;
    T.tmp.swap(B.tmp);
    return true;
}

Bvh build_bvh(const float* positions, const int32_t* prims, size_t count)
{
    Bvh out;
    BvhTree T;
    build_tree_host(positions, prims, count, T);
    emit_bvh(positions, T, out);
    return out;
}

}  // namespace mcpt
