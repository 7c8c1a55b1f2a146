// csrc/light_lists.cpp -- host builder of the per-primitive survivor lists (light_lists.h).
#include "light_lists.h"
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_map>

namespace mcpt {

namespace {
constexpr uint32_t kMinLights = 33;            // <= 32 lights: k_lightsel's per-lane loop, no sweep to shorten
constexpr uint64_t kMaxPairs = 48ull << 20;    // primitive x light tests done on the host at scene creation
constexpr uint32_t kMaxLists = 1u << 16;       // distinct lists the counting sort of k_lsel_* is sized for
constexpr uint64_t kMaxEntries = 24ull << 20;  // 96 MB of list entries
inline uint32_t lsoa_offset(uint32_t l) { return (l >> 5) * (uint32_t)kLsoaBlock + (l & 31); }  // lights.cuh lsoa_index
}  // namespace

LightLists build_light_lists(const float* positions, const float* normals, const int32_t* light_of_prim, uint32_t P,
                             const int32_t* light_prim, const double* geom_normals, const double* light_area, uint32_t L, uint32_t Lpad)
{
    LightLists R;
    if (const char* e = std::getenv("MCPT_LIGHT_LISTS"))
        if (std::atoi(e) == 0) return R;
    if (L < kMinLights || Lpad <= L) return R;
    uint64_t listed = 0;
    for (uint32_t p = 0; p < P; p++) listed += light_of_prim[p] < 0;
    if (listed == 0 || listed * (uint64_t)L > kMaxPairs) return R;
    const auto t0 = std::chrono::steady_clock::now();

    // lights in SoA, exactly the FP32 values the device table holds (api.cu: lsoa planes), widened to FP64
    std::vector<double> lv(9 * (size_t)L), lq(3 * (size_t)L), lc(L), lthr(L), lqa(3 * (size_t)L), lva(3 * (size_t)L);
    for (uint32_t l = 0; l < L; l++) {
        const float* v = positions + (size_t)light_prim[l] * 9;
        const double* g = geom_normals + (size_t)light_prim[l] * 3;
        const double a2 = 2.0 * light_area[l];
        double c = 0;
        for (int k = 0; k < 3; k++) {
            const double q = (double)(-(float)(g[k] * a2));  // plane 3 of lsoa: -2 area ng
            lq[(size_t)k * L + l] = q;
            lqa[(size_t)k * L + l] = std::fabs(q);
            c += q * (double)v[k];
            double m = 0;
            for (int j = 0; j < 3; j++) {
                lv[(size_t)(3 * j + k) * L + l] = (double)v[3 * j + k];
                m = std::max(m, std::fabs((double)v[3 * j + k]));
            }
            lva[(size_t)k * L + l] = m;
        }
        lc[l] = c;                                  // q . v0
        lthr[l] = (double)(float)(1e-8 * a2);       // plane 2 .w: the cull threshold of `num`
    }

    const uint32_t words = (L + 63) / 64;
    std::vector<uint64_t> mask(words);
    std::unordered_map<std::string, int32_t> ids;
    std::vector<std::vector<uint32_t>> lists;
    R.prim_list.assign(P, -1);
    std::vector<uint8_t> keep(L);
    double frac_sum = 0;
    uint64_t total_entries = 0;
    for (uint32_t p = 0; p < P; p++) {
        if (light_of_prim[p] >= 0) continue;
        const float* pv = positions + (size_t)p * 9;
        const float* pn = normals + (size_t)p * 9;
        double pmax[3] = {0, 0, 0};
        for (int j = 0; j < 3; j++)
            for (int k = 0; k < 3; k++) pmax[k] = std::max(pmax[k], std::fabs((double)pv[3 * j + k]));
        // min_j n_i . p_j per corner normal i
        double nb[3];
        for (int i = 0; i < 3; i++) {
            double m = 1e300;
            for (int j = 0; j < 3; j++) m = std::min(m, (double)pn[3 * i] * pv[3 * j] + (double)pn[3 * i + 1] * pv[3 * j + 1] + (double)pn[3 * i + 2] * pv[3 * j + 2]);
            nb[i] = m;
        }
        const double n0x = pn[0], n0y = pn[1], n0z = pn[2], n1x = pn[3], n1y = pn[4], n1z = pn[5], n2x = pn[6], n2y = pn[7], n2z = pn[8];
        const double na_x = std::max({std::fabs(n0x), std::fabs(n1x), std::fabs(n2x)}), na_y = std::max({std::fabs(n0y), std::fabs(n1y), std::fabs(n2y)}),
                     na_z = std::max({std::fabs(n0z), std::fabs(n1z), std::fabs(n2z)});
        const double p0x = pv[0], p0y = pv[1], p0z = pv[2], p1x = pv[3], p1y = pv[4], p1z = pv[5], p2x = pv[6], p2y = pv[7], p2z = pv[8];
        const double* qx = &lq[0]; const double* qy = &lq[L]; const double* qz = &lq[2 * (size_t)L];
        const double* qax = &lqa[0]; const double* qay = &lqa[L]; const double* qaz = &lqa[2 * (size_t)L];
        const double* vax = &lva[0]; const double* vay = &lva[L]; const double* vaz = &lva[2 * (size_t)L];
        for (uint32_t l = 0; l < L; l++) {
            // behind the light everywhere on the primitive?  num(x) = q . (v0 - x), culled when num < thr  (Mylight.cpp:178-185)
            const double m0 = qx[l] * p0x + qy[l] * p0y + qz[l] * p0z, m1 = qx[l] * p1x + qy[l] * p1y + qz[l] * p1z, m2 = qx[l] * p2x + qy[l] * p2y + qz[l] * p2z;
            const double num_max = lc[l] - std::min(m0, std::min(m1, m2));
            const double margin_b = 1e-5 * (qax[l] * (vax[l] + pmax[0]) + qay[l] * (vay[l] + pmax[1]) + qaz[l] * (vaz[l] + pmax[2]));
            // below the horizon everywhere?  sign of N(x) . (v_k - x) <= max_ijk n_i . (v_k - p_j)      (Mylight.cpp:187-197)
            double h = -1e300;
            for (int k = 0; k < 3; k++) {
                const double vx = lv[(size_t)(3 * k) * L + l], vy = lv[(size_t)(3 * k + 1) * L + l], vz = lv[(size_t)(3 * k + 2) * L + l];
                h = std::max(h, n0x * vx + n0y * vy + n0z * vz - nb[0]);
                h = std::max(h, n1x * vx + n1y * vy + n1z * vz - nb[1]);
                h = std::max(h, n2x * vx + n2y * vy + n2z * vz - nb[2]);
            }
            const double margin_h = 1e-5 * (na_x * (vax[l] + pmax[0]) + na_y * (vay[l] + pmax[1]) + na_z * (vaz[l] + pmax[2]));
            keep[l] = (num_max >= lthr[l] - margin_b) && (h >= -margin_h);
        }
        std::fill(mask.begin(), mask.end(), 0);
        uint32_t n = 0;
        for (uint32_t l = 0; l < L; l++)
            if (keep[l]) { mask[l >> 6] |= 1ull << (l & 63); n++; }
        frac_sum += (double)n / L;
        std::string key((const char*)mask.data(), words * 8);
        auto it = ids.find(key);
        if (it == ids.end()) {
            if (lists.size() >= kMaxLists) return LightLists();
            std::vector<uint32_t> ls;
            ls.reserve((n + 31) / 32 * 32);
            for (uint32_t l = 0; l < L; l++)
                if (keep[l]) ls.push_back(lsoa_offset(l));
            while (ls.size() % 32) ls.push_back(lsoa_offset(L));  // a padding light of the table: weight 0
            R.max_len = std::max<uint32_t>(R.max_len, (uint32_t)ls.size());
            total_entries += ls.size();
            if (total_entries > kMaxEntries) return LightLists();
            it = ids.emplace(std::move(key), (int32_t)lists.size()).first;
            lists.push_back(std::move(ls));
        }
        R.prim_list[p] = it->second;
    }
    R.nlists = (uint32_t)lists.size();
    R.range.resize(2 * (size_t)R.nlists);
    R.entries.reserve(total_entries);
    for (uint32_t i = 0; i < R.nlists; i++) {
        R.range[2 * i] = (int32_t)R.entries.size();
        R.range[2 * i + 1] = (int32_t)lists[i].size();
        R.entries.insert(R.entries.end(), lists[i].begin(), lists[i].end());
    }
    R.mean_fraction = frac_sum / (double)listed;
    R.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return R;
}

}  // namespace mcpt
