#include "clusters.hpp"

#include <algorithm>
#include <cmath>
#include <limits>
#include <numeric>
#include <unordered_map>

namespace perch {
namespace {

constexpr int kGreedyMaxTris = 2000000;  // above this the O(T * neighbourhood) grouping is replaced by Morton chunks

uint32_t spread10(uint32_t v) {  // 10 bits -> every third bit
    v &= 0x3FF;
    v = (v | (v << 16)) & 0x030000FF;
    v = (v | (v << 8)) & 0x0300F00F;
    v = (v | (v << 4)) & 0x030C30C3;
    v = (v | (v << 2)) & 0x09249249;
    return v;
}

}  // namespace

ClusterSet build_clusters(const float* tri9, const std::vector<int>& low, const std::vector<int>& high) {
    ClusterSet out;
    const int n_models = (int)low.size();
    out.model_low.resize(n_models);
    out.model_high.resize(n_models);
    for (int m = 0; m < n_models; ++m) {
        const int t0 = low[m], nt = high[m] - low[m];
        out.model_low[m] = (int)(out.bounds.size() / 4);
        if (nt <= 0) {
            out.model_high[m] = out.model_low[m];
            continue;
        }
        float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (int t = 0; t < nt; ++t)
            for (int k = 0; k < 9; ++k) {
                float p = tri9[(size_t)(t0 + t) * 9 + k];
                if (std::isfinite(p)) { mn[k % 3] = std::min(mn[k % 3], p); mx[k % 3] = std::max(mx[k % 3], p); }
            }
        std::vector<uint32_t> code(nt);
        for (int t = 0; t < nt; ++t) {
            const float* p = tri9 + (size_t)(t0 + t) * 9;
            uint32_t q[3];
            for (int c = 0; c < 3; ++c) {
                float cen = (p[c] + p[3 + c] + p[6 + c]) / 3.0f;
                float ext = mx[c] - mn[c];
                float f = (ext > 0 && std::isfinite(cen)) ? (cen - mn[c]) / ext : 0.f;
                q[c] = (uint32_t)std::min(1023.0f, std::max(0.0f, f * 1023.0f));
            }
            code[t] = spread10(q[0]) | (spread10(q[1]) << 1) | (spread10(q[2]) << 2);
        }
        std::vector<int> order(nt);
        std::iota(order.begin(), order.end(), 0);
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return code[a] < code[b]; });
        // Greedy grouping: seeds in Morton order; a cluster grows by the unassigned triangle of the 27 surrounding
        // hash cells that keeps the merged AABB diagonal smallest.  Triangles with no free neighbour stay in short
        // clusters (padded with copies of their last member) rather than being merged with distant ones.
        double edge_sum = 0;
        int edge_n = 0;
        for (int t = 0; t < nt; ++t) {
            const float* p = tri9 + (size_t)(t0 + t) * 9;
            double e = std::sqrt(((double)p[3] - p[0]) * ((double)p[3] - p[0]) + ((double)p[4] - p[1]) * ((double)p[4] - p[1]) +
                                 ((double)p[5] - p[2]) * ((double)p[5] - p[2]));
            if (std::isfinite(e)) { edge_sum += e; ++edge_n; }
        }
        const double cell = std::max(2.5 * (edge_n ? edge_sum / edge_n : 0.0), 1e-12);
        const bool use_grid = nt <= kGreedyMaxTris && std::isfinite(cell) && std::isfinite((double)mn[0] + mn[1] + mn[2]);
        auto cell_of = [&](const double c[3], int64_t ijk[3]) {
            for (int a = 0; a < 3; ++a) {
                double f = std::isfinite(c[a]) ? (c[a] - mn[a]) / cell : 0.0;
                ijk[a] = (int64_t)std::min(std::max(f, 0.0), 2097150.0);
            }
        };
        auto key_of = [](const int64_t ijk[3]) { return (uint64_t)ijk[0] | ((uint64_t)ijk[1] << 21) | ((uint64_t)ijk[2] << 42); };
        std::unordered_map<uint64_t, std::vector<int>> grid;
        if (use_grid) {
            grid.reserve((size_t)nt);
            for (int t = 0; t < nt; ++t) {
                const float* p = tri9 + (size_t)(t0 + t) * 9;
                double c[3] = {((double)p[0] + p[3] + p[6]) / 3, ((double)p[1] + p[4] + p[7]) / 3, ((double)p[2] + p[5] + p[8]) / 3};
                int64_t ijk[3];
                cell_of(c, ijk);
                grid[key_of(ijk)].push_back(t);
            }
        }
        std::vector<char> used(nt, 0);
        for (int oi = 0; oi < nt; ++oi) {
            const int seed = order[oi];
            if (used[seed]) continue;
            used[seed] = 1;
            int members[kClusterTris];
            int n_mem = 0;
            members[n_mem++] = t0 + seed;
            if (use_grid) {
                double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
                auto grow = [&](const float* p, double l[3], double h[3]) {
                    for (int k = 0; k < 9; ++k) { l[k % 3] = std::min(l[k % 3], (double)p[k]); h[k % 3] = std::max(h[k % 3], (double)p[k]); }
                };
                grow(tri9 + (size_t)(t0 + seed) * 9, lo, hi);
                while (n_mem < kClusterTris) {
                    double c[3] = {0.5 * (lo[0] + hi[0]), 0.5 * (lo[1] + hi[1]), 0.5 * (lo[2] + hi[2])};
                    int64_t ijk[3];
                    cell_of(c, ijk);
                    double best = INFINITY;
                    int best_t = -1;
                    for (int64_t dz = -1; dz <= 1; ++dz)
                        for (int64_t dy = -1; dy <= 1; ++dy)
                            for (int64_t dx = -1; dx <= 1; ++dx) {
                                int64_t q[3] = {ijk[0] + dx, ijk[1] + dy, ijk[2] + dz};
                                if (q[0] < 0 || q[1] < 0 || q[2] < 0) continue;
                                auto it = grid.find(key_of(q));
                                if (it == grid.end()) continue;
                                for (int t : it->second) {
                                    if (used[t]) continue;
                                    double l[3] = {lo[0], lo[1], lo[2]}, h[3] = {hi[0], hi[1], hi[2]};
                                    grow(tri9 + (size_t)(t0 + t) * 9, l, h);
                                    double d = (h[0] - l[0]) * (h[0] - l[0]) + (h[1] - l[1]) * (h[1] - l[1]) + (h[2] - l[2]) * (h[2] - l[2]);
                                    if (d < best) { best = d; best_t = t; }
                                }
                            }
                    if (best_t < 0) break;
                    used[best_t] = 1;
                    members[n_mem++] = t0 + best_t;
                    grow(tri9 + (size_t)(t0 + best_t) * 9, lo, hi);
                }
            } else {
                for (int oj = oi + 1; oj < nt && n_mem < kClusterTris; ++oj)
                    if (!used[order[oj]]) { used[order[oj]] = 1; members[n_mem++] = t0 + order[oj]; }
            }
            for (int j = n_mem; j < kClusterTris; ++j) members[j] = members[n_mem - 1];
            const size_t rec = out.recs.size();
            out.recs.resize(rec + kClusterFloats);
            bool finite = true;
            double pts[3 * kClusterTris][3];
            for (int j = 0; j < kClusterTris; ++j) {
                const float* p = tri9 + (size_t)members[j] * 9;
                out.orig.push_back(members[j]);
                for (int k = 0; k < 9; ++k) {
                    out.recs[rec + (size_t)k * kClusterTris + j] = p[k];
                    if (!std::isfinite(p[k])) finite = false;
                    pts[3 * j + k / 3][k % 3] = p[k];
                }
            }
            float cx = 0.f, cy = 0.f, cz = 0.f, rad = INFINITY;
            if (finite) {
                // near-minimal enclosing ball: start at the AABB centre, walk towards the farthest point (Badoiu-Clarkson)
                const int np = 3 * kClusterTris;
                double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
                for (int i = 0; i < np; ++i)
                    for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], pts[i][a]); hi[a] = std::max(hi[a], pts[i][a]); }
                auto radius2 = [&](const double c[3], int* far) {
                    double r2 = -1;
                    for (int i = 0; i < np; ++i) {
                        double d = (pts[i][0] - c[0]) * (pts[i][0] - c[0]) + (pts[i][1] - c[1]) * (pts[i][1] - c[1]) + (pts[i][2] - c[2]) * (pts[i][2] - c[2]);
                        if (d > r2) { r2 = d; if (far) *far = i; }
                    }
                    return r2;
                };
                double c[3] = {0.5 * (lo[0] + hi[0]), 0.5 * (lo[1] + hi[1]), 0.5 * (lo[2] + hi[2])};
                double bestc[3] = {c[0], c[1], c[2]};
                double best2 = radius2(c, nullptr);
                for (int it = 0; it < 96; ++it) {
                    int far = 0;
                    radius2(c, &far);
                    for (int a = 0; a < 3; ++a) c[a] += (pts[far][a] - c[a]) / (it + 2.0);
                    double r2 = radius2(c, nullptr);
                    if (r2 < best2) { best2 = r2; bestc[0] = c[0]; bestc[1] = c[1]; bestc[2] = c[2]; }
                }
                cx = (float)bestc[0]; cy = (float)bestc[1]; cz = (float)bestc[2];
                const double cf[3] = {cx, cy, cz};  // radius about the ROUNDED centre
                // round up: the device test relies on |vertex - centre| <= radius
                rad = std::nextafter((float)(std::sqrt(radius2(cf, nullptr)) * (1.0 + 1e-6)), std::numeric_limits<float>::infinity());
                if (!std::isfinite(rad)) rad = INFINITY;
            }
            out.bounds.insert(out.bounds.end(), {cx, cy, cz, rad});
        }
        out.model_high[m] = (int)(out.bounds.size() / 4);
    }
    return out;
}

}  // namespace perch
