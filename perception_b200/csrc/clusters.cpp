#include "clusters.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <numeric>
#include <thread>
#include <unordered_map>

namespace perch {
namespace {

constexpr int kGreedyMaxUnits = 2000000;  // above this the O(N * neighbourhood) grouping is replaced by Morton chunks

uint32_t spread10(uint32_t v) {  // 10 bits -> every third bit
    v &= 0x3FF;
    v = (v | (v << 16)) & 0x030000FF;
    v = (v | (v << 8)) & 0x0300F00F;
    v = (v | (v << 4)) & 0x030C30C3;
    v = (v | (v << 2)) & 0x09249249;
    return v;
}

struct VKey {
    uint32_t a, b, c;
    bool operator==(const VKey& o) const { return a == o.a && b == o.b && c == o.c; }
};
struct VKeyHash {
    size_t operator()(const VKey& k) const {
        uint64_t h = (uint64_t)k.a * 0x9E3779B97F4A7C15ull;
        h ^= ((uint64_t)k.b + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full;
        h ^= ((uint64_t)k.c + 0x165667B1ull) * 0x9FB21C651E98DF25ull;
        return (size_t)(h ^ (h >> 29));
    }
};

// A quad: one or two triangles over (at most) four distinct vertices.
struct Unit {
    int tri[2];        // model-local triangle indices; tri[1] < 0 for a single triangle
    float v[4][3];     // quad vertices (see ClusterSet::recs)
    uint32_t perm[2];  // per triangle: quad vertex of its v0, v1, v2
};

// Near-minimal enclosing ball of np points: start at the AABB centre, walk towards the farthest point
// (Badoiu-Clarkson); the radius is taken about the ROUNDED (float) centre and rounded up, so that the device test may
// rely on |vertex - centre| <= radius.  Non-finite input: radius = inf.
void enclosing_ball(const double (*pts)[3], int np, float out4[4]) {
    out4[0] = out4[1] = out4[2] = 0.f;
    out4[3] = INFINITY;
    for (int i = 0; i < np; ++i)
        for (int a = 0; a < 3; ++a)
            if (!std::isfinite(pts[i][a])) return;
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
    out4[0] = (float)bestc[0]; out4[1] = (float)bestc[1]; out4[2] = (float)bestc[2];
    const double cf[3] = {out4[0], out4[1], out4[2]};
    float rad = std::nextafter((float)(std::sqrt(radius2(cf, nullptr)) * (1.0 + 1e-6)), std::numeric_limits<float>::infinity());
    out4[3] = std::isfinite(rad) ? rad : INFINITY;
}

}  // namespace

ClusterSet build_clusters(const float* tri9, const std::vector<int>& low, const std::vector<int>& high) {
    const int n_models = (int)low.size();
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    // one model: triangles paired into quads, quads into clusters, records written (bounding spheres: pass 4 below)
    auto build_one = [&](int m, ClusterSet& out) {
        const int t0 = low[m], nt = high[m] - low[m];
        if (nt <= 0) return;
        const float* T9 = tri9 + (size_t)t0 * 9;
        float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (int t = 0; t < nt; ++t)
            for (int k = 0; k < 9; ++k) {
                float p = T9[(size_t)t * 9 + k];
                if (std::isfinite(p)) { mn[k % 3] = std::min(mn[k % 3], p); mx[k % 3] = std::max(mx[k % 3], p); }
            }
        auto morton_of = [&](const double c[3]) {
            uint32_t q[3];
            for (int a = 0; a < 3; ++a) {
                double ext = (double)mx[a] - mn[a];
                double f = (ext > 0 && std::isfinite(c[a])) ? (c[a] - mn[a]) / ext : 0.0;
                q[a] = (uint32_t)std::min(1023.0, std::max(0.0, f * 1023.0));
            }
            return spread10(q[0]) | (spread10(q[1]) << 1) | (spread10(q[2]) << 2);
        };
        // ---- 1. vertex ids by exact bit pattern; triangles by edge ------------------------------------------------
        std::vector<int> vid((size_t)nt * 3);
        {
            std::unordered_map<VKey, int, VKeyHash> seen;
            seen.reserve((size_t)nt * 2);
            for (int t = 0; t < nt; ++t)
                for (int v = 0; v < 3; ++v) {
                    VKey k;
                    memcpy(&k, T9 + (size_t)t * 9 + 3 * v, 12);
                    auto it = seen.find(k);
                    if (it == seen.end()) it = seen.emplace(k, (int)seen.size()).first;
                    vid[(size_t)t * 3 + v] = it->second;
                }
        }
        auto distinct = [&](int t) {
            const int* v = &vid[(size_t)t * 3];
            return v[0] != v[1] && v[1] != v[2] && v[0] != v[2];
        };
        auto edge_key = [](int a, int b) { return ((uint64_t)(uint32_t)std::min(a, b) << 32) | (uint32_t)std::max(a, b); };
        std::unordered_map<uint64_t, std::vector<int>> by_edge;
        by_edge.reserve((size_t)nt * 2);
        for (int t = 0; t < nt; ++t) {
            if (!distinct(t)) continue;
            const int* v = &vid[(size_t)t * 3];
            for (int e = 0; e < 3; ++e) by_edge[edge_key(v[e], v[(e + 1) % 3])].push_back(t);
        }
        std::vector<uint32_t> tcode(nt);
        for (int t = 0; t < nt; ++t) {
            const float* p = T9 + (size_t)t * 9;
            double c[3] = {((double)p[0] + p[3] + p[6]) / 3, ((double)p[1] + p[4] + p[7]) / 3, ((double)p[2] + p[5] + p[8]) / 3};
            tcode[t] = morton_of(c);
        }
        std::vector<int> torder(nt);
        std::iota(torder.begin(), torder.end(), 0);
        std::stable_sort(torder.begin(), torder.end(), [&](int a, int b) { return tcode[a] < tcode[b]; });
        // ---- 2. greedy pairing into quads: the free edge neighbour with the smallest joint extent -------------------
        std::vector<char> paired(nt, 0);
        std::vector<Unit> units;
        units.reserve((size_t)nt / 2 + 16);
        auto extent2 = [&](int a, int b) {
            double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
            for (int t : {a, b})
                for (int k = 0; k < 9; ++k) {
                    double p = T9[(size_t)t * 9 + k];
                    lo[k % 3] = std::min(lo[k % 3], p); hi[k % 3] = std::max(hi[k % 3], p);
                }
            double d = 0;
            for (int a3 = 0; a3 < 3; ++a3) d += (hi[a3] - lo[a3]) * (hi[a3] - lo[a3]);
            return std::isfinite(d) ? d : INFINITY;
        };
        for (int oi = 0; oi < nt; ++oi) {
            const int t = torder[oi];
            if (paired[t]) continue;
            paired[t] = 1;
            Unit u;
            u.tri[0] = t; u.tri[1] = -1;
            const float* p = T9 + (size_t)t * 9;
            const int* v = &vid[(size_t)t * 3];
            int mate = -1, mate_edge = -1;
            if (distinct(t)) {
                double best = INFINITY;
                for (int e = 0; e < 3; ++e) {
                    auto it = by_edge.find(edge_key(v[e], v[(e + 1) % 3]));
                    if (it == by_edge.end()) continue;
                    for (int o : it->second) {
                        if (paired[o]) continue;
                        double d = extent2(t, o);
                        if (mate < 0 || d < best) { best = d; mate = o; mate_edge = e; }
                    }
                }
            }
            if (mate < 0) {
                // single triangle: vertices in their own order, vertex 3 repeats vertex 2
                for (int i = 0; i < 3; ++i) memcpy(u.v[i], p + 3 * i, 12);
                memcpy(u.v[3], p + 6, 12);
                u.perm[0] = 0u | (1u << 2) | (2u << 4);
                u.perm[1] = kPermInvalid;
            } else {
                paired[mate] = 1;
                u.tri[1] = mate;
                const int e0 = mate_edge, e1 = (mate_edge + 1) % 3, e2 = (mate_edge + 2) % 3;  // t: shared edge (e0,e1), own e2
                memcpy(u.v[0], p + 3 * e0, 12);
                memcpy(u.v[1], p + 3 * e1, 12);
                memcpy(u.v[2], p + 3 * e2, 12);
                uint32_t pa = 0;
                const int slot_of_t[3] = {e0, e1, e2};
                for (int i = 0; i < 3; ++i)
                    for (int k = 0; k < 3; ++k)
                        if (slot_of_t[k] == i) pa |= (uint32_t)k << (2 * i);
                u.perm[0] = pa;
                // mate: its vertices equal to v[e0] / v[e1] map to quad vertices 0 / 1, the remaining one is vertex 3
                // (or vertex 2 when the mate has the same three vertices)
                const float* q = T9 + (size_t)mate * 9;
                const int* w = &vid[(size_t)mate * 3];
                memcpy(u.v[3], u.v[2], 12);
                uint32_t pb = 0;
                for (int i = 0; i < 3; ++i) {
                    uint32_t k;
                    if (w[i] == v[e0]) k = 0;
                    else if (w[i] == v[e1]) k = 1;
                    else if (w[i] == v[e2]) k = 2;
                    else { k = 3; memcpy(u.v[3], q + 3 * i, 12); }
                    pb |= k << (2 * i);
                }
                u.perm[1] = pb;
            }
            units.push_back(u);
        }
        // ---- 3. clusters of two neighbouring quads: seeds in Morton order, mate = smallest merged AABB ------------------
        const int nu = (int)units.size();
        std::vector<double> ucen((size_t)nu * 3);
        std::vector<uint32_t> ucode(nu);
        double edge_sum = 0;
        int edge_n = 0;
        for (int i = 0; i < nu; ++i) {
            double c[3] = {0, 0, 0};
            for (int k = 0; k < 4; ++k)
                for (int a = 0; a < 3; ++a) c[a] += units[i].v[k][a] / 4.0;
            for (int a = 0; a < 3; ++a) ucen[(size_t)i * 3 + a] = c[a];
            ucode[i] = morton_of(c);
            double e = 0;
            for (int a = 0; a < 3; ++a) e += ((double)units[i].v[1][a] - units[i].v[0][a]) * ((double)units[i].v[1][a] - units[i].v[0][a]);
            e = std::sqrt(e);
            if (std::isfinite(e)) { edge_sum += e; ++edge_n; }
        }
        std::vector<int> uorder(nu);
        std::iota(uorder.begin(), uorder.end(), 0);
        std::stable_sort(uorder.begin(), uorder.end(), [&](int a, int b) { return ucode[a] < ucode[b]; });
        const double cell = std::max(2.5 * (edge_n ? edge_sum / edge_n : 0.0), 1e-12);
        const bool use_grid = nu <= kGreedyMaxUnits && std::isfinite(cell) && std::isfinite((double)mn[0] + mn[1] + mn[2]);
        auto cell_of = [&](const double c[3], int64_t ijk[3]) {
            for (int a = 0; a < 3; ++a) {
                double f = std::isfinite(c[a]) ? (c[a] - mn[a]) / cell : 0.0;
                ijk[a] = (int64_t)std::min(std::max(f, 0.0), 2097150.0);
            }
        };
        auto key_of = [](const int64_t ijk[3]) { return (uint64_t)ijk[0] | ((uint64_t)ijk[1] << 21) | ((uint64_t)ijk[2] << 42); };
        std::unordered_map<uint64_t, std::vector<int>> grid;
        if (use_grid) {
            grid.reserve((size_t)nu);
            for (int i = 0; i < nu; ++i) {
                int64_t ijk[3];
                cell_of(&ucen[(size_t)i * 3], ijk);
                grid[key_of(ijk)].push_back(i);
            }
        }
        auto grow = [&](const Unit& u, double l[3], double h[3]) {
            for (int k = 0; k < 4; ++k)
                for (int a = 0; a < 3; ++a) { l[a] = std::min(l[a], (double)u.v[k][a]); h[a] = std::max(h[a], (double)u.v[k][a]); }
        };
        std::vector<char> used(nu, 0);
        for (int oi = 0; oi < nu; ++oi) {
            const int seed = uorder[oi];
            if (used[seed]) continue;
            used[seed] = 1;
            int second = -1;
            if (use_grid) {
                double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
                grow(units[seed], lo, hi);
                int64_t ijk[3];
                cell_of(&ucen[(size_t)seed * 3], ijk);
                double best = INFINITY;
                for (int64_t dz = -1; dz <= 1; ++dz)
                    for (int64_t dy = -1; dy <= 1; ++dy)
                        for (int64_t dx = -1; dx <= 1; ++dx) {
                            int64_t q[3] = {ijk[0] + dx, ijk[1] + dy, ijk[2] + dz};
                            if (q[0] < 0 || q[1] < 0 || q[2] < 0) continue;
                            auto it = grid.find(key_of(q));
                            if (it == grid.end()) continue;
                            for (int o : it->second) {
                                if (used[o]) continue;
                                double l[3] = {lo[0], lo[1], lo[2]}, h[3] = {hi[0], hi[1], hi[2]};
                                grow(units[o], l, h);
                                double d = (h[0] - l[0]) * (h[0] - l[0]) + (h[1] - l[1]) * (h[1] - l[1]) + (h[2] - l[2]) * (h[2] - l[2]);
                                if (d < best) { best = d; second = o; }
                            }
                        }
            } else {
                for (int oj = oi + 1; oj < nu; ++oj)
                    if (!used[uorder[oj]]) { second = uorder[oj]; break; }
            }
            if (second >= 0) used[second] = 1;
            // ---- record: 2 quads x 12 coordinates; a cluster without a second quad repeats the first, invalid
            const Unit& qa = units[seed];
            const Unit& qb = units[second >= 0 ? second : seed];
            const size_t rec = out.recs.size();
            out.recs.resize(rec + kClusterFloats);
            uint32_t perm = 0;
            for (int q = 0; q < kClusterQuads; ++q) {
                const Unit& u = q ? qb : qa;
                for (int i = 0; i < kQuadVerts; ++i)
                    for (int a = 0; a < 3; ++a) {
                        const float p = u.v[i][a];
                        out.recs[rec + (size_t)q * kQuadFloats + 3 * i + a] = p;
                    }
                const bool repeat = q == 1 && second < 0;
                for (int j = 0; j < 2; ++j) {
                    const bool own = !repeat && u.tri[j] >= 0;
                    out.orig.push_back(t0 + (u.tri[j] >= 0 ? u.tri[j] : u.tri[0]));
                    perm |= (own ? u.perm[j] : kPermInvalid) << (6 * (2 * q + j));
                }
            }
            out.perm.push_back(perm);
            // bounding spheres: filled in by the parallel pass below (they are 90 % of this function's time)
            out.qbounds.resize(out.qbounds.size() + 4 * kClusterQuads, 0.f);
            out.bounds.resize(out.bounds.size() + 4, 0.f);
        }
    };
    // models are independent: filed by a pool of threads, concatenated in model order (same result as one after another)
    std::vector<ClusterSet> parts(n_models);
    {
        const int n_thr = (int)std::min<size_t>(std::min<unsigned>(hw, 16), (size_t)n_models);
        if (n_thr <= 1) {
            for (int m = 0; m < n_models; ++m) build_one(m, parts[m]);
        } else {
            std::vector<std::thread> pool;
            for (int i = 0; i < n_thr; ++i)
                pool.emplace_back([&, i] { for (int m = i; m < n_models; m += n_thr) build_one(m, parts[m]); });
            for (auto& th : pool) th.join();
        }
    }
    ClusterSet out;
    out.model_low.resize(n_models);
    out.model_high.resize(n_models);
    for (int m = 0; m < n_models; ++m) {
        ClusterSet& pm = parts[m];
        out.model_low[m] = (int)(out.bounds.size() / 4);
        out.recs.insert(out.recs.end(), pm.recs.begin(), pm.recs.end());
        out.orig.insert(out.orig.end(), pm.orig.begin(), pm.orig.end());
        out.perm.insert(out.perm.end(), pm.perm.begin(), pm.perm.end());
        out.bounds.insert(out.bounds.end(), pm.bounds.begin(), pm.bounds.end());
        out.qbounds.insert(out.qbounds.end(), pm.qbounds.begin(), pm.qbounds.end());
        out.model_high[m] = (int)(out.bounds.size() / 4);
        pm = ClusterSet();
    }
    // ---- 4. bounding spheres of every cluster and of its quads, from the records (threads over cluster ranges; each
    // sphere depends on its own record only, so the result does not depend on the thread count) -----------------------
    const size_t n_clusters = out.bounds.size() / 4;
    auto balls = [&](size_t c0, size_t c1) {
        for (size_t c = c0; c < c1; ++c) {
            double pts[kQuadVerts * kClusterQuads][3];
            for (int i = 0; i < kQuadVerts * kClusterQuads; ++i)
                for (int a = 0; a < 3; ++a) pts[i][a] = out.recs[c * kClusterFloats + 3 * (size_t)i + a];
            enclosing_ball(pts, kQuadVerts * kClusterQuads, &out.bounds[4 * c]);
            // a quad slot without a triangle of its own repeats the cluster's first quad (its record does): harmless (level
            // 2 drops invalid slots) and keeps level 1 free of a special case
            for (int q = 0; q < kClusterQuads; ++q) enclosing_ball(pts + q * kQuadVerts, kQuadVerts, &out.qbounds[4 * (kClusterQuads * c + q)]);
        }
    };
    const size_t n_thr = std::min<size_t>(std::min<size_t>(hw, 16), n_clusters / 512 + 1);
    if (n_thr <= 1) balls(0, n_clusters);
    else {
        std::vector<std::thread> pool;
        for (size_t i = 0; i < n_thr; ++i) pool.emplace_back(balls, n_clusters * i / n_thr, n_clusters * (i + 1) / n_thr);
        for (auto& th : pool) th.join();
    }
    return out;
}

}  // namespace perch
