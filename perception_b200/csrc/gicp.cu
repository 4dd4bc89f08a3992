// Batched Generalized-ICP refinement of candidate poses: the `do_icp` block of render_cuda_multi_unified
// (renderer.cu:1688-1817), which hands every pose's rendered cloud (source) and the observed cloud (target, restricted
// to the pose's segmentation label in 6-DoF) to fast_gicp::FastGICPCudaCore::optimize_multi with
// k_correspondences = 10, max 150 iterations, rotation epsilon 2e-3, translation epsilon 5e-4 (renderer.cu:1696-1705),
// and composes the result with the pose (concatenate_transforms, renderer.cu:1412-1429).
//
// fast_gicp is an un-vendored dependency of the reference with no pinned version, so this is a
// restatement of the published algorithm (Segal et al., "Generalized-ICP"; Koide et al., "Voxelized GICP"), pinned
// against oracle/perch_oracle.c:oracle_gicp (double precision) rather than against reference outputs:
//   * covariances from the k nearest neighbours of every point within its own cloud (the point itself included),
//     regularised to eigenvalues (1, 1, 1e-3) -- i.e. C = I - (1 - 1e-3) n n^T with n the direction of least variance,
//     so only the normal n is stored (3 floats instead of a 6-float covariance);
//   * per iteration: exact nearest target point of every transformed source point (no distance gate),
//     M = (C_B + R C_A R^T)^-1, residual r = b - (R a + t), Jacobian [skew(R a + t), -I] (rotation first),
//     H = sum J^T M J, g = sum J^T M r, delta = -H^-1 g (Cholesky, double), T <- exp(delta) T;
//     stop when every |exp(delta).R - I| entry < rotation_epsilon and every |exp(delta).t| entry < translation_epsilon.
//
// One CTA per pose: source points and normals live in shared memory for the whole optimisation, nearest neighbours go
// through the observed-cloud grid (ring expansion, exact), the 27 + 1 sums are reduced in a fixed order (warp shuffles,
// then one thread over the warps) so results are run-to-run deterministic, and the 6x6 solve stays on the device:
// no host round trip per iteration (the reference's cuda_icp needs two, icp.cu:156-218).
#include <cuda_runtime.h>

#include <cfloat>

#include "kernels.cuh"

namespace perch {
namespace {

#ifndef PERCH_GICP_THREADS
#define PERCH_GICP_THREADS 32
#endif
constexpr int kGicpThreads = PERCH_GICP_THREADS;
constexpr int kGicpWarps = kGicpThreads / 32;
constexpr int kMaxRings = 12;        // nearest-neighbour shells of the coarse grid searched before a linear scan
constexpr int kFineRings = 6;        // shells of the fine grid searched before moving to the coarse one
// Measured on the 233k-point scene cloud (tools/gicp_dense_probe.py, 4096 poses x 20 linearisations): rows per group
// 1 / 2 / 4 / 8 -> 8.86 / 8.80 / 9.42 / 10.8 ms (pruning by the running best beats loads in flight); point loads per
// round 2 / 4 / 8 -> 10.5 / 8.80 / 8.86 ms; 16 / 20 / 24 resident warps per SM (124 / 96 / 80 registers) -> 9.97 / 9.42 / 9.29 ms.
#ifndef PERCH_GICP_BALL_ROWS
#define PERCH_GICP_BALL_ROWS 2
#endif
constexpr int kBallRows = PERCH_GICP_BALL_ROWS;
#ifndef PERCH_GICP_SCAN_UNROLL
#define PERCH_GICP_SCAN_UNROLL 4
#endif
constexpr int kScanUnroll = PERCH_GICP_SCAN_UNROLL;  // point loads in flight per lane while scanning a grid run
constexpr float kPlaneEps = 1e-3f;  // fast_gicp RegularizationMethod::PLANE: singular values (1, 1, 1e-3)

// Eigenvector of the smallest eigenvalue of the symmetric matrix [a00 a01 a02; a01 a11 a12; a02 a12 a22]
// (cyclic Jacobi, 6 sweeps).  T = float on the device, double in the oracle's copy of this routine.
template <class T>
__host__ __device__ inline void smallest_eigenvector(T a00, T a01, T a02, T a11, T a12, T a22, T n[3]) {
    T A[3][3] = {{a00, a01, a02}, {a01, a11, a12}, {a02, a12, a22}};
    T V[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int sweep = 0; sweep < 6; ++sweep) {
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                const T apq = A[p][q];
                if (apq == (T)0) continue;
                const T theta = (A[q][q] - A[p][p]) / ((T)2 * apq);
                const T sgn = theta >= (T)0 ? (T)1 : (T)-1;
                const T t = sgn / (sgn * theta + sqrt(theta * theta + (T)1));
                const T c = (T)1 / sqrt(t * t + (T)1), s = t * c;
                for (int k = 0; k < 3; ++k) {  // A <- A J
                    const T akp = A[k][p], akq = A[k][q];
                    A[k][p] = c * akp - s * akq;
                    A[k][q] = s * akp + c * akq;
                }
                for (int k = 0; k < 3; ++k) {  // A <- J^T A
                    const T apk = A[p][k], aqk = A[q][k];
                    A[p][k] = c * apk - s * aqk;
                    A[q][k] = s * apk + c * aqk;
                }
                for (int k = 0; k < 3; ++k) {
                    const T vkp = V[k][p], vkq = V[k][q];
                    V[k][p] = c * vkp - s * vkq;
                    V[k][q] = s * vkp + c * vkq;
                }
            }
    }
    int m = 0;
    if (A[1][1] < A[m][m]) m = 1;
    if (A[2][2] < A[m][m]) m = 2;
    n[0] = V[0][m]; n[1] = V[1][m]; n[2] = V[2][m];
}

// normal of point i from its k nearest neighbours (given as indices into xyz AoS); fewer than 3 neighbours => n = 0
// (covariance = identity, plain point-to-point weighting)
__device__ inline void normal_from_neighbours(const float* xyz_aos, const int* nbr, int count, float n[3]) {
    n[0] = n[1] = n[2] = 0.f;
    if (count < 3) return;
    float mx = 0, my = 0, mz = 0;
    for (int j = 0; j < count; ++j) {
        const float* p = xyz_aos + 3 * (size_t)nbr[j];
        mx += p[0]; my += p[1]; mz += p[2];
    }
    const float inv = 1.0f / (float)count;
    mx *= inv; my *= inv; mz *= inv;
    float c00 = 0, c01 = 0, c02 = 0, c11 = 0, c12 = 0, c22 = 0;
    for (int j = 0; j < count; ++j) {
        const float* p = xyz_aos + 3 * (size_t)nbr[j];
        const float x = p[0] - mx, y = p[1] - my, z = p[2] - mz;
        c00 += x * x; c01 += x * y; c02 += x * z; c11 += y * y; c12 += y * z; c22 += z * z;
    }
    smallest_eigenvector<float>(c00 * inv, c01 * inv, c02 * inv, c11 * inv, c12 * inv, c22 * inv, n);
}

// the same on the float4 cloud (x, y, z, .): identical operations in identical order
__device__ inline void normal_from_neighbours4(const float4* xyz4, const int* nbr, int count, float n[3]) {
    n[0] = n[1] = n[2] = 0.f;
    if (count < 3) return;
    float mx = 0, my = 0, mz = 0;
    for (int j = 0; j < count; ++j) {
        const float4 p = __ldg(xyz4 + nbr[j]);
        mx += p.x; my += p.y; mz += p.z;
    }
    const float inv = 1.0f / (float)count;
    mx *= inv; my *= inv; mz *= inv;
    float c00 = 0, c01 = 0, c02 = 0, c11 = 0, c12 = 0, c22 = 0;
    for (int j = 0; j < count; ++j) {
        const float4 p = __ldg(xyz4 + nbr[j]);
        const float x = p.x - mx, y = p.y - my, z = p.z - mz;
        c00 += x * x; c01 += x * y; c02 += x * z; c11 += y * y; c12 += y * z; c22 += z * z;
    }
    smallest_eigenvector<float>(c00 * inv, c01 * inv, c02 * inv, c11 * inv, c12 * inv, c22 * inv, n);
}

// ---- target normals: neighbours come from launch_knn (k x N index matrix) -------------------------------------------
__global__ void target_normals_kernel(const float* xyz_aos, const int* knn_index, int n, int k, float* normals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int nbr[32];
    const int kk = min(k, 32);
    for (int j = 0; j < kk; ++j) nbr[j] = knn_index[(size_t)j * n + i];
    float nn[3];
    normal_from_neighbours(xyz_aos, nbr, kk, nn);
    normals[3 * (size_t)i] = nn[0]; normals[3 * (size_t)i + 1] = nn[1]; normals[3 * (size_t)i + 2] = nn[2];
}

// ---- target normals through the search grid (k = 10, the reference's hard-coded value) --------------------------------
// The same K nearest neighbours as the brute-force k-NN above finds -- ascending (squared distance, index), the point
// itself included, labels ignored -- but gathered by ring expansion around the point's cell, so a 300k-point scene cloud
// costs milliseconds instead of the ~80 ms of an N x N scan and never leaves the device.  The K best live in registers;
// insertion compares (distance, index) explicitly, so the list does not depend on the order cells are visited in.
// After ring r the unvisited points are at least r * h away: the search stops once the K-th distance is below that.  A
// point whose neighbourhood is still incomplete after kNormalRings rings (an isolated point) starts over with a linear
// scan of the cloud -- same result, rare.
constexpr int kNormalRings = 8;

template <int K>
struct KBest {
    float bd[K];
    int bi[K];
    __device__ void clear() {
#pragma unroll
        for (int s = 0; s < K; ++s) { bd[s] = FLT_MAX; bi[s] = 0x7fffffff; }
    }
    __device__ void offer(float d, int j) {
        if (d > bd[K - 1] || (d == bd[K - 1] && j >= bi[K - 1])) return;
#pragma unroll
        for (int s = K - 1; s > 0; --s) {
            const bool shift = bd[s - 1] > d || (bd[s - 1] == d && bi[s - 1] > j);   // slot s-1 is worse than (d, j)
            const bool here = !shift && (bd[s] > d || (bd[s] == d && bi[s] > j));   // first slot from the left that is worse
            const float nd = shift ? bd[s - 1] : (here ? d : bd[s]);
            const int ni = shift ? bi[s - 1] : (here ? j : bi[s]);
            bd[s] = nd; bi[s] = ni;
        }
        if (bd[0] > d || (bd[0] == d && bi[0] > j)) { bd[0] = d; bi[0] = j; }
    }
};

template <int K>
__global__ void target_normals_grid_kernel(ObsGridDev g, float h, const float4* __restrict__ pts_sorted, int n, float* normals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 q = __ldg(pts_sorted + i);
    KBest<K> kb;
    kb.clear();
    const int cx = min(max((int)floorf((q.x - g.ox) * g.inv_h), 0), g.dx - 1);
    const int cy = min(max((int)floorf((q.y - g.oy) * g.inv_h), 0), g.dy - 1);
    const int cz = min(max((int)floorf((q.z - g.oz) * g.inv_h), 0), g.dz - 1);
    const int rmax = max(max(max(cx, g.dx - 1 - cx), max(cy, g.dy - 1 - cy)), max(cz, g.dz - 1 - cz));
    bool done = false;
    for (int r = 0; r <= min(rmax, kNormalRings) && !done; ++r) {
        const int z0 = max(cz - r, 0), z1 = min(cz + r, g.dz - 1);
        const int y0 = max(cy - r, 0), y1 = min(cy + r, g.dy - 1);
        for (int z = z0; z <= z1; ++z) {
            const bool zface = (z == cz - r) || (z == cz + r);
            for (int y = y0; y <= y1; ++y) {
                const bool face = zface || (y == cy - r) || (y == cy + r);
                const int row = (z * g.dy + y) * g.dx;
                for (int part = 0; part < 2; ++part) {
                    int xa, xb;
                    if (face) { if (part) break; xa = max(cx - r, 0); xb = min(cx + r, g.dx - 1); }
                    else {
                        if (r == 0) break;
                        xa = xb = part ? cx + r : cx - r;
                        if (xa < 0 || xa > g.dx - 1) continue;
                    }
                    const uint32_t b = __ldg(g.cell_start + row + xa), e = __ldg(g.cell_start + row + xb + 1);
                    for (uint32_t j = b; j < e; ++j) {
                        const float4 p = __ldg(g.pts + j);
                        kb.offer(sqdist3(q.x, q.y, q.z, p.x, p.y, p.z), __float_as_int(p.w));
                    }
                }
            }
        }
        // rings 0..r complete: every unvisited point is at least r * h away (ring_search's bound, one ring later)
        const float lim = (float)r * h;
        done = (kb.bi[K - 1] != 0x7fffffff && kb.bd[K - 1] < lim * lim * 0.9999f) || r == rmax;
    }
    if (!done) {
        kb.clear();
        for (int j = 0; j < n; ++j) {
            const float4 p = __ldg(pts_sorted + j);
            kb.offer(sqdist3(q.x, q.y, q.z, p.x, p.y, p.z), j);
        }
    }
    int nbr[K], cnt = 0;
#pragma unroll
    for (int s = 0; s < K; ++s) {
        nbr[s] = kb.bi[s];
        cnt += kb.bi[s] != 0x7fffffff;
    }
    float nn[3];
    normal_from_neighbours4(pts_sorted, nbr, cnt, nn);
    normals[3 * (size_t)i] = nn[0]; normals[3 * (size_t)i + 1] = nn[1]; normals[3 * (size_t)i + 2] = nn[2];
}

// Distance test of the grid points [b, e) of one x-run: four independent 16-byte loads in flight (a lane's search is a
// chain of dependent latencies otherwise; profiles/r2_gicp_dense_v2.md: 55 % of the stall samples on this load).  The tail
// re-reads point e - 1, which cannot change the result (strict `<`, equal distance only wins with a lower index).
__device__ __forceinline__ void scan_run(const ObsGridDev& g, uint32_t b, uint32_t e, float qx, float qy, float qz, bool use_label,
                                         int label, float& best, int& best_ref, float4& best_pt) {
    for (uint32_t j = b; j < e; j += kScanUnroll) {
        float4 p[kScanUnroll];
        int lab[kScanUnroll];
#pragma unroll
        for (int u = 0; u < kScanUnroll; ++u) {
            const uint32_t jj = min(j + u, e - 1);
            p[u] = __ldg(g.pts + jj);
            lab[u] = use_label ? __ldg(g.label + jj) : label;
        }
#pragma unroll
        for (int u = 0; u < kScanUnroll; ++u) {
            if (lab[u] != label) continue;
            const float d = sqdist3(qx, qy, qz, p[u].x, p[u].y, p[u].z);
            const int ref = __float_as_int(p[u].w);
            if (d < best || (d == best && ref < best_ref)) { best = d; best_ref = ref; best_pt = p[u]; }
        }
    }
}

// Ring expansion over one uniform grid, carrying the best candidate in and out.  Returns true when the result is final:
// the best squared distance is below the bound of the unvisited cells, or every ring of the grid has been visited.
// Returns false after max_rings rings without that certainty.  Same distance arithmetic and tie-break (lowest reference
// index) as the cost stage (exact_math.cuh:sqdist3).
__device__ inline bool ring_search(const ObsGridDev& g, float h, int max_rings, float qx, float qy, float qz, bool use_label,
                                   int label, float& best, int& best_ref, float4& best_pt) {
    const int cx = min(max((int)floorf((qx - g.ox) * g.inv_h), 0), g.dx - 1);
    const int cy = min(max((int)floorf((qy - g.oy) * g.inv_h), 0), g.dy - 1);
    const int cz = min(max((int)floorf((qz - g.oz) * g.inv_h), 0), g.dz - 1);
    const int rmax = max(max(max(cx, g.dx - 1 - cx), max(cy, g.dy - 1 - cy)), max(cz, g.dz - 1 - cz));
    for (int r = 0; r <= rmax; ++r) {
        if (best_ref >= 0) {
            const float lim = (float)(r - 1) * h;  // ring r-1 complete => unvisited points are >= (r-1)*h away ... (*)
            if (r >= 1 && best < lim * lim * 0.9999f) return true;
        }
        if (r > max_rings) return false;
        const int z0 = max(cz - r, 0), z1 = min(cz + r, g.dz - 1);
        const int y0 = max(cy - r, 0), y1 = min(cy + r, g.dy - 1);
        // Cells are pruned by their exact distance from the query: a point of cell (x, y, z) is at least
        // gap(qx, x)^2 + gap(qy, y)^2 + gap(qz, z)^2 away, gap = distance from the coordinate to the cell's interval
        // [o + i h, o + (i + 1) h].  A row (y, z) whose bound exceeds the best distance so far is skipped without a load,
        // and the x-run of a visited row is shortened from both ends the same way: after the query's own cell (ring 0)
        // a dense cloud needs a handful of the 26 cells of ring 1 instead of all of them.  (1.0002: the bound is computed
        // in float; `>` keeps equal distances, which the lowest-index tie-break needs; every gap is shortened by 0.1 % of
        // a cell, far more than the rounding of a point's cell assignment (floor((p - o) * inv_h), inv_h ~ 1 / h).)
        const float slack = 1e-3f * h;
        auto gap2 = [&](float q, float o, int i) {
            const float lo = fmaf((float)i, h, o), d = fmaxf(fmaxf(lo - q, q - (lo + h)) - slack, 0.f);
            return d * d;
        };
        for (int z = z0; z <= z1; ++z) {
            const bool zface = (z == cz - r) || (z == cz + r);
            const float gz = gap2(qz, g.oz, z);
            if (gz > best * 1.0002f) continue;
            for (int y = y0; y <= y1; ++y) {
                const bool face = zface || (y == cy - r) || (y == cy + r);
                const float gyz = gz + gap2(qy, g.oy, y);
                if (gyz > best * 1.0002f) continue;
                const int row = (z * g.dy + y) * g.dx;
                // on a face of the ring cube the whole x-run belongs to the ring, otherwise only its two end cells
                for (int part = 0; part < 2; ++part) {
                    int xa, xb;
                    if (face) { if (part) break; xa = max(cx - r, 0); xb = min(cx + r, g.dx - 1); }
                    else {
                        if (r == 0) break;
                        xa = xb = part ? cx + r : cx - r;
                        if (xa < 0 || xa > g.dx - 1) continue;
                    }
                    while (xa <= xb && gyz + gap2(qx, g.ox, xa) > best * 1.0002f) ++xa;
                    while (xb >= xa && gyz + gap2(qx, g.ox, xb) > best * 1.0002f) --xb;
                    if (xa > xb) continue;
                    const uint32_t b = __ldg(g.cell_start + row + xa), e = __ldg(g.cell_start + row + xb + 1);
                    scan_run(g, b, e, qx, qy, qz, use_label, label, best, best_ref, best_pt);
                }
            }
        }
    }
    return true;
    // (*) a query inside cell c is at least (rho - 1)*h from any point of a cell at Chebyshev distance rho (also when
    // the query lies outside the grid and c is the clamped cell); after rings 0..r-1 the unvisited cells have rho >= r,
    // hence the bound (r-1)*h used before visiting ring r.
}

// Exact nearest neighbour when a candidate within a few cell widths is already known (the correspondence of the previous
// linearisation: poses move little between iterations): only the cells that intersect the ball of the candidate's distance
// around the query can hold a nearer point (or an equally near one with a lower reference index) -- typically 1-4 cells
// instead of the 27 a cold search must visit, and one dependent load chain instead of three rings of them.  `best` /
// `best_ref` / `best_pt` carry the candidate in and the nearest neighbour out; same distance arithmetic and tie-break as
// ring_search.
__device__ inline void ball_search(const ObsGridDev& g, float h, float qx, float qy, float qz, bool use_label, int label,
                                   float& best, int& best_ref, float4& best_pt) {
    const float slack = 1e-3f * h;
    const float r = sqrtf(best) * 1.0001f + slack;
    const int x0 = min(max((int)floorf((qx - r - g.ox) * g.inv_h), 0), g.dx - 1), x1 = min(max((int)floorf((qx + r - g.ox) * g.inv_h), 0), g.dx - 1);
    const int y0 = min(max((int)floorf((qy - r - g.oy) * g.inv_h), 0), g.dy - 1), y1 = min(max((int)floorf((qy + r - g.oy) * g.inv_h), 0), g.dy - 1);
    const int z0 = min(max((int)floorf((qz - r - g.oz) * g.inv_h), 0), g.dz - 1), z1 = min(max((int)floorf((qz + r - g.oz) * g.inv_h), 0), g.dz - 1);
    auto gap2 = [&](float q, float o, int i) {
        const float lo = fmaf((float)i, h, o), d = fmaxf(fmaxf(lo - q, q - (lo + h)) - slack, 0.f);
        return d * d;
    };
    for (int z = z0; z <= z1; ++z) {
        const float gz = gap2(qz, g.oz, z);
        if (gz > best * 1.0002f) continue;
        // rows in groups of kBallRows: their range bounds are requested together, then scanned (a row pruned by a bound that
        // an earlier row of its group would have tightened is scanned needlessly, never wrongly)
        for (int yb = y0; yb <= y1; yb += kBallRows) {
            uint32_t rb[kBallRows], re[kBallRows];
#pragma unroll
            for (int u = 0; u < kBallRows; ++u) {
                const int y = yb + u;
                const float gyz = gz + gap2(qy, g.oy, min(y, y1));
                const float wx = sqrtf(fmaxf(best * 1.0002f - gyz, 0.f)) * 1.0001f + slack;  // the ball's chord along x in this row
                const int xa = max(x0, (int)floorf((qx - wx - g.ox) * g.inv_h)), xb = min(x1, (int)floorf((qx + wx - g.ox) * g.inv_h));
                const bool live = y <= y1 && !(gyz > best * 1.0002f) && xa <= xb;
                const int row = (z * g.dy + min(y, y1)) * g.dx;
                rb[u] = live ? __ldg(g.cell_start + row + xa) : 0u;
                re[u] = live ? __ldg(g.cell_start + row + xb + 1) : 0u;
            }
#pragma unroll
            for (int u = 0; u < kBallRows; ++u) scan_run(g, rb[u], re[u], qx, qy, qz, use_label, label, best, best_ref, best_pt);
        }
    }
}

// Exact nearest observed point (optionally of one label), no distance gate: ring expansion over the fine grid (cells
// sized for a few points each), then -- for queries farther from every candidate than its rings reach -- over the
// coarse grid of the cost stage with the best candidate carried along, and only for outliers beyond that (a pose far
// from the scene, a label with few points) a linear scan of the point array.  All three give the same answer.
__device__ inline void nn_unbounded(const ObsGridDev& fine, float h_fine, const ObsGridDev& coarse, float h_coarse,
                                    float qx, float qy, float qz, bool use_label, int label, float& best, int& best_ref,
                                    float4& best_pt) {
    best = FLT_MAX;
    best_ref = -1;
    const bool two_level = coarse.pts != fine.pts;
    if (ring_search(fine, h_fine, two_level ? kFineRings : kMaxRings, qx, qy, qz, use_label, label, best, best_ref, best_pt)) return;
    if (two_level && ring_search(coarse, h_coarse, kMaxRings, qx, qy, qz, use_label, label, best, best_ref, best_pt))
        return;
    for (int j = 0; j < fine.n_obs; ++j) {
        if (use_label && __ldg(fine.label + j) != label) continue;
        const float4 p = __ldg(fine.pts + j);
        const float d = sqdist3(qx, qy, qz, p.x, p.y, p.z);
        const int ref = __float_as_int(p.w);
        if (d < best || (d == best && ref < best_ref)) { best = d; best_ref = ref; best_pt = p; }
    }
}

// ball_search for all points of a pose at once, by one warp, with every lane busy.  A per-lane ball_search leaves most of
// the warp idle: the queries of a round need very different numbers of rows and points (ncu, 233k-point cloud: 8 of 32
// lanes active on average).  Here the search is a flat loop in which every lane performs ONE step per trip -- open the next
// grid row of its ball, or test up to kScanUnroll points of the row it has open -- and a lane whose query is finished takes
// the next unassigned point of the pose in the same trip, so the warp stays convergent and lanes run dry only at the very
// end.  Per query the visiting order, the pruning rule and the arithmetic are ball_search's (rows pruned one at a time by
// the running best), so refs[] is exactly what the per-lane search returns.  Points whose previous correspondence is -1
// (no target point carries the label) are searched cold, by their lane alone.
__device__ inline void warm_search_warp(const GicpArgs& a, const float* pts, int n, const float* R, const float* t, int* refs,
                                        bool use_label, int label, int lane, int warp) {
    const ObsGridDev& g = a.grid;
    const float h = a.grid_h, slack = 1e-3f * h;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int n_mine = (n - warp + kGicpWarps - 1) / kGicpWarps;  // this warp owns points warp, warp + kGicpWarps, ...
    auto gap2 = [&](float q, float o, int i) {
        const float lo = fmaf((float)i, h, o), d = fmaxf(fmaxf(lo - q, q - (lo + h)) - slack, 0.f);
        return d * d;
    };
    int next = 0;  // warp-uniform: points handed out so far
    bool active = false;
    int i = 0, best_ref = -1;
    float qx = 0.f, qy = 0.f, qz = 0.f, best = 0.f, gz = 0.f;
    int x0 = 0, x1 = 0, y0 = 0, y1 = 0, z1 = 0, y = 0, z = 0;
    uint32_t j = 0, e = 0;
    for (;;) {
        const unsigned idle = __ballot_sync(0xffffffffu, !active);
        if (idle) {
            const int my = next + __popc(idle & lt_mask);
            next += __popc(idle);
            if (!active && my < n_mine) {
                i = warp + kGicpWarps * my;
                const float ax = pts[3 * i], ay = pts[3 * i + 1], az = pts[3 * i + 2];
                qx = R[0] * ax + R[1] * ay + R[2] * az + t[0];
                qy = R[3] * ax + R[4] * ay + R[5] * az + t[1];
                qz = R[6] * ax + R[7] * ay + R[8] * az + t[2];
                const int r0 = refs[i];
                if (r0 < 0) {
                    float4 bp;
                    nn_unbounded(a.grid, a.grid_h, a.coarse, a.coarse_h, qx, qy, qz, use_label, label, best, best_ref, bp);
                    refs[i] = best_ref;
                } else {
                    const float4 p0 = __ldg(a.tgt_sorted + r0);
                    best = sqdist3(qx, qy, qz, p0.x, p0.y, p0.z);
                    best_ref = r0;
                    const float r = sqrtf(best) * 1.0001f + slack;
                    x0 = min(max((int)floorf((qx - r - g.ox) * g.inv_h), 0), g.dx - 1); x1 = min(max((int)floorf((qx + r - g.ox) * g.inv_h), 0), g.dx - 1);
                    y0 = min(max((int)floorf((qy - r - g.oy) * g.inv_h), 0), g.dy - 1); y1 = min(max((int)floorf((qy + r - g.oy) * g.inv_h), 0), g.dy - 1);
                    z = min(max((int)floorf((qz - r - g.oz) * g.inv_h), 0), g.dz - 1); z1 = min(max((int)floorf((qz + r - g.oz) * g.inv_h), 0), g.dz - 1);
                    y = y0 - 1;
                    gz = gap2(qz, g.oz, z);
                    j = e = 0;
                    active = true;
                }
            }
        }
        if (next >= n_mine && __ballot_sync(0xffffffffu, active) == 0u) break;
        if (active && j >= e) {
            // open the next row of the ball (one candidate per trip; a pruned row costs a trip without a load)
            ++y;
            if (y > y1) {
                ++z;
                y = y0;
                if (z <= z1) gz = gap2(qz, g.oz, z);
            }
            if (z > z1) {
                refs[i] = best_ref;
                active = false;
            } else if (gz > best * 1.0002f) {
                y = y1;  // the whole slice is out of reach: the next trip moves on to z + 1
            } else {
                const float gyz = gz + gap2(qy, g.oy, y);
                if (!(gyz > best * 1.0002f)) {
                    // the ball's chord along x in this row: points of the row are at least sqrt(gyz) away in (y, z)
                    const float wx = sqrtf(fmaxf(best * 1.0002f - gyz, 0.f)) * 1.0001f + slack;
                    const int xa = max(x0, (int)floorf((qx - wx - g.ox) * g.inv_h)), xb = min(x1, (int)floorf((qx + wx - g.ox) * g.inv_h));
                    if (xa <= xb) {
                        const int row = (z * g.dy + y) * g.dx;
                        j = __ldg(g.cell_start + row + xa);
                        e = __ldg(g.cell_start + row + xb + 1);
                    }
                }
            }
        }
        if (active && j < e) {
            float4 p[kScanUnroll];
            int lab[kScanUnroll];
#pragma unroll
            for (int u = 0; u < kScanUnroll; ++u) {
                const uint32_t jj = min(j + u, e - 1);
                p[u] = __ldg(g.pts + jj);
                lab[u] = use_label ? __ldg(g.label + jj) : label;
            }
#pragma unroll
            for (int u = 0; u < kScanUnroll; ++u) {
                if (lab[u] != label) continue;
                const float d = sqdist3(qx, qy, qz, p[u].x, p[u].y, p[u].z);
                const int ref = __float_as_int(p[u].w);
                if (d < best || (d == best && ref < best_ref)) { best = d; best_ref = ref; }
            }
            j += kScanUnroll;
        }
    }
}

// warm_search_warp on the pyramid of grids (GicpArgs::levels), for every linearisation including the first.
// * Level: a ball of radius r opens ~ (r / h)^2 rows and tests ~ density * (2 r h + h^2) points on cells of size h, so the
//   best cell size grows like cbrt(r / density): GicpArgs::level_k.  On the finest grid alone a query 5 cm from a sparse
//   target opens ~ 700 rows (C1: 626 source points per pose, most of them centimetres from the observed cloud: 46 ms per
//   2 000 poses x 20); on cells of r / 2 a dense target (40 points per cm^2) costs thousands of point tests per query.
// * Seed: the previous correspondence when there is one; otherwise the first point (of the pose's label) in the query's own
//   cell on the finest level where that cell holds one -- all levels are probed at once, the loads are independent -- and
//   as a last resort the first point of the label's range.  Any seed gives the exact result: the ball search visits every
//   cell that can hold a nearer point (or an equally near one with a lower index).
// Same per-query arithmetic, pruning and tie-break as ball_search; the level only decides which cells are enumerated.
__device__ inline void pyramid_search_warp(const GicpArgs& a, const float* pts, int n, const float* R, const float* t, int* refs,
                                           bool use_label, int label, int lb, int le, bool use_prev, int lane, int warp) {
    const unsigned lt_mask = (1u << lane) - 1u;
    const int n_mine = (n - warp + kGicpWarps - 1) / kGicpWarps;
    const int L = a.n_levels;
    int next = 0;
    bool active = false;
    int i = 0, best_ref = -1;
    float qx = 0.f, qy = 0.f, qz = 0.f, best = 0.f, gz = 0.f;
    float ox = 0.f, oy = 0.f, oz = 0.f, h = 1.f, slack = 0.f;
    int dx = 1, dy = 1;
    const uint32_t* cell_start = nullptr;
    const float4* gpts = nullptr;
    const int* glabel = nullptr;
    int x0 = 0, x1 = 0, y0 = 0, y1 = 0, z1 = 0, y = 0, z = 0;
    uint32_t j = 0, e = 0;
    auto gap2 = [&](float q, float o, int c) {
        const float lo = fmaf((float)c, h, o), d = fmaxf(fmaxf(lo - q, q - (lo + h)) - slack, 0.f);
        return d * d;
    };
    for (;;) {
        const unsigned idle = __ballot_sync(0xffffffffu, !active);
        if (idle) {
            const int my = next + __popc(idle & lt_mask);
            next += __popc(idle);
            if (!active && my < n_mine) {
                i = warp + kGicpWarps * my;
                const float ax = pts[3 * i], ay = pts[3 * i + 1], az = pts[3 * i + 2];
                qx = R[0] * ax + R[1] * ay + R[2] * az + t[0];
                qy = R[3] * ax + R[4] * ay + R[5] * az + t[1];
                qz = R[6] * ax + R[7] * ay + R[8] * az + t[2];
                int seed = use_prev ? refs[i] : -1;
                if (seed < 0) {
                    // the query's own cell on every level, finest first
                    for (int l = 0; l < L && seed < 0; ++l) {
                        const ObsGridDev& g = a.levels[l];
                        const int cx = min(max((int)floorf((qx - g.ox) * g.inv_h), 0), g.dx - 1);
                        const int cy = min(max((int)floorf((qy - g.oy) * g.inv_h), 0), g.dy - 1);
                        const int cz = min(max((int)floorf((qz - g.oz) * g.inv_h), 0), g.dz - 1);
                        const int c = (cz * g.dy + cy) * g.dx + cx;
                        const uint32_t b0 = __ldg(g.cell_start + c), e0 = __ldg(g.cell_start + c + 1);
                        // the nearest of the cell's first few points (of the pose's label): a closer seed is a smaller ball
                        float sd = FLT_MAX;
                        int seen = 0;
                        for (uint32_t k = b0; k < e0 && seen < 8; ++k) {
                            if (use_label && __ldg(g.label + k) != label) continue;
                            const float4 p = __ldg(g.pts + k);
                            const float d = sqdist3(qx, qy, qz, p.x, p.y, p.z);
                            if (d < sd) { sd = d; seed = __float_as_int(p.w); }
                            ++seen;
                        }
                    }
                    if (seed < 0 && le > lb) seed = lb;  // any point of the label (the cloud is sorted by label)
                }
                if (seed < 0) {
                    refs[i] = -1;  // no target point carries this label
                } else {
                    const float4 p0 = __ldg(a.tgt_sorted + seed);
                    best = sqdist3(qx, qy, qz, p0.x, p0.y, p0.z);
                    best_ref = seed;
                    const float r0 = sqrtf(best);
                    // level whose cells are nearest (in ratio) to h* = cbrt(level_k * r), but never so fine that the ball spans
                    // more than ~16 cells per axis
                    const float h_want = fmaxf(cbrtf(a.level_k * r0), r0 * 0.125f) * 0.70710678f;
                    int l = 0;
                    while (l + 1 < L && 1.0f / a.levels[l].inv_h < h_want) ++l;
                    const ObsGridDev& g = a.levels[l];
                    ox = g.ox; oy = g.oy; oz = g.oz; h = 1.0f / g.inv_h; slack = 1e-3f * h;
                    dx = g.dx; dy = g.dy;
                    cell_start = g.cell_start; gpts = g.pts; glabel = g.label;
                    const float r = r0 * 1.0001f + slack;
                    x0 = min(max((int)floorf((qx - r - ox) * g.inv_h), 0), g.dx - 1); x1 = min(max((int)floorf((qx + r - ox) * g.inv_h), 0), g.dx - 1);
                    y0 = min(max((int)floorf((qy - r - oy) * g.inv_h), 0), g.dy - 1); y1 = min(max((int)floorf((qy + r - oy) * g.inv_h), 0), g.dy - 1);
                    z = min(max((int)floorf((qz - r - oz) * g.inv_h), 0), g.dz - 1); z1 = min(max((int)floorf((qz + r - oz) * g.inv_h), 0), g.dz - 1);
                    y = y0 - 1;
                    gz = gap2(qz, oz, z);
                    j = e = 0;
                    active = true;
                }
            }
        }
        if (next >= n_mine && __ballot_sync(0xffffffffu, active) == 0u) break;
        if (active && j >= e) {
            ++y;
            if (y > y1) {
                ++z;
                y = y0;
                if (z <= z1) gz = gap2(qz, oz, z);
            }
            if (z > z1) {
                refs[i] = best_ref;
                active = false;
            } else if (gz > best * 1.0002f) {
                y = y1;
            } else {
                const float gyz = gz + gap2(qy, oy, y);
                if (!(gyz > best * 1.0002f)) {
                    const float wx = sqrtf(fmaxf(best * 1.0002f - gyz, 0.f)) * 1.0001f + slack;
                    const float inv_h = 1.0f / h;
                    const int xa = max(x0, (int)floorf((qx - wx - ox) * inv_h)), xb = min(x1, (int)floorf((qx + wx - ox) * inv_h));
                    if (xa <= xb) {
                        const int row = (z * dy + y) * dx;
                        j = __ldg(cell_start + row + xa);
                        e = __ldg(cell_start + row + xb + 1);
                    }
                }
            }
        }
        if (active && j < e) {
            float4 p[kScanUnroll];
            int lab[kScanUnroll];
#pragma unroll
            for (int u = 0; u < kScanUnroll; ++u) {
                const uint32_t jj = min(j + u, e - 1);
                p[u] = __ldg(gpts + jj);
                lab[u] = use_label ? __ldg(glabel + jj) : label;
            }
#pragma unroll
            for (int u = 0; u < kScanUnroll; ++u) {
                if (lab[u] != label) continue;
                const float d = sqdist3(qx, qy, qz, p[u].x, p[u].y, p[u].z);
                const int ref = __float_as_int(p[u].w);
                if (d < best || (d == best && ref < best_ref)) { best = d; best_ref = ref; }
            }
            j += kScanUnroll;
        }
    }
}

// The same search over a short contiguous range of the label-sorted cloud (a label's points in 6-DoF, the whole cloud
// in 3-DoF): every lane streams the same points (one broadcast LDG.128 each), no divergence.  Ascending index with a
// strict `<` keeps the lowest index on ties, so the result is identical to nn_unbounded's.
__device__ inline void nn_range(const float4* __restrict__ tgt, int lb, int le, float qx, float qy, float qz, float& best,
                                int& best_ref, float4& best_pt) {
    best = FLT_MAX;
    best_ref = -1;
    int j = lb;
    for (; j + 4 <= le; j += 4) {
        const float4 p0 = __ldg(tgt + j), p1 = __ldg(tgt + j + 1), p2 = __ldg(tgt + j + 2), p3 = __ldg(tgt + j + 3);
        const float d0 = sqdist3(qx, qy, qz, p0.x, p0.y, p0.z), d1 = sqdist3(qx, qy, qz, p1.x, p1.y, p1.z);
        const float d2 = sqdist3(qx, qy, qz, p2.x, p2.y, p2.z), d3 = sqdist3(qx, qy, qz, p3.x, p3.y, p3.z);
        if (d0 < best) { best = d0; best_ref = j; }
        if (d1 < best) { best = d1; best_ref = j + 1; }
        if (d2 < best) { best = d2; best_ref = j + 2; }
        if (d3 < best) { best = d3; best_ref = j + 3; }
    }
    for (; j < le; ++j) {
        const float4 p = __ldg(tgt + j);
        const float d = sqdist3(qx, qy, qz, p.x, p.y, p.z);
        if (d < best) { best = d; best_ref = j; }
    }
    if (best_ref >= 0) best_pt = __ldg(tgt + best_ref);
}

// k nearest neighbours of point i inside the pose's own cloud (brute force, the point itself included), ascending by
// (distance, index): the K best live in registers (K is the reference's hard-coded 10 in practice, renderer.cu:1697).
template <int K>
__device__ inline int self_knn(const float* pts, int n, int i, int* out_idx) {
    float bd[K];
    int bi[K];
#pragma unroll
    for (int s = 0; s < K; ++s) { bd[s] = FLT_MAX; bi[s] = -1; }
    const float px = pts[3 * i], py = pts[3 * i + 1], pz = pts[3 * i + 2];
    for (int j = 0; j < n; ++j) {
        const float d = sqdist3(px, py, pz, pts[3 * j], pts[3 * j + 1], pts[3 * j + 2]);
        if (d >= bd[K - 1]) continue;      // ties keep the earlier (lower) index, like the generic insertion below
        // insert: shift every slot that is worse than d one place down
#pragma unroll
        for (int s = K - 1; s > 0; --s) {
            const bool shift = bd[s - 1] > d;
            const bool here = !shift && bd[s] > d;  // first slot from the left whose value exceeds d
            const float nd = shift ? bd[s - 1] : (here ? d : bd[s]);
            const int ni = shift ? bi[s - 1] : (here ? j : bi[s]);
            bd[s] = nd; bi[s] = ni;
        }
        if (bd[0] > d) { bd[0] = d; bi[0] = j; }
    }
    int cnt = 0;
#pragma unroll
    for (int s = 0; s < K; ++s) {
        out_idx[s] = bi[s];
        cnt += bi[s] >= 0;
    }
    return cnt;
}

struct Sym3 { float xx, xy, xz, yy, yz, zz; };

// inverse of the symmetric positive definite 2I - w(na na^T + nb nb^T)
__device__ inline Sym3 inverse_sum_cov(const float na[3], const float nb[3], float w) {
    const float a = 2.f - w * (na[0] * na[0] + nb[0] * nb[0]);
    const float b = -w * (na[0] * na[1] + nb[0] * nb[1]);
    const float c = -w * (na[0] * na[2] + nb[0] * nb[2]);
    const float d = 2.f - w * (na[1] * na[1] + nb[1] * nb[1]);
    const float e = -w * (na[1] * na[2] + nb[1] * nb[2]);
    const float f = 2.f - w * (na[2] * na[2] + nb[2] * nb[2]);
    const float A = d * f - e * e, B = c * e - b * f, C = b * e - c * d;
    const float inv = 1.0f / (a * A + b * B + c * C);
    Sym3 m;
    m.xx = A * inv; m.xy = B * inv; m.xz = C * inv;
    m.yy = (a * f - c * c) * inv; m.yz = (b * c - a * e) * inv; m.zz = (a * d - b * b) * inv;
    return m;
}

// Solve H x = -g for symmetric positive definite H (6x6, row-major upper part used) by Cholesky in double.
// Returns false when H is not positive definite (degenerate geometry): caller stops with delta = 0.
__device__ inline bool solve6(const double* H, const double* g, double* x) {
    double L[6][6];
    for (int i = 0; i < 6; ++i)
        for (int j = 0; j <= i; ++j) {
            double s = H[i * 6 + j];
            for (int k = 0; k < j; ++k) s -= L[i][k] * L[j][k];
            if (i == j) {
                if (!(s > 1e-300)) return false;
                L[i][i] = sqrt(s);
            } else L[i][j] = s / L[j][j];
        }
    double y[6];
    for (int i = 0; i < 6; ++i) {
        double s = -g[i];
        for (int k = 0; k < i; ++k) s -= L[i][k] * y[k];
        y[i] = s / L[i][i];
    }
    for (int i = 5; i >= 0; --i) {
        double s = y[i];
        for (int k = i + 1; k < 6; ++k) s -= L[k][i] * x[k];
        x[i] = s / L[i][i];
    }
    return true;
}

// exp of the twist (omega, v): R = exp([omega]x), t = V v
__device__ inline void se3_exp(const double* d, double R[9], double t[3]) {
    const double wx = d[0], wy = d[1], wz = d[2];
    const double th2 = wx * wx + wy * wy + wz * wz, th = sqrt(th2);
    double A, B, Cc;  // sin(th)/th, (1-cos th)/th^2, (th - sin th)/th^3
    if (th < 1e-6) { A = 1.0 - th2 / 6.0; B = 0.5 - th2 / 24.0; Cc = 1.0 / 6.0 - th2 / 120.0; }
    else { A = sin(th) / th; B = (1.0 - cos(th)) / th2; Cc = (th - sin(th)) / (th2 * th); }
    const double K[9] = {0, -wz, wy, wz, 0, -wx, -wy, wx, 0};
    double K2[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) K2[i * 3 + j] = K[i * 3] * K[j] + K[i * 3 + 1] * K[3 + j] + K[i * 3 + 2] * K[6 + j];
    double V[9];
    for (int i = 0; i < 9; ++i) {
        const double I = (i % 4 == 0) ? 1.0 : 0.0;
        R[i] = I + A * K[i] + B * K2[i];
        V[i] = I + B * K[i] + Cc * K2[i];
    }
    for (int i = 0; i < 3; ++i) t[i] = V[i * 3] * d[3] + V[i * 3 + 1] * d[4] + V[i * 3 + 2] * d[5];
}

__global__ void pose_count_max_kernel(const int* pose_begin, int rows_per_pose, int n_poses, int* out) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    int n = 0;
    if (p < n_poses) n = pose_begin[(size_t)(p + 1) * rows_per_pose] - pose_begin[(size_t)p * rows_per_pose];
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) n = max(n, __shfl_xor_sync(0xffffffffu, n, o));
    if ((threadIdx.x & 31) == 0 && n > 0) atomicMax(out, n);
}

#ifndef PERCH_GICP_MIN_BLOCKS
#define PERCH_GICP_MIN_BLOCKS 20
#endif
__global__ void __launch_bounds__(kGicpThreads, PERCH_GICP_MIN_BLOCKS) gicp_kernel(GicpArgs a) {
    extern __shared__ __align__(16) unsigned char gicp_smem[];
    float* s_pts = reinterpret_cast<float*>(gicp_smem);              // 3 x cap
    float* s_nrm = s_pts + 3 * (size_t)a.smem_points;                // 3 x cap
    int* s_ref = reinterpret_cast<int*>(s_nrm + 3 * (size_t)a.smem_points);  // correspondence of the last linearisation
    __shared__ double s_T[12];                                       // R row-major (9) + t (3), metres
    __shared__ double s_Tt[12];                                      // trial transform of the Levenberg-Marquardt step
    __shared__ double s_H[36], s_g[6], s_d[6];
    __shared__ double s_y0, s_den, s_lambda, s_nu;
    __shared__ double s_best_y, s_best_T[12];                        // stall detection (GicpArgs::stall_window)
    __shared__ int s_since_best;
    __shared__ float s_part[kGicpWarps][28];
    __shared__ int s_state;                                          // 0 = iterate, 1 = converged / stop
    __shared__ int s_lm;                                             // trial loop: 0 = try again, 1 = step decided, 2 = abort
    __shared__ int s_iters;

    const int pose_i = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int begin = a.pose_begin[(size_t)pose_i * a.rows_per_pose];
    const int end = a.pose_begin[(size_t)(pose_i + 1) * a.rows_per_pose];
    const int n = end - begin;
    const bool in_smem = n <= a.smem_points;
    const float* g_pts = a.cloud_aos + 3 * (size_t)begin;
    float* g_nrm = a.src_normals + 3 * (size_t)begin;
    const float* pts = in_smem ? s_pts : g_pts;
    float* nrm = in_smem ? s_nrm : g_nrm;
    int* refs = in_smem ? s_ref : a.src_ref + begin;
    const bool use_label = a.pose_label != nullptr;
    const int label = use_label ? a.pose_label[pose_i] : 0;
    // target range of this pose in the label-sorted cloud; short ranges are searched by streaming (nn_range)
    int lb = 0, le = a.grid.n_obs;
    if (use_label) {
        const bool known = label >= 0 && label < a.n_label_ranges;
        lb = known ? a.label_range[2 * label] : 0;
        le = known ? a.label_range[2 * label + 1] : 0;
    }
    const bool stream_target = a.tgt_sorted != nullptr && (le - lb) <= a.stream_max;

    if (tid < 12) s_T[tid] = (tid == 0 || tid == 4 || tid == 8) ? 1.0 : 0.0;
    if (tid == 0) { s_state = (n == 0) ? 1 : 0; s_iters = 0; s_lambda = -1.0; s_best_y = 1e300; s_since_best = 0; }
    if (in_smem)
        for (int i = tid; i < 3 * n; i += kGicpThreads) s_pts[i] = g_pts[i];
    __syncthreads();

    // ---- source normals: k nearest neighbours within the pose's own cloud (brute force, point itself included) -----
    const int k = min(a.k_correspondences, 32);
    for (int i = tid; i < n; i += kGicpThreads) {
        int bi[32];
        int cnt = 0;
        if (k == 10) {
            cnt = self_knn<10>(pts, n, i, bi);
        } else {
            float bd[32];
            const float px = pts[3 * i], py = pts[3 * i + 1], pz = pts[3 * i + 2];
            for (int j = 0; j < n; ++j) {
                const float d = sqdist3(px, py, pz, pts[3 * j], pts[3 * j + 1], pts[3 * j + 2]);
                if (cnt == k && d >= bd[k - 1]) continue;
                int pos = cnt < k ? cnt++ : k - 1;
                while (pos > 0 && bd[pos - 1] > d) { bd[pos] = bd[pos - 1]; bi[pos] = bi[pos - 1]; --pos; }
                bd[pos] = d; bi[pos] = j;
            }
        }
        float nn[3];
        normal_from_neighbours(pts, bi, cnt, nn);
        nrm[3 * i] = nn[0]; nrm[3 * i + 1] = nn[1]; nrm[3 * i + 2] = nn[2];
    }
    __syncthreads();

    const float w = 1.0f - kPlaneEps;
    for (int it = 0; it < a.max_iterations; ++it) {
        if (s_state) break;
        float R[9], t[3];
#pragma unroll
        for (int i = 0; i < 9; ++i) R[i] = (float)s_T[i];
#pragma unroll
        for (int i = 0; i < 3; ++i) t[i] = (float)s_T[9 + i];
        // ---- phase 1: correspondences of this linearisation -> refs[] ------------------------------------------------
        const bool flat = !stream_target && it > 0 && a.tgt_sorted != nullptr && a.warm_cells >= 1e5f;
        if (!stream_target && a.levels != nullptr && a.tgt_sorted != nullptr && a.warm_cells >= 1e5f) {
            pyramid_search_warp(a, pts, n, R, t, refs, use_label, label, lb, le, it > 0, lane, warp);
        } else if (flat) {
            warm_search_warp(a, pts, n, R, t, refs, use_label, label, lane, warp);
        } else {
            for (int i = tid; i < n; i += kGicpThreads) {
                const float ax = pts[3 * i], ay = pts[3 * i + 1], az = pts[3 * i + 2];
                const float px = R[0] * ax + R[1] * ay + R[2] * az + t[0];
                const float py = R[3] * ax + R[4] * ay + R[5] * az + t[1];
                const float pz = R[6] * ax + R[7] * ay + R[8] * az + t[2];
                float best; int ref; float4 b;
                if (stream_target) nn_range(a.tgt_sorted, lb, le, px, py, pz, best, ref, b);
                else {
                    // warm start from the previous linearisation's correspondence when it is within warm_cells cell widths
                    bool warm = false;
                    if (it > 0 && refs[i] >= 0) {
                        const int r0 = refs[i];
                        const float4 p0 = __ldg(a.tgt_sorted + r0);
                        const float d0 = sqdist3(px, py, pz, p0.x, p0.y, p0.z);
                        if (d0 < a.grid_h * a.grid_h * a.warm_cells * a.warm_cells) {
                            best = d0; ref = r0; b = p0;
                            ball_search(a.grid, a.grid_h, px, py, pz, use_label, label, best, ref, b);
                            warm = true;
                        }
                    }
                    if (!warm) nn_unbounded(a.grid, a.grid_h, a.coarse, a.coarse_h, px, py, pz, use_label, label, best, ref, b);
                }
                refs[i] = ref;
            }
        }
        __syncthreads();
        // ---- phase 2: the 21 + 6 + 1 sums of the linearisation (fixed assignment of points to lanes, fixed order) ---------
        float acc[28];
#pragma unroll
        for (int i = 0; i < 28; ++i) acc[i] = 0.f;
        for (int i = tid; i < n; i += kGicpThreads) {
            const int ref = refs[i];
            if (ref < 0) continue;  // no target point carries this label
            const float ax = pts[3 * i], ay = pts[3 * i + 1], az = pts[3 * i + 2];
            const float px = R[0] * ax + R[1] * ay + R[2] * az + t[0];
            const float py = R[3] * ax + R[4] * ay + R[5] * az + t[1];
            const float pz = R[6] * ax + R[7] * ay + R[8] * az + t[2];
            const float4 b = __ldg((a.tgt_sorted ? a.tgt_sorted : a.grid.pts) + ref);
            const float nax = nrm[3 * i], nay = nrm[3 * i + 1], naz = nrm[3 * i + 2];
            const float na[3] = {R[0] * nax + R[1] * nay + R[2] * naz, R[3] * nax + R[4] * nay + R[5] * naz,
                                 R[6] * nax + R[7] * nay + R[8] * naz};
            const float nb[3] = {a.tgt_normals[3 * (size_t)ref], a.tgt_normals[3 * (size_t)ref + 1], a.tgt_normals[3 * (size_t)ref + 2]};
            const Sym3 M = inverse_sum_cov(na, nb, w);
            const float rx = b.x - px, ry = b.y - py, rz = b.z - pz;
            // J = [skew(p), -I]: columns 0..2 = skew(p) columns, 3..5 = -e_x,-e_y,-e_z.  skew(p) = [0 -pz py; pz 0 -px; -py px 0]
            const float J[6][3] = {{0.f, pz, -py}, {-pz, 0.f, px}, {py, -px, 0.f}, {-1.f, 0.f, 0.f}, {0.f, -1.f, 0.f}, {0.f, 0.f, -1.f}};
            float MJ[6][3];
#pragma unroll
            for (int c = 0; c < 6; ++c) {
                MJ[c][0] = M.xx * J[c][0] + M.xy * J[c][1] + M.xz * J[c][2];
                MJ[c][1] = M.xy * J[c][0] + M.yy * J[c][1] + M.yz * J[c][2];
                MJ[c][2] = M.xz * J[c][0] + M.yz * J[c][1] + M.zz * J[c][2];
            }
            int o = 0;
#pragma unroll
            for (int r2 = 0; r2 < 6; ++r2)
#pragma unroll
                for (int c = r2; c < 6; ++c) acc[o++] += J[r2][0] * MJ[c][0] + J[r2][1] * MJ[c][1] + J[r2][2] * MJ[c][2];
#pragma unroll
            for (int c = 0; c < 6; ++c) acc[21 + c] += MJ[c][0] * rx + MJ[c][1] * ry + MJ[c][2] * rz;
            acc[27] += rx * (M.xx * rx + M.xy * ry + M.xz * rz) + ry * (M.xy * rx + M.yy * ry + M.yz * rz) +
                       rz * (M.xz * rx + M.yz * ry + M.zz * rz);
        }
#pragma unroll
        for (int i = 0; i < 28; ++i) {
            float v = acc[i];
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0) s_part[warp][i] = v;
        }
        __syncthreads();
        if (tid == 0) {
            int o = 0;
            for (int r2 = 0; r2 < 6; ++r2)
                for (int c = r2; c < 6; ++c) {
                    double sum = 0;
                    for (int wv = 0; wv < kGicpWarps; ++wv) sum += (double)s_part[wv][o];
                    s_H[r2 * 6 + c] = sum; s_H[c * 6 + r2] = sum;
                    ++o;
                }
            for (int c = 0; c < 6; ++c) {
                double sum = 0;
                for (int wv = 0; wv < kGicpWarps; ++wv) sum += (double)s_part[wv][21 + c];
                s_g[c] = sum;
            }
            double err = 0;
            for (int wv = 0; wv < kGicpWarps; ++wv) err += (double)s_part[wv][27];
            s_y0 = err;
            s_iters = it + 1;
            if (s_lambda < 0.0) {  // LsqRegistration: lm_lambda = lm_init_lambda_factor (1e-9) * H.diagonal().maxCoeff()
                double mx = 0;
                for (int c = 0; c < 6; ++c) mx = fmax(mx, s_H[c * 7]);
                s_lambda = 1e-9 * mx;
            }
            s_nu = 2.0;
            s_lm = 0;
            if (a.out_error) a.out_error[pose_i] = (float)err;
            if (a.stall_window > 0) {
                if (err < s_best_y * (1.0 - (double)a.stall_rel)) {
                    s_best_y = err;
                    for (int i = 0; i < 12; ++i) s_best_T[i] = s_T[i];
                    s_since_best = 0;
                } else if (++s_since_best >= a.stall_window) {
                    for (int i = 0; i < 12; ++i) s_T[i] = s_best_T[i];  // cycling: back to the best linearisation point
                    if (a.out_error) a.out_error[pose_i] = (float)s_best_y;
                    s_state = 1;
                    s_lm = 2;
                }
            }
        }
        __syncthreads();
        if (s_lm == 2) break;
        // ---- Levenberg-Marquardt step (fast_gicp LsqRegistration::step_lm): up to 10 damped trials.  A trial solves
        // (H + lambda I) d = -g, evaluates the error of exp(d) T with the correspondences and Mahalanobis matrices of the
        // linearisation FROZEN (FastGICP::compute_error), and is accepted when the gain ratio
        // rho = (y0 - y') / (d . (lambda d - g)) is not negative; a rejected trial multiplies lambda by nu (2, 4, 8 ...).
        for (int trial = 0; trial < 10; ++trial) {
            if (tid == 0) {
                double Hd[36], x[6];
                for (int i = 0; i < 36; ++i) Hd[i] = s_H[i];
                for (int c = 0; c < 6; ++c) Hd[c * 7] += s_lambda;
                if (!solve6(Hd, s_g, x)) {
                    s_state = 1;
                    s_lm = 2;
                } else {
                    double dR[9], dt[3];
                    se3_exp(x, dR, dt);
                    for (int i = 0; i < 3; ++i) {
                        for (int j = 0; j < 3; ++j)
                            s_Tt[i * 3 + j] = dR[i * 3] * s_T[j] + dR[i * 3 + 1] * s_T[3 + j] + dR[i * 3 + 2] * s_T[6 + j];
                        s_Tt[9 + i] = dR[i * 3] * s_T[9] + dR[i * 3 + 1] * s_T[10] + dR[i * 3 + 2] * s_T[11] + dt[i];
                    }
                    double den = 0;
                    for (int c = 0; c < 6; ++c) { s_d[c] = x[c]; den += x[c] * (s_lambda * x[c] - s_g[c]); }
                    s_den = den;
                }
            }
            __syncthreads();
            if (s_lm == 2) break;
            float Rt[9], tt[3];
#pragma unroll
            for (int i = 0; i < 9; ++i) Rt[i] = (float)s_Tt[i];
#pragma unroll
            for (int i = 0; i < 3; ++i) tt[i] = (float)s_Tt[9 + i];
            float yt = 0.f;
            for (int i = tid; i < n; i += kGicpThreads) {
                const int ref = refs[i];
                if (ref < 0) continue;
                const float ax = pts[3 * i], ay = pts[3 * i + 1], az = pts[3 * i + 2];
                const float4 b = __ldg(a.tgt_sorted + ref);
                const float rx = b.x - (Rt[0] * ax + Rt[1] * ay + Rt[2] * az + tt[0]);
                const float ry = b.y - (Rt[3] * ax + Rt[4] * ay + Rt[5] * az + tt[1]);
                const float rz = b.z - (Rt[6] * ax + Rt[7] * ay + Rt[8] * az + tt[2]);
                const float nax = nrm[3 * i], nay = nrm[3 * i + 1], naz = nrm[3 * i + 2];
                const float na[3] = {R[0] * nax + R[1] * nay + R[2] * naz, R[3] * nax + R[4] * nay + R[5] * naz,
                                     R[6] * nax + R[7] * nay + R[8] * naz};  // R of the linearisation: M is frozen
                const float nb[3] = {a.tgt_normals[3 * (size_t)ref], a.tgt_normals[3 * (size_t)ref + 1], a.tgt_normals[3 * (size_t)ref + 2]};
                const Sym3 M = inverse_sum_cov(na, nb, w);
                yt += rx * (M.xx * rx + M.xy * ry + M.xz * rz) + ry * (M.xy * rx + M.yy * ry + M.yz * rz) +
                      rz * (M.xz * rx + M.yz * ry + M.zz * rz);
            }
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) yt += __shfl_xor_sync(0xffffffffu, yt, o);
            if (lane == 0) s_part[warp][0] = yt;
            __syncthreads();
            if (tid == 0) {
                double yi = 0;
                for (int wv = 0; wv < kGicpWarps; ++wv) yi += (double)s_part[wv][0];
                double dR[9], dt[3];
                se3_exp(s_d, dR, dt);
                double mr = 0, mt = 0;
                for (int i = 0; i < 9; ++i) mr = fmax(mr, fabs(dR[i] - ((i % 4 == 0) ? 1.0 : 0.0)));
                for (int i = 0; i < 3; ++i) mt = fmax(mt, fabs(dt[i]));
                const bool small = mr < (double)a.rotation_epsilon && mt < (double)a.translation_epsilon;
                const double rho = (s_y0 - yi) / s_den;
                if (rho < 0.0) {
                    if (small) { s_state = 1; s_lm = 1; }        // rejected, but the step is below the epsilons: converged
                    else { s_lambda *= s_nu; s_nu *= 2.0; }
                } else {
                    for (int i = 0; i < 12; ++i) s_T[i] = s_Tt[i];
                    const double q = 2.0 * rho - 1.0;
                    s_lambda *= fmax(1.0 / 3.0, 1.0 - q * q * q);
                    s_lm = 1;
                    if (small) s_state = 1;
                }
            }
            __syncthreads();
            const int lm = s_lm;
            __syncthreads();  // everyone has read s_lm before thread 0 may overwrite it in the next trial
            if (lm) break;
        }
        if (tid == 0 && s_lm == 0) s_state = 1;  // ten rejected trials: "lm not converged", the pose keeps its last accepted T
        __syncthreads();
    }

    // ---- outputs: transform (metres) and the adjusted pose (x100, renderer.cu:1412-1429) ---------------------------
    if (tid == 0) {
        if (a.out_transforms)
            for (int i = 0; i < 12; ++i) a.out_transforms[12 * (size_t)pose_i + i] = (float)s_T[i];
        if (a.out_iterations) a.out_iterations[pose_i] = s_iters;
        if (n == 0 && a.out_error) a.out_error[pose_i] = 0.f;
        // pose' = T * to_eigen(pose, 100), re-scaled by 100: float arithmetic like the reference's Matrix4f product
        const float* P = reinterpret_cast<const float*>(a.poses_in + pose_i);
        float* O = reinterpret_cast<float*>(a.poses_out + pose_i);
        float A[12];
        for (int i = 0; i < 12; ++i) A[i] = P[i] / 100.f;  // rows a, b, c of the mat4x4 (model.h:108-127)
        float Tf[12];
        for (int i = 0; i < 12; ++i) Tf[i] = (float)s_T[i];
        for (int r2 = 0; r2 < 3; ++r2) {
            for (int c = 0; c < 4; ++c) {
                float v = Tf[r2 * 3] * A[c] + Tf[r2 * 3 + 1] * A[4 + c] + Tf[r2 * 3 + 2] * A[8 + c];
                if (c == 3) v += Tf[9 + r2];
                O[r2 * 4 + c] = (float)((double)v * 100.0);
            }
        }
        O[12] = P[12]; O[13] = P[13]; O[14] = P[14]; O[15] = P[15];
    }
}

}  // namespace

cudaError_t launch_target_normals(const float* d_xyz_aos, const int* d_knn_index, int n, int k, float* d_normals,
                                  cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    target_normals_kernel<<<(n + 127) / 128, 128, 0, stream>>>(d_xyz_aos, d_knn_index, n, k, d_normals);
    return cudaGetLastError();
}

cudaError_t launch_target_normals_grid(const ObsGridDev& grid, float cell, const float4* d_pts_sorted, int n, float* d_normals,
                                       cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    target_normals_grid_kernel<10><<<(n + 127) / 128, 128, 0, stream>>>(grid, cell, d_pts_sorted, n, d_normals);
    return cudaGetLastError();
}

cudaError_t launch_pose_count_max(const int* d_pose_begin, int rows_per_pose, int n_poses, int* d_out, cudaStream_t stream) {
    if (n_poses <= 0) return cudaSuccess;
    pose_count_max_kernel<<<(n_poses + 255) / 256, 256, 0, stream>>>(d_pose_begin, rows_per_pose, n_poses, d_out);
    return cudaGetLastError();
}

size_t gicp_smem_bytes(int smem_points) { return (size_t)smem_points * 7 * sizeof(float); }  // xyz, normal, correspondence

cudaError_t launch_gicp(const GicpArgs& a, int n_poses, cudaStream_t stream) {
    if (n_poses <= 0) return cudaSuccess;
    const size_t smem = gicp_smem_bytes(a.smem_points);
    // per function and per device, not per context: always the device's opt-in limit (see launch_pose_kernel)
    int dev = 0, limit = 0;
    cudaError_t err = cudaGetDevice(&dev);
    if (err == cudaSuccess) err = cudaDeviceGetAttribute(&limit, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (err != cudaSuccess) return err;
    cudaFuncAttributes fa;
    if ((err = cudaFuncGetAttributes(&fa, gicp_kernel)) != cudaSuccess) return err;
    limit -= (int)fa.sharedSizeBytes;  // the opt-in limit covers static + dynamic shared memory
    if (smem > (size_t)limit) return cudaErrorInvalidValue;
    err = cudaFuncSetAttribute(gicp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, limit);
    if (err != cudaSuccess) return err;
    gicp_kernel<<<n_poses, kGicpThreads, smem, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace perch
