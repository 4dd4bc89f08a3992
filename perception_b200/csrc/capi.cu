// C-ABI of libperch_b200.so (include/perch_b200.h): host orchestration of the sm_100a kernels.
// Mirrors the stage machine of cuda_renderer::render_cuda_multi_unified (renderer.cu:1431-1934) and
// depth2cloud_global (renderer.cu:1936-2070), with all scene/model state resident in HBM between calls.
#include <cuda_runtime.h>

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <string>
#include <vector>

#include "context.hpp"
#include "kernels.cuh"
#include "clusters.hpp"
#include "capi_common.hpp"
#include "perch_b200.h"

namespace {
thread_local std::string g_err;
}

namespace perch_detail {
int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
}  // namespace perch_detail


namespace perch_detail {
int ensure_scratch(perch_ctx* c, size_t bytes) {
    if (bytes <= c->scratch_bytes) return PERCH_OK;
    if (c->d_scratch) cudaFree(c->d_scratch);
    c->d_scratch = nullptr;
    c->scratch_bytes = 0;
    PERCH_CU(cudaMalloc(&c->d_scratch, bytes));
    c->scratch_bytes = bytes;
    size_t free_b = 0, total_b = 0;  // high-water mark moves when the scratch grows, not on every call (utils.cuh:18-23)
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess)
        c->stats.peak_memory_usage = std::max(c->stats.peak_memory_usage, (double)(total_b - free_b) / 1024.0 / 1024.0);
    return PERCH_OK;
}

int ensure_pinned(perch_ctx* c, size_t bytes) {
    if (bytes <= c->pinned_bytes) return PERCH_OK;
    if (c->h_pinned) cudaFreeHost(c->h_pinned);
    c->h_pinned = nullptr;
    c->pinned_bytes = 0;
    PERCH_CU(cudaMallocHost(&c->h_pinned, bytes));
    c->pinned_bytes = bytes;
    return PERCH_OK;
}
}  // namespace perch_detail

namespace {
using perch_detail::fail;

#define CU(call)                                                                                           \
    do {                                                                                                   \
        cudaError_t e__ = (call);                                                                          \
        if (e__ != cudaSuccess)                                                                            \
            return fail(PERCH_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, \
                        __LINE__);                                                                         \
    } while (0)

using perch_detail::DeviceGuard;

template <class T>
int upload(T** dst, const T* src, size_t n, cudaStream_t s) {
    if (*dst) {
        cudaFree(*dst);
        *dst = nullptr;
    }
    if (n == 0) return PERCH_OK;
    CU(cudaMalloc(dst, n * sizeof(T)));
    if (src) CU(cudaMemcpyAsync(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice, s));
    return PERCH_OK;
}

using perch_detail::Scratch;

void note_memory(perch_ctx* c) {
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
        double used = (double)(total_b - free_b) / 1024.0 / 1024.0;  // MiB like utils.cuh:18-23
        c->stats.peak_memory_usage = std::max(c->stats.peak_memory_usage, used);
    }
}

using perch_detail::ensure_scratch;

int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

// ---- observed-cloud grid ------------------------------------------------------------------------------------
// Uniform grid over a host cloud (xyz: n x 3; ids: the reference index stored with point i, or null for i itself, must
// ascend; labels: per point or null) with cells at least `h` wide (widened until it has at most max_cells cells): points
// sorted by cell, x fastest, ascending reference index inside a cell.  *avg_per_cell = points per non-empty cell.
int build_point_grid_from(perch_ctx* c, const float* xyz, int n, const int* ids, const int* labels, float h, double max_cells,
                          float4** d_pts, int** d_label, uint32_t** d_cell_start, perch::ObsGridDev* out, double* avg_per_cell) {
    using namespace perch;
    cudaFree(*d_pts); *d_pts = nullptr;
    cudaFree(*d_label); *d_label = nullptr;
    cudaFree(*d_cell_start); *d_cell_start = nullptr;
    *out = ObsGridDev{};
    out->n_obs = n;
    if (avg_per_cell) *avg_per_cell = 0.0;
    if (n == 0) return PERCH_OK;
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = 0; i < n; ++i)
        for (int k = 0; k < 3; ++k) {
            float v = xyz[3 * (size_t)i + k];
            if (!std::isfinite(v)) return fail(PERCH_ERR_INVALID, "cloud contains a non-finite coordinate");
            mn[k] = std::min(mn[k], v);
            mx[k] = std::max(mx[k], v);
        }
    int dim[3];
    float org[3], inv_h;
    for (;;) {
        inv_h = 1.0f / h;
        double cells = 1;
        for (int k = 0; k < 3; ++k) {
            org[k] = mn[k] - h;
            dim[k] = (int)std::floor((mx[k] - org[k]) * inv_h) + 2;
            cells *= dim[k];
        }
        if (cells <= max_cells) break;
        h *= 1.26f;  // ~2x fewer cells per step; larger cells only add candidates, never lose one
    }
    const size_t ncells = (size_t)dim[0] * dim[1] * dim[2];
    std::vector<uint32_t> cell_of(n), start(ncells + 1, 0);
    for (int i = 0; i < n; ++i) {
        int cc[3];
        for (int k = 0; k < 3; ++k) {
            float d = xyz[3 * (size_t)i + k] - org[k];
            float f = std::floor(d * inv_h);
            cc[k] = std::min(std::max((int)f, 0), dim[k] - 1);
        }
        cell_of[i] = (uint32_t)(((size_t)cc[2] * dim[1] + cc[1]) * dim[0] + cc[0]);
        start[cell_of[i] + 1]++;
    }
    size_t nonempty = 0;
    for (size_t k = 0; k < ncells; ++k) {
        nonempty += start[k + 1] != 0;
        start[k + 1] += start[k];
    }
    if (avg_per_cell) *avg_per_cell = (double)n / (double)std::max<size_t>(nonempty, 1);
    std::vector<uint32_t> cursor(start.begin(), start.end() - 1);
    std::vector<float4> pts(n);
    std::vector<int> lab(labels ? n : 0);
    for (int i = 0; i < n; ++i) {  // ascending i => ascending reference index inside every cell
        uint32_t slot = cursor[cell_of[i]]++;
        float4 p;
        p.x = xyz[3 * (size_t)i];
        p.y = xyz[3 * (size_t)i + 1];
        p.z = xyz[3 * (size_t)i + 2];
        int ref = ids ? ids[i] : i;
        memcpy(&p.w, &ref, 4);
        pts[slot] = p;
        if (labels) lab[slot] = labels[i];
    }
    CU(cudaMalloc(d_pts, n * sizeof(float4)));
    CU(cudaMemcpyAsync(*d_pts, pts.data(), n * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    if (labels) {
        CU(cudaMalloc(d_label, n * sizeof(int)));
        CU(cudaMemcpyAsync(*d_label, lab.data(), n * sizeof(int), cudaMemcpyHostToDevice, c->stream));
    }
    CU(cudaMalloc(d_cell_start, (ncells + 1) * sizeof(uint32_t)));
    CU(cudaMemcpyAsync(*d_cell_start, start.data(), (ncells + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));  // host vectors go out of scope
    out->pts = *d_pts;
    out->label = *d_label;
    out->cell_start = *d_cell_start;
    out->ox = org[0]; out->oy = org[1]; out->oz = org[2];
    out->inv_h = inv_h;
    out->dx = dim[0]; out->dy = dim[1]; out->dz = dim[2];
    return PERCH_OK;
}

int ensure_obs_tmp(perch_ctx* c, size_t n_ints) {
    if (n_ints <= c->obs_tmp_n) return PERCH_OK;
    if (c->d_obs_tmp) cudaFree(c->d_obs_tmp);
    c->d_obs_tmp = nullptr;
    c->obs_tmp_n = 0;
    CU(cudaMalloc(&c->d_obs_tmp, n_ints * sizeof(int)));
    c->obs_tmp_n = n_ints;
    return PERCH_OK;
}

// Uniform grid over the resident observed cloud (d_obs_xyz4, label-sorted) with cells at least `h` wide (widened until it
// has at most max_cells cells), built ON THE DEVICE (obs_grid.cu: count -> scan -> scatter); only the number of non-empty
// cells comes back (4 bytes), for *avg_per_cell = points per non-empty cell.
int build_point_grid(perch_ctx* c, float h, double max_cells, float4** d_pts, int** d_label, uint32_t** d_cell_start,
                     perch::ObsGridDev* out, double* avg_per_cell) {
    using namespace perch;
    cudaFree(*d_pts); *d_pts = nullptr;
    cudaFree(*d_label); *d_label = nullptr;
    cudaFree(*d_cell_start); *d_cell_start = nullptr;
    *out = ObsGridDev{};
    const int n = c->n_obs;
    out->n_obs = n;
    if (avg_per_cell) *avg_per_cell = 0.0;
    if (n == 0) return PERCH_OK;
    if (c->obs_nonfinite) return fail(PERCH_ERR_INVALID, "cloud contains a non-finite coordinate");
    int dim[3];
    float org[3], inv_h;
    for (;;) {
        inv_h = 1.0f / h;
        double cells = 1;
        for (int k = 0; k < 3; ++k) {
            org[k] = c->obs_min[k] - h;
            dim[k] = (int)std::floor((c->obs_max[k] - org[k]) * inv_h) + 2;
            cells *= dim[k];
        }
        if (cells <= max_cells) break;
        h *= 1.26f;  // ~2x fewer cells per step; larger cells only add candidates, never lose one
    }
    const size_t ncells = (size_t)dim[0] * dim[1] * dim[2];
    ObsGridDev g;
    g.ox = org[0]; g.oy = org[1]; g.oz = org[2];
    g.inv_h = inv_h;
    g.dx = dim[0]; g.dy = dim[1]; g.dz = dim[2];
    g.n_obs = n;
    int rc;
    if ((rc = ensure_obs_tmp(c, ncells + 64))) return rc;
    CU(cudaMalloc(d_pts, (size_t)n * sizeof(float4)));
    if (c->obs_has_label) CU(cudaMalloc(d_label, (size_t)n * sizeof(int)));
    CU(cudaMalloc(d_cell_start, (ncells + 1) * sizeof(uint32_t)));
    int* d_nonempty = c->d_obs_tmp + ncells;
    CU(perch::launch_obs_grid_build(c->d_obs_xyz4, c->obs_has_label ? c->d_obs_label_sorted : nullptr, n, g, *d_cell_start,
                                    c->d_obs_tmp, *d_pts, *d_label, avg_per_cell ? d_nonempty : nullptr, c->stream));
    if (avg_per_cell) {
        int nonempty = 0;
        CU(cudaMemcpyAsync(&nonempty, d_nonempty, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        *avg_per_cell = (double)n / (double)std::max(nonempty, 1);
    }
    g.pts = *d_pts;
    g.label = *d_label;
    g.cell_start = *d_cell_start;
    *out = g;
    return PERCH_OK;
}

// The grid of the cost stage: cells just wider than sensor_resolution, so that the 27 cells around a query hold every
// observed point within sensor_resolution of it.
int build_grid(perch_ctx* c, float res) {
    c->grid_res = res;
    c->gicp_grid_valid = false;
    return build_point_grid(c, (res > 0.f ? res : 1e-3f) * 1.01f, 32.0 * 1024 * 1024, &c->d_grid_pts, &c->d_grid_label,
                            &c->d_cell_start, &c->grid, &c->grid_avg_per_cell);
}

// The grid GICP searches: its nearest-neighbour queries have no distance gate and always complete rings 0 and 1
// (27 cells), so on a dense cloud (a full-resolution scene: ~40 points per 1 cm cell) the cost grid makes every query
// scan hundreds of points.  When the cost grid holds more than 8 points per non-empty cell a second grid is built with
// cells sized for ~6 points each (points per cell of a surface scale with h^2; measured optimum of the warp-wide warm
// search on a 233k-point cloud: 1 / 2 / 3 / 5 / 8 / 12 points per cell -> 12.9 / 8.8 / 7.5 / 6.6 / 6.5 / 7.2 ms); otherwise the
// cost grid is shared.
int prepare_gicp_grid(perch_ctx* c) {
    if (c->gicp_grid_valid) return PERCH_OK;
    const float h_cost = c->grid.inv_h > 0.f ? 1.0f / c->grid.inv_h : 1.f;
    const double dense = env_int("PERCH_GICP_GRID_DENSITY", 8);
    if (c->n_obs == 0 || c->grid_avg_per_cell <= dense) {
        c->gicp_grid = c->grid;
        c->gicp_grid_h = h_cost;
        c->gicp_grid_avg = c->grid_avg_per_cell;
    } else {
        const double target = std::max(1, env_int("PERCH_GICP_GRID_TARGET_X10", 60)) / 10.0;
        const float h = h_cost * (float)std::sqrt(target / c->grid_avg_per_cell);
        int rc = build_point_grid(c, h, 64.0 * 1024 * 1024, &c->d_ggrid_pts, &c->d_ggrid_label, &c->d_ggrid_cell_start,
                                  &c->gicp_grid, &c->gicp_grid_avg);
        if (rc) return rc;
        c->gicp_grid_h = 1.0f / c->gicp_grid.inv_h;
    }
    // pyramid: cells doubling per level until a level has at most two cells per axis (gicp.cu: pyramid_search_warp)
    c->gicp_n_levels = 0;
    // Built for sparse targets only (the cost grid doubles as level 0): there a far query opens hundreds of near-empty rows
    // on the finest grid (C1's label cloud: 47 -> 25 ms per 2 000 poses x 20).  On a dense cloud with its own fine grid the
    // coarser levels' cells hold too many points to pay (233k points: 6.5 ms without, 6.8-7.1 ms with): PERCH_GICP_PYRAMID=2
    // forces it on, 0 off.
    const int pyramid = env_int("PERCH_GICP_PYRAMID", 1);
    if (c->n_obs > 0 && (pyramid == 2 || (pyramid == 1 && c->gicp_grid.pts == c->grid.pts))) {
        std::vector<perch::ObsGridDev> lv{c->gicp_grid};
        float h = c->gicp_grid_h;
        while ((int)lv.size() < perch_ctx::kGicpMaxLevels && std::max(lv.back().dx, std::max(lv.back().dy, lv.back().dz)) > 3) {
            h *= 2.f;
            const int l = (int)lv.size();
            perch::ObsGridDev g;
            int rc = build_point_grid(c, h, 64.0 * 1024 * 1024, &c->d_lvl_pts[l], &c->d_lvl_label[l], &c->d_lvl_cell_start[l], &g, nullptr);
            if (rc) return rc;
            h = 1.0f / g.inv_h;
            lv.push_back(g);
        }
        cudaFree(c->d_gicp_levels);
        c->d_gicp_levels = nullptr;
        CU(cudaMalloc(&c->d_gicp_levels, lv.size() * sizeof(perch::ObsGridDev)));
        CU(cudaMemcpyAsync(c->d_gicp_levels, lv.data(), lv.size() * sizeof(perch::ObsGridDev), cudaMemcpyHostToDevice, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        c->gicp_n_levels = (int)lv.size();
    }
    c->gicp_grid_valid = true;
    return PERCH_OK;
}

struct ResolvedParams {
    int stride, cost_type, calc_observed, single_image, do_icp;
    float res, color_thr, occ_thr;
};

int check_ready(perch_ctx* c, bool need_obs) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    if (!c->have_cam) return fail(PERCH_ERR_INVALID, "perch_set_camera has not been called");
    if (!c->have_models) return fail(PERCH_ERR_INVALID, "perch_set_models has not been called");
    if (!c->have_scene) return fail(PERCH_ERR_INVALID, "perch_set_scene has not been called");
    if (need_obs && !c->have_obs) return fail(PERCH_ERR_INVALID, "perch_set_observed has not been called");
    return PERCH_OK;
}

int fill_pose_args(perch_ctx* c, const perch_params* p, perch::PoseArgs& a) {
    a.models = c->models;
    a.cam = c->cam;
    a.scene = c->scene;
    a.stride = p->stride;
    a.calc_observed = p->calculate_observed_cost;
    a.res2 = p->sensor_resolution * p->sensor_resolution;  // renderer.cu:1877
    a.color_thr = p->color_distance_threshold;
    a.occlusion_threshold = p->occlusion_threshold;
    return PERCH_OK;
}

// 6-DoF calls restrict the nearest-neighbour search to the pose's label and read the observed labels on the device: refuse
// them when perch_set_observed was given none (the kernels would dereference a null label array).
int fail_unlabelled() {
    return fail(PERCH_ERR_INVALID, "pose_segmentation_label given but perch_set_observed was called without observed labels");
}

// `a.pose_label` must already say whether the call is 6-DoF (null = 3-DoF); everything else of `a` is filled here.
int prepare_cost_args(perch_ctx* c, const perch_params* p, perch::PoseArgs& a, bool& color) {
    if (p->stride <= 0 || c->cam.width % p->stride != 0)
        return fail(PERCH_ERR_INVALID, "stride %d must divide the image width %d (compute_point_clouds.cuh:271)",
                    p->stride, c->cam.width);
    if (p->cost_type < 0 || p->cost_type > 2) return fail(PERCH_ERR_INVALID, "cost_type %d not in {0,1,2}", p->cost_type);
    if (a.pose_label && c->n_obs > 0 && !c->obs_has_label) return fail_unlabelled();
    if (c->grid_res != p->sensor_resolution) {
        int rc = build_grid(c, p->sensor_resolution);
        if (rc) return rc;
    }
    fill_pose_args(c, p, a);
    a.grid = c->grid;
    color = p->cost_type == 1;
    a.tri_lab = c->d_tri_lab;
    a.obs_lab = c->d_obs_lab;
    a.mask_words = (c->n_obs + 31) / 32;
    a.obs_sorted = c->d_obs_xyz4;
    a.label_range = c->d_label_range;
    a.n_label_ranges = c->n_label_ranges;
    const int Lw = c->cam.width / p->stride;
    int cap = env_int("PERCH_COST_TILE_CAP", 2048);
    a.tile_cap = std::max(cap, Lw);
    size_t smem = perch::pose_kernel_smem_bytes(a, false, color);
    if (smem > (size_t)c->max_smem_optin)
        return fail(PERCH_ERR_CAPACITY, "observed cloud of %d points needs %zu B of shared memory per CTA (limit %d)",
                    c->n_obs, smem, c->max_smem_optin);
    a.err_flag = c->d_err_flag;
    return PERCH_OK;
}

int ensure_gicp_buffer(perch_ctx* c, size_t bytes) {
    if (bytes <= c->gicp_bytes) return PERCH_OK;
    if (c->d_gicp) cudaFree(c->d_gicp);
    c->d_gicp = nullptr;
    c->gicp_bytes = 0;
    CU(cudaMalloc(&c->d_gicp, bytes));
    c->gicp_bytes = bytes;
    return PERCH_OK;
}

// Normals of the observed cloud (target covariances of GICP), cached until the cloud or k changes.
int prepare_target_normals(perch_ctx* c, int k) {
    const int n = c->n_obs;
    if (n == 0 || c->tgt_normals_k == k) return PERCH_OK;
    int rc;
    // k = 10 (the reference's value): ring search through GICP's own grid, on the device (gicp.cu).  Other k, or
    // PERCH_GICP_NORMALS_BRUTE=1 (A/B tests): the N x N k-NN kernel on a host-transposed copy of the cloud.
    if (k == 10 && c->d_obs_xyz4 && !env_int("PERCH_GICP_NORMALS_BRUTE", 0)) {
        if ((rc = prepare_gicp_grid(c))) return rc;
        if ((rc = upload<float>(&c->d_tgt_normals, nullptr, 3 * (size_t)n, c->stream))) return rc;
        CU(perch::launch_target_normals_grid(c->gicp_grid, c->gicp_grid_h, c->d_obs_xyz4, n, c->d_tgt_normals, c->stream));
        c->tgt_normals_k = k;
        return PERCH_OK;
    }
    if ((rc = perch_detail::ensure_obs_host(c))) return rc;
    if ((rc = upload(&c->d_obs_xyz_aos, c->h_obs_xyz.data(), 3 * (size_t)n, c->stream))) return rc;
    if ((rc = upload<float>(&c->d_tgt_normals, nullptr, 3 * (size_t)n, c->stream))) return rc;
    const int kk = std::min(k, n);
    std::vector<float> soa(3 * (size_t)n);
    for (int i = 0; i < n; ++i)
        for (int a = 0; a < 3; ++a) soa[(size_t)a * n + i] = c->h_obs_xyz[3 * (size_t)i + a];
    if ((rc = ensure_gicp_buffer(c, Scratch::align(3 * (size_t)n * 4) + 2 * Scratch::align((size_t)kk * n * 4) + 1024))) return rc;
    char* base = static_cast<char*>(c->d_gicp);
    float* d_soa = reinterpret_cast<float*>(base);
    float* d_dist = reinterpret_cast<float*>(base + Scratch::align(3 * (size_t)n * 4));
    int* d_idx = reinterpret_cast<int*>(base + Scratch::align(3 * (size_t)n * 4) + Scratch::align((size_t)kk * n * 4));
    CU(cudaMemcpyAsync(d_soa, soa.data(), soa.size() * 4, cudaMemcpyHostToDevice, c->stream));
    CU(perch::launch_knn(d_soa, n, d_soa, n, 3, kk, d_dist, d_idx, c->sm_count, c->stream));
    CU(perch::launch_target_normals(c->d_obs_xyz_aos, d_idx, n, kk, c->d_tgt_normals, c->stream));
    CU(cudaStreamSynchronize(c->stream));  // `soa` goes out of scope
    c->tgt_normals_k = k;
    return PERCH_OK;
}

// The do_icp block (renderer.cu:1688-1817): lattice render of every pose -> compacted clouds -> batched GICP against
// the observed cloud -> adjusted poses in d_poses_out.  `base` carries the device pose arrays.  Adds to *launches.
int run_gicp(perch_ctx* c, const perch_params* p, const perch::PoseArgs& base, int n_poses, float* d_poses_out,
             float* d_transforms, int* d_iterations, float* d_error, int* launches) {
    if (n_poses <= 0) return PERCH_OK;
    if (base.pose_label && c->n_obs > 0 && !c->obs_has_label) return fail_unlabelled();
    const int W = c->cam.width, H = c->cam.height, s = p->stride;
    if (s <= 0 || W % s != 0)
        return fail(PERCH_ERR_INVALID, "stride %d must divide the image width %d (compute_point_clouds.cuh:271)", s, W);
    int rc;
    if (c->grid_res != p->sensor_resolution && (rc = build_grid(c, p->sensor_resolution))) return rc;
    const int k = p->icp_k_correspondences > 0 ? p->icp_k_correspondences : 10;           // renderer.cu:1697
    if (k > 32) return fail(PERCH_ERR_CAPACITY, "icp_k_correspondences %d exceeds 32", k);
    if ((rc = prepare_target_normals(c, k))) return rc;
    const int Lw = W / s, Lh = (H + s - 1) / s;
    const size_t P = (size_t)n_poses, lat = (size_t)Lw * Lh;
    // the lattice depth images and row counts live in the context scratch AFTER the caller's slices: use a private buffer
    const size_t img_bytes = Scratch::align(P * lat * 4), rows_bytes = Scratch::align((P * Lh + 1) * 4 + 64);
    if ((rc = ensure_gicp_buffer(c, img_bytes + rows_bytes))) return rc;
    int32_t* l_depth = reinterpret_cast<int32_t*>(c->d_gicp);
    int* d_rows = reinterpret_cast<int*>(static_cast<char*>(c->d_gicp) + img_bytes);
    int* d_max = d_rows + P * Lh + 1;
    perch::PoseArgs a = base;
    a.stride = s;
    a.img_depth = l_depth; a.img_r = nullptr; a.img_g = nullptr; a.img_b = nullptr;
    a.single_keys = nullptr;
    a.tile_cap = std::max(env_int("PERCH_IMAGE_TILE_CAP", 6144), Lw);
    CU(perch::launch_raster_images(a, c->stream));
    perch::CloudArgs ca;
    ca.depth = l_depth;
    ca.pose_label = base.pose_label;
    ca.img_w = Lw; ca.img_h = Lh; ca.full_h = H; ca.num_images = n_poses; ca.stride = 1; ca.coord_scale = s;
    ca.cx = c->cam.cx; ca.cy = c->cam.cy; ca.fx = c->cam.fx; ca.fy = c->cam.fy; ca.depth_factor = c->cam.depth_factor;
    ca.row_counts = d_rows;
    CU(perch::launch_cloud_count(ca, c->stream));
    CU(perch::launch_cloud_scan(ca, c->stream));
    CU(cudaMemsetAsync(d_max, 0, 4, c->stream));
    CU(perch::launch_pose_count_max(d_rows, Lh, n_poses, d_max, c->stream));
    int total = 0, max_n = 0;
    CU(cudaMemcpyAsync(&total, d_rows + P * Lh, 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(&max_n, d_max, 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    *launches += 4;
    // grow the private buffer for cloud + normals without losing the images/rows: they are re-created if it moves
    const size_t cloud_bytes = Scratch::align((size_t)std::max(total, 1) * 12);
    const size_t ref_bytes = Scratch::align((size_t)std::max(total, 1) * 4);
    const size_t need = img_bytes + rows_bytes + 2 * cloud_bytes + ref_bytes;
    if (need > c->gicp_bytes) {
        if ((rc = ensure_gicp_buffer(c, need))) return rc;
        l_depth = reinterpret_cast<int32_t*>(c->d_gicp);
        d_rows = reinterpret_cast<int*>(static_cast<char*>(c->d_gicp) + img_bytes);
        a.img_depth = l_depth;
        ca.depth = l_depth; ca.row_counts = d_rows;
        CU(perch::launch_raster_images(a, c->stream));
        CU(perch::launch_cloud_count(ca, c->stream));
        CU(perch::launch_cloud_scan(ca, c->stream));
        *launches += 3;
    }
    float* d_cloud = reinterpret_cast<float*>(static_cast<char*>(c->d_gicp) + img_bytes + rows_bytes);
    float* d_src_normals = reinterpret_cast<float*>(static_cast<char*>(c->d_gicp) + img_bytes + rows_bytes + cloud_bytes);
    ca.cloud_aos = d_cloud;
    ca.capacity = total;
    if (total > 0) {
        CU(perch::launch_cloud_write(ca, c->stream));
        ++*launches;
    }
    perch::GicpArgs ga;
    ga.cloud_aos = d_cloud;
    ga.pose_begin = d_rows;
    ga.rows_per_pose = Lh;
    ga.src_normals = d_src_normals;
    ga.src_ref = reinterpret_cast<int*>(static_cast<char*>(c->d_gicp) + img_bytes + rows_bytes + 2 * cloud_bytes);
    if ((rc = prepare_gicp_grid(c))) return rc;
    ga.grid = c->gicp_grid;
    ga.grid_h = c->gicp_grid_h;
    ga.levels = c->gicp_n_levels > 1 ? c->d_gicp_levels : nullptr;
    ga.n_levels = c->gicp_n_levels;
    ga.level_k = (float)(env_int("PERCH_GICP_LEVEL_KAPPA_X10", 40) / 10.0 * c->gicp_grid_h * c->gicp_grid_h / std::max(c->gicp_grid_avg, 1.0));
    ga.coarse = c->grid;
    ga.coarse_h = c->grid.inv_h > 0.f ? 1.0f / c->grid.inv_h : 1.f;
    ga.tgt_normals = c->d_tgt_normals;
    ga.tgt_sorted = c->d_obs_xyz4;
    ga.label_range = c->d_label_range;
    ga.n_label_ranges = c->n_label_ranges;
    ga.stream_max = env_int("PERCH_GICP_STREAM_MAX", 512);
    ga.pose_label = base.pose_label;
    ga.poses_in = base.poses;
    ga.poses_out = reinterpret_cast<perch::Mat4*>(d_poses_out);
    ga.out_transforms = d_transforms;
    ga.out_iterations = d_iterations;
    ga.out_error = d_error;
    // Shared memory for the largest pose cloud (rounded up), but never more than lets the register-limited 20 one-warp
    // CTAs share an SM (28 B per point => 384 points): the kernel is latency-bound and lives on resident warps; a pose
    // with more points keeps its arrays in global memory (L1 / L2) instead.  (C1, 626 points per pose on average, 1 216 at
    // most: 34 KB per CTA left 6 warps per SM -- ncu sm__warps_active 6.9 % -- and cost 46 ms per 2 000 poses x 20.)
    const int cap_pts = std::min((int)std::min<size_t>(((size_t)c->max_smem_optin / 2 - 2048) / 28, 4096),
                                 std::max(env_int("PERCH_GICP_SMEM_POINTS", 384), 64));
    ga.smem_points = std::min(std::max((max_n + 63) & ~63, 64), cap_pts);
    ga.k_correspondences = k;
    ga.max_iterations = p->icp_max_iterations > 0 ? p->icp_max_iterations : 150;                     // renderer.cu:1696
    ga.rotation_epsilon = p->icp_rotation_epsilon >= 0.f ? p->icp_rotation_epsilon : 2e-3f;         // renderer.cu:1698
    ga.translation_epsilon = p->icp_transformation_epsilon >= 0.f ? p->icp_transformation_epsilon : 5e-4f;  // :1699
    ga.warm_cells = (float)std::max(env_int("PERCH_GICP_WARM_CELLS", 1000000), 0);
    ga.stall_window = std::max(env_int("PERCH_GICP_STALL_WINDOW", 8), 0);  // 0 = off: iterate to the cap like fast_gicp
    if (c->n_obs == 0) ga.max_iterations = 0;  // no target: poses pass through unchanged
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    CU(cudaEventRecord(e0, c->stream));
    cudaError_t le = perch::launch_gicp(ga, n_poses, c->stream);
    CU(cudaEventRecord(e1, c->stream));
    ++*launches;
    cudaError_t se = cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    CU(le);
    CU(se);
    c->stats.icp_runtime = std::max(c->stats.icp_runtime, ms * 1e-3f);  // seconds, max over calls (renderer.cu:1739)
    c->last_gicp_ms = ms;
    return PERCH_OK;
}

int check_err_flag(perch_ctx* c) {
    int flag = 0;
    CU(cudaMemcpyAsync(&flag, c->d_err_flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (flag) {
        cudaMemsetAsync(c->d_err_flag, 0, sizeof(int), c->stream);
        if (flag & 2) return fail(PERCH_ERR_INVALID, "a state names a model_id outside [0, n_models)");
        if (flag & 4) return fail(PERCH_ERR_INVALID, "a state's segmentation_label_id has no entry in the observed labels");
        return fail(PERCH_ERR_CUDA, "internal error: a fragment fell outside its pose tile (flag %d)", flag);
    }
    return PERCH_OK;
}

}  // namespace

namespace perch_detail {
int prepare_fused_cost(perch_ctx* c, const perch_params* p, perch::PoseArgs& a, bool& color) {
    int rc = check_ready(c, true);
    if (rc) return rc;
    return prepare_cost_args(c, p, a, color);
}
int report_err_flag(int flag) {
    if (!flag) return PERCH_OK;
    if (flag & 2) return fail(PERCH_ERR_INVALID, "a state names a model_id outside [0, n_models)");
    if (flag & 4) return fail(PERCH_ERR_INVALID, "a state's segmentation_label_id has no entry in the observed labels");
    if (flag & 256) return fail(PERCH_ERR_CUDA, "internal error: a shared-memory index of the per-pose kernel was out of range (bounds-check build, flag %d)", flag);
    return fail(PERCH_ERR_CUDA, "internal error: a fragment fell outside its pose tile (flag %d)", flag);
}
}  // namespace perch_detail

extern "C" {

const char* perch_last_error(void) { return g_err.c_str(); }
int perch_abi_version(void) { return PERCH_B200_ABI_VERSION; }

int perch_create(int device, perch_ctx** out) {
    if (!out) return fail(PERCH_ERR_INVALID, "out is null");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(PERCH_ERR_NO_DEVICE, "no CUDA device available (%s); libperch_b200 has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= count) return fail(PERCH_ERR_INVALID, "device %d out of range [0,%d)", device, count);
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail(PERCH_ERR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major,
                    prop.minor);
    DeviceGuard g(device);
    perch_ctx* c = new perch_ctx();
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    c->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
    CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&c->ev0));
    CU(cudaEventCreate(&c->ev1));
    CU(cudaEventCreate(&c->ev_t0));
    CU(cudaEventCreate(&c->ev_t1));
    CU(cudaMalloc(&c->d_err_flag, sizeof(int)));
    CU(cudaMemset(c->d_err_flag, 0, sizeof(int)));
    *out = c;
    return PERCH_OK;
}

void perch_destroy(perch_ctx* c) {
    if (!c) return;
    DeviceGuard g(c->device);
    cudaStreamSynchronize(c->stream);
    for (int l = 0; l < perch_ctx::kGicpMaxLevels; ++l) {
        cudaFree(c->d_lvl_pts[l]); cudaFree(c->d_lvl_label[l]); cudaFree(c->d_lvl_cell_start[l]);
    }
    void* ptrs[] = {c->d_tri_color, c->d_model_low, c->d_model_high, c->d_model_aabb, c->d_model_qrmax, c->d_tri_lab, c->d_cl_recs, c->d_cl_bounds, c->d_q_bounds,
                    c->d_cl_orig, c->d_cl_perm, c->d_cl_low, c->d_cl_high,
                    c->d_src_depth, c->d_src_label, c->d_src_rgb, c->d_grid_pts, c->d_grid_label, c->d_cell_start,
                    c->d_icp_grid_pts, c->d_icp_grid_label, c->d_icp_grid_cells, c->d_cmp_depth, c->d_cmp_label, c->d_full_keys, c->d_full_big, c->d_full_count, c->d_full_acc, c->d_full_rects, c->d_ggrid_pts, c->d_ggrid_label, c->d_ggrid_cell_start, c->d_gicp_levels,
                    c->d_obs_lab, c->d_obs_rgb, c->d_scratch, c->d_err_flag, c->d_flush, c->d_obs_xyz_aos,
                    c->d_tgt_normals, c->d_obs_xyz4, c->d_obs_label_sorted, c->d_obs_tmp, c->d_in_xyz, c->d_in_rgb, c->d_in_label, c->d_label_range, c->d_gicp, c->d_icp_pcd, c->d_icp_normals};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    if (c->h_pinned) cudaFreeHost(c->h_pinned);
    perch_detail::destroy_comm(c);
    perch_detail::destroy_states(c);
    cudaEventDestroy(c->ev0);
    cudaEventDestroy(c->ev1);
    cudaEventDestroy(c->ev_t0);
    cudaEventDestroy(c->ev_t1);
    cudaStreamDestroy(c->stream);
    delete c;
}

int perch_get_stats(perch_ctx* c, perch_gpu_stats* out) {
    if (!c || !out) return fail(PERCH_ERR_INVALID, "null argument");
    *out = c->stats;
    return PERCH_OK;
}

int perch_last_timing(perch_ctx* c, float* ms, int* launches) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    if (ms) *ms = c->last_ms;
    if (launches) *launches = c->last_launches;
    return PERCH_OK;
}

int perch_compute_proj(const float K[9], int width, int height, float near_, float far_, float out16[16]) {
    if (!K || !out16 || width <= 0 || height <= 0) return fail(PERCH_ERR_INVALID, "bad compute_proj argument");
    // renderer.cu:1386-1410.  The reference writes each flipped entry as -(expr); negation is exact, so the values
    // below are the same floats.
    volatile float t;
    float p[16] = {0};
    t = 2 * K[0] / width; p[0] = t;
    t = -2 * K[1] / width; p[1] = -t;
    t = -2 * K[2] / width; t = t + 1; p[2] = -t;
    t = 2 * K[4] / height; p[5] = -t;
    t = 2 * K[5] / height; t = t - 1; p[6] = -t;
    t = -(far_ + near_) / (far_ - near_); p[10] = -t;
    t = -2 * far_ * near_ / (far_ - near_); p[11] = t;
    p[14] = 1;
    memcpy(out16, p, sizeof(p));
    return PERCH_OK;
}

int perch_set_camera(perch_ctx* c, int width, int height, const float proj16[16], float cx, float cy, float fx, float fy,
                     float depth_factor) {
    if (!c || !proj16) return fail(PERCH_ERR_INVALID, "null argument");
    if (width <= 0 || height <= 0 || (int64_t)width * height > (1 << 26))
        return fail(PERCH_ERR_INVALID, "bad image size %dx%d", width, height);
    if (width != c->cam.width || height != c->cam.height) {
        c->have_scene = false;
        // the composed source images are sized for the camera
        DeviceGuard g(c->device);
        if (c->d_cmp_depth) cudaFree(c->d_cmp_depth);
        if (c->d_cmp_label) cudaFree(c->d_cmp_label);
        c->d_cmp_depth = nullptr;
        c->d_cmp_label = nullptr;
        c->composed = false;
        // ... and so are the resident scene images (perch_set_scene / perch_set_input re-allocate for the new size)
        cudaFree(c->d_src_depth); cudaFree(c->d_src_label); cudaFree(c->d_src_rgb);
        c->d_src_depth = nullptr; c->d_src_label = nullptr; c->d_src_rgb = nullptr;
        c->src_px = 0;
        c->scene = perch::SceneDev{};
    }
    c->cam.width = width;
    c->cam.height = height;
    memcpy(&c->cam.proj, proj16, 64);
    c->cam.cx = cx; c->cam.cy = cy; c->cam.fx = fx; c->cam.fy = fy;
    c->cam.depth_factor = depth_factor;
    c->have_cam = true;
    return PERCH_OK;
}

}  // extern "C"

namespace {
// Two independent multiply-rotate lanes over a byte range (8 bytes at a time, the tail zero-padded).
void digest_bytes(const void* data, size_t bytes, uint64_t h[2]) {
    const unsigned char* base = static_cast<const unsigned char*>(data);
    auto mix = [&](uint64_t w) {
        h[0] = (h[0] ^ w) * 0xFF51AFD7ED558CCDull; h[0] = (h[0] << 31) | (h[0] >> 33);
        h[1] = (h[1] + w) * 0xC4CEB9FE1A85EC53ull; h[1] ^= h[1] >> 29;
    };
    size_t i = 0;
    for (; i + 8 <= bytes; i += 8) {
        uint64_t w;
        memcpy(&w, base + i, 8);
        mix(w);
    }
    if (i < bytes) {
        uint64_t w = 0;
        memcpy(&w, base + i, bytes - i);
        mix(w);
    }
    mix((uint64_t)bytes);
}

// Two independent multiply-rotate lanes over the 8-byte words of the triangle array, then the counts.
void models_digest(const perch_triangle* tris, int64_t n_tris, const int* counts, int n_models, uint64_t out[2]) {
    uint64_t h0 = 0x9E3779B97F4A7C15ull ^ (uint64_t)n_tris, h1 = 0xC2B2AE3D27D4EB4Full ^ (uint64_t)n_models;
    auto mix = [&](uint64_t w) {
        h0 = (h0 ^ w) * 0xFF51AFD7ED558CCDull; h0 = (h0 << 31) | (h0 >> 33);
        h1 = (h1 + w) * 0xC4CEB9FE1A85EC53ull; h1 ^= h1 >> 29;
    };
    const size_t words = (size_t)n_tris * sizeof(perch_triangle) / 8;
    const unsigned char* base = reinterpret_cast<const unsigned char*>(tris);
    for (size_t i = 0; i < words; ++i) {
        uint64_t w;
        memcpy(&w, base + 8 * i, 8);
        mix(w);
    }
    for (int m = 0; m < n_models; ++m) mix((uint64_t)(uint32_t)counts[m] | ((uint64_t)m << 32));
    out[0] = h0; out[1] = h1;
}
}  // namespace

extern "C" {

int perch_set_models(perch_ctx* c, const perch_triangle* tris, int64_t n_tris, const int* tris_model_count, int n_models) {
    if (!c || (!tris && n_tris > 0) || !tris_model_count || n_models <= 0 || n_tris < 0)
        return fail(PERCH_ERR_INVALID, "bad set_models argument");
    // candidate entries carry the triangle slot in 26 bits (pose_kernels.cu) and a model without shared edges needs two
    // slots per triangle
    if (n_tris > (1 << 24)) return fail(PERCH_ERR_CAPACITY, "more than 2^24 triangles in one context");
    DeviceGuard g(c->device);
    // The reference passes the meshes with every render call (renderer.cu:1532-1544) and so does the drop-in backend
    // (compat/cuda_renderer_compat.cpp); re-filing them costs ~2 us per triangle on the host.  An input identical to the
    // previous one (128-bit digest of all triangle bytes and counts, ~0.1 ms per MB) leaves the resident copy in place.
    uint64_t digest[2];
    models_digest(tris, n_tris, tris_model_count, n_models, digest);
    if (c->have_models && c->models.n_tris == n_tris && c->models.n_models == n_models && digest[0] == c->models_digest[0] &&
        digest[1] == c->models_digest[1] && !env_int("PERCH_SET_MODELS_ALWAYS", 0))
        return PERCH_OK;
    c->have_models = false;
    std::vector<int> low(n_models), high(n_models);
    int64_t run = 0;
    for (int m = 0; m < n_models; ++m) {
        if (tris_model_count[m] < 0) return fail(PERCH_ERR_INVALID, "negative tris_model_count[%d]", m);
        low[m] = (int)run;
        run += tris_model_count[m];
        high[m] = (int)std::min<int64_t>(run, n_tris);  // the reference also bounds by device_tris_size (image_renderer.cuh:240)
        low[m] = std::min(low[m], high[m]);
    }
    std::vector<uint32_t> color((size_t)std::max<int64_t>(n_tris, 1), 0);
    std::vector<float> aabb((size_t)n_models * 6);
    for (int m = 0; m < n_models; ++m) {
        float* bb = &aabb[(size_t)m * 6];
        bb[0] = bb[1] = bb[2] = INFINITY;
        bb[3] = bb[4] = bb[5] = -INFINITY;
        for (int t = low[m]; t < high[m]; ++t) {
            const float* v = tris[t].v0;  // v0,v1,v2 are contiguous: 9 floats
            for (int k = 0; k < 9; ++k) {
                bb[k % 3] = std::min(bb[k % 3], v[k]);
                bb[3 + k % 3] = std::max(bb[3 + k % 3], v[k]);
            }
            // (uint8_t)(color.v0) etc. (image_renderer.cuh:152-154)
            color[t] = (uint32_t)(uint8_t)tris[t].color[0] | ((uint32_t)(uint8_t)tris[t].color[1] << 8) |
                       ((uint32_t)(uint8_t)tris[t].color[2] << 16);
        }
    }
    int rc;
    if ((rc = upload(&c->d_tri_color, color.data(), color.size(), c->stream))) return rc;
    if ((rc = upload(&c->d_model_low, low.data(), low.size(), c->stream))) return rc;
    if ((rc = upload(&c->d_model_high, high.data(), high.size(), c->stream))) return rc;
    if ((rc = upload(&c->d_model_aabb, aabb.data(), aabb.size(), c->stream))) return rc;
    if ((rc = upload<float>(&c->d_tri_lab, nullptr, 3 * color.size(), c->stream))) return rc;
    CU(perch::launch_rgb2lab_packed(c->d_tri_color, (int)n_tris, c->d_tri_lab, c->stream));
    // clusters for the hierarchical candidate test
    {
        std::vector<float> tri9((size_t)std::max<int64_t>(n_tris, 1) * 9);
        for (int64_t t = 0; t < n_tris; ++t) memcpy(&tri9[(size_t)t * 9], tris[t].v0, 36);
        perch::ClusterSet cs = perch::build_clusters(tri9.data(), low, high);
        if (env_int("PERCH_DEBUG", 0)) {
            double rs = 0; size_t nf = 0;
            for (size_t i = 0; i < cs.bounds.size() / 4; ++i)
                if (std::isfinite(cs.bounds[4 * i + 3])) { rs += cs.bounds[4 * i + 3]; ++nf; }
            fprintf(stderr, "[perch] clusters: %zu of %d triangles, mean radius %.5f\n", cs.bounds.size() / 4,
                    perch::kClusterTris, nf ? rs / nf : 0.0);
        }
        if (cs.bounds.empty()) {
            cs.bounds.assign(4, 0.f);
            cs.qbounds.assign(8, 0.f);
            cs.recs.assign(perch::kClusterFloats, 0.f);
            cs.orig.assign(perch::kClusterTris, 0);
            cs.perm.assign(1, 0xFFFFFFu);  // four invalid slots
        }
        if ((rc = upload(&c->d_cl_recs, cs.recs.data(), cs.recs.size(), c->stream))) return rc;
        if ((rc = upload(&c->d_cl_bounds, reinterpret_cast<const float4*>(cs.bounds.data()), cs.bounds.size() / 4, c->stream))) return rc;
        if ((rc = upload(&c->d_q_bounds, reinterpret_cast<const float4*>(cs.qbounds.data()), cs.qbounds.size() / 4, c->stream))) return rc;
        if ((rc = upload(&c->d_cl_orig, cs.orig.data(), cs.orig.size(), c->stream))) return rc;
        if ((rc = upload(&c->d_cl_perm, cs.perm.data(), cs.perm.size(), c->stream))) return rc;
        if ((rc = upload(&c->d_cl_low, cs.model_low.data(), cs.model_low.size(), c->stream))) return rc;
        if ((rc = upload(&c->d_cl_high, cs.model_high.data(), cs.model_high.size(), c->stream))) return rc;
        CU(cudaStreamSynchronize(c->stream));  // host vectors go out of scope
        c->models.cl_recs = c->d_cl_recs;
        c->models.cl_bounds = c->d_cl_bounds;
        c->models.q_bounds = c->d_q_bounds;
        {
            std::vector<float> qrmax(cs.model_low.size(), 0.f);
            for (size_t m = 0; m < cs.model_low.size(); ++m)
                for (int q = 2 * cs.model_low[m]; q < 2 * cs.model_high[m]; ++q) {
                    const float r = cs.qbounds[4 * (size_t)q + 3];
                    if (std::isfinite(r)) qrmax[m] = std::max(qrmax[m], r);
                }
            if ((rc = upload(&c->d_model_qrmax, qrmax.data(), qrmax.size(), c->stream))) return rc;
            CU(cudaStreamSynchronize(c->stream));
            c->models.model_qrmax = c->d_model_qrmax;
        }
        c->models.cl_orig = c->d_cl_orig;
        c->models.cl_perm = c->d_cl_perm;
        c->models.cl_low = c->d_cl_low;
        c->models.cl_high = c->d_cl_high;
        c->max_model_quads = 0;
        for (size_t m = 0; m < cs.model_low.size(); ++m)
            c->max_model_quads = std::max(c->max_model_quads, perch::kClusterQuads * (cs.model_high[m] - cs.model_low[m]));
    }
    CU(cudaStreamSynchronize(c->stream));
    c->models.tri_color = c->d_tri_color;
    c->models.model_low = c->d_model_low;
    c->models.model_high = c->d_model_high;
    c->models.model_aabb = c->d_model_aabb;
    c->models.n_models = n_models;
    c->models.n_tris = n_tris;
    c->h_model_low = low;
    c->h_model_high = high;
    c->models_digest[0] = digest[0];
    c->models_digest[1] = digest[1];
    c->have_models = true;
    note_memory(c);
    return PERCH_OK;
}

int perch_cluster_stats(const perch_triangle* tris, int64_t n_tris, const int* tris_model_count, int n_models,
                        int64_t out[6]) {
    if ((!tris && n_tris > 0) || !tris_model_count || n_models <= 0 || n_tris < 0 || !out || n_tris > INT_MAX)
        return fail(PERCH_ERR_INVALID, "bad cluster_stats argument");
    std::vector<int> low(n_models), high(n_models);
    int64_t run = 0;
    for (int m = 0; m < n_models; ++m) {
        low[m] = (int)std::min<int64_t>(run, n_tris);
        run += std::max(tris_model_count[m], 0);
        high[m] = (int)std::min<int64_t>(run, n_tris);
    }
    std::vector<float> tri9((size_t)std::max<int64_t>(n_tris, 1) * 9);
    for (int64_t t = 0; t < n_tris; ++t) memcpy(&tri9[(size_t)t * 9], tris[t].v0, 36);
    perch::ClusterSet cs = perch::build_clusters(tri9.data(), low, high);
    const size_t nc = cs.bounds.size() / 4;
    std::vector<int> seen((size_t)std::max<int64_t>(n_tris, 1), 0);
    int64_t paired = 0, single = 0, mismatch = 0, outside = 0;
    for (size_t c = 0; c < nc; ++c) {
        int model = -1;
        for (int m = 0; m < n_models; ++m)
            if ((int)c >= cs.model_low[m] && (int)c < cs.model_high[m]) model = m;
        for (int j = 0; j < perch::kClusterTris; ++j) {
            const uint32_t perm = (cs.perm[c] >> (6 * j)) & 63u;
            if (perm == perch::kPermInvalid) continue;
            const int t = cs.orig[c * perch::kClusterTris + j];
            if (t < 0 || t >= n_tris || model < 0 || t < low[model] || t >= high[model]) { ++outside; continue; }
            ++seen[t];
            const uint32_t other = (cs.perm[c] >> (6 * (j ^ 1))) & 63u;
            (other == perch::kPermInvalid ? single : paired) += 1;
            const float* rec = &cs.recs[c * perch::kClusterFloats + (j >> 1) * perch::kQuadFloats];
            for (int k = 0; k < 3; ++k)
                for (int a = 0; a < 3; ++a) {
                    const float got = rec[((perm >> (2 * k)) & 3u) * 3 + a];
                    if (memcmp(&got, &tri9[(size_t)t * 9 + 3 * k + a], 4) != 0) ++mismatch;
                }
            // every vertex must lie inside its quad's bounding sphere (level 1 of the candidate test) ...
            const float* qb = &cs.qbounds[4 * (c * perch::kClusterQuads + (j >> 1))];
            for (int k = 0; k < 3 && std::isfinite(qb[3]); ++k) {
                double d2 = 0;
                for (int a = 0; a < 3; ++a) d2 += ((double)tri9[(size_t)t * 9 + 3 * k + a] - qb[a]) * ((double)tri9[(size_t)t * 9 + 3 * k + a] - qb[a]);
                if (std::sqrt(d2) > (double)qb[3]) ++outside;
            }
            // ... and inside the cluster's
            const float* b = &cs.bounds[4 * c];
            for (int k = 0; k < 3 && std::isfinite(b[3]); ++k) {
                double d2 = 0;
                for (int a = 0; a < 3; ++a) d2 += ((double)tri9[(size_t)t * 9 + 3 * k + a] - b[a]) * ((double)tri9[(size_t)t * 9 + 3 * k + a] - b[a]);
                if (std::sqrt(d2) > (double)b[3]) ++outside;
            }
        }
    }
    int64_t missing = 0;
    for (int m = 0; m < n_models; ++m)
        for (int t = low[m]; t < high[m]; ++t) missing += seen[t] != 1;
    out[0] = (int64_t)nc; out[1] = paired; out[2] = single; out[3] = mismatch; out[4] = missing; out[5] = outside;
    return PERCH_OK;
}

int perch_set_scene(perch_ctx* c, const int32_t* depth, const uint8_t* red, const uint8_t* green, const uint8_t* blue,
                    const uint8_t* label) {
    if (!c || !depth) return fail(PERCH_ERR_INVALID, "null argument");
    if (!c->have_cam) return fail(PERCH_ERR_INVALID, "perch_set_camera must be called before perch_set_scene");
    DeviceGuard g(c->device);
    const size_t npx = (size_t)c->cam.width * c->cam.height;
    // The buffers are kept across calls of the same image size (the reference's calling pattern sets the scene with every
    // render call: a cudaFree + cudaMalloc per buffer per call cost more than the copies).
    bool grew = false;
    if (c->src_px != npx) {
        cudaFree(c->d_src_depth); cudaFree(c->d_src_label); cudaFree(c->d_src_rgb);
        c->d_src_depth = nullptr; c->d_src_label = nullptr; c->d_src_rgb = nullptr;
        c->src_px = 0;
        CU(cudaMalloc(&c->d_src_depth, npx * sizeof(int32_t)));
        CU(cudaMalloc(&c->d_src_rgb, 3 * npx));
        c->src_px = npx;
        grew = true;
    }
    if (label && !c->d_src_label) { CU(cudaMalloc(&c->d_src_label, npx)); grew = true; }
    if (!label && c->d_src_label) { cudaFree(c->d_src_label); c->d_src_label = nullptr; }  // d_src_label != NULL means "labelled scene"
    CU(cudaMemcpyAsync(c->d_src_depth, depth, npx * sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
    if (label) CU(cudaMemcpyAsync(c->d_src_label, label, npx, cudaMemcpyHostToDevice, c->stream));
    const uint8_t* planes[3] = {red, green, blue};
    for (int k = 0; k < 3; ++k) {
        if (planes[k]) CU(cudaMemcpyAsync(c->d_src_rgb + k * npx, planes[k], npx, cudaMemcpyHostToDevice, c->stream));
        else CU(cudaMemsetAsync(c->d_src_rgb + k * npx, 0, npx, c->stream));
    }
    CU(cudaStreamSynchronize(c->stream));
    c->scene.depth = c->d_src_depth;
    c->scene.label = c->d_src_label;
    c->composed = false;
    c->scene.r = c->d_src_rgb;
    c->scene.g = c->d_src_rgb + npx;
    c->scene.b = c->d_src_rgb + 2 * npx;
    c->have_scene = true;
    if (grew) note_memory(c);
    return PERCH_OK;
}

int perch_set_observed(perch_ctx* c, const float* xyz_aos, const uint8_t* rgb_soa, const int* label, int n_obs) {
    if (!c || n_obs < 0 || (n_obs > 0 && !xyz_aos)) return fail(PERCH_ERR_INVALID, "bad set_observed argument");
    if (n_obs > (1 << 24)) return fail(PERCH_ERR_CAPACITY, "observed cloud larger than 16M points");
    DeviceGuard g(c->device);
    // An observed cloud identical to the resident one (the reference passes it with every render call, renderer.cu:1668,
    // 1881) keeps everything built from it: sorted copy, colours, nearest-neighbour grids, target normals.
    uint64_t digest[2] = {0x9E3779B97F4A7C15ull ^ (uint64_t)n_obs, 0xC2B2AE3D27D4EB4Full ^ (uint64_t)(label != nullptr) ^ ((uint64_t)(rgb_soa != nullptr) << 1)};
    digest_bytes(xyz_aos, 12 * (size_t)n_obs, digest);
    if (rgb_soa) digest_bytes(rgb_soa, 3 * (size_t)n_obs, digest);
    if (label) digest_bytes(label, 4 * (size_t)n_obs, digest);
    if (c->have_obs && c->obs_digest_valid && c->n_obs == n_obs && digest[0] == c->obs_digest[0] && digest[1] == c->obs_digest[1] &&
        !env_int("PERCH_SET_OBSERVED_ALWAYS", 0))
        return PERCH_OK;
    c->obs_digest_valid = false;
    c->n_obs = n_obs;
    c->obs_has_label = label != nullptr;
    c->h_obs_xyz.assign(xyz_aos, xyz_aos + 3 * (size_t)n_obs);
    c->h_obs_label.clear();
    if (label) {
        // renderer.cu:1681: sort the cloud by label (stable here; thrust's radix sort on int keys is stable too), so
        // that reference indices -- and with them nearest-neighbour tie-breaks -- follow the sorted order.
        std::vector<int> order(n_obs);
        std::iota(order.begin(), order.end(), 0);
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return label[a] < label[b]; });
        c->h_obs_label.resize(n_obs);
        for (int i = 0; i < n_obs; ++i) {
            c->h_obs_label[i] = label[order[i]];
            for (int k = 0; k < 3; ++k) c->h_obs_xyz[3 * (size_t)i + k] = xyz_aos[3 * (size_t)order[i] + k];
        }
    }
    int rc;
    // the label-sorted cloud, its labels and the per-label ranges live on the device from here on
    {
        std::vector<float4> xyz4((size_t)std::max(n_obs, 1));
        c->obs_nonfinite = false;
        for (int k = 0; k < 3; ++k) { c->obs_min[k] = INFINITY; c->obs_max[k] = -INFINITY; }
        for (int i = 0; i < n_obs; ++i) {
            const float* v = &c->h_obs_xyz[3 * (size_t)i];
            xyz4[i] = make_float4(v[0], v[1], v[2], 0.f);
            for (int k = 0; k < 3; ++k) {
                if (!std::isfinite(v[k])) c->obs_nonfinite = true;
                c->obs_min[k] = std::min(c->obs_min[k], v[k]);
                c->obs_max[k] = std::max(c->obs_max[k], v[k]);
            }
        }
        int n_labels = 0;
        for (int l : c->h_obs_label) n_labels = std::max(n_labels, l + 1);
        std::vector<int> range(2 * (size_t)std::max(n_labels, 1), 0);
        for (int i = 0; i < (int)c->h_obs_label.size(); ++i) {  // labels are sorted ascending
            const int l = c->h_obs_label[i];
            if (l < 0) continue;
            if (range[2 * l + 1] == 0) range[2 * l] = i;
            range[2 * l + 1] = i + 1;
        }
        if ((rc = upload(&c->d_obs_xyz4, xyz4.data(), xyz4.size(), c->stream))) return rc;
        if ((rc = upload(&c->d_obs_label_sorted, label ? c->h_obs_label.data() : nullptr, label ? (size_t)std::max(n_obs, 1) : 0, c->stream))) return rc;
        if ((rc = upload(&c->d_label_range, range.data(), range.size(), c->stream))) return rc;
        CU(cudaStreamSynchronize(c->stream));  // host vectors go out of scope
        c->n_label_ranges = n_labels;
        c->obs_host_valid = true;
        c->in_on_device = false;
    }
    if ((rc = upload<uint8_t>(&c->d_obs_rgb, nullptr, 3 * (size_t)std::max(n_obs, 1), c->stream))) return rc;
    if ((rc = upload<float>(&c->d_obs_lab, nullptr, 3 * (size_t)std::max(n_obs, 1), c->stream))) return rc;
    if (n_obs > 0) {
        if (rgb_soa) CU(cudaMemcpyAsync(c->d_obs_rgb, rgb_soa, 3 * (size_t)n_obs, cudaMemcpyHostToDevice, c->stream));
        else CU(cudaMemsetAsync(c->d_obs_rgb, 0, 3 * (size_t)n_obs, c->stream));
        // colours stay in the ORIGINAL order and are looked up with (sorted) NN indices, as renderer.cu:1881 does
        CU(perch::launch_rgb2lab_planes(c->d_obs_rgb, c->d_obs_rgb + n_obs, c->d_obs_rgb + 2 * (size_t)n_obs, n_obs,
                                        c->d_obs_lab, c->stream));
    }
    CU(cudaStreamSynchronize(c->stream));
    c->grid_res = -1.f;  // grid is (re)built on the next cost call, when sensor_resolution is known
    c->tgt_normals_k = 0;
    c->have_obs = true;
    c->obs_digest[0] = digest[0];
    c->obs_digest[1] = digest[1];
    c->obs_digest_valid = true;
    ++c->obs_version;
    note_memory(c);
    return PERCH_OK;
}

}  // extern "C"

// The host mirrors of the label-sorted cloud (h_obs_xyz, h_obs_label): valid after perch_set_observed, fetched from the
// device on first use after perch_set_input (GICP's target-normal preparation, the unfused check path, label counts).
int perch_detail::ensure_obs_host(perch_ctx* c) {
    if (c->obs_host_valid) return PERCH_OK;
    const int n = c->n_obs;
    std::vector<float4> xyz4((size_t)std::max(n, 1));
    c->h_obs_xyz.assign(3 * (size_t)n, 0.f);
    c->h_obs_label.assign(c->obs_has_label ? n : 0, 0);
    if (n > 0) {
        CU(cudaMemcpyAsync(xyz4.data(), c->d_obs_xyz4, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
        if (c->obs_has_label)
            CU(cudaMemcpyAsync(c->h_obs_label.data(), c->d_obs_label_sorted, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        for (int i = 0; i < n; ++i) {
            c->h_obs_xyz[3 * (size_t)i] = xyz4[i].x; c->h_obs_xyz[3 * (size_t)i + 1] = xyz4[i].y; c->h_obs_xyz[3 * (size_t)i + 2] = xyz4[i].z;
        }
    }
    c->obs_host_valid = true;
    return PERCH_OK;
}

// perch_set_observed for a cloud that is already on the device (perch_set_input): d_aos n x 3, colours as three planes
// `rgb_stride` bytes apart, d_label (may be null) in [0, 256).  Stable sort by label, colours / CIELAB in the original
// order, bounding box, per-label ranges -- all on the context's stream; 24 bytes (the box) + 4 (the label check) come back.
int perch_detail::install_observed_device(perch_ctx* c, const float* d_aos, const uint8_t* d_rgb, size_t rgb_stride,
                                          const int* d_label, int n) {
    if (n > (1 << 24)) return fail(PERCH_ERR_CAPACITY, "observed cloud larger than 16M points");
    c->obs_digest_valid = false;
    c->n_obs = n;
    c->obs_has_label = d_label != nullptr;
    c->obs_nonfinite = false;
    const size_t N = (size_t)std::max(n, 1);
    int rc;
    if ((rc = upload<float4>(&c->d_obs_xyz4, nullptr, N, c->stream))) return rc;
    if ((rc = upload<int>(&c->d_obs_label_sorted, nullptr, d_label ? N : 0, c->stream))) return rc;
    if ((rc = upload<int>(&c->d_label_range, nullptr, 512, c->stream))) return rc;
    if ((rc = upload<uint8_t>(&c->d_obs_rgb, nullptr, 3 * N, c->stream))) return rc;
    if ((rc = upload<float>(&c->d_obs_lab, nullptr, 3 * N, c->stream))) return rc;
    const size_t hist_n = (size_t)perch::label_sort_chunks(n) * 256;
    if ((rc = ensure_obs_tmp(c, hist_n + 64))) return rc;
    int* d_small = c->d_obs_tmp + hist_n;  // [0..5] bounding box, [6] bad-label flag
    CU(perch::launch_obs_bbox(d_aos, n, d_small, c->stream));
    if (d_label) {
        CU(perch::launch_label_sort(d_aos, d_label, n, c->d_obs_tmp, c->d_label_range, d_small + 6, c->d_obs_xyz4,
                                    c->d_obs_label_sorted, c->stream));
        c->n_label_ranges = 256;
    } else {
        CU(perch::launch_aos_to_xyz4(d_aos, n, c->d_obs_xyz4, c->stream));
        CU(cudaMemsetAsync(c->d_label_range, 0, 512 * sizeof(int), c->stream));
        CU(cudaMemsetAsync(d_small + 6, 0, sizeof(int), c->stream));
        c->n_label_ranges = 0;
    }
    if (n > 0) {
        for (int ch = 0; ch < 3; ++ch)
            CU(cudaMemcpyAsync(c->d_obs_rgb + (size_t)ch * n, d_rgb + (size_t)ch * rgb_stride, (size_t)n, cudaMemcpyDeviceToDevice, c->stream));
        // colours stay in the ORIGINAL order and are looked up with (sorted) NN indices, as renderer.cu:1881 does
        CU(perch::launch_rgb2lab_planes(c->d_obs_rgb, c->d_obs_rgb + n, c->d_obs_rgb + 2 * (size_t)n, n, c->d_obs_lab, c->stream));
    }
    int small[7] = {0, 0, 0, 0, 0, 0, 0};
    CU(cudaMemcpyAsync(small, d_small, sizeof small, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (small[6]) return fail(PERCH_ERR_INVALID, "an observed label lies outside [0, 256)");
    for (int k = 0; k < 3; ++k) {
        c->obs_min[k] = n > 0 ? perch::obs_bbox_decode(small[k]) : 0.f;
        c->obs_max[k] = n > 0 ? perch::obs_bbox_decode(small[3 + k]) : 0.f;
    }
    c->h_obs_xyz.clear();
    c->h_obs_label.clear();
    c->obs_host_valid = n == 0;
    c->grid_res = -1.f;  // grid is (re)built on the next cost call, when sensor_resolution is known
    c->tgt_normals_k = 0;
    c->have_obs = true;
    ++c->obs_version;
    note_memory(c);
    return PERCH_OK;
}

extern "C" {

int perch_depth2cloud(perch_ctx* c, const int32_t* depth, const uint8_t* red, const uint8_t* green, const uint8_t* blue,
                      int width, int height, int num_poses, float cx, float cy, float fx, float fy, float depth_factor,
                      int stride, const uint8_t* label_mask, const double* bounds6, const float* camera_transform16,
                      int* point_num, float* out_cloud_soa, float* out_cloud_aos, uint8_t* out_cloud_color,
                      int* out_dc_index, int* out_cloud_pose_map, int* out_cloud_label) {
    if (!c || !depth || !point_num) return fail(PERCH_ERR_INVALID, "null argument");
    if (width <= 0 || height <= 0 || num_poses <= 0 || stride <= 0) return fail(PERCH_ERR_INVALID, "bad size argument");
    if (width % stride != 0)
        return fail(PERCH_ERR_INVALID, "stride %d must divide the image width %d (compute_point_clouds.cuh:271)", stride, width);
    DeviceGuard g(c->device);
    const size_t npx = (size_t)width * height * num_poses;
    if (npx > (size_t)INT_MAX) return fail(PERCH_ERR_CAPACITY, "num_poses*width*height exceeds INT_MAX");
    const int Lh = (height + stride - 1) / stride;
    const size_t rows = (size_t)num_poses * Lh;
    const bool want_out = out_cloud_soa || out_cloud_aos || out_cloud_color || out_cloud_pose_map || out_cloud_label;
    const int capacity = want_out ? *point_num : 0;
    if (want_out && capacity < 0) return fail(PERCH_ERR_INVALID, "point_num must hold the buffer capacity on pass 2");
    // scratch: depth + 3 colour planes + label + row counts + outputs
    size_t need = Scratch::align(npx * 4) + 4 * Scratch::align(npx) + Scratch::align((rows + 1) * 4) +
                  2 * Scratch::align((size_t)capacity * 12) + Scratch::align((size_t)capacity * 3) +
                  2 * Scratch::align((size_t)capacity * 4) + (out_dc_index ? Scratch::align(npx * 4) : 0) + 4096;
    int rc = ensure_scratch(c, need);
    if (rc) return rc;
    Scratch s(c);
    perch::CloudArgs a;
    int32_t* d_depth = s.take<int32_t>(npx);
    CU(cudaMemcpyAsync(d_depth, depth, npx * 4, cudaMemcpyHostToDevice, c->stream));
    a.depth = d_depth;
    const uint8_t* planes[3] = {red, green, blue};
    const uint8_t* d_planes[3] = {nullptr, nullptr, nullptr};
    for (int k = 0; k < 3; ++k) {
        uint8_t* p = s.take<uint8_t>(npx);
        if (planes[k]) {
            CU(cudaMemcpyAsync(p, planes[k], npx, cudaMemcpyHostToDevice, c->stream));
            d_planes[k] = p;
        }
    }
    a.red = d_planes[0]; a.green = d_planes[1]; a.blue = d_planes[2];
    uint8_t* d_label = s.take<uint8_t>(npx);
    if (label_mask) {
        CU(cudaMemcpyAsync(d_label, label_mask, npx, cudaMemcpyHostToDevice, c->stream));
        a.label_mask = d_label;
    }
    a.img_w = width; a.img_h = height; a.full_h = height; a.num_images = num_poses;
    a.stride = stride; a.coord_scale = 1;
    a.cx = cx; a.cy = cy; a.fx = fx; a.fy = fy; a.depth_factor = depth_factor;
    a.use_bounds = bounds6 != nullptr && camera_transform16 != nullptr;
    if (a.use_bounds) {
        for (int k = 0; k < 6; ++k) a.bounds[k] = (float)bounds6[k];
        memcpy(a.cam_tf, camera_transform16, 64);
    }
    a.row_counts = s.take<int>(rows + 1);
    a.capacity = capacity;
    CU(perch::launch_cloud_count(a, c->stream));
    CU(perch::launch_cloud_scan(a, c->stream));
    int total = 0;
    CU(cudaMemcpyAsync(&total, a.row_counts + rows, 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (!want_out && !out_dc_index) {
        *point_num = total;
        return PERCH_OK;
    }
    if (want_out && capacity < total)
        return fail(PERCH_ERR_CAPACITY, "cloud has %d points but the buffers hold %d", total, capacity);
    float* d_soa = s.take<float>((size_t)capacity * 3);
    float* d_aos = s.take<float>((size_t)capacity * 3);
    uint8_t* d_rgb = s.take<uint8_t>((size_t)capacity * 3);
    int* d_map = s.take<int>(capacity);
    int* d_lab = s.take<int>(capacity);
    int* d_dc = out_dc_index ? s.take<int>(npx) : nullptr;
    // the reference zero-initialises its outputs (compute_point_clouds.cuh:302-305,317); slots whose pixel fails the
    // second bounds test stay zero
    if (capacity > 0) {
        CU(cudaMemsetAsync(d_soa, 0, (size_t)capacity * 12, c->stream));
        CU(cudaMemsetAsync(d_aos, 0, (size_t)capacity * 12, c->stream));
        CU(cudaMemsetAsync(d_rgb, 0, (size_t)capacity * 3, c->stream));
        CU(cudaMemsetAsync(d_map, 0, (size_t)capacity * 4, c->stream));
        CU(cudaMemsetAsync(d_lab, 0, (size_t)capacity * 4, c->stream));
    }
    a.cloud_soa = out_cloud_soa ? d_soa : nullptr;
    a.cloud_aos = out_cloud_aos ? d_aos : nullptr;
    a.cloud_rgb = out_cloud_color ? d_rgb : nullptr;
    a.pose_map = out_cloud_pose_map ? d_map : nullptr;
    a.cloud_label = (out_cloud_label && label_mask) ? d_lab : nullptr;
    a.dc_index = d_dc;
    CU(perch::launch_cloud_write(a, c->stream));
    // out_cloud_soa rows are `capacity` floats apart on the device; the host buffer is 3 x total (point_num) like the
    // reference's cudaMemcpy2D (renderer.cu:2052-2053)
    for (int k = 0; k < 3 && total > 0; ++k) {
        if (out_cloud_soa)
            CU(cudaMemcpyAsync(out_cloud_soa + (size_t)k * total, d_soa + (size_t)k * capacity, (size_t)total * 4,
                               cudaMemcpyDeviceToHost, c->stream));
        if (out_cloud_color)
            CU(cudaMemcpyAsync(out_cloud_color + (size_t)k * total, d_rgb + (size_t)k * capacity, (size_t)total,
                               cudaMemcpyDeviceToHost, c->stream));
    }
    if (out_cloud_aos && total > 0)
        CU(cudaMemcpyAsync(out_cloud_aos, d_aos, (size_t)total * 12, cudaMemcpyDeviceToHost, c->stream));
    if (out_cloud_pose_map && total > 0)
        CU(cudaMemcpyAsync(out_cloud_pose_map, d_map, (size_t)total * 4, cudaMemcpyDeviceToHost, c->stream));
    if (out_cloud_label && label_mask && total > 0)
        CU(cudaMemcpyAsync(out_cloud_label, d_lab, (size_t)total * 4, cudaMemcpyDeviceToHost, c->stream));
    if (out_dc_index) CU(cudaMemcpyAsync(out_dc_index, d_dc, npx * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    *point_num = total;
    note_memory(c);
    return PERCH_OK;
}

int perch_render_cost_device(perch_ctx* c, const float* d_poses16, int n_poses, const int* d_pose_model_map,
                             const int* d_pose_label, const float* d_pose_obs_total, const perch_params* p,
                             float* d_costs3p, int32_t* d_counts3p) {
    int rc = check_ready(c, true);
    if (rc) return rc;
    if (!p || !d_poses16 || !d_pose_model_map || !d_costs3p || n_poses < 0)
        return fail(PERCH_ERR_INVALID, "bad render_cost_device argument");
    DeviceGuard g(c->device);
    perch::PoseArgs a;
    bool color = false;
    a.pose_label = d_pose_label;
    if ((rc = prepare_cost_args(c, p, a, color))) return rc;
    a.poses = reinterpret_cast<const perch::Mat4*>(d_poses16);
    a.pose_model_map = d_pose_model_map;
    a.pose_obs_total = d_pose_obs_total;
    a.n_poses = n_poses;
    a.out_costs = d_costs3p;
    a.out_counts = d_counts3p;
    CU(cudaEventRecord(c->ev0, c->stream));
    CU(perch::launch_fused_cost(a, color, c->stream));
    CU(cudaEventRecord(c->ev1, c->stream));
    c->last_launches = n_poses > 0 ? 1 : 0;
    c->timing_pending = true;
    return PERCH_OK;
}

int perch_device_alloc(perch_ctx* c, size_t bytes, void** d_ptr) {
    if (!c || !d_ptr) return fail(PERCH_ERR_INVALID, "null argument");
    DeviceGuard g(c->device);
    CU(cudaMalloc(d_ptr, bytes ? bytes : 1));
    return PERCH_OK;
}
int perch_device_free(perch_ctx* c, void* d_ptr) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    DeviceGuard g(c->device);
    CU(cudaFree(d_ptr));
    return PERCH_OK;
}
int perch_memcpy_h2d(perch_ctx* c, void* d_dst, const void* h_src, size_t bytes) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    DeviceGuard g(c->device);
    CU(cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return PERCH_OK;
}
int perch_memcpy_d2h(perch_ctx* c, void* h_dst, const void* d_src, size_t bytes) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    DeviceGuard g(c->device);
    CU(cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return PERCH_OK;
}
int perch_synchronize(perch_ctx* c) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    DeviceGuard g(c->device);
    CU(cudaStreamSynchronize(c->stream));
    if (c->timing_pending) {
        cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1);
        c->timing_pending = false;
    }
    return check_err_flag(c);
}

int perch_timer_start(perch_ctx* c) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    DeviceGuard g(c->device);
    CU(cudaEventRecord(c->ev_t0, c->stream));
    return PERCH_OK;
}
int perch_timer_stop(perch_ctx* c) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    DeviceGuard g(c->device);
    CU(cudaEventRecord(c->ev_t1, c->stream));
    return PERCH_OK;
}
int perch_timer_elapsed(perch_ctx* c, float* ms) {
    if (!c || !ms) return fail(PERCH_ERR_INVALID, "null argument");
    DeviceGuard g(c->device);
    CU(cudaEventSynchronize(c->ev_t1));
    CU(cudaEventElapsedTime(ms, c->ev_t0, c->ev_t1));
    return PERCH_OK;
}

int perch_l2_flush(perch_ctx* c) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    DeviceGuard g(c->device);
    const size_t bytes = (size_t)256 << 20;
    if (!c->d_flush) CU(cudaMalloc(&c->d_flush, bytes));
    c->flush_value ^= 0xFF;
    CU(cudaMemsetAsync(c->d_flush, c->flush_value, bytes, c->stream));
    return PERCH_OK;
}

int perch_gicp(perch_ctx* c, const float* poses16, int n_poses, const int* pose_model_map, const int* pose_label,
               const perch_params* p, float* out_poses16, float* out_transforms12, int* out_iterations, float* out_error) {
    if (!p) return fail(PERCH_ERR_INVALID, "null params");
    int rc = check_ready(c, true);
    if (rc) return rc;
    if (n_poses < 0 || (n_poses > 0 && (!poses16 || !pose_model_map))) return fail(PERCH_ERR_INVALID, "bad pose arguments");
    if (n_poses == 0) return PERCH_OK;
    DeviceGuard g(c->device);
    const size_t P = (size_t)n_poses;
    if ((rc = ensure_scratch(c, 2 * Scratch::align(P * 64) + 4 * Scratch::align(P * 4) + Scratch::align(P * 48) + 4096))) return rc;
    Scratch sc(c);
    float* d_poses = sc.take<float>(P * 16);
    float* d_adj = sc.take<float>(P * 16);
    int* d_pmm = sc.take<int>(P);
    int* d_plabel = sc.take<int>(P);
    int* d_it = sc.take<int>(P);
    float* d_err = sc.take<float>(P);
    float* d_tf = sc.take<float>(P * 12);
    CU(cudaMemcpyAsync(d_poses, poses16, P * 64, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_pmm, pose_model_map, P * 4, cudaMemcpyHostToDevice, c->stream));
    if (pose_label) CU(cudaMemcpyAsync(d_plabel, pose_label, P * 4, cudaMemcpyHostToDevice, c->stream));
    perch::PoseArgs base;
    fill_pose_args(c, p, base);
    base.poses = reinterpret_cast<const perch::Mat4*>(d_poses);
    base.pose_model_map = d_pmm;
    base.pose_label = pose_label ? d_plabel : nullptr;
    base.n_poses = n_poses;
    base.err_flag = c->d_err_flag;
    int launches = 0;
    if ((rc = run_gicp(c, p, base, n_poses, d_adj, d_tf, d_it, d_err, &launches))) return rc;
    if (out_poses16) CU(cudaMemcpyAsync(out_poses16, d_adj, P * 64, cudaMemcpyDeviceToHost, c->stream));
    if (out_transforms12) CU(cudaMemcpyAsync(out_transforms12, d_tf, P * 48, cudaMemcpyDeviceToHost, c->stream));
    if (out_iterations) CU(cudaMemcpyAsync(out_iterations, d_it, P * 4, cudaMemcpyDeviceToHost, c->stream));
    if (out_error) CU(cudaMemcpyAsync(out_error, d_err, P * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->last_ms = c->last_gicp_ms;
    c->last_launches = launches;
    c->timing_pending = false;
    note_memory(c);
    return check_err_flag(c);
}

int perch_icp_set_scene(perch_ctx* c, const int32_t* scene_depth, int width, int height, const float K9[9],
                        float max_dist_diff, float* out_pcd, float* out_normals) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    if (!scene_depth || !K9 || width <= 0 || height <= 0) return fail(PERCH_ERR_INVALID, "bad icp scene argument");
    DeviceGuard g(c->device);
    const size_t npx = (size_t)width * height;
    int rc;
    if ((rc = upload<float>(&c->d_icp_pcd, nullptr, 3 * npx, c->stream))) return rc;
    if ((rc = upload<float>(&c->d_icp_normals, nullptr, 3 * npx, c->stream))) return rc;
    if ((rc = ensure_scratch(c, Scratch::align(npx * 4) + 1024))) return rc;
    Scratch s(c);
    int32_t* d_depth = s.take<int32_t>(npx);
    CU(cudaMemcpyAsync(d_depth, scene_depth, npx * 4, cudaMemcpyHostToDevice, c->stream));
    CU(perch::launch_icp_scene(d_depth, width, height, K9, c->d_icp_pcd, c->d_icp_normals, c->stream));
    if (out_pcd) CU(cudaMemcpyAsync(out_pcd, c->d_icp_pcd, 3 * npx * 4, cudaMemcpyDeviceToHost, c->stream));
    if (out_normals) CU(cudaMemcpyAsync(out_normals, c->d_icp_normals, 3 * npx * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->icp_w = width; c->icp_h = height;
    memcpy(c->icp_K, K9, sizeof c->icp_K);
    c->icp_max_dist = max_dist_diff;
    c->icp_grid_dist = -1.f;
    note_memory(c);
    return PERCH_OK;
}

namespace {
// Scene_nn's cloud (Scene_nn::init_Scene_nn_cpu, pcd_scene.cpp:4-38): every pixel with a depth, row-major; here the
// pixel index doubles as the point's reference index into the per-pixel pcd / normal buffers.
int prepare_icp_nn_grid(perch_ctx* c, float max_dist) {
    if (c->icp_grid_dist == max_dist) return PERCH_OK;
    const size_t npx = (size_t)c->icp_w * c->icp_h;
    std::vector<float> pcd(3 * npx);
    CU(cudaMemcpyAsync(pcd.data(), c->d_icp_pcd, 3 * npx * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    std::vector<float> xyz;
    std::vector<int> ids;
    xyz.reserve(3 * npx);
    ids.reserve(npx);
    for (size_t i = 0; i < npx; ++i)
        if (pcd[3 * i + 2] > 0.f) {
            xyz.insert(xyz.end(), {pcd[3 * i], pcd[3 * i + 1], pcd[3 * i + 2]});
            ids.push_back((int)i);
        }
    const int n = (int)ids.size();
    int rc = build_point_grid_from(c, xyz.data(), n, ids.data(), nullptr, std::max(max_dist, 1e-4f) * 1.001f, 64.0 * 1024 * 1024,
                                   &c->d_icp_grid_pts, &c->d_icp_grid_label, &c->d_icp_grid_cells, &c->icp_grid, nullptr);
    if (rc) return rc;
    c->icp_grid_dist = max_dist;
    return PERCH_OK;
}
}  // namespace

namespace {
int icp_point2plane_impl(perch_ctx* c, bool nn, float nn_max_dist, const float* model_pcd, const int* cloud_offsets, int n_clouds,
                         float relative_fitness, float relative_rmse, int max_iteration, float* out_transforms16,
                         float* out_fitness, float* out_inlier_rmse, int* out_iterations, float* out_pcd) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    if (c->icp_w == 0) return fail(PERCH_ERR_INVALID, "perch_icp_set_scene has not been called");
    if (n_clouds < 0 || (n_clouds > 0 && (!cloud_offsets || !out_transforms16 || !out_fitness || !out_inlier_rmse)) || max_iteration < 0)
        return fail(PERCH_ERR_INVALID, "bad icp argument");
    if (n_clouds == 0) return PERCH_OK;
    for (int i = 0; i < n_clouds; ++i)
        if (cloud_offsets[i + 1] < cloud_offsets[i] || cloud_offsets[0] != 0) return fail(PERCH_ERR_INVALID, "cloud_offsets must ascend from 0");
    const size_t n = (size_t)cloud_offsets[n_clouds], B = (size_t)n_clouds;
    if (n > 0 && !model_pcd) return fail(PERCH_ERR_INVALID, "null model_pcd");
    DeviceGuard g(c->device);
    int rc;
    if ((rc = ensure_scratch(c, Scratch::align(std::max<size_t>(n, 1) * 12) + Scratch::align((B + 1) * 4) + Scratch::align(B * 64) +
                                    3 * Scratch::align(B * 4) + 1024))) return rc;
    Scratch s(c);
    perch::IcpArgs a;
    a.points = s.take<float>(std::max<size_t>(n, 1) * 3);
    int* d_off = s.take<int>(B + 1);
    a.offsets = d_off;
    a.out_T = s.take<float>(B * 16);
    a.out_fitness = s.take<float>(B);
    a.out_rmse = s.take<float>(B);
    a.out_iterations = s.take<int>(B);
    if (n > 0) CU(cudaMemcpyAsync(a.points, model_pcd, n * 12, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_off, cloud_offsets, (B + 1) * 4, cudaMemcpyHostToDevice, c->stream));
    a.scene_pcd = c->d_icp_pcd; a.scene_normals = c->d_icp_normals;
    a.width = c->icp_w; a.height = c->icp_h;
    a.k00 = c->icp_K[0]; a.k02 = c->icp_K[2]; a.k11 = c->icp_K[4]; a.k12 = c->icp_K[5];
    a.max_dist_diff = c->icp_max_dist;
    if (nn) {
        if (!(nn_max_dist > 0.f)) return fail(PERCH_ERR_INVALID, "max_dist_diff must be positive");
        if ((rc = prepare_icp_nn_grid(c, nn_max_dist))) return rc;
        a.nn = 1;
        a.nn_grid = c->icp_grid;
        a.max_dist_diff = nn_max_dist;
    }
    a.relative_fitness = relative_fitness; a.relative_rmse = relative_rmse; a.max_iteration = max_iteration;
    CU(cudaEventRecord(c->ev0, c->stream));
    CU(perch::launch_icp_point2plane(a, n_clouds, c->stream));
    CU(cudaEventRecord(c->ev1, c->stream));
    CU(cudaMemcpyAsync(out_transforms16, a.out_T, B * 64, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(out_fitness, a.out_fitness, B * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(out_inlier_rmse, a.out_rmse, B * 4, cudaMemcpyDeviceToHost, c->stream));
    if (out_iterations) CU(cudaMemcpyAsync(out_iterations, a.out_iterations, B * 4, cudaMemcpyDeviceToHost, c->stream));
    if (out_pcd && n > 0) CU(cudaMemcpyAsync(out_pcd, a.points, n * 12, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1);
    c->last_launches = 1;
    c->timing_pending = false;
    note_memory(c);
    return PERCH_OK;
}
}  // namespace

int perch_icp_get_scene_device(perch_ctx* c, const float** d_pcd, const float** d_normals, int* width, int* height) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    if (c->icp_w == 0) return fail(PERCH_ERR_INVALID, "perch_icp_set_scene has not been called");
    if (d_pcd) *d_pcd = c->d_icp_pcd;
    if (d_normals) *d_normals = c->d_icp_normals;
    if (width) *width = c->icp_w;
    if (height) *height = c->icp_h;
    return PERCH_OK;
}

int perch_icp_point2plane_device(perch_ctx* c, float* d_model_pcd, int n_points, const float* d_scene_pcd,
                                 const float* d_scene_normals, int width, int height, const float K9[9], float max_dist_diff,
                                 int nearest_neighbour, float relative_fitness, float relative_rmse, int max_iteration,
                                 float out_transform16[16], float* out_fitness, float* out_inlier_rmse, int* out_iterations) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    if (n_points < 0 || (n_points > 0 && !d_model_pcd) || !out_transform16 || !out_fitness || !out_inlier_rmse || max_iteration < 0)
        return fail(PERCH_ERR_INVALID, "bad icp argument");
    const bool own_scene = d_scene_pcd == nullptr;
    if (own_scene && c->icp_w == 0) return fail(PERCH_ERR_INVALID, "perch_icp_set_scene has not been called");
    if (!own_scene && (!d_scene_normals || !K9 || width <= 0 || height <= 0)) return fail(PERCH_ERR_INVALID, "bad icp scene argument");
    if (nearest_neighbour && !own_scene && d_scene_pcd != c->d_icp_pcd)
        return fail(PERCH_ERR_INVALID, "nearest-neighbour association runs against the scene of perch_icp_set_scene");
    if (nearest_neighbour && c->icp_w == 0) return fail(PERCH_ERR_INVALID, "perch_icp_set_scene has not been called");
    DeviceGuard g(c->device);
    int rc;
    if ((rc = ensure_scratch(c, Scratch::align(2 * 4) + Scratch::align(64) + 3 * Scratch::align(4) + 1024))) return rc;
    Scratch s(c);
    perch::IcpArgs a;
    a.points = d_model_pcd;
    int* d_off = s.take<int>(2);
    a.offsets = d_off;
    a.out_T = s.take<float>(16);
    a.out_fitness = s.take<float>(1);
    a.out_rmse = s.take<float>(1);
    a.out_iterations = s.take<int>(1);
    const int off[2] = {0, n_points};
    CU(cudaMemcpyAsync(d_off, off, sizeof off, cudaMemcpyHostToDevice, c->stream));
    const float* K = own_scene || nearest_neighbour ? c->icp_K : K9;
    a.scene_pcd = own_scene ? c->d_icp_pcd : d_scene_pcd;
    a.scene_normals = own_scene ? c->d_icp_normals : d_scene_normals;
    a.width = own_scene ? c->icp_w : width;
    a.height = own_scene ? c->icp_h : height;
    a.k00 = K[0]; a.k02 = K[2]; a.k11 = K[4]; a.k12 = K[5];
    a.max_dist_diff = max_dist_diff;
    if (nearest_neighbour) {
        if (!(max_dist_diff > 0.f)) return fail(PERCH_ERR_INVALID, "max_dist_diff must be positive");
        if ((rc = prepare_icp_nn_grid(c, max_dist_diff))) return rc;
        a.nn = 1;
        a.nn_grid = c->icp_grid;
    }
    a.relative_fitness = relative_fitness; a.relative_rmse = relative_rmse; a.max_iteration = max_iteration;
    CU(cudaEventRecord(c->ev0, c->stream));
    CU(perch::launch_icp_point2plane(a, 1, c->stream));
    CU(cudaEventRecord(c->ev1, c->stream));
    CU(cudaMemcpyAsync(out_transform16, a.out_T, 64, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(out_fitness, a.out_fitness, 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(out_inlier_rmse, a.out_rmse, 4, cudaMemcpyDeviceToHost, c->stream));
    if (out_iterations) CU(cudaMemcpyAsync(out_iterations, a.out_iterations, 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1);
    c->last_launches = 1;
    c->timing_pending = false;
    return PERCH_OK;
}

int perch_icp_depth2cloud_device(perch_ctx* c, const void* d_depth, int depth_bits, int width, int height, const float K9[9],
                                 int stride, int tl_x, int tl_y, float* d_out_xyz, int capacity, int* n_points) {
    if (!c || !d_depth || !K9 || !n_points) return fail(PERCH_ERR_INVALID, "null argument");
    if ((depth_bits != 16 && depth_bits != 32) || width <= 0 || height <= 0 || stride <= 0 || capacity < 0 || width % stride != 0)
        return fail(PERCH_ERR_INVALID, "bad depth2cloud argument (depth_bits 16 | 32; stride must divide the width)");
    DeviceGuard g(c->device);
    const size_t npx = (size_t)width * height;
    const int Lh = (height + stride - 1) / stride;
    int rc;
    if ((rc = ensure_scratch(c, Scratch::align(npx * 4) + Scratch::align(((size_t)Lh + 1) * 4) + 1024))) return rc;
    Scratch s(c);
    int32_t* d32 = s.take<int32_t>(npx);
    perch::CloudArgs a;
    if (depth_bits == 16) {
        CU(perch::launch_depth_to_i32(d_depth, 16, d32, (int)npx, c->stream));
        a.depth = d32;
    } else {
        a.depth = static_cast<const int32_t*>(d_depth);
    }
    a.img_w = width; a.img_h = height; a.full_h = height; a.num_images = 1; a.stride = stride; a.coord_scale = 1;
    a.lattice_coords = 1; a.u_off = tl_x; a.v_off = tl_y;                 // icp.cu:252-253: (x + tl_x - K02) / K00 * z
    a.cx = K9[2]; a.cy = K9[5]; a.fx = K9[0]; a.fy = K9[4]; a.depth_factor = 100.f;  // icp.cu:251: depth / 100.0f
    a.row_counts = s.take<int>((size_t)Lh + 1);
    a.capacity = capacity;
    a.cloud_aos = d_out_xyz;
    CU(perch::launch_cloud_count(a, c->stream));
    CU(perch::launch_cloud_scan(a, c->stream));
    if (d_out_xyz && capacity > 0) CU(perch::launch_cloud_write(a, c->stream));
    int total = 0;
    CU(cudaMemcpyAsync(&total, a.row_counts + Lh, 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    *n_points = total;
    if (d_out_xyz && total > capacity) return fail(PERCH_ERR_CAPACITY, "cloud has %d points but the buffer holds %d", total, capacity);
    return PERCH_OK;
}

int perch_icp_point2plane(perch_ctx* c, const float* model_pcd, const int* cloud_offsets, int n_clouds,
                          float relative_fitness, float relative_rmse, int max_iteration, float* out_transforms16,
                          float* out_fitness, float* out_inlier_rmse, int* out_iterations, float* out_pcd) {
    return icp_point2plane_impl(c, false, 0.f, model_pcd, cloud_offsets, n_clouds, relative_fitness, relative_rmse, max_iteration,
                                out_transforms16, out_fitness, out_inlier_rmse, out_iterations, out_pcd);
}

int perch_icp_point2plane_nn(perch_ctx* c, const float* model_pcd, const int* cloud_offsets, int n_clouds, float max_dist_diff,
                             float relative_fitness, float relative_rmse, int max_iteration, float* out_transforms16,
                             float* out_fitness, float* out_inlier_rmse, int* out_iterations, float* out_pcd) {
    return icp_point2plane_impl(c, true, max_dist_diff, model_pcd, cloud_offsets, n_clouds, relative_fitness, relative_rmse,
                                max_iteration, out_transforms16, out_fitness, out_inlier_rmse, out_iterations, out_pcd);
}

int perch_knn(perch_ctx* c, const float* ref, int ref_nb, const float* query, int query_nb, int dim, int k,
              float* knn_dist, int* knn_index) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    if (!ref || !query || !knn_dist || !knn_index) return fail(PERCH_ERR_INVALID, "null knn argument");
    if (ref_nb <= 0 || query_nb < 0 || dim <= 0 || k <= 0) return fail(PERCH_ERR_INVALID, "bad knn sizes");
    if (k > ref_nb) return fail(PERCH_ERR_INVALID, "k = %d exceeds the number of reference points %d", k, ref_nb);
    if (query_nb == 0) return PERCH_OK;
    DeviceGuard g(c->device);
    const size_t nr = (size_t)dim * ref_nb, nq = (size_t)dim * query_nb, no = (size_t)k * query_nb;
    int rc = ensure_scratch(c, Scratch::align(nr * 4) + Scratch::align(nq * 4) + 2 * Scratch::align(no * 4) + 1024);
    if (rc) return rc;
    Scratch s(c);
    float* d_ref = s.take<float>(nr);
    float* d_query = s.take<float>(nq);
    float* d_dist = s.take<float>(no);
    int* d_index = s.take<int>(no);
    CU(cudaMemcpyAsync(d_ref, ref, nr * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_query, query, nq * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaEventRecord(c->ev0, c->stream));
    CU(perch::launch_knn(d_ref, ref_nb, d_query, query_nb, dim, k, d_dist, d_index, c->sm_count, c->stream));
    CU(cudaEventRecord(c->ev1, c->stream));
    CU(cudaMemcpyAsync(knn_dist, d_dist, no * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(knn_index, d_index, no * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1);
    c->last_launches = 1;
    c->timing_pending = false;
    note_memory(c);
    return PERCH_OK;
}

int perch_render_cost(perch_ctx* c, const char* stage, const float* poses16, int n_poses, const int* pose_model_map,
                      const int* pose_label, const float* pose_obs_total, const perch_params* p, perch_outputs* out) {
    perch_detail::PoseInputs in{poses16, pose_model_map, pose_label, pose_obs_total, false};
    return perch_detail::render_cost_impl(c, stage, in, n_poses, p, out);
}

}  // extern "C"

// The body of perch_render_cost; `in` may already be device-resident (perch_score_states, capi_states.cu).
int perch_detail::render_cost_impl(perch_ctx* c, const char* stage, const PoseInputs& in, int n_poses, const perch_params* p,
                                   perch_outputs* out) {
    const float* poses16 = in.poses16;
    const int* pose_model_map = in.model_map;
    const int* pose_label = in.label;
    const float* pose_obs_total = in.obs_total;
    if (!stage || !p || !out) return fail(PERCH_ERR_INVALID, "null argument");
    const std::string st(stage);
    const bool s_render = st == "RENDER", s_debug = st == "DEBUG";
    const bool s_cloud = st.find("CLOUD") != std::string::npos;  // renderer.cu:1821
    const bool s_cost = st.find("COST") != std::string::npos;    // renderer.cu:1908
    if (!s_render && !s_debug && !s_cloud && !s_cost) return fail(PERCH_ERR_INVALID, "unknown stage '%s'", stage);
    const bool do_cost = s_cost || s_debug;
    int rc = check_ready(c, do_cost);
    if (rc) return rc;
    if (n_poses < 0 || (n_poses > 0 && (!poses16 || !pose_model_map))) return fail(PERCH_ERR_INVALID, "bad pose arguments");
    if (p->do_icp && p->single_result_image) return fail(PERCH_ERR_INVALID, "do_icp with single_result_image is undefined");
    if (p->do_icp && !c->have_obs) return fail(PERCH_ERR_INVALID, "do_icp needs perch_set_observed (the GICP target)");
    DeviceGuard g(c->device);
    const int W = c->cam.width, H = c->cam.height;
    const size_t npx = (size_t)W * H;
    const bool do_images = s_render || s_debug;
    const bool do_cloud = s_cloud || s_debug;
    if (p->stride <= 0 || ((do_cloud || do_cost) && W % p->stride != 0))
        return fail(PERCH_ERR_INVALID, "stride %d must divide the image width %d (compute_point_clouds.cuh:271)", p->stride, W);
    if (p->single_result_image && (do_cloud || do_cost))
        return fail(PERCH_ERR_INVALID, "single_result_image is only supported for stage RENDER");
    const int s = p->stride;
    const int Lw = W / (do_cloud || do_cost ? s : 1), Lh = (H + s - 1) / s;
    const size_t P = (size_t)n_poses;
    const size_t n_img = p->single_result_image ? 1 : P;

    size_t need = 2 * Scratch::align(P * 64) + 2 * Scratch::align(P * 4) + Scratch::align(P * 4) + 2 * Scratch::align(P * 12) + 4096;
    if (do_images) need += Scratch::align(n_img * npx * 4) + 3 * Scratch::align(n_img * npx) + Scratch::align(npx * 8);
    const size_t lat = (size_t)Lw * Lh;
    // row counters, the scan total, cloud_capacity and dc_index entries are ints (like the reference's, which overflow at
    // >= 6 991 poses of 640x480, compute_point_clouds.cuh:60,119): refuse instead of wrapping
    if (do_cloud && (P * lat > (size_t)INT_MAX || (out->dc_index && P * npx > (size_t)INT_MAX)))
        return fail(PERCH_ERR_CAPACITY, "stage %s: n_poses * lattice (or image, with dc_index) size exceeds INT_MAX; use smaller batches", stage);
    if (do_images && n_img * npx > (size_t)INT_MAX * 4)
        return fail(PERCH_ERR_CAPACITY, "stage %s: n_poses * width * height exceeds 2^33 pixels; use smaller batches", stage);
    if (do_cloud) need += Scratch::align(P * lat * 4) + 3 * Scratch::align(P * lat) + Scratch::align((P * Lh + 1) * 4);
    const int cap = do_cloud ? std::max(out->cloud_capacity, 0) : 0;
    if (do_cloud) need += Scratch::align((size_t)cap * 12) + Scratch::align((size_t)cap * 3) + Scratch::align((size_t)cap * 4) +
                          (out->dc_index ? Scratch::align(P * npx * 4) : 0);
    if ((rc = ensure_scratch(c, need))) return rc;
    Scratch sc(c);
    float* d_poses = sc.take<float>(P * 16);
    int* d_pmm = sc.take<int>(P);
    int* d_plabel = sc.take<int>(P);
    float* d_tot = sc.take<float>(P);
    float* d_costs = sc.take<float>(P * 3);
    int32_t* d_counts = sc.take<int32_t>(P * 3);
    float* d_poses_adj = sc.take<float>(P * 16);
    if (P > 0) {
        const cudaMemcpyKind kind = in.on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
        CU(cudaMemcpyAsync(d_poses, poses16, P * 64, kind, c->stream));
        CU(cudaMemcpyAsync(d_pmm, pose_model_map, P * 4, kind, c->stream));
        if (pose_label) CU(cudaMemcpyAsync(d_plabel, pose_label, P * 4, kind, c->stream));
        if (pose_obs_total) CU(cudaMemcpyAsync(d_tot, pose_obs_total, P * 4, kind, c->stream));
    }
    int launches = 0;
    CU(cudaEventRecord(c->ev0, c->stream));

    perch::PoseArgs base;
    fill_pose_args(c, p, base);
    base.poses = reinterpret_cast<const perch::Mat4*>(d_poses);
    base.pose_model_map = d_pmm;
    base.pose_label = pose_label ? d_plabel : nullptr;
    base.pose_obs_total = pose_obs_total ? d_tot : nullptr;
    base.n_poses = n_poses;
    base.err_flag = c->d_err_flag;

    // ---- images (stage RENDER / DEBUG): renderer.cu:1594-1615 ------------------------------------------------
    int32_t* d_depth = nullptr;
    uint8_t *d_r = nullptr, *d_g = nullptr, *d_b = nullptr;
    if (do_images) {
        d_depth = sc.take<int32_t>(n_img * npx);
        d_r = sc.take<uint8_t>(n_img * npx);
        d_g = sc.take<uint8_t>(n_img * npx);
        d_b = sc.take<uint8_t>(n_img * npx);
        unsigned long long* d_keys = sc.take<unsigned long long>(npx);
        perch::PoseArgs a = base;
        a.stride = 1;
        a.img_depth = d_depth; a.img_r = d_r; a.img_g = d_g; a.img_b = d_b;
        a.tile_cap = std::max(env_int("PERCH_IMAGE_TILE_CAP", 6144), W);
        if (p->single_result_image) {
            a.single_keys = d_keys;
            CU(perch::launch_single_clear(d_keys, (int)npx, c->stream));
            ++launches;
        }
        if (n_poses > 0 && !env_int("PERCH_IMAGE_LEGACY", 0)) {
            // every pixel: the dedicated rasteriser (raster_full.cu); PERCH_IMAGE_LEGACY=1 keeps the per-pose kernel (A/B)
            perch::FullJob j;
            j.poses = base.poses; j.pose_model_map = base.pose_model_map; j.pose_label = base.pose_label;
            j.n_poses = n_poses; j.occlusion_threshold = base.occlusion_threshold;
            j.single_keys = p->single_result_image ? d_keys : nullptr;
            j.img_depth = d_depth; j.img_r = d_r; j.img_g = d_g; j.img_b = d_b;
            CU(perch::full_raster_job(c, j, &launches));
        } else if (n_poses > 0) {
            CU(perch::launch_raster_images(a, c->stream));
            ++launches;
        }
        if (p->single_result_image) {
            CU(perch::launch_resolve_single(a, c->stream));
            ++launches;
        }
    }
    // ---- GICP refinement (do_icp, renderer.cu:1688-1817): clouds and costs below use the adjusted poses -------------
    if (p->do_icp && (do_cloud || do_cost) && n_poses > 0) {
        if ((rc = run_gicp(c, p, base, n_poses, d_poses_adj, nullptr, nullptr, nullptr, &launches))) return rc;
        base.poses = reinterpret_cast<const perch::Mat4*>(d_poses_adj);
        if (out->adjusted_poses)
            CU(cudaMemcpyAsync(out->adjusted_poses, d_poses_adj, P * 64, cudaMemcpyDeviceToHost, c->stream));
    }
    // ---- clouds (stage CLOUD / DEBUG): renderer.cu:1617-1659,1821-1847 ------------------------------------------
    int* d_rows = nullptr;
    float* d_cloud = nullptr;
    uint8_t* d_cloud_rgb = nullptr;
    int* d_cloud_map = nullptr;
    int* d_dc = nullptr;
    if (do_cloud) {
        int32_t* l_depth = sc.take<int32_t>(P * lat);
        uint8_t* l_r = sc.take<uint8_t>(P * lat);
        uint8_t* l_g = sc.take<uint8_t>(P * lat);
        uint8_t* l_b = sc.take<uint8_t>(P * lat);
        d_rows = sc.take<int>(P * Lh + 1);
        d_cloud = sc.take<float>((size_t)cap * 3);
        d_cloud_rgb = sc.take<uint8_t>((size_t)cap * 3);
        d_cloud_map = sc.take<int>(cap);
        if (out->dc_index) d_dc = sc.take<int>(P * npx);
        perch::PoseArgs a = base;
        a.img_depth = l_depth; a.img_r = l_r; a.img_g = l_g; a.img_b = l_b;
        a.tile_cap = std::max(env_int("PERCH_IMAGE_TILE_CAP", 6144), Lw);
        if (n_poses > 0) {
            CU(perch::launch_raster_images(a, c->stream));
            ++launches;
        }
        perch::CloudArgs ca;
        ca.depth = l_depth; ca.red = l_r; ca.green = l_g; ca.blue = l_b;
        ca.pose_label = base.pose_label;
        ca.img_w = Lw; ca.img_h = Lh; ca.full_h = H; ca.num_images = n_poses; ca.stride = 1; ca.coord_scale = s;
        ca.cx = c->cam.cx; ca.cy = c->cam.cy; ca.fx = c->cam.fx; ca.fy = c->cam.fy; ca.depth_factor = c->cam.depth_factor;
        ca.row_counts = d_rows;
        ca.capacity = cap;
        ca.cloud_soa = out->cloud ? d_cloud : nullptr;
        ca.cloud_rgb = out->cloud_color ? d_cloud_rgb : nullptr;
        ca.pose_map = out->cloud_pose_map ? d_cloud_map : nullptr;
        ca.dc_index = d_dc;
        if (n_poses > 0) {
            CU(perch::launch_cloud_count(ca, c->stream));
            CU(perch::launch_cloud_scan(ca, c->stream));
            CU(perch::launch_cloud_write(ca, c->stream));
            launches += 3 + (d_dc ? 1 : 0);
        } else {
            CU(cudaMemsetAsync(d_rows, 0, 4, c->stream));
        }
    }
    // ---- costs (stage COST / DEBUG): renderer.cu:1852-1922 ------------------------------------------------------
    if (do_cost) {
        perch::PoseArgs a;
        bool color = false;
        a.pose_label = base.pose_label;
        if ((rc = prepare_cost_args(c, p, a, color))) return rc;
        a.poses = base.poses; a.pose_model_map = base.pose_model_map;
        a.pose_obs_total = base.pose_obs_total; a.n_poses = n_poses;
        a.out_costs = d_costs;
        a.out_counts = d_counts;
        if (n_poses > 0) {
            CU(perch::launch_fused_cost(a, color, c->stream));
            ++launches;
        }
    }
    CU(cudaEventRecord(c->ev1, c->stream));

    // ---- results to the host -------------------------------------------------------------------------------------
    if (do_images) {
        if (out->depth) CU(cudaMemcpyAsync(out->depth, d_depth, n_img * npx * 4, cudaMemcpyDeviceToHost, c->stream));
        if (out->red) CU(cudaMemcpyAsync(out->red, d_r, n_img * npx, cudaMemcpyDeviceToHost, c->stream));
        if (out->green) CU(cudaMemcpyAsync(out->green, d_g, n_img * npx, cudaMemcpyDeviceToHost, c->stream));
        if (out->blue) CU(cudaMemcpyAsync(out->blue, d_b, n_img * npx, cudaMemcpyDeviceToHost, c->stream));
    }
    if (do_cloud) {
        int total = 0;
        CU(cudaMemcpyAsync(&total, d_rows + P * Lh, 4, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        out->cloud_point_num = total;
        const size_t ncopy = (size_t)std::min(total, cap);
        if (ncopy > 0) {
            for (int k = 0; k < 3; ++k) {
                if (out->cloud)
                    CU(cudaMemcpyAsync(out->cloud + (size_t)k * cap, d_cloud + (size_t)k * cap, ncopy * 4,
                                       cudaMemcpyDeviceToHost, c->stream));
                if (out->cloud_color)
                    CU(cudaMemcpyAsync(out->cloud_color + (size_t)k * cap, d_cloud_rgb + (size_t)k * cap, ncopy,
                                       cudaMemcpyDeviceToHost, c->stream));
            }
            if (out->cloud_pose_map)
                CU(cudaMemcpyAsync(out->cloud_pose_map, d_cloud_map, ncopy * 4, cudaMemcpyDeviceToHost, c->stream));
        }
        if (out->dc_index && P > 0)
            CU(cudaMemcpyAsync(out->dc_index, d_dc, P * npx * 4, cudaMemcpyDeviceToHost, c->stream));
    }
    if (do_cost && P > 0) {
        if (out->rendered_cost) CU(cudaMemcpyAsync(out->rendered_cost, d_costs, P * 4, cudaMemcpyDeviceToHost, c->stream));
        if (p->calculate_observed_cost) {
            if (out->observed_cost)
                CU(cudaMemcpyAsync(out->observed_cost, d_costs + P, P * 4, cudaMemcpyDeviceToHost, c->stream));
            if (out->points_diff_cost)
                CU(cudaMemcpyAsync(out->points_diff_cost, d_costs + 2 * P, P * 4, cudaMemcpyDeviceToHost, c->stream));
        }
        if (out->counts) CU(cudaMemcpyAsync(out->counts, d_counts, P * 12, cudaMemcpyDeviceToHost, c->stream));
    }
    int flag = 0;
    CU(cudaMemcpyAsync(&flag, c->d_err_flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1);
    c->last_launches = launches;
    c->timing_pending = false;
    if (flag) {
        cudaMemsetAsync(c->d_err_flag, 0, sizeof(int), c->stream);
        if (flag & 2) return fail(PERCH_ERR_INVALID, "a state names a model_id outside [0, n_models)");
        if (flag & 4) return fail(PERCH_ERR_INVALID, "a state's segmentation_label_id has no entry in the observed labels");
        return fail(PERCH_ERR_CUDA, "internal error: a fragment fell outside its pose tile (flag %d)", flag);
    }
    return PERCH_OK;
}

extern "C" {

int perch_topk(const float* rendered, const float* observed, const int* pose_model_map, int n_poses, int pose_id_offset,
               int n_models, int k, perch_hit* out) {
    if (!rendered || !observed || !pose_model_map || !out || n_models <= 0 || k <= 0 || n_poses < 0)
        return fail(PERCH_ERR_INVALID, "bad topk argument");
    // bounded insertion per model: one pass, no allocation beyond the k-entry lists
    const perch_hit empty{INT32_MAX, -1, 0.f, 0.f};
    for (size_t j = 0; j < (size_t)n_models * k; ++j) out[j] = empty;
    auto before = [](const perch_hit& a, const perch_hit& b) { return a.cost != b.cost ? a.cost < b.cost : a.pose_id < b.pose_id; };
    for (int i = 0; i < n_poses; ++i) {
        int m = pose_model_map[i];
        if (m < 0 || m >= n_models) continue;
        float r = rendered[i], o = observed[i];
        if (!(r == r) || !(o == o) || !(std::fabs(r) < 2e9f) || !(std::fabs(o) < 2e9f)) continue;  // NaN/inf: (int) is UB
        // search_env.cpp:2021-2036: cost = -1 if (int)rendered < 0, else (int)(rendered + observed)
        if ((int)r < 0) continue;
        int cost = (int)(r + o);
        if (cost == -1 || cost == -2) continue;  // search_env.cpp:2555 "Invalid successor"
        // search_env.cpp:2560-2561: abs(target_cost - source_cost) < 30 on the truncated ints
        if (!(std::abs((int)r - (int)o) < 30)) continue;
        const perch_hit h{cost, pose_id_offset + i, r, o};
        perch_hit* list = out + (size_t)m * k;
        if (!before(h, list[k - 1])) continue;
        int pos = k - 1;
        while (pos > 0 && before(h, list[pos - 1])) { list[pos] = list[pos - 1]; --pos; }
        list[pos] = h;
    }
    return PERCH_OK;
}

}  // extern "C"
