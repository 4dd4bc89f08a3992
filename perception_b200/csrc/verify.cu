// perch_render_cost_unfused: the COST stage computed the way the reference structures it -- render every pixel, build the
// clouds, brute-force nearest neighbour, count -- out of this library's OTHER kernels, none of which the fused per-pose
// kernel (pose_kernels.cu) shares its culling or its lattice rasterisation with:
//   raster_full.cu   every triangle of every pose through the exact path into full 640x480 images (no culling at all)
//   cloud.cu         stride sampling + ordered compaction (compute_point_clouds.cuh:37-184)
//   brute_cost_kernel (here)  for every rendered point a strict-`<` linear scan over the observed points of its label
//                    (what fast_gicp::brute_force_knn_search does, renderer.cu:1853-1865), then compute_costs' counters
//                    (compute_costs.cuh:161-290)
// It exists so that EVERY pose of a full-size batch (16 384 x 20 000 triangles) can be cross-checked on the device in a
// few hundred milliseconds (tests/test_parity_gpu.py, bench.py --verify), and as the measured "what fusion buys" baseline.
// cost_type 1 (colour) is not offered here.
#include <cuda_runtime.h>

#include <algorithm>
#include <vector>

#include "capi_common.hpp"
#include "context.hpp"
#include "kernels.cuh"
#include "perch_b200.h"

namespace perch {
namespace {

struct BruteArgs {
    const float* cloud_aos;   // rendered points of this chunk, n x 3
    const int* pose_map;      // chunk-local pose index per point
    int n_points;
    const float4* obs;        // label-sorted observed cloud
    int n_obs;
    const int* label_range;   // [begin, end) per label, or null (3-DoF: whole cloud)
    int n_labels;
    const int* pose_label;    // chunk-local, or null
    float res2;
    int* n;                   // per chunk pose: rendered points
    int* c;                   // unexplained rendered points
    uint32_t* mask;           // chunk poses x mask_words explained bits
    int mask_words;
};

__global__ void __launch_bounds__(256) brute_cost_kernel(BruteArgs a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n_points) return;
    const int p = a.pose_map[i];
    const float qx = a.cloud_aos[3 * (size_t)i], qy = a.cloud_aos[3 * (size_t)i + 1], qz = a.cloud_aos[3 * (size_t)i + 2];
    int b = 0, e = a.n_obs;
    if (a.pose_label) {
        const int l = a.pose_label[p];
        if (l >= 0 && l < a.n_labels) { b = a.label_range[2 * l]; e = a.label_range[2 * l + 1]; }
        else { b = 0; e = 0; }
    }
    float best = FLT_MAX;
    int ref = -1;
    for (int j = b; j < e; ++j) {
        const float4 o = __ldg(a.obs + j);
        const float d = sqdist3(qx, qy, qz, o.x, o.y, o.z);
        if (d < best) { best = d; ref = j; }
    }
    atomicAdd(&a.n[p], 1);                                         // compute_costs.cuh:199
    if (best > a.res2) atomicAdd(&a.c[p], 1);                      // :206-209
    else atomicOr(&a.mask[(size_t)p * a.mask_words + (ref >> 5)], 1u << (ref & 31));  // :238,244,268
}

struct FinishArgs {
    const int* n; const int* c; const uint32_t* mask; int mask_words;
    const float* pose_obs_total;  // chunk-local or null
    int calc_observed, n_chunk, P, p0;
    float* out_costs;   // 3 x P
    int32_t* out_counts;  // 3 x P
};

__global__ void finish_costs_kernel(FinishArgs a) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= a.n_chunk) return;
    int ei = 0;
    for (int w = 0; w < a.mask_words; ++w) ei += __popc(a.mask[(size_t)q * a.mask_words + w]);
    const int p = a.p0 + q, P = a.P;
    const float n = (float)a.n[q], c = (float)a.c[q], e = (float)ei;
    // compute_costs.cuh:364-378, 407-446 -- the same float sequence as the fused kernel's epilogue
    const float rendered_explained = __fsub_rn(n, c);
    float rc = (n == 0.f) ? -1.f : __fdiv_rn(c, n);
    rc = (rc == -1.f) ? -1.f : __fmul_rn(rc, 100.f);
    a.out_costs[p] = rc;
    if (a.calc_observed) {
        const float tot = a.pose_obs_total ? a.pose_obs_total[q] : 0.f;
        a.out_costs[P + p] = __fmul_rn(__fdiv_rn(__fsub_rn(tot, e), tot), 100.f);
        a.out_costs[2 * P + p] = __fsub_rn(rendered_explained, e);
    }
    a.out_counts[p] = a.n[q];
    a.out_counts[P + p] = a.c[q];
    a.out_counts[2 * P + p] = ei;
}

struct DevBufs {  // frees everything it handed out
    std::vector<void*> ptrs;
    ~DevBufs() { for (void* p : ptrs) cudaFree(p); }
    template <class T> cudaError_t get(T** out, size_t n) {
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T));
        if (e == cudaSuccess) ptrs.push_back(p);
        *out = static_cast<T*>(p);
        return e;
    }
};

}  // namespace
}  // namespace perch

extern "C" int perch_render_cost_unfused(perch_ctx* c, const float* poses16, int n_poses, const int* pose_model_map,
                                         const int* pose_label, const float* pose_obs_total, const perch_params* p,
                                         perch_outputs* out) {
    using perch_detail::fail;
    using namespace perch;
    if (!c || !p || !out) return fail(PERCH_ERR_INVALID, "null argument");
    if (!c->have_cam || !c->have_models || !c->have_scene || !c->have_obs)
        return fail(PERCH_ERR_INVALID, "camera, models, scene and observed cloud must be set first");
    if (n_poses < 0 || (n_poses > 0 && (!poses16 || !pose_model_map))) return fail(PERCH_ERR_INVALID, "bad pose arguments");
    if (p->cost_type != 0 && p->cost_type != 2) return fail(PERCH_ERR_INVALID, "the unfused check path offers cost_type 0 and 2 only");
    const int W = c->cam.width, H = c->cam.height, s = p->stride;
    if (s <= 0 || W % s != 0) return fail(PERCH_ERR_INVALID, "stride %d must divide the image width %d", s, W);
    if (pose_label && c->n_obs > 0 && !c->obs_has_label)
        return fail(PERCH_ERR_INVALID, "pose_segmentation_label given but the observed cloud was set without labels");
    if (n_poses == 0) return PERCH_OK;
    perch_detail::DeviceGuard g(c->device);
    {
        const int rc = perch_detail::ensure_obs_host(c);
        if (rc) return rc;
    }
    const size_t P = (size_t)n_poses, npx = (size_t)W * H;
    const int Lw = W / s, Lh = (H + s - 1) / s;
    const int n_obs = c->n_obs, words = std::max((n_obs + 31) / 32, 1);
    const int chunk = (int)std::min<size_t>(P, 512);
    const size_t cap = (size_t)chunk * Lw * Lh;
    DevBufs bufs;
    float *d_poses, *d_tot, *d_costs, *d_cloud;
    int *d_pmm, *d_lab, *d_rows, *d_map, *d_n, *d_c, *d_range;
    int32_t *d_counts, *d_depth;
    uint32_t* d_mask;
    float4* d_obs;
    PERCH_CU(bufs.get(&d_poses, P * 16));
    PERCH_CU(bufs.get(&d_pmm, P));
    PERCH_CU(bufs.get(&d_lab, P));
    PERCH_CU(bufs.get(&d_tot, P));
    PERCH_CU(bufs.get(&d_costs, 3 * P));
    PERCH_CU(bufs.get(&d_counts, 3 * P));
    PERCH_CU(bufs.get(&d_depth, (size_t)chunk * npx));
    PERCH_CU(bufs.get(&d_rows, (size_t)chunk * Lh + 1));
    PERCH_CU(bufs.get(&d_cloud, cap * 3));
    PERCH_CU(bufs.get(&d_map, cap));
    PERCH_CU(bufs.get(&d_n, (size_t)chunk));
    PERCH_CU(bufs.get(&d_c, (size_t)chunk));
    PERCH_CU(bufs.get(&d_mask, (size_t)chunk * words));
    PERCH_CU(bufs.get(&d_obs, (size_t)n_obs));
    // label-sorted observed cloud + per-label ranges (renderer.cu:1668-1686), from the context's host copies
    std::vector<float4> obs4((size_t)std::max(n_obs, 1));
    for (int i = 0; i < n_obs; ++i)
        obs4[i] = make_float4(c->h_obs_xyz[3 * (size_t)i], c->h_obs_xyz[3 * (size_t)i + 1], c->h_obs_xyz[3 * (size_t)i + 2], 0.f);
    int n_labels = 0;
    for (int l : c->h_obs_label) n_labels = std::max(n_labels, l + 1);
    std::vector<int> range(2 * (size_t)std::max(n_labels, 1), 0);
    for (int i = 0; i < (int)c->h_obs_label.size(); ++i) {
        const int l = c->h_obs_label[i];
        if (l < 0) continue;
        if (range[2 * l + 1] == 0) range[2 * l] = i;
        range[2 * l + 1] = i + 1;
    }
    PERCH_CU(bufs.get(&d_range, range.size()));
    cudaStream_t st = c->stream;
    PERCH_CU(cudaMemcpyAsync(d_obs, obs4.data(), (size_t)n_obs * sizeof(float4), cudaMemcpyHostToDevice, st));
    PERCH_CU(cudaMemcpyAsync(d_range, range.data(), range.size() * 4, cudaMemcpyHostToDevice, st));
    PERCH_CU(cudaMemcpyAsync(d_poses, poses16, P * 64, cudaMemcpyHostToDevice, st));
    PERCH_CU(cudaMemcpyAsync(d_pmm, pose_model_map, P * 4, cudaMemcpyHostToDevice, st));
    if (pose_label) PERCH_CU(cudaMemcpyAsync(d_lab, pose_label, P * 4, cudaMemcpyHostToDevice, st));
    if (pose_obs_total) PERCH_CU(cudaMemcpyAsync(d_tot, pose_obs_total, P * 4, cudaMemcpyHostToDevice, st));
    PERCH_CU(cudaEventRecord(c->ev0, st));
    int launches = 0;
    for (size_t p0 = 0; p0 < P; p0 += (size_t)chunk) {
        const int n = (int)std::min<size_t>((size_t)chunk, P - p0);
        FullJob j;
        j.poses = reinterpret_cast<const Mat4*>(d_poses) + p0;
        j.pose_model_map = d_pmm + p0;
        j.pose_label = pose_label ? d_lab + p0 : nullptr;
        j.n_poses = n;
        j.occlusion_threshold = p->occlusion_threshold;
        j.img_depth = d_depth;
        PERCH_CU(full_raster_job(c, j, &launches));
        CloudArgs ca;
        ca.depth = d_depth;
        ca.img_w = W; ca.img_h = H; ca.full_h = H; ca.num_images = n; ca.stride = s; ca.coord_scale = 1;
        ca.cx = c->cam.cx; ca.cy = c->cam.cy; ca.fx = c->cam.fx; ca.fy = c->cam.fy; ca.depth_factor = c->cam.depth_factor;
        ca.row_counts = d_rows;
        ca.capacity = (int)cap;
        ca.cloud_aos = d_cloud;
        ca.pose_map = d_map;
        PERCH_CU(launch_cloud_count(ca, st));
        PERCH_CU(launch_cloud_scan(ca, st));
        PERCH_CU(launch_cloud_write(ca, st));
        int total = 0;
        PERCH_CU(cudaMemcpyAsync(&total, d_rows + (size_t)n * Lh, 4, cudaMemcpyDeviceToHost, st));
        PERCH_CU(cudaMemsetAsync(d_n, 0, (size_t)n * 4, st));
        PERCH_CU(cudaMemsetAsync(d_c, 0, (size_t)n * 4, st));
        PERCH_CU(cudaMemsetAsync(d_mask, 0, (size_t)n * words * 4, st));
        PERCH_CU(cudaStreamSynchronize(st));
        if (total > 0) {
            BruteArgs b{d_cloud, d_map, total, d_obs, n_obs, pose_label ? d_range : nullptr, n_labels,
                        pose_label ? d_lab + p0 : nullptr, p->sensor_resolution * p->sensor_resolution, d_n, d_c, d_mask, words};
            brute_cost_kernel<<<(total + 255) / 256, 256, 0, st>>>(b);
            PERCH_CU(cudaGetLastError());
        }
        FinishArgs f{d_n, d_c, d_mask, words, pose_obs_total ? d_tot + p0 : nullptr, p->calculate_observed_cost, n, (int)P, (int)p0,
                     d_costs, d_counts};
        finish_costs_kernel<<<(n + 127) / 128, 128, 0, st>>>(f);
        PERCH_CU(cudaGetLastError());
        launches += 5;
    }
    PERCH_CU(cudaEventRecord(c->ev1, st));
    if (out->rendered_cost) PERCH_CU(cudaMemcpyAsync(out->rendered_cost, d_costs, P * 4, cudaMemcpyDeviceToHost, st));
    if (p->calculate_observed_cost) {
        if (out->observed_cost) PERCH_CU(cudaMemcpyAsync(out->observed_cost, d_costs + P, P * 4, cudaMemcpyDeviceToHost, st));
        if (out->points_diff_cost) PERCH_CU(cudaMemcpyAsync(out->points_diff_cost, d_costs + 2 * P, P * 4, cudaMemcpyDeviceToHost, st));
    }
    if (out->counts) PERCH_CU(cudaMemcpyAsync(out->counts, d_counts, P * 12, cudaMemcpyDeviceToHost, st));
    PERCH_CU(cudaStreamSynchronize(st));
    cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1);
    c->last_launches = launches;
    c->timing_pending = false;
    return PERCH_OK;
}
