// TEST INFRASTRUCTURE ONLY -- never linked into or called from the product library.
//
// Builds the reference's OWN CUDA kernels for this path, unmodified, from where they lie under
// /root/reference (never copied into this repo), against the POD stub headers in oracle/ref_shims/,
// into oracle/_ref/libperch_ref.so.  It is the bit-exact parity anchor for the `-m gpu` tests and the
// "reference CUDA path" timing arm.
//
//   #include'd reference sources:
//     cuda_renderer/include/cuda_renderer/cuda/image_renderer.cuh         (image_render, render_triangle_multi)
//     cuda_renderer/include/cuda_renderer/cuda/compute_point_clouds.cuh   (compute_point_clouds)
//     cuda_renderer/include/cuda_renderer/cuda/compute_costs.cuh          (compute_costs)
//
// cuda_renderer/src/cuda/renderer.cu itself cannot be built here (it includes the un-vendored fast_gicp
// headers, renderer.cu:7-9), so the ~120 lines of orchestration in render_cuda_multi_unified
// (renderer.cu:1561-1659 render+cloud, 1668-1686 label sort, 1852-1907 KNN+cost) are restated below as
// calls to the reference's own functions.  The one arithmetic step that lives in fast_gicp
// (brute_force_knn_search, k=1; call site renderer.cu:1853-1865) is restated as `ref_knn_k1` -- its
// tie-break (lowest index) and rounding (fma chain) are OUR convention: that boundary is unpinned.
#include <cstdio>
#include <unistd.h>
#include <fcntl.h>
#include <cstdint>
#include <cfloat>
#include <climits>
#include <vector>
#include <thrust/sort.h>
#include <thrust/device_vector.h>

#ifdef PERCH_REF_TREE
// Second build of the same unmodified sources (_ref/libperch_ref_tree.so) with the two switches of model.h:17-18 turned
// on, so that the USE_TREE / USE_CLUTTER branches of image_renderer.cuh (187-206, 467-482) are compiled in: model.h is
// read first (its include guard keeps image_renderer.cuh from reading it again), then the macros are redefined.
#include "cuda_renderer/model.h"
#undef USE_TREE
#undef USE_CLUTTER
#define USE_TREE 1
#define USE_CLUTTER 1
#endif
#include "cuda_renderer/cuda/image_renderer.cuh"
#include "cuda_renderer/cuda/compute_point_clouds.cuh"
#include "cuda_renderer/cuda/compute_costs.cuh"

using cuda_renderer::Model;
using cuda_renderer::gpu_stats;

namespace {

// The reference printf()s progress on every call; silence stdout around its calls so test logs stay readable.
struct quiet_stdout {
    int saved;
    quiet_stdout() {
        fflush(stdout);
        saved = dup(1);
        int nul = open("/dev/null", O_WRONLY);
        dup2(nul, 1);
        close(nul);
    }
    ~quiet_stdout() {
        fflush(stdout);
        dup2(saved, 1);
        close(saved);
    }
};

// Restated fast_gicp::brute_force_knn_search(k=1) (call site renderer.cu:1853-1865): one thread per source
// point, linear scan of the target range of that point's label (whole cloud when no labels), strict `<`
// so the lowest target index wins ties; squared L2 as fma(dz,dz,fma(dy,dy,dx*dx)).
__global__ void ref_knn_k1(const Eigen::Vector3f* src, int n_src, const Eigen::Vector3f* tgt, int n_tgt,
                           const int* src_label, const int* label_begin, const int* label_end, int n_labels,
                           float* out_sqdist, int* out_index) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_src) return;
    int b = 0, e = n_tgt;
    if (src_label != NULL) {
        int l = src_label[i];
        if (l < 0 || l >= n_labels) { b = 0; e = 0; }
        else { b = label_begin[l]; e = label_end[l]; }
    }
    float px = src[i](0), py = src[i](1), pz = src[i](2);
    float best = FLT_MAX; int best_j = -1;
    for (int j = b; j < e; ++j) {
        float dx = px - tgt[j](0), dy = py - tgt[j](1), dz = pz - tgt[j](2);
        float d = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
        if (d < best) { best = d; best_j = j; }
    }
    out_sqdist[i] = best;
    out_index[i] = best_j;
}

}  // namespace




extern "C" {

int ref_sizeof_triangle() { return (int)sizeof(Model::Triangle); }
int ref_sizeof_mat4x4() { return (int)sizeof(Model::mat4x4); }

// image_render (image_renderer.cuh:336-496) on host arrays.  pose_label may be NULL (3-DoF rule).
// Outputs: depth int32[P*W*H] (0 = empty), rgb 3 x uint8[P*W*H] (each may be NULL).
int ref_render(const void* tris64, int n_tris, const float* poses16, int n_poses, const int* pose_model_map,
               const int* tris_model_count, int n_models, int width, int height, const float* proj16,
               const int32_t* source_depth, const uint8_t* source_r, const uint8_t* source_g,
               const uint8_t* source_b, const uint8_t* source_label, const int* pose_label,
               float occlusion_threshold, int single_result_image, int32_t* out_depth, uint8_t* out_r,
               uint8_t* out_g, uint8_t* out_b) {
    quiet_stdout q;
    const Model::Triangle* t = (const Model::Triangle*)tris64;
    const Model::mat4x4* p = (const Model::mat4x4*)poses16;
    size_t npx = (size_t)width * height;
    thrust::device_vector<Model::Triangle> d_tris(t, t + n_tris);
    thrust::device_vector<Model::mat4x4> d_poses(p, p + n_poses);
    thrust::device_vector<int> d_pmm(pose_model_map, pose_model_map + n_poses);
    thrust::device_vector<int> d_tmc(tris_model_count, tris_model_count + n_models);
    thrust::device_vector<int32_t> d_sd(source_depth, source_depth + npx);
    thrust::device_vector<uint8_t> d_sr(source_r, source_r + npx), d_sg(source_g, source_g + npx),
        d_sb(source_b, source_b + npx);
    thrust::device_vector<uint8_t> d_sl;
    if (source_label) d_sl.assign(source_label, source_label + npx);
    thrust::device_vector<int> d_pl;
    if (pose_label) d_pl.assign(pose_label, pose_label + n_poses);
    thrust::device_vector<int> occ, occ_other;
    thrust::device_vector<float> clutter, total;
    thrust::device_vector<int32_t> d_depth;
    thrust::device_vector<uint8_t> d_r, d_g, d_b;
    Model::mat4x4 proj;
    proj.init_from_ptr(proj16);
    gpu_stats stats{0, 0};
    cuda_renderer::image_render(d_tris, d_poses, d_pmm, d_tmc, d_sd, d_sr, d_sg, d_sb, d_sl, d_pl, n_poses, width,
                                height, proj, occlusion_threshold, single_result_image, occ, occ_other, clutter,
                                total, d_depth, d_r, d_g, d_b, stats);
    if (cudaDeviceSynchronize() != cudaSuccess) return 1;
    // in single-image mode the reference still allocates P images but renders everything into the first
    size_t n = npx * (single_result_image ? 1 : n_poses);
    if (out_depth) cudaMemcpy(out_depth, thrust::raw_pointer_cast(d_depth.data()), n * 4, cudaMemcpyDeviceToHost);
    if (out_r) cudaMemcpy(out_r, thrust::raw_pointer_cast(d_r.data()), n, cudaMemcpyDeviceToHost);
    if (out_g) cudaMemcpy(out_g, thrust::raw_pointer_cast(d_g.data()), n, cudaMemcpyDeviceToHost);
    if (out_b) cudaMemcpy(out_b, thrust::raw_pointer_cast(d_b.data()), n, cudaMemcpyDeviceToHost);
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}

int ref_tree_enabled() { return USE_TREE && USE_CLUTTER; }

// image_render with its four per-pose vectors handed back: pose_occluded, pose_occluded_other, the clutter vector
// (100 * clutter / total after image_renderer.cuh:467-482 when USE_CLUTTER) and the fragment count.
int ref_render_tree(const void* tris64, int n_tris, const float* poses16, int n_poses, const int* pose_model_map,
                    const int* tris_model_count, int n_models, int width, int height, const float* proj16,
                    const int32_t* source_depth, const uint8_t* source_label, const int* pose_label,
                    float occlusion_threshold, int32_t* out_depth, int* out_occluded, int* out_occluded_other,
                    float* out_clutter, float* out_total) {
    quiet_stdout q;
    const Model::Triangle* t = (const Model::Triangle*)tris64;
    const Model::mat4x4* p = (const Model::mat4x4*)poses16;
    size_t npx = (size_t)width * height;
    thrust::device_vector<Model::Triangle> d_tris(t, t + n_tris);
    thrust::device_vector<Model::mat4x4> d_poses(p, p + n_poses);
    thrust::device_vector<int> d_pmm(pose_model_map, pose_model_map + n_poses);
    thrust::device_vector<int> d_tmc(tris_model_count, tris_model_count + n_models);
    thrust::device_vector<int32_t> d_sd(source_depth, source_depth + npx);
    thrust::device_vector<uint8_t> d_sc(npx, 0);
    thrust::device_vector<uint8_t> d_sl;
    if (source_label) d_sl.assign(source_label, source_label + npx);
    thrust::device_vector<int> d_pl;
    if (pose_label) d_pl.assign(pose_label, pose_label + n_poses);
    thrust::device_vector<int> occ, occ_other;
    thrust::device_vector<float> clutter, total;
    thrust::device_vector<int32_t> d_depth;
    thrust::device_vector<uint8_t> d_r, d_g, d_b;
    Model::mat4x4 proj;
    proj.init_from_ptr(proj16);
    gpu_stats stats{0, 0};
    cuda_renderer::image_render(d_tris, d_poses, d_pmm, d_tmc, d_sd, d_sc, d_sc, d_sc, d_sl, d_pl, n_poses, width, height,
                                proj, occlusion_threshold, 0, occ, occ_other, clutter, total, d_depth, d_r, d_g, d_b, stats);
    if (cudaDeviceSynchronize() != cudaSuccess) return 1;
    if (out_depth) cudaMemcpy(out_depth, thrust::raw_pointer_cast(d_depth.data()), npx * n_poses * 4, cudaMemcpyDeviceToHost);
    thrust::copy(occ.begin(), occ.end(), out_occluded);
    thrust::copy(occ_other.begin(), occ_other.end(), out_occluded_other);
    thrust::copy(clutter.begin(), clutter.end(), out_clutter);
    thrust::copy(total.begin(), total.end(), out_total);
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}

// compute_point_clouds (compute_point_clouds.cuh:188-367) as driven by depth2cloud_global (renderer.cu:1936-2070).
// Two-call protocol: pass out_* = NULL to get the point count, then call again with buffers of that size.
// cam_tf_colmajor16 / bounds6 / label_mask may be NULL.  out_xyz_soa is the 3 x N SoA, out_xyz_aos the
// Eigen::Vector3f AoS; out_rgb the 3 x N SoA.
int ref_depth2cloud(const int32_t* depth, const uint8_t* r, const uint8_t* g, const uint8_t* b, int width, int height,
                    int num_poses, float cx, float cy, float fx, float fy, float depth_factor, int stride,
                    const uint8_t* label_mask, const double* bounds6, const float* cam_tf_colmajor16,
                    int* out_count, float* out_xyz_soa, float* out_xyz_aos, uint8_t* out_rgb, int* out_pose_map,
                    int* out_label, int* out_dc_index) {
    quiet_stdout q;
    size_t n = (size_t)width * height * num_poses;
    thrust::device_vector<int32_t> d_depth(depth, depth + n);
    thrust::device_vector<uint8_t> d_r(r, r + n), d_g(g, g + n), d_b(b, b + n);
    thrust::device_vector<int> d_occ(num_poses, 0);
    thrust::device_vector<uint8_t> d_label;
    if (label_mask) d_label.assign(label_mask, label_mask + n);
    thrust::device_vector<double> d_bounds;
    if (bounds6) d_bounds.assign(bounds6, bounds6 + 6);
    Eigen::Matrix4f* d_tf = NULL;
    if (cam_tf_colmajor16) {
        cudaMalloc(&d_tf, sizeof(Eigen::Matrix4f));
        cudaMemcpy(d_tf, cam_tf_colmajor16, sizeof(Eigen::Matrix4f), cudaMemcpyHostToDevice);
    }
    float* cloud2d = NULL; size_t pitch = 0;
    thrust::device_vector<Eigen::Vector3f> c_eigen;
    thrust::device_vector<float> c_flat;
    thrust::device_vector<uint8_t> c_rgb;
    thrust::device_vector<int> dc_index, pose_map, c_label;
    int count = 0;
    gpu_stats stats{0, 0};
    cuda_renderer::compute_point_clouds(d_depth, d_r, d_g, d_b, num_poses, width, height, cx, cy, fx, fy,
                                        depth_factor, stride, d_occ, cloud2d, pitch, c_eigen, c_flat, c_rgb, count,
                                        dc_index, pose_map, c_label, stats, d_tf, d_label, d_bounds);
    if (cudaDeviceSynchronize() != cudaSuccess) return 1;
    *out_count = count;
    if (out_xyz_soa)
        cudaMemcpy2D(out_xyz_soa, count * sizeof(float), cloud2d, pitch, count * sizeof(float), 3,
                     cudaMemcpyDeviceToHost);
    if (out_xyz_aos)
        cudaMemcpy(out_xyz_aos, thrust::raw_pointer_cast(c_eigen.data()), (size_t)count * 12, cudaMemcpyDeviceToHost);
    if (out_rgb) cudaMemcpy(out_rgb, thrust::raw_pointer_cast(c_rgb.data()), (size_t)count * 3, cudaMemcpyDeviceToHost);
    if (out_pose_map)
        cudaMemcpy(out_pose_map, thrust::raw_pointer_cast(pose_map.data()), (size_t)count * 4, cudaMemcpyDeviceToHost);
    if (out_label && label_mask)
        cudaMemcpy(out_label, thrust::raw_pointer_cast(c_label.data()), (size_t)count * 4, cudaMemcpyDeviceToHost);
    if (out_dc_index)
        cudaMemcpy(out_dc_index, thrust::raw_pointer_cast(dc_index.data()), n * 4, cudaMemcpyDeviceToHost);
    if (d_tf) cudaFree(d_tf);
    cudaFree(cloud2d);
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}

// The "COST" stage of render_cuda_multi_unified (renderer.cu:1431-1934) without ICP, restated as calls into the
// reference's image_render / compute_point_clouds / compute_costs, including its per-call uploads and
// allocations.  obs_xyz_aos: N_obs x 3 floats; obs_rgb: 3 x N_obs SoA; obs_label may be NULL (3-DoF).
// Outputs (each P floats, may be NULL): rendered_cost, observed_cost, points_diff_cost; plus optional per-pose
// integer-valued counters recovered before the percentage math is not possible without touching the reference,
// so tests recompute counts from the rendered cloud instead (out_n_rend = total rendered points).
// ms_out[4]: render, cloud, knn, cost wall-clock milliseconds (device-synchronised), like the reference's printf timers.
int ref_render_cost(const void* tris64, int n_tris, const float* poses16, int n_poses, const int* pose_model_map,
                    const int* tris_model_count, int n_models, int width, int height, const float* proj16,
                    const int32_t* source_depth, const uint8_t* source_r, const uint8_t* source_g,
                    const uint8_t* source_b, const uint8_t* source_label, const int* pose_label, int stride,
                    int depth_factor, float cx, float cy, float fx, float fy, const float* obs_xyz_aos,
                    const uint8_t* obs_rgb, const int* obs_label, int n_obs, const float* pose_obs_total,
                    int cost_type, int calc_observed, float sensor_resolution, float color_distance_threshold,
                    float occlusion_threshold, float* rendered_cost, float* observed_cost, float* points_diff_cost,
                    int* out_n_rend, float* out_cloud_soa, int* out_pose_map, int cloud_capacity, float* out_nn_sqdist,
                    int* out_nn_index, float* ms_out) {
    quiet_stdout q;
    cudaEvent_t ev[5];
    for (auto& e : ev) cudaEventCreate(&e);
    cudaEventRecord(ev[0]);
    const Model::Triangle* t = (const Model::Triangle*)tris64;
    const Model::mat4x4* p = (const Model::mat4x4*)poses16;
    size_t npx = (size_t)width * height;
    // per-call uploads, as renderer.cu:1532-1544
    thrust::device_vector<Model::Triangle> d_tris(t, t + n_tris);
    thrust::device_vector<Model::mat4x4> d_poses(p, p + n_poses);
    thrust::device_vector<int> d_pmm(pose_model_map, pose_model_map + n_poses);
    thrust::device_vector<int> d_tmc(tris_model_count, tris_model_count + n_models);
    thrust::device_vector<int> d_pl;
    if (pose_label) d_pl.assign(pose_label, pose_label + n_poses);
    thrust::device_vector<int32_t> d_sd(source_depth, source_depth + npx);
    thrust::device_vector<uint8_t> d_sr(source_r, source_r + npx), d_sg(source_g, source_g + npx),
        d_sb(source_b, source_b + npx);
    thrust::device_vector<uint8_t> d_sl;
    if (source_label) d_sl.assign(source_label, source_label + npx);
    thrust::device_vector<int> occ, occ_other;
    thrust::device_vector<float> clutter, total;
    thrust::device_vector<int32_t> d_depth;
    thrust::device_vector<uint8_t> d_r, d_g, d_b;
    Model::mat4x4 proj;
    proj.init_from_ptr(proj16);
    gpu_stats stats{0, 0};
    cuda_renderer::image_render(d_tris, d_poses, d_pmm, d_tmc, d_sd, d_sr, d_sg, d_sb, d_sl, d_pl, n_poses, width,
                                height, proj, occlusion_threshold, 0, occ, occ_other, clutter, total, d_depth, d_r, d_g,
                                d_b, stats);
    cudaEventRecord(ev[1]);
    // renderer.cu:1621-1659
    float* cloud2d = NULL; size_t pitch = 0;
    thrust::device_vector<Eigen::Vector3f> c_eigen;
    thrust::device_vector<float> c_flat;
    thrust::device_vector<uint8_t> c_rgb;
    thrust::device_vector<int> dc_index, pose_map, c_label;
    int n_rend = 0;
    cuda_renderer::compute_point_clouds(d_depth, d_r, d_g, d_b, n_poses, width, height, cx, cy, fx, fy,
                                        (float)depth_factor, stride, occ, cloud2d, pitch, c_eigen, c_flat, c_rgb,
                                        n_rend, dc_index, pose_map, c_label, stats, NULL,
                                        thrust::device_vector<uint8_t>(0), thrust::device_vector<double>(0), d_pl);
    cudaEventRecord(ev[2]);
    // renderer.cu:1668-1686: observed cloud upload, 6-DoF label sort + per-label ranges
    const Eigen::Vector3f* oe = (const Eigen::Vector3f*)obs_xyz_aos;
    thrust::device_vector<Eigen::Vector3f> d_obs(oe, oe + n_obs);
    thrust::device_vector<int> d_obs_label, d_lb, d_le;
    int n_labels = 0;
    if (pose_label) {
        d_obs_label.assign(obs_label, obs_label + n_obs);
        thrust::stable_sort_by_key(d_obs_label.begin(), d_obs_label.end(), d_obs.begin());
        std::vector<int> hl(n_obs);
        thrust::copy(d_obs_label.begin(), d_obs_label.end(), hl.begin());
        for (int i = 0; i < n_obs; ++i) n_labels = hl[i] + 1 > n_labels ? hl[i] + 1 : n_labels;
        std::vector<int> lb(n_labels, 0), le(n_labels, 0);
        for (int i = 0; i < n_obs; ++i) {
            if (hl[i] < 0) continue;
            if (i == 0 || hl[i - 1] != hl[i]) lb[hl[i]] = i;
            le[hl[i]] = i + 1;
        }
        d_lb = lb; d_le = le;
    }
    // renderer.cu:1852-1877
    thrust::device_vector<float> k_dist(n_rend);
    thrust::device_vector<int> k_idx(n_rend);
    if (n_rend > 0)
        ref_knn_k1<<<(n_rend + 255) / 256, 256>>>(
            thrust::raw_pointer_cast(c_eigen.data()), n_rend, thrust::raw_pointer_cast(d_obs.data()), n_obs,
            pose_label ? thrust::raw_pointer_cast(c_label.data()) : NULL,
            pose_label ? thrust::raw_pointer_cast(d_lb.data()) : NULL,
            pose_label ? thrust::raw_pointer_cast(d_le.data()) : NULL, n_labels,
            thrust::raw_pointer_cast(k_dist.data()), thrust::raw_pointer_cast(k_idx.data()));
    cudaEventRecord(ev[3]);
    sensor_resolution = sensor_resolution * sensor_resolution;
    // renderer.cu:1881-1907
    thrust::device_vector<uint8_t> d_obs_rgb(obs_rgb, obs_rgb + 3 * (size_t)n_obs);
    thrust::device_vector<float> d_tot(pose_obs_total, pose_obs_total + n_poses);
    thrust::device_vector<float> v_rend, v_obs, v_diff;
    cuda_renderer::compute_costs(n_poses, cost_type, calc_observed != 0, sensor_resolution, color_distance_threshold,
                                 d_obs_rgb, d_obs_label, n_obs, c_rgb, pose_map, occ, d_pl, d_tot, n_rend, k_dist,
                                 k_idx, v_rend, v_obs, v_diff, stats);
    if (rendered_cost)
        cudaMemcpy(rendered_cost, thrust::raw_pointer_cast(v_rend.data()), n_poses * 4, cudaMemcpyDeviceToHost);
    if (calc_observed && observed_cost)
        cudaMemcpy(observed_cost, thrust::raw_pointer_cast(v_obs.data()), n_poses * 4, cudaMemcpyDeviceToHost);
    if (calc_observed && points_diff_cost)
        cudaMemcpy(points_diff_cost, thrust::raw_pointer_cast(v_diff.data()), n_poses * 4, cudaMemcpyDeviceToHost);
    cudaEventRecord(ev[4]);
    if (cudaDeviceSynchronize() != cudaSuccess) return 1;
    if (out_n_rend) *out_n_rend = n_rend;
    int ncopy = n_rend < cloud_capacity ? n_rend : cloud_capacity;
    if (out_cloud_soa && ncopy > 0)
        cudaMemcpy2D(out_cloud_soa, (size_t)cloud_capacity * 4, cloud2d, pitch, (size_t)ncopy * 4, 3,
                     cudaMemcpyDeviceToHost);
    if (out_pose_map && ncopy > 0)
        cudaMemcpy(out_pose_map, thrust::raw_pointer_cast(pose_map.data()), (size_t)ncopy * 4, cudaMemcpyDeviceToHost);
    if (out_nn_sqdist && ncopy > 0)
        cudaMemcpy(out_nn_sqdist, thrust::raw_pointer_cast(k_dist.data()), (size_t)ncopy * 4, cudaMemcpyDeviceToHost);
    if (out_nn_index && ncopy > 0)
        cudaMemcpy(out_nn_index, thrust::raw_pointer_cast(k_idx.data()), (size_t)ncopy * 4, cudaMemcpyDeviceToHost);
    if (ms_out)
        for (int i = 0; i < 4; ++i) cudaEventElapsedTime(&ms_out[i], ev[i], ev[i + 1]);
    for (auto& e : ev) cudaEventDestroy(e);
    cudaFree(cloud2d);
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}

// compute_point_clouds + label sort + k=1 search + compute_costs of the reference (renderer.cu:1617-1659,1668-1686,
// 1852-1907) on GIVEN rendered images.  The reference rasterizer is not deterministic on sm_70+, so
// the cloud/cost kernels are pinned separately from it: feed them canonical images and every output is reproducible.
int ref_cost_from_images(const int32_t* depth, const uint8_t* r, const uint8_t* g, const uint8_t* b, int n_poses, int width,
                         int height, const int* pose_label, int stride, int depth_factor, float cx, float cy, float fx,
                         float fy, const float* obs_xyz_aos, const uint8_t* obs_rgb, const int* obs_label, int n_obs,
                         const float* pose_obs_total, int cost_type, int calc_observed, float sensor_resolution,
                         float color_distance_threshold, float* rendered_cost, float* observed_cost,
                         float* points_diff_cost, int* out_n_rend) {
    quiet_stdout q;
    size_t n = (size_t)width * height * n_poses;
    thrust::device_vector<int32_t> d_depth(depth, depth + n);
    thrust::device_vector<uint8_t> d_r(r, r + n), d_g(g, g + n), d_b(b, b + n);
    thrust::device_vector<int> occ(n_poses, 0);
    thrust::device_vector<int> d_pl;
    if (pose_label) d_pl.assign(pose_label, pose_label + n_poses);
    gpu_stats stats{0, 0};
    float* cloud2d = NULL; size_t pitch = 0;
    thrust::device_vector<Eigen::Vector3f> c_eigen;
    thrust::device_vector<float> c_flat;
    thrust::device_vector<uint8_t> c_rgb;
    thrust::device_vector<int> dc_index, pose_map, c_label;
    int n_rend = 0;
    cuda_renderer::compute_point_clouds(d_depth, d_r, d_g, d_b, n_poses, width, height, cx, cy, fx, fy,
                                        (float)depth_factor, stride, occ, cloud2d, pitch, c_eigen, c_flat, c_rgb,
                                        n_rend, dc_index, pose_map, c_label, stats, NULL,
                                        thrust::device_vector<uint8_t>(0), thrust::device_vector<double>(0), d_pl);
    const Eigen::Vector3f* oe = (const Eigen::Vector3f*)obs_xyz_aos;
    thrust::device_vector<Eigen::Vector3f> d_obs(oe, oe + n_obs);
    thrust::device_vector<int> d_obs_label, d_lb, d_le;
    int n_labels = 0;
    if (pose_label) {
        d_obs_label.assign(obs_label, obs_label + n_obs);
        thrust::stable_sort_by_key(d_obs_label.begin(), d_obs_label.end(), d_obs.begin());
        std::vector<int> hl(n_obs);
        thrust::copy(d_obs_label.begin(), d_obs_label.end(), hl.begin());
        for (int i = 0; i < n_obs; ++i) n_labels = hl[i] + 1 > n_labels ? hl[i] + 1 : n_labels;
        std::vector<int> lb(n_labels, 0), le(n_labels, 0);
        for (int i = 0; i < n_obs; ++i) {
            if (hl[i] < 0) continue;
            if (i == 0 || hl[i - 1] != hl[i]) lb[hl[i]] = i;
            le[hl[i]] = i + 1;
        }
        d_lb = lb; d_le = le;
    }
    thrust::device_vector<float> k_dist(n_rend);
    thrust::device_vector<int> k_idx(n_rend);
    if (n_rend > 0)
        ref_knn_k1<<<(n_rend + 255) / 256, 256>>>(
            thrust::raw_pointer_cast(c_eigen.data()), n_rend, thrust::raw_pointer_cast(d_obs.data()), n_obs,
            pose_label ? thrust::raw_pointer_cast(c_label.data()) : NULL,
            pose_label ? thrust::raw_pointer_cast(d_lb.data()) : NULL,
            pose_label ? thrust::raw_pointer_cast(d_le.data()) : NULL, n_labels,
            thrust::raw_pointer_cast(k_dist.data()), thrust::raw_pointer_cast(k_idx.data()));
    sensor_resolution = sensor_resolution * sensor_resolution;
    thrust::device_vector<uint8_t> d_obs_rgb(obs_rgb, obs_rgb + 3 * (size_t)n_obs);
    thrust::device_vector<float> d_tot(pose_obs_total, pose_obs_total + n_poses);
    thrust::device_vector<float> v_rend, v_obs, v_diff;
    cuda_renderer::compute_costs(n_poses, cost_type, calc_observed != 0, sensor_resolution, color_distance_threshold,
                                 d_obs_rgb, d_obs_label, n_obs, c_rgb, pose_map, occ, d_pl, d_tot, n_rend, k_dist,
                                 k_idx, v_rend, v_obs, v_diff, stats);
    if (cudaDeviceSynchronize() != cudaSuccess) return 1;
    if (rendered_cost)
        cudaMemcpy(rendered_cost, thrust::raw_pointer_cast(v_rend.data()), n_poses * 4, cudaMemcpyDeviceToHost);
    if (calc_observed && observed_cost)
        cudaMemcpy(observed_cost, thrust::raw_pointer_cast(v_obs.data()), n_poses * 4, cudaMemcpyDeviceToHost);
    if (calc_observed && points_diff_cost)
        cudaMemcpy(points_diff_cost, thrust::raw_pointer_cast(v_diff.data()), n_poses * 4, cudaMemcpyDeviceToHost);
    if (out_n_rend) *out_n_rend = n_rend;
    cudaFree(cloud2d);
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}

}  // extern "C"
