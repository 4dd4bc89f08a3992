// TEST INFRASTRUCTURE.  extern "C" trampolines that call the C++ drop-in backend
// (perception_b200/csrc/compat/cuda_renderer_compat.cpp) through the REFERENCE's own declarations
// (<cuda_renderer/renderer.h>: std::vector / Eigen / cv::Mat arguments), so that tests/test_compat_gpu.py can compare
// what a reference caller would receive with what the C-ABI returns directly.
#include "cuda_renderer/renderer.h"

#include <cstdlib>
#include <cstring>
#include <exception>
#include <string>
#include <vector>

static std::string g_err;

extern "C" {

const char* compat_last_error() { return g_err.c_str(); }

int compat_compute_proj(const float* K9, int w, int h, float* out16) {
    try {
        cv::Mat K;
        K.rows = 3; K.cols = 3; K.p = const_cast<float*>(K9);
        cuda_renderer::Model::mat4x4 p = cuda_renderer::compute_proj(K, w, h);
        std::memcpy(out16, &p, 64);
        return 0;
    } catch (const std::exception& e) { g_err = e.what(); return 1; }
}

// returns the point count; copies up to `cap` points into the caller's buffers and frees the callee-malloc'd ones
int compat_depth2cloud(const int32_t* depth, const uint8_t* rgb3, int w, int h, float cx, float cy, float fx, float fy,
                       float depth_factor, int stride, const uint8_t* label_mask, const double* bounds6,
                       const float* tf16_colmajor, int cap, float* out_soa, float* out_aos, uint8_t* out_rgb,
                       int* out_label, int* out_n) {
    try {
        const size_t n_px = (size_t)w * h;
        std::vector<int32_t> d(depth, depth + n_px);
        std::vector<std::vector<uint8_t>> color(3);
        for (int k = 0; k < 3; ++k) color[k].assign(rgb3 + k * n_px, rgb3 + (k + 1) * n_px);
        std::vector<uint8_t> mask;
        if (label_mask) mask.assign(label_mask, label_mask + n_px);
        std::vector<double> bounds;
        if (bounds6) bounds.assign(bounds6, bounds6 + 6);
        Eigen::Matrix4f tf;
        if (tf16_colmajor) std::memcpy(&tf, tf16_colmajor, 64);
        Eigen::Vector3f* aos = nullptr; float* soa = nullptr; uint8_t* rgb = nullptr;
        int *dc = nullptr, *pm = nullptr, *lab = nullptr, n = 0;
        std::vector<int> occluded;
        cuda_renderer::depth2cloud_global(d, color, aos, soa, rgb, dc, n, pm, lab, w, h, 1, occluded, cx, cy, fx, fy,
                                          depth_factor, stride, 3, mask, bounds, tf16_colmajor ? &tf : nullptr);
        *out_n = n;
        const int m = n < cap ? n : cap;
        for (int k = 0; k < 3; ++k) {
            std::memcpy(out_soa + (size_t)k * cap, soa + (size_t)k * n, (size_t)m * 4);
            std::memcpy(out_rgb + (size_t)k * cap, rgb + (size_t)k * n, (size_t)m);
        }
        std::memcpy(out_aos, aos, (size_t)m * 12);
        std::memcpy(out_label, lab, (size_t)m * 4);
        std::free(aos); std::free(soa); std::free(rgb); std::free(dc); std::free(pm); std::free(lab);
        return 0;
    } catch (const std::exception& e) { g_err = e.what(); return 1; }
}

int compat_render(const char* stage, const void* tris64, int n_tris, const float* poses16, int n_poses,
                  const int* pose_model_map, const int* tris_model_count, int n_models, int w, int h, const float* proj16,
                  const int32_t* source_depth, const uint8_t* source_rgb3, const uint8_t* source_label,
                  const int* pose_label, int stride, int depth_factor, float cx, float cy, float fx, float fy,
                  const float* obs_aos, const uint8_t* obs_rgb_soa, int n_obs, const float* pose_obs_total,
                  const int* obs_label, int cost_type, int calc_observed, float sensor_resolution, float color_thr,
                  float occlusion_thr,
                  /* out */ int32_t* out_depth, uint8_t* out_rgb3, int cloud_cap, float* out_cloud_soa, int* out_cloud_n,
                  float* out_costs3p) {
    try {
        using namespace cuda_renderer;
        const size_t n_px = (size_t)w * h;
        std::vector<Model::Triangle> tris(n_tris);
        std::memcpy(tris.data(), tris64, (size_t)n_tris * 64);
        std::vector<Model::mat4x4> poses(n_poses);
        std::memcpy(poses.data(), poses16, (size_t)n_poses * 64);
        Model::mat4x4 proj;
        std::memcpy(&proj, proj16, 64);
        std::vector<int> pmm(pose_model_map, pose_model_map + n_poses), tmc(tris_model_count, tris_model_count + n_models);
        std::vector<int32_t> sd(source_depth, source_depth + n_px);
        std::vector<std::vector<uint8_t>> sc(3);
        for (int k = 0; k < 3; ++k) sc[k].assign(source_rgb3 + k * n_px, source_rgb3 + (k + 1) * n_px);
        std::vector<uint8_t> sl;
        if (source_label) sl.assign(source_label, source_label + n_px);
        std::vector<int> pl;
        if (pose_label) pl.assign(pose_label, pose_label + n_poses);
        std::vector<float> tot;
        if (pose_obs_total) tot.assign(pose_obs_total, pose_obs_total + n_poses);
        std::vector<float> clutter;
        std::vector<int32_t> rd;
        std::vector<std::vector<uint8_t>> rc;
        float *cloud = nullptr, *c_r = nullptr, *c_o = nullptr, *c_d = nullptr;
        uint8_t* cloud_rgb = nullptr;
        int cloud_n = 0, *cloud_pm = nullptr, *dc = nullptr;
        std::vector<Model::mat4x4> adjusted;
        gpu_stats stats{};
        render_cuda_multi_unified(stage, tris, poses, pmm, tmc, (size_t)w, (size_t)h, proj, sd, sc, 0, clutter, sl, pl,
                                  stride, 3, depth_factor, cx, cy, fx, fy, nullptr,
                                  reinterpret_cast<Eigen::Vector3f*>(const_cast<float*>(obs_aos)),
                                  const_cast<uint8_t*>(obs_rgb_soa), n_obs, tot, const_cast<int*>(obs_label), cost_type,
                                  calc_observed != 0, sensor_resolution, color_thr, occlusion_thr, false, rd, rc, cloud,
                                  cloud_rgb, cloud_n, cloud_pm, dc, adjusted, c_r, c_o, c_d, stats);
        if (out_depth && !rd.empty()) {
            std::memcpy(out_depth, rd.data(), rd.size() * 4);
            for (int k = 0; k < 3; ++k) std::memcpy(out_rgb3 + k * rd.size(), rc[k].data(), rd.size());
        }
        if (out_cloud_n) *out_cloud_n = cloud_n;
        if (out_cloud_soa && cloud) {
            const int m = cloud_n < cloud_cap ? cloud_n : cloud_cap;
            for (int k = 0; k < 3; ++k) std::memcpy(out_cloud_soa + (size_t)k * cloud_cap, cloud + (size_t)k * cloud_n, (size_t)m * 4);
        }
        if (out_costs3p && c_r) {
            std::memcpy(out_costs3p, c_r, (size_t)n_poses * 4);
            if (c_o) std::memcpy(out_costs3p + n_poses, c_o, (size_t)n_poses * 4);
            if (c_d) std::memcpy(out_costs3p + 2 * (size_t)n_poses, c_d, (size_t)n_poses * 4);
        }
        std::free(cloud); std::free(cloud_rgb); std::free(cloud_pm); std::free(dc); std::free(c_r); std::free(c_o); std::free(c_d);
        return 0;
    } catch (const std::exception& e) { g_err = e.what(); return 1; }
}

}  // extern "C"

// ---- kNN package API through the reference's own <cuda_renderer/knncuda.h> declarations ---------------------------
#include "cuda_renderer/knncuda.h"

extern "C" int compat_knn(int variant, const float* ref, int ref_nb, const float* query, int query_nb, int dim, int k,
                          float* knn_dist, int* knn_index) {
    using namespace cuda_renderer;
    bool ok = false;
    switch (variant) {
        case 0: ok = knn_cuda_global(ref, ref_nb, query, query_nb, dim, k, knn_dist, knn_index); break;
        case 1: ok = knn_cuda_texture(ref, ref_nb, query, query_nb, dim, k, knn_dist, knn_index); break;
        case 2: ok = knn_cublas(ref, ref_nb, query, query_nb, dim, k, knn_dist, knn_index); break;
        default: ok = knn_test(ref, ref_nb, query, query_nb, dim, k, knn_dist, knn_index); break;
    }
    return ok ? 0 : 1;
}
