// Drop-in backend for the reference's cuda_renderer package: the three functions
// sbpl_perception::EnvObjectRecognition calls, with the reference's OWN declarations (this file includes the reference's
// unchanged <cuda_renderer/renderer.h>, renderer.h:87,131-153,221-268) implemented on top of the C-ABI of
// libperch_b200.so.  A maintainer compiles this file instead of cuda_renderer/src/cuda/renderer.cu
// (cuda_renderer/CMakeLists.txt:53-54) and links libperch_b200.so; nothing above the boundary changes.
//
// In this repository it is compile-checked against the reference header (with the POD shim headers of
// oracle/ref_shims standing in for Eigen/OpenCV/assimp, which are not installed here) by `make -C oracle compat`,
// and exercised on the GPU through oracle/compat_driver.cpp by tests/test_compat_gpu.py.
//
// Behaviour kept from renderer.cu: callee-malloc'd T*& outputs owned by the caller afterwards (1825-1828, 1911-1917,
// 2045-2050), result_color filled by push_back of three planes (1608-1610), stage strings matched with the same
// compare/find rules (1594, 1821, 1908), `adjusted_poses` left untouched without do_icp.  Differences: no printf
// chatter, CUDA/argument errors throw std::runtime_error instead of being ignored.
#include "cuda_renderer/renderer.h"

#include <cstdlib>
#include <climits>
#include <cstring>
#include <stdexcept>
#include <string>

#include "perch_b200.h"

namespace cuda_renderer {
namespace {

static_assert(sizeof(Model::Triangle) == sizeof(perch_triangle), "64-byte Triangle (model.h:62-73)");
static_assert(sizeof(Model::mat4x4) == 16 * sizeof(float), "mat4x4 = 16 floats a0..d3 (model.h:74-127)");

perch_ctx* context() {
    static perch_ctx* ctx = [] {
        perch_ctx* c = nullptr;
        if (perch_create(0, &c) != PERCH_OK) throw std::runtime_error(perch_last_error());
        return c;
    }();
    return ctx;
}

void check(int rc) {
    if (rc != PERCH_OK) throw std::runtime_error(perch_last_error());
}

template <class T> T* alloc(size_t n) { return static_cast<T*>(std::malloc(std::max<size_t>(n, 1) * sizeof(T))); }

}  // namespace

Model::mat4x4 compute_proj(const cv::Mat& K, int width, int height, float near, float far) {
    float k9[9];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) k9[3 * r + c] = K.at<float>(r, c);
    Model::mat4x4 p;
    check(perch_compute_proj(k9, width, height, near, far, reinterpret_cast<float*>(&p)));
    return p;
}

bool depth2cloud_global(const std::vector<int32_t>& depth_data, const std::vector<std::vector<uint8_t>>& color_data,
                        Eigen::Vector3f*& result_cloud_eigen, float*& result_cloud, uint8_t*& result_cloud_color,
                        int*& dc_index, int& point_num, int*& cloud_pose_map, int*& result_observed_cloud_label,
                        int width, int height, int num_poses, const std::vector<int>& /*pose_occluded*/, float kCameraCX,
                        float kCameraCY, float kCameraFX, float kCameraFY, float depth_factor, int stride,
                        int /*point_dim*/, const std::vector<uint8_t>& label_mask_data,
                        const std::vector<double>& observed_cloud_bounds, const Eigen::Matrix4f* camera_transform) {
    perch_ctx* ctx = context();
    const uint8_t* r = color_data.size() > 0 ? color_data[0].data() : nullptr;
    const uint8_t* g = color_data.size() > 1 ? color_data[1].data() : nullptr;
    const uint8_t* b = color_data.size() > 2 ? color_data[2].data() : nullptr;
    const uint8_t* mask = label_mask_data.empty() ? nullptr : label_mask_data.data();
    const double* bounds = observed_cloud_bounds.size() >= 6 ? observed_cloud_bounds.data() : nullptr;
    // Eigen::Matrix4f is column-major, which is what perch_depth2cloud expects
    const float* tf = camera_transform ? &(*camera_transform)(0, 0) : nullptr;
    int n = 0;
    check(perch_depth2cloud(ctx, depth_data.data(), r, g, b, width, height, num_poses, kCameraCX, kCameraCY, kCameraFX,
                            kCameraFY, depth_factor, stride, mask, bounds, tf, &n, nullptr, nullptr, nullptr, nullptr,
                            nullptr, nullptr));
    result_cloud = alloc<float>(3 * (size_t)n);
    result_cloud_eigen = alloc<Eigen::Vector3f>((size_t)n);
    result_cloud_color = alloc<uint8_t>(3 * (size_t)n);
    dc_index = alloc<int>((size_t)num_poses * width * height);
    cloud_pose_map = alloc<int>((size_t)n);
    result_observed_cloud_label = alloc<int>((size_t)n);
    check(perch_depth2cloud(ctx, depth_data.data(), r, g, b, width, height, num_poses, kCameraCX, kCameraCY, kCameraFX,
                            kCameraFY, depth_factor, stride, mask, bounds, tf, &n, result_cloud,
                            reinterpret_cast<float*>(result_cloud_eigen), result_cloud_color, dc_index, cloud_pose_map,
                            result_observed_cloud_label));
    point_num = n;
    return true;
}

void render_cuda_multi_unified(
    const std::string stage, const std::vector<Model::Triangle>& tris, const std::vector<Model::mat4x4>& poses,
    const std::vector<int> pose_model_map, const std::vector<int> tris_model_count, size_t width, size_t height,
    const Model::mat4x4& proj_mat, const std::vector<int32_t>& source_depth,
    const std::vector<std::vector<uint8_t>>& source_color, int single_result_image, std::vector<float>& /*clutter_cost*/,
    const std::vector<uint8_t>& source_mask_label, const std::vector<int>& pose_segmentation_label, int stride,
    int /*point_dim*/, int depth_factor, float kCameraCX, float kCameraCY, float kCameraFX, float kCameraFY,
    float* /*observed_depth (SoA copy of the same cloud; the AoS one is consumed, renderer.cu:1853-1865)*/,
    Eigen::Vector3f* observed_depth_eigen, uint8_t* observed_color, int observed_point_num,
    std::vector<float> pose_observed_points_total, int* result_observed_cloud_label, int cost_type,
    bool calculate_observed_cost, float sensor_resolution, float color_distance_threshold, float occlusion_threshold,
    bool do_icp, std::vector<int32_t>& result_depth, std::vector<std::vector<uint8_t>>& result_color,
    float*& result_cloud, uint8_t*& result_cloud_color, int& result_cloud_point_num, int*& result_cloud_pose_map,
    int*& result_dc_index, std::vector<Model::mat4x4>& adjusted_poses, float*& rendered_cost, float*& observed_cost,
    float*& points_diff_cost, gpu_stats& stats) {
    perch_ctx* ctx = context();
    const int P = (int)poses.size();
    const int W = (int)width, H = (int)height;
    const bool debug = stage.compare("DEBUG") == 0;
    const bool want_images = debug || stage.compare("RENDER") == 0;
    const bool want_cloud = debug || stage.find("CLOUD") != std::string::npos;
    const bool want_cost = debug || stage.find("COST") != std::string::npos;

    // The reference uploads all of this on every call (renderer.cu:1532-1544, 1668, 1881); the setters are idempotent
    // and a caller that knows its scene can hoist them out of the batch loop.
    check(perch_set_camera(ctx, W, H, reinterpret_cast<const float*>(&proj_mat), kCameraCX, kCameraCY, kCameraFX,
                           kCameraFY, (float)depth_factor));
    check(perch_set_models(ctx, reinterpret_cast<const perch_triangle*>(tris.data()), (int64_t)tris.size(),
                           tris_model_count.data(), (int)tris_model_count.size()));
    const bool have_rgb = source_color.size() >= 3;
    check(perch_set_scene(ctx, source_depth.data(), have_rgb ? source_color[0].data() : nullptr,
                          have_rgb ? source_color[1].data() : nullptr, have_rgb ? source_color[2].data() : nullptr,
                          source_mask_label.empty() ? nullptr : source_mask_label.data()));
    const bool six_dof = !pose_segmentation_label.empty();
    // the GICP target of do_icp is the observed cloud of THIS call (renderer.cu:1688-1720), cost stage or not
    if ((want_cost || do_icp) && observed_point_num > 0)
        check(perch_set_observed(ctx, reinterpret_cast<const float*>(observed_depth_eigen), observed_color,
                                 six_dof ? result_observed_cloud_label : nullptr, observed_point_num));

    perch_params prm;
    std::memset(&prm, 0, sizeof prm);
    prm.stride = stride;
    prm.cost_type = cost_type;
    prm.calculate_observed_cost = calculate_observed_cost ? 1 : 0;
    prm.sensor_resolution = sensor_resolution;
    prm.color_distance_threshold = color_distance_threshold;
    prm.occlusion_threshold = occlusion_threshold;
    prm.single_result_image = single_result_image;
    prm.do_icp = do_icp ? 1 : 0;

    perch_outputs out;
    std::memset(&out, 0, sizeof out);
    const size_t n_img = (size_t)(single_result_image ? 1 : P) * W * H;
    std::vector<uint8_t> red, green, blue;
    if (want_images) {
        result_depth.resize(n_img);
        red.resize(n_img); green.resize(n_img); blue.resize(n_img);
        out.depth = result_depth.data();
        out.red = red.data(); out.green = green.data(); out.blue = blue.data();
    }
    std::vector<float> cloud_tmp;
    std::vector<uint8_t> cloud_rgb_tmp;
    std::vector<int> cloud_map_tmp;
    size_t cap = 0;
    if (want_cloud) {
        cap = (size_t)P * (W / std::max(stride, 1)) * ((H + stride - 1) / std::max(stride, 1));
        if (cap > (size_t)INT_MAX || (size_t)P * W * H > (size_t)INT_MAX)
            throw std::runtime_error("render_cuda_multi_unified: poses * width * height exceeds INT_MAX (the reference "
                                     "overflows its int indices here too, compute_point_clouds.cuh:60,119): use smaller batches");
        cloud_tmp.resize(3 * cap); cloud_rgb_tmp.resize(3 * cap); cloud_map_tmp.resize(cap);
        result_dc_index = alloc<int>((size_t)P * W * H);
        out.cloud = cloud_tmp.data(); out.cloud_color = cloud_rgb_tmp.data(); out.cloud_pose_map = cloud_map_tmp.data();
        out.dc_index = result_dc_index;
        out.cloud_capacity = (int)cap;
    }
    if (want_cost) {
        rendered_cost = alloc<float>((size_t)P);
        out.rendered_cost = rendered_cost;
        if (calculate_observed_cost) {
            observed_cost = alloc<float>((size_t)P);
            points_diff_cost = alloc<float>((size_t)P);
            out.observed_cost = observed_cost;
            out.points_diff_cost = points_diff_cost;
        }
    }
    if (do_icp && (want_cloud || want_cost)) {
        adjusted_poses.resize(P);  // renderer.cu:1729-1730
        out.adjusted_poses = reinterpret_cast<float*>(adjusted_poses.data());
    }
    // the stage machine of the C-ABI has the reference's four stage names; composite strings fall back to DEBUG
    const char* st = debug ? "DEBUG" : (stage.compare("RENDER") == 0 ? "RENDER" : (stage.compare("CLOUD") == 0 ? "CLOUD" :
                     (stage.compare("COST") == 0 ? "COST" : "DEBUG")));
    check(perch_render_cost(ctx, st, reinterpret_cast<const float*>(poses.data()), P, pose_model_map.data(),
                            six_dof ? pose_segmentation_label.data() : nullptr,
                            pose_observed_points_total.empty() ? nullptr : pose_observed_points_total.data(), &prm, &out));
    if (want_images) {
        result_color.push_back(std::move(red));
        result_color.push_back(std::move(green));
        result_color.push_back(std::move(blue));
    }
    if (want_cloud) {
        const size_t n = (size_t)out.cloud_point_num;
        result_cloud_point_num = (int)n;
        result_cloud = alloc<float>(3 * n);
        result_cloud_color = alloc<uint8_t>(3 * n);
        result_cloud_pose_map = alloc<int>(n);
        for (int k = 0; k < 3; ++k) {  // SoA with row stride n, like the reference's cudaMemcpy2D (renderer.cu:1831-1832)
            std::memcpy(result_cloud + k * n, cloud_tmp.data() + k * cap, n * sizeof(float));
            std::memcpy(result_cloud_color + k * n, cloud_rgb_tmp.data() + k * cap, n);
        }
        std::memcpy(result_cloud_pose_map, cloud_map_tmp.data(), n * sizeof(int));
    }
    perch_gpu_stats gs;
    if (perch_get_stats(ctx, &gs) == PERCH_OK) {
        stats.peak_memory_usage = gs.peak_memory_usage;
        stats.icp_runtime = gs.icp_runtime;
    }
}

}  // namespace cuda_renderer
