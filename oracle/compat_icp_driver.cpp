// TEST INFRASTRUCTURE.  extern "C" trampolines that call the C++ drop-in backend for cuda_icp
// (perception_b200/csrc/compat/cuda_icp_compat.cpp) through the REFERENCE's own declarations (<cuda_icp/icp.h>:
// device_vector_holder / Scene_projective / Scene_nn / cv::Mat arguments), so that tests/test_icp.py can compare what a
// reference caller would receive with what the C-ABI returns directly.  Same shapes as oracle/ref_icp_harness.cu.
#include <cstring>
#include <exception>
#include <string>
#include <vector>

#include "cuda_icp/icp.h"
#include "cuda_icp/pcd_scene.h"

static std::string g_err;

extern "C" {

const char* compat_icp_last_error() { return g_err.c_str(); }

static cv::Mat mat_of(const int32_t* depth, int width, int height) {
    cv::Mat m;
    m.rows = height; m.cols = width; m.t = CV_32S; m.p = (float*)const_cast<int32_t*>(depth);
    return m;
}

int compat_icp_projective(float* model_pcd, int n, const int32_t* depth, int width, int height, const float* K9, float max_dist_diff,
                          float relative_fitness, float relative_rmse, int max_iteration, float* out_T16, float* out_fitness,
                          float* out_rmse, float* out_scene_pcd, float* out_scene_normals) {
    try {
        cv::Mat m = mat_of(depth, width, height);
        Mat3x3f K;
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) K[i][j] = K9[i * 3 + j];
        device_vector_holder<Vec3f> spcd, snrm;
        Scene_projective scene;
        scene.init_Scene_projective_cuda(m, K, spcd, snrm, (size_t)width, (size_t)height, max_dist_diff);
        if (out_scene_pcd) cudaMemcpy(out_scene_pcd, spcd.data(), spcd.size() * 12, cudaMemcpyDeviceToHost);
        if (out_scene_normals) cudaMemcpy(out_scene_normals, snrm.data(), snrm.size() * 12, cudaMemcpyDeviceToHost);
        device_vector_holder<Vec3f> model((size_t)n);
        cudaMemcpy(model.data(), model_pcd, (size_t)n * 12, cudaMemcpyHostToDevice);
        cuda_icp::RegistrationResult r = cuda_icp::ICP_Point2Plane_cuda(model, scene,
            cuda_icp::ICPConvergenceCriteria(relative_fitness, relative_rmse, max_iteration));
        cudaMemcpy(model_pcd, model.data(), (size_t)n * 12, cudaMemcpyDeviceToHost);
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out_T16[i * 4 + j] = r.transformation_[i][j];
        *out_fitness = r.fitness_;
        *out_rmse = r.inlier_rmse_;
        return 0;
    } catch (const std::exception& e) { g_err = e.what(); return 1; }
}

int compat_icp_nn(float* model_pcd, int n, const int32_t* depth, int width, int height, const float* K9, float relative_fitness,
                  float relative_rmse, int max_iteration, float* out_T16, float* out_fitness, float* out_rmse, int* out_scene_points) {
    try {
        cv::Mat m = mat_of(depth, width, height);
        Mat3x3f K;
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) K[i][j] = K9[i * 3 + j];
        KDTree_cuda tree;
        Scene_nn scene;
        scene.init_Scene_nn_cuda(m, K, tree);
        if (out_scene_points) *out_scene_points = (int)tree.pcd_buffer.size();
        device_vector_holder<Vec3f> model((size_t)n);
        cudaMemcpy(model.data(), model_pcd, (size_t)n * 12, cudaMemcpyHostToDevice);
        cuda_icp::RegistrationResult r = cuda_icp::ICP_Point2Plane_cuda(model, scene,
            cuda_icp::ICPConvergenceCriteria(relative_fitness, relative_rmse, max_iteration));
        cudaMemcpy(model_pcd, model.data(), (size_t)n * 12, cudaMemcpyDeviceToHost);
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out_T16[i * 4 + j] = r.transformation_[i][j];
        *out_fitness = r.fitness_;
        *out_rmse = r.inlier_rmse_;
        return 0;
    } catch (const std::exception& e) { g_err = e.what(); return 1; }
}

// depth (host int32 or uint16) -> device -> cuda_icp::depth2cloud_cuda -> host; returns the point count
int compat_icp_depth2cloud(const void* depth, int bits, int width, int height, const float* K9, int stride, int tl_x, int tl_y,
                           float* out_xyz, int capacity, int* out_n) {
    try {
        Mat3x3f K;
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) K[i][j] = K9[i * 3 + j];
        void* d = nullptr;
        const size_t bytes = (size_t)width * height * (bits / 8);
        cudaMalloc(&d, bytes);
        cudaMemcpy(d, depth, bytes, cudaMemcpyHostToDevice);
        device_vector_holder<Vec3f> cloud = bits == 16
            ? cuda_icp::depth2cloud_cuda((uint16_t*)d, (uint32_t)width, (uint32_t)height, K, (uint32_t)stride, (uint32_t)tl_x, (uint32_t)tl_y)
            : cuda_icp::depth2cloud_cuda((int32_t*)d, (uint32_t)width, (uint32_t)height, K, (uint32_t)stride, (uint32_t)tl_x, (uint32_t)tl_y);
        *out_n = (int)cloud.size();
        const size_t m = cloud.size() < (size_t)capacity ? cloud.size() : (size_t)capacity;
        if (m) cudaMemcpy(out_xyz, cloud.data(), m * 12, cudaMemcpyDeviceToHost);
        cudaFree(d);
        return 0;
    } catch (const std::exception& e) { g_err = e.what(); return 1; }
}

}  // extern "C"
