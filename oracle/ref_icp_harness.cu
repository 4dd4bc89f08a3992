// TEST INFRASTRUCTURE.  The reference's own cuda_icp sources, compiled unmodified from where they lie:
//   cuda_icp/icp.cu                      ICP_Point2Plane_cuda<Scene_projective> (thrust transform_reduce loop), transform_pcd_cuda
//   cuda_icp/scene/common.cu             device_vector_holder
//   cuda_icp/scene/common.cpp            get_normal (host)
//   cuda_icp/scene/depth_scene/depth_scene.cpp   Scene_projective::init_Scene_projective_cpu (host)
//   cuda_icp/scene/pcd_scene/pcd_scene.cpp, .cu  Scene_nn: scene cloud, KDTree_cpu::build_tree, upload (Scene_nn::query is in
//                                        pcd_scene.h and runs inside ICP_Point2Plane_cuda<Scene_nn>)
// behind extern "C" trampolines.  ONE substitution: cuda_icp::eigen_slover_666 (icp.cpp:29-36) needs Eigen's LDLT and
// AngleAxis, which are not installed here; it is restated below (LDL^T in double, R = Rz Ry Rx, icp.cpp:7-17).
// cuda_icp/icp.cpp itself (the CPU loop) cannot be built for the same reason.
#include <cstring>
#include <cmath>
#include <vector>

#include "cuda_icp/icp.h"
#include "cuda_icp/pcd_scene.h"

namespace cuda_icp {
Mat4x4f eigen_slover_666(float* A, float* b) {
    double L[6][6] = {{0}}, D[6], y[6], x[6] = {0, 0, 0, 0, 0, 0};
    bool ok = true;
    for (int j = 0; j < 6 && ok; ++j) {
        double s = A[j * 6 + j];
        for (int k = 0; k < j; ++k) s -= L[j][k] * L[j][k] * D[k];
        D[j] = s;
        if (!(std::fabs(s) > 1e-300)) { ok = false; break; }
        for (int i = j + 1; i < 6; ++i) {
            double t = A[i * 6 + j];
            for (int k = 0; k < j; ++k) t -= L[i][k] * L[j][k] * D[k];
            L[i][j] = t / s;
        }
    }
    if (ok) {
        for (int i = 0; i < 6; ++i) { double s = b[i]; for (int k = 0; k < i; ++k) s -= L[i][k] * y[k]; y[i] = s; }
        for (int i = 5; i >= 0; --i) { double s = y[i] / D[i]; for (int k = i + 1; k < 6; ++k) s -= L[k][i] * x[k]; x[i] = s; }
    }
    const double ca = std::cos(x[0]), sa = std::sin(x[0]), cb = std::cos(x[1]), sb = std::sin(x[1]), cg = std::cos(x[2]), sg = std::sin(x[2]);
    const double E[16] = {cg * cb, cg * sb * sa - sg * ca, cg * sb * ca + sg * sa, x[3],
                          sg * cb, sg * sb * sa + cg * ca, sg * sb * ca - cg * sa, x[4],
                          -sb, cb * sa, cb * ca, x[5], 0, 0, 0, 1};
    Mat4x4f out;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out[i][j] = (float)E[i * 4 + j];
    return out;
}
}  // namespace cuda_icp

extern "C" {

// reference scene preparation on the host: pcd + normals (width*height*3 floats each)
int ref_icp_scene(const int32_t* depth, int width, int height, const float* K9, float* out_pcd, float* out_normals) {
    cv::Mat m;
    m.rows = height; m.cols = width; m.t = CV_32S; m.p = (float*)const_cast<int32_t*>(depth);
    Mat3x3f K;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) K[i][j] = K9[i * 3 + j];
    std::vector<Vec3f> pcd, normals;
    Scene_projective scene;
    scene.init_Scene_projective_cpu(m, K, pcd, normals, (size_t)width, (size_t)height, 0.1f);
    std::memcpy(out_pcd, pcd.data(), pcd.size() * sizeof(Vec3f));
    std::memcpy(out_normals, normals.data(), normals.size() * sizeof(Vec3f));
    return 0;
}

// reference ICP loop on the GPU against scene buffers prepared by the caller
int ref_icp_point2plane(float* model_pcd /* n x 3, transformed in place */, int n, const float* scene_pcd,
                        const float* scene_normals, int width, int height, const float* K9, float max_dist_diff,
                        float relative_fitness, float relative_rmse, int max_iteration, float* out_T16, float* out_fitness,
                        float* out_rmse) {
    device_vector_holder<Vec3f> model((size_t)n), spcd((size_t)width * height), snrm((size_t)width * height);
    cudaMemcpy(model.data(), model_pcd, (size_t)n * 12, cudaMemcpyHostToDevice);
    cudaMemcpy(spcd.data(), scene_pcd, (size_t)width * height * 12, cudaMemcpyHostToDevice);
    cudaMemcpy(snrm.data(), scene_normals, (size_t)width * height * 12, cudaMemcpyHostToDevice);
    Scene_projective scene;
    scene.width = width; scene.height = height; scene.max_dist_diff = max_dist_diff;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) scene.K[i][j] = K9[i * 3 + j];
    scene.pcd_ptr = spcd.data();
    scene.normal_ptr = snrm.data();
    cuda_icp::RegistrationResult r = cuda_icp::ICP_Point2Plane_cuda(model, scene,
        cuda_icp::ICPConvergenceCriteria(relative_fitness, relative_rmse, max_iteration));
    cudaDeviceSynchronize();
    cudaMemcpy(model_pcd, model.data(), (size_t)n * 12, cudaMemcpyDeviceToHost);
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out_T16[i * 4 + j] = r.transformation_[i][j];
    *out_fitness = r.fitness_;
    *out_rmse = r.inlier_rmse_;
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}

// reference ICP loop with kd-tree association (Scene_nn, pcd_scene.h:47-136): the scene cloud, its normals and the tree
// are built by the reference's own init_Scene_nn_cuda from the depth image
int ref_icp_nn_point2plane(float* model_pcd /* n x 3, transformed in place */, int n, const int32_t* depth, int width,
                           int height, const float* K9, float relative_fitness, float relative_rmse, int max_iteration,
                           float* out_T16, float* out_fitness, float* out_rmse) {
    cv::Mat m;
    m.rows = height; m.cols = width; m.t = CV_32S; m.p = (float*)const_cast<int32_t*>(depth);
    Mat3x3f K;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) K[i][j] = K9[i * 3 + j];
    KDTree_cuda tree;
    Scene_nn scene;
    scene.init_Scene_nn_cuda(m, K, tree);
    device_vector_holder<Vec3f> model((size_t)n);
    cudaMemcpy(model.data(), model_pcd, (size_t)n * 12, cudaMemcpyHostToDevice);
    cuda_icp::RegistrationResult r = cuda_icp::ICP_Point2Plane_cuda(model, scene,
        cuda_icp::ICPConvergenceCriteria(relative_fitness, relative_rmse, max_iteration));
    cudaDeviceSynchronize();
    cudaMemcpy(model_pcd, model.data(), (size_t)n * 12, cudaMemcpyDeviceToHost);
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out_T16[i * 4 + j] = r.transformation_[i][j];
    *out_fitness = r.fitness_;
    *out_rmse = r.inlier_rmse_;
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}

// reference cuda_icp::depth2cloud_cuda<int32_t | uint16_t> (icp.cu:225-297) on a host depth image; returns the count.
// Only stride 1 is meaningful: for stride > 1 the reference indexes its width*height/stride^2 mask with x + y*width.
int ref_icp_depth2cloud(const void* depth, int bits, int width, int height, const float* K9, int stride, int tl_x, int tl_y,
                        float* out_xyz, int capacity, int* out_n) {
    Mat3x3f K;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) K[i][j] = K9[i * 3 + j];
    void* d = NULL;
    const size_t bytes = (size_t)width * height * (bits / 8);
    cudaMalloc(&d, bytes);
    cudaMemcpy(d, depth, bytes, cudaMemcpyHostToDevice);
    {
        device_vector_holder<Vec3f> cloud = bits == 16
            ? cuda_icp::depth2cloud_cuda((uint16_t*)d, (uint32_t)width, (uint32_t)height, K, (uint32_t)stride, (uint32_t)tl_x, (uint32_t)tl_y)
            : cuda_icp::depth2cloud_cuda((int32_t*)d, (uint32_t)width, (uint32_t)height, K, (uint32_t)stride, (uint32_t)tl_x, (uint32_t)tl_y);
        cudaDeviceSynchronize();
        *out_n = (int)cloud.size();
        const size_t m = cloud.size() < (size_t)capacity ? cloud.size() : (size_t)capacity;
        if (m) cudaMemcpy(out_xyz, cloud.data(), m * 12, cudaMemcpyDeviceToHost);
    }
    cudaFree(d);
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}

}  // extern "C"
