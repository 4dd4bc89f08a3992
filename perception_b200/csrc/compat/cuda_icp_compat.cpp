// Drop-in backend for the reference's cuda_icp package: the CUDA entry points declared in the reference's unchanged
// <cuda_icp/icp.h> (83-98), <cuda_icp/depth_scene.h> (19-25) and <cuda_icp/pcd_scene.h> (55-58), implemented on top of
// the C-ABI of libperch_b200.so.  A maintainer compiles this file instead of cuda_icp/icp.cu, scene/common.cu,
// scene/depth_scene/depth_scene.cu and scene/pcd_scene/pcd_scene.cu (cuda_icp/CMakeLists.txt) with -DCUDA_ON and links
// libperch_b200.so; callers of cuda_icp::ICP_Point2Plane / depth2cloud keep their code.
//
//   device_vector_holder<T>                    common.h:21-43        plain cudaMalloc / cudaFree owner (no thrust)
//   cuda_icp::depth2cloud_cuda<uint16_t|int32_t>   icp.h:83-90      perch_icp_depth2cloud_device
//   Scene_projective::init_Scene_projective_cuda   depth_scene.h:19-25   scene points + LINEMOD-style normals computed ON
//                                              THE DEVICE (the reference computes them on the host and uploads)
//   cuda_icp::ICP_Point2Plane_cuda<Scene_projective>   icp.h:92-96   perch_icp_point2plane_device: the whole loop
//                                              (association, 29 sums, 6x6 solve, update) in one launch, no host round trips
//   Scene_nn::init_Scene_nn_cuda / ICP_Point2Plane_cuda<Scene_nn>   pcd_scene.h:55-58, icp.h:97-98: exact grid
//                                              nearest neighbour instead of the host-built kd-tree; KDTree_cuda receives
//                                              the valid scene points and their normals (row-major pixel order), its
//                                              `nodes` one leaf that spans them.  Scene_nn keeps its pointers private, so
//                                              the search runs against the scene of the LAST init_Scene_nn_cuda of the
//                                              process (one live Scene_nn at a time).
// Compile-checked against the reference headers by `make -C oracle ref` (_ref/libcuda_icp_compat.so) and exercised on
// the GPU through oracle/compat_icp_driver.cpp by tests/test_icp.py.  Errors throw std::runtime_error.
#include "cuda_icp/icp.h"

#include <cstring>
#include <stdexcept>
#include <vector>

#include "perch_b200.h"

// ---- device_vector_holder (common.h:21-43) ---------------------------------------------------------------------------------
template <typename T>
void device_vector_holder<T>::__free() {
    if (valid) {
        cudaFree(__gpu_memory);
        valid = false;
        __size = 0;
    }
}
template <typename T>
void device_vector_holder<T>::__malloc(size_t size_) {
    if (valid) __free();
    if (cudaMalloc(reinterpret_cast<void**>(&__gpu_memory), (size_ ? size_ : 1) * sizeof(T)) != cudaSuccess)
        throw std::runtime_error("device_vector_holder: cudaMalloc failed");
    __size = size_;
    valid = true;
}
template <typename T>
device_vector_holder<T>::device_vector_holder(size_t size_) { __malloc(size_); }
template <typename T>
device_vector_holder<T>::device_vector_holder(size_t size_, T init) {
    __malloc(size_);
    std::vector<T> host(size_, init);
    if (size_) cudaMemcpy(__gpu_memory, host.data(), size_ * sizeof(T), cudaMemcpyHostToDevice);
}
template <typename T>
device_vector_holder<T>::~device_vector_holder() { __free(); }
template class device_vector_holder<Vec3f>;
template class device_vector_holder<Node_kdtree>;

namespace {

perch_ctx* icp_context() {
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

void K9_of(const Mat3x3f& K, float out[9]) {
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) out[3 * i + j] = K[i][j];
}

// the depth image as int32 centimetres, whatever the cv::Mat holds (CV_16U or CV_32S, depth_scene.cpp:11-12)
std::vector<int32_t> depth_i32(cv::Mat& m, size_t width, size_t height) {
    std::vector<int32_t> d(width * height);
    const int type = m.type();
    if (type != CV_16U && type != CV_32S) throw std::runtime_error("scene depth must be CV_16U or CV_32S");
    for (size_t r = 0; r < height; ++r)
        for (size_t c = 0; c < width; ++c)
            d[c + r * width] = type == CV_16U ? (int32_t)m.at<uint16_t>((int)r, (int)c) : m.at<int32_t>((int)r, (int)c);
    return d;
}

// The reference runs on the caller's default / per-thread stream, so device buffers the caller filled just before the
// call (a pageable cudaMemcpy returns before its DMA has landed) are ordered with its kernels.  The library's stream is
// its own: wait for the caller's outstanding device work before touching caller-provided DEVICE pointers.
void caller_work_done() {
    if (cudaDeviceSynchronize() != cudaSuccess) throw std::runtime_error("cudaDeviceSynchronize failed before a cuda_icp call");
}

float g_nn_max_dist = 0.01f;  // Scene_nn::max_dist_diff (pcd_scene.h:48; private, no setter in the reference)

cuda_icp::RegistrationResult run_icp(device_vector_holder<Vec3f>& model, const float* d_scene_pcd, const float* d_scene_normals,
                                     int width, int height, const float K9[9], float max_dist, int nn,
                                     const cuda_icp::ICPConvergenceCriteria& cr) {
    float T[16], fit = 0.f, rmse = 0.f;
    caller_work_done();
    check(perch_icp_point2plane_device(icp_context(), reinterpret_cast<float*>(model.data()), (int)model.size(), d_scene_pcd,
                                       d_scene_normals, width, height, K9, max_dist, nn, cr.relative_fitness_,
                                       cr.relative_rmse_, cr.max_iteration_, T, &fit, &rmse, nullptr));
    cuda_icp::RegistrationResult res;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) res.transformation_[i][j] = T[4 * i + j];
    res.fitness_ = fit;
    res.inlier_rmse_ = rmse;
    return res;
}

}  // namespace

// ---- scenes ------------------------------------------------------------------------------------------------------------------
void Scene_projective::init_Scene_projective_cuda(cv::Mat& scene_depth, Mat3x3f& scene_K, device_vector_holder<Vec3f>& pcd_buffer,
                                                  device_vector_holder<Vec3f>& normal_buffer, size_t width_, size_t height_,
                                                  float max_dist_diff_) {
    K = scene_K;
    width = width_;
    height = height_;
    max_dist_diff = max_dist_diff_;
    float K9[9];
    K9_of(scene_K, K9);
    const std::vector<int32_t> d = depth_i32(scene_depth, width_, height_);
    perch_ctx* ctx = icp_context();
    check(perch_icp_set_scene(ctx, d.data(), (int)width_, (int)height_, K9, max_dist_diff_, nullptr, nullptr));
    const float *d_pcd = nullptr, *d_nrm = nullptr;
    check(perch_icp_get_scene_device(ctx, &d_pcd, &d_nrm, nullptr, nullptr));
    const size_t n = width_ * height_;
    pcd_buffer.__malloc(n);
    normal_buffer.__malloc(n);
    cudaMemcpy(pcd_buffer.data(), d_pcd, n * sizeof(Vec3f), cudaMemcpyDeviceToDevice);
    cudaMemcpy(normal_buffer.data(), d_nrm, n * sizeof(Vec3f), cudaMemcpyDeviceToDevice);
    pcd_ptr = pcd_buffer.data();
    normal_ptr = normal_buffer.data();
}

void Scene_nn::init_Scene_nn_cuda(cv::Mat& scene_depth, Mat3x3f& scene_K, KDTree_cuda& kdtree) {
    const size_t W = (size_t)scene_depth.cols, H = (size_t)scene_depth.rows;
    float K9[9];
    K9_of(scene_K, K9);
    std::vector<int32_t> d = depth_i32(scene_depth, W, H);
    // init_Scene_nn_cpu converts CV_32S depth to CV_16U first (pcd_scene.cpp:12-16): saturate like cv::Mat::convertTo
    for (int32_t& v : d) v = v < 0 ? 0 : (v > 65535 ? 65535 : v);
    perch_ctx* ctx = icp_context();
    std::vector<float> pcd(3 * W * H), nrm(3 * W * H);
    check(perch_icp_set_scene(ctx, d.data(), (int)W, (int)H, K9, g_nn_max_dist, pcd.data(), nrm.data()));
    // what the caller's KDTree_cuda holds afterwards: the valid scene points and their normals (pcd_scene.cpp:24-32)
    std::vector<Vec3f> pts, nrms;
    for (size_t i = 0; i < W * H; ++i)
        if (d[i] > 0) {
            pts.push_back({pcd[3 * i], pcd[3 * i + 1], pcd[3 * i + 2]});
            nrms.push_back({nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]});
        }
    kdtree.pcd_buffer.__malloc(pts.size());
    kdtree.normal_buffer.__malloc(nrms.size());
    if (!pts.empty()) {
        cudaMemcpy(kdtree.pcd_buffer.data(), pts.data(), pts.size() * sizeof(Vec3f), cudaMemcpyHostToDevice);
        cudaMemcpy(kdtree.normal_buffer.data(), nrms.data(), nrms.size() * sizeof(Vec3f), cudaMemcpyHostToDevice);
    }
    Node_kdtree leaf;  // one leaf spanning every point: the kd-tree is not what finds neighbours here
    leaf.left = 0;
    leaf.right = (int)pts.size();
    kdtree.nodes.__malloc(1);
    cudaMemcpy(kdtree.nodes.data(), &leaf, sizeof leaf, cudaMemcpyHostToDevice);
    pcd_ptr = kdtree.pcd_buffer.data();
    normal_ptr = kdtree.normal_buffer.data();
    node_ptr = kdtree.nodes.data();
}

namespace cuda_icp {

// ---- depth2cloud_cuda (icp.h:83-90): `depth` is a DEVICE pointer, as in the reference ------------------------------------------
template <class T>
device_vector_holder<Vec3f> depth2cloud_cuda(T* depth, uint32_t width, uint32_t height, Mat3x3f& K, uint32_t stride, uint32_t tl_x,
                                             uint32_t tl_y) {
    static_assert(sizeof(T) == 2 || sizeof(T) == 4, "uint16_t or int32_t depth");
    float K9[9];
    K9_of(K, K9);
    perch_ctx* ctx = icp_context();
    int n = 0;
    caller_work_done();
    check(perch_icp_depth2cloud_device(ctx, depth, 8 * (int)sizeof(T), (int)width, (int)height, K9, (int)stride, (int)tl_x,
                                       (int)tl_y, nullptr, 0, &n));
    device_vector_holder<Vec3f> cloud((size_t)n);
    if (n > 0)
        check(perch_icp_depth2cloud_device(ctx, depth, 8 * (int)sizeof(T), (int)width, (int)height, K9, (int)stride, (int)tl_x,
                                           (int)tl_y, reinterpret_cast<float*>(cloud.data()), n, &n));
    return cloud;
}
template device_vector_holder<Vec3f> depth2cloud_cuda(uint16_t* depth, uint32_t width, uint32_t height, Mat3x3f& K,
                                                      uint32_t stride, uint32_t tl_x, uint32_t tl_y);
template device_vector_holder<Vec3f> depth2cloud_cuda(int32_t* depth, uint32_t width, uint32_t height, Mat3x3f& K, uint32_t stride,
                                                      uint32_t tl_x, uint32_t tl_y);

// ---- ICP_Point2Plane_cuda (icp.h:92-98): the primary template + the two instantiations the header declares extern ---------------
namespace {
RegistrationResult icp_on(device_vector_holder<Vec3f>& model_pcd, const Scene_projective& scene, const ICPConvergenceCriteria& cr) {
    float K9[9];
    K9_of(scene.K, K9);
    return run_icp(model_pcd, reinterpret_cast<const float*>(scene.pcd_ptr), reinterpret_cast<const float*>(scene.normal_ptr),
                   (int)scene.width, (int)scene.height, K9, scene.max_dist_diff, 0, cr);
}
RegistrationResult icp_on(device_vector_holder<Vec3f>& model_pcd, const Scene_nn& /*pointers are private*/, const ICPConvergenceCriteria& cr) {
    return run_icp(model_pcd, nullptr, nullptr, 0, 0, nullptr, g_nn_max_dist, 1, cr);
}
}  // namespace

template <class Scene>
RegistrationResult ICP_Point2Plane_cuda(device_vector_holder<Vec3f>& model_pcd, const Scene scene, const ICPConvergenceCriteria criteria) {
    return icp_on(model_pcd, scene, criteria);
}
template RegistrationResult ICP_Point2Plane_cuda(device_vector_holder<Vec3f>&, const Scene_projective, const ICPConvergenceCriteria);
template RegistrationResult ICP_Point2Plane_cuda(device_vector_holder<Vec3f>&, const Scene_nn, const ICPConvergenceCriteria);

}  // namespace cuda_icp
