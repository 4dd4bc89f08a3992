// Point-to-plane ICP against a depth-image scene with projective association: the package API of the reference's
// cuda_icp (ICP_Point2Plane_cuda<Scene_projective>, icp.h:69-99, icp.cu:156-218; Scene_projective::query,
// depth_scene.h:29-50; scene preparation depth_scene.cpp:3-37 + get_normal, scene/common.cpp:17-107; per-point normal
// equations thrust__pcd2Ab, icp.h:129-210; update eigen_slover_666, icp.cpp:7-36).
//
// The reference registers ONE cloud per call and goes through the host twice per iteration (thrust reduce -> D2H ->
// Eigen LDLT on the CPU -> kernel to move the cloud).  Here a launch registers a BATCH of clouds, one CTA each, and the
// whole loop -- association, the 29 sums, the fitness / rmse stop rule, the 6x6 solve in double, the Euler-ZYX update
// and the transform of the cloud -- stays on the device.  Sums are reduced in a fixed order (deterministic), whereas
// the reference's cub reduction order is unspecified; results agree within the last bits of float accumulation.
#include <cuda_runtime.h>

#include <cfloat>
#include <cstdint>

#include "kernels.cuh"

namespace perch {
namespace {

constexpr int kIcpThreads = 256;
constexpr int kIcpWarps = kIcpThreads / 32;

// ---- scene preparation ----------------------------------------------------------------------------------------------
// pcd: dep2pcd (common.h:50-66), depth in cm read as unsigned like scene_depth.at<uint32_t> (depth_scene.cpp:26).
// normal: LINEMOD-style integer bilateral gradient on the depth saturated to 16 bits (common.cpp:23-29: convertTo CV_16U),
// radius 5, distance threshold 2000, difference threshold 50, 64-bit integer accumulation (common.cpp:3-15, 56-98).
__device__ __forceinline__ long long sat16(int32_t d) { return d < 0 ? 0 : (d > 65535 ? 65535 : d); }

__global__ void icp_scene_kernel(const int32_t* __restrict__ depth, int W, int H, float k00, float k02, float k11, float k12,
                                 float* __restrict__ pcd, float* __restrict__ normals) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= W || y >= H) return;
    const size_t i = (size_t)y * W + x;
    const uint32_t dep = (uint32_t)depth[i];
    float px = 0.f, py = 0.f, pz = 0.f;
    if (dep != 0) {
        pz = __fdiv_rn((float)dep, 100.0f);
        px = __fmul_rn(__fdiv_rn(__fsub_rn((float)x, k02), k00), pz);
        py = __fmul_rn(__fdiv_rn(__fsub_rn((float)y, k12), k11), pz);
    }
    pcd[3 * i] = px; pcd[3 * i + 1] = py; pcd[3 * i + 2] = pz;
    float nx = 0.f, ny = 0.f, nz = 0.f;
    const int r = 5;
    if (y >= r && y < H - r - 1 && x >= r && x < W - r - 1) {
        const long long d = sat16(depth[i]);
        if (d < 2000) {
            long long A0 = 0, A1 = 0, A3 = 0, b0 = 0, b1 = 0;
#pragma unroll
            for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
                for (int dx = -1; dx <= 1; ++dx) {
                    if (dx == 0 && dy == 0) continue;
                    const long long ii = dx * r, jj = dy * r;
                    const long long delta = sat16(depth[(size_t)(y + dy * r) * W + (x + dx * r)]) - d;
                    const long long f = (delta < 0 ? -delta : delta) < 50 ? 1 : 0;
                    A0 += f * ii * ii; A1 += f * ii * jj; A3 += f * jj * jj;
                    b0 += f * ii * delta; b1 += f * jj * delta;
                }
            const long long det = A0 * A3 - A1 * A1;
            const long long ddx = A3 * b0 - A1 * b1;
            const long long ddy = -A1 * b0 + A0 * b1;
            // the reference runs this on the HOST (plain IEEE float, no contraction): explicit _rn operations
            const float lx = __fmul_rn(k00, (float)ddx), ly = __fmul_rn(k11, (float)ddy), lz = (float)(-det * d);
            const float len = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(lx, lx), __fmul_rn(ly, ly)), __fmul_rn(lz, lz)));
            if (len > 0.f) {
                const float inv = __fdiv_rn(1.0f, len);
                nx = __fmul_rn(lx, inv); ny = __fmul_rn(ly, inv); nz = __fmul_rn(lz, inv);
            }
        }
    }
    normals[3 * i] = nx; normals[3 * i + 1] = ny; normals[3 * i + 2] = nz;
}

// LDL^T solve of the symmetric 6x6 system A x = b in double (Eigen's ldlt() pivots; for the positive semi-definite
// normal matrices of ICP an unpivoted factorisation gives the same solution up to rounding).
__device__ inline bool ldlt_solve6(const double* A, const double* b, double* x) {
    double L[6][6], D[6];
    for (int j = 0; j < 6; ++j) {
        double s = A[j * 6 + j];
        for (int k = 0; k < j; ++k) s -= L[j][k] * L[j][k] * D[k];
        D[j] = s;
        if (!(fabs(s) > 1e-300)) return false;
        for (int i = j + 1; i < 6; ++i) {
            double t = A[i * 6 + j];
            for (int k = 0; k < j; ++k) t -= L[i][k] * L[j][k] * D[k];
            L[i][j] = t / s;
        }
    }
    double y[6];
    for (int i = 0; i < 6; ++i) {
        double s = b[i];
        for (int k = 0; k < i; ++k) s -= L[i][k] * y[k];
        y[i] = s;
    }
    for (int i = 5; i >= 0; --i) {
        double s = y[i] / D[i];
        for (int k = i + 1; k < 6; ++k) s -= L[k][i] * x[k];
        x[i] = s;
    }
    return true;
}

// Scene_nn::query (pcd_scene.h:60-136): the nearest scene point, accepted iff its squared distance is below
// max_dist_diff^2.  The reference walks a kd-tree; any exact search gives the same point (equal distances aside: the
// lowest pixel index wins here, the tree's visiting order there), and a point closer than max_dist_diff lies in the 27
// cells around the query because cells are wider than that.
__device__ __forceinline__ bool nn_associate(const ObsGridDev& g, float max_d2, float qx, float qy, float qz, int& idx) {
    const float fx = floorf((qx - g.ox) * g.inv_h), fy = floorf((qy - g.oy) * g.inv_h), fz = floorf((qz - g.oz) * g.inv_h);
    if (!(fx >= -1.f && fx <= (float)g.dx && fy >= -1.f && fy <= (float)g.dy && fz >= -1.f && fz <= (float)g.dz)) return false;
    const int cx = (int)fx, cy = (int)fy, cz = (int)fz;
    const int x0 = max(cx - 1, 0), x1 = min(cx + 1, g.dx - 1);
    if (x0 > x1) return false;
    float best = FLT_MAX;
    idx = -1;
    for (int z = max(cz - 1, 0); z <= min(cz + 1, g.dz - 1); ++z)
        for (int y = max(cy - 1, 0); y <= min(cy + 1, g.dy - 1); ++y) {
            const int row = (z * g.dy + y) * g.dx;
            const uint32_t b = __ldg(g.cell_start + row + x0), e = __ldg(g.cell_start + row + x1 + 1);
            for (uint32_t j = b; j < e; ++j) {
                const float4 p = __ldg(g.pts + j);
                const float d = sqdist3(qx, qy, qz, p.x, p.y, p.z);
                const int ref = __float_as_int(p.w);
                if (d < best || (d == best && ref < idx)) { best = d; idx = ref; }
            }
        }
    return idx >= 0 && best < max_d2;
}

template <bool NN>
__global__ void __launch_bounds__(kIcpThreads) icp_point2plane_kernel(IcpArgs a) {
    __shared__ float s_part[kIcpWarps][29];
    __shared__ float s_T[16];        // accumulated transformation, row-major
    __shared__ float s_E[12];        // this iteration's update (rows 0..2)
    __shared__ float s_fit, s_rmse;
    __shared__ int s_stop, s_iters;
    const int cloud = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int begin = a.offsets[cloud], end = a.offsets[cloud + 1], n = end - begin;
    float* pts = a.points + 3 * (size_t)begin;   // transformed in place, like the reference's model_pcd
    if (tid < 16) s_T[tid] = (tid % 5 == 0) ? 1.f : 0.f;
    if (tid == 0) { s_fit = 0.f; s_rmse = 0.f; s_stop = (n == 0) ? 1 : 0; s_iters = 0; }
    __syncthreads();
    for (int iter = 0; iter <= a.max_iteration; ++iter) {
        if (s_stop) break;
        float acc[29];
#pragma unroll
        for (int i = 0; i < 29; ++i) acc[i] = 0.f;
        for (int i = tid; i < n; i += kIcpThreads) {
            const float sx = pts[3 * i], sy = pts[3 * i + 1], sz = pts[3 * i + 2];
            size_t idx;
            if (NN) {
                int ref;
                if (!nn_associate(a.nn_grid, a.max_dist_diff * a.max_dist_diff, sx, sy, sz, ref)) continue;
                idx = (size_t)ref;
            } else {
                // Scene_projective::query (depth_scene.h:29-50) with pcd2dep (common.h:68-78)
                const int u = (int)(sx / sz * a.k00 + a.k02 + 0.5f);
                const int v = (int)(sy / sz * a.k11 + a.k12 + 0.5f);
                if (u >= a.width || v >= a.height || u < 0 || v < 0) continue;
                idx = (size_t)u + (size_t)v * a.width;
            }
            const float dx = a.scene_pcd[3 * idx], dy = a.scene_pcd[3 * idx + 1], dz = a.scene_pcd[3 * idx + 2];
            if (!NN && (dz <= 0.f || fabsf(sz - dz) > a.max_dist_diff)) continue;
            const float nx = a.scene_normals[3 * idx], ny = a.scene_normals[3 * idx + 1], nz = a.scene_normals[3 * idx + 2];
            // thrust__pcd2Ab (icp.h:129-210)
            const float bt = (dx - sx) * nx + (dy - sy) * ny + (dz - sz) * nz;
            float A[6];
            A[0] = nz * sy - ny * sz; A[1] = nx * sz - nz * sx; A[2] = ny * sx - nx * sy;
            A[3] = nx; A[4] = ny; A[5] = nz;
            int o = 0;
#pragma unroll
            for (int r = 0; r < 6; ++r)
#pragma unroll
                for (int c = r; c < 6; ++c) acc[o++] += A[r] * A[c];
#pragma unroll
            for (int r = 0; r < 6; ++r) acc[21 + r] += A[r] * bt;
            acc[27] += bt * bt;
            acc[28] += 1.f;
        }
#pragma unroll
        for (int i = 0; i < 29; ++i) {
            float v = acc[i];
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0) s_part[warp][i] = v;
        }
        __syncthreads();
        if (tid == 0) {
            float sum[29];
            for (int i = 0; i < 29; ++i) {
                float s = 0.f;
                for (int w = 0; w < kIcpWarps; ++w) s += s_part[w][i];
                sum[i] = s;
            }
            const float count = sum[28], total_error = sum[27];
            s_iters = iter;
            if (count == 0.f) s_stop = 1;  // icp.cu:186 "avoid divid 0": the previous result stands
            else {
                const float prev_fit = s_fit, prev_rmse = s_rmse;
                s_fit = count / (float)n;
                s_rmse = sqrtf(total_error / count);
                if (iter == a.max_iteration) s_stop = 1;  // the extra turn only measures (icp.cu:193)
                else if (fabsf(s_fit - prev_fit) < a.relative_fitness && fabsf(s_rmse - prev_rmse) < a.relative_rmse) s_stop = 1;
                else {
                    double Am[36], bv[6], x[6];
                    int o = 0;
                    for (int r = 0; r < 6; ++r)
                        for (int c = r; c < 6; ++c) { Am[r * 6 + c] = sum[o]; Am[c * 6 + r] = sum[o]; ++o; }
                    for (int r = 0; r < 6; ++r) bv[r] = sum[21 + r];
                    if (!ldlt_solve6(Am, bv, x)) s_stop = 1;
                    else {
                        // TransformVector6dToMatrix4d (icp.cpp:7-17): R = Rz(x[2]) Ry(x[1]) Rx(x[0]), t = x[3..5]
                        const double ca = cos(x[0]), sa = sin(x[0]), cb = cos(x[1]), sb = sin(x[1]), cg = cos(x[2]), sg = sin(x[2]);
                        const double E[12] = {cg * cb, cg * sb * sa - sg * ca, cg * sb * ca + sg * sa, x[3],
                                              sg * cb, sg * sb * sa + cg * ca, sg * sb * ca - cg * sa, x[4],
                                              -sb, cb * sa, cb * ca, x[5]};
                        float Ef[12];
                        for (int i = 0; i < 12; ++i) { Ef[i] = (float)E[i]; s_E[i] = Ef[i]; }
                        // result.transformation_ = extrinsic * result.transformation_ (icp.cu:213), float
                        float nT[12];
                        for (int r = 0; r < 3; ++r)
                            for (int c = 0; c < 4; ++c)
                                nT[r * 4 + c] = Ef[r * 4] * s_T[c] + Ef[r * 4 + 1] * s_T[4 + c] + Ef[r * 4 + 2] * s_T[8 + c] +
                                                Ef[r * 4 + 3] * s_T[12 + c];
                        for (int i = 0; i < 12; ++i) s_T[i] = nT[i];
                    }
                }
            }
        }
        __syncthreads();
        if (s_stop) break;
        // transform_pcd_cuda (icp.cu:142-153)
        for (int i = tid; i < n; i += kIcpThreads) {
            const float x = pts[3 * i], y = pts[3 * i + 1], z = pts[3 * i + 2];
            pts[3 * i] = s_E[0] * x + s_E[1] * y + s_E[2] * z + s_E[3];
            pts[3 * i + 1] = s_E[4] * x + s_E[5] * y + s_E[6] * z + s_E[7];
            pts[3 * i + 2] = s_E[8] * x + s_E[9] * y + s_E[10] * z + s_E[11];
        }
        __syncthreads();
    }
    if (tid < 16) a.out_T[16 * (size_t)cloud + tid] = s_T[tid];
    if (tid == 0) {
        a.out_fitness[cloud] = s_fit;
        a.out_rmse[cloud] = s_rmse;
        if (a.out_iterations) a.out_iterations[cloud] = s_iters;
    }
}

}  // namespace

cudaError_t launch_icp_scene(const int32_t* d_depth, int W, int H, const float K9[9], float* d_pcd, float* d_normals,
                             cudaStream_t stream) {
    dim3 block(32, 8), grid((W + 31) / 32, (H + 7) / 8);
    icp_scene_kernel<<<grid, block, 0, stream>>>(d_depth, W, H, K9[0], K9[2], K9[4], K9[5], d_pcd, d_normals);
    return cudaGetLastError();
}

cudaError_t launch_icp_point2plane(const IcpArgs& a, int n_clouds, cudaStream_t stream) {
    if (n_clouds <= 0) return cudaSuccess;
    if (a.nn) icp_point2plane_kernel<true><<<n_clouds, kIcpThreads, 0, stream>>>(a);
    else icp_point2plane_kernel<false><<<n_clouds, kIcpThreads, 0, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace perch
