// Measures |approximate - exact| screen coordinates of pass A's culling transform (pose_kernels.cu: s_apx / tri_hits)
// against the pinned exact transform (exact_math.cuh: transform_vertex), over random in-view vertices and poses.
// The culling margin kEps must stay well above the maximum this prints.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -Iperception_b200/csrc -Iinclude tools/approx_error.cu -o tools/approx_error.bin
#include <cstdio>
#include <cmath>
#include <cuda_runtime.h>
#include "exact_math.cuh"
using namespace perch;

__device__ unsigned hash(unsigned x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }
__device__ float urand(unsigned& s) { s = hash(s + 0x9e3779b9u); return (s >> 8) * (1.0f / 16777216.0f); }
__device__ float frcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

__global__ void measure(int n_poses, int verts_per_pose, int stride, float depth_lo, float depth_hi, float* max_err, unsigned long long* counted) {
    const int W = 640, H = 480;
    const float wf = W, hf = H, fx = 576.09757860f, fy = fx, cx = 321.06398107f, cy = 242.97676897f;
    const int p = blockIdx.x;
    if (p >= n_poses) return;
    unsigned seed = hash(p * 2654435761u + 17u);
    // compute_proj (renderer.cu:1386-1410)
    Mat4 P = {};
    P.a0 = 2 * fx / wf; P.a2 = 2 * cx / wf - 1; P.b1 = -2 * fy / hf; P.b2 = -(2 * cy / hf - 1);
    // random rotation (unit quaternion), translation inside the frustum at a log-uniform depth, cm units (x100)
    float q[4]; float n = 0;
    for (int i = 0; i < 4; ++i) { q[i] = urand(seed) * 2 - 1; n += q[i] * q[i]; }
    n = rsqrtf(n); for (int i = 0; i < 4; ++i) q[i] *= n;
    const float w = q[0], x = q[1], y = q[2], z = q[3];
    float R[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y), 2 * (x * y + w * z), 1 - 2 * (x * x + z * z),
                  2 * (y * z - w * x), 2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)};
    const float depth = depth_lo * powf(depth_hi / depth_lo, urand(seed));  // cm
    const float tx = (urand(seed) * 2 - 1) * depth * (0.5f * wf / fx), ty = (urand(seed) * 2 - 1) * depth * (0.5f * hf / fy);
    Mat4 pose = {100 * R[0], 100 * R[1], 100 * R[2], tx, 100 * R[3], 100 * R[4], 100 * R[5], ty, 100 * R[6], 100 * R[7], 100 * R[8], depth, 0, 0, 0, 1};
    const float inv_s = 1.0f / stride, half_w = 0.5f * wf, half_h = 0.5f * hf, row_phase = (float)((H - 1) % stride);
    const float kx = half_w * inv_s, ky = half_h * inv_s;
    float U[4], V[4], C[4];
    U[0] = kx * (P.a0 * pose.a0 + P.a1 * pose.b0); U[1] = kx * (P.a0 * pose.a1 + P.a1 * pose.b1);
    U[2] = kx * (P.a0 * pose.a2 + P.a1 * pose.b2); U[3] = kx * (P.a0 * pose.a3 + P.a1 * pose.b3 + P.a3);
    V[0] = ky * (P.b0 * pose.a0 + P.b1 * pose.b0); V[1] = ky * (P.b0 * pose.a1 + P.b1 * pose.b1);
    V[2] = ky * (P.b0 * pose.a2 + P.b1 * pose.b2); V[3] = ky * (P.b0 * pose.a3 + P.b1 * pose.b3 + P.b3);
    C[0] = pose.c0; C[1] = pose.c1; C[2] = pose.c2; C[3] = pose.c3;
    const float offx = (half_w + half_w * P.a2) * inv_s, offy = (half_h + half_h * P.b2 - row_phase) * inv_s;
    float worst = 0; unsigned long long cnt = 0;
    for (int i = threadIdx.x; i < verts_per_pose; i += blockDim.x) {
        unsigned s2 = hash(seed ^ (i * 747796405u));
        const float vx = (urand(s2) * 2 - 1) * 0.15f, vy = (urand(s2) * 2 - 1) * 0.15f, vz = (urand(s2) * 2 - 1) * 0.15f;
        ScreenVertex e = transform_vertex(pose, P, wf, hf, vx, vy, vz);
        if (!(e.z > 1e-3f) || !(e.sx >= -64.f && e.sx <= wf + 64.f && e.sy >= -64.f && e.sy <= hf + 64.f)) continue;
        const float n0 = fmaf(vz, U[2], fmaf(vy, U[1], fmaf(vx, U[0], U[3])));
        const float n1 = fmaf(vz, V[2], fmaf(vy, V[1], fmaf(vx, V[0], V[3])));
        const float iz = frcp_approx(fmaf(vz, C[2], fmaf(vy, C[1], fmaf(vx, C[0], C[3]))));
        const float ux = fmaf(n0, iz, offx), uy = fmaf(n1, iz, offy);
        const float ax = ux * stride, ay = uy * stride + row_phase;
        worst = fmaxf(worst, fmaxf(fabsf(ax - e.sx), fabsf(ay - e.sy)));
        ++cnt;
    }
    for (int o = 16; o >= 1; o >>= 1) { worst = fmaxf(worst, __shfl_xor_sync(0xffffffffu, worst, o)); cnt += __shfl_xor_sync(0xffffffffu, cnt, o); }
    if ((threadIdx.x & 31) == 0) { atomicMax((int*)max_err, __float_as_int(worst)); atomicAdd(counted, cnt); }
}

int main() {
    float* d_err; unsigned long long* d_cnt;
    cudaMalloc(&d_err, 4); cudaMalloc(&d_cnt, 8);
    const int strides[3] = {1, 4, 8};
    const float ranges[3][2] = {{20.f, 300.f}, {300.f, 3000.f}, {3000.f, 100000.f}};
    for (int si = 0; si < 3; ++si)
        for (int ri = 0; ri < 3; ++ri) {
            cudaMemset(d_err, 0, 4); cudaMemset(d_cnt, 0, 8);
            measure<<<65536, 128>>>(65536, 4096, strides[si], ranges[ri][0], ranges[ri][1], d_err, d_cnt);
            float err; unsigned long long cnt;
            cudaMemcpy(&err, d_err, 4, cudaMemcpyDeviceToHost); cudaMemcpy(&cnt, d_cnt, 8, cudaMemcpyDeviceToHost);
            printf("stride %d depth %.0f-%.0f cm: max |approx - exact| = %.6f px over %llu in-view vertices (%s)\n", strides[si],
                   ranges[ri][0], ranges[ri][1], err, cnt, cudaGetErrorString(cudaGetLastError()));
        }
    return 0;
}
