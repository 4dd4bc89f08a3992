// Float arithmetic of the reference rasterizer / back-projection, pinned operation by operation.
//
// The reference (cuda_renderer/include/cuda_renderer/cuda/image_renderer.cuh:21-57,79-129 and
// compute_point_clouds.cuh:20-22) is built with nvcc's default -fmad=true (cuda_renderer/CMakeLists.txt:27),
// so which a*b+c become FMAs is a property of the compiler, not of the source: NVVM fuses some (visible in PTX)
// and ptxas fuses more (`mul.f32`/`add.f32` without .rn may be contracted; visible only in SASS).  The sequences
// below are the ones in the sm_100a SASS nvcc 12.9 -O3 produces for those expressions (oracle/_ref/libperch_ref.so,
// `cuobjdump -sass`; see DESIGN.md "arithmetic contract"); they are written with round-to-nearest intrinsics so
// that this file compiles to the same operations regardless of surrounding code, inlining or flags.  Integer depth
// parity depends on every one (tests: frag_single_tri golden, one triangle per image => no z-buffer contention).
#pragma once
#include <cfloat>
#include <climits>
#include <cstdint>

namespace perch {

struct Mat4 {  // row-major a0..d3 like cuda_renderer::Model::mat4x4 (model.h:74-78)
    float a0, a1, a2, a3, b0, b1, b2, b3, c0, c1, c2, c3, d0, d1, d2, d3;
};

// mat_mul_v row (image_renderer.cuh:21-27): m0*x + m1*y + m2*z + m3  ==>  m3 + fma(m2,z, fma(m0,x, m1*y))
__device__ __forceinline__ float xf_row(float m0, float m1, float m2, float m3, float x, float y, float z) {
    return __fadd_rn(m3, __fmaf_rn(m2, z, __fmaf_rn(m0, x, __fmul_rn(m1, y))));
}

struct ScreenVertex {
    float sx, sy;  // viewport coordinates (image_renderer.cuh:82-84), y not yet flipped
    float z;       // camera-space depth = the reference's `last_row` entry (image_renderer.cuh:299-303)
};

// Model -> camera (pose) -> clip (proj rows a,b) -> viewport.
// viewport: x/z*W/2 + W/2  ==>  fma(Wf, 0.5, ((x/z)*Wf)*0.5)
__device__ __forceinline__ ScreenVertex transform_vertex(const Mat4& pose, const Mat4& proj, float wf, float hf,
                                                         float x, float y, float z) {
    float X = xf_row(pose.a0, pose.a1, pose.a2, pose.a3, x, y, z);
    float Y = xf_row(pose.b0, pose.b1, pose.b2, pose.b3, x, y, z);
    float Z = xf_row(pose.c0, pose.c1, pose.c2, pose.c3, x, y, z);
    float px = xf_row(proj.a0, proj.a1, proj.a2, proj.a3, X, Y, Z);
    float py = xf_row(proj.b0, proj.b1, proj.b2, proj.b3, X, Y, Z);
    ScreenVertex v;
    v.sx = __fmaf_rn(wf, 0.5f, __fmul_rn(__fmul_rn(__fdiv_rn(px, Z), wf), 0.5f));
    v.sy = __fmaf_rn(hf, 0.5f, __fmul_rn(__fmul_rn(__fdiv_rn(py, Z), hf), 0.5f));
    v.z = Z;
    return v;
}

__device__ __forceinline__ float sel_min(float a, float b) { return (a < b) ? a : b; }  // std__min (image_renderer.cuh:17-18)
__device__ __forceinline__ float sel_max(float a, float b) { return (a > b) ? a : b; }  // std__max (image_renderer.cuh:14-15)

struct TriSetup {
    float A0, A1;              // screen vertex 0
    float e_c0, e_b1, e_b0, e_c1;  // C0-A0, B1-A1, B0-A0, C1-A1
    float base_inv;            // 1 / signed area
    float z0, z1, z2;
    float bminx, bmaxx, bminy, bmaxy;  // clamped float bbox (image_renderer.cuh:86-107)
};

// bbox clamp + edge setup for one triangle.  clamp_max = (W-1, H-1), clamp_min = (0,0) (no ROI: image_render passes
// roi = {0,0,0,0}, image_renderer.cuh:369).
__device__ __forceinline__ void bbox_of(const ScreenVertex& a, const ScreenVertex& b, const ScreenVertex& c, float cmaxx,
                                        float cmaxy, float& bminx, float& bmaxx, float& bminy, float& bmaxy) {
    float mnx = FLT_MAX, mny = FLT_MAX, mxx = -FLT_MAX, mxy = -FLT_MAX;
    mnx = sel_max(0.f, sel_min(mnx, a.sx)); mxx = sel_min(cmaxx, sel_max(mxx, a.sx));
    mny = sel_max(0.f, sel_min(mny, a.sy)); mxy = sel_min(cmaxy, sel_max(mxy, a.sy));
    mnx = sel_max(0.f, sel_min(mnx, b.sx)); mxx = sel_min(cmaxx, sel_max(mxx, b.sx));
    mny = sel_max(0.f, sel_min(mny, b.sy)); mxy = sel_min(cmaxy, sel_max(mxy, b.sy));
    mnx = sel_max(0.f, sel_min(mnx, c.sx)); mxx = sel_min(cmaxx, sel_max(mxx, c.sx));
    mny = sel_max(0.f, sel_min(mny, c.sy)); mxy = sel_min(cmaxy, sel_max(mxy, c.sy));
    bminx = mnx; bmaxx = mxx; bminy = mny; bmaxy = mxy;
}

__device__ __forceinline__ void setup_edges(const ScreenVertex& a, const ScreenVertex& b, const ScreenVertex& c,
                                            TriSetup& t) {
    t.A0 = a.sx; t.A1 = a.sy;
    t.e_c0 = __fsub_rn(c.sx, a.sx);
    t.e_b1 = __fsub_rn(b.sy, a.sy);
    t.e_b0 = __fsub_rn(b.sx, a.sx);
    t.e_c1 = __fsub_rn(c.sy, a.sy);
    // calculateSignedArea(A,B,C) (image_renderer.cuh:40-42): 0.5*((C0-A0)*(B1-A1) - (B0-A0)*(C1-A1));
    // SASS: FMUL t = (B0-A0)*(C1-A1); FFMA (C0-A0)*(B1-A1) - t; FMUL 0.5
    float area = __fmul_rn(__fmaf_rn(t.e_c0, t.e_b1, -__fmul_rn(t.e_b0, t.e_c1)), 0.5f);
    t.base_inv = __frcp_rn(area);  // `1/area` (image_renderer.cuh:48)
    t.z0 = a.z; t.z1 = b.z; t.z2 = c.z;
}

// One fragment at integer pixel (P0,P1) (P1 in the un-flipped raster row).  Returns false if outside.
// image_renderer.cuh:44-57 (barycentric), 114-129 (inside test, perspective depth, rounding).
__device__ __forceinline__ bool fragment_depth(const TriSetup& t, float P0f, float P1f, int32_t& depth) {
    float dy = __fsub_rn(P1f, t.A1);
    float r0 = __fmul_rn(t.e_c0, dy);
    float r1 = __fmul_rn(dy, t.e_b0);
    float dx = __fsub_rn(P0f, t.A0);
    // SASS: FFMA (C1-A1)*(-dx) + r0 ; FFMA (B1-A1)*dx - r1 ; each then FMUL 0.5, FMUL base_inv
    float beta = __fmul_rn(t.base_inv, __fmul_rn(__fmaf_rn(t.e_c1, -dx, r0), 0.5f));
    float gamma = __fmul_rn(t.base_inv, __fmul_rn(__fmaf_rn(t.e_b1, dx, -r1), 0.5f));
    float alpha = __fsub_rn(__fsub_rn(1.0f, beta), gamma);
    float mn = fminf(gamma, fminf(beta, alpha));
    float mx = fmaxf(fmaxf(alpha, beta), gamma);
    if (mn < 0.0f || mx > 1.0f) return false;
    float num = __fadd_rn(gamma, __fadd_rn(beta, alpha));
    float den = __fadd_rn(__fdiv_rn(gamma, t.z2), __fadd_rn(__fdiv_rn(beta, t.z1), __fdiv_rn(alpha, t.z0)));
    depth = (int32_t)__fadd_rn(__fdiv_rn(num, den), 0.5f);  // cvt.rzi.s32.f32: saturating, NaN -> 0
    return true;
}

// Source-occlusion rule applied once per pixel to the final min depth d (> source means the observed scene is in
// front).  Equivalent to the reference's per-fragment test under any serial order (image_renderer.cuh:161-186;
// see DESIGN.md "canonical raster semantics"):
//   3-DoF (no pose labels):  (float)|d-s| > occlusion_threshold  &&  d > s  &&  s > 0
//   6-DoF (pose labels):     pose_label != source_label-1  &&  |d-s| > 0.5  &&  d > s  &&  s > 0
__device__ __forceinline__ bool source_occludes(int32_t d, int32_t s, bool use_label, int pose_label, int src_label,
                                                float occlusion_threshold) {
    if (!(d > s && s > 0)) return false;
    int ad = abs(d - s);
    if (use_label) return (pose_label != src_label - 1) && (ad >= 1);
    return (float)ad > occlusion_threshold;
}

// compute_point_clouds.cuh:20-22: z = float(d)/depth_factor; x = (float(u)-cx)/fx*z; y = (float(v)-cy)/fy*z  (no FMA)
__device__ __forceinline__ void back_project(int u, int v, int32_t d, float cx, float cy, float fx, float fy,
                                             float depth_factor, float& x, float& y, float& z) {
    z = __fdiv_rn((float)d, depth_factor);
    x = __fmul_rn(__fdiv_rn(__fsub_rn((float)u, cx), fx), z);
    y = __fmul_rn(__fdiv_rn(__fsub_rn((float)v, cy), fy), z);
}

// Squared L2 in the convention of oracle knn (fast_gicp boundary is unpinned): fma(dz,dz, fma(dy,dy, dx*dx)).
__device__ __forceinline__ float sqdist3(float px, float py, float pz, float qx, float qy, float qz) {
    float dx = __fsub_rn(px, qx), dy = __fsub_rn(py, qy), dz = __fsub_rn(pz, qz);
    return __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
}

}  // namespace perch
