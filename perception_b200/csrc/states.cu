// Caller glue on the device.  Reference behaviour restated from sbpl_perception:
//   GetStateImagesUnifiedGPU   search_env.cpp:1505-1622   (pose matrices, pose_observed_points_total)
//   IsValidPose                search_env.cpp:309-528
//   GenerateSuccessorStates    search_env.cpp:7268-7472   (3-DoF grid)
//   ContPose::GetTransform     object_state.cpp:83-97
//   ObjectModel::GetFootprint / PointsInsideFootprint / isXYPointIn2DXYPolygonCustom   object_model.cpp:183-219, 571-651
//   Model::mat4x4::init_from_eigen   cuda_renderer model.h:89-107
// The reference evaluates these with Eigen (double) / PCL (float) on the host; every operation below is written with
// explicit round-to-nearest intrinsics in the reference's evaluation order so that no FMA contraction changes a bit.
// Eigen, PCL and FLANN are not part of the reference tree: the order of additions inside their products is restated
// from their published sources and the CPU oracle (oracle/perch_states_oracle.c) follows the same
// order with the host's libm for sin/cos.
#include <cfloat>
#include <cmath>

#include "states.cuh"

namespace perch {
namespace {

constexpr int kWarp = 32;

__device__ __forceinline__ double dmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double dsub(double a, double b) { return __dsub_rn(a, b); }

// Eigen's quaternion product (Geometry/Quaternion.h, generic quat_product); q = (w, x, y, z)
__device__ void quat_mul(const double a[4], const double b[4], double o[4]) {
    o[0] = dsub(dsub(dsub(dmul(a[0], b[0]), dmul(a[1], b[1])), dmul(a[2], b[2])), dmul(a[3], b[3]));
    o[1] = dsub(dadd(dadd(dmul(a[0], b[1]), dmul(a[1], b[0])), dmul(a[2], b[3])), dmul(a[3], b[2]));
    o[2] = dsub(dadd(dadd(dmul(a[0], b[2]), dmul(a[2], b[0])), dmul(a[3], b[1])), dmul(a[1], b[3]));
    o[3] = dsub(dadd(dadd(dmul(a[0], b[3]), dmul(a[3], b[0])), dmul(a[1], b[2])), dmul(a[2], b[1]));
}

// ContPose::GetTransform (object_state.cpp:83-97): rotation R (row-major 3x3) of the normalised quaternion
__device__ void state_rotation(const perch_state& s, double R[9]) {
    double q[4];
    if (s.qw == 0 && s.qx == 0 && s.qy == 0 && s.qz == 0) {
        // quaternion = yaw_angle * pitch_angle * roll_angle; AngleAxis -> Quaternion: w = cos(a/2), vec = sin(a/2)*axis
        double sr, cr, sp, cp, sy, cy;
        sincos(dmul(0.5, s.roll), &sr, &cr);
        sincos(dmul(0.5, s.pitch), &sp, &cp);
        sincos(dmul(0.5, s.yaw), &sy, &cy);
        // sin * unit axis: the zero components are +0 (sin * 0.0) or -0; their products vanish either way
        const double qy_[4] = {cy, dmul(sy, 0.0), dmul(sy, 0.0), sy};
        const double qp_[4] = {cp, dmul(sp, 0.0), sp, dmul(sp, 0.0)};
        const double qr_[4] = {cr, sr, dmul(sr, 0.0), dmul(sr, 0.0)};
        double t[4];
        quat_mul(qy_, qp_, t);
        quat_mul(t, qr_, q);
    } else {
        q[0] = s.qw; q[1] = s.qx; q[2] = s.qy; q[3] = s.qz;
    }
    // normalize(): coeffs (x, y, z, w) / sqrt(x^2 + y^2 + z^2 + w^2)
    const double n2 = dadd(dadd(dadd(dmul(q[1], q[1]), dmul(q[2], q[2])), dmul(q[3], q[3])), dmul(q[0], q[0]));
    const double n = __dsqrt_rn(n2);
    const double w = __ddiv_rn(q[0], n), x = __ddiv_rn(q[1], n), y = __ddiv_rn(q[2], n), z = __ddiv_rn(q[3], n);
    // QuaternionBase::toRotationMatrix
    const double tx = dmul(2.0, x), ty = dmul(2.0, y), tz = dmul(2.0, z);
    const double twx = dmul(tx, w), twy = dmul(ty, w), twz = dmul(tz, w);
    const double txx = dmul(tx, x), txy = dmul(ty, x), txz = dmul(tz, x);
    const double tyy = dmul(ty, y), tyz = dmul(tz, y), tzz = dmul(tz, z);
    R[0] = dsub(1.0, dadd(tyy, tzz)); R[1] = dsub(txy, twz);            R[2] = dadd(txz, twy);
    R[3] = dadd(txy, twz);            R[4] = dsub(1.0, dadd(txx, tzz)); R[5] = dsub(tyz, twx);
    R[6] = dsub(txz, twy);            R[7] = dadd(tyz, twx);            R[8] = dsub(1.0, dadd(txx, tyy));
}

// 4x4 product, coefficient (i, j) = ((a_i0*b_0j + a_i1*b_1j) + a_i2*b_2j) + a_i3*b_3j
__device__ void mat4_mul(const double* a, const double* b, double* o) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
            o[4 * i + j] = dadd(dadd(dadd(dmul(a[4 * i], b[j]), dmul(a[4 * i + 1], b[4 + j])), dmul(a[4 * i + 2], b[8 + j])),
                                dmul(a[4 * i + 3], b[12 + j]));
}

__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// pcl::search::KdTree::radiusSearch (FLANN kd-tree, L2_Simple<float>) as a count: points of `range` with
//   ((dx*dx) + dy*dy) + dz*dz < r2        (float, no contraction; r2 = float(radius*radius))
// All lanes of the warp call this with the same arguments; every lane receives the total.
__device__ int radius_count(const SearchGridDev& g, int range, float qx, float qy, float qz, float r2, float r) {
    if (range < 0 || range >= g.n_ranges || g.pts == nullptr) return 0;
    const int lane = threadIdx.x & (kWarp - 1);
    const float fx0 = (qx - r - g.x0) * g.inv_cell - 1.f, fx1 = (qx + r - g.x0) * g.inv_cell + 1.f;
    const float fy0 = (qy - r - g.y0) * g.inv_cell - 1.f, fy1 = (qy + r - g.y0) * g.inv_cell + 1.f;
    int ix0 = 0, ix1 = g.nx - 1, iy0 = 0, iy1 = g.ny - 1;
    if (fx0 == fx0 && fx1 == fx1 && fy0 == fy0 && fy1 == fy1) {  // NaN anywhere: scan everything (the test fails anyway)
        ix0 = (int)fminf(fmaxf(floorf(fx0), 0.f), (float)(g.nx - 1));
        ix1 = (int)fminf(fmaxf(floorf(fx1), 0.f), (float)(g.nx - 1));
        iy0 = (int)fminf(fmaxf(floorf(fy0), 0.f), (float)(g.ny - 1));
        iy1 = (int)fminf(fmaxf(floorf(fy1), 0.f), (float)(g.ny - 1));
    }
    int cnt = 0;
    for (int iy = iy0; iy <= iy1; ++iy) {
        const int* row = g.cell_start + ((size_t)range * g.ny + iy) * g.nx;
        const int b = __ldg(row + ix0), e = __ldg(row + ix1 + 1);
        for (int i = b + lane; i < e; i += kWarp) {
            const float4 p = __ldg(g.pts + i);
            const float dx = __fsub_rn(qx, p.x), dy = __fsub_rn(qy, p.y), dz = __fsub_rn(qz, p.z);
            const float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            cnt += (d < r2) ? 1 : 0;
        }
    }
    return warp_sum(cnt);
}

// radiusSearch's max_nn rule (pcl kdtree_flann.hpp): 0 or more than the cloud => the cloud size
__device__ __forceinline__ int cap_max_nn(int count, int max_nn, int total) {
    if (max_nn <= 0 || max_nn > total) max_nn = total;
    return count < max_nn ? count : max_nn;
}

// ---- grid build --------------------------------------------------------------------------------------------------
__global__ void grid_key_kernel(const float* xyz, int n, const int* ranges, int n_ranges, float x0, float y0, float inv_cell,
                                int nx, int ny, int* key, int* hist) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int r = n_ranges;  // points outside every range go to a bucket behind the last range
    for (int k = 0; k < n_ranges; ++k)
        if (i >= ranges[2 * k] && i < ranges[2 * k + 1]) { r = k; break; }
    int k = n_ranges * ny * nx;
    if (r < n_ranges) {
        const float x = xyz[3 * (size_t)i], y = xyz[3 * (size_t)i + 1];
        const int ix = (int)fminf(fmaxf(floorf((x - x0) * inv_cell), 0.f), (float)(nx - 1));
        const int iy = (int)fminf(fmaxf(floorf((y - y0) * inv_cell), 0.f), (float)(ny - 1));
        k = (r * ny + iy) * nx + ix;
    }
    key[i] = k;
    atomicAdd(hist + k, 1);
}

// exclusive scan of `n` ints in place by one block (chunks of blockDim.x with a running carry); data[n] = total
__global__ void scan_kernel(int* data, long long n) {
    __shared__ int warp_tot[32];
    __shared__ int carry_s;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (long long base = 0; base < n; base += blockDim.x) {
        const long long i = base + tid;
        const int v = i < n ? data[i] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) warp_tot[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            int w = lane < (int)(blockDim.x >> 5) ? warp_tot[lane] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += t;
            }
            warp_tot[lane] = w;  // inclusive over warps
        }
        __syncthreads();
        const int carry = carry_s;
        const int before = carry + (wid > 0 ? warp_tot[wid - 1] : 0) + inc - v;
        if (i < n) data[i] = before;
        __syncthreads();
        if (tid == blockDim.x - 1) carry_s = carry + warp_tot[(blockDim.x >> 5) - 1];
        __syncthreads();
    }
    if (tid == 0) data[n] = carry_s;
}

__global__ void grid_scatter_kernel(const float* xyz, int n, const int* key, const int* cell_start, int* cursor, float4* sorted,
                                    int n_keys) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int k = key[i];
    if (k >= n_keys) return;
    const int pos = cell_start[k] + atomicAdd(cursor + k, 1);
    sorted[pos] = make_float4(xyz[3 * (size_t)i], xyz[3 * (size_t)i + 1], xyz[3 * (size_t)i + 2], 0.f);
}

__global__ void range_size_kernel(const int* ranges, int n_ranges, int* range_size) {
    const int k = threadIdx.x;
    if (k < n_ranges) range_size[k] = max(ranges[2 * k + 1] - ranges[2 * k], 0);
}

// ---- assemble ------------------------------------------------------------------------------------------------------
// One warp per state: lane 0 does the double-precision matrix chain, the warp shares the cylinder count.
__global__ void assemble_kernel(StateArgs a) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) / kWarp;
    const int lane = threadIdx.x & (kWarp - 1);
    if (w >= a.n_states) return;
    const perch_state s = a.states[w];
    const bool model_ok = s.model_id >= 0 && s.model_id < a.env.n_models;
    if (!model_ok && lane == 0) atomicOr(a.err_flag, 2);
    if (lane == 0) {
        float* out = a.poses16 + 16 * (size_t)w;
        if (model_ok) {
            double R[9], T[16], M[16], Pc[16];
            state_rotation(s, R);
            T[0] = R[0]; T[1] = R[1]; T[2] = R[2]; T[3] = s.x;
            T[4] = R[3]; T[5] = R[4]; T[6] = R[5]; T[7] = s.y;
            T[8] = R[6]; T[9] = R[7]; T[10] = R[8]; T[11] = s.z;
            T[12] = 0.0; T[13] = 0.0; T[14] = 0.0; T[15] = 1.0;
            mat4_mul(T, a.metas[s.model_id].pre, M);   // transform * preprocess_transform (search_env.cpp:1567)
            mat4_mul(a.env.cam_matrix, M, Pc);         // cam_matrix * transform (1571)
            // init_from_eigen(pose_in_cam, 100): float(double * int) on the top three rows (model.h:89-107)
#pragma unroll
            for (int k = 0; k < 12; ++k) out[k] = (float)dmul(Pc[k], 100.0);
#pragma unroll
            for (int k = 12; k < 16; ++k) out[k] = (float)Pc[k];
        } else {
            for (int k = 0; k < 16; ++k) out[k] = 0.f;
        }
        a.model_map[w] = s.model_id;
    }
    float total = a.env.observed_point_num;
    if (a.env.use_external_pose_list == 1) {
        const int required_object_id = s.segmentation_label_id - 1;  // search_env.cpp:1582
        const bool ok = required_object_id >= 0 && required_object_id < a.env.n_labels;
        if (!ok && lane == 0) atomicOr(a.err_flag, 4);
        total = ok ? a.seg_counts[required_object_id] : 0.f;
        if (lane == 0) a.label[w] = required_object_id;
    } else {
        if (lane == 0) a.label[w] = -1;
        if (a.env.use_cylinder_observed && model_ok) {
            // search_env.cpp:1593-1613: points of the projected cloud within inflation * circumscribed radius
            const ModelMetaDev& m = a.metas[s.model_id];
            const double rad = dmul(m.inflation, m.circ);
            const float r2 = (float)dmul(rad, rad);
            const float rf = (float)rad * 1.00001f + 1e-30f;
            const int cnt = radius_count(a.grid, 0, (float)s.x, (float)s.y, (float)s.z, r2, rf);
            const int tot0 = a.grid.n_ranges > 0 ? a.grid.range_size[0] : 0;
            total = (float)cap_max_nn(cnt, a.env.max_nn_pixels, tot0);
        }
    }
    if (lane == 0) a.obs_total[w] = total;
}

// ---- IsValidPose ---------------------------------------------------------------------------------------------------
// Shared by the list validator and the grid generator.  Rf = pose.GetTransform().matrix().cast<float>() rotation part,
// (tx, ty, tz) its translation.  All lanes call with the same arguments.
struct PoseF {
    float r[9];
    float tx, ty, tz;
};

// search_env.cpp:484-499 with GetFootprint (object_model.cpp:637-651) and pcl::transformPointCloud's
// x' = ((m00*x + m01*y) + m02*z) + m03
__device__ bool footprint_in_bounds(const StateArgs& a, const ModelMetaDev& m, const PoseF& p) {
    const int lane = threadIdx.x & (kWarp - 1);
    const double lo_x = dsub(a.env.x_min, a.env.footprint_tolerance), hi_x = dadd(a.env.x_max, a.env.footprint_tolerance);
    const double lo_y = dsub(a.env.y_min, a.env.footprint_tolerance), hi_y = dadd(a.env.y_max, a.env.footprint_tolerance);
    bool bad = false;
    for (int i = lane; i < m.fp_cnt; i += kWarp) {
        const float* h = a.footprints + 3 * (size_t)(m.fp_off + i);
        const float hx = h[0], hy = h[1], hz = h[2];
        const float x = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(p.r[0], hx), __fmul_rn(p.r[1], hy)), __fmul_rn(p.r[2], hz)), p.tx);
        const float y = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(p.r[3], hx), __fmul_rn(p.r[4], hy)), __fmul_rn(p.r[5], hz)), p.ty);
        bad |= ((double)x < lo_x) || ((double)x > hi_x) || ((double)y < lo_y) || ((double)y > hi_y);
    }
    return !__any_sync(0xffffffffu, bad);
}

// search_env.cpp:504-525 with PointsInsideFootprint (object_model.cpp:571-623): constraint points are taken into the
// model frame by the inverse of the inflated pose (Eigen Affine3f::inverse(): cofactor inverse of the linear part, then
// -inv * t) and tested against the un-transformed hull with isXYPointIn2DXYPolygonCustom (183-219).
__device__ bool constraint_ok(const StateArgs& a, const ModelMetaDev& m, const PoseF& p) {
    if (a.n_constraint <= 0) return true;
    const int lane = threadIdx.x & (kWarp - 1);
    const float f = (float)m.inflation;
    float L[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) L[k] = __fmul_rn(f, p.r[k]);
    // Eigen compute_inverse_size3: cofactor(i, j) = m(i1,j1)*m(i2,j2) - m(i1,j2)*m(i2,j1), inverse(i, j) = cofactor(j, i) / det
    auto cof = [&](int i, int j) {
        const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
        return __fsub_rn(__fmul_rn(L[3 * i1 + j1], L[3 * i2 + j2]), __fmul_rn(L[3 * i1 + j2], L[3 * i2 + j1]));
    };
    const float c00 = cof(0, 0), c10 = cof(1, 0), c20 = cof(2, 0);
    const float det = __fadd_rn(__fadd_rn(__fmul_rn(c00, L[0]), __fmul_rn(c10, L[3])), __fmul_rn(c20, L[6]));
    const float invdet = __fdiv_rn(1.0f, det);
    float I[9];
    I[0] = __fmul_rn(c00, invdet); I[1] = __fmul_rn(c10, invdet); I[2] = __fmul_rn(c20, invdet);
    I[3] = __fmul_rn(cof(0, 1), invdet); I[4] = __fmul_rn(cof(1, 1), invdet); I[5] = __fmul_rn(cof(2, 1), invdet);
    I[6] = __fmul_rn(cof(0, 2), invdet); I[7] = __fmul_rn(cof(1, 2), invdet); I[8] = __fmul_rn(cof(2, 2), invdet);
    // translation of the inverse: -(I * t)
    const float itx = -__fadd_rn(__fadd_rn(__fmul_rn(I[0], p.tx), __fmul_rn(I[1], p.ty)), __fmul_rn(I[2], p.tz));
    const float ity = -__fadd_rn(__fadd_rn(__fmul_rn(I[3], p.tx), __fmul_rn(I[4], p.ty)), __fmul_rn(I[5], p.tz));
    int inside = 0;
    for (int i = lane; i < a.n_constraint; i += kWarp) {
        // homogeneous_point << x, y, 0, 1 (floats); inverse_transform * homogeneous_point, columns accumulated in order
        const float cx = a.constraint[3 * (size_t)i], cy = a.constraint[3 * (size_t)i + 1];
        const float px = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(I[0], cx), __fmul_rn(I[1], cy)), __fmul_rn(I[2], 0.f)), itx);
        const float py = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(I[3], cx), __fmul_rn(I[4], cy)), __fmul_rn(I[5], 0.f)), ity);
        // transformed_points are Vector2d, then narrowed into a PointT again: the values stay floats
        const double qx = (double)px, qy = (double)py;
        bool in_poly = false;
        const int nv = m.fp_cnt;
        if (nv > 0) {
            const float* hull = a.footprints + 3 * (size_t)m.fp_off;
            double xold = (double)hull[3 * (nv - 1)], yold = (double)hull[3 * (nv - 1) + 1];
            for (int v = 0; v < nv; ++v) {
                const double xnew = (double)hull[3 * v], ynew = (double)hull[3 * v + 1];
                double x1, x2, y1, y2;
                if (xnew > xold) { x1 = xold; x2 = xnew; y1 = yold; y2 = ynew; }
                else { x1 = xnew; x2 = xold; y1 = ynew; y2 = yold; }
                if ((xnew < qx) == (qx <= xold) && dmul(dsub(qy, y1), dsub(x2, x1)) < dmul(dsub(y2, y1), dsub(qx, x1)))
                    in_poly = !in_poly;
                xold = xnew; yold = ynew;
            }
        }
        inside += in_poly ? 1 : 0;
    }
    inside = warp_sum(inside);
    const double min_points = (double)min(a.env.min_points_for_constraint_cloud, a.n_constraint);
    return !((double)inside < min_points);
}

// search_env.cpp:331-404: neighbour count inside the search radius
__device__ bool enough_neighbours(const StateArgs& a, const ModelMetaDev& m, float px, float py, float pz, int seg_label_id) {
    // grid_cell_circumscribing_radius = hypot(search_resolution / 2, search_resolution / 2) (344-349), evaluated once
    // per model on the host (perch_set_env) so that libm's hypot is the one the reference uses
    const double cell_rad = a.after_refinement ? 0.0 : m.cell_rad;
    double search_rad;
    int range = 0;
    if (a.env.use_external_pose_list != 1) {
        search_rad = fmax(m.circ, cell_rad);
    } else {
        search_rad = fmax(dmul(m.inflation, m.circ3d), cell_rad);
        const int required_object_id = seg_label_id - 1;   // GenerateSuccessorStates passes the label index (7136-7146)
        range = required_object_id >= 0 ? 1 + required_object_id : 0;
    }
    if (a.env.min_neighbor_points <= 0) return true;
    const float r2 = (float)dmul(search_rad, search_rad);
    const float rf = (float)search_rad * 1.00001f + 1e-30f;
    const int cnt = radius_count(a.grid, range, px, py, pz, r2, rf);
    return cnt >= a.env.min_neighbor_points;
}

// search_env.cpp:455-474: inscribed circles of placed objects
__device__ bool clear_of_placed(const StateArgs& a, const ModelMetaDev& m, double x, double y) {
    if (a.env.use_external_pose_list == 1) return true;
    bool hit = false;
    const int lane = threadIdx.x & (kWarp - 1);
    for (int i = lane; i < a.n_placed; i += kWarp) {
        const perch_state o = a.placed[i];
        if (o.model_id < 0 || o.model_id >= a.env.n_models) continue;
        const double rad = dadd(m.insc, a.metas[o.model_id].insc);
        const double dx = dsub(x, o.x), dy = dsub(y, o.y);
        hit |= dadd(dmul(dx, dx), dmul(dy, dy)) < dmul(rad, rad);
    }
    return !__any_sync(0xffffffffu, hit);
}

__device__ void pose_to_float(const perch_state& s, PoseF& p) {
    double R[9];
    state_rotation(s, R);
#pragma unroll
    for (int k = 0; k < 9; ++k) p.r[k] = (float)R[k];
    p.tx = (float)s.x; p.ty = (float)s.y; p.tz = (float)s.z;
}

__global__ void validate_kernel(StateArgs a) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) / kWarp;
    const int lane = threadIdx.x & (kWarp - 1);
    if (w >= a.n_states) return;
    const perch_state s = a.states[w];
    bool ok = s.model_id >= 0 && s.model_id < a.env.n_models;
    if (!ok) {
        if (lane == 0) { a.valid[w] = 0; atomicOr(a.err_flag, 2); }
        return;
    }
    const ModelMetaDev& m = a.metas[s.model_id];
    ok = enough_neighbours(a, m, (float)s.x, (float)s.y, (float)s.z, s.segmentation_label_id);
    if (ok) ok = clear_of_placed(a, m, s.x, s.y);
    if (ok && (a.env.use_external_pose_list != 1 || a.n_constraint > 0)) {
        PoseF p;
        pose_to_float(s, p);
        if (a.env.use_external_pose_list != 1) ok = footprint_in_bounds(a, m, p);
        if (ok) ok = constraint_ok(a, m, p);
    }
    if (lane == 0) a.valid[w] = ok ? 1 : 0;
}

// ---- 3-DoF successor generation ------------------------------------------------------------------------------------
__device__ __forceinline__ void locate_cell(const GenerateArgs& g, long long cell, int& gm, int& ix, int& iy) {
    gm = 0;
    while (gm + 1 < g.n_gen_models && cell >= g.cell_off[gm + 1]) ++gm;
    const long long local = cell - g.cell_off[gm];
    const int nyg = g.y_off[gm + 1] - g.y_off[gm];
    ix = (int)(local / nyg);
    iy = (int)(local % nyg);
}

// One warp per (model, x, y) cell: IsValidPose for every yaw, then the loop rules of search_env.cpp:7283-7469.
__global__ void generate_flags_kernel(GenerateArgs g) {
    const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) / kWarp;
    const int lane = threadIdx.x & (kWarp - 1);
    if (w >= g.n_cells) return;
    const StateArgs& a = g.base;
    int gm, ix, iy;
    locate_cell(g, w, gm, ix, iy);
    const int model = g.gen_models[gm];
    const ModelMetaDev& m = a.metas[model];
    const double x = g.xs[g.x_off[gm] + ix], y = g.ys[g.y_off[gm] + iy], z = a.env.table_height;
    uint8_t* emit = g.emit + (size_t)w * g.n_thetas;
    bool cell_ok = enough_neighbours(a, m, (float)x, (float)y, (float)z, -1);
    if (cell_ok) cell_ok = clear_of_placed(a, m, x, y);
    int count = 0;
    bool stop = false;
    for (int t = 0; t < g.n_thetas; ++t) {
        bool e = false;
        if (!stop && cell_ok) {
            PoseF p;  // rotation of ContPose(x, y, z, 0, 0, theta), shared by every cell (theta_rot_kernel)
#pragma unroll
            for (int k = 0; k < 9; ++k) p.r[k] = g.theta_rot[9 * t + k];
            p.tx = (float)x; p.ty = (float)y; p.tz = (float)z;
            bool ok = footprint_in_bounds(a, m, p);
            if (ok) ok = constraint_ok(a, m, p);
            if (ok) {
                // 7302-7306: semi-symmetric objects stop above pi + theta_res (checked after IsValidPose, before adding)
                if (m.symmetry_mode == 1 && g.thetas[t] > dadd(M_PI, g.theta_res)) {
                    stop = true;
                } else {
                    e = true;
                    // 7466-7468: rotationally symmetric objects keep the first valid yaw only
                    if (m.symmetric || m.symmetry_mode == 2) stop = true;
                }
            }
        }
        if (lane == 0) emit[t] = e ? 1 : 0;
        count += e ? 1 : 0;
    }
    if (lane == 0) g.cell_count[w] = count;
}

__global__ void theta_rot_kernel(const double* yaws, int n, float* rot) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    perch_state s;
    s.x = s.y = s.z = 0.0; s.roll = 0.0; s.pitch = 0.0; s.yaw = yaws[t];
    s.qx = s.qy = s.qz = s.qw = 0.0;
    PoseF p;
    pose_to_float(s, p);
    for (int k = 0; k < 9; ++k) rot[9 * t + k] = p.r[k];
}

__global__ void generate_write_kernel(GenerateArgs g) {
    const long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= g.n_cells) return;
    int gm, ix, iy;
    locate_cell(g, w, gm, ix, iy);
    const uint8_t* emit = g.emit + (size_t)w * g.n_thetas;
    int pos = g.cell_count[w];
    perch_state s;
    s.x = g.xs[g.x_off[gm] + ix]; s.y = g.ys[g.y_off[gm] + iy]; s.z = g.base.env.table_height;
    s.roll = 0.0; s.pitch = 0.0;
    s.qx = s.qy = s.qz = s.qw = 0.0;
    s.model_id = g.gen_models[gm];
    s.segmentation_label_id = -1;
    for (int t = 0; t < g.n_thetas; ++t)
        if (emit[t]) {
            s.yaw = g.yaws[t];
            g.out[pos++] = s;
        }
}

}  // namespace

cudaError_t launch_grid_build(const float* d_xyz, int n, const int* d_ranges, int n_ranges, float x0, float y0,
                              float inv_cell, int nx, int ny, int* d_point_key, int* d_cell_start, int* d_cursor,
                              float4* d_sorted, int* d_range_size, cudaStream_t stream) {
    const int n_keys = n_ranges * ny * nx;
    cudaError_t e;
    if ((e = cudaMemsetAsync(d_cell_start, 0, ((size_t)n_keys + 2) * sizeof(int), stream))) return e;
    if ((e = cudaMemsetAsync(d_cursor, 0, ((size_t)n_keys + 1) * sizeof(int), stream))) return e;
    if (n > 0) grid_key_kernel<<<(n + 255) / 256, 256, 0, stream>>>(d_xyz, n, d_ranges, n_ranges, x0, y0, inv_cell, nx, ny, d_point_key, d_cell_start);
    scan_kernel<<<1, 1024, 0, stream>>>(d_cell_start, (long long)n_keys + 1);
    if (n > 0) grid_scatter_kernel<<<(n + 255) / 256, 256, 0, stream>>>(d_xyz, n, d_point_key, d_cell_start, d_cursor, d_sorted, n_keys);
    range_size_kernel<<<1, 64, 0, stream>>>(d_ranges, n_ranges, d_range_size);
    return cudaGetLastError();
}

cudaError_t launch_assemble_states(const StateArgs& a, cudaStream_t stream) {
    if (a.n_states <= 0) return cudaSuccess;
    const int threads = 128, per_block = threads / kWarp;
    assemble_kernel<<<(a.n_states + per_block - 1) / per_block, threads, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_validate_states(const StateArgs& a, cudaStream_t stream) {
    if (a.n_states <= 0) return cudaSuccess;
    const int threads = 128, per_block = threads / kWarp;
    validate_kernel<<<(a.n_states + per_block - 1) / per_block, threads, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_generate_states(const GenerateArgs& g, cudaStream_t stream) {
    if (g.n_cells <= 0) return cudaMemsetAsync(g.cell_count, 0, sizeof(int), stream);
    const int threads = 128, per_block = threads / kWarp;
    theta_rot_kernel<<<(g.n_thetas + 63) / 64, 64, 0, stream>>>(g.yaws, g.n_thetas, g.theta_rot);
    generate_flags_kernel<<<(unsigned)((g.n_cells + per_block - 1) / per_block), threads, 0, stream>>>(g);
    scan_kernel<<<1, 1024, 0, stream>>>(g.cell_count, g.n_cells);
    return cudaGetLastError();
}

cudaError_t launch_generate_write(const GenerateArgs& g, cudaStream_t stream) {
    if (g.n_cells <= 0) return cudaSuccess;
    generate_write_kernel<<<(unsigned)((g.n_cells + 127) / 128), 128, 0, stream>>>(g);
    return cudaGetLastError();
}

}  // namespace perch
