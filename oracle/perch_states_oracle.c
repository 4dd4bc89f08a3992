/*
 * perch_states_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into libperch_b200.so).
 *
 * CPU restatement of the caller glue around render_cuda_multi_unified in sbpl_perception (reference paths relative to
 * /root/reference/sbpl_perception):
 *   oracle_camera_matrix      src/search_env.cpp:1535-1541  (cam_to_world * cam_to_body).matrix().inverse()
 *   oracle_assemble_states    src/search_env.cpp:1555-1622  per-pose matrix, model map, labels, observed totals
 *   oracle_validate_states    src/search_env.cpp:309-528    IsValidPose
 *   oracle_generate_states    src/search_env.cpp:7268-7472  GenerateSuccessorStates, 3-DoF branch
 *   ContPose::GetTransform    src/object_state.cpp:83-97
 *   GetFootprint / PointsInsideFootprint / isXYPointIn2DXYPolygonCustom   src/object_model.cpp:183-219, 571-651
 *   Model::mat4x4::init_from_eigen   ../cuda_renderer/include/cuda_renderer/model.h:89-107
 *
 * PARITY UNPINNED at the third-party boundary: the reference evaluates these steps with Eigen (quaternion product,
 * 4x4 products and inverse), PCL (transformPointCloud) and FLANN (kd-tree radius search), none of which is part of the
 * reference tree or pinned by it (perch.rosinstall, docker/Dockerfile install "latest Eigen").  The arithmetic below
 * restates their published generic (non-SIMD) code paths: Eigen 3.3 Geometry/Quaternion.h quat_product and
 * toRotationMatrix, LU/InverseImpl.h compute_inverse_size3/size4, PCL 1.8 common/impl/transforms.hpp, FLANN
 * L2_Simple<float>.  The reference has no test or golden vector for any of it.
 *
 * Radius searches are brute force over the whole range; build with -ffp-contract=off.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/perch_b200.h"

/* ---- Eigen restatements ------------------------------------------------------------------------------------------ */
static void quat_mul(const double a[4], const double b[4], double o[4]) { /* (w, x, y, z) */
    o[0] = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
    o[1] = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
    o[2] = a[0] * b[2] + a[2] * b[0] + a[3] * b[1] - a[1] * b[3];
    o[3] = a[0] * b[3] + a[3] * b[0] + a[1] * b[2] - a[2] * b[1];
}

/* ContPose::GetTransform (object_state.cpp:83-97): rotation of the normalised quaternion, row-major */
static void state_rotation(const perch_state* s, double R[9]) {
    double q[4];
    if (s->qw == 0 && s->qx == 0 && s->qy == 0 && s->qz == 0) {
        const double hr = 0.5 * s->roll, hp = 0.5 * s->pitch, hy = 0.5 * s->yaw;
        const double sr = sin(hr), cr = cos(hr), sp = sin(hp), cp = cos(hp), sy = sin(hy), cy = cos(hy);
        const double qy_[4] = {cy, sy * 0.0, sy * 0.0, sy};
        const double qp_[4] = {cp, sp * 0.0, sp, sp * 0.0};
        const double qr_[4] = {cr, sr, sr * 0.0, sr * 0.0};
        double t[4];
        quat_mul(qy_, qp_, t); /* yaw_angle * pitch_angle */
        quat_mul(t, qr_, q);   /* ... * roll_angle */
    } else {
        q[0] = s->qw; q[1] = s->qx; q[2] = s->qy; q[3] = s->qz;
    }
    const double n2 = q[1] * q[1] + q[2] * q[2] + q[3] * q[3] + q[0] * q[0];
    const double n = sqrt(n2);
    const double w = q[0] / n, x = q[1] / n, y = q[2] / n, z = q[3] / n;
    const double tx = 2.0 * x, ty = 2.0 * y, tz = 2.0 * z;
    const double twx = tx * w, twy = ty * w, twz = tz * w;
    const double txx = tx * x, txy = ty * x, txz = tz * x;
    const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
    R[0] = 1.0 - (tyy + tzz); R[1] = txy - twz;         R[2] = txz + twy;
    R[3] = txy + twz;         R[4] = 1.0 - (txx + tzz); R[5] = tyz - twx;
    R[6] = txz - twy;         R[7] = tyz + twx;         R[8] = 1.0 - (txx + tyy);
}

static void mat4_mul(const double* a, const double* b, double* o) {
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            o[4 * i + j] = ((a[4 * i] * b[j] + a[4 * i + 1] * b[4 + j]) + a[4 * i + 2] * b[8 + j]) + a[4 * i + 3] * b[12 + j];
}

static double det3_helper(const double* m, int i1, int i2, int i3, int j1, int j2, int j3) {
    return m[4 * i1 + j1] * (m[4 * i2 + j2] * m[4 * i3 + j3] - m[4 * i2 + j3] * m[4 * i3 + j2]);
}
static double cofactor4(const double* m, int i, int j) {
    const int i1 = (i + 1) % 4, i2 = (i + 2) % 4, i3 = (i + 3) % 4, j1 = (j + 1) % 4, j2 = (j + 2) % 4, j3 = (j + 3) % 4;
    return det3_helper(m, i1, i2, i3, j1, j2, j3) + det3_helper(m, i2, i3, i1, j1, j2, j3) + det3_helper(m, i3, i1, i2, j1, j2, j3);
}

/* search_env.cpp:1535-1541; row-major in and out */
void oracle_camera_matrix(const double* cam_to_world, double* out) {
    static const double cam_to_body[16] = {0, 0, 1, 0, -1, 0, 0, 0, 0, -1, 0, 0, 0, 0, 0, 1};
    double zf[16];
    mat4_mul(cam_to_world, cam_to_body, zf);
    for (int row = 0; row < 4; ++row)
        for (int col = 0; col < 4; ++col) {
            const double cf = cofactor4(zf, col, row);
            out[4 * row + col] = ((row + col) & 1) ? -cf : cf;
        }
    const double det = ((zf[0] * out[0] + zf[4] * out[1]) + zf[8] * out[2]) + zf[12] * out[3];
    for (int k = 0; k < 16; ++k) out[k] /= det;
}

/* ---- FLANN radius search as a count ------------------------------------------------------------------------------- */
static int radius_count(const float* xyz, int begin, int end, float qx, float qy, float qz, float r2) {
    int cnt = 0;
    for (int i = begin; i < end; ++i) {
        const float dx = qx - xyz[3 * (size_t)i], dy = qy - xyz[3 * (size_t)i + 1], dz = qz - xyz[3 * (size_t)i + 2];
        float d = dx * dx;
        d += dy * dy;
        d += dz * dz;
        cnt += d < r2;
    }
    return cnt;
}
static int cap_max_nn(int count, int max_nn, int total) {
    if (max_nn <= 0 || max_nn > total) max_nn = total;
    return count < max_nn ? count : max_nn;
}

typedef struct oracle_search {
    const float* xyz;      /* n points */
    const int* ranges;     /* 2 * n_ranges */
    int n_ranges;
    const float* constraint;
    int n_constraint;
} oracle_search;

/* ---- GetStateImagesUnifiedGPU loop ---------------------------------------------------------------------------------- */
int oracle_assemble_states(const perch_env* env, const perch_model_meta* metas, int n_models, const perch_state* states,
                           int n_states, const float* seg_counts, int n_labels, float observed_point_num, int num_pixels,
                           const float* search_xyz, const int* ranges, int n_ranges, float* poses16, int* model_map,
                           int* label, float* obs_total) {
    double cam[16];
    oracle_camera_matrix(env->cam_to_world, cam);
    for (int i = 0; i < n_states; ++i) {
        const perch_state* s = &states[i];
        if (s->model_id < 0 || s->model_id >= n_models) return 2;
        const perch_model_meta* m = &metas[s->model_id];
        double R[9], T[16], pre[16], M[16], P[16];
        state_rotation(s, R);
        T[0] = R[0]; T[1] = R[1]; T[2] = R[2]; T[3] = s->x;
        T[4] = R[3]; T[5] = R[4]; T[6] = R[5]; T[7] = s->y;
        T[8] = R[6]; T[9] = R[7]; T[10] = R[8]; T[11] = s->z;
        T[12] = 0; T[13] = 0; T[14] = 0; T[15] = 1;
        for (int k = 0; k < 16; ++k) pre[k] = (double)m->preprocessing_transform[k];
        mat4_mul(T, pre, M);
        mat4_mul(cam, M, P);
        for (int k = 0; k < 12; ++k) poses16[16 * (size_t)i + k] = (float)(P[k] * 100);
        for (int k = 12; k < 16; ++k) poses16[16 * (size_t)i + k] = (float)P[k];
        model_map[i] = s->model_id;
        if (env->use_external_pose_list == 1) {
            const int required_object_id = s->segmentation_label_id - 1;
            if (required_object_id < 0 || required_object_id >= n_labels) return 4;
            label[i] = required_object_id;
            obs_total[i] = seg_counts[required_object_id];
        } else {
            label[i] = -1;
            if (env->use_cylinder_observed) {
                const double rad = m->inflation_factor * m->circumscribed_radius;
                const float r2 = (float)(rad * rad);
                const int b = n_ranges > 0 ? ranges[0] : 0, e = n_ranges > 0 ? ranges[1] : 0;
                const int cnt = radius_count(search_xyz, b, e, (float)s->x, (float)s->y, (float)s->z, r2);
                obs_total[i] = (float)cap_max_nn(cnt, num_pixels, e - b);
            } else {
                obs_total[i] = observed_point_num;
            }
        }
    }
    return 0;
}

/* ---- IsValidPose ----------------------------------------------------------------------------------------------------- */
typedef struct pose_f {
    float r[9];
    float tx, ty, tz;
} pose_f;

static void pose_to_float(const perch_state* s, pose_f* p) {
    double R[9];
    state_rotation(s, R);
    for (int k = 0; k < 9; ++k) p->r[k] = (float)R[k];
    p->tx = (float)s->x; p->ty = (float)s->y; p->tz = (float)s->z;
}

static int footprint_in_bounds(const perch_env* env, const perch_model_meta* m, const float* footprints, const pose_f* p) {
    const double lo_x = env->x_min - env->footprint_tolerance, hi_x = env->x_max + env->footprint_tolerance;
    const double lo_y = env->y_min - env->footprint_tolerance, hi_y = env->y_max + env->footprint_tolerance;
    for (int i = 0; i < m->footprint_count; ++i) {
        const float* h = footprints + 3 * (size_t)(m->footprint_offset + i);
        /* pcl::transformPointCloud (PCL 1.8 transforms.hpp): t(0,0)*x + t(0,1)*y + t(0,2)*z + t(0,3), left to right */
        const float x = p->r[0] * h[0] + p->r[1] * h[1] + p->r[2] * h[2] + p->tx;
        const float y = p->r[3] * h[0] + p->r[4] * h[1] + p->r[5] * h[2] + p->ty;
        if ((double)x < lo_x || (double)x > hi_x || (double)y < lo_y || (double)y > hi_y) return 0;
    }
    return 1;
}

static float cof3(const float* L, int i, int j) {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return L[3 * i1 + j1] * L[3 * i2 + j2] - L[3 * i1 + j2] * L[3 * i2 + j1];
}

static int constraint_ok(const perch_env* env, const perch_model_meta* m, const float* footprints, const float* constraint,
                         int n_constraint, const pose_f* p) {
    if (n_constraint <= 0) return 1;
    const float f = (float)m->inflation_factor;
    float L[9], I[9];
    for (int k = 0; k < 9; ++k) L[k] = f * p->r[k];
    const float c00 = cof3(L, 0, 0), c10 = cof3(L, 1, 0), c20 = cof3(L, 2, 0);
    const float det = c00 * L[0] + c10 * L[3] + c20 * L[6];
    const float invdet = 1.0f / det;
    I[0] = c00 * invdet; I[1] = c10 * invdet; I[2] = c20 * invdet;
    I[3] = cof3(L, 0, 1) * invdet; I[4] = cof3(L, 1, 1) * invdet; I[5] = cof3(L, 2, 1) * invdet;
    I[6] = cof3(L, 0, 2) * invdet; I[7] = cof3(L, 1, 2) * invdet; I[8] = cof3(L, 2, 2) * invdet;
    const float itx = -(I[0] * p->tx + I[1] * p->ty + I[2] * p->tz);
    const float ity = -(I[3] * p->tx + I[4] * p->ty + I[5] * p->tz);
    int inside = 0;
    const int nv = m->footprint_count;
    const float* hull = footprints + 3 * (size_t)m->footprint_offset;
    for (int i = 0; i < n_constraint; ++i) {
        const float cx = constraint[3 * (size_t)i], cy = constraint[3 * (size_t)i + 1];
        const float px = I[0] * cx + I[1] * cy + I[2] * 0.f + itx;
        const float py = I[3] * cx + I[4] * cy + I[5] * 0.f + ity;
        const double qx = (double)px, qy = (double)py;
        int in_poly = 0;
        if (nv > 0) {
            double xold = hull[3 * (nv - 1)], yold = hull[3 * (nv - 1) + 1];
            for (int v = 0; v < nv; ++v) {
                const double xnew = hull[3 * v], ynew = hull[3 * v + 1];
                double x1, x2, y1, y2;
                if (xnew > xold) { x1 = xold; x2 = xnew; y1 = yold; y2 = ynew; }
                else { x1 = xnew; x2 = xold; y1 = ynew; y2 = yold; }
                if ((xnew < qx) == (qx <= xold) && (qy - y1) * (x2 - x1) < (y2 - y1) * (qx - x1)) in_poly = !in_poly;
                xold = xnew; yold = ynew;
            }
        }
        inside += in_poly;
    }
    const int mp = env->min_points_for_constraint_cloud < n_constraint ? env->min_points_for_constraint_cloud : n_constraint;
    return !((double)inside < (double)mp);
}

static int enough_neighbours(const perch_env* env, const perch_model_meta* m, const oracle_search* sc, float px, float py,
                             float pz, int seg_label_id, int after_refinement) {
    const double sr = env->use_model_specific_search_resolution ? m->search_resolution : env->res;
    const double cell_rad = after_refinement ? 0.0 : hypot(sr / 2.0, sr / 2.0);
    double search_rad;
    int range = 0;
    if (env->use_external_pose_list != 1) {
        search_rad = fmax(m->circumscribed_radius, cell_rad);
    } else {
        search_rad = fmax(m->inflation_factor * m->circumscribed_radius_3d, cell_rad);
        const int required_object_id = seg_label_id - 1;
        range = required_object_id >= 0 ? 1 + required_object_id : 0;
    }
    if (env->min_neighbor_points_for_valid_pose <= 0) return 1;
    if (range >= sc->n_ranges) return 0;
    const float r2 = (float)(search_rad * search_rad);
    const int cnt = radius_count(sc->xyz, sc->ranges[2 * range], sc->ranges[2 * range + 1], px, py, pz, r2);
    return cnt >= env->min_neighbor_points_for_valid_pose;
}

static int clear_of_placed(const perch_env* env, const perch_model_meta* metas, int n_models, const perch_model_meta* m,
                           const perch_state* placed, int n_placed, double x, double y) {
    if (env->use_external_pose_list == 1) return 1;
    for (int i = 0; i < n_placed; ++i) {
        if (placed[i].model_id < 0 || placed[i].model_id >= n_models) continue;
        const double rad = m->inscribed_radius + metas[placed[i].model_id].inscribed_radius;
        const double dx = x - placed[i].x, dy = y - placed[i].y;
        if (dx * dx + dy * dy < rad * rad) return 0;
    }
    return 1;
}

static int is_valid_pose(const perch_env* env, const perch_model_meta* metas, int n_models, const float* footprints,
                         const oracle_search* sc, const perch_state* placed, int n_placed, const perch_state* s,
                         int after_refinement) {
    const perch_model_meta* m = &metas[s->model_id];
    if (!enough_neighbours(env, m, sc, (float)s->x, (float)s->y, (float)s->z, s->segmentation_label_id, after_refinement)) return 0;
    if (!clear_of_placed(env, metas, n_models, m, placed, n_placed, s->x, s->y)) return 0;
    if (env->use_external_pose_list != 1 || sc->n_constraint > 0) {
        pose_f p;
        pose_to_float(s, &p);
        if (env->use_external_pose_list != 1 && !footprint_in_bounds(env, m, footprints, &p)) return 0;
        if (!constraint_ok(env, m, footprints, sc->constraint, sc->n_constraint, &p)) return 0;
    }
    return 1;
}

int oracle_validate_states(const perch_env* env, const perch_model_meta* metas, int n_models, const float* footprints,
                           const float* search_xyz, const int* ranges, int n_ranges, const float* constraint, int n_constraint,
                           const perch_state* states, int n_states, const perch_state* placed, int n_placed,
                           int after_refinement, uint8_t* valid) {
    const oracle_search sc = {search_xyz, ranges, n_ranges, constraint, n_constraint};
    for (int i = 0; i < n_states; ++i) {
        if (states[i].model_id < 0 || states[i].model_id >= n_models) return 2;
        valid[i] = (uint8_t)is_valid_pose(env, metas, n_models, footprints, &sc, placed, n_placed, &states[i], after_refinement);
    }
    return 0;
}

/* angles::normalize_angle_positive (ROS angles.h) as applied by ContPose's Euler constructor (object_state.cpp:17-22) */
double oracle_normalize_angle_positive(double a) { return fmod(fmod(a, 2.0 * M_PI) + 2.0 * M_PI, 2.0 * M_PI); }

/* ---- GenerateSuccessorStates, 3-DoF branch (search_env.cpp:7268-7472) ------------------------------------------------ */
/* Returns the number of successors; writes at most `capacity` of them. */
long long oracle_generate_states(const perch_env* env, const perch_model_meta* metas, int n_models, const float* footprints,
                                 const float* search_xyz, const int* ranges, int n_ranges, const float* constraint,
                                 int n_constraint, const perch_state* placed, int n_placed, perch_state* out,
                                 long long capacity) {
    const oracle_search sc = {search_xyz, ranges, n_ranges, constraint, n_constraint};
    long long n = 0;
    for (int ii = 0; ii < n_models; ++ii) {
        int already = 0;
        for (int i = 0; i < n_placed; ++i) already |= placed[i].model_id == ii;
        if (already) continue; /* 7067-7075 */
        const perch_model_meta* m = &metas[ii];
        const double sr = env->use_model_specific_search_resolution ? m->search_resolution : env->res;
        const double res = env->use_adaptive_resolution ? m->inscribed_radius : sr;
        if (!(res > 0.0) || !(env->theta_res > 0.0)) return -1;
        for (double x = env->x_min; x <= env->x_max; x += res) {
            for (double y = env->y_min; y <= env->y_max; y += res) {
                for (double theta = 0; theta < 2 * M_PI; theta += env->theta_res) {
                    perch_state s;
                    memset(&s, 0, sizeof(s));
                    /* ContPose p(x, y, table_height, 0.0, 0.00, theta) */
                    s.x = x; s.y = y; s.z = env->table_height;
                    s.roll = oracle_normalize_angle_positive(0.0);
                    s.pitch = oracle_normalize_angle_positive(0.0);
                    s.yaw = oracle_normalize_angle_positive(theta);
                    s.model_id = ii;
                    s.segmentation_label_id = -1;
                    if (!is_valid_pose(env, metas, n_models, footprints, &sc, placed, n_placed, &s, 0)) continue;
                    if (m->symmetry_mode == 1 && theta > (M_PI + env->theta_res)) break;
                    if (n < capacity) out[n] = s;
                    ++n;
                    if (m->symmetric || m->symmetry_mode == 2) break;
                }
            }
        }
    }
    return n;
}

/* ---- SetInput's image steps (search_env.cpp:5911-5929, 2487-2498) ---------------------------------------------------- */
/* cv::medianBlur on CV_8UC1 (search_env.cpp:5922): exact median of the k x k window, BORDER_REPLICATE.  OpenCV is not
 * part of the reference tree; restated from its documented behaviour (tests pin this against scipy.ndimage). */
static int cmp_u8(const void* a, const void* b) { return (int)*(const uint8_t*)a - (int)*(const uint8_t*)b; }
void oracle_median_blur_u8(const uint8_t* in, uint8_t* out, int w, int h, int k) {
    if (k <= 1) { memcpy(out, in, (size_t)w * h); return; }
    const int r = k / 2;
    uint8_t* win = (uint8_t*)malloc((size_t)k * k);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            int n = 0;
            for (int dy = -r; dy <= r; ++dy)
                for (int dx = -r; dx <= r; ++dx) {
                    int sx = x + dx, sy = y + dy;
                    sx = sx < 0 ? 0 : (sx >= w ? w - 1 : sx);
                    sy = sy < 0 ? 0 : (sy >= h ? h - 1 : sy);
                    win[n++] = in[(size_t)sy * w + sx];
                }
            qsort(win, n, 1, cmp_u8);
            out[(size_t)y * w + x] = win[n / 2];
        }
    free(win);
}

/* input_depth_image_vec[i] /= division_factor: int32 element, float factor (search_env.cpp:2489-2496) */
void oracle_divide_depth(const int32_t* in, float factor, int32_t* out, int n) {
    for (int i = 0; i < n; ++i) out[i] = (int32_t)((float)in[i] / factor);
}
