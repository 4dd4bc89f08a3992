/*
 * perch_oracle.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's render + cloud + KNN + cost path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this file's
 * library; the product (libperch_b200.so) never does and has no CPU fallback.
 *
 * Parity status: PINNED for the rasterizer, depth->cloud and cost kernels against the reference's own CUDA kernels
 * compiled unmodified from /root/reference (oracle/ref_harness.cu -> oracle/_ref/libperch_ref.so, compared on a B200
 * in tests/test_parity_gpu.py).  The reference repository ships no golden vectors for this path.
 * UNPINNED: the nearest-neighbour tie-break / distance rounding (lives in the un-vendored fast_gicp dependency,
 * call site renderer.cu:1853-1865) -- lowest index, fma(dz,dz,fma(dy,dy,dx*dx)) is this repo's convention.
 *
 * Each function cites the reference lines it follows (paths relative to the reference repository).
 * Build: gcc -O2 -ffp-contract=off (the float sequences below spell out where the reference's nvcc 12.9 / sm_100a
 * build fuses multiply-adds -- by NVVM, visible in PTX, and by ptxas, visible only in `cuobjdump -sass` of
 * oracle/_ref/libperch_ref.so; the C compiler must not add or remove any).
 */
#include <float.h>
#include <limits.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <stdio.h>
#include <string.h>

typedef struct { float x, y, z; } vec3;
typedef struct { vec3 v0, v1, v2; uint64_t c0, c1, c2; } tri64; /* cuda_renderer::Model::Triangle, model.h:62-73 */

/* cvt.rzi.s32.f32: what `int32_t(x)` does on the GPU (saturating, NaN -> 0); a plain C cast is UB out of range. */
static int32_t f2i_sat(float f) {
    if (f != f) return 0;
    if (f >= 2147483648.0f) return INT32_MAX;
    if (f <= -2147483648.0f) return INT32_MIN;
    return (int32_t)f;
}
/* cvt.rzi.u64.f32 for `size_t(x)` */
static uint64_t f2u64_sat(float f) {
    if (f != f || f <= 0.0f) return 0;
    if (f >= 18446744073709551616.0f) return UINT64_MAX;
    return (uint64_t)f;
}

/* mat_mul_v, image_renderer.cuh:21-27.  nvcc evaluates m0*x + m1*y + m2*z + m3 as m3 + fma(m2,z,fma(m0,x,m1*y)). */
static float row_dot(const float* m, float x, float y, float z) {
    float t = m[1] * y;
    t = fmaf(m[0], x, t);
    t = fmaf(m[2], z, t);
    return m[3] + t;
}
static vec3 mat_mul_v(const float* m16, vec3 v) {
    vec3 r = {row_dot(m16, v.x, v.y, v.z), row_dot(m16 + 4, v.x, v.y, v.z), row_dot(m16 + 8, v.x, v.y, v.z)};
    return r;
}
static float sel_max(float a, float b) { return (a > b) ? a : b; } /* std__max, image_renderer.cuh:14-15 */
static float sel_min(float a, float b) { return (a < b) ? a : b; } /* std__min, image_renderer.cuh:17-18 */

/* compute_proj, renderer.cu:1386-1410.  K row-major 3x3. */
void oracle_compute_proj(const float* K, int width, int height, float near_, float far_, float* p) {
    memset(p, 0, 16 * sizeof(float));
    p[0] = 2 * K[0] / width;
    p[1] = -2 * K[1] / width; p[1] = -p[1];
    p[2] = -2 * K[2] / width + 1; p[2] = -p[2];
    p[5] = 2 * K[4] / height; p[5] = -p[5];
    p[6] = 2 * K[5] / height - 1; p[6] = -p[6];
    p[10] = -(far_ + near_) / (far_ - near_); p[10] = -p[10];
    p[11] = -2 * far_ * near_ / (far_ - near_);
    p[14] = -1; p[14] = -p[14];
}

/*
 * image_render, image_renderer.cuh:336-496 with kernel render_triangle_multi (212-321) and
 * rasterization_with_source (59-210), executed SERIALLY in triangle order (the loop skeleton of the reference's own
 * CPU rasterizer, cuda_renderer/src/renderer.cpp:291-330).  Running the reference's per-fragment z-test +
 * source-occlusion step one fragment at a time is the race-free meaning of the kernel: the result is independent of
 * the order (DESIGN.md "canonical raster semantics"), colour ties go to the lowest triangle index (strict `<`, :149).
 *
 * out_* are P*W*H (W*H when single_result_image).  pose_label NULL => 3-DoF occlusion rule.
 */
/* Optional per-pose outputs of the USE_TREE / USE_CLUTTER branches (image_renderer.cuh:187-206), two readings:
 *  serial_*: exactly what the per-fragment code yields when fragments arrive in triangle order (ONE legal execution of
 *            the kernel; pose_occluded_other and the clutter count depend on that order);
 *  canon_*:  per covered pixel on its minimum depth -- order independent; what libperch_b200 computes.
 * The two agree wherever a pixel receives at most one fragment per pose.  Counts are raw (no division). */
typedef struct {
    int* serial_occluded; int* serial_occluded_other; float* serial_clutter; float* serial_total;
    int* canon_occluded; int* canon_occluded_other; float* canon_clutter; float* canon_total;
} tree_out;

static void render_impl(const tri64* tris, int n_tris, const float* poses16, int n_poses, const int* pose_model_map,
                        const int* tris_model_count, int n_models, int width, int height, const float* proj16,
                        const int32_t* source_depth, const uint8_t* source_label, const int* pose_label,
                        float occlusion_threshold, int single_result_image, int32_t* out_depth, uint8_t* out_r,
                        uint8_t* out_g, uint8_t* out_b, const tree_out* tree) {
    const size_t npx = (size_t)width * height;
    int32_t* min_entry = tree ? (int32_t*)malloc(sizeof(int32_t) * npx) : NULL; /* min over fragments, never cleared */
    const size_t n_img = single_result_image ? 1 : (size_t)n_poses;
    /* image_renderer.cuh:395-398 */
    for (size_t i = 0; i < n_img * npx; ++i) { out_depth[i] = INT_MAX; out_r[i] = 0; out_g[i] = 0; out_b[i] = 0; }
    /* per-model triangle ranges, image_renderer.cuh:371-380 */
    int* low = (int*)malloc(sizeof(int) * (size_t)n_models);
    int* high = (int*)malloc(sizeof(int) * (size_t)n_models);
    int run = 0;
    for (int m = 0; m < n_models; ++m) { low[m] = run; run += tris_model_count[m]; high[m] = run; }
    const float wf = (float)(uint64_t)width, hf = (float)(uint64_t)height;
    const float clamp_max[2] = {(float)(width - 1), (float)(height - 1)};
    const float clamp_min[2] = {0, 0};
    const int use_label = pose_label != NULL;

    for (int pi = 0; pi < n_poses; ++pi) {
        const int model = pose_model_map[pi];
        const float* pose = poses16 + 16 * (size_t)pi;
        int32_t* depth_entry = out_depth + (single_result_image ? 0 : (size_t)pi * npx);
        uint8_t* r_entry = out_r + (single_result_image ? 0 : (size_t)pi * npx);
        uint8_t* g_entry = out_g + (single_result_image ? 0 : (size_t)pi * npx);
        uint8_t* b_entry = out_b + (single_result_image ? 0 : (size_t)pi * npx);
        /* image_renderer.cuh:273-274,287-288: in single-image mode every pose reads label entry 0 */
        const int this_pose_label = use_label ? pose_label[single_result_image ? 0 : pi] : 0;
        const int ti_out = single_result_image ? 0 : pi; /* image_renderer.cuh:262-291: entry 0 in single-image mode */
        if (tree) {
            for (size_t i = 0; i < npx; ++i) min_entry[i] = INT_MAX;
            if (!single_result_image || pi == 0) {
                tree->serial_occluded[ti_out] = 0; tree->serial_occluded_other[ti_out] = 0;
                tree->serial_clutter[ti_out] = 0; tree->serial_total[ti_out] = 0;
                tree->canon_occluded[ti_out] = 0; tree->canon_occluded_other[ti_out] = 0;
                tree->canon_clutter[ti_out] = 0; tree->canon_total[ti_out] = 0;
            }
        }
        for (int ti = 0; ti < n_tris; ++ti) {
            if (!(ti < high[model] && ti >= low[model])) continue; /* :242-243 */
            const tri64* T = tris + ti;
            /* :296-305 model transform, last_row, projection */
            vec3 c0 = mat_mul_v(pose, T->v0), c1 = mat_mul_v(pose, T->v1), c2 = mat_mul_v(pose, T->v2);
            const float z0 = c0.z, z1 = c1.z, z2 = c2.z;
            vec3 q0 = mat_mul_v(proj16, c0), q1 = mat_mul_v(proj16, c1), q2 = mat_mul_v(proj16, c2);
            /* :82-84 viewport: x/z*W/2 + W/2 -> fma(W, 0.5, ((x/z)*W)*0.5) in the nvcc build */
            float pts[3][2];
            pts[0][0] = fmaf(wf, 0.5f, ((q0.x / z0) * wf) * 0.5f); pts[0][1] = fmaf(hf, 0.5f, ((q0.y / z0) * hf) * 0.5f);
            pts[1][0] = fmaf(wf, 0.5f, ((q1.x / z1) * wf) * 0.5f); pts[1][1] = fmaf(hf, 0.5f, ((q1.y / z1) * hf) * 0.5f);
            pts[2][0] = fmaf(wf, 0.5f, ((q2.x / z2) * wf) * 0.5f); pts[2][1] = fmaf(hf, 0.5f, ((q2.y / z2) * hf) * 0.5f);
            /* :86-107 */
            float bboxmin[2] = {FLT_MAX, FLT_MAX}, bboxmax[2] = {-FLT_MAX, -FLT_MAX};
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 2; ++j) {
                    bboxmin[j] = sel_max(clamp_min[j], sel_min(bboxmin[j], pts[i][j]));
                    bboxmax[j] = sel_min(clamp_max[j], sel_max(bboxmax[j], pts[i][j]));
                }
            /* calculateSignedArea(A,B,C) :40-42; in the sm_100a build the first product is fused into the subtraction
             * (ptxas contraction, visible in SASS only): fma(C0-A0, B1-A1, -((B0-A0)*(C1-A1))) * 0.5; base_inv = 1/area :48 */
            const float* A = pts[0]; const float* B = pts[1]; const float* C = pts[2];
            const float area = fmaf(C[0] - A[0], B[1] - A[1], -((B[0] - A[0]) * (C[1] - A[1]))) * 0.5f;
            const float base_inv = 1.0f / area;
            /* :110-111 */
            for (uint64_t P1 = f2u64_sat(bboxmin[1] + 0.5f); (float)P1 <= bboxmax[1]; ++P1) {
                for (uint64_t P0 = f2u64_sat(bboxmin[0] + 0.5f); (float)P0 <= bboxmax[0]; ++P0) {
                    const float fp0 = (float)P0, fp1 = (float)P1;
                    /* barycentric :44-57: beta = area(A,P,C)*base_inv, gamma = area(A,B,P)*base_inv.  SASS of the sm_100a
                     * build: the row products (C0-A0)*(P1-A1) and (P1-A1)*(B0-A0) are hoisted and rounded, the
                     * per-pixel products are fused: fma(C1-A1, -(P0-A0), row0) and fma(B1-A1, P0-A0, -row1). */
                    const float aAPC = fmaf(C[1] - A[1], -(fp0 - A[0]), (C[0] - A[0]) * (fp1 - A[1])) * 0.5f;
                    const float aABP = fmaf(B[1] - A[1], fp0 - A[0], -((fp1 - A[1]) * (B[0] - A[0]))) * 0.5f;
                    const float beta = base_inv * aAPC;
                    const float gamma = base_inv * aABP;
                    const float alpha = (1.0f - beta) - gamma;
                    /* :114-115 (a NaN coordinate fails every comparison and is therefore NOT rejected) */
                    if (alpha < -0.0f || beta < -0.0f || gamma < -0.0f || alpha > 1.0f || beta > 1.0f || gamma > 1.0f)
                        continue;
                    /* :117-129 */
                    const float oz0 = alpha / z0, oz1 = beta / z1, oz2 = gamma / z2;
                    const float frag = (gamma + (beta + alpha)) / (oz2 + (oz1 + oz0));
                    const int32_t curr = f2i_sat(frag + 0.5f);
                    const size_t x_to_write = (size_t)P0, y_to_write = (size_t)(height - 1) - (size_t)P1;
                    const size_t o = x_to_write + y_to_write * (size_t)width;
                    const int32_t src = source_depth[o];
                    const int src_label = (use_label && source_label) ? source_label[o] : 0;
                    /* :146-159 z-test */
                    if (curr < depth_entry[o]) {
                        depth_entry[o] = curr;
                        r_entry[o] = (uint8_t)T->c0; g_entry[o] = (uint8_t)T->c1; b_entry[o] = (uint8_t)T->c2;
                    }
                    /* :161-186 source occlusion, evaluated on the CURRENT buffer value */
                    const int32_t nd = depth_entry[o];
                    const int64_t diff = (int64_t)nd - (int64_t)src;
                    const int32_t ad = (int32_t)(diff < 0 ? -diff : diff > INT32_MAX ? INT32_MAX : diff);
                    int test;
                    if (!use_label) test = (float)ad > occlusion_threshold;
                    else test = (this_pose_label != src_label - 1) && ((double)ad > 0.5);
                    if (test && nd > src && src > 0) {
                        r_entry[o] = 0; g_entry[o] = 0; b_entry[o] = 0;
                        depth_entry[o] = INT_MAX; /* atomicMax(&new_depth, INT_MAX) */
                        if (tree) {
                            tree->serial_occluded_other[ti_out] |= 1;                       /* :187-188 */
                            /* :189-195 `source_depth <= new_depth - 5`: new_depth is a REFERENCE to the pixel, which
                             * the atomicMax above has just set to INT_MAX, so the margin is vacuous and every cleared
                             * fragment counts (pinned by the golden vectors of the USE_CLUTTER build) */
                            if ((int64_t)src <= (int64_t)INT_MAX - 5) tree->serial_clutter[ti_out] += 1.0f;
                        }
                    } else if (test && nd <= src && src > 0) {
                        if (tree) tree->serial_occluded[ti_out] |= 1;                       /* :198-203 */
                    }
                    if (tree) {
                        tree->serial_total[ti_out] += 1.0f;                                 /* :205-206 */
                        if (curr < min_entry[o]) min_entry[o] = curr;
                    }
                }
            }
        }
        if (tree) {
            /* canonical reading: the same branches on each covered pixel's minimum depth */
            for (size_t o = 0; o < npx; ++o) {
                const int32_t m = min_entry[o];
                if (m == INT_MAX) continue;
                const int32_t src = source_depth[o];
                const int src_label = (use_label && source_label) ? source_label[o] : 0;
                const int64_t diff = (int64_t)m - (int64_t)src;
                const int32_t ad = (int32_t)(diff < 0 ? -diff : diff > INT32_MAX ? INT32_MAX : diff);
                const int test = use_label ? ((this_pose_label != src_label - 1) && ((double)ad > 0.5))
                                           : ((float)ad > occlusion_threshold);
                tree->canon_total[ti_out] += 1.0f;
                if (test && src > 0) {
                    if (m > src) {
                        tree->canon_occluded_other[ti_out] |= 1;
                        tree->canon_clutter[ti_out] += 1.0f; /* every cleared pixel, see the serial reading */
                    } else {
                        tree->canon_occluded[ti_out] |= 1;
                    }
                }
            }
        }
    }
    /* max2zero, :324-333,465-466 */
    for (size_t i = 0; i < n_img * npx; ++i)
        if (out_depth[i] == INT_MAX) out_depth[i] = 0;
    free(low);
    free(high);
    free(min_entry);
}

void oracle_render(const tri64* tris, int n_tris, const float* poses16, int n_poses, const int* pose_model_map,
                   const int* tris_model_count, int n_models, int width, int height, const float* proj16,
                   const int32_t* source_depth, const uint8_t* source_label, const int* pose_label,
                   float occlusion_threshold, int single_result_image, int32_t* out_depth, uint8_t* out_r,
                   uint8_t* out_g, uint8_t* out_b) {
    render_impl(tris, n_tris, poses16, n_poses, pose_model_map, tris_model_count, n_models, width, height, proj16,
                source_depth, source_label, pose_label, occlusion_threshold, single_result_image, out_depth, out_r, out_g,
                out_b, NULL);
}

/* oracle_render + the tree / clutter vectors (n_poses entries each; raw counts in the float arrays). */
void oracle_render_tree(const tri64* tris, int n_tris, const float* poses16, int n_poses, const int* pose_model_map,
                        const int* tris_model_count, int n_models, int width, int height, const float* proj16,
                        const int32_t* source_depth, const uint8_t* source_label, const int* pose_label,
                        float occlusion_threshold, int32_t* out_depth, uint8_t* out_r, uint8_t* out_g, uint8_t* out_b,
                        int* serial_occluded, int* serial_occluded_other, float* serial_clutter, float* serial_total,
                        int* canon_occluded, int* canon_occluded_other, float* canon_clutter, float* canon_total) {
    tree_out t = {serial_occluded, serial_occluded_other, serial_clutter, serial_total,
                  canon_occluded, canon_occluded_other, canon_clutter, canon_total};
    render_impl(tris, n_tris, poses16, n_poses, pose_model_map, tris_model_count, n_models, width, height, proj16,
                source_depth, source_label, pose_label, occlusion_threshold, 0, out_depth, out_r, out_g, out_b, &t);
}

/* transform_point, compute_point_clouds.cuh:14-35 (camera-frame part :20-22; no fused ops in the nvcc build) */
static void back_project(int x, int y, int32_t d, float cx, float cy, float fx, float fy, float depth_factor, float* px,
                         float* py, float* pz) {
    const float z = (float)d / depth_factor;
    *pz = z;
    *px = (((float)x - cx) / fx) * z;
    *py = (((float)y - cy) / fy) * z;
}

/* world-frame part :25-32 as the nvcc build of the oracle's Eigen stand-in evaluates it:
 * row = fma(m2,z,fma(m1,y,m0*x)) + t.  (Real Eigen's association is not pinned by any reference test.) */
static void to_world(const float* m /* column-major 4x4 */, float x, float y, float z, float* w) {
    for (int r = 0; r < 3; ++r) {
        float t = m[r] * x;
        t = fmaf(m[4 + r], y, t);
        t = fmaf(m[8 + r], z, t);
        w[r] = t + m[12 + r];
    }
}

/*
 * compute_point_clouds, compute_point_clouds.cuh:188-367 (kernels depth_to_mask :37-93, depth_to_2d_cloud :95-184,
 * index = exclusive scan of the mask in (image, y, x) order :290-292).  Returns the point count; output pointers may
 * be NULL.  out_xyz_soa / out_rgb are 3 x count with row stride `count`; dc_index is num_images*W*H.
 * Buffers the reference zero-fills (:302-305,317) must be zero-filled by the caller.
 */
int oracle_depth2cloud(const int32_t* depth, const uint8_t* r, const uint8_t* g, const uint8_t* b, int width, int height,
                       int num_images, float cx, float cy, float fx, float fy, float depth_factor, int stride,
                       const uint8_t* label_mask, const int* pose_label, const double* bounds6, const float* cam_tf16,
                       float* out_xyz_soa, float* out_xyz_aos, uint8_t* out_rgb, int* out_pose_map, int* out_label,
                       int* out_dc_index, int count_hint) {
    const int use_bounds = bounds6 != NULL && cam_tf16 != NULL;
    float fb[6] = {0, 0, 0, 0, 0, 0};
    if (use_bounds) for (int k = 0; k < 6; ++k) fb[k] = (float)bounds6[k];
    int count = 0;
    const size_t npx = (size_t)width * height;
    for (int n = 0; n < num_images; ++n)
        for (int y = 0; y < height; ++y)
            for (int x = 0; x < width; ++x) {
                const size_t idx = (size_t)n * npx + (size_t)y * width + x;
                if (out_dc_index) out_dc_index[idx] = count;
                if (x % stride != 0 || y % stride != 0) continue; /* :47-58 samples x*stride, y*stride */
                const int32_t d = depth[idx];
                if (!(d > 0)) continue; /* :64 */
                int in_bounds = 1;
                if (use_bounds) {
                    float px, py, pz, w[3];
                    back_project(x, y, d, cx, cy, fx, fy, depth_factor, &px, &py, &pz);
                    to_world(cam_tf16, px, py, pz, w);
                    if (w[0] > fb[0] || w[0] < fb[1]) in_bounds = 0;
                    if (w[1] > fb[2] || w[1] < fb[3]) in_bounds = 0;
                    if (w[2] > fb[4] || w[2] < fb[5]) in_bounds = 0;
                }
                int in_mask; /* depth_to_mask :66-91 */
                if (label_mask == NULL && !use_bounds) in_mask = 1;
                else if (label_mask != NULL) in_mask = label_mask[idx] > 0;
                else in_mask = in_bounds;
                if (!in_mask) continue;
                const int slot = count++;
                /* depth_to_2d_cloud :123-139 re-tests label AND bounds before writing */
                if (label_mask != NULL && !(label_mask[idx] > 0)) continue;
                if (use_bounds && !in_bounds) continue;
                if (count_hint <= 0 || slot >= count_hint) continue;
                float px, py, pz;
                back_project(x, y, d, cx, cy, fx, fy, depth_factor, &px, &py, &pz);
                if (out_xyz_soa) {
                    out_xyz_soa[slot] = px; out_xyz_soa[(size_t)count_hint + slot] = py;
                    out_xyz_soa[2 * (size_t)count_hint + slot] = pz;
                }
                if (out_xyz_aos) { out_xyz_aos[3 * (size_t)slot] = px; out_xyz_aos[3 * (size_t)slot + 1] = py; out_xyz_aos[3 * (size_t)slot + 2] = pz; }
                if (out_rgb) {
                    out_rgb[slot] = r ? r[idx] : 0; out_rgb[(size_t)count_hint + slot] = g ? g[idx] : 0;
                    out_rgb[2 * (size_t)count_hint + slot] = b ? b[idx] : 0;
                }
                if (out_pose_map) out_pose_map[slot] = n;
                if (out_label) { /* :168-178 */
                    if (label_mask != NULL) out_label[slot] = (int)label_mask[idx] - 1;
                    else if (pose_label != NULL) out_label[slot] = pose_label[n];
                }
            }
    return count;
}

/*
 * fast_gicp::brute_force_knn_search(k=1), call site renderer.cu:1853-1865 (source not in the reference tree; see the
 * header).  Linear scan, strict `<` => lowest index on ties; restricted to [begin,end) of the source point's label.
 */
void oracle_knn_k1(const float* src_aos, int n_src, const float* tgt_aos, int n_tgt, const int* src_label,
                   const int* label_begin, const int* label_end, int n_labels, float* out_sqdist, int* out_index) {
    for (int i = 0; i < n_src; ++i) {
        int b = 0, e = n_tgt;
        if (src_label) {
            const int l = src_label[i];
            if (l < 0 || l >= n_labels) { b = 0; e = 0; } else { b = label_begin[l]; e = label_end[l]; }
        }
        const float px = src_aos[3 * (size_t)i], py = src_aos[3 * (size_t)i + 1], pz = src_aos[3 * (size_t)i + 2];
        float best = FLT_MAX; int best_j = -1;
        for (int j = b; j < e; ++j) {
            const float dx = px - tgt_aos[3 * (size_t)j], dy = py - tgt_aos[3 * (size_t)j + 1], dz = pz - tgt_aos[3 * (size_t)j + 2];
            const float d = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
            if (d < best) { best = d; best_j = j; }
        }
        out_sqdist[i] = best;
        out_index[i] = best_j;
    }
}

/* rgb2lab, compute_costs.cuh:57-88 (double precision; host libm pow/cbrt may differ from CUDA's in the last ulp) */
static void rgb2lab(uint8_t rr, uint8_t gg, uint8_t bbb, float* lab) {
    double r = rr / 255.0, g = gg / 255.0, b = bbb / 255.0, x, y, z;
    r = ((r > 0.04045) ? pow((r + 0.055) / 1.055, 2.4) : (r / 12.92)) * 100.0;
    g = ((g > 0.04045) ? pow((g + 0.055) / 1.055, 2.4) : (g / 12.92)) * 100.0;
    b = ((b > 0.04045) ? pow((b + 0.055) / 1.055, 2.4) : (b / 12.92)) * 100.0;
    x = r * 0.4124564 + g * 0.3575761 + b * 0.1804375;
    y = r * 0.2126729 + g * 0.7151522 + b * 0.0721750;
    z = r * 0.0193339 + g * 0.1191920 + b * 0.9503041;
    x = x / 95.047; y = y / 100.00; z = z / 108.883;
    x = (x > 0.008856) ? cbrt(x) : (7.787 * x + 16.0 / 116.0);
    y = (y > 0.008856) ? cbrt(y) : (7.787 * y + 16.0 / 116.0);
    z = (z > 0.008856) ? cbrt(z) : (7.787 * z + 16.0 / 116.0);
    lab[0] = (float)((116.0 * y) - 16);
    lab[1] = (float)(500 * (x - y));
    lab[2] = (float)(200 * (y - z));
}

#define SQR(x) ((x) * (x))
#define POW7(x) (((x) * (x) * (x)) * ((x) * (x) * (x)) * (x))
#define DEG2RAD(x) ((x) * M_PI / 180)
#define RAD2DEG(x) ((x) / M_PI * 180)
/* color_distance (CIEDE2000), compute_costs.cuh:90-159, including its two non-exclusive `if`s (:112-123, :133-144):
 * the low-chroma assignments are always overwritten by the following if/else chain. */
static double color_distance(float l1, float a1, float b1, float l2, float a2, float b2) {
    const double eps = 1e-5;
    double c1 = sqrtf(SQR(a1) + SQR(b1));
    double c2 = sqrtf(SQR(a2) + SQR(b2));
    double meanC = (c1 + c2) / 2.0;
    double meanC7 = POW7(meanC);
    double g = 0.5 * (1 - sqrtf(meanC7 / (meanC7 + 6103515625.)));
    double a1p = a1 * (1 + g);
    double a2p = a2 * (1 + g);
    c1 = sqrtf(SQR(a1p) + SQR(b1));
    c2 = sqrtf(SQR(a2p) + SQR(b2));
    double h1 = fmodf(atan2f(b1, a1p) + 2 * M_PI, 2 * M_PI);
    double h2 = fmodf(atan2f(b2, a2p) + 2 * M_PI, 2 * M_PI);
    double deltaL = l2 - l1;
    double deltaC = c2 - c1;
    double deltah = 0;
    if (fabs(h2 - h1) <= M_PI) deltah = h2 - h1;
    else if (h2 > h1) deltah = h2 - h1 - 2 * M_PI;
    else deltah = h2 - h1 + 2 * M_PI;
    double deltaH = 2 * sqrtf(c1 * c2) * sinf(deltah / 2);
    double meanL = (l1 + l2) / 2;
    meanC = (c1 + c2) / 2.0;
    meanC7 = POW7(meanC);
    double meanH;
    if (fabs(h1 - h2) <= M_PI + eps) meanH = (h1 + h2) / 2;
    else if (h1 + h2 < 2 * M_PI) meanH = (h1 + h2 + 2 * M_PI) / 2;
    else meanH = (h1 + h2 - 2 * M_PI) / 2;
    double T = 1 - 0.17 * cosf(meanH - DEG2RAD(30)) + 0.24 * cosf(2 * meanH) + 0.32 * cosf(3 * meanH + DEG2RAD(6)) -
               0.2 * cosf(4 * meanH - DEG2RAD(63));
    double sl = 1 + (0.015 * SQR(meanL - 50)) / sqrtf(20 + SQR(meanL - 50));
    double sc = 1 + 0.045 * meanC;
    double sh = 1 + 0.015 * meanC * T;
    double rc = 2 * sqrtf(meanC7 / (meanC7 + 6103515625.));
    double rt = -sinf(DEG2RAD(60 * expf(-SQR((RAD2DEG(meanH) - 275) / 25)))) * rc;
    return sqrtf(SQR(deltaL / sl) + SQR(deltaC / sc) + SQR(deltaH / sh) + rt * deltaC / sc * deltaH / sh);
}

/* exported so tests can locate colour pairs whose distance sits next to the threshold (flip-tolerant parity class) */
double oracle_color_distance_rgb(uint8_t r1, uint8_t g1, uint8_t b1, uint8_t r2, uint8_t g2, uint8_t b2) {
    float A[3], B[3];
    rgb2lab(r1, g1, b1, A);
    rgb2lab(r2, g2, b2, B);
    return color_distance(A[0], A[1], A[2], B[0], B[1], B[2]);
}

/*
 * The COST stage of render_cuda_multi_unified from already-rendered images: renderer.cu:1617-1659 (clouds),
 * 1668-1686 (6-DoF: stable sort of the observed cloud by label + per-label ranges), 1852-1877 (k=1 search, squared
 * resolution), 1881-1907 + compute_costs.cuh:293-457 (kernels compute_render_cost :161-273, compute_observed_cost
 * :274-290, percentage math :364-378,407-446).
 *
 * depth/r/g/b: P rendered images.  obs_xyz_aos / obs_rgb_soa / obs_label: observed cloud as handed to the reference
 * (colours are NOT permuted by the label sort, renderer.cu:1881).  out_counts (3*P ints, optional): rendered points,
 * unexplained rendered points, explained observed points.  near_thr_flags (P ints, optional): number of colour
 * comparisons of that pose that fall within color_eps of the threshold (cost_type 1, flip-tolerant class).
 */
void oracle_costs(const int32_t* depth, const uint8_t* r, const uint8_t* g, const uint8_t* b, int n_poses, int width,
                  int height, float cx, float cy, float fx, float fy, float depth_factor, int stride,
                  const int* pose_label, const float* obs_xyz_aos, const uint8_t* obs_rgb_soa, const int* obs_label,
                  int n_obs, const float* pose_obs_total, int cost_type, int calc_observed, float sensor_resolution,
                  float color_distance_threshold, float* rendered_cost, float* observed_cost, float* points_diff_cost,
                  int32_t* out_counts, int* near_thr_flags, double color_eps) {
    /* clouds */
    const int n_rend = oracle_depth2cloud(depth, r, g, b, width, height, n_poses, cx, cy, fx, fy, depth_factor, stride,
                                          NULL, pose_label, NULL, NULL, NULL, NULL, NULL, NULL, NULL, NULL, 0);
    const size_t nr = (size_t)(n_rend > 0 ? n_rend : 1);
    float* c_aos = (float*)calloc(nr * 3, sizeof(float));
    uint8_t* c_rgb = (uint8_t*)calloc(nr * 3, 1);
    int* c_map = (int*)calloc(nr, sizeof(int));
    int* c_label = (int*)calloc(nr, sizeof(int));
    oracle_depth2cloud(depth, r, g, b, width, height, n_poses, cx, cy, fx, fy, depth_factor, stride, NULL, pose_label,
                       NULL, NULL, NULL, c_aos, c_rgb, c_map, c_label, NULL, n_rend);
    /* observed cloud: label sort + ranges (6-DoF only) */
    const size_t no = (size_t)(n_obs > 0 ? n_obs : 1);
    float* o_aos = (float*)malloc(no * 3 * sizeof(float));
    memcpy(o_aos, obs_xyz_aos, (size_t)n_obs * 3 * sizeof(float));
    int *lb = NULL, *le = NULL, n_labels = 0;
    if (pose_label) {
        /* stable counting sort by label */
        int maxl = -1;
        for (int i = 0; i < n_obs; ++i) if (obs_label[i] > maxl) maxl = obs_label[i];
        n_labels = maxl + 1;
        lb = (int*)calloc((size_t)(n_labels > 0 ? n_labels : 1), sizeof(int));
        le = (int*)calloc((size_t)(n_labels > 0 ? n_labels : 1), sizeof(int));
        int* cnt = (int*)calloc((size_t)n_labels + 1, sizeof(int));
        for (int i = 0; i < n_obs; ++i) cnt[obs_label[i] + 1]++; /* labels are >= 0 here (mask-1 of mask>0 pixels) */
        for (int l = 0; l < n_labels; ++l) { lb[l] = cnt[l]; cnt[l + 1] += cnt[l]; le[l] = cnt[l + 1]; }
        int* cur = (int*)malloc(sizeof(int) * (size_t)(n_labels > 0 ? n_labels : 1));
        for (int l = 0; l < n_labels; ++l) cur[l] = lb[l];
        for (int i = 0; i < n_obs; ++i) {
            const int s = cur[obs_label[i]]++;
            o_aos[3 * (size_t)s] = obs_xyz_aos[3 * (size_t)i];
            o_aos[3 * (size_t)s + 1] = obs_xyz_aos[3 * (size_t)i + 1];
            o_aos[3 * (size_t)s + 2] = obs_xyz_aos[3 * (size_t)i + 2];
        }
        free(cnt);
        free(cur);
    }
    float* k_dist = (float*)malloc(nr * sizeof(float));
    int* k_idx = (int*)malloc(nr * sizeof(int));
    oracle_knn_k1(c_aos, n_rend, o_aos, n_obs, pose_label ? c_label : NULL, lb, le, n_labels, k_dist, k_idx);
    const float res2 = sensor_resolution * sensor_resolution; /* renderer.cu:1877 */

    const size_t P = (size_t)(n_poses > 0 ? n_poses : 1);
    float* cost = (float*)calloc(P, sizeof(float));
    float* num = (float*)calloc(P, sizeof(float));
    uint8_t* explained = (uint8_t*)calloc(P * no, 1);
    if (near_thr_flags) memset(near_thr_flags, 0, sizeof(int) * (size_t)n_poses);
    /* compute_render_cost, compute_costs.cuh:188-272 */
    for (int i = 0; i < n_rend; ++i) {
        const int p = c_map[i], o = k_idx[i];
        num[p] += 1;
        if (k_dist[i] > res2) { cost[p] += 1.0f; continue; }
        if (cost_type == 1) {
            /* :214-229: plane 2 is passed as "red" for both clouds */
            float lab2[3], lab1[3];
            rgb2lab(c_rgb[(size_t)i + 2 * n_rend], c_rgb[(size_t)i + n_rend], c_rgb[i], lab2);
            rgb2lab(obs_rgb_soa[(size_t)o + 2 * n_obs], obs_rgb_soa[(size_t)o + n_obs], obs_rgb_soa[o], lab1);
            const double cd = color_distance(lab1[0], lab1[1], lab1[2], lab2[0], lab2[1], lab2[2]);
            if (near_thr_flags && fabs(cd - (double)color_distance_threshold) <= color_eps) near_thr_flags[p]++;
            if (cd > color_distance_threshold) cost[p] += 1.0f;
            else explained[(size_t)p * no + o] = 1;
        } else {
            explained[(size_t)p * no + o] = 1; /* type 0 and type 2 */
        }
    }
    for (int p = 0; p < n_poses; ++p) {
        /* :364-378 */
        const float rendered_explained = num[p] - cost[p];
        float rc = (num[p] == 0) ? -1 : cost[p] / num[p];
        rc = (rc == -1) ? -1 : rc * 100;
        rendered_cost[p] = rc;
        float expl = 0;
        for (int o = 0; o < n_obs; ++o) expl += (float)explained[(size_t)p * no + o]; /* compute_observed_cost :284-288 */
        if (calc_observed) {
            /* :407-446 */
            const float tot = pose_obs_total ? pose_obs_total[p] : 0.f;
            if (points_diff_cost) points_diff_cost[p] = rendered_explained - expl;
            if (observed_cost) observed_cost[p] = ((tot - expl) / tot) * 100;
        }
        if (out_counts) {
            out_counts[p] = (int32_t)num[p];
            out_counts[n_poses + p] = (int32_t)cost[p];
            out_counts[2 * n_poses + p] = (int32_t)expl;
        }
    }
    free(c_aos); free(c_rgb); free(c_map); free(c_label); free(o_aos); free(lb); free(le);
    free(k_dist); free(k_idx); free(cost); free(num); free(explained);
}

/*
 * render_cuda_multi_unified("COST") for a LIST of poses without keeping their images: per pose, oracle_render into a
 * thread-local 640x480 image (image_renderer.cuh:59-321 as restated above) followed by oracle_costs on that one image
 * (compute_point_clouds.cuh:37-184, renderer.cu:1852-1877, compute_costs.cuh:161-290).  Poses are independent in the
 * reference too (one image, one cloud segment and one `explained` row each), so running them one at a time -- here
 * with OpenMP over poses -- yields exactly what the batched call yields.  Lets the parity tests and bench.py's
 * parity_check compare the fused CUDA kernel with the oracle at the full C2 / C3 / C5 sizes (a 20 000-triangle pose
 * costs ~5 ms of one core).  out_counts is 3 x P like oracle_costs.
 */
void oracle_score_poses(const tri64* tris, int n_tris, const float* poses16, int n_poses, const int* pose_model_map,
                        const int* tris_model_count, int n_models, int width, int height, const float* proj16,
                        const int32_t* source_depth, const uint8_t* source_label, const int* pose_label,
                        float occlusion_threshold, float cx, float cy, float fx, float fy, float depth_factor, int stride,
                        const float* obs_xyz_aos, const uint8_t* obs_rgb_soa, const int* obs_label, int n_obs,
                        const float* pose_obs_total, int cost_type, int calc_observed, float sensor_resolution,
                        float color_distance_threshold, float* rendered_cost, float* observed_cost,
                        float* points_diff_cost, int32_t* out_counts) {
    const size_t npx = (size_t)width * height;
#pragma omp parallel
    {
        int32_t* d = (int32_t*)malloc(npx * sizeof(int32_t));
        uint8_t* rgb = (uint8_t*)malloc(3 * npx);
#pragma omp for schedule(dynamic, 1)
        for (int p = 0; p < n_poses; ++p) {
            render_impl(tris, n_tris, poses16 + 16 * (size_t)p, 1, pose_model_map + p, tris_model_count, n_models, width, height,
                        proj16, source_depth, source_label, pose_label ? pose_label + p : NULL, occlusion_threshold, 0, d,
                        rgb, rgb + npx, rgb + 2 * npx, NULL);
            float rc = 0, oc = 0, dc = 0;
            int32_t cnt[3] = {0, 0, 0};
            oracle_costs(d, rgb, rgb + npx, rgb + 2 * npx, 1, width, height, cx, cy, fx, fy, depth_factor, stride,
                         pose_label ? pose_label + p : NULL, obs_xyz_aos, obs_rgb_soa, obs_label, n_obs,
                         pose_obs_total ? pose_obs_total + p : NULL, cost_type, calc_observed, sensor_resolution,
                         color_distance_threshold, &rc, &oc, &dc, cnt, NULL, 0.0);
            rendered_cost[p] = rc;
            if (calc_observed) {
                if (observed_cost) observed_cost[p] = oc;
                if (points_diff_cost) points_diff_cost[p] = dc;
            }
            if (out_counts) {
                out_counts[p] = cnt[0];
                out_counts[n_poses + p] = cnt[1];
                out_counts[2 * (size_t)n_poses + p] = cnt[2];
            }
        }
        free(d);
        free(rgb);
    }
}

/*
 * kNN-CUDA package API (cuda_renderer/knncuda.h:33-100).  Layout: ref is dim x ref_nb, query dim x query_nb (one row
 * per dimension); outputs k x query_nb, ascending distance, square root applied, 0-based indices.
 * Selection = knn_c's modified insertion sort (knncuda.cu:882-912, same as the device kernel :149-195): element i is
 * skipped when i >= k and dist >= dist[k-1]; otherwise it moves left past STRICTLY larger entries only, so equal
 * distances stay in ascending index order.
 * fused != 0: squared distance as the device kernel computes it under nvcc's default -fmad, ssd = fma(diff, diff, ssd)
 *             over the dimensions in order (compute_distances, knncuda.cu:26-96) -> bit-comparable with
 *             knn_cuda_global / knn_cuda_texture.
 * fused == 0: knn_c's host arithmetic (compute_distance, knncuda.cu:852-865): separate multiply and add.
 * Needs k <= ref_nb.
 */
int oracle_knn(const float* ref, int ref_nb, const float* query, int query_nb, int dim, int k, int fused,
               float* knn_dist, int* knn_index) {
    if (k > ref_nb || k <= 0 || dim <= 0) return 1;
    float* dist = (float*)malloc((size_t)ref_nb * sizeof(float));
    int* index = (int*)malloc((size_t)ref_nb * sizeof(int));
    if (!dist || !index) { free(dist); free(index); return 2; }
    for (int i = 0; i < query_nb; ++i) {
        for (int j = 0; j < ref_nb; ++j) {
            float sum = 0.f;
            for (int d = 0; d < dim; ++d) {
                const float diff = ref[(size_t)d * ref_nb + j] - query[(size_t)d * query_nb + i];
                if (fused) sum = fmaf(diff, diff, sum);
                else { const float sq = diff * diff; sum = sum + sq; }
            }
            /* knn_c takes the root BEFORE the selection (compute_distance :864), the device path after it
               (compute_sqrt :206-211): two different squared distances that round to the same root are a tie (index
               order) for the former and ordered for the latter. */
            dist[j] = fused ? sum : sqrtf(sum);
            index[j] = j;
        }
        index[0] = 0;
        for (int t = 1; t < ref_nb; ++t) {
            const float cur = dist[t];
            if (t >= k && cur >= dist[k - 1]) continue;
            int j = t < k - 1 ? t : k - 1;
            while (j > 0 && dist[j - 1] > cur) { dist[j] = dist[j - 1]; index[j] = index[j - 1]; --j; }
            dist[j] = cur;
            index[j] = t;
        }
        for (int j = 0; j < k; ++j) {
            knn_dist[(size_t)j * query_nb + i] = fused ? sqrtf(dist[j]) : dist[j];
            knn_index[(size_t)j * query_nb + i] = index[j];
        }
    }
    free(dist);
    free(index);
    return 0;
}

/* =====================================================================================================================
 * Generalized-ICP of one source cloud onto a target cloud -- the per-pose work of the reference's do_icp block
 * (renderer.cu:1688-1817: fast_gicp::FastGICPCudaCore with k_correspondences 10, <= 150 iterations, rotation epsilon
 * 2e-3, translation epsilon 5e-4).  PARITY UNPINNED: fast_gicp is not in the reference tree and has no pinned version;
 * this is a double-precision restatement of the published algorithm (Segal et al. 2009; Koide et al. 2021), the
 * yardstick for perception_b200/csrc/gicp.cu (tolerance 0.1 mm / 0.01 deg, SURVEY.md 8c class 4):
 *   covariances: k nearest neighbours within the own cloud (self included), regularised to eigenvalues (1,1,1e-3)
 *                => C = I - (1 - 1e-3) n n^T, n = direction of least variance (n = 0 when fewer than 3 neighbours);
 *   iteration:   nearest target point (same label only, if labels are given) of every transformed source point, float
 *                distance fma(dz,dz,fma(dy,dy,dx*dx)), lowest index on ties; M = (C_B + R C_A R^T)^-1;
 *                r = b - (R a + t); J = [skew(R a + t), -I]; H = sum J^T M J, g = sum J^T M r, y0 = sum r^T M r;
 *   step:        Levenberg-Marquardt as in fast_gicp's LsqRegistration::step_lm (the optimiser FastGICP runs by
 *                default; lambda starts at 1e-9 max diag H): up to 10 trials, delta = -(H + lambda I)^-1 g, trial error
 *                y' of exp(delta) T with the correspondences and M of the linearisation frozen, gain ratio
 *                rho = (y0 - y') / (delta . (lambda delta - g)); rho >= 0: accept, lambda *= max(1/3, 1 - (2 rho - 1)^3);
 *                rho < 0: lambda *= nu, nu *= 2 (or stop as converged when exp(delta) is below the epsilons);
 *   stop:        all |exp(delta).R - I| < rotation_epsilon and all |exp(delta).t| < translation_epsilon; or (extension,
 *                not in fast_gicp: hard assignment makes some poses cycle through a few correspondence sets for ever) the
 *                linearisation error has not improved on its best value by 1e-3 for 8 linearisations: the transform of
 *                that best linearisation is returned (PERCH_GICP_STALL_WINDOW, 0 = off).
 * out_T12 = R row-major (9) then t (3), metres.  Returns the number of linearisations performed.
 */
static void jacobi_smallest(double a00, double a01, double a02, double a11, double a12, double a22, double n[3]) {
    double A[3][3] = {{a00, a01, a02}, {a01, a11, a12}, {a02, a12, a22}};
    double V[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int sweep = 0; sweep < 12; ++sweep)
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                const double apq = A[p][q];
                if (apq == 0) continue;
                const double theta = (A[q][q] - A[p][p]) / (2 * apq);
                const double sgn = theta >= 0 ? 1 : -1;
                const double t = sgn / (sgn * theta + sqrt(theta * theta + 1));
                const double c = 1 / sqrt(t * t + 1), s = t * c;
                for (int k = 0; k < 3; ++k) { const double x = A[k][p], y = A[k][q]; A[k][p] = c * x - s * y; A[k][q] = s * x + c * y; }
                for (int k = 0; k < 3; ++k) { const double x = A[p][k], y = A[q][k]; A[p][k] = c * x - s * y; A[q][k] = s * x + c * y; }
                for (int k = 0; k < 3; ++k) { const double x = V[k][p], y = V[k][q]; V[k][p] = c * x - s * y; V[k][q] = s * x + c * y; }
            }
    int m = 0;
    if (A[1][1] < A[m][m]) m = 1;
    if (A[2][2] < A[m][m]) m = 2;
    n[0] = V[0][m]; n[1] = V[1][m]; n[2] = V[2][m];
}

/* normals[3*i..] of every point of a cloud from its k nearest neighbours (self included) */
void oracle_cloud_normals(const float* aos, int n, int k, double* normals) {
    float* bd = (float*)malloc(sizeof(float) * (size_t)(k > 0 ? k : 1));
    int* bi = (int*)malloc(sizeof(int) * (size_t)(k > 0 ? k : 1));
    for (int i = 0; i < n; ++i) {
        int cnt = 0;
        for (int j = 0; j < n; ++j) {
            const float dx = aos[3 * (size_t)i] - aos[3 * (size_t)j], dy = aos[3 * (size_t)i + 1] - aos[3 * (size_t)j + 1],
                        dz = aos[3 * (size_t)i + 2] - aos[3 * (size_t)j + 2];
            const float d = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
            if (cnt == k && d >= bd[k - 1]) continue;
            int pos = cnt < k ? cnt++ : k - 1;
            while (pos > 0 && bd[pos - 1] > d) { bd[pos] = bd[pos - 1]; bi[pos] = bi[pos - 1]; --pos; }
            bd[pos] = d; bi[pos] = j;
        }
        double* nn = normals + 3 * (size_t)i;
        nn[0] = nn[1] = nn[2] = 0;
        if (cnt < 3) continue;
        double m[3] = {0, 0, 0};
        for (int j = 0; j < cnt; ++j) for (int a = 0; a < 3; ++a) m[a] += aos[3 * (size_t)bi[j] + a];
        for (int a = 0; a < 3; ++a) m[a] /= cnt;
        double c[6] = {0, 0, 0, 0, 0, 0};
        for (int j = 0; j < cnt; ++j) {
            const double x = aos[3 * (size_t)bi[j]] - m[0], y = aos[3 * (size_t)bi[j] + 1] - m[1], z = aos[3 * (size_t)bi[j] + 2] - m[2];
            c[0] += x * x; c[1] += x * y; c[2] += x * z; c[3] += y * y; c[4] += y * z; c[5] += z * z;
        }
        jacobi_smallest(c[0] / cnt, c[1] / cnt, c[2] / cnt, c[3] / cnt, c[4] / cnt, c[5] / cnt, nn);
    }
    free(bd);
    free(bi);
}

static int chol_solve6(const double H[36], const double g[6], double x[6]) {
    double L[6][6];
    for (int i = 0; i < 6; ++i)
        for (int j = 0; j <= i; ++j) {
            double s = H[i * 6 + j];
            for (int k = 0; k < j; ++k) s -= L[i][k] * L[j][k];
            if (i == j) { if (!(s > 1e-300)) return 0; L[i][i] = sqrt(s); } else L[i][j] = s / L[j][j];
        }
    double y[6];
    for (int i = 0; i < 6; ++i) { double s = -g[i]; for (int k = 0; k < i; ++k) s -= L[i][k] * y[k]; y[i] = s / L[i][i]; }
    for (int i = 5; i >= 0; --i) { double s = y[i]; for (int k = i + 1; k < 6; ++k) s -= L[k][i] * x[k]; x[i] = s / L[i][i]; }
    return 1;
}

static void se3_exp_d(const double d[6], double R[9], double t[3]) {
    const double wx = d[0], wy = d[1], wz = d[2], th2 = wx * wx + wy * wy + wz * wz, th = sqrt(th2);
    double A, B, C;
    if (th < 1e-6) { A = 1 - th2 / 6; B = 0.5 - th2 / 24; C = 1.0 / 6 - th2 / 120; }
    else { A = sin(th) / th; B = (1 - cos(th)) / th2; C = (th - sin(th)) / (th2 * th); }
    const double K[9] = {0, -wz, wy, wz, 0, -wx, -wy, wx, 0};
    double K2[9];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) K2[i * 3 + j] = K[i * 3] * K[j] + K[i * 3 + 1] * K[3 + j] + K[i * 3 + 2] * K[6 + j];
    double V[9];
    for (int i = 0; i < 9; ++i) { const double I = (i % 4 == 0); R[i] = I + A * K[i] + B * K2[i]; V[i] = I + B * K[i] + C * K2[i]; }
    for (int i = 0; i < 3; ++i) t[i] = V[i * 3] * d[3] + V[i * 3 + 1] * d[4] + V[i * 3 + 2] * d[5];
}

int oracle_gicp(const float* src_aos, int n_src, const float* tgt_aos, int n_tgt, const int* tgt_label, int label,
                int k, int max_iterations, double rotation_epsilon, double translation_epsilon, double* out_T12,
                double* out_error) {
    double T[12] = {1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0};
    int iters = 0;
    double err = 0;
    if (n_src > 0 && n_tgt > 0) {
        double* na = (double*)malloc(sizeof(double) * 3 * (size_t)n_src);
        double* nb = (double*)malloc(sizeof(double) * 3 * (size_t)n_tgt);
        int* corr = (int*)malloc(sizeof(int) * (size_t)n_src);
        double* Ms = (double*)malloc(sizeof(double) * 9 * (size_t)n_src);
        oracle_cloud_normals(src_aos, n_src, k, na);
        oracle_cloud_normals(tgt_aos, n_tgt, k, nb);
        const double w = 1.0 - 1e-3;
        double lambda = -1.0, best_y = 1e300, best_T[12];
        int since_best = 0;
        const char* sw = getenv("PERCH_GICP_STALL_WINDOW");
        const int stall_window = sw && *sw ? atoi(sw) : 8;
        for (int it = 0; it < max_iterations; ++it) {
            /* ---- linearise at T: correspondences, Mahalanobis matrices, H, g, y0 (LsqRegistration::linearize) */
            double H[36], g[6];
            memset(H, 0, sizeof H); memset(g, 0, sizeof g);
            err = 0;
            for (int i = 0; i < n_src; ++i) {
                const double ax = src_aos[3 * (size_t)i], ay = src_aos[3 * (size_t)i + 1], az = src_aos[3 * (size_t)i + 2];
                const double p[3] = {T[0] * ax + T[1] * ay + T[2] * az + T[9], T[3] * ax + T[4] * ay + T[5] * az + T[10],
                                     T[6] * ax + T[7] * ay + T[8] * az + T[11]};
                const float qx = (float)p[0], qy = (float)p[1], qz = (float)p[2];
                float best = FLT_MAX; int bj = -1;
                for (int j = 0; j < n_tgt; ++j) {
                    if (tgt_label && tgt_label[j] != label) continue;
                    const float dx = qx - tgt_aos[3 * (size_t)j], dy = qy - tgt_aos[3 * (size_t)j + 1], dz = qz - tgt_aos[3 * (size_t)j + 2];
                    const float d = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                    if (d < best) { best = d; bj = j; }
                }
                corr[i] = bj;
                if (bj < 0) continue;
                const double* a0 = na + 3 * (size_t)i;
                const double ra[3] = {T[0] * a0[0] + T[1] * a0[1] + T[2] * a0[2], T[3] * a0[0] + T[4] * a0[1] + T[5] * a0[2],
                                      T[6] * a0[0] + T[7] * a0[1] + T[8] * a0[2]};
                const double* b0 = nb + 3 * (size_t)bj;
                /* S = C_B + R C_A R^T = 2I - w (ra ra^T + b0 b0^T); M = S^-1 */
                const double a = 2 - w * (ra[0] * ra[0] + b0[0] * b0[0]), b = -w * (ra[0] * ra[1] + b0[0] * b0[1]),
                             c = -w * (ra[0] * ra[2] + b0[0] * b0[2]), d = 2 - w * (ra[1] * ra[1] + b0[1] * b0[1]),
                             e = -w * (ra[1] * ra[2] + b0[1] * b0[2]), f = 2 - w * (ra[2] * ra[2] + b0[2] * b0[2]);
                const double Aa = d * f - e * e, Bb = c * e - b * f, Cc = b * e - c * d, inv = 1.0 / (a * Aa + b * Bb + c * Cc);
                const double M[3][3] = {{Aa * inv, Bb * inv, Cc * inv}, {Bb * inv, (a * f - c * c) * inv, (b * c - a * e) * inv},
                                        {Cc * inv, (b * c - a * e) * inv, (a * d - b * b) * inv}};
                memcpy(Ms + 9 * (size_t)i, M, sizeof M);
                const double r[3] = {tgt_aos[3 * (size_t)bj] - p[0], tgt_aos[3 * (size_t)bj + 1] - p[1], tgt_aos[3 * (size_t)bj + 2] - p[2]};
                const double J[6][3] = {{0, p[2], -p[1]}, {-p[2], 0, p[0]}, {p[1], -p[0], 0}, {-1, 0, 0}, {0, -1, 0}, {0, 0, -1}};
                double MJ[6][3];
                for (int q = 0; q < 6; ++q) for (int x = 0; x < 3; ++x) MJ[q][x] = M[x][0] * J[q][0] + M[x][1] * J[q][1] + M[x][2] * J[q][2];
                for (int q = 0; q < 6; ++q) {
                    for (int s = 0; s < 6; ++s) H[q * 6 + s] += J[q][0] * MJ[s][0] + J[q][1] * MJ[s][1] + J[q][2] * MJ[s][2];
                    g[q] += MJ[q][0] * r[0] + MJ[q][1] * r[1] + MJ[q][2] * r[2];
                }
                for (int x = 0; x < 3; ++x) err += r[x] * (M[x][0] * r[0] + M[x][1] * r[1] + M[x][2] * r[2]);
            }
            iters = it + 1;
            const double y0 = err;
            /* stall detection (extension, see GicpArgs::stall_window): no improvement of the linearisation error on its best
             * value by 1e-3 (relative) for `stall_window` linearisations => back to that best point and stop */
            if (stall_window > 0) {
                if (y0 < best_y * (1.0 - 1e-3)) { best_y = y0; memcpy(best_T, T, sizeof T); since_best = 0; }
                else if (++since_best >= stall_window) { memcpy(T, best_T, sizeof T); err = best_y; break; }
            }
            /* ---- Levenberg-Marquardt step (fast_gicp LsqRegistration::step_lm; defaults lm_init_lambda_factor 1e-9,
             * lm_max_iterations 10): trial error with the correspondences and M of the linearisation frozen
             * (FastGICP::compute_error), gain ratio rho = (y0 - yi) / (d . (lambda d - b)) */
            if (lambda < 0) {
                double mx = 0;
                for (int c = 0; c < 6; ++c) mx = fmax(mx, H[c * 7]);
                lambda = 1e-9 * mx;
            }
            double nu = 2.0;
            int decided = 0, converged = 0;
            for (int trial = 0; trial < 10 && !decided; ++trial) {
                double Hd[36], x[6];
                memcpy(Hd, H, sizeof H);
                for (int c = 0; c < 6; ++c) Hd[c * 7] += lambda;
                if (!chol_solve6(Hd, g, x)) { decided = 1; converged = 1; break; }
                double dR[9], dt[3], nT[12];
                se3_exp_d(x, dR, dt);
                for (int i = 0; i < 3; ++i) {
                    for (int j = 0; j < 3; ++j) nT[i * 3 + j] = dR[i * 3] * T[j] + dR[i * 3 + 1] * T[3 + j] + dR[i * 3 + 2] * T[6 + j];
                    nT[9 + i] = dR[i * 3] * T[9] + dR[i * 3 + 1] * T[10] + dR[i * 3 + 2] * T[11] + dt[i];
                }
                double yi = 0;
                for (int i = 0; i < n_src; ++i) {
                    const int bj = corr[i];
                    if (bj < 0) continue;
                    const double ax = src_aos[3 * (size_t)i], ay = src_aos[3 * (size_t)i + 1], az = src_aos[3 * (size_t)i + 2];
                    const double r[3] = {tgt_aos[3 * (size_t)bj] - (nT[0] * ax + nT[1] * ay + nT[2] * az + nT[9]),
                                         tgt_aos[3 * (size_t)bj + 1] - (nT[3] * ax + nT[4] * ay + nT[5] * az + nT[10]),
                                         tgt_aos[3 * (size_t)bj + 2] - (nT[6] * ax + nT[7] * ay + nT[8] * az + nT[11])};
                    const double* M = Ms + 9 * (size_t)i;
                    for (int x2 = 0; x2 < 3; ++x2) yi += r[x2] * (M[3 * x2] * r[0] + M[3 * x2 + 1] * r[1] + M[3 * x2 + 2] * r[2]);
                }
                double den = 0, mr = 0, mt = 0;
                for (int c = 0; c < 6; ++c) den += x[c] * (lambda * x[c] - g[c]);
                for (int i = 0; i < 9; ++i) mr = fmax(mr, fabs(dR[i] - (i % 4 == 0)));
                for (int i = 0; i < 3; ++i) mt = fmax(mt, fabs(dt[i]));
                const int small = mr < rotation_epsilon && mt < translation_epsilon;
                const double rho = (y0 - yi) / den;
                if (rho < 0) {
                    if (small) { decided = 1; converged = 1; }
                    else { lambda *= nu; nu *= 2; }
                } else {
                    memcpy(T, nT, sizeof T);
                    const double q = 2 * rho - 1;
                    lambda *= fmax(1.0 / 3.0, 1 - q * q * q);
                    decided = 1;
                    converged = small;
                }
            }
            if (getenv("PERCH_ORACLE_GICP_TRACE"))
                fprintf(stderr, "  it %3d y0 %.6e lambda %.3e decided %d converged %d t=(%.5f %.5f %.5f)\n", it, y0, lambda, decided,
                        converged, T[9], T[10], T[11]);
            if (!decided || converged) break;  /* ten rejected trials ("lm not converged") or a step below the epsilons */
        }
        free(na);
        free(nb);
        free(corr);
        free(Ms);
    }
    memcpy(out_T12, T, sizeof T);
    if (out_error) *out_error = err;
    return iters;
}

/* =====================================================================================================================
 * cuda_icp package API: point-to-plane ICP of a model cloud against a depth-image scene (projective association).
 * Scene preparation: Scene_projective::init_Scene_projective_cpu (cuda_icp/scene/depth_scene/depth_scene.cpp:3-37) =
 * dep2pcd per pixel (include/cuda_icp/common.h:50-66, depth in cm read as unsigned) + get_normal
 * (cuda_icp/scene/common.cpp:3-15,17-107: depth saturated to 16 bits, 8 neighbours at radius 5, integer bilateral
 * accumulation in `long`, thresholds 2000 / 50).  Pinned bit-exact against that host code (oracle/_ref/libicp_ref.so).
 */
static long sat16l(int32_t d) { return d < 0 ? 0 : (d > 65535 ? 65535 : d); }

void oracle_icp_scene(const int32_t* depth, int W, int H, const float* K9, float* pcd, float* normals) {
    const float k00 = K9[0], k02 = K9[2], k11 = K9[4], k12 = K9[5];
    memset(normals, 0, sizeof(float) * 3 * (size_t)W * H);
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const size_t i = (size_t)y * W + x;
            const uint32_t dep = (uint32_t)depth[i];
            float px = 0, py = 0, pz = 0;
            if (dep != 0) {
                pz = dep / 100.0f;
                px = ((float)(size_t)x - k02) / k00 * pz;
                py = ((float)(size_t)y - k12) / k11 * pz;
            }
            pcd[3 * i] = px; pcd[3 * i + 1] = py; pcd[3 * i + 2] = pz;
        }
    const int r = 5;
    for (int y = r; y < H - r - 1; ++y)
        for (int x = r; x < W - r - 1; ++x) {
            const size_t i = (size_t)y * W + x;
            const long d = sat16l(depth[i]);
            if (!(d < 2000)) continue;
            long A0 = 0, A1 = 0, A3 = 0, b0 = 0, b1 = 0;
            for (int dy = -1; dy <= 1; ++dy)
                for (int dx = -1; dx <= 1; ++dx) {
                    if (!dx && !dy) continue;
                    const long ii = dx * r, jj = dy * r;
                    const long delta = sat16l(depth[(size_t)(y + dy * r) * W + (x + dx * r)]) - d;
                    const long f = labs(delta) < 50 ? 1 : 0;
                    A0 += f * ii * ii; A1 += f * ii * jj; A3 += f * jj * jj; b0 += f * ii * delta; b1 += f * jj * delta;
                }
            const long det = A0 * A3 - A1 * A1, ddx = A3 * b0 - A1 * b1, ddy = -A1 * b0 + A0 * b1;
            float nx = (float)(k00 * ddx), ny = (float)(k11 * ddy), nz = (float)(-det * d);
            const float len = sqrtf(nx * nx + ny * ny + nz * nz);
            if (len > 0) {
                const float inv = 1.0f / len;
                normals[3 * i] = nx * inv; normals[3 * i + 1] = ny * inv; normals[3 * i + 2] = nz * inv;
            }
        }
}

/*
 * ICP_Point2Plane_{cpu,cuda}<Scene_projective> (icp.cpp:117-179 / icp.cu:156-218): per iteration the 29 sums of
 * thrust__pcd2Ab (icp.h:129-210) over the points with a valid projective correspondence (Scene_projective::query,
 * depth_scene.h:29-50; pcd2dep, common.h:68-78), fitness = count / N, inlier_rmse = sqrt(sum b^2 / count); stop when
 * both change by less than the criteria, or at the extra turn max_iteration; else solve the 6x6 system in double and
 * apply R = Rz Ry Rx, t (eigen_slover_666 + TransformVector6dToMatrix4d, icp.cpp:7-36) to the cloud in place.
 * Sums are accumulated serially in float (the reference's order is unspecified: OpenMP / cub reduction), so agreement
 * with the reference is to float-accumulation accuracy, not bit for bit.  Returns the index of the last turn.
 */
/* nn = 0: Scene_projective::query (depth_scene.h:29-50).  nn = 1: Scene_nn::query (pcd_scene.h:60-136) restated as what
 * its kd-tree walk computes -- the nearest scene point (every pixel with a depth, pcd_scene.cpp:24-33), by brute force,
 * accepted iff its squared distance is below max_dist_diff^2; equal distances go to the lowest pixel index. */
static int icp_loop(int nn, float* pts, int n, const float* scene_pcd, const float* scene_normals, int W, int H,
                    const float* K9, float max_dist_diff, float relative_fitness, float relative_rmse,
                    int max_iteration, float* out_T16, float* out_fitness, float* out_rmse) {
    const float k00 = K9[0], k02 = K9[2], k11 = K9[4], k12 = K9[5];
    long* assoc = nn ? (long*)malloc(sizeof(long) * (size_t)(n > 0 ? n : 1)) : NULL;
    float T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    float fit = 0, rmse = 0;
    int last = 0;
    for (int iter = 0; iter <= max_iteration && n > 0; ++iter) {
        float sum[29];
        memset(sum, 0, sizeof sum);
        if (nn) {
#pragma omp parallel for schedule(static)
            for (int i = 0; i < n; ++i) {
                const float sx = pts[3 * i], sy = pts[3 * i + 1], sz = pts[3 * i + 2];
                float best = FLT_MAX;
                long bi = -1;
                for (long j = 0; j < (long)W * H; ++j) {
                    if (!(scene_pcd[3 * j + 2] > 0)) continue;
                    const float ex = sx - scene_pcd[3 * j], ey = sy - scene_pcd[3 * j + 1], ez = sz - scene_pcd[3 * j + 2];
                    const float d = fmaf(ez, ez, fmaf(ey, ey, ex * ex));
                    if (d < best) { best = d; bi = j; }
                }
                assoc[i] = (bi >= 0 && best < max_dist_diff * max_dist_diff) ? bi : -1;
            }
        }
        for (int i = 0; i < n; ++i) {
            const float sx = pts[3 * i], sy = pts[3 * i + 1], sz = pts[3 * i + 2];
            size_t idx;
            if (nn) {
                if (assoc[i] < 0) continue;
                idx = (size_t)assoc[i];
            } else {
                /* device arithmetic of the path being replaced (nvcc -fmad contracts x/z*K + c): fma, then + 0.5 */
                const float fu = fmaf(sx / sz, k00, k02) + 0.5f, fv = fmaf(sy / sz, k11, k12) + 0.5f;
                if (!(fu > -2147483648.0f && fu < 2147483648.0f && fv > -2147483648.0f && fv < 2147483648.0f)) continue;
                const int u = (int)fu, v = (int)fv;
                if (u >= W || v >= H || u < 0 || v < 0) continue;
                idx = (size_t)u + (size_t)v * W;
            }
            const float dx = scene_pcd[3 * idx], dy = scene_pcd[3 * idx + 1], dz = scene_pcd[3 * idx + 2];
            if (!nn && (dz <= 0 || fabsf(sz - dz) > max_dist_diff)) continue;
            const float nx = scene_normals[3 * idx], ny = scene_normals[3 * idx + 1], nz = scene_normals[3 * idx + 2];
            const float bt = (dx - sx) * nx + (dy - sy) * ny + (dz - sz) * nz;
            const float A[6] = {nz * sy - ny * sz, nx * sz - nz * sx, ny * sx - nx * sy, nx, ny, nz};
            int o = 0;
            for (int r = 0; r < 6; ++r) for (int c = r; c < 6; ++c) sum[o++] += A[r] * A[c];
            for (int r = 0; r < 6; ++r) sum[21 + r] += A[r] * bt;
            sum[27] += bt * bt;
            sum[28] += 1.0f;
        }
        last = iter;
        if (sum[28] == 0) break;
        const float pf = fit, pr = rmse;
        fit = sum[28] / (float)n;
        rmse = sqrtf(sum[27] / sum[28]);
        if (iter == max_iteration) break;
        if (fabsf(fit - pf) < relative_fitness && fabsf(rmse - pr) < relative_rmse) break;
        double Am[36], b[6], L[6][6], D[6], y[6], x[6] = {0, 0, 0, 0, 0, 0};
        int o = 0, ok = 1;
        for (int r = 0; r < 6; ++r) for (int c = r; c < 6; ++c) { Am[r * 6 + c] = sum[o]; Am[c * 6 + r] = sum[o]; ++o; }
        for (int r = 0; r < 6; ++r) b[r] = sum[21 + r];
        memset(L, 0, sizeof L);
        for (int j = 0; j < 6 && ok; ++j) {
            double s = Am[j * 6 + j];
            for (int k = 0; k < j; ++k) s -= L[j][k] * L[j][k] * D[k];
            D[j] = s;
            if (!(fabs(s) > 1e-300)) { ok = 0; break; }
            for (int i = j + 1; i < 6; ++i) { double t = Am[i * 6 + j]; for (int k = 0; k < j; ++k) t -= L[i][k] * L[j][k] * D[k]; L[i][j] = t / s; }
        }
        if (!ok) break;
        for (int i = 0; i < 6; ++i) { double s = b[i]; for (int k = 0; k < i; ++k) s -= L[i][k] * y[k]; y[i] = s; }
        for (int i = 5; i >= 0; --i) { double s = y[i] / D[i]; for (int k = i + 1; k < 6; ++k) s -= L[k][i] * x[k]; x[i] = s; }
        const double ca = cos(x[0]), sa = sin(x[0]), cb = cos(x[1]), sb = sin(x[1]), cg = cos(x[2]), sg = sin(x[2]);
        const float E[12] = {(float)(cg * cb), (float)(cg * sb * sa - sg * ca), (float)(cg * sb * ca + sg * sa), (float)x[3],
                             (float)(sg * cb), (float)(sg * sb * sa + cg * ca), (float)(sg * sb * ca - cg * sa), (float)x[4],
                             (float)(-sb), (float)(cb * sa), (float)(cb * ca), (float)x[5]};
        for (int i = 0; i < n; ++i) {
            const float px = pts[3 * i], py = pts[3 * i + 1], pz = pts[3 * i + 2];
            pts[3 * i] = E[0] * px + E[1] * py + E[2] * pz + E[3];
            pts[3 * i + 1] = E[4] * px + E[5] * py + E[6] * pz + E[7];
            pts[3 * i + 2] = E[8] * px + E[9] * py + E[10] * pz + E[11];
        }
        float nT[12];
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 4; ++c) nT[r * 4 + c] = E[r * 4] * T[c] + E[r * 4 + 1] * T[4 + c] + E[r * 4 + 2] * T[8 + c] + E[r * 4 + 3] * T[12 + c];
        memcpy(T, nT, sizeof nT);
    }
    memcpy(out_T16, T, sizeof T);
    *out_fitness = fit;
    *out_rmse = rmse;
    free(assoc);
    return last;
}

int oracle_icp_point2plane(float* pts, int n, const float* scene_pcd, const float* scene_normals, int W, int H,
                           const float* K9, float max_dist_diff, float relative_fitness, float relative_rmse,
                           int max_iteration, float* out_T16, float* out_fitness, float* out_rmse) {
    return icp_loop(0, pts, n, scene_pcd, scene_normals, W, H, K9, max_dist_diff, relative_fitness, relative_rmse,
                    max_iteration, out_T16, out_fitness, out_rmse);
}

int oracle_icp_point2plane_nn(float* pts, int n, const float* scene_pcd, const float* scene_normals, int W, int H,
                              const float* K9, float max_dist_diff, float relative_fitness, float relative_rmse,
                              int max_iteration, float* out_T16, float* out_fitness, float* out_rmse) {
    return icp_loop(1, pts, n, scene_pcd, scene_normals, W, H, K9, max_dist_diff, relative_fitness, relative_rmse,
                    max_iteration, out_T16, out_fitness, out_rmse);
}
