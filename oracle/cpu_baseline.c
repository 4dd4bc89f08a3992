/*
 * cpu_baseline.c -- TEST / BENCH INFRASTRUCTURE ONLY (timed by bench.py as `cpu_baseline` and `--impl reference`).
 *
 * CPU restatement ("port") of the reference's NON-GPU scoring path for one candidate pose, run over a batch of poses
 * with OpenMP (the reference distributes candidates over MPI ranks, search_env.cpp:920-1023, `mpirun -n 6` in
 * fat_dataset/perch.py:177-179).  PCL, OpenGL (kinect_sim) and ROS are absent from this image, so:
 *   - the render is the reference's own CPU rasterizer structure (cuda_renderer/src/renderer.cpp:228-330) via
 *     oracle_render, including the source-occlusion rule, instead of the kinect_sim OpenGL render
 *     (search_env.cpp:5267-5373; graphics: unavailable on this box);
 *   - pcl::search::KdTree::radiusSearch(r = sensor_resolution, max 1 neighbour) is an exact uniform-grid "any point
 *     within r" test.
 * Cost semantics follow GetTargetCost (search_env.cpp:4037-4135: a rendered point is unexplained if no observed point
 * lies within sensor_resolution; cost = (int)(100 * unexplained / N), 100 if N < min_points) and GetSourceCost
 * (search_env.cpp:4137-4361: observed points inside the candidate's footprint that have no rendered point within
 * sensor_resolution; same normalisation).  It is a reported baseline, not a parity oracle: its observed-cost
 * definition differs from the GPU path's nearest-neighbour-index rule (hard part 3).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { float x, y, z; } bvec3;
typedef struct { bvec3 v0, v1, v2; uint64_t c0, c1, c2; } btri64;

void oracle_render(const btri64* tris, int n_tris, const float* poses16, int n_poses, const int* pose_model_map,
                   const int* tris_model_count, int n_models, int width, int height, const float* proj16,
                   const int32_t* source_depth, const uint8_t* source_label, const int* pose_label,
                   float occlusion_threshold, int single_result_image, int32_t* out_depth, uint8_t* out_r,
                   uint8_t* out_g, uint8_t* out_b);

typedef struct {
    float org[3], inv_h;
    int dim[3];
    int* start;  /* ncells + 1 */
    float* pts;  /* sorted xyz */
    int* label;  /* sorted labels or NULL */
    int n;
} grid_t;

static void grid_cell(const grid_t* g, const float* p, int* c) {
    for (int k = 0; k < 3; ++k) {
        int v = (int)floorf((p[k] - g->org[k]) * g->inv_h);
        c[k] = v;
    }
}

static void grid_build(grid_t* g, const float* xyz, const int* label, int n, float h) {
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = 0; i < n; ++i)
        for (int k = 0; k < 3; ++k) {
            if (xyz[3 * i + k] < mn[k]) mn[k] = xyz[3 * i + k];
            if (xyz[3 * i + k] > mx[k]) mx[k] = xyz[3 * i + k];
        }
    g->n = n;
    g->inv_h = 1.0f / h;
    size_t cells = 1;
    for (int k = 0; k < 3; ++k) {
        g->org[k] = n ? mn[k] - h : 0;
        g->dim[k] = n ? (int)floorf((mx[k] - g->org[k]) * g->inv_h) + 2 : 1;
        cells *= (size_t)g->dim[k];
    }
    g->start = (int*)calloc(cells + 1, sizeof(int));
    g->pts = (float*)malloc(sizeof(float) * 3 * (size_t)(n > 0 ? n : 1));
    g->label = label ? (int*)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1)) : NULL;
    int* cell = (int*)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; ++i) {
        int c[3];
        grid_cell(g, xyz + 3 * i, c);
        cell[i] = (c[2] * g->dim[1] + c[1]) * g->dim[0] + c[0];
        g->start[cell[i] + 1]++;
    }
    for (size_t k = 0; k < cells; ++k) g->start[k + 1] += g->start[k];
    int* cur = (int*)malloc(sizeof(int) * cells);
    memcpy(cur, g->start, sizeof(int) * cells);
    for (int i = 0; i < n; ++i) {
        int s = cur[cell[i]]++;
        memcpy(g->pts + 3 * s, xyz + 3 * i, 12);
        if (label) g->label[s] = label[i];
    }
    free(cur);
    free(cell);
}

static void grid_free(grid_t* g) { free(g->start); free(g->pts); free(g->label); }

/* radiusSearch(point, r, ..., max_nn = 1) > 0 */
static int grid_any_within(const grid_t* g, const float* q, float r2, int want_label) {
    if (g->n == 0) return 0;
    int c[3];
    grid_cell(g, q, c);
    for (int z = c[2] - 1; z <= c[2] + 1; ++z) {
        if (z < 0 || z >= g->dim[2]) continue;
        for (int y = c[1] - 1; y <= c[1] + 1; ++y) {
            if (y < 0 || y >= g->dim[1]) continue;
            int x0 = c[0] - 1 < 0 ? 0 : c[0] - 1, x1 = c[0] + 1 >= g->dim[0] ? g->dim[0] - 1 : c[0] + 1;
            if (x0 > x1) continue;
            int row = (z * g->dim[1] + y) * g->dim[0];
            for (int j = g->start[row + x0]; j < g->start[row + x1 + 1]; ++j) {
                if (g->label && want_label >= 0 && g->label[j] != want_label) continue;
                float dx = q[0] - g->pts[3 * j], dy = q[1] - g->pts[3 * j + 1], dz = q[2] - g->pts[3 * j + 2];
                if (dx * dx + dy * dy + dz * dz <= r2) return 1;
            }
        }
    }
    return 0;
}

/* torchrun exports OMP_NUM_THREADS=1 to every rank; the baseline is quoted on ALL host cores, so bench.py sets the
 * thread count explicitly (n <= 0: leave the OpenMP default). */
void cpu_baseline_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int cpu_baseline_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/*
 * Scores n_poses candidates.  footprint: for 3-DoF pass cyl_center_xy (2 floats per pose, table frame), cyl_radius
 * (per pose) and table_from_cam (row-major 3x4) -- the observed points considered are those inside the cylinder
 * (search_env.cpp:4183-4254); for 6-DoF pass pose_label/obs_label -- the pose's segment.  Outputs ints like the
 * reference (target_cost, source_cost).
 */
void cpu_baseline_score(const btri64* tris, int n_tris, const float* poses16, int n_poses, const int* pose_model_map,
                        const int* tris_model_count, int n_models, int width, int height, const float* proj16,
                        const int32_t* source_depth, const uint8_t* source_label, const int* pose_label, float cx,
                        float cy, float fx, float fy, float depth_factor, int stride, const float* obs_xyz, const int* obs_label,
                        int n_obs, const float* cyl_center_xy, const float* cyl_radius, const float* table_from_cam,
                        float sensor_resolution, float occlusion_threshold, int min_points, int* target_cost,
                        int* source_cost) {
    const float r2 = sensor_resolution * sensor_resolution;
    grid_t og;
    grid_build(&og, obs_xyz, obs_label, n_obs, sensor_resolution * 1.01f);
    /* observed points in the table frame (xy only) for the cylinder test */
    float* obs_xy = NULL;
    if (table_from_cam) {
        obs_xy = (float*)malloc(sizeof(float) * 2 * (size_t)(n_obs > 0 ? n_obs : 1));
        for (int i = 0; i < n_obs; ++i)
            for (int k = 0; k < 2; ++k)
                obs_xy[2 * i + k] = table_from_cam[4 * k] * obs_xyz[3 * i] + table_from_cam[4 * k + 1] * obs_xyz[3 * i + 1] +
                                    table_from_cam[4 * k + 2] * obs_xyz[3 * i + 2] + table_from_cam[4 * k + 3];
    }
    const size_t npx = (size_t)width * height;
#pragma omp parallel
    {
        int32_t* depth = (int32_t*)malloc(npx * 4);
        uint8_t* rgb = (uint8_t*)malloc(npx * 3);
        float* cloud = (float*)malloc(sizeof(float) * 3 * (npx / (size_t)(stride * stride) + (size_t)width + 16));
#pragma omp for schedule(dynamic, 4)
        for (int p = 0; p < n_poses; ++p) {
            const int one_label = pose_label ? pose_label[p] : 0;
            oracle_render(tris, n_tris, poses16 + 16 * (size_t)p, 1, pose_model_map + p, tris_model_count, n_models, width,
                          height, proj16, source_depth, source_label, pose_label ? &one_label : NULL, occlusion_threshold,
                          0, depth, rgb, rgb + npx, rgb + 2 * npx);
            /* rendered cloud (camera frame), stride-sampled like the GPU path */
            int n = 0;
            for (int y = 0; y < height; y += stride)
                for (int x = 0; x < width; x += stride) {
                    int32_t d = depth[(size_t)y * width + x];
                    if (d <= 0) continue;
                    float z = (float)d / depth_factor;
                    cloud[3 * n] = ((float)x - cx) / fx * z;
                    cloud[3 * n + 1] = ((float)y - cy) / fy * z;
                    cloud[3 * n + 2] = z;
                    ++n;
                }
            /* GetTargetCost */
            int unexplained = 0;
            for (int i = 0; i < n; ++i)
                if (!grid_any_within(&og, cloud + 3 * i, r2, pose_label ? one_label : -1)) ++unexplained;
            target_cost[p] = (n < min_points) ? 100 : (int)((double)unexplained * 100.0 / (double)n);
            /* GetSourceCost: kd-tree on the rendered cloud, rebuilt per candidate (search_env.cpp:4146-4148) */
            if (n == 0) { source_cost[p] = 100; continue; }
            grid_t rg;
            grid_build(&rg, cloud, NULL, n, sensor_resolution * 1.01f);
            int considered = 0, src_unexpl = 0;
            for (int i = 0; i < n_obs; ++i) {
                int inside;
                if (pose_label) inside = obs_label && obs_label[i] == one_label;
                else if (obs_xy) {
                    float dx = obs_xy[2 * i] - cyl_center_xy[2 * p], dy = obs_xy[2 * i + 1] - cyl_center_xy[2 * p + 1];
                    inside = dx * dx + dy * dy <= cyl_radius[p] * cyl_radius[p];
                } else inside = 1;
                if (!inside) continue;
                ++considered;
                if (!grid_any_within(&rg, obs_xyz + 3 * i, r2, -1)) ++src_unexpl;
            }
            source_cost[p] = (considered < min_points) ? 100 : (int)((double)src_unexpl * 100.0 / (double)considered);
            grid_free(&rg);
        }
        free(depth);
        free(rgb);
        free(cloud);
    }
    grid_free(&og);
    free(obs_xy);
}
