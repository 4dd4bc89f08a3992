// Device-side caller glue: candidate ObjectStates -> pose matrices, observed-point totals,
// IsValidPose, 3-DoF successor generation.  Kernels in states.cu; the C-ABI in capi.cu.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "perch_b200.h"

namespace perch {

// Uniform (x, y) grid over the search clouds.  Points are sorted by (range, row, column); a radius query visits the
// rows its circle's bounding box touches, each row being one contiguous run of points, and tests every point of the
// run with the float arithmetic of the reference's kd-tree leaf test (see radius_count in states.cu).
struct SearchGridDev {
    const float4* pts = nullptr;     // sorted points (w unused)
    const int* cell_start = nullptr; // n_ranges * ny * nx + 1
    const int* range_size = nullptr; // points per range
    int nx = 0, ny = 0, n_ranges = 0;
    float x0 = 0.f, y0 = 0.f, inv_cell = 0.f;
};

struct ModelMetaDev {
    double pre[16];  // preprocessing transform, row-major, cast to double (search_env.cpp:1559-1560)
    float pre_f[16];
    double circ, circ3d, insc, inflation, search_res;
    double cell_rad;  // hypot(sr/2, sr/2) of the model's search resolution (search_env.cpp:338-349)
    int symmetric, symmetry_mode, fp_off, fp_cnt;
};

struct EnvDev {
    double cam_matrix[16];  // (cam_to_world * cam_to_body)^-1, row-major (search_env.cpp:1535-1541)
    double x_min, x_max, y_min, y_max, table_height, res, footprint_tolerance;
    int use_external_pose_list, use_cylinder_observed, min_neighbor_points, use_model_specific_search_resolution;
    int min_points_for_constraint_cloud;
    int n_models;
    int max_nn_pixels;          // kNumPixels cap of the cylinder search (search_env.cpp:1607-1610)
    float observed_point_num;   // pose_obs_points_total when the cylinder is off (1618)
    int n_labels;               // entries of seg_counts
};

struct StateArgs {
    EnvDev env;
    SearchGridDev grid;
    const ModelMetaDev* metas = nullptr;
    const float* footprints = nullptr;   // xyz per hull point
    const float* constraint = nullptr;   // xyz per projected constraint point
    int n_constraint = 0;
    const float* seg_counts = nullptr;   // segmented_observed_point_count (search_env.cpp:6053-6059)
    const perch_state* states = nullptr;
    int n_states = 0;
    const perch_state* placed = nullptr;
    int n_placed = 0;
    int after_refinement = 0;
    // outputs
    float* poses16 = nullptr;
    int* model_map = nullptr;
    int* label = nullptr;
    float* obs_total = nullptr;
    uint8_t* valid = nullptr;
    int* err_flag = nullptr;
};

struct GenerateArgs {
    StateArgs base;
    const int* gen_models = nullptr;   // models to generate for (not yet placed)
    int n_gen_models = 0;
    const double* xs = nullptr;        // per generated model: offsets into the axis arrays (res may be per model)
    const double* ys = nullptr;
    const int* x_off = nullptr;        // n_gen_models + 1
    const int* y_off = nullptr;        // n_gen_models + 1
    const long long* cell_off = nullptr;  // n_gen_models + 1: first (x, y) cell of every model
    const double* thetas = nullptr;    // loop values of theta
    const double* yaws = nullptr;      // normalize_angle_positive(theta): what ContPose stores
    float* theta_rot = nullptr;        // n_thetas x 9: float rotation of ContPose(.., 0, 0, theta)
    int n_thetas = 0;
    double theta_res = 0.0;
    long long n_cells = 0;
    uint8_t* emit = nullptr;           // n_cells x n_thetas
    int* cell_count = nullptr;         // n_cells + 1 (scanned in place)
    perch_state* out = nullptr;
};

cudaError_t launch_grid_build(const float* d_xyz, int n, const int* d_ranges, int n_ranges, float x0, float y0,
                              float inv_cell, int nx, int ny, int* d_point_key, int* d_cell_start, int* d_cursor,
                              float4* d_sorted, int* d_range_size, cudaStream_t stream);
cudaError_t launch_assemble_states(const StateArgs& a, cudaStream_t stream);
cudaError_t launch_validate_states(const StateArgs& a, cudaStream_t stream);
cudaError_t launch_generate_states(const GenerateArgs& a, cudaStream_t stream);  // flags + counts + scan
cudaError_t launch_generate_write(const GenerateArgs& a, cudaStream_t stream);

}  // namespace perch
