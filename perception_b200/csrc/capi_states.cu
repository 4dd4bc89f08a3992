// C-ABI of the caller glue (include/perch_b200.h, "caller glue" section): host orchestration of states.cu.
// Reference: sbpl_perception/src/search_env.cpp:309-528, 1505-1622, 7056-7490 (see states.cu).
#include <cuda_runtime.h>

#include <cstring>
#include <string>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "capi_common.hpp"
#include "states.cuh"

namespace perch {

struct StatesCtx {
    perch_env env{};
    bool have_env = false;
    std::vector<perch_model_meta> metas;
    EnvDev env_dev{};
    ModelMetaDev* d_metas = nullptr;
    float* d_footprints = nullptr;
    // search clouds
    bool have_search = false;
    int n_search = 0;
    float* d_search_xyz = nullptr;
    float4* d_sorted = nullptr;
    int* d_cell_start = nullptr;
    int* d_cursor = nullptr;
    int* d_key = nullptr;
    int* d_ranges = nullptr;
    int* d_range_size = nullptr;
    SearchGridDev grid;
    float* d_constraint = nullptr;
    int n_constraint = 0;
    // segmented_observed_point_count, derived from the labels of perch_set_observed
    float* d_seg_counts = nullptr;
    int n_labels = 0;
    int seg_version = -1;
    // one state batch: states, poses16, model map, label, observed totals, validity
    void* d_batch = nullptr;
    size_t batch_bytes = 0;
    // generated successor list and its work buffers
    perch_state* d_gen = nullptr;
    size_t gen_capacity = 0;
    int n_gen = 0;
    void* d_gen_work = nullptr;
    size_t gen_work_bytes = 0;
};

}  // namespace perch

using perch_detail::DeviceGuard;
using perch_detail::fail;

namespace {

size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }

perch::StatesCtx* states_of(perch_ctx* c) {
    if (!c->states) c->states = new perch::StatesCtx();
    return c->states;
}

template <class T>
int replace_device(T** dst, const T* src, size_t n, cudaStream_t s) {
    if (*dst) {
        cudaFree(*dst);
        *dst = nullptr;
    }
    PERCH_CU(cudaMalloc(dst, std::max<size_t>(n, 1) * sizeof(T)));
    if (src && n) PERCH_CU(cudaMemcpyAsync(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice, s));
    return PERCH_OK;
}

int grow(void** p, size_t* have, size_t need) {
    if (need <= *have) return PERCH_OK;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *have = 0;
    PERCH_CU(cudaMalloc(p, need));
    *have = need;
    return PERCH_OK;
}

// Eigen's generic 4x4 inverse (LU/InverseImpl.h, compute_inverse_size4): cofactors, then division by
// det = sum_i m(i,0) * result(0,i).  Row-major in and out.
double det3_helper(const double* m, int i1, int i2, int i3, int j1, int j2, int j3) {
    return m[4 * i1 + j1] * (m[4 * i2 + j2] * m[4 * i3 + j3] - m[4 * i2 + j3] * m[4 * i3 + j2]);
}
double cofactor4(const double* m, int i, int j) {
    const int i1 = (i + 1) % 4, i2 = (i + 2) % 4, i3 = (i + 3) % 4, j1 = (j + 1) % 4, j2 = (j + 2) % 4, j3 = (j + 3) % 4;
    return det3_helper(m, i1, i2, i3, j1, j2, j3) + det3_helper(m, i2, i3, i1, j1, j2, j3) + det3_helper(m, i3, i1, i2, j1, j2, j3);
}
void inverse4(const double* m, double* r) {
    for (int row = 0; row < 4; ++row)
        for (int col = 0; col < 4; ++col) {
            const double cf = cofactor4(m, col, row);
            r[4 * row + col] = ((row + col) & 1) ? -cf : cf;
        }
    const double det = ((m[0] * r[0] + m[4] * r[1]) + m[8] * r[2]) + m[12] * r[3];
    for (int k = 0; k < 16; ++k) r[k] /= det;
}

// search_env.cpp:1535-1541: cam_matrix = (cam_to_world_ * cam_to_body).matrix().inverse()
void camera_matrix(const double* cam_to_world, double* out) {
    static const double cam_to_body[16] = {0, 0, 1, 0, -1, 0, 0, 0, 0, -1, 0, 0, 0, 0, 0, 1};
    double z_front[16];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            z_front[4 * i + j] = ((cam_to_world[4 * i] * cam_to_body[j] + cam_to_world[4 * i + 1] * cam_to_body[4 + j]) +
                                  cam_to_world[4 * i + 2] * cam_to_body[8 + j]) + cam_to_world[4 * i + 3] * cam_to_body[12 + j];
    inverse4(z_front, out);
}

// angles::normalize_angle_positive (ROS angles.h): fmod(fmod(angle, 2pi) + 2pi, 2pi); what ContPose's Euler constructor
// applies to roll, pitch, yaw (object_state.cpp:17-22)
double normalize_angle_positive(double a) { return std::fmod(std::fmod(a, 2.0 * M_PI) + 2.0 * M_PI, 2.0 * M_PI); }

int refresh_seg_counts(perch_ctx* c, perch::StatesCtx* st) {
    if (st->seg_version == c->obs_version) return PERCH_OK;
    // search_env.cpp:6024-6059: number of observed points per label
    {
        const int rc0 = perch_detail::ensure_obs_host(c);
        if (rc0) return rc0;
    }
    int n_labels = 0;
    if (c->obs_has_label)
        for (int l : c->h_obs_label) n_labels = std::max(n_labels, l + 1);
    std::vector<float> counts(std::max(n_labels, 1), 0.f);
    if (c->obs_has_label)
        for (int l : c->h_obs_label)
            if (l >= 0) counts[l] += 1.f;
    int rc = replace_device<float>(&st->d_seg_counts, counts.data(), counts.size(), c->stream);
    if (rc) return rc;
    PERCH_CU(cudaStreamSynchronize(c->stream));
    st->n_labels = n_labels;
    st->seg_version = c->obs_version;
    return PERCH_OK;
}

int check_states_ready(perch_ctx* c, bool need_search) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    if (!c->states || !c->states->have_env) return fail(PERCH_ERR_INVALID, "perch_set_env has not been called");
    if (need_search && !c->states->have_search) return fail(PERCH_ERR_INVALID, "perch_set_search_cloud has not been called");
    return PERCH_OK;
}

struct Batch {
    perch_state* states;
    float* poses16;
    int* model_map;
    int* label;
    float* obs_total;
    uint8_t* valid;
    perch_state* placed;
};

int take_batch(perch_ctx* c, perch::StatesCtx* st, size_t n, size_t n_placed, Batch& b) {
    const size_t need = align256(n * sizeof(perch_state)) + align256(n * 64) + 3 * align256(n * 4) + align256(n) +
                        align256(n_placed * sizeof(perch_state)) + 256;
    int rc = grow(&st->d_batch, &st->batch_bytes, need);
    if (rc) return rc;
    char* p = static_cast<char*>(st->d_batch);
    b.states = reinterpret_cast<perch_state*>(p); p += align256(n * sizeof(perch_state));
    b.poses16 = reinterpret_cast<float*>(p); p += align256(n * 64);
    b.model_map = reinterpret_cast<int*>(p); p += align256(n * 4);
    b.label = reinterpret_cast<int*>(p); p += align256(n * 4);
    b.obs_total = reinterpret_cast<float*>(p); p += align256(n * 4);
    b.valid = reinterpret_cast<uint8_t*>(p); p += align256(n);
    b.placed = reinterpret_cast<perch_state*>(p);
    return PERCH_OK;
}

int fill_args(perch_ctx* c, perch::StatesCtx* st, perch::StateArgs& a) {
    int rc = refresh_seg_counts(c, st);
    if (rc) return rc;
    a.env = st->env_dev;
    a.env.n_labels = st->n_labels;
    a.env.observed_point_num = (float)c->n_obs;
    a.env.max_nn_pixels = c->have_cam ? c->cam.width * c->cam.height : 640 * 480;  // kNumPixels
    a.grid = st->grid;
    a.metas = st->d_metas;
    a.footprints = st->d_footprints;
    a.constraint = st->d_constraint;
    a.n_constraint = st->n_constraint;
    a.seg_counts = st->d_seg_counts;
    a.err_flag = c->d_err_flag;
    return PERCH_OK;
}

int read_flag(perch_ctx* c) {
    int flag = 0;
    PERCH_CU(cudaMemcpyAsync(&flag, c->d_err_flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    PERCH_CU(cudaStreamSynchronize(c->stream));
    if (flag) {
        cudaMemsetAsync(c->d_err_flag, 0, sizeof(int), c->stream);
        if (flag & 2) return fail(PERCH_ERR_INVALID, "a state names a model_id outside [0, n_models)");
        if (flag & 4) return fail(PERCH_ERR_INVALID, "a state's segmentation_label_id has no entry in the observed labels");
        return fail(PERCH_ERR_CUDA, "internal error flag %d", flag);
    }
    return PERCH_OK;
}

// assemble `n` device-resident states into b (device)
int assemble_on_device(perch_ctx* c, perch::StatesCtx* st, const perch_state* d_states, int n, const Batch& b) {
    perch::StateArgs a;
    int rc = fill_args(c, st, a);
    if (rc) return rc;
    if (a.env.use_external_pose_list != 1 && a.env.use_cylinder_observed && !st->have_search)
        return fail(PERCH_ERR_INVALID, "use_cylinder_observed needs perch_set_search_cloud (the projected cloud)");
    a.states = d_states;
    a.n_states = n;
    a.poses16 = b.poses16; a.model_map = b.model_map; a.label = b.label; a.obs_total = b.obs_total;
    PERCH_CU(perch::launch_assemble_states(a, c->stream));
    return PERCH_OK;
}

}  // namespace

void perch_detail::destroy_states(perch_ctx* c) {
    perch::StatesCtx* st = c->states;
    if (!st) return;
    void* ptrs[] = {st->d_metas, st->d_footprints, st->d_search_xyz, st->d_sorted, st->d_cell_start, st->d_cursor, st->d_key,
                    st->d_ranges, st->d_range_size, st->d_constraint, st->d_seg_counts, st->d_batch, st->d_gen, st->d_gen_work};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    delete st;
    c->states = nullptr;
}

extern "C" {

int perch_set_env(perch_ctx* c, const perch_env* env, const perch_model_meta* metas, int n_models, const float* footprint_xyz,
                  int n_footprint_points) {
    if (!c || !env || !metas || n_models <= 0 || n_footprint_points < 0 || (n_footprint_points > 0 && !footprint_xyz))
        return fail(PERCH_ERR_INVALID, "bad perch_set_env argument");
    if (c->have_models && c->models.n_models != n_models)
        return fail(PERCH_ERR_INVALID, "perch_set_env: %d models, but perch_set_models holds %d", n_models, c->models.n_models);
    DeviceGuard g(c->device);
    perch::StatesCtx* st = states_of(c);
    std::vector<perch::ModelMetaDev> dev(n_models);
    for (int m = 0; m < n_models; ++m) {
        const perch_model_meta& s = metas[m];
        if (s.footprint_offset < 0 || s.footprint_count < 0 || s.footprint_offset + s.footprint_count > n_footprint_points)
            return fail(PERCH_ERR_INVALID, "model %d: footprint range outside the footprint array", m);
        perch::ModelMetaDev& d = dev[m];
        for (int k = 0; k < 16; ++k) {
            d.pre[k] = (double)s.preprocessing_transform[k];  // .matrix().cast<double>() (search_env.cpp:1559-1560)
            d.pre_f[k] = s.preprocessing_transform[k];
        }
        d.circ = s.circumscribed_radius; d.circ3d = s.circumscribed_radius_3d; d.insc = s.inscribed_radius;
        d.inflation = s.inflation_factor; d.search_res = s.search_resolution;
        const double sr = env->use_model_specific_search_resolution ? s.search_resolution : env->res;
        d.cell_rad = std::hypot(sr / 2.0, sr / 2.0);  // search_env.cpp:347-348
        d.symmetric = s.symmetric; d.symmetry_mode = s.symmetry_mode;
        d.fp_off = s.footprint_offset; d.fp_cnt = s.footprint_count;
    }
    int rc;
    if ((rc = replace_device<perch::ModelMetaDev>(&st->d_metas, dev.data(), dev.size(), c->stream))) return rc;
    if ((rc = replace_device<float>(&st->d_footprints, footprint_xyz, 3 * (size_t)n_footprint_points, c->stream))) return rc;
    PERCH_CU(cudaStreamSynchronize(c->stream));
    st->env = *env;
    st->metas.assign(metas, metas + n_models);
    perch::EnvDev& e = st->env_dev;
    camera_matrix(env->cam_to_world, e.cam_matrix);
    e.x_min = env->x_min; e.x_max = env->x_max; e.y_min = env->y_min; e.y_max = env->y_max;
    e.table_height = env->table_height; e.res = env->res; e.footprint_tolerance = env->footprint_tolerance;
    e.use_external_pose_list = env->use_external_pose_list;
    e.use_cylinder_observed = env->use_cylinder_observed;
    e.min_neighbor_points = env->min_neighbor_points_for_valid_pose;
    e.use_model_specific_search_resolution = env->use_model_specific_search_resolution;
    e.min_points_for_constraint_cloud = env->min_points_for_constraint_cloud;
    e.n_models = n_models;
    st->have_env = true;
    st->n_gen = 0;
    return PERCH_OK;
}

int perch_set_search_cloud(perch_ctx* c, const float* xyz_aos, int n_points, const int* ranges, int n_ranges,
                           const float* constraint_xyz, int n_constraint) {
    if (!c || n_points < 0 || (n_points > 0 && !xyz_aos) || n_constraint < 0 || (n_constraint > 0 && !constraint_xyz))
        return fail(PERCH_ERR_INVALID, "bad perch_set_search_cloud argument");
    if (ranges && (n_ranges <= 0 || n_ranges > 64)) return fail(PERCH_ERR_INVALID, "n_ranges must be in [1, 64]");
    DeviceGuard g(c->device);
    perch::StatesCtx* st = states_of(c);
    std::vector<int> rg;
    if (ranges) rg.assign(ranges, ranges + 2 * (size_t)n_ranges);
    else { rg = {0, n_points}; n_ranges = 1; }
    for (int k = 0; k < n_ranges; ++k)
        if (rg[2 * k] < 0 || rg[2 * k + 1] < rg[2 * k] || rg[2 * k + 1] > n_points)
            return fail(PERCH_ERR_INVALID, "range %d = [%d, %d) is outside the cloud", k, rg[2 * k], rg[2 * k + 1]);
    // grid extent over the finite points; at most 128 x 128 cells, at least 1 mm wide
    float mn[2] = {INFINITY, INFINITY}, mx[2] = {-INFINITY, -INFINITY};
    for (int i = 0; i < n_points; ++i)
        for (int k = 0; k < 2; ++k) {
            const float v = xyz_aos[3 * (size_t)i + k];
            if (std::isfinite(v)) { mn[k] = std::min(mn[k], v); mx[k] = std::max(mx[k], v); }
        }
    if (!(mn[0] <= mx[0]) || !(mn[1] <= mx[1])) { mn[0] = mn[1] = 0.f; mx[0] = mx[1] = 0.f; }
    const float extent = std::max(mx[0] - mn[0], mx[1] - mn[1]);
    const float cell = std::max(extent / 128.f, 1e-3f);
    const int nx = std::min(128, (int)((mx[0] - mn[0]) / cell) + 1), ny = std::min(128, (int)((mx[1] - mn[1]) / cell) + 1);
    const size_t n_keys = (size_t)n_ranges * nx * ny;
    int rc;
    if ((rc = replace_device<float>(&st->d_search_xyz, xyz_aos, 3 * (size_t)n_points, c->stream))) return rc;
    if ((rc = replace_device<float4>(&st->d_sorted, nullptr, n_points, c->stream))) return rc;
    if ((rc = replace_device<int>(&st->d_cell_start, nullptr, n_keys + 2, c->stream))) return rc;
    if ((rc = replace_device<int>(&st->d_cursor, nullptr, n_keys + 1, c->stream))) return rc;
    if ((rc = replace_device<int>(&st->d_key, nullptr, n_points, c->stream))) return rc;
    if ((rc = replace_device<int>(&st->d_ranges, rg.data(), rg.size(), c->stream))) return rc;
    if ((rc = replace_device<int>(&st->d_range_size, nullptr, n_ranges, c->stream))) return rc;
    if ((rc = replace_device<float>(&st->d_constraint, constraint_xyz, 3 * (size_t)n_constraint, c->stream))) return rc;
    PERCH_CU(perch::launch_grid_build(st->d_search_xyz, n_points, st->d_ranges, n_ranges, mn[0], mn[1], 1.0f / cell, nx, ny,
                                      st->d_key, st->d_cell_start, st->d_cursor, st->d_sorted, st->d_range_size, c->stream));
    PERCH_CU(cudaStreamSynchronize(c->stream));
    st->grid.pts = st->d_sorted;
    st->grid.cell_start = st->d_cell_start;
    st->grid.range_size = st->d_range_size;
    st->grid.nx = nx; st->grid.ny = ny; st->grid.n_ranges = n_ranges;
    st->grid.x0 = mn[0]; st->grid.y0 = mn[1]; st->grid.inv_cell = 1.0f / cell;
    st->n_search = n_points;
    st->n_constraint = n_constraint;
    st->have_search = true;
    st->n_gen = 0;
    return PERCH_OK;
}

int perch_assemble_states(perch_ctx* c, const perch_state* states, int n_states, float* out_poses16, int* out_model_map,
                          int* out_label, float* out_obs_total) {
    int rc = check_states_ready(c, false);
    if (rc) return rc;
    if (n_states < 0 || (n_states > 0 && !states)) return fail(PERCH_ERR_INVALID, "bad state list");
    DeviceGuard g(c->device);
    perch::StatesCtx* st = c->states;
    Batch b;
    if ((rc = take_batch(c, st, n_states, 0, b))) return rc;
    if (n_states == 0) return PERCH_OK;
    PERCH_CU(cudaMemcpyAsync(b.states, states, (size_t)n_states * sizeof(perch_state), cudaMemcpyHostToDevice, c->stream));
    if ((rc = assemble_on_device(c, st, b.states, n_states, b))) return rc;
    const size_t P = n_states;
    if (out_poses16) PERCH_CU(cudaMemcpyAsync(out_poses16, b.poses16, P * 64, cudaMemcpyDeviceToHost, c->stream));
    if (out_model_map) PERCH_CU(cudaMemcpyAsync(out_model_map, b.model_map, P * 4, cudaMemcpyDeviceToHost, c->stream));
    if (out_label) PERCH_CU(cudaMemcpyAsync(out_label, b.label, P * 4, cudaMemcpyDeviceToHost, c->stream));
    if (out_obs_total) PERCH_CU(cudaMemcpyAsync(out_obs_total, b.obs_total, P * 4, cudaMemcpyDeviceToHost, c->stream));
    return read_flag(c);
}

int perch_score_states(perch_ctx* c, const char* stage, const perch_state* states, int n_states, const perch_params* params,
                       perch_outputs* out) {
    int rc = check_states_ready(c, false);
    if (rc) return rc;
    DeviceGuard g(c->device);
    perch::StatesCtx* st = c->states;
    const perch_state* d_states = nullptr;
    Batch b;
    if (states) {
        if (n_states < 0) return fail(PERCH_ERR_INVALID, "bad state list");
        if ((rc = take_batch(c, st, n_states, 0, b))) return rc;
        if (n_states > 0)
            PERCH_CU(cudaMemcpyAsync(b.states, states, (size_t)n_states * sizeof(perch_state), cudaMemcpyHostToDevice, c->stream));
        d_states = b.states;
    } else {
        n_states = st->n_gen;
        if ((rc = take_batch(c, st, n_states, 0, b))) return rc;
        d_states = st->d_gen;
    }
    if (n_states > 0 && (rc = assemble_on_device(c, st, d_states, n_states, b))) return rc;
    // 6-DoF passes the segmentation labels, 3-DoF an empty vector (search_env.cpp:1580-1590, 1683)
    perch_detail::PoseInputs in{b.poses16, b.model_map, st->env_dev.use_external_pose_list == 1 ? b.label : nullptr, b.obs_total, true};
    rc = perch_detail::render_cost_impl(c, stage, in, n_states, params, out);
    if (rc) return rc;
    return PERCH_OK;
}

int perch_validate_states(perch_ctx* c, const perch_state* states, int n_states, const perch_state* placed, int n_placed,
                          int after_refinement, uint8_t* out_valid) {
    int rc = check_states_ready(c, true);
    if (rc) return rc;
    if (n_states < 0 || n_placed < 0 || (n_states > 0 && (!states || !out_valid)) || (n_placed > 0 && !placed))
        return fail(PERCH_ERR_INVALID, "bad perch_validate_states argument");
    if (n_states == 0) return PERCH_OK;
    DeviceGuard g(c->device);
    perch::StatesCtx* st = c->states;
    Batch b;
    if ((rc = take_batch(c, st, n_states, n_placed, b))) return rc;
    PERCH_CU(cudaMemcpyAsync(b.states, states, (size_t)n_states * sizeof(perch_state), cudaMemcpyHostToDevice, c->stream));
    if (n_placed > 0)
        PERCH_CU(cudaMemcpyAsync(b.placed, placed, (size_t)n_placed * sizeof(perch_state), cudaMemcpyHostToDevice, c->stream));
    perch::StateArgs a;
    if ((rc = fill_args(c, st, a))) return rc;
    a.states = b.states; a.n_states = n_states;
    a.placed = b.placed; a.n_placed = n_placed;
    a.after_refinement = after_refinement;
    a.valid = b.valid;
    PERCH_CU(perch::launch_validate_states(a, c->stream));
    PERCH_CU(cudaMemcpyAsync(out_valid, b.valid, (size_t)n_states, cudaMemcpyDeviceToHost, c->stream));
    return read_flag(c);
}

int perch_generate_states(perch_ctx* c, const perch_state* placed, int n_placed, perch_state* out_states, int capacity,
                          int* n_out) {
    int rc = check_states_ready(c, true);
    if (rc) return rc;
    if (!n_out || n_placed < 0 || (n_placed > 0 && !placed) || capacity < 0 || (capacity > 0 && !out_states))
        return fail(PERCH_ERR_INVALID, "bad perch_generate_states argument");
    DeviceGuard g(c->device);
    perch::StatesCtx* st = c->states;
    const perch_env& env = st->env;
    if (env.use_external_pose_list == 1)
        return fail(PERCH_ERR_INVALID, "6-DoF candidates come from poses.txt (search_env.cpp:7091-7266): use perch_validate_states");
    if (!(env.theta_res > 0.0)) return fail(PERCH_ERR_INVALID, "theta_res must be positive");
    const int n_models = st->env_dev.n_models;
    // models already in the source state generate nothing (search_env.cpp:7067-7075)
    std::vector<int> gen_models;
    for (int m = 0; m < n_models; ++m) {
        bool placed_already = false;
        for (int i = 0; i < n_placed; ++i) placed_already |= placed[i].model_id == m;
        if (!placed_already) gen_models.push_back(m);
    }
    // the loop variables, accumulated in double exactly like `for (double x = x_min; x <= x_max; x += res)` (7276-7283)
    std::vector<double> xs, ys, thetas, yaws;
    std::vector<int> x_off{0}, y_off{0};
    std::vector<long long> cell_off{0};
    const size_t kAxisCap = 1u << 20;
    for (int m : gen_models) {
        const perch_model_meta& mm = st->metas[m];
        const double sr = env.use_model_specific_search_resolution ? mm.search_resolution : env.res;
        const double res = env.use_adaptive_resolution ? mm.inscribed_radius : sr;  // 7088-7089
        if (!(res > 0.0)) return fail(PERCH_ERR_INVALID, "model %d: grid resolution %g is not positive", m, res);
        size_t nx = 0, ny = 0;
        for (double x = env.x_min; x <= env.x_max; x += res) {
            xs.push_back(x);
            if (++nx > kAxisCap) return fail(PERCH_ERR_CAPACITY, "more than 2^20 grid steps along x");
        }
        for (double y = env.y_min; y <= env.y_max; y += res) {
            ys.push_back(y);
            if (++ny > kAxisCap) return fail(PERCH_ERR_CAPACITY, "more than 2^20 grid steps along y");
        }
        x_off.push_back((int)xs.size());
        y_off.push_back((int)ys.size());
        cell_off.push_back(cell_off.back() + (long long)nx * (long long)ny);
    }
    for (double theta = 0; theta < 2 * M_PI; theta += env.theta_res) {
        thetas.push_back(theta);
        yaws.push_back(normalize_angle_positive(theta));
        if (thetas.size() > 4096) return fail(PERCH_ERR_CAPACITY, "more than 4096 yaw steps");
    }
    const long long n_cells = cell_off.back();
    const size_t n_thetas = thetas.size();
    if ((double)n_cells * (double)n_thetas > 2e9) return fail(PERCH_ERR_CAPACITY, "more than 2e9 grid candidates");
    const size_t G = gen_models.size();
    // work buffer layout
    size_t need = align256(G * 4 + 4) + align256(xs.size() * 8 + 8) + align256(ys.size() * 8 + 8) + 2 * align256((G + 1) * 4) +
                  align256((G + 1) * 8) + 2 * align256(n_thetas * 8 + 8) + align256(n_thetas * 36 + 4) +
                  align256((size_t)n_cells * n_thetas + 1) + align256(((size_t)n_cells + 1) * 4) +
                  align256((size_t)n_placed * sizeof(perch_state) + 8);
    if ((rc = grow(&st->d_gen_work, &st->gen_work_bytes, need))) return rc;
    char* p = static_cast<char*>(st->d_gen_work);
    auto take = [&](size_t bytes) { char* q = p; p += align256(bytes); return q; };
    int* d_gen_models = (int*)take(G * 4 + 4);
    double* d_xs = (double*)take(xs.size() * 8 + 8);
    double* d_ys = (double*)take(ys.size() * 8 + 8);
    int* d_x_off = (int*)take((G + 1) * 4);
    int* d_y_off = (int*)take((G + 1) * 4);
    long long* d_cell_off = (long long*)take((G + 1) * 8);
    double* d_thetas = (double*)take(n_thetas * 8 + 8);
    double* d_yaws = (double*)take(n_thetas * 8 + 8);
    float* d_theta_rot = (float*)take(n_thetas * 36 + 4);
    uint8_t* d_emit = (uint8_t*)take((size_t)n_cells * n_thetas + 1);
    int* d_cell_count = (int*)take(((size_t)n_cells + 1) * 4);
    perch_state* d_placed = (perch_state*)take((size_t)n_placed * sizeof(perch_state) + 8);
    auto h2d = [&](void* dst, const void* src, size_t bytes) -> cudaError_t {
        return bytes ? cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->stream) : cudaSuccess;
    };
    PERCH_CU(h2d(d_gen_models, gen_models.data(), G * 4));
    PERCH_CU(h2d(d_xs, xs.data(), xs.size() * 8));
    PERCH_CU(h2d(d_ys, ys.data(), ys.size() * 8));
    PERCH_CU(h2d(d_x_off, x_off.data(), (G + 1) * 4));
    PERCH_CU(h2d(d_y_off, y_off.data(), (G + 1) * 4));
    PERCH_CU(h2d(d_cell_off, cell_off.data(), (G + 1) * 8));
    PERCH_CU(h2d(d_thetas, thetas.data(), n_thetas * 8));
    PERCH_CU(h2d(d_yaws, yaws.data(), n_thetas * 8));
    PERCH_CU(h2d(d_placed, placed, (size_t)n_placed * sizeof(perch_state)));
    perch::GenerateArgs ga;
    if ((rc = fill_args(c, st, ga.base))) return rc;
    ga.base.placed = d_placed; ga.base.n_placed = n_placed; ga.base.after_refinement = 0;
    ga.gen_models = d_gen_models; ga.n_gen_models = (int)G;
    ga.xs = d_xs; ga.ys = d_ys; ga.x_off = d_x_off; ga.y_off = d_y_off; ga.cell_off = d_cell_off;
    ga.thetas = d_thetas; ga.yaws = d_yaws; ga.theta_rot = d_theta_rot; ga.n_thetas = (int)n_thetas;
    ga.theta_res = env.theta_res;
    ga.n_cells = n_cells;
    ga.emit = d_emit; ga.cell_count = d_cell_count;
    PERCH_CU(perch::launch_generate_states(ga, c->stream));
    int total = 0;
    PERCH_CU(cudaMemcpyAsync(&total, d_cell_count + n_cells, 4, cudaMemcpyDeviceToHost, c->stream));
    PERCH_CU(cudaStreamSynchronize(c->stream));
    if ((size_t)total > st->gen_capacity) {
        if (st->d_gen) cudaFree(st->d_gen);
        st->d_gen = nullptr;
        st->gen_capacity = 0;
        PERCH_CU(cudaMalloc(&st->d_gen, (size_t)total * sizeof(perch_state)));
        st->gen_capacity = total;
    }
    ga.out = st->d_gen;
    if (total > 0) PERCH_CU(perch::launch_generate_write(ga, c->stream));
    st->n_gen = total;
    *n_out = total;
    const size_t ncopy = std::min<size_t>(total, capacity);
    if (ncopy > 0 && out_states)
        PERCH_CU(cudaMemcpyAsync(out_states, st->d_gen, ncopy * sizeof(perch_state), cudaMemcpyDeviceToHost, c->stream));
    return read_flag(c);
}

}  // extern "C"

// ---- 6-DoF candidate intake: the poses.txt reader of GenerateSuccessorStates (search_env.cpp:7109-7146) ---------------------
// Each line holds `x y z qx qy qz qw` separated by single spaces; the reference splits on ' ' (boost::is_any_of(" "), no token
// compression), converts every token with std::stod and builds ContPose(x, y, z, qx, qy, qz, qw) from the first seven
// (object_state.cpp:24-41).  A token std::stod would throw on (empty after a double space, not a number) is an error here
// instead of an uncaught exception; a trailing '\r' is ignored like stod ignores it.  Host-only.
extern "C" int perch_parse_pose_list(const char* text, size_t length, int model_id, int segmentation_label_id, perch_state* out_states,
                                     int capacity, int* n_out) {
    using perch_detail::fail;
    if ((!text && length > 0) || !n_out || capacity < 0 || (capacity > 0 && !out_states))
        return fail(PERCH_ERR_INVALID, "bad parse_pose_list argument");
    int n = 0, line_no = 0;
    size_t pos = 0;
    while (pos < length) {
        size_t eol = pos;
        while (eol < length && text[eol] != '\n') ++eol;
        ++line_no;
        double v[7];
        int nv = 0;
        size_t tok = pos;
        bool any = eol > pos;
        while (any && tok <= eol) {
            size_t end = tok;
            while (end < eol && text[end] != ' ') ++end;
            if (nv < 7) {
                std::string t(text + tok, end - tok);
                size_t used = 0;
                try {
                    v[nv] = std::stod(t, &used);
                } catch (...) {
                    return fail(PERCH_ERR_INVALID, "pose list line %d: token %d ('%s') is not a number", line_no, nv + 1, t.c_str());
                }
            }
            ++nv;
            if (end >= eol) break;
            tok = end + 1;
        }
        if (any) {
            if (nv < 7) return fail(PERCH_ERR_INVALID, "pose list line %d: %d values, expected x y z qx qy qz qw", line_no, nv);
            if (n < capacity) {
                perch_state st;
                memset(&st, 0, sizeof st);
                st.x = v[0]; st.y = v[1]; st.z = v[2];
                st.qx = v[3]; st.qy = v[4]; st.qz = v[5]; st.qw = v[6];
                st.model_id = model_id;
                st.segmentation_label_id = segmentation_label_id;
                out_states[n] = st;
            }
            ++n;
        }
        pos = eol + 1;
    }
    *n_out = n;
    return PERCH_OK;
}
