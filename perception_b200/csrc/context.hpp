// Internal state behind perch_ctx (include/perch_b200.h).  Everything the hot path consults lives in HBM for the
// lifetime of the context; the reference re-uploads all of it on every call (renderer.cu:1532-1544).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

#include "exact_math.cuh"
#include "perch_b200.h"

namespace perch {

// Uniform grid over the observed cloud for exact radius-bounded nearest neighbour (replaces the
// N_rend x N_obs brute force of fast_gicp::brute_force_knn_search at renderer.cu:1853-1865).
// Cells are h >= 1.001*sensor_resolution wide, so any observed point within sensor_resolution of a query lies in the
// 3x3x3 cell neighbourhood; points are sorted by cell (x fastest), so each (y,z) row of 3 cells is one range.
struct ObsGridDev {
    const float4* pts = nullptr;        // xyz + bit-cast reference index (index in the label-sorted observed cloud)
    const int* label = nullptr;         // per sorted point; null in 3-DoF mode
    const uint32_t* cell_start = nullptr;  // ncells + 1
    float ox = 0, oy = 0, oz = 0;       // grid origin
    float inv_h = 0;
    int dx = 0, dy = 0, dz = 0;
    int n_obs = 0;
};

struct CameraDev {
    int width = 0, height = 0;
    Mat4 proj;
    float cx = 0, cy = 0, fx = 0, fy = 0, depth_factor = 100.f;
};

struct ModelsDev {
    // Triangle soup as cluster records (below) + packed colour (r | g<<8 | b<<16) per ORIGINAL triangle index:
    // 40 bytes per triangle instead of the reference's 64-byte AoS Triangle (model.h:62-73).
    const uint32_t* tri_color = nullptr;
    const int* model_low = nullptr;   // exclusive scan of tris_model_count (image_renderer.cuh:373-380)
    const int* model_high = nullptr;  // inclusive scan
    const float* model_aabb = nullptr;  // 6 floats per model: min xyz, max xyz (object-space)
    const float* model_qrmax = nullptr; // per model: largest finite quad bounding-sphere radius (level-1 scale bound)
    // clusters (clusters.hpp): Morton-ordered groups of kClusterTris triangles with a bounding sphere
    const float* cl_recs = nullptr;     // n_clusters x 24: rec[c][q][3 * i + a] = coordinate a of vertex i of quad q
    const uint32_t* cl_perm = nullptr;  // n_clusters: 6 bits per triangle slot, its three quad vertices in original order
    const float4* cl_bounds = nullptr;  // centre xyz, radius
    const float4* q_bounds = nullptr;   // 2 per cluster: bounding sphere of each quad
    const int* cl_orig = nullptr;       // n_clusters x 4: original triangle index (colour tie-break, colour lookup)
    const int* cl_low = nullptr;        // cluster range per model
    const int* cl_high = nullptr;
    int n_models = 0;
    int64_t n_tris = 0;
};

struct SceneDev {
    const int32_t* depth = nullptr;  // W*H, centimetres
    const uint8_t* label = nullptr;  // W*H or null
    const uint8_t* r = nullptr;      // kept for API completeness; the reference reads but never uses them
    const uint8_t* g = nullptr;
    const uint8_t* b = nullptr;
};

struct StatesCtx;  // capi_states.cu

}  // namespace perch

struct perch_ctx {
    // 128-bit digest of the last perch_set_models input (triangles + counts): an identical call returns at once
    uint64_t models_digest[2] = {0, 0};
    // the same for perch_set_observed (host cloud); cleared whenever the observed cloud is replaced any other way
    uint64_t obs_digest[2] = {0, 0};
    bool obs_digest_valid = false;
    int device = 0;
    int sm_count = 0;
    int max_smem_optin = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;  // perch_timer_start / perch_timer_stop
    float last_ms = 0.f;
    float last_gicp_ms = 0.f;          // device time of the last GICP kernel
    bool timing_pending = false;
    int* d_err_flag = nullptr;
    int last_launches = 0;
    perch_gpu_stats stats{0.f, 0.0};

    perch::CameraDev cam;
    bool have_cam = false;

    // models
    uint32_t* d_tri_color = nullptr;
    int* d_model_low = nullptr;
    int* d_model_high = nullptr;
    float* d_model_aabb = nullptr;
    float* d_model_qrmax = nullptr;
    float* d_tri_lab = nullptr;        // 3 x T CIELAB of the triangle colours (cost_type 1)
    float* d_cl_recs = nullptr;
    uint32_t* d_cl_perm = nullptr;
    float4* d_cl_bounds = nullptr;
    float4* d_q_bounds = nullptr;
    int* d_cl_orig = nullptr;
    int* d_cl_low = nullptr;
    int* d_cl_high = nullptr;
    std::vector<int> h_model_low, h_model_high;
    perch::ModelsDev models;
    bool have_models = false;

    // scene
    size_t src_px = 0;  // pixels d_src_depth / d_src_rgb (and d_src_label, when present) are allocated for
    int32_t* d_src_depth = nullptr;
    uint8_t* d_src_label = nullptr;
    uint8_t* d_src_rgb = nullptr;  // 3 planes
    perch::SceneDev scene;
    bool have_scene = false;
    // composition mode (capi_compose.cu): while `composed`, scene.depth / scene.label point at these instead of the
    // observed images above; perch_set_scene / perch_set_input / perch_compose_source(mode NONE) switch back
    int32_t* d_cmp_depth = nullptr;
    uint8_t* d_cmp_label = nullptr;
    bool composed = false;
    // full-resolution rasteriser (raster_full.cu): key images of one chunk of poses, deferred-triangle list, counters
    unsigned long long* d_full_keys = nullptr;
    size_t full_keys_n = 0;            // keys allocated (all empty between calls)
    int2* d_full_big = nullptr;
    int* d_full_count = nullptr;
    int* d_full_acc = nullptr;
    int* d_full_rects = nullptr;       // 4 ints per pose of a chunk: key-image rectangle the pose touched
    size_t full_rects_n = 0;
    size_t full_acc_n = 0;
    int max_model_quads = 0;

    // observed cloud (host copies kept so the grid can be rebuilt when sensor_resolution changes)
    // The label-sorted cloud lives on the device (d_obs_xyz4 / d_obs_label_sorted below); the host mirrors are filled by
    // perch_set_observed (host input) and only on demand after perch_set_input (perch_detail::ensure_obs_host)
    bool obs_host_valid = true;
    bool obs_nonfinite = false;        // a coordinate of the host cloud was not finite (reported when the grid is built)
    float obs_min[3] = {0, 0, 0}, obs_max[3] = {0, 0, 0};  // bounding box of the cloud
    int* d_obs_label_sorted = nullptr; // sorted labels (null in 3-DoF)
    int* d_obs_tmp = nullptr;          // label-sort histograms / grid cursors / small readbacks
    size_t obs_tmp_n = 0;
    // perch_set_input: the compact cloud in the order depth2cloud_global returns it, kept on the device for
    // perch_get_observed (xyz in_n x 3, colours 3 x in_n, labels in_n)
    float* d_in_xyz = nullptr;
    uint8_t* d_in_rgb = nullptr;
    int* d_in_label = nullptr;
    int in_n = 0;
    bool in_on_device = false;
    std::vector<float> h_obs_xyz;      // label-sorted, AoS
    std::vector<uint8_t> h_obs_rgb;    // original order, 3 x N SoA (reference never permutes colours, renderer.cu:1881)
    std::vector<int> h_obs_label;      // sorted labels (empty in 3-DoF)
    bool obs_has_label = false;
    int n_obs = 0;
    float grid_res = -1.f;             // sensor_resolution the current grid was built for
    float4* d_grid_pts = nullptr;
    int* d_grid_label = nullptr;
    uint32_t* d_cell_start = nullptr;
    uint8_t* d_obs_rgb = nullptr;      // 3 x N, original order
    float* d_obs_lab = nullptr;        // 3 x N CIELAB of observed colours (cost_type 1), indexed by reference index
    perch::ObsGridDev grid;
    double grid_avg_per_cell = 0.0;    // points per non-empty cell of `grid`
    bool have_obs = false;
    // GICP's own grid over the same cloud (finer cells when the cloud is dense; else a copy of `grid`), capi.cu
    float4* d_ggrid_pts = nullptr;
    int* d_ggrid_label = nullptr;
    uint32_t* d_ggrid_cell_start = nullptr;
    perch::ObsGridDev gicp_grid;
    float gicp_grid_h = 0.f;
    double gicp_grid_avg = 1.0;  // points per non-empty cell of gicp_grid
    bool gicp_grid_valid = false;
    // pyramid above gicp_grid (levels 1..): own buffers per level; d_gicp_levels = the ObsGridDev array incl. level 0
    static constexpr int kGicpMaxLevels = 14;
    float4* d_lvl_pts[kGicpMaxLevels] = {};
    int* d_lvl_label[kGicpMaxLevels] = {};
    uint32_t* d_lvl_cell_start[kGicpMaxLevels] = {};
    perch::ObsGridDev* d_gicp_levels = nullptr;
    int gicp_n_levels = 0;

    // per-call scratch (grown on demand, never shrunk)
    void* d_scratch = nullptr;
    size_t scratch_bytes = 0;
    // GICP (gicp.cu): target normals are cached per observed cloud; the work buffer grows on demand
    float* d_obs_xyz_aos = nullptr;    // label-sorted observed cloud, N x 3
    float* d_tgt_normals = nullptr;    // N x 3
    float4* d_obs_xyz4 = nullptr;      // the same cloud as float4 (streaming nearest-neighbour search of short ranges)
    int* d_label_range = nullptr;      // [begin, end) per label in the sorted cloud
    int n_label_ranges = 0;
    int tgt_normals_k = 0;             // k the cached normals were computed with (0 = stale)
    void* d_gicp = nullptr;
    size_t gicp_bytes = 0;
    // cuda_icp package API (icp.cu): the prepared depth-image scene
    float* d_icp_pcd = nullptr;        // W*H x 3
    float* d_icp_normals = nullptr;    // W*H x 3
    int icp_w = 0, icp_h = 0;
    float icp_K[9] = {0};
    float icp_max_dist = 0.1f;
    // Scene_nn association: grid over the valid scene points, built on first use for a given max_dist_diff
    float4* d_icp_grid_pts = nullptr;
    int* d_icp_grid_label = nullptr;   // unused (the builder's slot)
    uint32_t* d_icp_grid_cells = nullptr;
    perch::ObsGridDev icp_grid;
    float icp_grid_dist = -1.f;        // max_dist_diff the grid was built for (< 0: stale)
    // caller glue (capi_states.cu): environment, model metadata, search clouds, assembled / generated state lists
    perch::StatesCtx* states = nullptr;
    int obs_version = 0;               // bumped by perch_set_observed (per-label point counts are cached against it)
    // perch_set_input (capi_input.cu): the observed cloud in the order depth2cloud_global returns it (for the caller's
    // host-side copies, search_env.cpp:6031-6060) and the label counts
    std::vector<float> h_in_xyz;
    std::vector<uint8_t> h_in_rgb;
    std::vector<int> h_in_label;
    bool in_has_label = false;
    void* d_flush = nullptr;           // 256 MiB L2-flush target (perch_l2_flush)
    int flush_value = 0;
    void* h_pinned = nullptr;
    size_t pinned_bytes = 0;
    // NCCL communicator of a sharded pose list (capi_multi.cu): one rank per context, either one process per GPU
    // (perch_comm_init_rank) or all devices of one process (perch_create_multi); null = single GPU, no gather
    void* comm = nullptr;              // ncclComm_t
    int comm_rank = 0, comm_size = 1;
    bool comm_owned = false;
};
