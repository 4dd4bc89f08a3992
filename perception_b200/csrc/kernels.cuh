// Launch interfaces of the sm_100a kernels behind libperch_b200.so.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "context.hpp"

namespace perch {

// ---- per-pose kernels (pose_kernels.cu): fused render -> cost, and the rasterizer alone ------------------------
struct PoseArgs {
    ModelsDev models;
    CameraDev cam;
    SceneDev scene;
    ObsGridDev grid;
    const Mat4* poses = nullptr;
    const int* pose_model_map = nullptr;
    const int* pose_label = nullptr;        // null => 3-DoF occlusion rule, unrestricted NN
    const float* pose_obs_total = nullptr;  // may be null
    int n_poses = 0;
    int stride = 1;                   // lattice pitch in pixels (1 = every pixel)
    int calc_observed = 0;
    float res2 = 0.f;                 // sensor_resolution^2 (renderer.cu:1877)
    float color_thr = 0.f;
    float occlusion_threshold = 0.f;
    const float* tri_lab = nullptr;   // 3 floats per triangle (cost_type 1)
    const float* obs_lab = nullptr;   // 3 floats per observed point (cost_type 1)
    // MODE_COST outputs
    float* out_costs = nullptr;       // 3 x P: rendered, observed, points_diff
    int32_t* out_counts = nullptr;    // 3 x P or null
    // MODE_IMAGE outputs: P x Lh x Lw lattice images (Lw = W/stride, Lh = ceil(H/stride)); stride 1 = full images
    int32_t* img_depth = nullptr;
    uint8_t* img_r = nullptr;
    uint8_t* img_g = nullptr;
    uint8_t* img_b = nullptr;
    // single_result_image (image_renderer.cuh:262-275): all poses z-tested into one W x H key image, resolved by
    // launch_resolve_single afterwards
    unsigned long long* single_keys = nullptr;
    int key_pose_id = 0;              // single_keys only: the key's low word is the pose index instead of the triangle
                                      // index (perch_compose_source labels each pixel by the placed object that owns it)
    // Composition-mode outputs per pose (the reference's USE_TREE / USE_CLUTTER branches, image_renderer.cuh:163-206),
    // evaluated per covered pixel on its minimum depth; all null unless perch_tree_flags asked for them.  With these
    // set img_depth may be null (nothing is written but the four vectors).
    int* tree_occluded = nullptr;        // the render lies in front of a source pixel (by more than the rule's gate)
    int* tree_occluded_other = nullptr;  // a source pixel hides (clears) a rendered pixel
    float* tree_clutter = nullptr;       // 100 * cleared pixels / covered pixels
    float* tree_total = nullptr;         // covered pixels
    int* err_flag = nullptr;          // set to 1 if a fragment fell outside its pose tile (must never happen)
    int tile_cap = 0;                 // lattice cells of shared-memory z-buffer per band
    int mask_words = 0;               // ceil(n_obs / 32) (MODE_COST)
    // A/B variant PERCH_COST_SMEM_OBS only (the mechanism BASELINE.json's north_star names): the observed points of the
    // pose's label are bulk-copied (cp.async.bulk) into shared memory and searched linearly from there
    const float4* obs_sorted = nullptr;   // label-sorted observed cloud (index = reference index)
    const int* label_range = nullptr;     // [begin, end) per label
    int n_label_ranges = 0;
};
size_t pose_kernel_smem_bytes(const PoseArgs& a, bool image_mode, bool color);
cudaError_t launch_fused_cost(const PoseArgs& a, bool color, cudaStream_t stream);
cudaError_t launch_raster_images(const PoseArgs& a, cudaStream_t stream);
// single_result_image: clear keys / resolve keys -> depth + rgb with the occlusion rule (label of pose 0, as the
// reference's pointer arithmetic does, image_renderer.cuh:273-274).
cudaError_t launch_single_clear(unsigned long long* keys, int n, cudaStream_t stream);
cudaError_t launch_resolve_single(const PoseArgs& a, cudaStream_t stream);

// ---- per-model top-k of a scored batch (topk.cu) -----------------------------------------------------------------------
constexpr int kTopkMaxModels = 255;       // the key's top byte carries the model; 0xFF marks a skipped pose
constexpr int kTopkMaxPoses = 1 << 24;    // ... and its low 24 bits the pose's index inside the call
struct TopkArgs {
    const float* costs = nullptr;          // 3 x n_poses as the fused kernel writes them: rendered, observed, points_diff
    const int* pose_model_map = nullptr;
    int n_poses = 0;
    int pose_id_offset = 0;                // global id of the call's first pose (rank offset of a sharded list)
    int n_models = 0;
    int k = 0;
    unsigned long long* keys = nullptr;    // n_poses scratch
    perch_hit* hits = nullptr;             // n_models x k, unordered within a model, cost = INT32_MAX pads
};
cudaError_t launch_topk(const TopkArgs& a, cudaStream_t stream);

// ---- full-resolution rasteriser (raster_full.cu): stages RENDER / DEBUG, composition, perch_tree_flags ---------------
struct FullRasterArgs {
    ModelsDev models;
    CameraDev cam;
    SceneDev scene;
    const Mat4* poses = nullptr;
    const int* pose_model_map = nullptr;
    const int* pose_label = nullptr;        // null => 3-DoF occlusion rule
    int n_poses = 0;
    float occlusion_threshold = 0.f;
    unsigned long long* keys = nullptr;     // n_poses x W*H (depth, triangle) keys (the resolve pass leaves every key
                                            // empty again), or ONE image when `single`
    int single = 0;                         // single_result_image: every pose z-tests into keys[0 .. W*H)
    int key_pose_id = 0;                    // low word of the key = pose index instead of original triangle index
    int2* big_list = nullptr;               // (pose, triangle slot | perm << 26) of triangles with large bboxes
    int* big_count = nullptr;
    int big_cap = 0;
    int big_ctas = 0;
    int big_cells = 4096;                   // bbox pixels above which a triangle is deferred to the warp-per-triangle kernel
    int lane_cells = 16;                    // bbox pixels up to which a lane walks its own triangle (above: the whole warp)
    // resolve outputs: n_poses x W*H images (any may be null), and the per-pose vectors of PoseArgs::tree_*
    int32_t* img_depth = nullptr;
    uint8_t* img_r = nullptr;
    uint8_t* img_g = nullptr;
    uint8_t* img_b = nullptr;
    int wide_stores = 0;                    // W*H % 4 == 0 and every image plane 16/4-byte aligned
    int* tree_occluded = nullptr;
    int* tree_occluded_other = nullptr;
    float* tree_clutter = nullptr;
    float* tree_total = nullptr;
    int* tree_acc = nullptr;                // 3 x n_poses scratch (flags, cleared, covered)
    // bounding rectangle (image columns / rows) of everything the rasterising kernels z-tested for each pose of the chunk:
    // rect_min = 2 ints per pose (x, y; initialised 0x7f7f7f7f), rect_max = 2 per pose (initialised 0x80808080).  The
    // resolve pass reads keys only inside it (all keys outside are empty) -- or everywhere when null.
    int* rect_min = nullptr;
    int* rect_max = nullptr;
};
cudaError_t launch_full_raster(const FullRasterArgs& a, int max_model_quads, cudaStream_t stream);
cudaError_t launch_full_resolve(const FullRasterArgs& a, int n_images, cudaStream_t stream);
// Host driver: poses in chunks whose key images stay L2-resident (or all poses into `single_keys`, raster only).
struct FullJob {
    const Mat4* poses = nullptr;
    const int* pose_model_map = nullptr;
    const int* pose_label = nullptr;
    int n_poses = 0;
    float occlusion_threshold = 0.f;
    unsigned long long* single_keys = nullptr;  // cleared by the caller; resolved by the caller (resolve_single / compose)
    int key_pose_id = 0;
    int32_t* img_depth = nullptr;
    uint8_t* img_r = nullptr;
    uint8_t* img_g = nullptr;
    uint8_t* img_b = nullptr;
    int* tree_occluded = nullptr;
    int* tree_occluded_other = nullptr;
    float* tree_clutter = nullptr;
    float* tree_total = nullptr;
};
cudaError_t full_raster_job(::perch_ctx* c, const FullJob& j, int* launches);

// ---- composition mode (compose.cu): key image of the placed objects -> source depth (+ label) image ----------------
struct ComposeArgs {
    const unsigned long long* keys = nullptr;  // W*H single_result_image keys with key_pose_id
    const int* pose_label = nullptr;           // per placed pose (0-based, like pose_segmentation_label), or null
    const int32_t* base_depth = nullptr;       // the observed source image (merge mode), or null
    const uint8_t* base_label = nullptr;
    int32_t* out_depth = nullptr;
    uint8_t* out_label = nullptr;              // or null
    int n_pixels = 0;
    int merge = 0;
};
cudaError_t launch_compose(const ComposeArgs& a, cudaStream_t stream);

// ---- depth image(s) -> compacted cloud (cloud.cu) -----------------------------------------------------------
struct CloudArgs {
    // Input images are num_images x img_h x img_w.  Samples are taken every `stride` pixels of the input image; a
    // sample at input pixel (ix,iy) is camera pixel (ix*coord_scale, iy*coord_scale).  Observed image: coord_scale 1,
    // stride = gpu_stride.  Lattice images written by the rasterizer (stage CLOUD): stride 1, coord_scale = gpu_stride.
    const int32_t* depth = nullptr;
    const uint8_t* red = nullptr;    // may be null
    const uint8_t* green = nullptr;
    const uint8_t* blue = nullptr;
    const uint8_t* label_mask = nullptr;  // observed image labels or null
    const int* pose_label = nullptr;      // rendered: per-pose label or null
    int img_w = 0, img_h = 0, num_images = 0, stride = 1, coord_scale = 1;
    int full_h = 0;  // camera image height (for dc_index)
    // cuda_icp::depth2cloud_cuda's coordinates (icp.cu:239-257): camera pixel = LATTICE index + offset (x + tl_x, not
    // x * stride + tl_x) -- lattice_coords = 1; u_off / v_off are added in either mode
    int lattice_coords = 0, u_off = 0, v_off = 0;
    float cx = 0, cy = 0, fx = 0, fy = 0, depth_factor = 100.f;
    bool use_bounds = false;
    float bounds[6] = {0, 0, 0, 0, 0, 0};  // already cast to float like compute_point_clouds.cuh:86-88
    float cam_tf[16] = {0};                 // column-major 4x4
    // scratch / outputs
    int* row_counts = nullptr;   // num_images * Lh + 1 (exclusive-scanned in place)
    float* cloud_soa = nullptr;  // 3 x capacity
    float* cloud_aos = nullptr;  // capacity x 3
    uint8_t* cloud_rgb = nullptr;
    int* pose_map = nullptr;
    int* cloud_label = nullptr;
    int* dc_index = nullptr;     // num_images*H*W or null
    int capacity = 0;
};
cudaError_t launch_cloud_count(const CloudArgs& a, cudaStream_t stream);
cudaError_t launch_cloud_scan(const CloudArgs& a, cudaStream_t stream);  // exclusive scan of row_counts, total in last
cudaError_t launch_cloud_write(const CloudArgs& a, cudaStream_t stream);

// in-place exclusive scan of n ints by one CTA; element n receives the total (cloud.cu)
cudaError_t launch_exclusive_scan(int* d_data, int n, cudaStream_t stream);

// ---- the observed cloud on the device (obs_grid.cu): bounding box, stable label sort, uniform grid ------------------------
cudaError_t launch_obs_bbox(const float* d_aos, int n, int* d_out6, cudaStream_t stream);  // 6 order-encoded ints
float obs_bbox_decode(int ord);
int label_sort_chunks(int n);
// stable sort by label in [0, 256): d_hist = label_sort_chunks(n) x 256 ints of scratch, d_ranges = 512 ints ([begin, end) per
// label), *d_bad is set when a label is out of range
cudaError_t launch_label_sort(const float* d_aos, const int* d_label, int n, int* d_hist, int* d_ranges, int* d_bad,
                              float4* d_out_xyz4, int* d_out_label, cudaStream_t stream);
cudaError_t launch_aos_to_xyz4(const float* d_aos, int n, float4* d_out, cudaStream_t stream);
// uniform grid over the (label-sorted) cloud; g carries origin / inv_h / dims; d_cell_start ncells + 1, d_cursor ncells scratch
cudaError_t launch_obs_grid_build(const float4* d_sorted, const int* d_sorted_label, int n, const ObsGridDev& g,
                                  uint32_t* d_cell_start, int* d_cursor, float4* d_out_pts, int* d_out_label, int* d_nonempty,
                                  cudaStream_t stream);  // *d_nonempty (may be null) = number of non-empty cells

// ---- colour preparation (cost_type 1) ------------------------------------------------------------------------
// lab[3*i..] = rgb2lab(c2[i], c1[i], c0[i]) -- the reference feeds plane 2 as "red" (compute_costs.cuh:214-220).
cudaError_t launch_rgb2lab_planes(const uint8_t* c0, const uint8_t* c1, const uint8_t* c2, int n, float* lab,
                                  cudaStream_t stream);
cudaError_t launch_rgb2lab_packed(const uint32_t* packed_rgb, int n, float* lab, cudaStream_t stream);

// ---- exact kNN with the knncuda.h contract (knn.cu) -----------------------------------------------------------------
// d_ref: dim x ref_nb, d_query: dim x query_nb (row per dimension); outputs k x query_nb.  Requires k <= ref_nb.
cudaError_t launch_knn(const float* d_ref, int ref_nb, const float* d_query, int query_nb, int dim, int k, float* d_dist,
                       int* d_index, int sm_count, cudaStream_t stream);

// ---- batched GICP refinement (gicp.cu) ---------------------------------------------------------------------------------
struct GicpArgs {
    const float* cloud_aos = nullptr;    // rendered clouds of all poses, N x 3 (metres, camera frame), pose-major
    const int* pose_begin = nullptr;     // exclusive scan of per-lattice-row point counts (cloud.cu): pose p owns
    int rows_per_pose = 1;               //   points [pose_begin[p*rows_per_pose], pose_begin[(p+1)*rows_per_pose])
    float* src_normals = nullptr;        // N x 3 scratch (used when a pose's cloud exceeds smem_points)
    int* src_ref = nullptr;              // N scratch: correspondences of the last linearisation (same condition)
    ObsGridDev grid;                     // observed cloud (target): cells sized for a few points each
    float grid_h = 0.f;                  // cell size of `grid`
    ObsGridDev coarse;                   // the cost stage's grid over the same cloud (== grid when the cloud is sparse)
    float coarse_h = 0.f;
    const float* tgt_normals = nullptr;  // n_obs x 3, indexed like the grid's reference indices
    const float4* tgt_sorted = nullptr;  // the label-sorted observed cloud itself (index = reference index), or null
    const int* label_range = nullptr;    // [begin, end) of every label in tgt_sorted (6-DoF)
    int n_label_ranges = 0;
    int stream_max = 0;                  // ranges up to this many points are searched by streaming instead of the grid
    const int* pose_label = nullptr;     // null => target unrestricted
    const Mat4* poses_in = nullptr;
    Mat4* poses_out = nullptr;           // T * pose, x100-scaled like the inputs (renderer.cu:1412-1429)
    float* out_transforms = nullptr;     // P x 12 (R row-major, t) in metres, or null
    int* out_iterations = nullptr;       // or null
    float* out_error = nullptr;          // sum r^T M r of the last linearisation, or null
    int smem_points = 0;                 // source points (and normals) a CTA keeps in shared memory
    int k_correspondences = 10;          // renderer.cu:1697
    int max_iterations = 150;            // renderer.cu:1696
    float rotation_epsilon = 2e-3f;      // renderer.cu:1698
    float translation_epsilon = 5e-4f;   // renderer.cu:1699
    // Extension (not in fast_gicp): hard nearest-neighbour assignment makes some poses cycle through the same few
    // correspondence sets for ever, each step above the epsilons, until the iteration cap.  When the error of the
    // linearisation point has not improved on its best value by stall_rel (relative) for stall_window consecutive
    // linearisations, the pose stops and returns the transform of that best linearisation.  stall_window = 0: off.
    int stall_window = 8;
    // warm start of the nearest-neighbour search from the previous linearisation's correspondence when it lies within
    // this many (fine) cell widths of the moved point; beyond it the search starts cold (PERCH_GICP_WARM_CELLS)
    float warm_cells = 1e6f;
    float stall_rel = 1e-3f;
    // Pyramid of grids over the target (level 0 = `grid`, cells doubling per level, the last level has at most two cells per
    // axis): a search whose ball has radius r runs on the level whose cells are about r / 2 wide, so it opens a bounded
    // number of rows whatever r is (a far query on the finest grid opens (2r / h)^2 of them).  Device array, or null.
    const ObsGridDev* levels = nullptr;
    int n_levels = 0;
    // level choice: rows opened ~ (r / h)^2, points tested ~ density * (2 r h + h^2) => best cell size h* = cbrt(level_k * r)
    // with level_k = kappa * h0^2 / (points per non-empty level-0 cell); kappa = cost of a row step in point tests
    float level_k = 0.f;
};
size_t gicp_smem_bytes(int smem_points);
cudaError_t launch_gicp(const GicpArgs& a, int n_poses, cudaStream_t stream);
// normals (direction of least variance of the k nearest neighbours, launch_knn's k x n index matrix) of a cloud
cudaError_t launch_target_normals(const float* d_xyz_aos, const int* d_knn_index, int n, int k, float* d_normals,
                                  cudaStream_t stream);
// k = 10 through the search grid (same neighbours, same order as the brute-force k-NN; nothing visits the host)
cudaError_t launch_target_normals_grid(const ObsGridDev& grid, float cell, const float4* d_pts_sorted, int n, float* d_normals,
                                       cudaStream_t stream);
// max over poses of the per-pose point count (pose_begin as above) -> *d_out (must be zeroed)
cudaError_t launch_pose_count_max(const int* d_pose_begin, int rows_per_pose, int n_poses, int* d_out, cudaStream_t stream);

// ---- point-to-plane ICP against a depth-image scene, cuda_icp package API (icp.cu) ------------------------------------------
struct IcpArgs {
    float* points = nullptr;          // all model clouds, N x 3, transformed IN PLACE (like the reference's model_pcd)
    const int* offsets = nullptr;     // n_clouds + 1
    const float* scene_pcd = nullptr; // W*H x 3 (dep2pcd of the scene depth)
    const float* scene_normals = nullptr;
    int width = 0, height = 0;
    float k00 = 0, k02 = 0, k11 = 0, k12 = 0;
    float max_dist_diff = 0.1f;       // depth_scene.h:9 (projective) / pcd_scene.h:48 (nearest neighbour, 0.01)
    // Scene_nn (pcd_scene.h:47-136): association by exact nearest scene point within max_dist_diff instead of by
    // projection; nn_grid holds the valid scene points (w = pixel index into scene_pcd / scene_normals) in cells just
    // wider than max_dist_diff.  nn = 0 => Scene_projective.
    int nn = 0;
    ObsGridDev nn_grid;
    float relative_fitness = 1e-5f, relative_rmse = 1e-5f;  // ICPConvergenceCriteria, icp.h:40-51
    int max_iteration = 30;
    float* out_T = nullptr;           // n_clouds x 16, row-major 4x4
    float* out_fitness = nullptr;
    float* out_rmse = nullptr;
    int* out_iterations = nullptr;    // or null
};
cudaError_t launch_icp_scene(const int32_t* d_depth, int W, int H, const float K9[9], float* d_pcd, float* d_normals,
                             cudaStream_t stream);
cudaError_t launch_icp_point2plane(const IcpArgs& a, int n_clouds, cudaStream_t stream);

// ---- observed-scene preparation (input.cu) -----------------------------------------------------------------------------
constexpr int kMedianMaxKsize = 31;
cudaError_t launch_median_blur_u8(const uint8_t* in, uint8_t* out, int w, int h, int ksize, cudaStream_t stream);
cudaError_t launch_depth_to_i32(const void* in, int bits, int32_t* out, int n, cudaStream_t stream);
cudaError_t launch_divide_depth(const int32_t* in, float factor, int32_t* out, int n, cudaStream_t stream);

}  // namespace perch
