/*
 * perch_b200.h -- C-ABI of libperch_b200.so: the B200-native (sm_100a) implementation of PERCH 2.0's
 * data-parallel hot path (batched candidate-pose rendering + per-pose depth/colour cost, depth->cloud,
 * KNN and batched GICP refinement).
 *
 * This is the drop-in boundary for what sbpl_perception::EnvObjectRecognition calls in the reference's
 * cuda_renderer package.  Every entry point cites the reference interface it replaces (paths relative to
 * the reference repo).  Plain pointers and sizes only; all functions return 0 on success and a non-zero
 * perch_status otherwise (perch_last_error() gives the text); nothing here exit()s or writes to stdout.
 *
 * Unlike the reference, which re-uploads meshes and the scene on every call (renderer.cu:1532-1544), state is
 * held in a context: set the models / camera / scene / observed cloud once, then score batches of poses.
 *
 * Units follow the reference: meshes in metres, poses carry a x100 scale in their top three rows
 * (model.h:89-107) so rendered depth is integer centimetres; depth_factor converts back (100 => metres).
 */
#ifndef PERCH_B200_H
#define PERCH_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PERCH_B200_ABI_VERSION 1

typedef enum perch_status {
    PERCH_OK = 0,
    PERCH_ERR_INVALID = 1,   /* bad argument / call order */
    PERCH_ERR_CUDA = 2,      /* CUDA runtime error (text in perch_last_error) */
    PERCH_ERR_NO_DEVICE = 3, /* no usable sm_100 device: the library has no CPU fallback */
    PERCH_ERR_CAPACITY = 4   /* a documented size limit was exceeded */
} perch_status;

typedef struct perch_ctx perch_ctx;

/* Layout-compatible with cuda_renderer::Model::Triangle (model.h:39-43,52-73): 3 x float3 + 3 x size_t colour. */
typedef struct perch_triangle {
    float v0[3], v1[3], v2[3];
    uint64_t color[3];
} perch_triangle; /* 64 bytes */

/* Replaces cuda_renderer::gpu_stats (model.h:24-27). */
typedef struct perch_gpu_stats {
    float icp_runtime;         /* seconds, max over calls like renderer.cu:1739 */
    double peak_memory_usage;  /* MiB in use on the device, high-water (utils.cuh:6-24) */
} perch_gpu_stats;

/* The scalar arguments of render_cuda_multi_unified (renderer.h:221-268 / renderer.cu:1431-1474). */
typedef struct perch_params {
    int stride;                      /* @stride: pixel subsampling of the rendered images for clouds/costs */
    int cost_type;                   /* 0 depth-only (3-DoF), 1 depth+CIEDE2000 colour, 2 6-DoF depth-only */
    int calculate_observed_cost;     /* @calculate_observed_cost */
    float sensor_resolution;         /* metres; compared squared against squared NN distance (renderer.cu:1877) */
    float color_distance_threshold;  /* CIEDE2000 threshold for cost_type 1 */
    float occlusion_threshold;       /* cm; 3-DoF source-occlusion rule (image_renderer.cuh:163) */
    int single_result_image;         /* 1: all poses z-tested into one image (RENDER stage only) */
    int do_icp;                      /* 1: batched GICP refinement before costing (renderer.cu:1688-1817) */
    int icp_max_iterations;          /* reference hard-codes 150 (renderer.cu:1696); <=0 selects that */
    int icp_k_correspondences;       /* reference hard-codes 10 (renderer.cu:1697); <=0 selects that */
    float icp_rotation_epsilon;      /* reference 2e-3 (renderer.cu:1698); <0 selects that */
    float icp_transformation_epsilon;/* reference 5e-4 (renderer.cu:1699); <0 selects that */
} perch_params;

/* Host output buffers of one perch_render_cost call.  Every pointer is optional (NULL = not wanted) and
 * CALLER-provided, sized by the capacities below.  (The C++ backend perception_b200/csrc/compat/cuda_renderer_compat.cpp
 * reproduces the reference's callee-malloc ownership on top of this.) */
typedef struct perch_outputs {
    /* stage RENDER / DEBUG (renderer.cu:1594-1615): P*W*H each (W*H when single_result_image) */
    int32_t* depth;
    uint8_t* red;
    uint8_t* green;
    uint8_t* blue;
    /* stage CLOUD / DEBUG (renderer.cu:1821-1847) */
    float* cloud;          /* 3 x cloud_capacity SoA (row stride = cloud_capacity floats) */
    uint8_t* cloud_color;  /* 3 x cloud_capacity SoA */
    int* cloud_pose_map;   /* cloud_capacity */
    int* dc_index;         /* P*W*H exclusive scan of the sample mask (compute_point_clouds.cuh:290-292) */
    int cloud_capacity;    /* in: points the three cloud buffers can hold */
    int cloud_point_num;   /* out: result_cloud_point_num (total points; may exceed capacity => truncated copy) */
    /* stage COST / DEBUG (renderer.cu:1908-1922): P floats each */
    float* rendered_cost;
    float* observed_cost;
    float* points_diff_cost;
    /* extra (not in the reference): the integer counters the costs are functions of, 3 x P:
     * [0..P) rendered points n, [P..2P) unexplained rendered points c, [2P..3P) explained observed points */
    int32_t* counts;
    /* do_icp: refined poses, 16 floats each, x100-scaled like the inputs (renderer.cu:1729-1730) */
    float* adjusted_poses;
} perch_outputs;

/* ---- context ------------------------------------------------------------------------------------------- */
/* One context per device (the reference uses the implicit current device, renderer.cu; knncuda.cu:302 forces 0).
 * Fails with PERCH_ERR_NO_DEVICE when no sm_100-class GPU is present: there is no CPU fallback. */
int perch_create(int device, perch_ctx** out);
void perch_destroy(perch_ctx* ctx);
const char* perch_last_error(void);
int perch_abi_version(void);
/* Replaces gpu_stats& stats (renderer.h:268). */
int perch_get_stats(perch_ctx* ctx, perch_gpu_stats* out);
/* Device milliseconds (CUDA events on the context's stream) of the last perch_render_cost call and the number
 * of kernels it launched -- for bench.py's roofline / gpu_launches fields. */
int perch_last_timing(perch_ctx* ctx, float* kernel_ms, int* kernel_launches);

/* ---- camera -------------------------------------------------------------------------------------------- */
/* Replaces cuda_renderer::compute_proj (renderer.h:87, renderer.cu:1386-1410).  K is row-major 3x3. Host-only. */
int perch_compute_proj(const float K[9], int width, int height, float near_, float far_, float out16[16]);
/* width,height,proj_mat + kCameraCX..FY, depth_factor arguments of render_cuda_multi_unified (renderer.cu:1437,1446-1450). */
int perch_set_camera(perch_ctx* ctx, int width, int height, const float proj16[16], float cx, float cy, float fx,
                     float fy, float depth_factor);

/* ---- models / scene ------------------------------------------------------------------------------------- */
/* @tris, @tris_model_count of render_cuda_multi_unified (renderer.cu:1433,1436): all models' triangles
 * concatenated, in the reference's 64-byte Triangle layout.  The three setters below may be called with every render
 * call, as the reference passes these arguments every time: perch_set_models and perch_set_observed return at once when the
 * input equals the resident one (128-bit digest of all bytes), perch_set_scene re-uploads into buffers it keeps. */
int perch_set_models(perch_ctx* ctx, const perch_triangle* tris, int64_t n_tris, const int* tris_model_count,
                     int n_models);
/* @source_depth, @source_color (3 planes), @source_mask_label (may be NULL) (renderer.cu:1438-1442). */
int perch_set_scene(perch_ctx* ctx, const int32_t* source_depth, const uint8_t* source_red,
                    const uint8_t* source_green, const uint8_t* source_blue, const uint8_t* source_mask_label);
/* @observed_depth_eigen (N x 3 AoS), @observed_color (3 x N SoA), @result_observed_cloud_label (may be NULL),
 * @observed_point_num (renderer.cu:1451-1456).  In 6-DoF mode the cloud is stably sorted by label like
 * renderer.cu:1681 (nearest-neighbour indices and their tie-breaks follow the sorted order) and the search is restricted
 * to observed points carrying the pose's label (the reference searches the label's range, renderer.cu:1668-1686). */
int perch_set_observed(perch_ctx* ctx, const float* observed_xyz_aos, const uint8_t* observed_color_soa,
                       const int* observed_label, int n_obs);

/* ---- depth image -> cloud -------------------------------------------------------------------------------- */
/* Replaces cuda_renderer::depth2cloud_global (renderer.h:131-153, renderer.cu:1936-2070).
 * Pass 1: all out_* NULL => *point_num receives the count.  Pass 2: buffers sized for that count.
 * color planes may be NULL (zeros are reported).  camera_transform16 is COLUMN-major (Eigen::Matrix4f storage).
 * bounds6 = {xmax,xmin,ymax,ymin,zmax,zmin} (compute_point_clouds.cuh:86-88). */
int perch_depth2cloud(perch_ctx* ctx, const int32_t* depth, const uint8_t* red, const uint8_t* green,
                      const uint8_t* blue, int width, int height, int num_poses, float cx, float cy, float fx,
                      float fy, float depth_factor, int stride, const uint8_t* label_mask, const double* bounds6,
                      const float* camera_transform16, int* point_num, float* out_cloud_soa, float* out_cloud_aos,
                      uint8_t* out_cloud_color, int* out_dc_index, int* out_cloud_pose_map, int* out_cloud_label);

/* Allocation-free variant of the depth image -> cloud kernels.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  perch_depth2cloud keeps its per-row counters in the context's cudaMalloc'd scratch; here
 * they live in a file-scope device array inside the library (PERCH_STATIC_CLOUD_MAX_ROWS + 64 ints = 4 MiB, held by
 * every GPU that loads the library until exit), so the call performs no cudaMalloc / cudaFree / allocating library call.
 * Inputs and outputs are CALLER-provided DEVICE pointers on the context's device: d_depth (num_images x height x width
 * int32), d_label_mask (same shape uint8, may be NULL); outputs (each may be NULL) d_out_cloud_aos (capacity x 3 floats),
 * d_out_cloud_soa (3 x capacity, row stride = capacity), d_out_cloud_pose_map, d_out_cloud_label (capacity ints; label
 * only with a mask).  *point_num (host) receives the number of points; values, order and the label rule are those of
 * perch_depth2cloud without bounds (compute_point_clouds.cuh:37-184).
 * Limits: num_images * ceil(height / stride) <= PERCH_STATIC_CLOUD_MAX_ROWS, else PERCH_ERR_CAPACITY (nothing launched);
 * more points than `capacity` => the first `capacity` are written and PERCH_ERR_CAPACITY is returned with *point_num set.
 * All calls on one device share the array and it keeps the last call's contents (nothing in it needs to start at zero:
 * every slot a call reads it has written first); calls that OVERLAP on the GPU -- two contexts / streams on the same
 * device -- race on it.  Runs on the context's stream and synchronises it. */
#define PERCH_STATIC_CLOUD_MAX_ROWS (1 << 20)
int perch_depth2cloud_static(perch_ctx* ctx, const int32_t* d_depth, const uint8_t* d_label_mask, int width, int height,
                             int num_images, float cx, float cy, float fx, float fy, float depth_factor, int stride,
                             int capacity, float* d_out_cloud_aos, float* d_out_cloud_soa, int* d_out_cloud_pose_map,
                             int* d_out_cloud_label, int* point_num);

/* ---- the hot path ----------------------------------------------------------------------------------------- */
/* Replaces cuda_renderer::render_cuda_multi_unified (renderer.h:221-268, renderer.cu:1431-1934).
 * stage: "RENDER" | "CLOUD" | "COST" | "DEBUG" with the reference's meaning (renderer.cu:1479).
 * poses16: P row-major 4x4 (Model::mat4x4 a0..d3).  pose_segmentation_label NULL => 3-DoF occlusion rule.
 * pose_observed_points_total: P floats (renderer.cu:1455); may be NULL when observed cost is off.
 * "COST" runs the fused render->cost kernel: no rendered image or cloud is materialised in HBM. */
int perch_render_cost(perch_ctx* ctx, const char* stage, const float* poses16, int n_poses,
                      const int* pose_model_map, const int* pose_segmentation_label,
                      const float* pose_observed_points_total, const perch_params* params, perch_outputs* out);

/* Device-resident variant for benchmarking / pipelining: inputs and the 3xP float + 3xP int outputs are
 * DEVICE pointers, work is enqueued on the context's stream and NOT synchronised. Stage is COST. */
int perch_render_cost_device(perch_ctx* ctx, const float* d_poses16, int n_poses, const int* d_pose_model_map,
                             const int* d_pose_segmentation_label, const float* d_pose_observed_points_total,
                             const perch_params* params, float* d_costs3p, int32_t* d_counts3p);
/* The COST stage computed WITHOUT the fused kernel, in the reference's own structure (renderer.cu:1561-1659, 1852-1907):
 * every pixel of every pose rasterised (no culling), stride-sampled into clouds, a strict-`<` linear nearest-neighbour
 * scan over the observed points of the pose's label, then compute_costs' counters.  Arguments and outputs as
 * perch_render_cost(stage "COST") (rendered_cost, observed_cost, points_diff_cost, counts); cost_type 0 and 2 only.
 * A self-check and A/B path (every pose of a full-size batch in a few hundred ms), not the product path. */
int perch_render_cost_unfused(perch_ctx* ctx, const float* poses16, int n_poses, const int* pose_model_map,
                              const int* pose_segmentation_label, const float* pose_observed_points_total,
                              const perch_params* params, perch_outputs* out);
/* Device memory helpers so hosts without a CUDA binding (ctypes) can stage persistent buffers. */
int perch_device_alloc(perch_ctx* ctx, size_t bytes, void** d_ptr);
int perch_device_free(perch_ctx* ctx, void* d_ptr);
int perch_memcpy_h2d(perch_ctx* ctx, void* d_dst, const void* h_src, size_t bytes);
int perch_memcpy_d2h(perch_ctx* ctx, void* h_dst, const void* d_src, size_t bytes);
int perch_synchronize(perch_ctx* ctx);
/* Measurement: CUDA events on the context's stream around whatever the caller enqueues in between (several
 * perch_render_cost_device launches back to back); perch_timer_elapsed waits for the stop event. */
int perch_timer_start(perch_ctx* ctx);
int perch_timer_stop(perch_ctx* ctx);
int perch_timer_elapsed(perch_ctx* ctx, float* ms);
/* Measurement hygiene: overwrite a buffer larger than the 126 MB L2 on the context's stream (between timed steps). */
int perch_l2_flush(perch_ctx* ctx);

/* ---- batched GICP refinement ------------------------------------------------------------------------------------- */
/* The do_icp block of render_cuda_multi_unified (renderer.cu:1688-1817) as a call of its own: every pose is rendered
 * on the stride lattice, its cloud is registered against the observed cloud (restricted to pose_segmentation_label in
 * 6-DoF) by Generalized-ICP, and out_poses16 receives T * pose (x100-scaled like the inputs, renderer.cu:1412-1429).
 * Optional outputs: out_transforms12 = T per pose (R row-major, then t; metres), out_iterations, out_error (sum of
 * r^T M r at the last linearisation).  params: stride, sensor_resolution (grid pitch only), occlusion_threshold, icp_*.
 * perch_render_cost with params->do_icp = 1 runs the same refinement and then clouds / costs the ADJUSTED poses.
 * fast_gicp, which the reference calls here, is not part of the reference tree and unpinned: see DESIGN.md 5. */
int perch_gicp(perch_ctx* ctx, const float* poses16, int n_poses, const int* pose_model_map,
               const int* pose_segmentation_label, const perch_params* params, float* out_poses16,
               float* out_transforms12, int* out_iterations, float* out_error);

/* ---- point-to-plane ICP against a depth image (package API of cuda_icp) -------------------------------------------- */
/* Replaces Scene_projective::init_Scene_projective_cuda (depth_scene.h:19-25, depth_scene.cpp:3-37): scene_depth is
 * width*height int32 in cm (the renderer's depth unit, "depth can be int32, if we use our cuda renderer", icp.h:55), K9
 * the row-major pinhole matrix.  Builds the per-pixel scene points (dep2pcd, common.h:50-66) and normals (get_normal,
 * scene/common.cpp:17-107) on the device and keeps them in the context; out_pcd / out_normals (width*height*3 floats,
 * may be NULL) return them to the host. */
int perch_icp_set_scene(perch_ctx* ctx, const int32_t* scene_depth, int width, int height, const float K9[9],
                        float max_dist_diff, float* out_pcd, float* out_normals);
/* Replaces cuda_icp::ICP_Point2Plane_cuda<Scene_projective> (icp.h:92-99, icp.cu:156-218) for a BATCH of clouds:
 * model_pcd holds all clouds back to back (x,y,z per point, metres), cloud i = points [cloud_offsets[i],
 * cloud_offsets[i+1]).  Per cloud: RegistrationResult (icp.h:27-38) = transformation_ (row-major 4x4 into
 * out_transforms16), fitness_, inlier_rmse_; criteria = ICPConvergenceCriteria (icp.h:40-51; reference defaults 1e-5,
 * 1e-5, 30).  out_pcd (may be NULL) receives the clouds as the reference leaves model_pcd: transformed in place.
 * Scene_nn (kd-tree association, pcd_scene.h) is perch_icp_point2plane_nn below. */
int perch_icp_point2plane(perch_ctx* ctx, const float* model_pcd, const int* cloud_offsets, int n_clouds,
                          float relative_fitness, float relative_rmse, int max_iteration, float* out_transforms16,
                          float* out_fitness, float* out_inlier_rmse, int* out_iterations, float* out_pcd);

/* Replaces cuda_icp::ICP_Point2Plane_cuda<Scene_nn> (icp.h:92-99; Scene_nn::query, pcd_scene.h:60-136): the same loop with
 * every model point associated to its NEAREST scene point (the reference walks a kd-tree built on the host,
 * pcd_scene.cpp:40-182; here an exact grid search over the scene set by perch_icp_set_scene), accepted iff the squared
 * distance is below max_dist_diff^2 (reference: 0.01 m, pcd_scene.h:48).  Arguments as perch_icp_point2plane. */
int perch_icp_point2plane_nn(perch_ctx* ctx, const float* model_pcd, const int* cloud_offsets, int n_clouds,
                             float max_dist_diff, float relative_fitness, float relative_rmse, int max_iteration,
                             float* out_transforms16, float* out_fitness, float* out_inlier_rmse, int* out_iterations,
                             float* out_pcd);

/* Device-pointer forms behind the C++ backend for the reference's unchanged <cuda_icp/icp.h>
 * (perception_b200/csrc/compat/cuda_icp_compat.cpp), where clouds and scenes live in device_vector_holder<Vec3f>:
 * perch_icp_get_scene_device  the per-pixel scene points / normals perch_icp_set_scene left on the device;
 * perch_icp_point2plane_device  ICP_Point2Plane_cuda (icp.h:92-99, icp.cu:156-218) of ONE device-resident cloud
 *     (n_points x 3 floats, transformed in place) against device-resident scene buffers (d_scene_pcd = NULL: the resident
 *     scene); nearest_neighbour = 1 selects Scene_nn's association, which needs the resident scene;
 * perch_icp_depth2cloud_device  cuda_icp::depth2cloud_cuda<uint16_t | int32_t> (icp.h:83-90, icp.cu:225-297): valid
 *     pixels (depth > 0) of every stride-th row / column, row-major, z = depth / 100, x = (column_index + tl_x - K02) /
 *     K00 * z with the LATTICE column index like the reference.  Pass d_out_xyz = NULL to count.  (The reference's
 *     mask index `x + y * width` overruns its `width * height / stride^2` mask for stride > 1; here the rows are
 *     compacted in the order that index implies.) */
int perch_icp_get_scene_device(perch_ctx* ctx, const float** d_pcd, const float** d_normals, int* width, int* height);
int perch_icp_point2plane_device(perch_ctx* ctx, float* d_model_pcd, int n_points, const float* d_scene_pcd,
                                 const float* d_scene_normals, int width, int height, const float K9[9],
                                 float max_dist_diff, int nearest_neighbour, float relative_fitness, float relative_rmse,
                                 int max_iteration, float out_transform16[16], float* out_fitness, float* out_inlier_rmse,
                                 int* out_iterations);
int perch_icp_depth2cloud_device(perch_ctx* ctx, const void* d_depth, int depth_bits, int width, int height,
                                 const float K9[9], int stride, int tl_x, int tl_y, float* d_out_xyz, int capacity,
                                 int* n_points);

/* ---- k nearest neighbours (package API of cuda_renderer/knncuda.h) ------------------------------------------ */
/* Replaces knn_cuda_global / knn_cuda_texture / knn_cublas (knncuda.h:33-100, knncuda.cu:276-421, 424-616, 618-838):
 * ref is dim x ref_nb and query dim x query_nb, one ROW per dimension (element (d, i) at d*nb + i); outputs are
 * k x query_nb: knn_dist[j*query_nb + i] = Euclidean distance (square root applied) of query i to its j-th nearest
 * reference point, knn_index the 0-based index of that point; ascending distance, equal distances by ascending
 * index.  All four pointers are HOST memory like in the reference.  Distances and indices are bit-identical to
 * knn_cuda_global's; knn_cublas's differ from both in the last bits (||r||^2+||q||^2-2rq formulation) and are
 * covered by the package's own acceptance rule (knn_test, knncuda.cu:1001-1090).  k must not exceed ref_nb. */
int perch_knn(perch_ctx* ctx, const float* ref, int ref_nb, const float* query, int query_nb, int dim, int k,
              float* knn_dist, int* knn_index);

/* Per-model argmin the caller performs after scoring (search_env.cpp:2542-2583), generalised to top-k:
 * cost = (int)(rendered + observed), -1 if rendered < 0 (search_env.cpp:2022-2027); candidates with
 * |rendered - observed| >= 30 or cost < 0 are skipped; ties resolve to the lowest pose id (strict `<`).
 * Host-side helper; used per GPU before the NCCL gather of (cost, pose-id) records. */
typedef struct perch_hit {
    int32_t cost;
    int32_t pose_id;
    float rendered_cost;
    float observed_cost;
} perch_hit;
int perch_topk(const float* rendered_cost, const float* observed_cost, const int* pose_model_map, int n_poses,
               int pose_id_offset, int n_models, int k, perch_hit* out_hits /* n_models*k, cost=INT32_MAX pads */);

/* ==== a scene in one call, pose list sharded over GPUs =================================================== */
/* The reference scores a scene's candidates in gpu_batch_size batches on one GPU (search_env.cpp:2504-2525) and takes the
 * per-model argmin on the host afterwards (2542-2583).  perch_score_scene is that loop plus the argmin as ONE call: host
 * pose batch in (pinned memory makes the copies asynchronous), fused render+cost kernel, per-model top-k ON THE DEVICE
 * (same rule as perch_topk), and -- when the context belongs to a communicator -- an ncclAllGather of the k x 16 B per
 * model over NVLink, all on the context's stream with one host synchronisation; the gathered records are merged by
 * (cost, pose id) on the host, so out_hits (n_models x k, cost = INT32_MAX pads) is the SCENE's result on every rank.
 * pose_id_offset = global id of this call's first pose (the rank's block offset).  out (may be NULL) receives this
 * rank's rendered_cost / observed_cost / points_diff_cost / counts as in perch_render_cost; the other fields are ignored.
 * Limits: n_models <= 255, n_poses <= 2^24 per call, do_icp = 0.  Every rank of a communicator must make the call (an
 * empty block is fine: n_poses = 0) with the same n_models and k; like any collective, a rank whose call fails validation
 * before the gather leaves the others waiting in it -- validate the scene-level arguments (n_models, k, params) once,
 * before the ranks split the pose list. */
int perch_score_scene(perch_ctx* ctx, const float* poses16, int n_poses, const int* pose_model_map,
                      const int* pose_segmentation_label, const float* pose_observed_points_total,
                      const perch_params* params, int pose_id_offset, int n_models, int k, perch_outputs* out,
                      perch_hit* out_hits);
/* One process per GPU (torchrun, MPI -- the reference distributes work over MPI ranks, search_env.cpp:920-1023): rank 0
 * obtains an id, the launcher's own channel broadcasts the PERCH_COMM_ID_BYTES bytes, every rank attaches its context.
 * n_ranks = 1 detaches.  Replaces nothing in the reference (it has no multi-GPU path). */
#define PERCH_COMM_ID_BYTES 128
int perch_comm_unique_id(void* out_id, size_t bytes);
int perch_comm_init_rank(perch_ctx* ctx, int n_ranks, int rank, const void* id_bytes);
int perch_comm_destroy(perch_ctx* ctx);
/* One process, n devices: one context, one host thread and one stream per device, ncclCommInitAll.  The setters below
 * replicate models / scene / observed cloud on every device in parallel; perch_multi_context(m, i) exposes device i's
 * context for anything else (same call on each).  perch_multi_score cuts the pose list into contiguous blocks
 * [r*P/n, (r+1)*P/n) -- pose ids stay monotone, so the first of equal costs still wins -- scores them concurrently and
 * returns the merged per-model top-k; `out` arrays are sized for all n_poses (counts: 3 x n_poses). */
typedef struct perch_multi perch_multi;
int perch_create_multi(const int* devices, int n_devices, perch_multi** out);
void perch_destroy_multi(perch_multi* m);
int perch_multi_size(perch_multi* m);
perch_ctx* perch_multi_context(perch_multi* m, int i);
int perch_multi_set_camera(perch_multi* m, int width, int height, const float proj16[16], float cx, float cy, float fx,
                           float fy, float depth_factor);
int perch_multi_set_models(perch_multi* m, const perch_triangle* tris, int64_t n_tris, const int* tris_model_count,
                           int n_models);
int perch_multi_set_scene(perch_multi* m, const int32_t* source_depth, const uint8_t* source_red,
                          const uint8_t* source_green, const uint8_t* source_blue, const uint8_t* source_mask_label);
int perch_multi_set_observed(perch_multi* m, const float* observed_xyz_aos, const uint8_t* observed_color_soa,
                             const int* observed_label, int n_obs);
int perch_multi_score(perch_multi* m, const float* poses16, int n_poses, const int* pose_model_map,
                      const int* pose_segmentation_label, const float* pose_observed_points_total,
                      const perch_params* params, int n_models, int k, perch_outputs* out, perch_hit* out_hits);
/* Max over devices of the fused kernel's device time in the last perch_multi_score. */
int perch_multi_last_timing(perch_multi* m, float* max_kernel_ms);

/* ==== caller glue: candidate ObjectStates instead of matrices ============================== */
/* The steps sbpl_perception::EnvObjectRecognition performs on the host, serially, on either side of
 * render_cuda_multi_unified -- GenerateSuccessorStates + IsValidPose before it (search_env.cpp:7056-7490, 309-528),
 * the per-pose matrix assembly and pose_observed_points_total inside GetStateImagesUnifiedGPU (1505-1622) -- run here
 * on the device, so that a scene goes from candidate generation to costs without the pose list visiting the host. */

/* ObjectState (object_state.h) with its ContPose flattened.  The fields hold what ContPose's constructors leave in the
 * object (object_state.cpp:17-41): angles already normalised to [0, 2pi) for the Euler constructor; qx..qw all zero
 * means "use the Euler angles" (object_state.cpp:88-93). */
typedef struct perch_state {
    double x, y, z;
    double roll, pitch, yaw;
    double qx, qy, qz, qw;
    int32_t model_id;              /* ObjectState::id() */
    int32_t segmentation_label_id; /* ObjectState::segmentation_label_id(): 1-based label in the mask; -1 = none */
} perch_state;                     /* 88 bytes */

/* What the glue reads from ObjectModel (object_model.h) and ModelMetaData. */
typedef struct perch_model_meta {
    float preprocessing_transform[16]; /* ObjectModel::preprocessing_transform().matrix(), ROW-major */
    double circumscribed_radius;       /* GetCircumscribedRadius()   (object_model.cpp:460-462) */
    double circumscribed_radius_3d;    /* GetCircumscribedRadius3D() (464-466) */
    double inscribed_radius;           /* GetInscribedRadius()       (456-458) */
    double inflation_factor;           /* GetInflationFactor()       (383) */
    double search_resolution;          /* ModelMetaData::search_resolution */
    int32_t symmetric;                 /* ObjectModel::symmetric() */
    int32_t symmetry_mode;             /* ModelMetaData::symmetry_mode: 0 none, 1 semi (180 deg), 2 full */
    int32_t footprint_offset;          /* convex_hull_footprint_ (object_model.cpp:361-372): first point ... */
    int32_t footprint_count;           /* ... and number of points in perch_set_env's footprint array */
} perch_model_meta;

/* The EnvParams / PerchParams fields the glue reads (search_env.cpp:140-216). */
typedef struct perch_env {
    double cam_to_world[16];           /* cam_to_world_.matrix(), ROW-major (the body-frame camera pose) */
    double x_min, x_max, y_min, y_max; /* table bounds */
    double table_height;
    double res;                        /* env_params_.res: translation step of the 3-DoF grid */
    double theta_res;                  /* env_params_.theta_res: yaw step */
    double footprint_tolerance;        /* perch_params_.footprint_tolerance */
    int32_t use_external_pose_list;    /* 1 = 6-DoF mode */
    int32_t use_cylinder_observed;     /* perch_params_.use_cylinder_observed */
    int32_t min_neighbor_points_for_valid_pose;
    int32_t use_adaptive_resolution;
    int32_t use_model_specific_search_resolution;
    int32_t min_points_for_constraint_cloud;
} perch_env;

/* Environment and per-model metadata; footprint_xyz holds every model's convex-hull footprint points (x,y,z floats,
 * implicitly closed polygons) back to back.  n_models must match perch_set_models when both are set. */
int perch_set_env(perch_ctx* ctx, const perch_env* env, const perch_model_meta* metas, int n_models,
                  const float* footprint_xyz, int n_footprint_points);
/* The clouds the radius searches run in (search_env.cpp:5624-5662): xyz_aos holds n_points points; ranges (may be NULL =
 * one range holding everything) is 2*n_ranges ints, [begin,end) per range: range 0 is the cloud searched when a state
 * carries no label (projected_knn_ in 3-DoF, where the caller has already set z = table_height, or knn in 6-DoF), range
 * 1+l the segmented cloud of label l (segmented_object_knn[l]).  constraint_xyz (may be NULL / 0) = projected
 * constraint cloud (search_env.cpp:5654-5658).  A uniform grid over (x, y) is built on the device. */
int perch_set_search_cloud(perch_ctx* ctx, const float* xyz_aos, int n_points, const int* ranges, int n_ranges,
                           const float* constraint_xyz, int n_constraint);
/* The loop of GetStateImagesUnifiedGPU (search_env.cpp:1535-1622): per state the camera-frame pose matrix
 * (cam_to_world * cam_to_body)^-1 * ContPose::GetTransform() * preprocessing_transform in double, x100 on the top three
 * rows, cast to float (model.h:89-107); pose_model_map; in 6-DoF mode the segmentation label (id - 1) and the label's
 * observed point count; in 3-DoF mode the number of projected-cloud points inside the model's inflated circumscribed
 * circle (use_cylinder_observed) or the observed point count.  Outputs are host pointers, each optional. */
int perch_assemble_states(perch_ctx* ctx, const perch_state* states, int n_states, float* out_poses16,
                          int* out_pose_model_map, int* out_pose_segmentation_label,
                          float* out_pose_observed_points_total);
/* perch_assemble_states followed by perch_render_cost on the assembled, device-resident batch: what
 * GetStateImagesUnifiedGPU does in one call.  states == NULL scores the list left on the device by the last
 * perch_generate_states (n_states is then ignored). */
int perch_score_states(perch_ctx* ctx, const char* stage, const perch_state* states, int n_states,
                       const perch_params* params, perch_outputs* out);
/* EnvObjectRecognition::IsValidPose (search_env.cpp:309-528) for a list of candidates against the objects already
 * placed: enough observed points within the search radius, no overlap of inscribed circles with placed objects,
 * footprint inside the table bounds (3-DoF), enough constraint-cloud points inside the footprint.  out_valid: one byte
 * per state. */
int perch_validate_states(perch_ctx* ctx, const perch_state* states, int n_states, const perch_state* placed,
                          int n_placed, int after_refinement, uint8_t* out_valid);
/* The 3-DoF branch of GenerateSuccessorStates (search_env.cpp:7268-7472) for every model not yet in `placed`: the
 * x / y / yaw grid (steps accumulated in double like the reference's loops), IsValidPose, the symmetry rules
 * (symmetric / symmetry_mode 2: first valid yaw only; mode 1: stop above pi + theta_res), output in the reference's
 * order (model, x, y, yaw).  The list stays on the device for perch_score_states(states = NULL); out_states (may be
 * NULL) receives at most `capacity` of them, *n_out the full count. */
int perch_generate_states(perch_ctx* ctx, const perch_state* placed, int n_placed, perch_state* out_states,
                          int capacity, int* n_out);

/* The 6-DoF branch of GenerateSuccessorStates reads its candidates from <rendered_root_dir>/<model>/poses.txt
 * (search_env.cpp:7099-7146): one pose per line, `x y z qx qy qz qw`, single spaces, parsed with std::stod into
 * ContPose(x, y, z, qx, qy, qz, qw) and then ObjectState(model, symmetric, pose, required_object_id + 1).  This parses the
 * text of such a file into perch_state records (Euler fields 0: ContPose::GetTransform uses the quaternion,
 * object_state.cpp:88-93) ready for perch_validate_states / perch_score_states.  *n_out = poses in the text; at most
 * `capacity` are written (pass 0 / NULL to count).  A malformed line is PERCH_ERR_INVALID with its line number (the
 * reference dies with an uncaught std::invalid_argument there).  Host-only. */
int perch_parse_pose_list(const char* text, size_t length, int model_id, int segmentation_label_id, perch_state* out_states,
                          int capacity, int* n_out);

/* ==== observed-scene preparation ================================================================ */
/* The GPU branch of EnvObjectRecognition::SetInput (search_env.cpp:5880-6069) as one call: the decoded input images go
 * in, the source depth image and the observed cloud stay resident in the context (what perch_set_scene +
 * perch_depth2cloud + perch_set_observed would do with a host round trip of every buffer in between). */
typedef struct perch_input {
    const void* depth;               /* width*height: uint8 when depth_bits == 8 (3-DoF, NDDS images, search_env.cpp:5919)
                                        or uint16 when 16 (6-DoF, 5890) */
    int depth_bits;
    const uint8_t* red;              /* input_color_image_vec[0..2] (5959-5972); each may be NULL */
    const uint8_t* green;
    const uint8_t* blue;
    const uint8_t* predicted_mask;   /* 6-DoF: label mask image (5894-5898), 0 = no object; NULL in 3-DoF */
    int median_blur_ksize;           /* perch_params_.depth_median_blur (5922): odd, <= 31, applied to 8-bit depth;
                                        <= 1 = none */
    float depth_factor;              /* input.depth_factor: depth units per metre of the cloud (6011) */
    int stride;                      /* gpu_stride */
    const double* bounds6;           /* 3-DoF: {x_max, x_min, y_max, y_min, table_height + 0.5, table_height}
                                        (5931-5938); NULL = no bounds filter */
    const float* camera_transform16; /* with bounds6: (cam_to_world * cam_to_body).cast<float>(), COLUMN-major (5947-5950) */
    float source_division_factor;    /* input_depth_factor / gpu_depth_factor: the source depth the renderer tests
                                        occlusion against is depth / factor, truncated (2487-2498); <= 0 = depth as is */
} perch_input;
/* Requires perch_set_camera.  *observed_point_num receives the size of the observed cloud. */
int perch_set_input(perch_ctx* ctx, const perch_input* in, int* observed_point_num);
/* The observed cloud of the last perch_set_input in the order depth2cloud_global returns it, for the caller's host-side
 * copies (search_env.cpp:6031-6060): xyz N x 3, colours 3 x N (row stride = *point_num), labels N (NULL allowed;
 * -1 everywhere in 3-DoF).  capacity = points the buffers hold. */
int perch_get_observed(perch_ctx* ctx, float* xyz_aos, uint8_t* rgb_soa, int* label, int capacity, int* point_num);
/* The source depth image the occlusion rule currently tests against: as resident after perch_set_input /
 * perch_set_scene, or as composed by perch_compose_source (width*height int32). */
int perch_get_source_depth(perch_ctx* ctx, int32_t* out_depth);

/* Diagnostic (host only, no device needed): runs the mesh preprocessing of perch_set_models (edge-sharing triangles
 * paired into quads, quads into clusters with bounding spheres) and checks it.  out = {clusters, triangles stored in
 * pairs, triangles stored alone, coordinates that differ from the input (must be 0), triangles not stored exactly once
 * (must be 0), vertices outside their cluster's sphere or triangles filed under the wrong model (must be 0)}. */
int perch_cluster_stats(const perch_triangle* tris, int64_t n_tris, const int* tris_model_count, int n_models,
                        int64_t out[6]);

/* The label image the occlusion rule currently tests against (width*height uint8); *has_label = 0 and the buffer
 * untouched when there is none. */
int perch_get_source_label(perch_ctx* ctx, uint8_t* out_label, int* has_label);

/* ==== composition mode =========================================================================== */
/* Multi-object states: the objects already placed are rendered into the source image the occlusion rule of
 * perch_render_cost tests against (image_renderer.cuh:163-186), so the next object is scored with the placed ones
 * occluding it.  The reference does this through the host: the state's objects are rendered with
 * single_result_image = 1 against an empty source (PrintStateGPU, search_env.cpp:1718-1781; the "GPU + Tree" flow of
 * the deprecated ComputeCostsInParallelGPU, 2054-2460) and the image goes back in as the next call's source_depth.
 * Here the composed image never leaves the device.
 *   PERCH_COMPOSE_REPLACE  source := the placed objects alone, 0 elsewhere (the reference's tree flow)
 *   PERCH_COMPOSE_MERGE    source := per pixel the nearer of the OBSERVED source image and the placed objects
 *   PERCH_COMPOSE_NONE     back to the observed image (perch_set_scene / perch_set_input do the same)
 * pose_label (may be NULL): 0-based segmentation label per placed pose; a pixel owned by pose i then carries source
 * label pose_label[i] + 1, otherwise 0 (a label no candidate shares: the pixel occludes every candidate).  The base of
 * a MERGE is always the observed image, so composing again replaces the previous composition. */
enum { PERCH_COMPOSE_NONE = 0, PERCH_COMPOSE_REPLACE = 1, PERCH_COMPOSE_MERGE = 2 };
int perch_compose_source(perch_ctx* ctx, const float* poses16, int n_poses, const int* pose_model_map,
                         const int* pose_label, int mode);
/* Per-pose outputs of the reference's USE_TREE / USE_CLUTTER branches (image_renderer.cuh:187-206, 467-482; compiled
 * out in the reference as shipped, model.h:17-18), on the full-resolution render of every pose against the current
 * source image:
 *   pose_occluded[p]        1 iff the render lies in front of a source pixel by more than the rule's gate
 *   pose_occluded_other[p]  1 iff a source pixel hides (clears) a rendered pixel
 *   clutter_cost[p]         100 * cleared pixels / covered pixels, NaN if nothing is covered (the reference's extra
 *                           condition `source <= new_depth - 5` reads the pixel after it was set to INT_MAX and
 *                           therefore always holds, image_renderer.cuh:183-195; pinned by tests/golden/tree_single_tri)
 *   total_points[p]         covered pixels.  DELIBERATE DEVIATION: the reference's USE_CLUTTER branch adds 1 per inside
 *                           FRAGMENT (image_renderer.cuh:190-205), i.e. at least 2 per pixel of a closed mesh, so its
 *                           clutter ratio is systematically smaller on multi-triangle meshes; the per-pixel count is
 *                           the one that does not depend on tessellation or fragment order.
 * evaluated per covered pixel on its minimum depth.  The reference evaluates them per fragment on the running minimum,
 * which depends on the order fragments arrive in (and races); both agree when a pixel receives one fragment.
 * pose_label NULL = 3-DoF gate (|d - s| > occlusion_threshold), else the 6-DoF gate.  Any output may be NULL. */
int perch_tree_flags(perch_ctx* ctx, const float* poses16, int n_poses, const int* pose_model_map, const int* pose_label,
                     float occlusion_threshold, int* pose_occluded, int* pose_occluded_other, float* clutter_cost,
                     float* total_points);

#ifdef __cplusplus
}
#endif
#endif /* PERCH_B200_H */
