// C-ABI of the composition mode (include/perch_b200.h): perch_compose_source, perch_tree_flags.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdlib>

#include "capi_common.hpp"
#include "kernels.cuh"

using perch_detail::DeviceGuard;
using perch_detail::fail;

namespace {

size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }

int ensure_scratch(perch_ctx* c, size_t need) {
    if (need <= c->scratch_bytes) return PERCH_OK;
    if (c->d_scratch) cudaFree(c->d_scratch);
    c->d_scratch = nullptr;
    c->scratch_bytes = 0;
    PERCH_CU(cudaMalloc(&c->d_scratch, need));
    c->scratch_bytes = need;
    return PERCH_OK;
}

int image_tile_cap(int min_cap) {
    const char* e = getenv("PERCH_IMAGE_TILE_CAP");
    return std::max(e ? atoi(e) : 6144, min_cap);
}

int check_err_flag(perch_ctx* c) {
    int flag = 0;
    PERCH_CU(cudaMemcpyAsync(&flag, c->d_err_flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    PERCH_CU(cudaStreamSynchronize(c->stream));
    if (flag) {
        cudaMemsetAsync(c->d_err_flag, 0, sizeof(int), c->stream);
        return fail(PERCH_ERR_CUDA, "internal error: a fragment fell outside its pose tile");
    }
    return PERCH_OK;
}

// PERCH_IMAGE_LEGACY=1: full-resolution work on the per-pose kernel instead of raster_full.cu (A/B tests)
bool legacy_images() {
    const char* v = getenv("PERCH_IMAGE_LEGACY");
    return v && *v && atoi(v) != 0;
}

}  // namespace

extern "C" {

int perch_compose_source(perch_ctx* c, const float* poses16, int n_poses, const int* pose_model_map,
                         const int* pose_label, int mode) {
    if (!c) return fail(PERCH_ERR_INVALID, "null context");
    if (mode != PERCH_COMPOSE_NONE && mode != PERCH_COMPOSE_REPLACE && mode != PERCH_COMPOSE_MERGE)
        return fail(PERCH_ERR_INVALID, "compose mode %d not in {NONE, REPLACE, MERGE}", mode);
    if (!c->have_cam || !c->have_scene) return fail(PERCH_ERR_INVALID, "camera and scene must be set before composing");
    DeviceGuard g(c->device);
    const int W = c->cam.width, H = c->cam.height;
    const size_t npx = (size_t)W * H;
    if (mode == PERCH_COMPOSE_NONE) {
        c->scene.depth = c->d_src_depth;
        c->scene.label = c->d_src_label;
        c->composed = false;
        return PERCH_OK;
    }
    if (n_poses < 0 || (n_poses > 0 && (!poses16 || !pose_model_map))) return fail(PERCH_ERR_INVALID, "bad pose arguments");
    if (n_poses > 0 && !c->have_models) return fail(PERCH_ERR_INVALID, "perch_set_models must be called before composing");
    const size_t P = (size_t)n_poses;
    int rc;
    if ((rc = ensure_scratch(c, align256(npx * 8) + align256(P * 64) + 2 * align256(P * 4) + 1024))) return rc;
    char* p = static_cast<char*>(c->d_scratch);
    auto take = [&](size_t bytes) { char* q = p; p += align256(bytes); return q; };
    unsigned long long* d_keys = (unsigned long long*)take(npx * 8);
    float* d_poses = (float*)take(P * 64);
    int* d_pmm = (int*)take(P * 4);
    int* d_label = (int*)take(P * 4);
    if (!c->d_cmp_depth) PERCH_CU(cudaMalloc(&c->d_cmp_depth, npx * 4));
    if (!c->d_cmp_label) PERCH_CU(cudaMalloc(&c->d_cmp_label, npx));
    PERCH_CU(perch::launch_single_clear(d_keys, (int)npx, c->stream));
    if (n_poses > 0) {
        PERCH_CU(cudaMemcpyAsync(d_poses, poses16, P * 64, cudaMemcpyHostToDevice, c->stream));
        PERCH_CU(cudaMemcpyAsync(d_pmm, pose_model_map, P * 4, cudaMemcpyHostToDevice, c->stream));
        if (pose_label) PERCH_CU(cudaMemcpyAsync(d_label, pose_label, P * 4, cudaMemcpyHostToDevice, c->stream));
        perch::PoseArgs a;
        a.models = c->models;
        a.cam = c->cam;
        a.scene = c->scene;  // read only by the resolve step, which single_keys bypasses
        a.poses = reinterpret_cast<const perch::Mat4*>(d_poses);
        a.pose_model_map = d_pmm;
        a.n_poses = n_poses;
        a.stride = 1;
        a.single_keys = d_keys;
        a.key_pose_id = 1;
        a.err_flag = c->d_err_flag;
        a.tile_cap = image_tile_cap(W);
        if (legacy_images()) {
            PERCH_CU(perch::launch_raster_images(a, c->stream));
        } else {
            perch::FullJob j;
            j.poses = a.poses; j.pose_model_map = d_pmm; j.n_poses = n_poses;
            j.single_keys = d_keys; j.key_pose_id = 1;
            PERCH_CU(perch::full_raster_job(c, j, nullptr));
        }
    }
    perch::ComposeArgs ca;
    ca.keys = d_keys;
    ca.pose_label = (pose_label && n_poses > 0) ? d_label : nullptr;
    ca.base_depth = c->d_src_depth;  // always the OBSERVED image: composing twice does not accumulate
    ca.base_label = c->d_src_label;
    ca.out_depth = c->d_cmp_depth;
    ca.out_label = c->d_cmp_label;
    ca.n_pixels = (int)npx;
    ca.merge = mode == PERCH_COMPOSE_MERGE;
    PERCH_CU(perch::launch_compose(ca, c->stream));
    if ((rc = check_err_flag(c))) return rc;
    c->scene.depth = c->d_cmp_depth;
    // a label image exists from now on iff the observed scene has one or the placed objects carry labels
    c->scene.label = (c->d_src_label || pose_label) ? c->d_cmp_label : nullptr;
    c->composed = true;
    c->last_launches = 2 + (n_poses > 0 ? 1 : 0);
    return PERCH_OK;
}

int perch_tree_flags(perch_ctx* c, const float* poses16, int n_poses, const int* pose_model_map, const int* pose_label,
                     float occlusion_threshold, int* pose_occluded, int* pose_occluded_other, float* clutter_cost,
                     float* total_points) {
    if (!c || n_poses < 0 || (n_poses > 0 && (!poses16 || !pose_model_map)))
        return fail(PERCH_ERR_INVALID, "bad tree_flags argument");
    if (!c->have_cam || !c->have_scene || !c->have_models)
        return fail(PERCH_ERR_INVALID, "camera, models and scene must be set before perch_tree_flags");
    if (n_poses == 0) return PERCH_OK;
    DeviceGuard g(c->device);
    const size_t P = (size_t)n_poses;
    int rc;
    if ((rc = ensure_scratch(c, align256(P * 64) + 6 * align256(P * 4) + 1024))) return rc;
    char* p = static_cast<char*>(c->d_scratch);
    auto take = [&](size_t bytes) { char* q = p; p += align256(bytes); return q; };
    float* d_poses = (float*)take(P * 64);
    int* d_pmm = (int*)take(P * 4);
    int* d_label = (int*)take(P * 4);
    int* d_occ = (int*)take(P * 4);
    int* d_other = (int*)take(P * 4);
    float* d_clutter = (float*)take(P * 4);
    float* d_total = (float*)take(P * 4);
    PERCH_CU(cudaMemcpyAsync(d_poses, poses16, P * 64, cudaMemcpyHostToDevice, c->stream));
    PERCH_CU(cudaMemcpyAsync(d_pmm, pose_model_map, P * 4, cudaMemcpyHostToDevice, c->stream));
    if (pose_label) PERCH_CU(cudaMemcpyAsync(d_label, pose_label, P * 4, cudaMemcpyHostToDevice, c->stream));
    perch::PoseArgs a;
    a.models = c->models;
    a.cam = c->cam;
    a.scene = c->scene;
    a.poses = reinterpret_cast<const perch::Mat4*>(d_poses);
    a.pose_model_map = d_pmm;
    a.pose_label = pose_label ? d_label : nullptr;
    a.n_poses = n_poses;
    a.stride = 1;  // the reference evaluates these branches on every rendered pixel
    a.occlusion_threshold = occlusion_threshold;
    a.tree_occluded = d_occ;
    a.tree_occluded_other = d_other;
    a.tree_clutter = d_clutter;
    a.tree_total = d_total;
    a.err_flag = c->d_err_flag;
    a.tile_cap = image_tile_cap(c->cam.width);
    PERCH_CU(cudaEventRecord(c->ev0, c->stream));
    int launches = 0;
    if (legacy_images()) {
        PERCH_CU(perch::launch_raster_images(a, c->stream));
        launches = 1;
    } else {
        perch::FullJob j;
        j.poses = a.poses; j.pose_model_map = d_pmm; j.pose_label = a.pose_label; j.n_poses = n_poses;
        j.occlusion_threshold = occlusion_threshold;
        j.tree_occluded = d_occ; j.tree_occluded_other = d_other; j.tree_clutter = d_clutter; j.tree_total = d_total;
        PERCH_CU(perch::full_raster_job(c, j, &launches));
    }
    PERCH_CU(cudaEventRecord(c->ev1, c->stream));
    c->timing_pending = true;
    c->last_launches = launches;
    if (pose_occluded) PERCH_CU(cudaMemcpyAsync(pose_occluded, d_occ, P * 4, cudaMemcpyDeviceToHost, c->stream));
    if (pose_occluded_other) PERCH_CU(cudaMemcpyAsync(pose_occluded_other, d_other, P * 4, cudaMemcpyDeviceToHost, c->stream));
    if (clutter_cost) PERCH_CU(cudaMemcpyAsync(clutter_cost, d_clutter, P * 4, cudaMemcpyDeviceToHost, c->stream));
    if (total_points) PERCH_CU(cudaMemcpyAsync(total_points, d_total, P * 4, cudaMemcpyDeviceToHost, c->stream));
    rc = check_err_flag(c);  // synchronises the stream
    if (c->timing_pending && cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1) == cudaSuccess) c->timing_pending = false;
    return rc;
}

int perch_get_source_label(perch_ctx* c, uint8_t* out_label, int* has_label) {
    if (!c || !out_label) return fail(PERCH_ERR_INVALID, "null argument");
    if (!c->have_scene) return fail(PERCH_ERR_INVALID, "no scene has been set");
    DeviceGuard g(c->device);
    if (has_label) *has_label = c->scene.label != nullptr;
    if (c->scene.label) {
        PERCH_CU(cudaMemcpyAsync(out_label, c->scene.label, (size_t)c->cam.width * c->cam.height, cudaMemcpyDeviceToHost, c->stream));
        PERCH_CU(cudaStreamSynchronize(c->stream));
    }
    return PERCH_OK;
}

}  // extern "C"
