// C-ABI of the observed-scene preparation (include/perch_b200.h): SetInput's GPU branch, search_env.cpp:5880-6069.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <vector>

#include "capi_common.hpp"
#include "kernels.cuh"

using perch_detail::DeviceGuard;
using perch_detail::fail;

namespace {
size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }
}  // namespace

extern "C" {

int perch_set_input(perch_ctx* c, const perch_input* in, int* observed_point_num) {
    if (!c || !in || !in->depth || !observed_point_num) return fail(PERCH_ERR_INVALID, "null argument");
    if (!c->have_cam) return fail(PERCH_ERR_INVALID, "perch_set_camera must be called before perch_set_input");
    if (in->depth_bits != 8 && in->depth_bits != 16) return fail(PERCH_ERR_INVALID, "depth_bits must be 8 or 16");
    const int W = c->cam.width, H = c->cam.height, k = in->median_blur_ksize;
    if (in->stride <= 0 || W % in->stride != 0)
        return fail(PERCH_ERR_INVALID, "stride %d must divide the image width %d (compute_point_clouds.cuh:271)", in->stride, W);
    if (k > 1 && (k % 2 == 0 || k > perch::kMedianMaxKsize))
        return fail(PERCH_ERR_INVALID, "median_blur_ksize %d must be odd and <= %d", k, perch::kMedianMaxKsize);
    if (k > 1 && in->depth_bits != 8) return fail(PERCH_ERR_INVALID, "the median blur applies to 8-bit depth images only");
    if ((in->bounds6 != nullptr) != (in->camera_transform16 != nullptr))
        return fail(PERCH_ERR_INVALID, "bounds6 and camera_transform16 go together");
    DeviceGuard g(c->device);
    const size_t npx = (size_t)W * H;
    const int s = in->stride, Lh = (H + s - 1) / s, Lw = W / s;
    const size_t cap = (size_t)Lw * Lh;  // every lattice pixel may be valid
    const size_t need = align256(npx * 2) * 2 + align256(npx * 4) * 2 + 4 * align256(npx) + align256((Lh + 1) * 4) +
                        align256(cap * 12) + align256(cap * 3) + align256(cap * 4) + 4096;
    // the context scratch is free between calls
    if (need > c->scratch_bytes) {
        if (c->d_scratch) cudaFree(c->d_scratch);
        c->d_scratch = nullptr;
        c->scratch_bytes = 0;
        PERCH_CU(cudaMalloc(&c->d_scratch, need));
        c->scratch_bytes = need;
    }
    char* p = static_cast<char*>(c->d_scratch);
    auto take = [&](size_t bytes) { char* q = p; p += align256(bytes); return q; };
    void* d_raw = take(npx * 2);
    uint8_t* d_blur = (uint8_t*)take(npx * 2);
    int32_t* d_depth = (int32_t*)take(npx * 4);
    int32_t* d_source = (int32_t*)take(npx * 4);
    uint8_t* d_planes[3] = {(uint8_t*)take(npx), (uint8_t*)take(npx), (uint8_t*)take(npx)};
    uint8_t* d_mask = (uint8_t*)take(npx);
    int* d_rows = (int*)take((Lh + 1) * 4);
    float* d_aos = (float*)take(cap * 12);
    uint8_t* d_rgb = (uint8_t*)take(cap * 3);
    int* d_lab = (int*)take(cap * 4);

    PERCH_CU(cudaMemcpyAsync(d_raw, in->depth, npx * (in->depth_bits / 8), cudaMemcpyHostToDevice, c->stream));
    const uint8_t* planes[3] = {in->red, in->green, in->blue};
    for (int ch = 0; ch < 3; ++ch) {
        if (planes[ch]) PERCH_CU(cudaMemcpyAsync(d_planes[ch], planes[ch], npx, cudaMemcpyHostToDevice, c->stream));
        else PERCH_CU(cudaMemsetAsync(d_planes[ch], 0, npx, c->stream));
    }
    if (in->predicted_mask) PERCH_CU(cudaMemcpyAsync(d_mask, in->predicted_mask, npx, cudaMemcpyHostToDevice, c->stream));
    // 1. median blur (3-DoF, search_env.cpp:5922) and widening to int32 (5911-5914, 5926-5929)
    const void* d_img = d_raw;
    if (k > 1) {
        PERCH_CU(perch::launch_median_blur_u8((const uint8_t*)d_raw, d_blur, W, H, k, c->stream));
        d_img = d_blur;
    }
    PERCH_CU(perch::launch_depth_to_i32(d_img, in->depth_bits, d_depth, (int)npx, c->stream));
    // 2. observed cloud (depth2cloud_global, 5993-6017): bounds filter in 3-DoF, label mask in 6-DoF
    perch::CloudArgs a;
    a.depth = d_depth;
    a.red = d_planes[0]; a.green = d_planes[1]; a.blue = d_planes[2];
    a.label_mask = in->predicted_mask ? d_mask : nullptr;
    a.img_w = W; a.img_h = H; a.full_h = H; a.num_images = 1; a.stride = s; a.coord_scale = 1;
    a.cx = c->cam.cx; a.cy = c->cam.cy; a.fx = c->cam.fx; a.fy = c->cam.fy; a.depth_factor = in->depth_factor;
    a.use_bounds = in->bounds6 != nullptr;
    if (a.use_bounds) {
        for (int i = 0; i < 6; ++i) a.bounds[i] = (float)in->bounds6[i];
        memcpy(a.cam_tf, in->camera_transform16, 64);
    }
    a.row_counts = d_rows;
    a.capacity = (int)cap;
    a.cloud_aos = d_aos; a.cloud_rgb = d_rgb;
    a.cloud_label = in->predicted_mask ? d_lab : nullptr;
    PERCH_CU(cudaMemsetAsync(d_aos, 0, cap * 12, c->stream));   // slots failing the second bounds test stay zero
    PERCH_CU(cudaMemsetAsync(d_rgb, 0, cap * 3, c->stream));
    PERCH_CU(cudaMemsetAsync(d_lab, 0, cap * 4, c->stream));
    PERCH_CU(perch::launch_cloud_count(a, c->stream));
    PERCH_CU(perch::launch_cloud_scan(a, c->stream));
    PERCH_CU(perch::launch_cloud_write(a, c->stream));
    // 3. source depth for the occlusion rule (2487-2498) straight into the resident scene
    const int32_t* d_src = d_depth;
    if (in->source_division_factor > 0.f) {
        PERCH_CU(perch::launch_divide_depth(d_depth, in->source_division_factor, d_source, (int)npx, c->stream));
        d_src = d_source;
    }
    if (c->src_px != npx) {  // buffers of another image size (perch_set_camera frees them; kept as a second line of defence)
        cudaFree(c->d_src_depth); cudaFree(c->d_src_label); cudaFree(c->d_src_rgb);
        c->d_src_depth = nullptr; c->d_src_label = nullptr; c->d_src_rgb = nullptr;
        c->src_px = 0;
        PERCH_CU(cudaMalloc(&c->d_src_depth, npx * 4));
        PERCH_CU(cudaMalloc(&c->d_src_rgb, 3 * npx));
        c->src_px = npx;
    }
    if (in->predicted_mask && !c->d_src_label) PERCH_CU(cudaMalloc(&c->d_src_label, npx));
    if (!in->predicted_mask && c->d_src_label) { cudaFree(c->d_src_label); c->d_src_label = nullptr; }  // non-null = labelled scene (capi_compose.cu)
    PERCH_CU(cudaMemcpyAsync(c->d_src_depth, d_src, npx * 4, cudaMemcpyDeviceToDevice, c->stream));
    for (int ch = 0; ch < 3; ++ch)
        PERCH_CU(cudaMemcpyAsync(c->d_src_rgb + ch * npx, d_planes[ch], npx, cudaMemcpyDeviceToDevice, c->stream));
    if (in->predicted_mask) PERCH_CU(cudaMemcpyAsync(c->d_src_label, d_mask, npx, cudaMemcpyDeviceToDevice, c->stream));
    // 4. the observed cloud stays on the device: only its size comes back here; label sort, colours, bounding box and
    // (at the first cost call) the nearest-neighbour grid are device work (perch_detail::install_observed_device,
    // obs_grid.cu).  A copy in depth2cloud_global's order is kept on the device for perch_get_observed.
    int total = 0;
    PERCH_CU(cudaMemcpyAsync(&total, d_rows + Lh, 4, cudaMemcpyDeviceToHost, c->stream));
    PERCH_CU(cudaStreamSynchronize(c->stream));
    c->in_has_label = in->predicted_mask != nullptr;
    for (void* q : {(void*)c->d_in_xyz, (void*)c->d_in_rgb, (void*)c->d_in_label}) cudaFree(q);
    c->d_in_xyz = nullptr; c->d_in_rgb = nullptr; c->d_in_label = nullptr;
    const size_t N = (size_t)std::max(total, 1);
    PERCH_CU(cudaMalloc(&c->d_in_xyz, N * 12));
    PERCH_CU(cudaMalloc(&c->d_in_rgb, N * 3));
    PERCH_CU(cudaMalloc(&c->d_in_label, N * 4));
    if (total > 0) {
        PERCH_CU(cudaMemcpyAsync(c->d_in_xyz, d_aos, (size_t)total * 12, cudaMemcpyDeviceToDevice, c->stream));
        for (int ch = 0; ch < 3; ++ch)
            PERCH_CU(cudaMemcpyAsync(c->d_in_rgb + (size_t)ch * total, d_rgb + (size_t)ch * cap, total, cudaMemcpyDeviceToDevice, c->stream));
        if (c->in_has_label) PERCH_CU(cudaMemcpyAsync(c->d_in_label, d_lab, (size_t)total * 4, cudaMemcpyDeviceToDevice, c->stream));
    }
    c->in_n = total;
    c->in_on_device = true;
    c->h_in_xyz.clear(); c->h_in_rgb.clear(); c->h_in_label.clear();
    c->scene.depth = c->d_src_depth;
    c->composed = false;
    c->scene.label = in->predicted_mask ? c->d_src_label : nullptr;
    c->scene.r = c->d_src_rgb;
    c->scene.g = c->d_src_rgb + npx;
    c->scene.b = c->d_src_rgb + 2 * npx;
    c->have_scene = true;
    *observed_point_num = total;
    int rc = perch_detail::install_observed_device(c, c->d_in_xyz, c->d_in_rgb, (size_t)total, c->in_has_label ? c->d_in_label : nullptr, total);
    c->in_on_device = rc == PERCH_OK;
    return rc;
}

int perch_get_observed(perch_ctx* c, float* xyz_aos, uint8_t* rgb_soa, int* label, int capacity, int* point_num) {
    if (!c || !point_num) return fail(PERCH_ERR_INVALID, "null argument");
    if (!c->in_on_device) return fail(PERCH_ERR_INVALID, "perch_set_input has not been called (or perch_set_observed replaced its cloud)");
    const int n = c->in_n;
    *point_num = n;
    if (!xyz_aos && !rgb_soa && !label) return PERCH_OK;
    if (capacity < n) return fail(PERCH_ERR_CAPACITY, "the observed cloud has %d points, the buffers hold %d", n, capacity);
    if (n == 0) return PERCH_OK;
    DeviceGuard g(c->device);
    if (xyz_aos) PERCH_CU(cudaMemcpyAsync(xyz_aos, c->d_in_xyz, (size_t)n * 12, cudaMemcpyDeviceToHost, c->stream));
    if (rgb_soa) PERCH_CU(cudaMemcpyAsync(rgb_soa, c->d_in_rgb, (size_t)n * 3, cudaMemcpyDeviceToHost, c->stream));
    if (label) {
        if (c->in_has_label) PERCH_CU(cudaMemcpyAsync(label, c->d_in_label, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
        else std::fill(label, label + n, -1);
    }
    PERCH_CU(cudaStreamSynchronize(c->stream));
    return PERCH_OK;
}

int perch_get_source_depth(perch_ctx* c, int32_t* out_depth) {
    if (!c || !out_depth) return fail(PERCH_ERR_INVALID, "null argument");
    if (!c->have_scene) return fail(PERCH_ERR_INVALID, "no scene has been set");
    DeviceGuard g(c->device);
    PERCH_CU(cudaMemcpyAsync(out_depth, c->scene.depth, (size_t)c->cam.width * c->cam.height * 4, cudaMemcpyDeviceToHost, c->stream));
    PERCH_CU(cudaStreamSynchronize(c->stream));
    return PERCH_OK;
}

}  // extern "C"
