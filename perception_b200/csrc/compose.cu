// Composition mode: placed objects rendered into the source depth image the occlusion rule tests
// against, without leaving the device.  The reference does this through the host: the state's objects are rendered
// with single_result_image = 1 against an empty source (PrintStateGPU, search_env.cpp:1718-1781; the "GPU + Tree"
// flow of the deprecated ComputeCostsInParallelGPU, 2054-2460) and the image is handed back as the next call's
// source_depth.
#include <cuda_runtime.h>

#include <climits>

#include "kernels.cuh"

namespace perch {

namespace {

// keys: per pixel (depth ^ 0x80000000) << 32 | pose index, ~0 = empty (the single_result_image key image of the
// per-pose kernels, z-tested with atomicMin).
__global__ void compose_kernel(ComposeArgs a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n_pixels) return;
    const unsigned long long k = a.keys[i];
    int32_t d = 0;
    int label = 0;
    if (k != ~0ull) {
        d = (int32_t)((uint32_t)(k >> 32) ^ 0x80000000u);
        if (d <= 0 || d == INT_MAX) d = 0;  // not a depth a source pixel can carry (the rule requires source > 0)
        else if (a.pose_label) label = a.pose_label[(uint32_t)k] + 1;  // source labels start at 1 (image_renderer.cuh:165)
    }
    int32_t base = a.base_depth ? a.base_depth[i] : 0;
    int base_label = a.base_label ? (int)a.base_label[i] : 0;
    // PERCH_COMPOSE_MERGE keeps whichever surface is nearer; a pixel empty in one image takes the other
    const bool take_placed = d > 0 && !(a.merge && base > 0 && base <= d);
    a.out_depth[i] = take_placed ? d : (a.merge ? base : 0);
    if (a.out_label) a.out_label[i] = (uint8_t)(take_placed ? label : (a.merge ? base_label : 0));
}

}  // namespace

cudaError_t launch_compose(const ComposeArgs& a, cudaStream_t stream) {
    if (a.n_pixels <= 0) return cudaSuccess;
    compose_kernel<<<(a.n_pixels + 255) / 256, 256, 0, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace perch
