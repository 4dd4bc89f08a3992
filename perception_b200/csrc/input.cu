// Observed-scene preparation on the device: the image steps of EnvObjectRecognition::SetInput's GPU
// branch (sbpl_perception/src/search_env.cpp:5880-5952) and the source-depth rescale of ComputeGreedyRenderPoses
// (2487-2498).
#include <cstdint>

#include "kernels.cuh"

namespace perch {
namespace {

constexpr int kMedTileW = 32, kMedTileH = 8, kMedMaxK = 31;

// cv::medianBlur on an 8-bit single-channel image (search_env.cpp:5922): exact median of the k x k window with
// replicated borders.  Per pixel: binary search for the smallest value v with #(window <= v) >= (k*k + 1) / 2.
__global__ void median_blur_u8_kernel(const uint8_t* in, uint8_t* out, int w, int h, int k) {
    __shared__ uint8_t tile[(kMedTileH + kMedMaxK - 1) * (kMedTileW + kMedMaxK - 1)];
    const int r = k / 2;
    const int tw = kMedTileW + 2 * r, th = kMedTileH + 2 * r;
    const int bx = blockIdx.x * kMedTileW, by = blockIdx.y * kMedTileH;
    for (int i = threadIdx.y * kMedTileW + threadIdx.x; i < tw * th; i += kMedTileW * kMedTileH) {
        const int ty = i / tw, tx = i % tw;
        const int sx = min(max(bx + tx - r, 0), w - 1), sy = min(max(by + ty - r, 0), h - 1);  // BORDER_REPLICATE
        tile[i] = in[(size_t)sy * w + sx];
    }
    __syncthreads();
    const int x = bx + threadIdx.x, y = by + threadIdx.y;
    if (x >= w || y >= h) return;
    const int need = (k * k + 1) / 2;
    int lo = 0, hi = 255;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        int cnt = 0;
        for (int dy = 0; dy < k; ++dy) {
            const uint8_t* row = tile + (threadIdx.y + dy) * tw + threadIdx.x;
            for (int dx = 0; dx < k; ++dx) cnt += row[dx] <= mid;
        }
        if (cnt >= need) hi = mid;
        else lo = mid + 1;
    }
    out[(size_t)y * w + x] = (uint8_t)lo;
}

// input_depth_image_vec.assign(ptr<uchar | unsigned short>(0), ...) (search_env.cpp:5911-5914, 5926-5929)
__global__ void depth_to_i32_kernel(const void* in, int bits, int32_t* out, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = bits == 8 ? (int32_t) reinterpret_cast<const uint8_t*>(in)[i] : (int32_t) reinterpret_cast<const uint16_t*>(in)[i];
}

// `input_depth_image_vec[i] /= division_factor` with an int32 element and a float factor (search_env.cpp:2489-2496):
// int -> float, IEEE divide, truncation toward zero
__global__ void divide_depth_kernel(const int32_t* in, float factor, int32_t* out, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = (int32_t)__fdiv_rn((float)in[i], factor);
}

}  // namespace

cudaError_t launch_median_blur_u8(const uint8_t* in, uint8_t* out, int w, int h, int k, cudaStream_t stream) {
    dim3 block(kMedTileW, kMedTileH), grid((w + kMedTileW - 1) / kMedTileW, (h + kMedTileH - 1) / kMedTileH);
    median_blur_u8_kernel<<<grid, block, 0, stream>>>(in, out, w, h, k);
    return cudaGetLastError();
}

cudaError_t launch_depth_to_i32(const void* in, int bits, int32_t* out, int n, cudaStream_t stream) {
    if (n > 0) depth_to_i32_kernel<<<(n + 255) / 256, 256, 0, stream>>>(in, bits, out, n);
    return cudaGetLastError();
}

cudaError_t launch_divide_depth(const int32_t* in, float factor, int32_t* out, int n, cudaStream_t stream) {
    if (n > 0) divide_depth_kernel<<<(n + 255) / 256, 256, 0, stream>>>(in, factor, out, n);
    return cudaGetLastError();
}

}  // namespace perch
