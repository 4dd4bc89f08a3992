// sRGB (8-bit) -> CIELAB (D65) for the colour cost (cost_type 1).  The reference converts both colours of every
// rendered/observed point pair inside compute_render_cost (compute_costs.cuh:57-88,225-228); the conversion depends
// only on the 8-bit colour, so here it runs once per triangle (set_models) and once per observed point (set_observed).
// Double precision with the reference's constants and operation order, results narrowed to float like its `lab[]`.
#include <cuda_runtime.h>

#include "kernels.cuh"

namespace perch {
namespace {

__device__ __forceinline__ double srgb_to_linear_pct(double c) {
    return ((c > 0.04045) ? pow((c + 0.055) / 1.055, 2.4) : (c / 12.92)) * 100.0;
}

__device__ __forceinline__ double lab_f(double t) { return (t > 0.008856) ? cbrt(t) : (7.787 * t + 16.0 / 116.0); }

__device__ void rgb8_to_lab(uint8_t rr, uint8_t gg, uint8_t bb, float* lab) {
    double r = srgb_to_linear_pct(rr / 255.0);
    double g = srgb_to_linear_pct(gg / 255.0);
    double b = srgb_to_linear_pct(bb / 255.0);
    double x = r * 0.4124564 + g * 0.3575761 + b * 0.1804375;
    double y = r * 0.2126729 + g * 0.7151522 + b * 0.0721750;
    double z = r * 0.0193339 + g * 0.1191920 + b * 0.9503041;
    x = lab_f(x / 95.047);
    y = lab_f(y / 100.00);
    z = lab_f(z / 108.883);
    lab[0] = (float)((116.0 * y) - 16);
    lab[1] = (float)(500 * (x - y));
    lab[2] = (float)(200 * (y - z));
}

__global__ void rgb2lab_planes_kernel(const uint8_t* c0, const uint8_t* c1, const uint8_t* c2, int n, float* lab) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    rgb8_to_lab(c2[i], c1[i], c0[i], lab + 3 * (size_t)i);
}

__global__ void rgb2lab_packed_kernel(const uint32_t* packed, int n, float* lab) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t c = packed[i];  // r | g<<8 | b<<16 == colour planes 0,1,2 of a rendered point
    rgb8_to_lab((uint8_t)(c >> 16), (uint8_t)(c >> 8), (uint8_t)c, lab + 3 * (size_t)i);
}

}  // namespace

cudaError_t launch_rgb2lab_planes(const uint8_t* c0, const uint8_t* c1, const uint8_t* c2, int n, float* lab,
                                  cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    rgb2lab_planes_kernel<<<(n + 255) / 256, 256, 0, stream>>>(c0, c1, c2, n, lab);
    return cudaGetLastError();
}

cudaError_t launch_rgb2lab_packed(const uint32_t* packed_rgb, int n, float* lab, cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    rgb2lab_packed_kernel<<<(n + 255) / 256, 256, 0, stream>>>(packed_rgb, n, lab);
    return cudaGetLastError();
}

}  // namespace perch
