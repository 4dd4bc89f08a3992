// Depth image(s) -> compacted point cloud: the work of compute_point_clouds (compute_point_clouds.cuh:188-367)
// without its full-resolution P*W*H int mask + scan.  Only lattice rows are counted (one int per row), scanned, and
// written in the reference's order: image-major, then row, then column (the order its exclusive scan over the
// full-resolution mask produces, compute_point_clouds.cuh:290-292).
#include <cuda_runtime.h>

#include <climits>

#include "capi_common.hpp"
#include "kernels.cuh"
#include "perch_b200.h"

namespace perch {
namespace {

constexpr int kRowThreads = 128;

struct Sample {
    bool in_mask;   // depth_to_mask says this pixel owns a cloud slot (compute_point_clouds.cuh:64-91)
    bool written;   // depth_to_2d_cloud actually writes it (compute_point_clouds.cuh:123-139)
    int32_t d;
    int u, v;
};

// World-frame point for the bounds test (compute_point_clouds.cuh:25-32).  Row sums use the fma chain nvcc emits for the
// 3x3 * 3x1 product of the oracle's Eigen stand-in; real Eigen's association is not pinned by any reference test.
__device__ __forceinline__ void to_world(const CloudArgs& a, float x, float y, float z, float& wx, float& wy, float& wz) {
    const float* m = a.cam_tf;  // column-major
    wx = __fadd_rn(__fmaf_rn(m[8], z, __fmaf_rn(m[4], y, __fmul_rn(m[0], x))), m[12]);
    wy = __fadd_rn(__fmaf_rn(m[9], z, __fmaf_rn(m[5], y, __fmul_rn(m[1], x))), m[13]);
    wz = __fadd_rn(__fmaf_rn(m[10], z, __fmaf_rn(m[6], y, __fmul_rn(m[2], x))), m[14]);
}

__device__ __forceinline__ bool inside_bounds(const CloudArgs& a, int u, int v, int32_t d) {
    float x, y, z, wx, wy, wz;
    back_project(u, v, d, a.cx, a.cy, a.fx, a.fy, a.depth_factor, x, y, z);
    to_world(a, x, y, z, wx, wy, wz);
    if (wx > a.bounds[0] || wx < a.bounds[1]) return false;
    if (wy > a.bounds[2] || wy < a.bounds[3]) return false;
    if (wz > a.bounds[4] || wz < a.bounds[5]) return false;
    return true;
}

__device__ __forceinline__ Sample sample_at(const CloudArgs& a, int img, int lx, int ly) {
    Sample s;
    const int ix = lx * a.stride, iy = ly * a.stride;  // position inside the input image
    s.u = (a.lattice_coords ? lx : ix * a.coord_scale) + a.u_off;
    s.v = (a.lattice_coords ? ly : iy * a.coord_scale) + a.v_off;
    size_t idx = ((size_t)img * a.img_h + iy) * a.img_w + ix;
    s.d = a.depth[idx];
    s.in_mask = false;
    s.written = false;
    if (s.d > 0) {
        bool label_ok = a.label_mask ? a.label_mask[idx] > 0 : true;
        bool bounds_ok = a.use_bounds ? inside_bounds(a, s.u, s.v, s.d) : true;
        // depth_to_mask: label mask wins over bounds (else-if chain); depth_to_2d_cloud applies both
        s.in_mask = a.label_mask ? label_ok : bounds_ok;
        s.written = label_ok && bounds_ok;
    }
    return s;
}

__global__ void __launch_bounds__(kRowThreads) cloud_count_kernel(CloudArgs a) {
    const int ly = blockIdx.x, img = blockIdx.y;
    const int Lw = a.img_w / a.stride;
    int cnt = 0;
    for (int lx = threadIdx.x; lx < Lw; lx += kRowThreads) cnt += sample_at(a, img, lx, ly).in_mask ? 1 : 0;
    __shared__ int s_sum;
    if (threadIdx.x == 0) s_sum = 0;
    __syncthreads();
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(&s_sum, cnt);
    __syncthreads();
    if (threadIdx.x == 0) a.row_counts[img * gridDim.x + ly] = s_sum;
}

// Single-CTA exclusive scan of n ints in place; element n receives the total.
__global__ void __launch_bounds__(1024) scan_kernel(int* data, int n) {
    __shared__ int s_part[1024];
    const int tid = threadIdx.x;
    const int per = (n + 1023) / 1024;
    const int b = min(tid * per, n), e = min(b + per, n);
    int sum = 0;
    for (int i = b; i < e; ++i) sum += data[i];
    s_part[tid] = sum;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        int v = (tid >= o) ? s_part[tid - o] : 0;
        __syncthreads();
        s_part[tid] += v;
        __syncthreads();
    }
    int run = s_part[tid] - sum;
    for (int i = b; i < e; ++i) {
        int v = data[i];
        data[i] = run;
        run += v;
    }
    if (tid == 1023) data[n] = s_part[1023];
}

__global__ void __launch_bounds__(kRowThreads) cloud_write_kernel(CloudArgs a) {
    const int ly = blockIdx.x, img = blockIdx.y, Lh = gridDim.x;
    const int Lw = a.img_w / a.stride;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ int s_warp[kRowThreads / 32];
    __shared__ int s_base;
    if (threadIdx.x == 0) s_base = a.row_counts[img * Lh + ly];
    __syncthreads();
    for (int x0 = 0; x0 < Lw; x0 += kRowThreads) {
        const int lx = x0 + threadIdx.x;
        Sample s;
        s.in_mask = false; s.written = false;
        if (lx < Lw) s = sample_at(a, img, lx, ly);
        unsigned ballot = __ballot_sync(0xffffffffu, s.in_mask);
        if (lane == 0) s_warp[warp] = __popc(ballot);
        __syncthreads();
        int off = s_base;
        for (int w = 0; w < warp; ++w) off += s_warp[w];
        int slot = off + __popc(ballot & ((1u << lane) - 1u));
        if (s.in_mask && s.written && slot < a.capacity) {
            float x, y, z;
            back_project(s.u, s.v, s.d, a.cx, a.cy, a.fx, a.fy, a.depth_factor, x, y, z);
            if (a.cloud_soa) {
                a.cloud_soa[slot] = x;
                a.cloud_soa[(size_t)a.capacity + slot] = y;
                a.cloud_soa[2 * (size_t)a.capacity + slot] = z;
            }
            if (a.cloud_aos) {
                a.cloud_aos[3 * (size_t)slot + 0] = x;
                a.cloud_aos[3 * (size_t)slot + 1] = y;
                a.cloud_aos[3 * (size_t)slot + 2] = z;
            }
            if (a.cloud_rgb) {
                size_t idx = ((size_t)img * a.img_h + ly * a.stride) * a.img_w + lx * a.stride;
                a.cloud_rgb[slot] = a.red ? a.red[idx] : 0;
                a.cloud_rgb[(size_t)a.capacity + slot] = a.green ? a.green[idx] : 0;
                a.cloud_rgb[2 * (size_t)a.capacity + slot] = a.blue ? a.blue[idx] : 0;
            }
            if (a.pose_map) a.pose_map[slot] = img;
            if (a.cloud_label) {
                // compute_point_clouds.cuh:168-178: observed label = mask-1; rendered label = pose label
                if (a.label_mask) {
                    size_t idx = ((size_t)img * a.img_h + ly * a.stride) * a.img_w + lx * a.stride;
                    a.cloud_label[slot] = (int)a.label_mask[idx] - 1;
                } else if (a.pose_label) {
                    a.cloud_label[slot] = a.pose_label[img];
                }
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int tot = 0;
            for (int w = 0; w < kRowThreads / 32; ++w) tot += s_warp[w];
            s_base += tot;
        }
        __syncthreads();
    }
}

// result_dc_index (compute_point_clouds.cuh:267,290-292): exclusive scan of the sample mask over the full-resolution
// (image, y, x) order.  Only the sampled pixels carry information; this fills the whole array for drop-in parity.
__global__ void dc_index_kernel(CloudArgs a) {
    const int W = a.img_w * a.coord_scale, H = a.full_h;
    const int step = a.stride * a.coord_scale;
    const int Lh = (H + step - 1) / step;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t total = (size_t)a.num_images * W * H;
    if (i >= total) return;
    int x = (int)(i % W), y = (int)((i / W) % H), img = (int)(i / ((size_t)W * H));
    int ly = y / step;
    if (y % step != 0) {
        // between lattice rows: everything up to and including lattice row ly has been counted
        a.dc_index[i] = a.row_counts[img * Lh + ly + 1];
        return;
    }
    // on a lattice row: row prefix + valid samples strictly left of x
    int cnt = a.row_counts[img * Lh + ly];
    int nlx = (x + step - 1) / step;  // samples with lx*step < x
    for (int lx = 0; lx < nlx; ++lx) cnt += sample_at(a, img, lx, ly).in_mask ? 1 : 0;
    a.dc_index[i] = cnt;
}

}  // namespace

cudaError_t launch_cloud_count(const CloudArgs& a, cudaStream_t stream) {
    const int Lh = (a.img_h + a.stride - 1) / a.stride;
    if (a.num_images <= 0 || Lh <= 0) return cudaSuccess;
    cloud_count_kernel<<<dim3(Lh, a.num_images), kRowThreads, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_cloud_scan(const CloudArgs& a, cudaStream_t stream) {
    const int Lh = (a.img_h + a.stride - 1) / a.stride;
    scan_kernel<<<1, 1024, 0, stream>>>(a.row_counts, a.num_images * Lh);
    return cudaGetLastError();
}

cudaError_t launch_exclusive_scan(int* d_data, int n, cudaStream_t stream) {
    scan_kernel<<<1, 1024, 0, stream>>>(d_data, n);
    return cudaGetLastError();
}

cudaError_t launch_cloud_write(const CloudArgs& a, cudaStream_t stream) {
    const int Lh = (a.img_h + a.stride - 1) / a.stride;
    if (a.num_images <= 0 || Lh <= 0) return cudaSuccess;
    cloud_write_kernel<<<dim3(Lh, a.num_images), kRowThreads, 0, stream>>>(a);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) return err;
    if (a.dc_index) {
        size_t total = (size_t)a.num_images * a.img_w * a.coord_scale * a.full_h;
        dc_index_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(a);
        err = cudaGetLastError();
    }
    return err;
}

}  // namespace perch

// ---- allocation-free variant: perch_depth2cloud_static (include/perch_b200.h) -----------------------------------------
// No counterpart in the reference.  For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops):
// the only scratch the cloud kernels need -- one int per lattice row plus the total -- lives in a file-scope device
// array instead of the context's cudaMalloc'd scratch.  Every GPU that loads the library holds the array
// (PERCH_STATIC_CLOUD_MAX_ROWS + 64 ints = 4 MiB) until exit.  All calls on one device share it: calls that overlap on
// the GPU race.  Nothing in it has to start at zero (the count kernel writes every row slot of the call, the scan
// writes the total), so the previous call's contents are simply overwritten.
static __device__ __align__(256) int g_static_row_counts[PERCH_STATIC_CLOUD_MAX_ROWS + 64];

namespace {
constexpr int kMaxDevices = 64;
int* g_static_rows_dev[kMaxDevices];  // device address of g_static_row_counts per device (host code cannot use the name)

// Must be called with `device` current.  The symbol's name is a host-side stand-in, not a device address.
cudaError_t static_rows_ptr(int device, int** out) {
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    int* p = __atomic_load_n(&g_static_rows_dev[device], __ATOMIC_ACQUIRE);
    if (!p) {
        cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_row_counts);
        if (e != cudaSuccess) return e;
        __atomic_store_n(&g_static_rows_dev[device], p, __ATOMIC_RELEASE);  // same value from every thread
    }
    *out = p;
    return cudaSuccess;
}
}  // namespace

extern "C" int perch_depth2cloud_static(perch_ctx* c, const int32_t* d_depth, const uint8_t* d_label_mask, int width,
                                        int height, int num_images, float cx, float cy, float fx, float fy,
                                        float depth_factor, int stride, int capacity, float* d_out_cloud_aos,
                                        float* d_out_cloud_soa, int* d_out_cloud_pose_map, int* d_out_cloud_label,
                                        int* point_num) {
    using perch_detail::fail;
    if (!c || !d_depth || !point_num) return fail(PERCH_ERR_INVALID, "null argument");
    if (width <= 0 || height <= 0 || num_images <= 0 || stride <= 0 || capacity < 0)
        return fail(PERCH_ERR_INVALID, "bad size argument");
    if (width % stride != 0)
        return fail(PERCH_ERR_INVALID, "stride %d must divide the image width %d (compute_point_clouds.cuh:271)", stride, width);
    if ((size_t)width * height * num_images > (size_t)INT_MAX)
        return fail(PERCH_ERR_CAPACITY, "num_images*width*height exceeds INT_MAX");
    const int Lh = (height + stride - 1) / stride;
    const size_t rows = (size_t)num_images * Lh;
    if (rows > (size_t)PERCH_STATIC_CLOUD_MAX_ROWS)
        return fail(PERCH_ERR_CAPACITY,
                    "perch_depth2cloud_static: %d images x %d lattice rows = %zu rows exceed the static scratch of %d rows "
                    "(use perch_depth2cloud, which allocates)", num_images, Lh, rows, PERCH_STATIC_CLOUD_MAX_ROWS);
    perch_detail::DeviceGuard g(c->device);
    int* d_rows = nullptr;
    PERCH_CU(static_rows_ptr(c->device, &d_rows));
    perch::CloudArgs a;
    a.depth = d_depth;
    a.label_mask = d_label_mask;
    a.img_w = width; a.img_h = height; a.full_h = height; a.num_images = num_images;
    a.stride = stride; a.coord_scale = 1;
    a.cx = cx; a.cy = cy; a.fx = fx; a.fy = fy; a.depth_factor = depth_factor;
    a.row_counts = d_rows;
    a.capacity = capacity;
    a.cloud_aos = d_out_cloud_aos;
    a.cloud_soa = d_out_cloud_soa;
    a.pose_map = d_out_cloud_pose_map;
    a.cloud_label = d_label_mask ? d_out_cloud_label : nullptr;
    PERCH_CU(perch::launch_cloud_count(a, c->stream));
    PERCH_CU(perch::launch_cloud_scan(a, c->stream));
    if (d_out_cloud_aos || d_out_cloud_soa || d_out_cloud_pose_map || a.cloud_label) PERCH_CU(perch::launch_cloud_write(a, c->stream));
    int total = 0;
    PERCH_CU(cudaMemcpyAsync(&total, d_rows + rows, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    PERCH_CU(cudaStreamSynchronize(c->stream));
    *point_num = total;
    if (total > capacity && (d_out_cloud_aos || d_out_cloud_soa || d_out_cloud_pose_map || a.cloud_label))
        return fail(PERCH_ERR_CAPACITY, "cloud has %d points but the buffers hold %d (the first %d were written)", total,
                    capacity, capacity);
    return PERCH_OK;
}
