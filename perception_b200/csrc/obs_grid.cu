// The observed cloud on the device: bounding box, stable sort by segmentation label (renderer.cu:1668-1686) and the
// uniform nearest-neighbour grid of the cost stage / GICP (context.hpp: ObsGridDev), so that perch_set_input
// (search_env.cpp:5880-6069) leaves the scene ready without the cloud visiting the host.
//
//   obs_bbox_kernel        min / max of the coordinates (order-preserving int encoding, atomicMin / atomicMax)
//   label_hist_kernel   \  stable counting sort by label in [0, 256): per-chunk histograms, one single-CTA pass that turns
//   label_offsets_kernel >  them into (label, chunk) offsets + per-label ranges, and an in-order scatter in which a warp
//   label_scatter_kernel/  walks its chunk 32 points at a time (rank inside the 32 by __match_any_sync) -- the same order
//                          std::stable_sort gives on the host path of perch_set_observed
//   cell_count_kernel   \  histogram of points per grid cell -> exclusive scan (cloud.cu's single-CTA scan) -> scatter with
//   cell_scatter_kernel /  an atomic cursor per cell.  The order INSIDE a cell is therefore arbitrary; every query breaks
//                          distance ties on the reference index explicitly (pose_kernels.cu nn_query, gicp.cu), so results
//                          do not depend on it.
#include <cuda_runtime.h>

#include <algorithm>
#include <cfloat>
#include <cstring>

#include "kernels.cuh"

namespace perch {
namespace {

constexpr int kChunk = 1024;    // points per warp in the label sort
constexpr int kLabels = 256;

__device__ __forceinline__ int f2ord(float f) {  // order-preserving float -> int
    const int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7fffffff;
}

__global__ void obs_bbox_kernel(const float* aos, int n, int* out6) {
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float v = aos[3 * (size_t)i + k];
            mn[k] = fminf(mn[k], v);
            mx[k] = fmaxf(mx[k], v);
        }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(&out6[k], f2ord(mn[k]));
            atomicMax(&out6[3 + k], f2ord(mx[k]));
        }
    }
}

// one warp per chunk of kChunk points; hist[chunk * kLabels + l]
__global__ void __launch_bounds__(128) label_hist_kernel(const int* label, int n, int n_chunks, int* hist, int* bad) {
    __shared__ int s_cnt[4][kLabels];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int chunk = blockIdx.x * 4 + warp;
    for (int b = lane; b < kLabels; b += 32) s_cnt[warp][b] = 0;
    __syncwarp();
    if (chunk < n_chunks) {
        const int lo = chunk * kChunk, hi = min(lo + kChunk, n);
        for (int i = lo + lane; i < hi; i += 32) {
            const int l = label[i];
            if (l < 0 || l >= kLabels) { atomicOr(bad, 1); continue; }
            atomicAdd(&s_cnt[warp][l], 1);
        }
        __syncwarp();
        for (int b = lane; b < kLabels; b += 32) hist[(size_t)chunk * kLabels + b] = s_cnt[warp][b];
    }
}

// single CTA: hist[chunk][l] -> start offset of (l, chunk) in the sorted array; ranges[2l], ranges[2l+1] = [begin, end) of l
__global__ void __launch_bounds__(kLabels) label_offsets_kernel(int* hist, int n_chunks, int* ranges) {
    __shared__ int s_tot[kLabels];
    const int l = threadIdx.x;
    int sum = 0;
    for (int c = 0; c < n_chunks; ++c) sum += hist[(size_t)c * kLabels + l];
    s_tot[l] = sum;
    __syncthreads();
    int base = 0;
    for (int k = 0; k < l; ++k) base += s_tot[k];
    ranges[2 * l] = sum ? base : 0;
    ranges[2 * l + 1] = sum ? base + sum : 0;
    for (int c = 0; c < n_chunks; ++c) {
        const int v = hist[(size_t)c * kLabels + l];
        hist[(size_t)c * kLabels + l] = base;
        base += v;
    }
}

__global__ void __launch_bounds__(128) label_scatter_kernel(const float* aos, const int* label, int n, int n_chunks,
                                                            const int* offsets, float4* out_xyz4, int* out_label) {
    __shared__ int s_cur[4][kLabels];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int chunk = blockIdx.x * 4 + warp;
    if (chunk >= n_chunks) return;
    for (int b = lane; b < kLabels; b += 32) s_cur[warp][b] = offsets[(size_t)chunk * kLabels + b];
    __syncwarp();
    const int lo = chunk * kChunk, hi = min(lo + kChunk, n);
    for (int i0 = lo; i0 < hi; i0 += 32) {
        const int i = i0 + lane;
        const bool on = i < hi;
        const int l = on ? (label[i] & (kLabels - 1)) : 0;
        const unsigned active = __ballot_sync(0xffffffffu, on);
        if (on) {
            const unsigned peers = __match_any_sync(active, l);
            const int rank = __popc(peers & ((1u << lane) - 1u));
            const int dst = s_cur[warp][l] + rank;
            out_xyz4[dst] = make_float4(aos[3 * (size_t)i], aos[3 * (size_t)i + 1], aos[3 * (size_t)i + 2], 0.f);
            out_label[dst] = l;
            __syncwarp(active);
            if (rank == 0) s_cur[warp][l] += __popc(peers);
        }
        __syncwarp();
    }
}

__global__ void aos_to_xyz4_kernel(const float* aos, int n, float4* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = make_float4(aos[3 * (size_t)i], aos[3 * (size_t)i + 1], aos[3 * (size_t)i + 2], 0.f);
}

__device__ __forceinline__ uint32_t cell_of(const ObsGridDev& g, float4 p) {
    // the host builder's arithmetic (capi.cu build_point_grid_from) and the query's (pose_kernels.cu nn_query)
    const int cx = min(max((int)floorf(__fmul_rn(__fsub_rn(p.x, g.ox), g.inv_h)), 0), g.dx - 1);
    const int cy = min(max((int)floorf(__fmul_rn(__fsub_rn(p.y, g.oy), g.inv_h)), 0), g.dy - 1);
    const int cz = min(max((int)floorf(__fmul_rn(__fsub_rn(p.z, g.oz), g.inv_h)), 0), g.dz - 1);
    return (uint32_t)(((size_t)cz * g.dy + cy) * g.dx + cx);
}

__global__ void cell_count_kernel(const float4* pts, int n, ObsGridDev g, int* counts) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) atomicAdd(&counts[cell_of(g, pts[i])], 1);
}

__global__ void cell_nonempty_kernel(const int* counts, size_t ncells, int* out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool on = i < ncells && counts[i] > 0;
    const unsigned b = __ballot_sync(0xffffffffu, on);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(out, __popc(b));
}

__global__ void cell_scatter_kernel(const float4* pts, const int* labels, int n, ObsGridDev g, const uint32_t* cell_start,
                                    int* cursor, float4* out_pts, int* out_label) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 p = pts[i];
    const uint32_t c = cell_of(g, p);
    const uint32_t slot = cell_start[c] + (uint32_t)atomicAdd(&cursor[c], 1);
    p.w = __int_as_float(i);  // reference index = index in the label-sorted cloud
    out_pts[slot] = p;
    if (labels) out_label[slot] = labels[i];
}

}  // namespace

cudaError_t launch_obs_bbox(const float* d_aos, int n, int* d_out6, cudaStream_t stream) {
    const int init[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000};
    cudaError_t e = cudaMemcpyAsync(d_out6, init, sizeof init, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess || n <= 0) return e;
    obs_bbox_kernel<<<std::min((n + 255) / 256, 1024), 256, 0, stream>>>(d_aos, n, d_out6);
    return cudaGetLastError();
}

float obs_bbox_decode(int ord) {
    const int i = ord >= 0 ? ord : ord ^ 0x7fffffff;
    float f;
    memcpy(&f, &i, 4);
    return f;
}

int label_sort_chunks(int n) { return (n + kChunk - 1) / kChunk; }

cudaError_t launch_label_sort(const float* d_aos, const int* d_label, int n, int* d_hist /* chunks x 256 */, int* d_ranges /* 512 */,
                              int* d_bad, float4* d_out_xyz4, int* d_out_label, cudaStream_t stream) {
    const int n_chunks = label_sort_chunks(n);
    if (n <= 0) return cudaMemsetAsync(d_ranges, 0, 2 * kLabels * sizeof(int), stream);
    cudaError_t e = cudaMemsetAsync(d_bad, 0, sizeof(int), stream);
    if (e != cudaSuccess) return e;
    label_hist_kernel<<<(n_chunks + 3) / 4, 128, 0, stream>>>(d_label, n, n_chunks, d_hist, d_bad);
    label_offsets_kernel<<<1, kLabels, 0, stream>>>(d_hist, n_chunks, d_ranges);
    label_scatter_kernel<<<(n_chunks + 3) / 4, 128, 0, stream>>>(d_aos, d_label, n, n_chunks, d_hist, d_out_xyz4, d_out_label);
    return cudaGetLastError();
}

cudaError_t launch_aos_to_xyz4(const float* d_aos, int n, float4* d_out, cudaStream_t stream) {
    if (n > 0) aos_to_xyz4_kernel<<<(n + 255) / 256, 256, 0, stream>>>(d_aos, n, d_out);
    return cudaGetLastError();
}

// d_cell_start: ncells + 1 entries (zeroed here), d_cursor: ncells ints of scratch
cudaError_t launch_obs_grid_build(const float4* d_sorted, const int* d_sorted_label, int n, const ObsGridDev& g, uint32_t* d_cell_start,
                                  int* d_cursor, float4* d_out_pts, int* d_out_label, int* d_nonempty, cudaStream_t stream) {
    const size_t ncells = (size_t)g.dx * g.dy * g.dz;
    cudaError_t e = cudaMemsetAsync(d_cell_start, 0, (ncells + 1) * sizeof(uint32_t), stream);
    if (e != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(d_cursor, 0, ncells * sizeof(int), stream)) != cudaSuccess) return e;
    if (n <= 0) return cudaSuccess;
    cell_count_kernel<<<(n + 255) / 256, 256, 0, stream>>>(d_sorted, n, g, reinterpret_cast<int*>(d_cell_start));
    if (d_nonempty) {
        if ((e = cudaMemsetAsync(d_nonempty, 0, sizeof(int), stream)) != cudaSuccess) return e;
        cell_nonempty_kernel<<<(unsigned)((ncells + 255) / 256), 256, 0, stream>>>(reinterpret_cast<const int*>(d_cell_start), ncells, d_nonempty);
    }
    if ((e = launch_exclusive_scan(reinterpret_cast<int*>(d_cell_start), (int)ncells, stream)) != cudaSuccess) return e;
    cell_scatter_kernel<<<(n + 255) / 256, 256, 0, stream>>>(d_sorted, d_sorted_label, n, g, d_cell_start, d_cursor, d_out_pts, d_out_label);
    return cudaGetLastError();
}

}  // namespace perch
