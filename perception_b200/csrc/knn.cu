// Exact k-nearest-neighbour search with the contract of the reference's kNN-CUDA package API
// (cuda_renderer/knncuda.h:33-100; knn_cuda_global knncuda.cu:276-421): points are dim x N row-per-dimension arrays,
// outputs are k x query_nb arrays of Euclidean distances (square root applied) and 0-based reference indices, ascending
// in distance, equal distances in ascending index order (the reference's insertion sort moves an element only past
// strictly larger ones, knncuda.cu:149-195).
//
// The reference materialises the ref_nb x query_nb distance matrix in HBM (4 * Nr * Nq bytes written, then re-read
// by a per-column insertion sort).  Here nothing of that size exists: a thread owns kQpt queries, reference points
// stream through shared memory in tiles (every LDS is a warp-wide broadcast shared by kQpt distance evaluations), and
// the k best of each query live directly in the k x query_nb output arrays (column accesses are coalesced across the
// threads of a warp; after the first few tiles insertions are rare, the k-th best is cached in a register).
// Squared distances use the reference kernel's arithmetic -- ssd = fma(diff, diff, ssd) over the dimensions in order
// (compute_distances, knncuda.cu:26-96, contracted by nvcc's default -fmad) -- and sqrt.rn at the end
// (compute_sqrt, knncuda.cu:206-211), so distances and indices are bit-identical to knn_cuda_global.
#include <cuda_runtime.h>

#include <cfloat>

#include "kernels.cuh"

namespace perch {
namespace {

constexpr int kKnnThreads = 128;
constexpr int kQpt = 4;        // queries per thread (fixed-dimension kernels)
constexpr int kTile = 1024;    // reference points per shared-memory tile (x DIM floats)

// insert (d2, j) into the ascending list of column q; `count` entries are valid, `worst` caches entry k-1 when full
__device__ __forceinline__ void knn_insert(float* dist, int* index, size_t pitch, int k, int& count, float& worst,
                                           float d2, int j) {
    int pos;
    if (count < k) pos = count++;
    else if (d2 >= worst) return;  // same skip rule as knncuda.cu:174 (false for NaN, which then overwrites slot k-1)
    else pos = k - 1;
    while (pos > 0) {
        float prev = dist[(size_t)(pos - 1) * pitch];
        if (!(prev > d2)) break;
        dist[(size_t)pos * pitch] = prev;
        index[(size_t)pos * pitch] = index[(size_t)(pos - 1) * pitch];
        --pos;
    }
    dist[(size_t)pos * pitch] = d2;
    index[(size_t)pos * pitch] = j;
    if (count == k) worst = dist[(size_t)(k - 1) * pitch];
}

template <int DIM>
__global__ void __launch_bounds__(kKnnThreads) knn_fixed_kernel(const float* __restrict__ ref, int ref_nb,
                                                                const float* __restrict__ query, int query_nb, int k,
                                                                float* __restrict__ out_dist, int* __restrict__ out_index) {
    __shared__ float tile[DIM][kTile];
    const int q0 = (blockIdx.x * kKnnThreads + threadIdx.x);  // queries q0 + i * stride_q: coalesced columns
    const int stride_q = gridDim.x * kKnnThreads;
    float qv[kQpt][DIM];
    float worst[kQpt];
    int count[kQpt];
    bool live[kQpt];
#pragma unroll
    for (int i = 0; i < kQpt; ++i) {
        const int q = q0 + i * stride_q;
        live[i] = q < query_nb;
        worst[i] = FLT_MAX;
        count[i] = 0;
#pragma unroll
        for (int d = 0; d < DIM; ++d) qv[i][d] = live[i] ? __ldg(query + (size_t)d * query_nb + q) : 0.f;
    }
    for (int base = 0; base < ref_nb; base += kTile) {
        const int n = min(kTile, ref_nb - base);
        __syncthreads();
        for (int t = threadIdx.x; t < n; t += kKnnThreads)
#pragma unroll
            for (int d = 0; d < DIM; ++d) tile[d][t] = __ldg(ref + (size_t)d * ref_nb + base + t);
        __syncthreads();
        for (int t = 0; t < n; ++t) {
            float r[DIM];
#pragma unroll
            for (int d = 0; d < DIM; ++d) r[d] = tile[d][t];
#pragma unroll
            for (int i = 0; i < kQpt; ++i) {
                float ssd = 0.f;
#pragma unroll
                for (int d = 0; d < DIM; ++d) {
                    const float diff = __fsub_rn(r[d], qv[i][d]);
                    ssd = __fmaf_rn(diff, diff, ssd);
                }
                if (live[i] && !(count[i] == k && ssd >= worst[i])) {
                    const int q = q0 + i * stride_q;
                    knn_insert(out_dist + q, out_index + q, (size_t)query_nb, k, count[i], worst[i], ssd, base + t);
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < kQpt; ++i) {
        const int q = q0 + i * stride_q;
        if (!live[i]) continue;
        for (int j = 0; j < k; ++j) out_dist[(size_t)j * query_nb + q] = __fsqrt_rn(out_dist[(size_t)j * query_nb + q]);
    }
}

// k <= KMAX <= 16: the k best of a query live in REGISTERS (unrolled, predicated insertion: no memory traffic, no
// dependent global loads while a warp waits on the one lane that found a closer point) and are written out once, sorted.
// Same insertion rule as knn_insert, element by element: entry `pos` (= count or k-1) takes the new value, which then
// moves down past every strictly larger predecessor (NaN compares false and stays where it is, knncuda.cu:149-195).
template <int KMAX>
struct RegList {
    float d[KMAX];
    int i[KMAX];
    int count = 0;
    __device__ __forceinline__ void insert(int k, float d2, int j) {
        int pos;
        if (count < k) pos = count++;
        else pos = k - 1;  // the caller has already applied the skip rule against `worst`
        bool moving = true;
#pragma unroll
        for (int s = KMAX - 1; s >= 1; --s) {
            if (s > pos || !moving) continue;
            if (d[s - 1] > d2) {
                d[s] = d[s - 1];
                i[s] = i[s - 1];
            } else {
                d[s] = d2;
                i[s] = j;
                moving = false;
            }
        }
        if (moving) {
            d[0] = d2;
            i[0] = j;
        }
    }
    // entry k-1 once the list is full
    __device__ __forceinline__ float worst(int k) const {
        float w = d[0];
#pragma unroll
        for (int s = 1; s < KMAX; ++s) w = (s == k - 1) ? d[s] : w;
        return w;
    }
};

template <int DIM, int KMAX, int QPT>
__global__ void __launch_bounds__(kKnnThreads) knn_reg_kernel(const float* __restrict__ ref, int ref_nb,
                                                              const float* __restrict__ query, int query_nb, int k,
                                                              float* __restrict__ out_dist, int* __restrict__ out_index) {
    __shared__ float tile[DIM][kTile];
    const int q0 = (blockIdx.x * kKnnThreads + threadIdx.x);
    const int stride_q = gridDim.x * kKnnThreads;
    float qv[QPT][DIM];
    float worst[QPT];
    RegList<KMAX> best[QPT];
    bool live[QPT];
#pragma unroll
    for (int i = 0; i < QPT; ++i) {
        const int q = q0 + i * stride_q;
        live[i] = q < query_nb;
        worst[i] = FLT_MAX;
#pragma unroll
        for (int s = 0; s < KMAX; ++s) { best[i].d[s] = FLT_MAX; best[i].i[s] = 0; }
#pragma unroll
        for (int d = 0; d < DIM; ++d) qv[i][d] = live[i] ? __ldg(query + (size_t)d * query_nb + q) : 0.f;
    }
    for (int base = 0; base < ref_nb; base += kTile) {
        const int n = min(kTile, ref_nb - base);
        __syncthreads();
        for (int t = threadIdx.x; t < n; t += kKnnThreads)
#pragma unroll
            for (int d = 0; d < DIM; ++d) tile[d][t] = __ldg(ref + (size_t)d * ref_nb + base + t);
        __syncthreads();
        for (int t = 0; t < n; ++t) {
            float r[DIM];
#pragma unroll
            for (int d = 0; d < DIM; ++d) r[d] = tile[d][t];
#pragma unroll
            for (int i = 0; i < QPT; ++i) {
                float ssd = 0.f;
#pragma unroll
                for (int d = 0; d < DIM; ++d) {
                    const float diff = __fsub_rn(r[d], qv[i][d]);
                    ssd = __fmaf_rn(diff, diff, ssd);
                }
                if (live[i] && !(best[i].count == k && ssd >= worst[i])) {
                    best[i].insert(k, ssd, base + t);
                    if (best[i].count == k) worst[i] = best[i].worst(k);
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < QPT; ++i) {
        const int q = q0 + i * stride_q;
        if (!live[i]) continue;
#pragma unroll
        for (int s = 0; s < KMAX; ++s) {
            if (s >= k) break;
            out_dist[(size_t)s * query_nb + q] = __fsqrt_rn(best[i].d[s]);
            out_index[(size_t)s * query_nb + q] = best[i].i[s];
        }
    }
}

template <int DIM>
static bool launch_knn_reg(const float* d_ref, int ref_nb, const float* d_query, int query_nb, int k, float* d_dist,
                           int* d_index, int sm_count, cudaStream_t stream) {
    auto blocks_for = [&](int qpt) {
        int blocks = (query_nb + kKnnThreads * qpt - 1) / (kKnnThreads * qpt);
        const int min_blocks = min((query_nb + kKnnThreads - 1) / kKnnThreads, 2 * sm_count);
        return max(blocks, min_blocks);
    };
    // Queries per thread: more of them share each broadcast LDS, fewer keep more warps resident.  Measured
    // (tools/knn_probe.py): one per thread wins until the queries alone fill the machine several times over
    // (k = 1: 4096 x 262144 0.86 ms vs 1.06 ms with four; 813 x 3M 1.81 vs 1.75 ms; k = 4, 65536 queries: 1.27 vs 1.86 ms).
    if (k == 1 && query_nb >= (1 << 20)) knn_reg_kernel<DIM, 1, 4><<<blocks_for(4), kKnnThreads, 0, stream>>>(d_ref, ref_nb, d_query, query_nb, k, d_dist, d_index);
    else if (k == 1) knn_reg_kernel<DIM, 1, 1><<<blocks_for(1), kKnnThreads, 0, stream>>>(d_ref, ref_nb, d_query, query_nb, k, d_dist, d_index);
    else if (k <= 4) knn_reg_kernel<DIM, 4, 1><<<blocks_for(1), kKnnThreads, 0, stream>>>(d_ref, ref_nb, d_query, query_nb, k, d_dist, d_index);
    else if (k <= 8) knn_reg_kernel<DIM, 8, 1><<<blocks_for(1), kKnnThreads, 0, stream>>>(d_ref, ref_nb, d_query, query_nb, k, d_dist, d_index);
    else if (k <= 16) knn_reg_kernel<DIM, 16, 1><<<blocks_for(1), kKnnThreads, 0, stream>>>(d_ref, ref_nb, d_query, query_nb, k, d_dist, d_index);
    else return false;
    return true;
}

// any dimension: one query per thread, its coordinates re-read through L1 (dim x query_nb layout is coalesced)
__global__ void __launch_bounds__(kKnnThreads) knn_generic_kernel(const float* __restrict__ ref, int ref_nb,
                                                                  const float* __restrict__ query, int query_nb, int dim,
                                                                  int k, float* __restrict__ out_dist,
                                                                  int* __restrict__ out_index) {
    const int q = blockIdx.x * kKnnThreads + threadIdx.x;
    if (q >= query_nb) return;
    float worst = FLT_MAX;
    int count = 0;
    for (int j = 0; j < ref_nb; ++j) {
        float ssd = 0.f;
        for (int d = 0; d < dim; ++d) {
            const float diff = __fsub_rn(__ldg(ref + (size_t)d * ref_nb + j), __ldg(query + (size_t)d * query_nb + q));
            ssd = __fmaf_rn(diff, diff, ssd);
        }
        if (!(count == k && ssd >= worst)) knn_insert(out_dist + q, out_index + q, (size_t)query_nb, k, count, worst, ssd, j);
    }
    for (int j = 0; j < k; ++j) out_dist[(size_t)j * query_nb + q] = __fsqrt_rn(out_dist[(size_t)j * query_nb + q]);
}

}  // namespace

cudaError_t launch_knn(const float* d_ref, int ref_nb, const float* d_query, int query_nb, int dim, int k, float* d_dist,
                       int* d_index, int sm_count, cudaStream_t stream) {
    if (query_nb <= 0) return cudaSuccess;
    if (dim >= 1 && dim <= 4 && k <= 16) {
        bool done = false;
        switch (dim) {
            case 1: done = launch_knn_reg<1>(d_ref, ref_nb, d_query, query_nb, k, d_dist, d_index, sm_count, stream); break;
            case 2: done = launch_knn_reg<2>(d_ref, ref_nb, d_query, query_nb, k, d_dist, d_index, sm_count, stream); break;
            case 3: done = launch_knn_reg<3>(d_ref, ref_nb, d_query, query_nb, k, d_dist, d_index, sm_count, stream); break;
            default: done = launch_knn_reg<4>(d_ref, ref_nb, d_query, query_nb, k, d_dist, d_index, sm_count, stream); break;
        }
        if (done) return cudaGetLastError();
    }
    if (dim >= 1 && dim <= 4) {
        // enough CTAs to cover the queries kQpt per thread, but at least ~2 waves of the SMs when queries are few
        int blocks = (query_nb + kKnnThreads * kQpt - 1) / (kKnnThreads * kQpt);
        const int min_blocks = min((query_nb + kKnnThreads - 1) / kKnnThreads, 2 * sm_count);
        blocks = max(blocks, min_blocks);
        switch (dim) {
            case 1: knn_fixed_kernel<1><<<blocks, kKnnThreads, 0, stream>>>(d_ref, ref_nb, d_query, query_nb, k, d_dist, d_index); break;
            case 2: knn_fixed_kernel<2><<<blocks, kKnnThreads, 0, stream>>>(d_ref, ref_nb, d_query, query_nb, k, d_dist, d_index); break;
            case 3: knn_fixed_kernel<3><<<blocks, kKnnThreads, 0, stream>>>(d_ref, ref_nb, d_query, query_nb, k, d_dist, d_index); break;
            default: knn_fixed_kernel<4><<<blocks, kKnnThreads, 0, stream>>>(d_ref, ref_nb, d_query, query_nb, k, d_dist, d_index); break;
        }
    } else {
        knn_generic_kernel<<<(query_nb + kKnnThreads - 1) / kKnnThreads, kKnnThreads, 0, stream>>>(d_ref, ref_nb, d_query, query_nb, dim, k, d_dist, d_index);
    }
    return cudaGetLastError();
}

}  // namespace perch
