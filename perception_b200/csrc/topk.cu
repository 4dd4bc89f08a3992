// Per-model top-k of a scored pose batch on the device: the caller's argmin (search_env.cpp:2016-2036, 2542-2583),
// generalised to the k best per model, so that only k x 16 bytes per model leave a GPU (and cross NVLink in the NCCL
// all-gather) instead of the cost arrays.  Same rule as the host-side perch_topk (capi.cu):
//   cost = (int)(rendered + observed); a candidate is skipped when (int)rendered < 0, when either cost is NaN / beyond
//   +-2e9, when cost is -1 or -2 ("Invalid successor", search_env.cpp:2555) or when |(int)rendered - (int)observed| >= 30
//   (:2560-2561); order = (cost, pose id) ascending, i.e. the first of equal costs wins like the reference's strict `<`.
//
// topk_key_kernel   one 64-bit key per pose:  model (8 bits) | cost biased to unsigned (32) | local pose index (24);
//                   skipped poses get the all-ones key.  Keys are unique, and the k best of model m are the k smallest
//                   keys carrying m in their top byte.
// topk_select_kernel  one CTA per model: MSB-first radix select of the k-th smallest key of that model (7 passes over
//                   the keys, 8 bits each, warp-aggregated shared-memory histogram), then one pass that writes the (at
//                   most k) keys up to it as perch_hit records, in no particular order (the merge after the gather
//                   orders them).  Keys are 8 B per pose and L2-resident; the whole thing is microseconds beside the
//                   fused kernel.
#include <cuda_runtime.h>

#include <climits>

#include "kernels.cuh"
#include "perch_b200.h"

namespace perch {
namespace {

constexpr int kSelThreads = 1024;
constexpr unsigned long long kInvalidKey = ~0ull;

__device__ __forceinline__ unsigned long long hit_key(float r, float o, int model, int n_models, int idx) {
    if (model < 0 || model >= n_models) return kInvalidKey;
    if (!(r == r) || !(o == o) || !(fabsf(r) < 2e9f) || !(fabsf(o) < 2e9f)) return kInvalidKey;
    const int ri = (int)r;
    if (ri < 0) return kInvalidKey;                       // search_env.cpp:2021-2027
    const int cost = (int)__fadd_rn(r, o);
    if (cost == -1 || cost == -2) return kInvalidKey;     // :2555
    if (!(abs(ri - (int)o) < 30)) return kInvalidKey;     // :2560-2561
    return ((unsigned long long)model << 56) | ((unsigned long long)((unsigned)cost ^ 0x80000000u) << 24) |
           (unsigned long long)idx;
}

__global__ void topk_key_kernel(TopkArgs a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n_poses) return;
    a.keys[i] = hit_key(a.costs[i], a.costs[a.n_poses + i], a.pose_model_map[i], a.n_models, i);
}

__global__ void __launch_bounds__(kSelThreads) topk_select_kernel(TopkArgs a) {
    __shared__ unsigned hist[256];
    __shared__ unsigned long long s_known;
    __shared__ int s_remaining, s_take_all, s_count;
    const int m = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const int k = a.k;
    unsigned long long known = (unsigned long long)m << 56, mask = 0xFFull << 56;
    int remaining = k;
    bool take_all = false;
    for (int byte = 6; byte >= 0 && !take_all; --byte) {
        for (int b = tid; b < 256; b += kSelThreads) hist[b] = 0;
        __syncthreads();
        const int shift = 8 * byte;
        for (int i0 = 0; i0 < a.n_poses; i0 += kSelThreads) {
            const int i = i0 + tid;
            const unsigned long long key = i < a.n_poses ? a.keys[i] : kInvalidKey;
            const bool in = (key & mask) == known;  // kInvalidKey never matches (its top byte is 0xFF > any model)
            const unsigned bin = (unsigned)(key >> shift) & 255u;
            // keys of one model share their high bytes: aggregate equal bins inside the warp before the atomic
            const unsigned active = __ballot_sync(0xffffffffu, in);
            if (in) {
                const unsigned peers = __match_any_sync(active, bin);
                if (lane == (__ffs(peers) - 1)) atomicAdd(&hist[bin], (unsigned)__popc(peers));
            }
        }
        __syncthreads();
        if (tid == 0) {
            unsigned cum = 0;
            int b = 0;
            for (; b < 256; ++b) {
                if (cum + hist[b] >= (unsigned)remaining) break;
                cum += hist[b];
            }
            if (b == 256) {
                s_take_all = 1;  // fewer than `remaining` keys under this prefix: only possible on the first pass
            } else {
                s_take_all = 0;
                s_known = known | ((unsigned long long)b << shift);
                s_remaining = remaining - (int)cum;
            }
        }
        __syncthreads();
        take_all = s_take_all != 0;
        if (!take_all) {
            known = s_known;
            remaining = s_remaining;
            mask |= 0xFFull << shift;
        }
        __syncthreads();
    }
    // everything of model m up to the k-th smallest key (or all of it when the model has fewer than k valid poses)
    const unsigned long long lo = (unsigned long long)m << 56;
    const unsigned long long hi = take_all ? (lo | 0x00FFFFFFFFFFFFFFull) : known;
    if (tid == 0) s_count = 0;
    __syncthreads();
    perch_hit* out = a.hits + (size_t)m * k;
    for (int i = tid; i < a.n_poses; i += kSelThreads) {
        const unsigned long long key = a.keys[i];
        if (key < lo || key > hi) continue;
        const int slot = atomicAdd(&s_count, 1);
        if (slot >= k) continue;  // cannot happen: keys are unique
        const int idx = (int)(key & 0xFFFFFFull);
        perch_hit h;
        h.cost = (int)((unsigned)(key >> 24) ^ 0x80000000u);
        h.pose_id = a.pose_id_offset + idx;
        h.rendered_cost = a.costs[idx];
        h.observed_cost = a.costs[a.n_poses + idx];
        out[slot] = h;
    }
    __syncthreads();
    for (int slot = min(s_count, k) + tid; slot < k; slot += kSelThreads) {
        perch_hit h;
        h.cost = INT_MAX; h.pose_id = -1; h.rendered_cost = 0.f; h.observed_cost = 0.f;
        out[slot] = h;
    }
}

}  // namespace

cudaError_t launch_topk(const TopkArgs& a, cudaStream_t stream) {
    if (a.n_models <= 0 || a.k <= 0) return cudaSuccess;
    if (a.n_poses > 0) topk_key_kernel<<<(a.n_poses + 255) / 256, 256, 0, stream>>>(a);
    topk_select_kernel<<<a.n_models, kSelThreads, 0, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace perch
