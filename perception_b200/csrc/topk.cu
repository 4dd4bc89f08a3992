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
//                   the keys, 8 bits each, warp-aggregated shared-memory histogram, the bin found by one warp), then one pass that writes the (at
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
constexpr int kSelCache = 16;  // keys per thread kept in registers by the select kernel
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
    __shared__ unsigned s_and[2], s_or[2];
    __shared__ int s_remaining, s_take_all, s_count;
    const int m = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const int k = a.k;
    // Up to kSelCache * kSelThreads keys (16 384: the headline batch) stay in registers for all passes, loaded with
    // every request in flight at once; longer lists are re-read from L2 each pass, four loads in flight per thread.
    const bool cached = a.n_poses <= kSelCache * kSelThreads;
    unsigned long long kreg[kSelCache];
    if (cached) {
#pragma unroll
        for (int u = 0; u < kSelCache; ++u) {
            const int i = u * kSelThreads + tid;
            kreg[u] = i < a.n_poses ? a.keys[i] : kInvalidKey;
        }
    }
    // f(key) is called by every thread of the CTA the same number of times (it may use warp collectives)
    auto for_each_key = [&](auto&& f) {
        if (cached) {
#pragma unroll
            for (int u = 0; u < kSelCache; ++u) f(kreg[u]);
        } else {
            for (int i0 = 0; i0 < a.n_poses; i0 += 4 * kSelThreads) {
                unsigned long long key[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int i = i0 + u * kSelThreads + tid;
                    key[u] = i < a.n_poses ? a.keys[i] : kInvalidKey;
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) f(key[u]);
            }
        }
    };
    // Pre-pass: AND and OR over the model's keys.  A byte on which they agree is shared by every key of the model (the
    // high bytes of costs that span a small range, the high index byte of a short list): its radix pass is skipped.
    if (tid == 0) { s_and[0] = s_and[1] = 0xFFFFFFFFu; s_or[0] = s_or[1] = 0u; }
    __syncthreads();
    {
        unsigned long long k_and = ~0ull, k_or = 0ull;
        for_each_key([&](unsigned long long key) {
            if ((key >> 56) == (unsigned long long)m) { k_and &= key; k_or |= key; }
        });
        const unsigned a0 = __reduce_and_sync(0xffffffffu, (unsigned)k_and), a1 = __reduce_and_sync(0xffffffffu, (unsigned)(k_and >> 32));
        const unsigned o0 = __reduce_or_sync(0xffffffffu, (unsigned)k_or), o1 = __reduce_or_sync(0xffffffffu, (unsigned)(k_or >> 32));
        if (lane == 0) {
            atomicAnd(&s_and[0], a0); atomicAnd(&s_and[1], a1);
            atomicOr(&s_or[0], o0); atomicOr(&s_or[1], o1);
        }
    }
    __syncthreads();
    const unsigned long long all_and = ((unsigned long long)s_and[1] << 32) | s_and[0];
    const unsigned long long shared_bits = ~(all_and ^ (((unsigned long long)s_or[1] << 32) | s_or[0]));
    unsigned long long known = (unsigned long long)m << 56, mask = 0xFFull << 56;
    int remaining = k;
    bool take_all = false;
    for (int byte = 6; byte >= 0 && !take_all; --byte) {
        const int shift = 8 * byte;
        if (((shared_bits >> shift) & 0xFFull) == 0xFFull) {  // (no key of this model at all: AND = ~0, OR = 0, nothing shared)
            known |= all_and & (0xFFull << shift);
            mask |= 0xFFull << shift;
            continue;
        }
        for (int b = tid; b < 256; b += kSelThreads) hist[b] = 0;
        __syncthreads();
        for_each_key([&](unsigned long long key) {
            const bool in = (key & mask) == known;  // kInvalidKey never matches (its top byte is 0xFF > any model)
            const unsigned bin = (unsigned)(key >> shift) & 255u;
            // keys of one model share their high bytes: aggregate equal bins inside the warp before the atomic
            const unsigned active = __ballot_sync(0xffffffffu, in);
            if (in) {
                const unsigned peers = __match_any_sync(active, bin);
                if (lane == (__ffs(peers) - 1)) atomicAdd(&hist[bin], (unsigned)__popc(peers));
            }
        });
        __syncthreads();
        if (tid < 32) {
            // bin holding the `remaining`-th smallest key: lane l owns bins 8l..8l+7, exclusive prefix by shuffles
            unsigned h[8], sum = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) { h[j] = hist[8 * lane + j]; sum += h[j]; }
            unsigned incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            unsigned cum = incl - sum;
            const bool mine = cum < (unsigned)remaining && incl >= (unsigned)remaining;  // exactly one lane, if any
            if (mine) {
                int b = 0;
#pragma unroll
                for (; b < 7; ++b) {
                    if (cum + h[b] >= (unsigned)remaining) break;
                    cum += h[b];
                }
                s_take_all = 0;
                s_known = known | ((unsigned long long)(8 * lane + b) << shift);
                s_remaining = remaining - (int)cum;
            }
            // fewer than `remaining` keys under this prefix: the model has fewer than k valid poses
            if (__ballot_sync(0xffffffffu, mine) == 0u && lane == 0) s_take_all = 1;
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
    for_each_key([&](unsigned long long key) {
        if (key < lo || key > hi) return;
        const int slot = atomicAdd(&s_count, 1);
        if (slot >= k) return;  // cannot happen: keys are unique
        const int idx = (int)(key & 0xFFFFFFull);
        perch_hit h;
        h.cost = (int)((unsigned)(key >> 24) ^ 0x80000000u);
        h.pose_id = a.pose_id_offset + idx;
        h.rendered_cost = a.costs[idx];
        h.observed_cost = a.costs[a.n_poses + idx];
        out[slot] = h;
    });
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
