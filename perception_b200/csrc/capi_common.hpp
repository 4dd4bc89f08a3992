// Helpers shared by the translation units that implement the C-ABI (capi.cu, capi_states.cu).
#pragma once
#include <cuda_runtime.h>

#include "context.hpp"
#include "kernels.cuh"
#include "perch_b200.h"

namespace perch_detail {

// records the message perch_last_error() returns for the calling thread and passes `code` through
int fail(int code, const char* fmt, ...);

// The pose batch of one perch_render_cost call: host pointers, or device pointers when an earlier stage (state
// assembly) produced them on the device.
struct PoseInputs {
    const float* poses16;
    const int* model_map;
    const int* label;
    const float* obs_total;
    bool on_device;
};
int render_cost_impl(perch_ctx* c, const char* stage, const PoseInputs& in, int n_poses, const perch_params* p,
                     perch_outputs* out);

// Bump allocator over the context scratch buffer (256-byte aligned slices).
struct Scratch {
    perch_ctx* c;
    size_t off = 0;
    explicit Scratch(perch_ctx* ctx) : c(ctx) {}
    static size_t align(size_t b) { return (b + 255) & ~(size_t)255; }
    template <class T>
    T* take(size_t n) {
        T* p = reinterpret_cast<T*>(static_cast<char*>(c->d_scratch) + off);
        off += align(n * sizeof(T));
        return p;
    }
};
int ensure_scratch(perch_ctx* c, size_t bytes);  // per-call device scratch, grown on demand (capi.cu)
int ensure_pinned(perch_ctx* c, size_t bytes);   // the context's pinned host staging buffer (capi.cu)
// readiness checks + grid (re)build + PoseArgs of the fused cost kernel; a.pose_label must be set by the caller first
int prepare_fused_cost(perch_ctx* c, const perch_params* p, perch::PoseArgs& a, bool& color);
int report_err_flag(int flag);  // d_err_flag value -> status + message
void destroy_comm(perch_ctx* c);  // capi_multi.cu
int ensure_obs_host(perch_ctx* c);  // host mirrors of the label-sorted observed cloud, fetched on demand (capi.cu)
int install_observed_device(perch_ctx* c, const float* d_aos, const uint8_t* d_rgb, size_t rgb_stride, const int* d_label,
                            int n);  // perch_set_observed for a device-resident cloud (capi.cu)

void destroy_states(perch_ctx* c);  // capi_states.cu

struct DeviceGuard {
    int prev = 0;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        cudaSetDevice(dev);
    }
    ~DeviceGuard() { cudaSetDevice(prev); }
};

}  // namespace perch_detail

#define PERCH_CU(call)                                                                                              \
    do {                                                                                                            \
        cudaError_t e__ = (call);                                                                                   \
        if (e__ != cudaSuccess)                                                                                     \
            return perch_detail::fail(PERCH_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, \
                                      __LINE__);                                                                    \
    } while (0)
