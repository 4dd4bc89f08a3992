// Helpers shared by the translation units that implement the C-ABI (capi.cu, capi_states.cu).
#pragma once
#include <cuda_runtime.h>

#include "context.hpp"
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
