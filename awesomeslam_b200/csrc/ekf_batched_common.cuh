// Pieces shared by the two batched EKF kernels (ekf_batched.cu: thread-owns-columns DFMA kernel for any padded
// dimension; tools/probes/ekf_batched_mma.cu: the measured-and-rejected warp-per-filter DMMA variant, not built): kernel parameters, the
// shared-memory message buffer and its cp.async prefetch.
#pragma once
#include <cuda_runtime.h>

#include "frontend.cuh"
#include "internal.hpp"

namespace aslam {

// One message of one filter, staged in shared memory by cp.async one step ahead of its use.
struct __align__(16) MsgBuf {
    double ox, oy, vx, az;
    float yaw, dt;
    uint32_t base, cnt;
    float rng[ASLAM_MAX_OBS];
    float brg[ASLAM_MAX_OBS];
};

__device__ __forceinline__ void cp_async4(void* dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src)
                 : "memory");
}
__device__ __forceinline__ void cp_async8(void* dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src)
                 : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

struct EkfParams {
    BatchState st;
    BatchInput in;
    FrontParams fp;
    double kq, kr_pose, kr_lm, kp_lm;
};

// Start the asynchronous copy of message `step` of filter f into m: the six odometry fields by six
// threads, the observation list spread over the filter's threads. No register is held across the wait.
template <int TPF>
__device__ __forceinline__ void prefetch_message(MsgBuf& m, const BatchInput& in, uint32_t batch, uint32_t f,
                                                 uint32_t step, int ct) {
    const size_t si = (size_t)step * batch + f;
    uint32_t base, cnt;
    if (in.offsets) {
        base = in.offsets[f];
        cnt = in.offsets[f + 1] - base;
    } else {
        base = (uint32_t)(si * in.obs_per_filter);
        cnt = in.obs_per_filter;
    }
    if (ct == 6) {
        m.base = base;
        m.cnt = cnt;
    }
    if (cnt > ASLAM_MAX_OBS) cnt = ASLAM_MAX_OBS;
    if (ct < 4) {
        const double* src = ct == 0 ? in.ox : ct == 1 ? in.oy : ct == 2 ? in.vx : in.az;
        cp_async8(&m.ox + ct, src + si);
    } else if (ct < 6) {
        const float* src = ct == 4 ? in.yaw : in.dt;
        cp_async4(&m.yaw + (ct - 4), src + si);
    }
    for (uint32_t o = ct; o < cnt; o += TPF) {
        cp_async4(&m.rng[o], in.range + base + o);
        cp_async4(&m.brg[o], in.bearing + base + o);
    }
}


}  // namespace aslam
