// Internal definitions shared by the C-ABI layer and the kernels of libaslam_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <string>

#include "../../include/aslam_c.h"

namespace aslam {

void set_error(const std::string& msg);
bool cuda_ok(cudaError_t e, const char* what);
#define ASLAM_CUDA(call)                                  \
    do {                                                  \
        if (!::aslam::cuda_ok((call), #call)) return ASLAM_ERR_CUDA; \
    } while (0)

// cudaFuncSetAttribute applies to the current device only: a process-wide "already set" flag would leave the second
// device of a process that holds handles on several GPUs without the shared-memory opt-in. One bit per device.
struct PerDeviceOnce {
    std::atomic<uint64_t> done{0};
    bool pending(int* dev) {
        if (cudaGetDevice(dev) != cudaSuccess || *dev < 0 || *dev >= 64) return true;
        return ((done.load(std::memory_order_acquire) >> *dev) & 1ull) == 0;
    }
    void mark(int dev) {
        if (dev >= 0 && dev < 64) done.fetch_or(1ull << dev, std::memory_order_release);
    }
};
#define ASLAM_DYN_SMEM(kernel, bytes)                                                                              \
    do {                                                                                                           \
        static ::aslam::PerDeviceOnce once_;                                                                       \
        int dev_ = -1;                                                                                             \
        if (once_.pending(&dev_)) {                                                                                \
            ASLAM_CUDA(cudaFuncSetAttribute((kernel), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes))); \
            once_.mark(dev_);                                                                                      \
        }                                                                                                          \
    } while (0)

// Device-resident state of a batch of filters. Layout in HBM (ld = padded state dim NP):
//   X, Z   [batch][ld]           doubles, entries >= N are zero
//   P      [batch][ld][ld]       P(i,j) at i*ld + j: for a fixed row the ld columns are contiguous,
//                                so the ld/CPT threads of a filter (one or two columns each)
//                                read one row with a single coalesced 16-byte-vector access
//   A      [batch][2]            EKF: A(0,0), A(1,0)
//   N, status, flags [batch]     flags bit0: init_x pending (ekf.cpp:87-91), bit1: init_z pending
//   wait-list: wcount [batch]; wrange, wbearing, whits [batch][ASLAM_MAX_WAIT]
struct BatchState {
    uint32_t batch = 0;
    uint32_t ld = 0;
    double* X = nullptr;
    double* Z = nullptr;
    double* P = nullptr;
    double* A = nullptr;
    uint32_t* N = nullptr;
    uint32_t* status = nullptr;
    uint32_t* flags = nullptr;
    uint32_t* wcount = nullptr;
    float* wrange = nullptr;
    float* wbearing = nullptr;
    uint32_t* whits = nullptr;
    int32_t* llt_fail = nullptr;  // UKF: first failing pivot column of the last step, or -1
};

enum : uint32_t { FLAG_INIT_X = 1u, FLAG_INIT_Z = 2u };

// One replayed message per filter (or `steps` of them, step-major).
struct BatchInput {
    const double* ox = nullptr;
    const double* oy = nullptr;
    const float* yaw = nullptr;
    const double* vx = nullptr;
    const double* az = nullptr;
    const float* dt = nullptr;
    const float* range = nullptr;
    const float* bearing = nullptr;
    const uint32_t* offsets = nullptr;  // [batch+1] or null when obs_per_filter is used
    uint32_t obs_per_filter = 0;
    int32_t* assoc = nullptr;           // [total observations of the LAST step]
    uint32_t steps = 1;
};

aslam_status launch_ekf_batched(const aslam_config& cfg, const BatchState& st, const BatchInput& in,
                                cudaStream_t stream);
aslam_status launch_ukf_batched(const aslam_config& cfg, const BatchState& st, const BatchInput& in,
                                cudaStream_t stream);
// Padded state dimension the batched kernels use for a given cap, 0 when unsupported.
uint32_t batched_padded_dim(int32_t max_state_dim, int kind);
aslam_status launch_batch_init(const aslam_config& cfg, int kind, const BatchState& st, cudaStream_t stream);

}  // namespace aslam
