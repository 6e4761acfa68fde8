// C-ABI layer for the batched filters (include/aslam_c.h): handle lifetime, device memory,
// host<->device staging, state export. The filter math lives in ekf_batched.cu /
// ukf_batched.cu; nothing here computes on the CPU.
#include <cuda_runtime.h>

#include <cmath>
#include <algorithm>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include "internal.hpp"
#include "nvtx_range.hpp"

namespace aslam {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
bool cuda_ok(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return true;
    g_error = std::string(what) + ": " + cudaGetErrorString(e);
    return false;
}

uint32_t batched_padded_dim(int32_t max_state_dim, int kind) {
    // N is odd and must stay < max_state_dim (ekf.cpp:263), so N <= max_state_dim - 1.
    const int nmax = max_state_dim - 1;
    if (nmax < 3) return 0;
    if (kind == ASLAM_KIND_UKF) {
        // the kernel's shared-memory footprint (and with it the number of resident filters per SM, which is what sets its
        // throughput) follows the cap the handle asks for: 20 (<= 8 landmarks: 7 per SM), 24, 30, 36
        if (nmax <= 19) return 20;
        if (nmax <= 23) return 24;
        if (nmax <= 29) return 30;  // the reference's cap (config.h:45)
        if (nmax <= 35) return 36;  // the cfg3 shape: 16 landmarks, N = 35, 75 sigma points
        return 0;                   // larger UKF maps: aslam_large_ukf_*
    }
    if (nmax <= 29) return 30;
    if (nmax <= 35) return 36;
    if (nmax <= 47) return 48;
    if (nmax <= 63) return 64;
    return 0;
}

// EKFSlam::initialize (ekf.cpp:49-71) / UKFSlam::initialize (ukf.cpp:49-67) for every filter.
__global__ void batch_init_kernel(BatchState st, double kp_pose) {
    const uint32_t f = blockIdx.x;
    const uint32_t ld = st.ld;
    for (uint32_t k = threadIdx.x; k < ld * ld; k += blockDim.x) {
        const uint32_t i = k / ld, j = k % ld;
        st.P[(size_t)f * ld * ld + k] = (i == j && i < 3) ? kp_pose : 0.0;
    }
    for (uint32_t k = threadIdx.x; k < ld; k += blockDim.x) {
        st.X[(size_t)f * ld + k] = 0.0;
        st.Z[(size_t)f * ld + k] = 0.0;
    }
    for (uint32_t k = threadIdx.x; k < ASLAM_MAX_WAIT; k += blockDim.x) {
        st.wrange[(size_t)f * ASLAM_MAX_WAIT + k] = 0.f;
        st.wbearing[(size_t)f * ASLAM_MAX_WAIT + k] = 0.f;
        st.whits[(size_t)f * ASLAM_MAX_WAIT + k] = 0;
    }
    if (threadIdx.x == 0) {
        st.N[f] = 3;
        st.status[f] = 0;
        st.flags[f] = FLAG_INIT_X;
        st.wcount[f] = 0;
        st.A[2 * (size_t)f] = 1.0;
        st.A[2 * (size_t)f + 1] = 0.0;
        st.llt_fail[f] = -1;
    }
}

aslam_status launch_batch_init(const aslam_config& cfg, int kind, const BatchState& st, cudaStream_t stream) {
    const double kp = kind == ASLAM_KIND_EKF ? (double)cfg.ekf_kp_robot_pose : (double)cfg.ukf_kp_robot_pose;
    batch_init_kernel<<<st.batch, 128, 0, stream>>>(st, kp);
    ASLAM_CUDA(cudaGetLastError());
    return ASLAM_OK;
}

// out[f][j*ld + i] = in[f][i*ld + j]: device row-major P -> column-major export (and back).
__global__ void batch_transpose_kernel(const double* in, double* out, uint32_t ld) {
    const size_t base = (size_t)blockIdx.x * ld * ld;
    for (uint32_t k = threadIdx.x; k < ld * ld; k += blockDim.x) {
        const uint32_t i = k / ld, j = k % ld;
        out[base + (size_t)j * ld + i] = in[base + k];
    }
}

__global__ void batch_landmarks_kernel(BatchState st, double* lx, double* ly, uint32_t* count) {
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= st.batch) return;
    const uint32_t lmax = (st.ld - 3) / 2;
    const uint32_t N = st.N[f];
    const uint32_t L = (N - 3) / 2;
    for (uint32_t l = 0; l < lmax; ++l) {  // common.h:98-102
        lx[(size_t)f * lmax + l] = l < L ? st.X[(size_t)f * st.ld + 3 + 2 * l] : 0.0;
        ly[(size_t)f * lmax + l] = l < L ? st.X[(size_t)f * st.ld + 4 + 2 * l] : 0.0;
    }
    count[f] = L;
}

// FP64 FMA-pipe probe: 8 independent DFMA chains per thread, enough threads to fill the chip.
__global__ void fp64_probe_kernel(double* out, int iters) {
    double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6,
           a7 = a0 + 7;
    const double m = 0.999999, c = 1e-7;
    for (int i = 0; i < iters; ++i) {
        a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
        a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
    if (a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 == 12345.678) out[0] = a0;
}

}  // namespace aslam

using namespace aslam;

struct aslam_filter {
    int kind = 0;
    aslam_config cfg;
    int device = 0;
    cudaStream_t stream = nullptr;
    BatchState st;
    // device staging for host-fed messages
    double* d_odom64 = nullptr;  // [4][batch]: x, y, vx, az
    float* d_odom32 = nullptr;   // [2][batch]: yaw, dt
    float* d_range = nullptr;
    float* d_bearing = nullptr;
    uint32_t* d_offsets = nullptr;
    size_t obs_capacity = 0;
    // current observation list (own staging or caller's device arrays)
    const float* cur_range = nullptr;
    const float* cur_bearing = nullptr;
    const uint32_t* cur_offsets = nullptr;
    uint32_t cur_total = 0;
    bool have_obs = false;
    int32_t* d_assoc = nullptr;
    size_t assoc_capacity = 0;
    double* d_scratch = nullptr;  // export scratch, lazily sized
    size_t scratch_bytes = 0;
    // run_host staging: two message blocks filled on `copy_stream` so that the host->device copy of
    // call k+1 overlaps the kernel of call k (ev_copied / ev_consumed order the two streams)
    unsigned char* d_replay[2] = {nullptr, nullptr};
    size_t replay_bytes[2] = {0, 0};
    int replay_next = 0;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copied[2] = {nullptr, nullptr};
    cudaEvent_t ev_consumed[2] = {nullptr, nullptr};
    // read_state_async: X is snapshotted device-side on `stream`, then copied out on `out_stream`
    double* d_xsnap[2] = {nullptr, nullptr};
    cudaStream_t out_stream = nullptr;
    cudaEvent_t ev_snap[2] = {nullptr, nullptr};
    cudaEvent_t ev_read[2] = {nullptr, nullptr};
    uint64_t read_seq = 0;
    uint64_t launches = 0;  // filter-step kernels launched so far
};

namespace {

template <typename T>
aslam_status dev_alloc(T** p, size_t count) {
    ASLAM_CUDA(cudaMalloc((void**)p, count * sizeof(T)));
    return ASLAM_OK;
}

aslam_status ensure_scratch(aslam_filter* h, size_t bytes) {
    if (h->scratch_bytes >= bytes) return ASLAM_OK;
    if (h->d_scratch) cudaFree(h->d_scratch);
    h->d_scratch = nullptr;
    h->scratch_bytes = 0;
    ASLAM_CUDA(cudaMalloc((void**)&h->d_scratch, bytes));
    h->scratch_bytes = bytes;
    return ASLAM_OK;
}

aslam_status ensure_assoc(aslam_filter* h, size_t count) {
    if (h->assoc_capacity >= count) return ASLAM_OK;
    if (h->d_assoc) cudaFree(h->d_assoc);
    h->d_assoc = nullptr;
    h->assoc_capacity = 0;
    ASLAM_CUDA(cudaMalloc((void**)&h->d_assoc, count * sizeof(int32_t)));
    h->assoc_capacity = count;
    return ASLAM_OK;
}

aslam_status create_filter(int kind, const aslam_config* cfg, uint32_t batch, int32_t device, aslam_filter** out) {
    if (!out || batch == 0) {
        set_error("create: null out or empty batch");
        return ASLAM_ERR_ARG;
    }
    *out = nullptr;
    aslam_config c;
    if (cfg)
        c = *cfg;
    else
        aslam_config_default(&c);
    const uint32_t ld = batched_padded_dim(c.max_state_dim, kind);
    if (!ld) {
        set_error("create: max_state_dim outside the batched kernels' range (use aslam_large_*)");
        return ASLAM_ERR_UNSUPPORTED;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_error("create: no CUDA device (this library has no CPU path)");
        return ASLAM_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) {
        set_error("create: device index out of range");
        return ASLAM_ERR_ARG;
    }
    ASLAM_CUDA(cudaSetDevice(device));
    aslam_filter* h = new (std::nothrow) aslam_filter;
    if (!h) return ASLAM_ERR_ARG;
    h->kind = kind;
    h->cfg = c;
    h->device = device;
    h->st.batch = batch;
    h->st.ld = ld;
    aslam_status rc = ASLAM_OK;
    auto chk = [&](aslam_status s) {
        if (rc == ASLAM_OK) rc = s;
    };
    if (!cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate")) rc = ASLAM_ERR_CUDA;
    if (!cuda_ok(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking), "cudaStreamCreate")) rc = ASLAM_ERR_CUDA;
    if (!cuda_ok(cudaStreamCreateWithFlags(&h->out_stream, cudaStreamNonBlocking), "cudaStreamCreate")) rc = ASLAM_ERR_CUDA;
    for (int i = 0; i < 2; ++i) {
        cudaEvent_t* evs[] = {&h->ev_copied[i], &h->ev_consumed[i], &h->ev_snap[i], &h->ev_read[i]};
        for (cudaEvent_t* e : evs)
            if (!cuda_ok(cudaEventCreateWithFlags(e, cudaEventDisableTiming), "cudaEventCreate")) rc = ASLAM_ERR_CUDA;
    }
    chk(dev_alloc(&h->st.X, (size_t)batch * ld));
    chk(dev_alloc(&h->st.Z, (size_t)batch * ld));
    chk(dev_alloc(&h->st.P, (size_t)batch * ld * ld));
    chk(dev_alloc(&h->st.A, (size_t)batch * 2));
    chk(dev_alloc(&h->st.N, batch));
    chk(dev_alloc(&h->st.status, batch));
    chk(dev_alloc(&h->st.flags, batch));
    chk(dev_alloc(&h->st.wcount, batch));
    chk(dev_alloc(&h->st.wrange, (size_t)batch * ASLAM_MAX_WAIT));
    chk(dev_alloc(&h->st.wbearing, (size_t)batch * ASLAM_MAX_WAIT));
    chk(dev_alloc(&h->st.whits, (size_t)batch * ASLAM_MAX_WAIT));
    chk(dev_alloc(&h->st.llt_fail, batch));
    chk(dev_alloc(&h->d_odom64, (size_t)batch * 4));
    chk(dev_alloc(&h->d_odom32, (size_t)batch * 2));
    chk(dev_alloc(&h->d_offsets, (size_t)batch + 1));
    if (rc == ASLAM_OK) rc = launch_batch_init(c, kind, h->st, h->stream);
    if (rc != ASLAM_OK) {
        aslam_filter_destroy(h);
        return rc;
    }
    *out = h;
    return ASLAM_OK;
}

aslam_status run_kernel(aslam_filter* h, const BatchInput& in) {
    h->launches += 1;
    if (h->kind == ASLAM_KIND_EKF) return launch_ekf_batched(h->cfg, h->st, in, h->stream);
    return launch_ukf_batched(h->cfg, h->st, in, h->stream);
}

}  // namespace

extern "C" {

void aslam_config_default(aslam_config* c) {
    if (!c) return;
    c->pi = 3.141592654f;
    c->min_dist_thresh = 0.5f;
    c->min_landmark_occ = 10;
    c->max_state_dim = 30;
    c->ukf_std_a = 0.2f;
    c->ukf_std_yaw = 0.2f;
    c->ukf_kp_robot_pose = 0.001f;
    c->ukf_kp_landmark_pose = 1.0f;
    c->ukf_kr = 0.2f;
    c->ukf_kq = 0.001f;
    c->ekf_kp_robot_pose = 0.001f;
    c->ekf_kp_landmark_pose = 10000.0f;
    c->ekf_kr = 0.2f;
    c->ekf_kq = 0.001f;
}

const char* aslam_last_error(void) { return g_error.c_str(); }
const char* aslam_version(void) { return "aslam_b200 0.1 (sm_100a)"; }
int32_t aslam_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

/* Measured peak of the FP64 FMA pipe (TFLOP/s, 2 flop per DFMA): the denominator of the batched
 * kernels' compute roofline (MEASURED_PEAKS.json records none). */
aslam_status aslam_probe_fp64_tflops(int32_t device, double* tflops) {
    if (!tflops) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(device));
    int sms = 0;
    ASLAM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    double* d = nullptr;
    ASLAM_CUDA(cudaMalloc((void**)&d, 8));
    cudaEvent_t e0, e1;
    ASLAM_CUDA(cudaEventCreate(&e0));
    ASLAM_CUDA(cudaEventCreate(&e1));
    const int iters = 20000, blocks = sms * 8, threads = 256;
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {
        ASLAM_CUDA(cudaEventRecord(e0));
        fp64_probe_kernel<<<blocks, threads>>>(d, iters);
        ASLAM_CUDA(cudaEventRecord(e1));
        ASLAM_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        ASLAM_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double tf = 2.0 * 8.0 * iters * (double)blocks * threads / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    *tflops = best;
    return ASLAM_OK;
}

/* X only, into a caller buffer [batch][ld] (e.g. pinned memory); synchronises the stream. */
aslam_status aslam_filter_read_state(aslam_filter* h, double* X) {
    if (!h || !X) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    ASLAM_CUDA(cudaMemcpyAsync(X, h->st.X, (size_t)h->st.batch * h->st.ld * sizeof(double), cudaMemcpyDeviceToHost,
                               h->stream));
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}

/* Pipelined variant: X is snapshotted on the handle's stream (device to device, behind everything
 * enqueued so far) and copied to the caller's buffer on a second stream, so the next run_* call does
 * not wait for the PCIe transfer. *ticket identifies the read for aslam_filter_read_wait. */
aslam_status aslam_filter_read_state_async(aslam_filter* h, double* X, uint64_t* ticket) {
    if (!h || !X) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    const size_t bytes = (size_t)h->st.batch * h->st.ld * sizeof(double);
    const int s = (int)(h->read_seq & 1);
    if (!h->d_xsnap[s]) ASLAM_CUDA(cudaMalloc((void**)&h->d_xsnap[s], bytes));
    ASLAM_CUDA(cudaStreamWaitEvent(h->stream, h->ev_read[s], 0));  // snapshot s was last drained two reads ago
    ASLAM_CUDA(cudaMemcpyAsync(h->d_xsnap[s], h->st.X, bytes, cudaMemcpyDeviceToDevice, h->stream));
    ASLAM_CUDA(cudaEventRecord(h->ev_snap[s], h->stream));
    ASLAM_CUDA(cudaStreamWaitEvent(h->out_stream, h->ev_snap[s], 0));
    ASLAM_CUDA(cudaMemcpyAsync(X, h->d_xsnap[s], bytes, cudaMemcpyDeviceToHost, h->out_stream));
    ASLAM_CUDA(cudaEventRecord(h->ev_read[s], h->out_stream));
    if (ticket) *ticket = h->read_seq;
    h->read_seq += 1;
    return ASLAM_OK;
}

/* Blocks until the read with this ticket (and every earlier one) has landed in its host buffer. */
aslam_status aslam_filter_read_wait(aslam_filter* h, uint64_t ticket) {
    if (!h || ticket >= h->read_seq) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    ASLAM_CUDA(cudaEventSynchronize(h->ev_read[ticket & 1]));
    return ASLAM_OK;
}

aslam_status aslam_ekf_create(const aslam_config* cfg, uint32_t batch, int32_t device, aslam_filter** out) {
    return create_filter(ASLAM_KIND_EKF, cfg, batch, device, out);
}
aslam_status aslam_ukf_create(const aslam_config* cfg, uint32_t batch, int32_t device, aslam_filter** out) {
    return create_filter(ASLAM_KIND_UKF, cfg, batch, device, out);
}

aslam_status aslam_filter_destroy(aslam_filter* h) {
    if (!h) return ASLAM_OK;
    cudaSetDevice(h->device);
    cudaStream_t streams[] = {h->stream, h->copy_stream, h->out_stream};
    for (cudaStream_t st : streams)
        if (st) cudaStreamSynchronize(st);
    void* ptrs[] = {h->st.X,      h->st.Z,        h->st.P,        h->st.A,        h->st.N,     h->st.status,
                    h->st.flags,  h->st.wcount,   h->st.wrange,   h->st.wbearing, h->st.whits, h->st.llt_fail,
                    h->d_odom64,  h->d_odom32,    h->d_range,     h->d_bearing,   h->d_offsets, h->d_assoc,
                    h->d_scratch, h->d_replay[0], h->d_replay[1], h->d_xsnap[0],  h->d_xsnap[1]};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    for (int i = 0; i < 2; ++i) {
        cudaEvent_t evs[] = {h->ev_copied[i], h->ev_consumed[i], h->ev_snap[i], h->ev_read[i]};
        for (cudaEvent_t e : evs)
            if (e) cudaEventDestroy(e);
    }
    for (cudaStream_t st : streams)
        if (st) cudaStreamDestroy(st);
    delete h;
    return ASLAM_OK;
}

aslam_status aslam_filter_sync(aslam_filter* h) {
    if (!h) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaStreamSynchronize(h->copy_stream));
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    ASLAM_CUDA(cudaStreamSynchronize(h->out_stream));
    return ASLAM_OK;
}
void* aslam_filter_stream(aslam_filter* h) { return h ? (void*)h->stream : nullptr; }
uint32_t aslam_filter_batch(const aslam_filter* h) { return h ? h->st.batch : 0; }
uint32_t aslam_filter_ld(const aslam_filter* h) { return h ? h->st.ld : 0; }
int32_t aslam_filter_kind(const aslam_filter* h) { return h ? h->kind : -1; }

aslam_status aslam_filter_set_observations(aslam_filter* h, const float* range, const float* bearing,
                                           const uint32_t* offsets) {
    if (!h || !offsets) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    const uint32_t B = h->st.batch;
    const size_t total = offsets[B];
    if (total && (!range || !bearing)) return ASLAM_ERR_ARG;
    for (uint32_t b = 0; b < B; ++b)  // the kernels take offsets[f+1] - offsets[f] as a count: it must not wrap
        if (offsets[b] > offsets[b + 1]) {
            set_error("set_observations: offsets must be non-decreasing");
            return ASLAM_ERR_ARG;
        }
    if (total > h->obs_capacity) {
        ASLAM_CUDA(cudaStreamSynchronize(h->stream));
        if (h->d_range) cudaFree(h->d_range);
        if (h->d_bearing) cudaFree(h->d_bearing);
        h->d_range = h->d_bearing = nullptr;
        h->obs_capacity = 0;
        const size_t cap = total + total / 4 + 16;
        ASLAM_CUDA(cudaMalloc((void**)&h->d_range, cap * sizeof(float)));
        ASLAM_CUDA(cudaMalloc((void**)&h->d_bearing, cap * sizeof(float)));
        h->obs_capacity = cap;
    }
    if (total) {
        ASLAM_CUDA(cudaMemcpyAsync(h->d_range, range, total * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        ASLAM_CUDA(cudaMemcpyAsync(h->d_bearing, bearing, total * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    }
    ASLAM_CUDA(cudaMemcpyAsync(h->d_offsets, offsets, ((size_t)B + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice,
                               h->stream));
    h->cur_range = h->d_range;
    h->cur_bearing = h->d_bearing;
    h->cur_offsets = h->d_offsets;
    h->cur_total = (uint32_t)total;
    h->have_obs = true;
    return ASLAM_OK;
}

aslam_status aslam_filter_set_observations_device(aslam_filter* h, const float* d_range, const float* d_bearing,
                                                  const uint32_t* d_offsets) {
    if (!h || !d_offsets) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    uint32_t total = 0;
    ASLAM_CUDA(cudaMemcpyAsync(&total, d_offsets + h->st.batch, sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    h->cur_range = d_range;
    h->cur_bearing = d_bearing;
    h->cur_offsets = d_offsets;
    h->cur_total = total;
    h->have_obs = true;
    return ASLAM_OK;
}

aslam_status aslam_filter_step_device(aslam_filter* h, const aslam_odom* o) {
    if (!h || !o || !o->x || !o->y || !o->yaw || !o->vx || !o->az || !o->dt) return ASLAM_ERR_ARG;
    if (!h->have_obs) {
        set_error("step before the first set_observations (cbOdom returns while init_z, ekf.cpp:76)");
        return ASLAM_ERR_NOT_READY;
    }
    ASLAM_CUDA(cudaSetDevice(h->device));
    aslam_status rc = ensure_assoc(h, h->cur_total ? h->cur_total : 1);
    if (rc != ASLAM_OK) return rc;
    BatchInput in;
    in.ox = o->x;
    in.oy = o->y;
    in.yaw = o->yaw;
    in.vx = o->vx;
    in.az = o->az;
    in.dt = o->dt;
    in.range = h->cur_range;
    in.bearing = h->cur_bearing;
    in.offsets = h->cur_offsets;
    in.assoc = h->d_assoc;
    in.steps = 1;
    return run_kernel(h, in);
}

aslam_status aslam_filter_step(aslam_filter* h, const aslam_odom* o) {
    ASLAM_NVTX("aslam_filter_step");
    if (!h || !o || !o->x || !o->y || !o->yaw || !o->vx || !o->az || !o->dt) return ASLAM_ERR_ARG;
    if (!h->have_obs) {
        set_error("step before the first set_observations (cbOdom returns while init_z, ekf.cpp:76)");
        return ASLAM_ERR_NOT_READY;
    }
    ASLAM_CUDA(cudaSetDevice(h->device));
    const size_t B = h->st.batch;
    ASLAM_CUDA(cudaMemcpyAsync(h->d_odom64 + 0 * B, o->x, B * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    ASLAM_CUDA(cudaMemcpyAsync(h->d_odom64 + 1 * B, o->y, B * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    ASLAM_CUDA(cudaMemcpyAsync(h->d_odom64 + 2 * B, o->vx, B * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    ASLAM_CUDA(cudaMemcpyAsync(h->d_odom64 + 3 * B, o->az, B * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    ASLAM_CUDA(cudaMemcpyAsync(h->d_odom32 + 0 * B, o->yaw, B * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    ASLAM_CUDA(cudaMemcpyAsync(h->d_odom32 + 1 * B, o->dt, B * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    aslam_odom d;
    d.x = h->d_odom64;
    d.y = h->d_odom64 + B;
    d.vx = h->d_odom64 + 2 * B;
    d.az = h->d_odom64 + 3 * B;
    d.yaw = h->d_odom32;
    d.dt = h->d_odom32 + B;
    return aslam_filter_step_device(h, &d);
}

aslam_status aslam_filter_run_device(aslam_filter* h, uint32_t steps, const aslam_odom* o, const float* d_range,
                                     const float* d_bearing, uint32_t obs_per_filter) {
    ASLAM_NVTX("aslam_filter_run_device");
    if (!h || !o || !o->x || !o->y || !o->yaw || !o->vx || !o->az || !o->dt) return ASLAM_ERR_ARG;
    if (steps == 0) return ASLAM_OK;
    if (obs_per_filter && (!d_range || !d_bearing)) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    const size_t total = (size_t)h->st.batch * obs_per_filter;
    aslam_status rc = ensure_assoc(h, total ? total : 1);
    if (rc != ASLAM_OK) return rc;
    BatchInput in;
    in.ox = o->x;
    in.oy = o->y;
    in.yaw = o->yaw;
    in.vx = o->vx;
    in.az = o->az;
    in.dt = o->dt;
    in.range = d_range;
    in.bearing = d_bearing;
    in.offsets = nullptr;
    in.obs_per_filter = obs_per_filter;
    in.assoc = h->d_assoc;
    // The kernels index a message's observations with 32-bit offsets from `range` / `bearing`
    // (step * batch * obs_per_filter + ...): a replay longer than 2^32 observations is cut into launches that each
    // stay below it, with the pointers advanced on the host (same results: the state carries over in HBM).
    const size_t max_steps = total ? (size_t)0xffffffffu / total : (size_t)steps;
    if (max_steps == 0) {
        set_error("run: batch * obs_per_filter exceeds 2^32 observations per message");
        return ASLAM_ERR_UNSUPPORTED;
    }
    for (uint32_t s0 = 0; s0 < steps;) {
        const uint32_t ns = (uint32_t)std::min<size_t>(steps - s0, max_steps);
        const size_t mo = (size_t)s0 * h->st.batch;
        in.ox = o->x + mo;
        in.oy = o->y + mo;
        in.yaw = o->yaw + mo;
        in.vx = o->vx + mo;
        in.az = o->az + mo;
        in.dt = o->dt + mo;
        in.range = d_range ? d_range + (size_t)s0 * total : nullptr;
        in.bearing = d_bearing ? d_bearing + (size_t)s0 * total : nullptr;
        in.steps = ns;
        rc = run_kernel(h, in);
        if (rc != ASLAM_OK) return rc;
        s0 += ns;
    }
    // the last message stays "current" (ekf.cpp:105: the list persists until replaced)
    h->cur_range = d_range + (size_t)(steps - 1) * total;
    h->cur_bearing = d_bearing + (size_t)(steps - 1) * total;
    h->cur_offsets = nullptr;
    h->cur_total = (uint32_t)total;
    h->have_obs = false;  // no offsets array for it: a later step() needs a fresh set_observations
    return ASLAM_OK;
}

aslam_status aslam_filter_run_host(aslam_filter* h, uint32_t steps, const aslam_odom* o, const float* range,
                                   const float* bearing, uint32_t obs_per_filter) {
    ASLAM_NVTX("aslam_filter_run_host");
    if (!h || !o || !o->x || !o->y || !o->yaw || !o->vx || !o->az || !o->dt) return ASLAM_ERR_ARG;
    if (steps == 0) return ASLAM_OK;
    if (obs_per_filter && (!range || !bearing)) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    const size_t n = (size_t)steps * h->st.batch, no = n * obs_per_filter;
    // one staging block: 4 double arrays, then 2 + 2*obs float arrays (all 8-byte aligned starts)
    const size_t bytes = n * 4 * sizeof(double) + (n * 2 + no * 2) * sizeof(float) + 64;
    const int b = h->replay_next;
    h->replay_next ^= 1;
    if (bytes > h->replay_bytes[b]) {
        ASLAM_CUDA(cudaStreamSynchronize(h->stream));
        ASLAM_CUDA(cudaStreamSynchronize(h->copy_stream));
        if (h->d_replay[b]) cudaFree(h->d_replay[b]);
        h->d_replay[b] = nullptr;
        h->replay_bytes[b] = 0;
        ASLAM_CUDA(cudaMalloc((void**)&h->d_replay[b], bytes));
        h->replay_bytes[b] = bytes;
    }
    // block b was last read by the kernel of the call before the previous one
    cudaStream_t cs = h->copy_stream;
    ASLAM_CUDA(cudaStreamWaitEvent(cs, h->ev_consumed[b], 0));
    double* d64 = reinterpret_cast<double*>(h->d_replay[b]);
    float* d32 = reinterpret_cast<float*>(d64 + 4 * n);
    const cudaMemcpyKind k = cudaMemcpyHostToDevice;
    ASLAM_CUDA(cudaMemcpyAsync(d64 + 0 * n, o->x, n * sizeof(double), k, cs));
    ASLAM_CUDA(cudaMemcpyAsync(d64 + 1 * n, o->y, n * sizeof(double), k, cs));
    ASLAM_CUDA(cudaMemcpyAsync(d64 + 2 * n, o->vx, n * sizeof(double), k, cs));
    ASLAM_CUDA(cudaMemcpyAsync(d64 + 3 * n, o->az, n * sizeof(double), k, cs));
    ASLAM_CUDA(cudaMemcpyAsync(d32 + 0 * n, o->yaw, n * sizeof(float), k, cs));
    ASLAM_CUDA(cudaMemcpyAsync(d32 + 1 * n, o->dt, n * sizeof(float), k, cs));
    float* dr = d32 + 2 * n;
    float* db = dr + no;
    if (no) {
        ASLAM_CUDA(cudaMemcpyAsync(dr, range, no * sizeof(float), k, cs));
        ASLAM_CUDA(cudaMemcpyAsync(db, bearing, no * sizeof(float), k, cs));
    }
    ASLAM_CUDA(cudaEventRecord(h->ev_copied[b], cs));
    ASLAM_CUDA(cudaStreamWaitEvent(h->stream, h->ev_copied[b], 0));
    aslam_odom d;
    d.x = d64;
    d.y = d64 + n;
    d.vx = d64 + 2 * n;
    d.az = d64 + 3 * n;
    d.yaw = d32;
    d.dt = d32 + n;
    const aslam_status rc = aslam_filter_run_device(h, steps, &d, dr, db, obs_per_filter);
    ASLAM_CUDA(cudaEventRecord(h->ev_consumed[b], h->stream));
    return rc;
}

uint64_t aslam_filter_launch_count(aslam_filter* h) { return h ? h->launches : 0; }

aslam_status aslam_filter_get_llt_fail(aslam_filter* h, int32_t* fail_k) {
    if (!h || !fail_k) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    ASLAM_CUDA(cudaMemcpyAsync(fail_k, h->st.llt_fail, (size_t)h->st.batch * sizeof(int32_t), cudaMemcpyDeviceToHost,
                               h->stream));
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}

aslam_status aslam_filter_get_state(aslam_filter* h, double* X, double* P, uint32_t* N, int32_t* assoc,
                                    uint32_t* status) {
    if (!h) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    const size_t B = h->st.batch, ld = h->st.ld;
    if (X) ASLAM_CUDA(cudaMemcpyAsync(X, h->st.X, B * ld * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    if (P) {
        aslam_status rc = ensure_scratch(h, B * ld * ld * sizeof(double));
        if (rc != ASLAM_OK) return rc;
        batch_transpose_kernel<<<(unsigned)B, 128, 0, h->stream>>>(h->st.P, h->d_scratch, (uint32_t)ld);
        ASLAM_CUDA(cudaGetLastError());
        ASLAM_CUDA(cudaMemcpyAsync(P, h->d_scratch, B * ld * ld * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    }
    if (N) ASLAM_CUDA(cudaMemcpyAsync(N, h->st.N, B * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
    if (status) ASLAM_CUDA(cudaMemcpyAsync(status, h->st.status, B * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
    if (assoc && h->cur_total)
        ASLAM_CUDA(cudaMemcpyAsync(assoc, h->d_assoc, (size_t)h->cur_total * sizeof(int32_t), cudaMemcpyDeviceToHost,
                                   h->stream));
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}

aslam_status aslam_filter_get_aux(aslam_filter* h, double* Z, double* a00, double* a10) {
    if (!h) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    const size_t B = h->st.batch, ld = h->st.ld;
    if (Z) ASLAM_CUDA(cudaMemcpyAsync(Z, h->st.Z, B * ld * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    if (a00 || a10) {
        std::vector<double> a(2 * B);
        ASLAM_CUDA(cudaMemcpyAsync(a.data(), h->st.A, 2 * B * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
        ASLAM_CUDA(cudaStreamSynchronize(h->stream));
        for (size_t b = 0; b < B; ++b) {
            if (a00) a00[b] = a[2 * b];
            if (a10) a10[b] = a[2 * b + 1];
        }
    }
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}

aslam_status aslam_filter_get_wait(aslam_filter* h, uint32_t* count, float* range, float* bearing, uint32_t* hits) {
    if (!h) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    const size_t B = h->st.batch, W = ASLAM_MAX_WAIT;
    if (count) ASLAM_CUDA(cudaMemcpyAsync(count, h->st.wcount, B * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
    if (range) ASLAM_CUDA(cudaMemcpyAsync(range, h->st.wrange, B * W * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    if (bearing)
        ASLAM_CUDA(cudaMemcpyAsync(bearing, h->st.wbearing, B * W * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    if (hits) ASLAM_CUDA(cudaMemcpyAsync(hits, h->st.whits, B * W * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}

aslam_status aslam_filter_get_landmarks(aslam_filter* h, double* x, double* y, uint32_t* count) {
    if (!h || !x || !y || !count) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    const size_t B = h->st.batch, lmax = (h->st.ld - 3) / 2;
    const size_t bytes = 2 * B * lmax * sizeof(double) + B * sizeof(uint32_t);
    aslam_status rc = ensure_scratch(h, bytes);
    if (rc != ASLAM_OK) return rc;
    double* dx = h->d_scratch;
    double* dy = dx + B * lmax;
    uint32_t* dc = reinterpret_cast<uint32_t*>(dy + B * lmax);
    batch_landmarks_kernel<<<(unsigned)((B + 127) / 128), 128, 0, h->stream>>>(h->st, dx, dy, dc);
    ASLAM_CUDA(cudaGetLastError());
    ASLAM_CUDA(cudaMemcpyAsync(x, dx, B * lmax * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    ASLAM_CUDA(cudaMemcpyAsync(y, dy, B * lmax * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    ASLAM_CUDA(cudaMemcpyAsync(count, dc, B * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}

aslam_status aslam_filter_device_state(aslam_filter* h, const double** d_X, const uint32_t** d_N) {
    if (!h) return ASLAM_ERR_ARG;
    if (d_X) *d_X = h->st.X;
    if (d_N) *d_N = h->st.N;
    return ASLAM_OK;
}

aslam_status aslam_filter_load_state(aslam_filter* h, const uint32_t* N, const double* X, const double* Z,
                                     const double* P) {
    if (!h || !N || !X || !Z || !P) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    const size_t B = h->st.batch, ld = h->st.ld;
    for (size_t b = 0; b < B; ++b)
        if (N[b] < 3 || N[b] > ld || (int32_t)N[b] >= h->cfg.max_state_dim || (N[b] & 1u) == 0) {
            set_error("load_state: N must be odd, >= 3 and < max_state_dim");
            return ASLAM_ERR_ARG;
        }
    aslam_status rc = ensure_scratch(h, B * ld * ld * sizeof(double));
    if (rc != ASLAM_OK) return rc;
    ASLAM_CUDA(cudaMemcpyAsync(h->st.N, N, B * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
    ASLAM_CUDA(cudaMemcpyAsync(h->st.X, X, B * ld * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    ASLAM_CUDA(cudaMemcpyAsync(h->st.Z, Z, B * ld * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    ASLAM_CUDA(cudaMemcpyAsync(h->d_scratch, P, B * ld * ld * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    batch_transpose_kernel<<<(unsigned)B, 128, 0, h->stream>>>(h->d_scratch, h->st.P, (uint32_t)ld);
    ASLAM_CUDA(cudaGetLastError());
    ASLAM_CUDA(cudaMemsetAsync(h->st.flags, 0, B * sizeof(uint32_t), h->stream));
    // everything else that outlives a message is put back to its initial value, so that nothing of a previous run
    // leaks into the loaded state: empty wait-list, A = I (ekf.cpp:60), no status bits, no LLT failure
    ASLAM_CUDA(cudaMemsetAsync(h->st.wcount, 0, B * sizeof(uint32_t), h->stream));
    ASLAM_CUDA(cudaMemsetAsync(h->st.status, 0, B * sizeof(uint32_t), h->stream));
    ASLAM_CUDA(cudaMemsetAsync(h->st.llt_fail, 0xff, B * sizeof(int32_t), h->stream));
    if (h->st.A) {
        std::vector<double> a(2 * B);
        for (size_t b = 0; b < B; ++b) {
            a[2 * b] = 1.0;
            a[2 * b + 1] = 0.0;
        }
        ASLAM_CUDA(cudaMemcpyAsync(h->st.A, a.data(), 2 * B * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    }
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}

aslam_status aslam_filter_load_aux(aslam_filter* h, const double* a00, const double* a10, const uint32_t* wcount,
                                   const float* wrange, const float* wbearing, const uint32_t* whits,
                                   const uint32_t* status) {
    if (!h) return ASLAM_ERR_ARG;
    if ((a00 == nullptr) != (a10 == nullptr)) return ASLAM_ERR_ARG;
    if (wcount && (!wrange || !wbearing || !whits)) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    const size_t B = h->st.batch;
    if (a00 && h->st.A) {
        std::vector<double> a(2 * B);
        for (size_t b = 0; b < B; ++b) {
            a[2 * b] = a00[b];
            a[2 * b + 1] = a10[b];
        }
        ASLAM_CUDA(cudaMemcpyAsync(h->st.A, a.data(), 2 * B * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    }
    if (wcount) {
        for (size_t b = 0; b < B; ++b)
            if (wcount[b] > ASLAM_MAX_WAIT) {
                set_error("load_aux: wait-list longer than ASLAM_MAX_WAIT");
                return ASLAM_ERR_ARG;
            }
        const size_t W = (size_t)ASLAM_MAX_WAIT;
        ASLAM_CUDA(cudaMemcpyAsync(h->st.wcount, wcount, B * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
        ASLAM_CUDA(cudaMemcpyAsync(h->st.wrange, wrange, B * W * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        ASLAM_CUDA(cudaMemcpyAsync(h->st.wbearing, wbearing, B * W * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        ASLAM_CUDA(cudaMemcpyAsync(h->st.whits, whits, B * W * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
    }
    if (status)
        ASLAM_CUDA(cudaMemcpyAsync(h->st.status, status, B * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}

}  // extern "C"
