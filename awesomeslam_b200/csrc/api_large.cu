// C-ABI layer of the large-map filters (include/aslam_c.h, aslam_large_*): handle lifetime,
// device memory, state import/export and the per-step call sequence. All arithmetic is in
// ekf_large.cu / ukf_large.cu.
#include <cuda_runtime.h>

#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "dense_fp64.hpp"
#include "large.hpp"
#include "nvtx_range.hpp"

namespace aslam {
aslam_status large_ekf_phase1(aslam_large* h, double x, double y, float yaw, double vx, double az, float dt, uint32_t m,
                              int mode);
aslam_status large_ekf_phase2(aslam_large* h);
aslam_status large_ukf_create(aslam_large* h);
void large_ukf_destroy(aslam_large* h);
aslam_status large_ukf_step(aslam_large* h, double x, double y, float yaw, double vx, double az, float dt, uint32_t m);
aslam_status large_ukf_load_z(aslam_large* h, const double* Z);
aslam_status large_ukf_status(aslam_large* h, uint32_t* status, int32_t* llt_fail);
aslam_status large_ukf_set_precision(aslam_large* h, int mode);
aslam_status large_ukf_products_ms(aslam_large* h, float* ms);
}  // namespace aslam

using namespace aslam;

namespace {

aslam_status create_large(int kind, const aslam_config* cfg, uint32_t n, uint32_t rb, uint32_t re, int32_t device,
                          aslam_large** out) {
    if (!out) return ASLAM_ERR_ARG;
    *out = nullptr;
    if (n < 5 || (n & 1u) == 0 || rb >= re || re > n || (rb != 0 && rb < 3)) {
        set_error("large create: n must be odd >= 5 and 0 <= row_begin < row_end <= n (a shard that starts after row 0 "
                  "starts at row >= 3)");
        return ASLAM_ERR_ARG;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_error("large create: no CUDA device (this library has no CPU path)");
        return ASLAM_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) {
        set_error("large create: device index out of range");
        return ASLAM_ERR_ARG;
    }
    ASLAM_CUDA(cudaSetDevice(device));
    aslam_large* h = new (std::nothrow) aslam_large;
    if (!h) return ASLAM_ERR_ARG;
    h->kind = kind;
    if (cfg)
        h->cfg = *cfg;
    else
        aslam_config_default(&h->cfg);
    h->device = device;
    h->n = n;
    h->row_begin = rb;
    h->row_end = re;
    h->ldp = (((size_t)(re - rb) + 15) / 16) * 16;
    aslam_status rc = ASLAM_OK;
    auto A = [&](void** p, size_t bytes) {
        if (rc == ASLAM_OK && !cuda_ok(cudaMalloc(p, bytes), "cudaMalloc")) rc = ASLAM_ERR_CUDA;
    };
    if (!cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate")) rc = ASLAM_ERR_CUDA;
    A((void**)&h->P, (size_t)n * h->ldp * sizeof(double));
    A((void**)&h->X, (size_t)n * sizeof(double));
    A((void**)&h->A, 2 * sizeof(double));
    A((void**)&h->predA, 4 * sizeof(double));
    A((void**)&h->sync, 4 * sizeof(unsigned long long));
    A((void**)&h->ctx, sizeof(LargeCtx));
    A((void**)&h->d_obs, 2 * ASLAM_MAX_OBS * sizeof(float));
    if (kind == ASLAM_KIND_EKF) {
        A((void**)&h->exch, (size_t)2 * LARGE_QMAX * n * sizeof(double));
        A((void**)&h->hp, (size_t)n * 132 * sizeof(double));
        A((void**)&h->K, (size_t)132 * n * sizeof(double));
    }
    if (rc == ASLAM_OK && !cuda_ok(cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking), "cudaStreamCreate")) rc = ASLAM_ERR_CUDA;
    if (rc == ASLAM_OK && kind == ASLAM_KIND_EKF) {
        // highest priority: its few small CTAs are placed ahead of the next update's 148 when an SM has room for either
        int prio_lo = 0, prio_hi = 0;
        cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
        if (!cuda_ok(cudaStreamCreateWithPriority(&h->solve_stream, cudaStreamNonBlocking, prio_hi), "cudaStreamCreate")) rc = ASLAM_ERR_CUDA;
        cudaEvent_t* evs[] = {&h->ev_before_upd, &h->ev_now, &h->ev_solve};
        for (cudaEvent_t* e : evs)
            if (rc == ASLAM_OK && !cuda_ok(cudaEventCreateWithFlags(e, cudaEventDisableTiming), "cudaEventCreate")) rc = ASLAM_ERR_CUDA;
        h->no_overlap = std::getenv("ASLAM_LARGE_NO_OVERLAP") != nullptr;  // diagnostic switch, same kernels either way
        if (const char* e = std::getenv("ASLAM_LARGE_OVERLAP_MIN_BYTES")) h->overlap_min_bytes = (size_t)std::strtoull(e, nullptr, 10);
        if (const char* e = std::getenv("ASLAM_LARGE_RESERVE_MIN_BYTES")) h->reserve_min_bytes = (size_t)std::strtoull(e, nullptr, 10);
        if (const char* e = std::getenv("ASLAM_LARGE_RESERVE_SMS")) h->reserve_sms = std::atoi(e);
    }
    if (rc == ASLAM_OK && !cuda_ok(cudaEventCreateWithFlags(&h->ev_ctx_free, cudaEventDisableTiming), "cudaEventCreate")) rc = ASLAM_ERR_CUDA;
    if (rc == ASLAM_OK && !cuda_ok(cudaEventCreateWithFlags(&h->ev_front, cudaEventDisableTiming), "cudaEventCreate")) rc = ASLAM_ERR_CUDA;
    if (rc == ASLAM_OK && !cuda_ok(cudaEventCreate(&h->ev0), "cudaEventCreate")) rc = ASLAM_ERR_CUDA;
    if (rc == ASLAM_OK && !cuda_ok(cudaEventCreate(&h->ev1), "cudaEventCreate")) rc = ASLAM_ERR_CUDA;
    if (rc == ASLAM_OK) {
        const double a[2] = {1.0, 0.0};  // A = I (ekf.cpp:60)
        if (!cuda_ok(cudaMemcpyAsync(h->A, a, sizeof(a), cudaMemcpyHostToDevice, h->stream), "init A")) rc = ASLAM_ERR_CUDA;
        if (!cuda_ok(cudaMemsetAsync(h->P, 0, (size_t)n * h->ldp * sizeof(double), h->stream), "init P")) rc = ASLAM_ERR_CUDA;
        if (!cuda_ok(cudaMemsetAsync(h->X, 0, (size_t)n * sizeof(double), h->stream), "init X")) rc = ASLAM_ERR_CUDA;
        if (!cuda_ok(cudaMemsetAsync(h->ctx, 0, sizeof(LargeCtx), h->stream), "init ctx")) rc = ASLAM_ERR_CUDA;
        if (!cuda_ok(cudaMemsetAsync(h->sync, 0, 4 * sizeof(unsigned long long), h->stream), "init sync")) rc = ASLAM_ERR_CUDA;
        if (!cuda_ok(cudaStreamSynchronize(h->stream), "init sync")) rc = ASLAM_ERR_CUDA;
    }
    if (rc == ASLAM_OK && kind == ASLAM_KIND_UKF) rc = large_ukf_create(h);
    if (rc != ASLAM_OK) {
        aslam_large_destroy(h);
        return rc;
    }
    h->timed = kind == ASLAM_KIND_UKF;  // EKF: opt-in (aslam_large_set_timing)
    if (kind == ASLAM_KIND_EKF && std::getenv("ASLAM_LARGE_TRACE")) {
        if (cudaMalloc((void**)&h->trace, 64 * 16 * sizeof(unsigned long long)) != cudaSuccess) h->trace = nullptr;
        if (h->trace) cudaMemset(h->trace, 0, 64 * 16 * sizeof(unsigned long long));
    }
    h->pdl = std::getenv("ASLAM_LARGE_NO_PDL") == nullptr;  // diagnostic switch, same kernels either way
    *out = h;
    return ASLAM_OK;
}

// EKF handles stage the observations on the side stream (the front kernel that reads them runs there), behind the
// event that says the previous step no longer needs the context / observation buffers.
aslam_status stage_obs(aslam_large* h, const float* range, const float* bearing, uint32_t m) {
    if (m > ASLAM_MAX_OBS) {
        set_error("large step: more than ASLAM_MAX_OBS observations in one message");
        return ASLAM_ERR_ARG;
    }
    if (m && (!range || !bearing)) return ASLAM_ERR_ARG;
    if (h->kind == ASLAM_KIND_EKF) {  // the EKF front kernel takes the observations in its arguments
        for (uint32_t i = 0; i < m; ++i) {
            h->obs_host[i] = range[i];
            h->obs_host[ASLAM_MAX_OBS + i] = bearing[i];
        }
    } else if (m) {
        ASLAM_CUDA(cudaMemcpyAsync(h->d_obs, range, m * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        ASLAM_CUDA(cudaMemcpyAsync(h->d_obs + ASLAM_MAX_OBS, bearing, m * sizeof(float), cudaMemcpyHostToDevice,
                                   h->stream));
    }
    h->m_last = m;
    return ASLAM_OK;
}

aslam_status fetch_assoc(aslam_large* h, int32_t* assoc_out) {
    if (!assoc_out || !h->m_last) return ASLAM_OK;
    ASLAM_CUDA(cudaMemcpyAsync(assoc_out, (const char*)h->ctx + offsetof(LargeCtx, assoc), h->m_last * sizeof(int32_t),
                               cudaMemcpyDeviceToHost, h->stream));
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}

}  // namespace

extern "C" {

aslam_status aslam_large_ekf_create(const aslam_config* cfg, uint32_t n, uint32_t row_begin, uint32_t row_end,
                                    int32_t device, aslam_large** out) {
    return create_large(ASLAM_KIND_EKF, cfg, n, row_begin, row_end, device, out);
}

aslam_status aslam_large_ukf_create(const aslam_config* cfg, uint32_t n, int32_t device, aslam_large** out) {
    return create_large(ASLAM_KIND_UKF, cfg, n, 0, n, device, out);
}

aslam_status aslam_large_destroy(aslam_large* h) {
    if (!h) return ASLAM_OK;
    cudaSetDevice(h->device);
    if (h->side) cudaStreamSynchronize(h->side);
    if (h->solve_stream) cudaStreamSynchronize(h->solve_stream);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->ukf) large_ukf_destroy(h);
    if (h->solve_stream) cudaStreamDestroy(h->solve_stream);
    if (h->ev_before_upd) cudaEventDestroy(h->ev_before_upd);
    if (h->ev_now) cudaEventDestroy(h->ev_now);
    if (h->ev_solve) cudaEventDestroy(h->ev_solve);
    if (h->ev_ctx_free) cudaEventDestroy(h->ev_ctx_free);
    if (h->ev_front) cudaEventDestroy(h->ev_front);
    if (h->side) cudaStreamDestroy(h->side);
    for (int p = 0; p < LARGE_MAX_PEERS; ++p)
        if (h->peer_ipc[p] && h->peer_base[p]) cudaIpcCloseMemHandle(h->peer_base[p]);
    void* ptrs[] = {h->P, h->P_alt, h->X, h->A, h->predA, h->sync, h->trace, h->ctx, h->d_obs, h->exch, h->hp, h->K, h->peer_block};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return ASLAM_OK;
}

aslam_status aslam_large_sync(aslam_large* h) {
    if (!h) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}

void* aslam_large_stream(aslam_large* h) { return h ? (void*)h->stream : nullptr; }
uint64_t aslam_large_launch_count(aslam_large* h) { return h ? h->launches : 0; }

aslam_status aslam_large_load_state(aslam_large* h, const double* X, const double* Z, const double* P, size_t ldp) {
    if (!h || !X || !P || ldp < h->n) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    const size_t n = h->n, rows = h->row_end - h->row_begin;
    h->pending = false;  // (ordered on `stream` behind the last update, which ends after the last solve kernel)
    ASLAM_CUDA(cudaMemcpyAsync(h->X, X, n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    ASLAM_CUDA(cudaMemcpy2DAsync(h->P, h->ldp * sizeof(double), P + h->row_begin, ldp * sizeof(double),
                                 rows * sizeof(double), n, cudaMemcpyHostToDevice, h->stream));
    if (h->kind == ASLAM_KIND_UKF) {
        aslam_status rc = large_ukf_load_z(h, Z);
        if (rc != ASLAM_OK) return rc;
    }
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}

aslam_status aslam_large_get_state(aslam_large* h, double* X, double* P, size_t ldp) {
    if (!h || (P && ldp < (size_t)(h->row_end - h->row_begin))) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    const size_t n = h->n, rows = h->row_end - h->row_begin;
    if (X) ASLAM_CUDA(cudaMemcpyAsync(X, h->X, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    // P receives this handle's rows only: column j at P + j*ldp, rows row_begin..row_end-1
    if (P)
        ASLAM_CUDA(cudaMemcpy2DAsync(P, ldp * sizeof(double), h->P, h->ldp * sizeof(double), rows * sizeof(double), n,
                                     cudaMemcpyDeviceToHost, h->stream));
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}

aslam_status aslam_large_step(aslam_large* h, double x, double y, float yaw, double vx, double az, float dt,
                              const float* range, const float* bearing, uint32_t m, int32_t* assoc_out) {
    ASLAM_NVTX("aslam_large_step");
    if (!h) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    aslam_status rc = stage_obs(h, range, bearing, m);
    if (rc != ASLAM_OK) return rc;
    if (h->kind == ASLAM_KIND_UKF) {
        rc = large_ukf_step(h, x, y, yaw, vx, az, dt, m);
        if (rc != ASLAM_OK) return rc;
        return fetch_assoc(h, assoc_out);
    }
    rc = large_ekf_phase1(h, x, y, yaw, vx, az, dt, m, LARGE_STEP_DIRECT);
    if (rc != ASLAM_OK) return rc;
    rc = large_ekf_phase2(h);
    if (rc != ASLAM_OK) return rc;
    return fetch_assoc(h, assoc_out);
}

aslam_status aslam_large_step_phase1(aslam_large* h, double x, double y, float yaw, double vx, double az, float dt,
                                     const float* range, const float* bearing, uint32_t m, uint32_t* q_out) {
    ASLAM_NVTX("aslam_large_step_phase1");
    if (!h || h->kind != ASLAM_KIND_EKF) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    aslam_status rc = stage_obs(h, range, bearing, m);
    if (rc != ASLAM_OK) return rc;
    rc = large_ekf_phase1(h, x, y, yaw, vx, az, dt, m, LARGE_STEP_EXCH);
    if (q_out) *q_out = h->q_host;
    return rc;
}

aslam_status aslam_large_exchange_buffer(aslam_large* h, double** d_buf, size_t* count) {
    if (!h || h->kind != ASLAM_KIND_EKF) return ASLAM_ERR_ARG;
    if (d_buf) *d_buf = h->exch;
    // q <= 20: (P H_o') slices [q][n] + the q x q block (padded to 20 x 20); larger q: raw rows + slices (round 1)
    if (count) *count = h->q_host <= 20 ? (size_t)h->q_host * h->n + 400 : (size_t)2 * h->q_host * h->n;
    return ASLAM_OK;
}

aslam_status aslam_large_step_phase2(aslam_large* h, int32_t* assoc_out) {
    ASLAM_NVTX("aslam_large_step_phase2");
    if (!h || h->kind != ASLAM_KIND_EKF) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    aslam_status rc = large_ekf_phase2(h);
    if (rc != ASLAM_OK) return rc;
    return fetch_assoc(h, assoc_out);
}

// block = [E parity 0][E parity 1][LARGE_MAX_PEERS arrival counters][UUID of the GPU the block lives on, 16 bytes]
static size_t peer_uuid_offset(const aslam_large* h) {
    return (size_t)4 * LARGE_QMAX * h->n * sizeof(double) + LARGE_MAX_PEERS * sizeof(unsigned long long);
}
static size_t peer_block_bytes(const aslam_large* h) { return peer_uuid_offset(h) + 64; }
static bool device_uuid(int device, unsigned char (&out)[16]) {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return false;
    static_assert(sizeof(prop.uuid) == 16, "CUDA device UUIDs are 16 bytes");
    memcpy(out, &prop.uuid, 16);
    return true;
}

aslam_status aslam_large_peer_export(aslam_large* h, void* ipc_handle64, void** block) {
    if (!h || h->kind != ASLAM_KIND_EKF) return ASLAM_ERR_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "the ABI carries the IPC handle as 64 opaque bytes");
    ASLAM_CUDA(cudaSetDevice(h->device));
    if (!h->peer_block) {
        ASLAM_CUDA(cudaMalloc((void**)&h->peer_block, peer_block_bytes(h)));
        ASLAM_CUDA(cudaMemset(h->peer_block, 0, peer_block_bytes(h)));
        // which GPU this block lives on: a peer that maps it (CUDA IPC reports the MAPPING device, not the owner) reads
        // this to find out whether the two shards share a GPU
        unsigned char uuid[16];
        if (!device_uuid(h->device, uuid)) {
            set_error("peer_export: cudaGetDeviceProperties failed");
            return ASLAM_ERR_CUDA;
        }
        ASLAM_CUDA(cudaMemcpy(h->peer_block + peer_uuid_offset(h), uuid, 16, cudaMemcpyHostToDevice));
        ASLAM_CUDA(cudaDeviceSynchronize());
    }
    if (ipc_handle64) ASLAM_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(ipc_handle64), h->peer_block));
    if (block) *block = h->peer_block;
    return ASLAM_OK;
}

aslam_status aslam_large_peer_attach(aslam_large* h, uint32_t world, uint32_t rank, const void* ipc_handles,
                                     void* const* blocks) {
    if (!h || h->kind != ASLAM_KIND_EKF || !h->peer_block || world == 0 || world > (uint32_t)LARGE_MAX_PEERS ||
        rank >= world || (!ipc_handles && !blocks)) {
        set_error("peer_attach: export first; 1 <= world <= 8; pass ipc_handles or blocks");
        return ASLAM_ERR_ARG;
    }
    if (h->peer_world) {
        set_error("peer_attach: already attached");
        return ASLAM_ERR_ARG;
    }
    ASLAM_CUDA(cudaSetDevice(h->device));
    for (uint32_t p = 0; p < world; ++p) {
        if (p == rank) {
            h->peer_base[p] = h->peer_block;
        } else if (blocks) {
            h->peer_base[p] = static_cast<unsigned char*>(blocks[p]);
            cudaPointerAttributes at{};
            if (cudaPointerGetAttributes(&at, blocks[p]) == cudaSuccess && at.device != h->device) {
                const cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                    set_error("peer_attach: no peer access between the two devices");
                    return ASLAM_ERR_CUDA;
                }
                cudaGetLastError();
            }
        } else {
            cudaIpcMemHandle_t hd;
            memcpy(&hd, static_cast<const unsigned char*>(ipc_handles) + (size_t)p * 64, 64);
            void* ptr = nullptr;
            ASLAM_CUDA(cudaIpcOpenMemHandle(&ptr, hd, cudaIpcMemLazyEnablePeerAccess));
            h->peer_base[p] = static_cast<unsigned char*>(ptr);
            h->peer_ipc[p] = true;
        }
        // A peer shard on the SAME device (several handles of one test process on one GPU): its solve kernel must find
        // room while this handle's kernels wait for it, so the streaming update is not made resident early (see
        // large_ekf_phase2_fast).
        if (p != rank) {
            unsigned char mine[16], theirs[16];
            if (!device_uuid(h->device, mine) ||
                !cuda_ok(cudaMemcpy(theirs, h->peer_base[p] + peer_uuid_offset(h), 16, cudaMemcpyDeviceToHost),
                         "peer_attach: reading the peer block")) {
                set_error("peer_attach: cannot read the peer's exchange block (exported by this library version?)");
                return ASLAM_ERR_CUDA;
            }
            if (memcmp(mine, theirs, 16) == 0) h->peer_same_device = true;
        }
    }
    h->peer_world = world;
    h->peer_rank = rank;
    h->peer_epoch = 0;
    return ASLAM_OK;
}

aslam_status aslam_large_step_fused(aslam_large* h, double x, double y, float yaw, double vx, double az, float dt,
                                    const float* range, const float* bearing, uint32_t m, int32_t* assoc_out) {
    ASLAM_NVTX("aslam_large_step_fused");
    if (!h || h->kind != ASLAM_KIND_EKF) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    aslam_status rc = stage_obs(h, range, bearing, m);
    if (rc != ASLAM_OK) return rc;
    rc = large_ekf_phase1(h, x, y, yaw, vx, az, dt, m, LARGE_STEP_PEER);
    if (rc != ASLAM_OK) return rc;
    rc = large_ekf_phase2(h);
    if (rc != ASLAM_OK) return rc;
    return fetch_assoc(h, assoc_out);
}

/* diagnostics: copies the [64][8] timeline of the last 64 steps (ASLAM_LARGE_TRACE=1), else ASLAM_ERR_UNSUPPORTED */
aslam_status aslam_large_trace(aslam_large* h, unsigned long long* out1024) {
    if (!h || !out1024) return ASLAM_ERR_ARG;
    if (!h->trace) return ASLAM_ERR_UNSUPPORTED;
    ASLAM_CUDA(cudaSetDevice(h->device));
    ASLAM_CUDA(cudaStreamSynchronize(h->side));
    if (h->solve_stream) ASLAM_CUDA(cudaStreamSynchronize(h->solve_stream));
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    ASLAM_CUDA(cudaMemcpy(out1024, h->trace, 64 * 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    return ASLAM_OK;
}

aslam_status aslam_large_set_timing(aslam_large* h, int32_t on) {
    if (!h) return ASLAM_ERR_ARG;
    h->timed = on != 0;
    return ASLAM_OK;
}

aslam_status aslam_large_last_kernel_ms(aslam_large* h, float* ms) {
    if (!h || !ms || !h->timed) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaEventSynchronize(h->ev1));
    ASLAM_CUDA(cudaEventElapsedTime(ms, h->ev0, h->ev1));
    return ASLAM_OK;
}

aslam_status aslam_large_status(aslam_large* h, uint32_t* status, int32_t* llt_fail) {
    if (!h) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    if (h->kind == ASLAM_KIND_UKF) return large_ukf_status(h, status, llt_fail);
    uint32_t st = 0;
    {
        ASLAM_CUDA(cudaMemcpyAsync(&st, (const char*)h->ctx + offsetof(LargeCtx, status), sizeof(uint32_t),
                                   cudaMemcpyDeviceToHost, h->stream));
        ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    }
    if (status) *status = st;
    if (llt_fail) *llt_fail = -1;
    if (st & ASLAM_ST_PEER_TIMEOUT) {
        set_error("large step: a peer's rows did not arrive within the spin bound; that step applied the predict only");
        return ASLAM_ERR_TIMEOUT;
    }
    return ASLAM_OK;
}

aslam_status aslam_dense_gemm_nt(void* stream, int32_t M, int32_t N, int32_t K, const double* d_A, size_t lda,
                                 const double* d_B, size_t ldb, double* d_C, size_t ldc, double alpha, double beta,
                                 int32_t lower_only) {
    if (!d_A || !d_B || !d_C) return ASLAM_ERR_ARG;
    dense::Epilogue e;
    e.alpha = alpha;
    e.beta = beta;
    e.lower_only = lower_only != 0;
    return dense::dgemm_nt((cudaStream_t)stream, M, N, K, d_A, lda, d_B, ldb, d_C, ldc, e, nullptr);
}

aslam_status aslam_dense_cholesky(void* stream, double* d_A, size_t ld, int32_t n, double* d_inv_blocks,
                                  int32_t* d_fail) {
    if (!d_A || !d_inv_blocks || !d_fail || n <= 0) return ASLAM_ERR_ARG;
    return dense::cholesky_lower((cudaStream_t)stream, d_A, ld, n, d_inv_blocks, d_fail, nullptr);
}

aslam_status aslam_dense_trsm(void* stream, double* d_B, size_t ldb, int32_t m, const double* d_L, size_t ldl,
                              int32_t n, const double* d_inv_blocks) {
    if (!d_B || !d_L || !d_inv_blocks) return ASLAM_ERR_ARG;
    return dense::trsm_right_lower_t((cudaStream_t)stream, d_B, ldb, m, d_L, ldl, n, d_inv_blocks, nullptr);
}

aslam_status aslam_large_ukf_set_precision(aslam_large* h, int32_t mode) {
    if (!h || h->kind != ASLAM_KIND_UKF) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    return large_ukf_set_precision(h, mode);
}

aslam_status aslam_large_ukf_products_ms(aslam_large* h, float* ms) {
    if (!h || h->kind != ASLAM_KIND_UKF) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(h->device));
    return large_ukf_products_ms(h, ms);
}

aslam_status aslam_dense_gemm_nt_tf32(void* stream, int32_t M, int32_t N, int32_t K, const double* d_A, size_t lda,
                                      const double* d_B, size_t ldb, double* d_C, size_t ldc, double alpha, double beta,
                                      int32_t lower_only, int32_t passes) {
    if (!d_A || !d_B || !d_C || M <= 0 || N <= 0 || K <= 0 || (passes != 1 && passes != 3)) return ASLAM_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t ldf = dense::tf32_padded_k(K);
    float *ahi = nullptr, *alo = nullptr, *bhi = nullptr, *blo = nullptr;
    int* err = nullptr;
    ASLAM_CUDA(cudaMalloc((void**)&ahi, (size_t)M * ldf * sizeof(float)));
    ASLAM_CUDA(cudaMalloc((void**)&alo, (size_t)M * ldf * sizeof(float)));
    ASLAM_CUDA(cudaMalloc((void**)&bhi, (size_t)N * ldf * sizeof(float)));
    ASLAM_CUDA(cudaMalloc((void**)&blo, (size_t)N * ldf * sizeof(float)));
    ASLAM_CUDA(cudaMalloc((void**)&err, sizeof(int)));
    ASLAM_CUDA(cudaMemsetAsync(err, 0, sizeof(int), st));
    aslam_status rc = dense::split_tf32(st, d_A, lda, M, K, ahi, alo);
    if (rc == ASLAM_OK) rc = dense::split_tf32(st, d_B, ldb, N, K, bhi, blo);
    dense::Epilogue e;
    e.alpha = alpha;
    e.beta = beta;
    e.lower_only = lower_only != 0;
    if (rc == ASLAM_OK) rc = dense::tf32_gemm_nt(st, M, N, K, ahi, alo, bhi, blo, d_C, ldc, e, passes, err);
    int herr = 0;
    if (rc == ASLAM_OK && !cuda_ok(cudaMemcpyAsync(&herr, err, sizeof(int), cudaMemcpyDeviceToHost, st), "tf32 err")) rc = ASLAM_ERR_CUDA;
    if (!cuda_ok(cudaStreamSynchronize(st), "tf32 gemm")) rc = ASLAM_ERR_CUDA;
    cudaFree(ahi); cudaFree(alo); cudaFree(bhi); cudaFree(blo); cudaFree(err);
    if (rc == ASLAM_OK && herr) {
        set_error("tf32 gemm: a wait inside the tcgen05 pipeline expired (code " + std::to_string(herr) + ")");
        return ASLAM_ERR_CUDA;
    }
    return rc;
}

aslam_status aslam_probe_tf32_tflops(int32_t device, double* tflops) {
    if (!tflops) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(device));
    return dense::probe_tf32_tflops(tflops);
}

aslam_status aslam_probe_dmma_tflops(int32_t device, double* tflops) {
    if (!tflops) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaSetDevice(device));
    return dense::probe_dmma_tflops(tflops);
}

}  // extern "C"
