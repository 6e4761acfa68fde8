// Large-map filter (one filter, thousands of landmarks): shared definitions of the C-ABI
// handle and the per-step context that the small kernels hand to the streaming kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "internal.hpp"

namespace aslam {

constexpr int LARGE_QMAX = 3 + 2 * (int)ASLAM_MAX_OBS;  // rows of H in use: pose + 2 per observed landmark

// Written by large_front_kernel each step, read by the later kernels of the step.
struct LargeCtx {
    int32_t q;                      // H rows in use this step
    int32_t rows[LARGE_QMAX];       // state index of each row (0,1,2, then 3+2l, 4+2l ascending l)
    int32_t hnnz[LARGE_QMAX];       // non-zeros of the row (1 pose, 4 range, 5 bearing: ekf.cpp:124-132)
    int32_t hcol[LARGE_QMAX][5];    // their state columns ...
    int32_t hk[LARGE_QMAX][5];      // ... and the H-row index k with rows[k] == that column
    double hval[LARGE_QMAX][5];
    double y[LARGE_QMAX];           // innovation, theta/bearing rows wrapped (ekf.cpp:302-307)
    double rdiag[LARGE_QMAX];       // diagonal of R for these rows
    int32_t assoc[ASLAM_MAX_OBS];
    uint32_t status;
};

// Fused exchange of the row-sharded large-map EKF: every rank maps the exchange block of every
// other rank (CUDA IPC over NVLink, or plain pointers inside one process) and the gather kernel stores each
// value it owns straight into all of them; one arrival counter per (destination, source) pair orders the step.
constexpr int LARGE_MAX_PEERS = 8;
enum { LARGE_STEP_DIRECT = 0, LARGE_STEP_EXCH = 1, LARGE_STEP_PEER = 2 };
struct PeerSet {
    double* E[LARGE_MAX_PEERS];                 // this step's (parity-selected) exchange buffer of rank p
    unsigned long long* flag[LARGE_MAX_PEERS];  // rank p's arrival counters, indexed by source rank
    int32_t world, rank;
};

}  // namespace aslam

struct aslam_large {
    int kind = 0;
    aslam_config cfg;
    int device = 0;
    cudaStream_t stream = nullptr;
    uint32_t n = 0;                  // state dimension 3 + 2 L
    uint32_t row_begin = 0, row_end = 0;
    size_t ldp = 0;                  // leading dimension of the local column-major P (rows, padded to 16)
    double* P = nullptr;             // [n][ldp]: column j at P + j*ldp holds rows row_begin..row_end-1
    // EKF, overlapped steps (ekf_large.cu: large_ekf_phase2_fast): the streaming update writes OUT of place, P <-> P_alt
    // swap every step, and the solve kernel of the next step runs on `solve_stream` under the update in flight.
    // h->P is always the buffer that holds the latest state once everything enqueued on `stream` has completed.
    double* P_alt = nullptr;         // second covariance buffer (allocated on the first overlapped step)
    cudaStream_t solve_stream = nullptr;
    cudaEvent_t ev_before_upd = nullptr;  // recorded on `stream` just before the last update: everything older is done
    cudaEvent_t ev_now = nullptr, ev_solve = nullptr;
    bool pending = false;            // an out-of-place update is in flight: P_alt + K/hp/predA[kbuf] describe h->P
    int kbuf = 0;                    // which half of K / hp / predA the last solve kernel wrote
    bool no_overlap = false;         // ASLAM_LARGE_NO_OVERLAP: keep solve and update on one stream, in place
    size_t overlap_min_bytes = (size_t)192 << 20;  // shards whose update streams less than this do not share SMs ...
    size_t reserve_min_bytes = (size_t)32 << 20;   // ... but, above this, overlap with SMs set aside for the solve kernel
    int reserve_sms = -1;            // -1: ceil(n / 256) + 2; 0: never; > 0: that many (ASLAM_LARGE_RESERVE_SMS)
    double* X = nullptr;             // [n]
    double* A = nullptr;             // [2]: A(0,0), A(1,0) (ekf.cpp:210-211)
    double* predA = nullptr;         // [2][2]: the A entries of the last two steps, for the predict fused into the update
    aslam::LargeCtx* ctx = nullptr;
    float* d_obs = nullptr;          // [2][ASLAM_MAX_OBS] range, bearing (UKF; the EKF front takes them as arguments)
    float obs_host[2 * ASLAM_MAX_OBS] = {};   // EKF: the observations of the message in flight
    unsigned long long* sync = nullptr;       // EKF: step-ordering counters in device memory (ekf_large.cu SYNC_*)
    uint64_t step_epoch = 0;                  // EKF: steps enqueued so far
    unsigned long long* trace = nullptr;      // EKF diagnostics (ASLAM_LARGE_TRACE=1): [64][8] %globaltimer stamps
    double* exch = nullptr;          // [2*LARGE_QMAX][n] raw rows of P, then P H^T slices
    double* hp = nullptr;            // [n][QP]   H P, k contiguous per column
    double* K = nullptr;             // [QP][n]   gain, row k contiguous
    uint32_t q_host = 0;             // rows in use of the step in flight (after phase1)
    uint32_t m_last = 0;
    uint64_t launches = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;  // around the streaming update kernel of the last step
    // EKF: the front kernel of step k+1 (association, pose predict, H rows; it needs X and the context, not P) runs
    // on `side`, so it overlaps the streaming update of step k; ordering through `sync` (fast path) / ev_front (q > 20)
    cudaStream_t side = nullptr;
    cudaEvent_t ev_ctx_free = nullptr, ev_front = nullptr;
    bool timed = false;              // CUDA events around the streaming update (aslam_large_set_timing)
    bool pdl = true;                 // launch the streaming update with programmatic stream serialization
    int step_mode = 0;               // LARGE_STEP_* of the step in flight (phase1 -> phase2)
    // fused peer exchange: block = [E parity 0][E parity 1][LARGE_MAX_PEERS arrival counters]
    unsigned char* peer_block = nullptr;        // own block (cudaMalloc)
    unsigned char* peer_base[aslam::LARGE_MAX_PEERS] = {};  // every rank's block as mapped here
    bool peer_ipc[aslam::LARGE_MAX_PEERS] = {};  // mapped with cudaIpcOpenMemHandle (to be closed)
    uint32_t peer_world = 0, peer_rank = 0;
    bool peer_same_device = false;              // some peer shard lives on this handle's own GPU
    uint64_t peer_epoch = 0;                    // fused steps taken so far (all ranks step in lockstep)
    // UKF large-map state lives in ukf_large.cu's own struct, reached through `ukf`
    void* ukf = nullptr;
};
