/*
 * aslam_c.h -- C ABI of the B200-native EKF/UKF-SLAM filter core (libaslam_b200.so).
 *
 * This is the drop-in boundary for the filter hot path of iamarkaj/AwesomeSLAM. The
 * reference has no FFI layer: its "API" is two ROS node classes whose private methods are
 * driven by callbacks. Each entry point below names the reference member it replaces
 * (paths relative to awesome_slam/ in the reference tree).
 *
 *   aslam_config                 <-> include/awesome_slam/config.h:39-65 (compile-time consts)
 *   aslam_{ekf,ukf}_create       <-> EKFSlam::EKFSlam + initialize   src/ekf/ekf.cpp:38-71
 *                                    UKFSlam::UKFSlam + initialize   src/ukf/ukf.cpp:38-67
 *   aslam_filter_set_observations<-> cbSensorLandmark                ekf.cpp:102-114, ukf.cpp:98-110
 *   aslam_filter_step            <-> cbOdom (updateZandA/updateZ, X=Z once, slam)
 *                                                                    ekf.cpp:74-99,  ukf.cpp:70-95
 *   aslam_filter_get_state       <-> param.X / param.P / N / corr_id ekf.h:56-67, ukf.h:56-82
 *   aslam_filter_get_landmarks   <-> convertToLandmarkMsg            include/awesome_slam/common.h:93-108
 *   aslam_large_*                <-> the same slam() math (ekf.cpp:293-311, ukf.cpp:260-392) for
 *                                    one filter whose map is too large for the batched kernels
 *
 * Conventions
 *   - Plain C types only; no exceptions cross the boundary. Every call returns aslam_status.
 *   - Host pointers unless the name ends in _device. Matrices are COLUMN-MAJOR (Eigen's
 *     MatrixXd layout) with an explicit leading dimension.
 *   - A handle owns its device memory and one CUDA stream; calls are enqueued in order and
 *     are asynchronous unless they return data to host memory. A handle is not thread-safe;
 *     distinct handles may be used from distinct threads.
 *   - There is NO CPU fallback: without a usable CUDA device create() fails with
 *     ASLAM_ERR_CUDA.
 */
#ifndef ASLAM_C_H
#define ASLAM_C_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int32_t aslam_status;
enum {
    ASLAM_OK = 0,
    ASLAM_ERR_ARG = -1,         /* null pointer / size out of range */
    ASLAM_ERR_CUDA = -2,        /* CUDA runtime error: see aslam_last_error() */
    ASLAM_ERR_UNSUPPORTED = -3, /* e.g. max_state_dim too large for the batched kernels */
    ASLAM_ERR_NOT_READY = -4,   /* step before the first set_observations (ekf.cpp:76) */
    ASLAM_ERR_TIMEOUT = -5      /* aslam_large_status: a fused-exchange step gave up waiting for a peer */
};

/* Per-filter status bits (reference failure modes that are silent there) */
enum {
    ASLAM_ST_CAP_REJECTED = 1u << 0,  /* updateNewLandmark dropped a batch (ekf.cpp:263-268) */
    ASLAM_ST_LLT_FAILED = 1u << 1,    /* UKF: Paug not positive definite this step (ukf.cpp:280) */
    ASLAM_ST_NONFINITE = 1u << 2,     /* X or P holds a NaN/Inf */
    ASLAM_ST_WAIT_OVERFLOW = 1u << 3, /* wait-list exceeded ASLAM_MAX_WAIT (reference: unbounded) */
    ASLAM_ST_OBS_OVERFLOW = 1u << 4,  /* more than ASLAM_MAX_OBS observations in one message */
    ASLAM_ST_PEER_TIMEOUT = 1u << 5,  /* fused exchange: a peer's rows did not arrive within the spin bound */
    ASLAM_ST_STEP_TIMEOUT = 1u << 6   /* large-map EKF: a kernel gave up waiting for the previous kernel of the step chain */
};

#define ASLAM_MAX_WAIT 256u /* capacity of the per-filter new-landmark wait-list */
#define ASLAM_MAX_OBS 64u  /* observations per filter per message */
#define ASLAM_BATCH_MAX_DIM 64 /* largest max_state_dim the batched kernels accept */

/* config.h:39-65. Defaults (aslam_config_default) equal the reference's constants. */
typedef struct aslam_config {
    float pi;                   /* PI 3.141592654f                    config.h:39 */
    float min_dist_thresh;      /* MIN_DIST_THRESH 0.5                config.h:43 */
    int32_t min_landmark_occ;   /* MIN_LANDMARK_OCC 10                config.h:44 */
    int32_t max_state_dim;      /* MAX_LANDMARK_COUNT 30 (compared with the STATE dim N) :45 */
    float ukf_std_a;            /* config.h:54 */
    float ukf_std_yaw;          /* config.h:55 */
    float ukf_kp_robot_pose;    /* config.h:56 */
    float ukf_kp_landmark_pose; /* config.h:57 */
    float ukf_kr;               /* config.h:58 */
    float ukf_kq;               /* config.h:59 */
    float ekf_kp_robot_pose;    /* config.h:62 */
    float ekf_kp_landmark_pose; /* config.h:63 (dead in the reference) */
    float ekf_kr;               /* config.h:64 */
    float ekf_kq;               /* config.h:65 */
} aslam_config;

/* One nav_msgs/Odometry per filter, structure-of-arrays over the batch. Only the fields the
 * reference reads (ekf.cpp:94,139-142,206-209). `yaw` is quat2euler(w,x,y,z) (tools.h:62-66)
 * evaluated by the caller -- aslam_quat2euler() does it -- and `dt` is delta_time of
 * ekf.cpp:80, both already rounded to float as in the reference. */
typedef struct aslam_odom {
    const double* x;   /* pose.pose.position.x   [batch] */
    const double* y;   /* pose.pose.position.y   [batch] */
    const float* yaw;  /* quat2euler(orientation) [batch] */
    const double* vx;  /* twist.twist.linear.x   [batch] */
    const double* az;  /* twist.twist.angular.z  [batch] */
    const float* dt;   /* delta_time             [batch] */
} aslam_odom;

typedef struct aslam_filter aslam_filter; /* batched EKF or UKF filters */
typedef struct aslam_large aslam_large;   /* one large-map filter */

enum { ASLAM_KIND_EKF = 0, ASLAM_KIND_UKF = 1 };

void aslam_config_default(aslam_config* cfg);
const char* aslam_last_error(void);
const char* aslam_version(void);
int32_t aslam_device_count(void);

/* Measured throughput of the FP64 FMA pipe of `device` in TFLOP/s (roofline denominator of the
 * batched kernels; not part of the filter path). */
aslam_status aslam_probe_fp64_tflops(int32_t device, double* tflops);

/* tools.h:62-66 quat2euler: arguments rounded to float, atan2f. Host-side helper. */
float aslam_quat2euler(double w, double x, double y, double z);

/* ---------------------------------------------------------------- batched filters -------- */
aslam_status aslam_ekf_create(const aslam_config* cfg, uint32_t batch, int32_t device, aslam_filter** out);
aslam_status aslam_ukf_create(const aslam_config* cfg, uint32_t batch, int32_t device, aslam_filter** out);
aslam_status aslam_filter_destroy(aslam_filter* h);
aslam_status aslam_filter_sync(aslam_filter* h);
/* cudaStream_t of the handle, for callers that enqueue their own copies around the calls */
void* aslam_filter_stream(aslam_filter* h);
uint32_t aslam_filter_batch(const aslam_filter* h);
uint32_t aslam_filter_ld(const aslam_filter* h); /* padded state dim used for X/P export */
int32_t aslam_filter_kind(const aslam_filter* h);

/* cbSensorLandmark. offsets has batch+1 entries; filter b owns observations
 * [offsets[b], offsets[b+1]) of range[] / bearing[]. The list persists until replaced. */
aslam_status aslam_filter_set_observations(aslam_filter* h, const float* range, const float* bearing,
                                           const uint32_t* offsets);
/* Same with device-resident arrays (no copy; they must stay valid until the next call). */
aslam_status aslam_filter_set_observations_device(aslam_filter* h, const float* d_range,
                                                  const float* d_bearing, const uint32_t* d_offsets);

/* cbOdom for every filter of the batch: association, wait-list, growth, predict, update. */
aslam_status aslam_filter_step(aslam_filter* h, const aslam_odom* odom);
aslam_status aslam_filter_step_device(aslam_filter* h, const aslam_odom* d_odom);

/* Replay `steps` consecutive messages that are already resident in HBM, step-major:
 * odometry arrays [steps][batch], observations [steps][batch][obs_per_filter] (every filter
 * sees obs_per_filter observations per step). Equivalent to `steps` x (set_observations_device
 * + step_device) but keeps each filter's covariance on-chip between steps. */
aslam_status aslam_filter_run_device(aslam_filter* h, uint32_t steps, const aslam_odom* d_odom,
                                     const float* d_range, const float* d_bearing, uint32_t obs_per_filter);

/* Same replay from HOST buffers (the end-to-end entry point): copies the `steps` messages to the
 * device, then runs them as one launch. Asynchronous: the copy is enqueued on a second stream into
 * one of two staging blocks, so the transfer of call k+1 overlaps the kernel of call k; the host
 * buffers must stay valid until aslam_filter_sync (pinned memory for a truly asynchronous copy). */
aslam_status aslam_filter_run_host(aslam_filter* h, uint32_t steps, const aslam_odom* odom, const float* range,
                                   const float* bearing, uint32_t obs_per_filter);
/* Number of filter-step kernels this handle has launched (bench.py's gpu_launches). */
uint64_t aslam_filter_launch_count(aslam_filter* h);
/* UKF: fail_k [batch] = first non-positive pivot column of Paug in the LAST step, -1 if it
 * factorised (Eigen's silent LLT failure, ukf.cpp:280). */
aslam_status aslam_filter_get_llt_fail(aslam_filter* h, int32_t* fail_k);

/* State export. Any pointer may be NULL. ld = aslam_filter_ld().
 *   X [batch][ld], P [batch][ld*ld] column-major with leading dimension ld, entries beyond a
 *   filter's N are zero; N [batch]; status [batch];
 *   assoc [total observations of the last message]: landmark slot j (0,2,4,.. as corr_id in
 *   ekf.cpp:160-173) the observation was written to, or -1 when it went to the wait-list. */
aslam_status aslam_filter_get_state(aslam_filter* h, double* X, double* P, uint32_t* N, int32_t* assoc,
                                    uint32_t* status);
/* param.X only, into a caller buffer [batch][ld] (pinned memory for the end-to-end path). */
aslam_status aslam_filter_read_state(aslam_filter* h, double* X);
/* Pipelined form of the same read (the publish half of cbOdom, ekf.cpp:97-98, without stalling the
 * next message): returns at once; X is complete after aslam_filter_read_wait(h, *ticket) or
 * aslam_filter_sync. At most two reads may be outstanding per destination buffer pair. */
aslam_status aslam_filter_read_state_async(aslam_filter* h, double* X, uint64_t* ticket);
aslam_status aslam_filter_read_wait(aslam_filter* h, uint64_t ticket);
/* Z [batch][ld]; EKF only: a00/a10 [batch] = A(0,0), A(1,0) (ekf.cpp:210-211) */
aslam_status aslam_filter_get_aux(aslam_filter* h, double* Z, double* a00, double* a10);
/* wait-list export: count [batch]; range/bearing/hits [batch][ASLAM_MAX_WAIT] */
aslam_status aslam_filter_get_wait(aslam_filter* h, uint32_t* count, float* range, float* bearing,
                                   uint32_t* hits);
/* convertToLandmarkMsg: x,y [batch][(ld-3)/2]; count [batch] */
aslam_status aslam_filter_get_landmarks(aslam_filter* h, double* x, double* y, uint32_t* count);
/* Device pointers of the live state for zero-copy consumers: X [batch][ld], N [batch] */
aslam_status aslam_filter_device_state(aslam_filter* h, const double** d_X, const uint32_t** d_N);

/* Checkpoint / start from a supplied state (the reference has none).
 * load_state sets N, X, Z, P for every filter, marks it initialised (init_x = init_z = false) and RESETS everything
 * else that outlives a message: empty wait-list, A = I, status 0, no LLT failure -- nothing of an earlier run on the
 * same handle leaks into the loaded state.
 * load_aux restores the rest of a checkpoint taken with get_aux / get_wait / get_state (call it after load_state):
 * the two persistent entries of A (ekf.cpp:206-211 keeps them when vx or az is 0; EKF handles only), the wait-list
 * with its hit counts (ekf.h:104; arrays laid out as get_wait returns them) and the status bits. Groups passed as
 * NULL keep their reset value. A run resumed from load_state + load_aux continues bit-identically to the
 * uninterrupted one. */
aslam_status aslam_filter_load_state(aslam_filter* h, const uint32_t* N, const double* X, const double* Z,
                                     const double* P);
aslam_status aslam_filter_load_aux(aslam_filter* h, const double* a00, const double* a10, const uint32_t* wcount,
                                   const float* wrange, const float* wbearing, const uint32_t* whits,
                                   const uint32_t* status);

/* ---------------------------------------------------------------- large-map filter ------- */
/* One filter with n = 3 + 2 * landmarks state entries; P is n x n FP64 in HBM.
 * Sharding: this handle owns covariance rows [row_begin, row_end) (pass 0, n for one GPU). */
aslam_status aslam_large_ekf_create(const aslam_config* cfg, uint32_t n, uint32_t row_begin, uint32_t row_end,
                                    int32_t device, aslam_large** out);
aslam_status aslam_large_ukf_create(const aslam_config* cfg, uint32_t n, int32_t device, aslam_large** out);
aslam_status aslam_large_destroy(aslam_large* h);
aslam_status aslam_large_sync(aslam_large* h);
void* aslam_large_stream(aslam_large* h);
/* X [n], Z [n] (may be NULL: taken as h(X)), P column-major n x n with leading dimension ldp
 * (the handle copies its own rows). */
aslam_status aslam_large_load_state(aslam_large* h, const double* X, const double* Z, const double* P, size_t ldp);
aslam_status aslam_large_get_state(aslam_large* h, double* X, double* P, size_t ldp);
/* cbSensorLandmark + cbOdom for one message pair. m observations; assoc_out[m] (may be NULL)
 * receives the landmark slot or -1. EKF: row-subset joint update over the pose rows and the
 * rows of the landmarks observed in THIS message; with every
 * landmark observed it equals ekf.cpp:300-310. UKF: the full ukf.cpp:260-392 step.
 * EKF with at most 8 observed landmarks per message (q <= 20): consecutive steps OVERLAP -- the update streams P out
 * of place between two covariance buffers (the second one, as large as the first, is allocated on the first such step;
 * if that fails the steps run in place) and the small kernels of the next step run under it. The calls only enqueue;
 * everything that reads the state afterwards (get_state, sync, work on aslam_large_stream) is ordered behind the
 * last update. ASLAM_LARGE_NO_OVERLAP=1 in the environment keeps the steps in place on one stream (same kernels). */
aslam_status aslam_large_step(aslam_large* h, double x, double y, float yaw, double vx, double az, float dt,
                              const float* range, const float* bearing, uint32_t m, int32_t* assoc_out);
/* Sharded EKF only (row_end - row_begin < n): the step is split around the one exchange.
 *   phase1: association, predict, then this rank's contribution to the exchange buffer
 *           [2q][n] doubles (q = 3 + 2m rows laid out, returned in q_out): rows 0..q-1 hold the
 *           rows of P named by H (pose + observed landmarks) where this rank owns them, rows
 *           q..2q-1 its slice of P H^T; everything it does not own is zero. The caller sums the
 *           buffers (aslam_large_exchange_buffer) of all ranks in place -- NCCL all-reduce over
 *           NVLink; every element has exactly one non-zero owner, so the sum is exact.
 *   phase2: S, K for the owned rows, X update, P_local -= K_local * (H P). */
aslam_status aslam_large_step_phase1(aslam_large* h, double x, double y, float yaw, double vx, double az,
                                     float dt, const float* range, const float* bearing, uint32_t m,
                                     uint32_t* q_out);
aslam_status aslam_large_exchange_buffer(aslam_large* h, double** d_buf, size_t* count);
aslam_status aslam_large_step_phase2(aslam_large* h, int32_t* assoc_out);
/* Fused form of the same exchange (no NCCL call, no host round trip, ONE kernel between the front and the
 * streaming update): every rank's solve kernel stores the values it owns straight into the exchange block of EVERY
 * rank over NVLink peer memory and bumps an arrival counter there, then spins (bounded) until all sources have
 * arrived. When the bound expires the step is abandoned cleanly -- K = 0: the filter applies the predict only, the
 * state stays valid -- ASLAM_ST_PEER_TIMEOUT is raised and aslam_large_status returns ASLAM_ERR_TIMEOUT.
 * All argument and attachment checks are made before the first launch: an error return leaves the handle untouched.
 * Set-up, once per handle:
 *   peer_export: allocates this handle's exchange block, returns its CUDA IPC handle (64 bytes) and,
 *                for ranks living in the same process, the raw device pointer (either may be NULL);
 *   peer_attach: `ipc_handles` = world x 64 bytes gathered from all ranks (other processes), or
 *                `blocks` = world device pointers (same process); entry `rank` is this handle itself.
 * All ranks must have attached before the first fused step (barrier), step in lockstep, and must not
 * destroy a handle while a peer may still write to it. Supports m <= 18 observations per message. */
aslam_status aslam_large_peer_export(aslam_large* h, void* ipc_handle64, void** block);
aslam_status aslam_large_peer_attach(aslam_large* h, uint32_t world, uint32_t rank, const void* ipc_handles,
                                     void* const* blocks);
aslam_status aslam_large_step_fused(aslam_large* h, double x, double y, float yaw, double vx, double az, float dt,
                                    const float* range, const float* bearing, uint32_t m, int32_t* assoc_out);
/* Kernels this handle has launched so far. */
uint64_t aslam_large_launch_count(aslam_large* h);
/* Device time (CUDA events on the handle's stream) of the dominant kernel of the LAST step: the
 * streaming covariance update (EKF) / the whole step's GEMM chain (UKF). Waits for that kernel.
 * EKF handles record the events only after aslam_large_set_timing(h, 1): an event between the solve kernel and the
 * update keeps the update from being launched early (programmatic dependent launch), so a timed step is slower. */
aslam_status aslam_large_set_timing(aslam_large* h, int32_t on);
/* Diagnostics (handles created with ASLAM_LARGE_TRACE=1 in the environment, else ASLAM_ERR_UNSUPPORTED): the
 * %globaltimer stamps [64 steps][16] the EKF kernels left for the last 64 steps (csrc/ekf_large.cu, TRACE_STEPS). */
aslam_status aslam_large_trace(aslam_large* h, unsigned long long* out1024);
aslam_status aslam_large_last_kernel_ms(aslam_large* h, float* ms);
/* status bits of the filter (ASLAM_ST_*), and for the UKF the first failing LLT pivot or -1.
 * Returns ASLAM_ERR_TIMEOUT (status still written) when ASLAM_ST_PEER_TIMEOUT is set. */
aslam_status aslam_large_status(aslam_large* h, uint32_t* status, int32_t* llt_fail);

/* ---------------------------------------------------------------- FP64 dense blocks ------- */
/* The contractions of the large-map UKF (ukf.cpp:280, 307-391: Eigen LLT, rank-1 sums, inverse) as stand-alone
 * calls on DEVICE pointers, enqueued on `stream` (a cudaStream_t, NULL = default stream). Row-major, contraction
 * index contiguous; pointers 16-byte aligned, leading dimensions even.
 *   gemm_nt : C[i][j] = alpha sum_k A[i][k] B[j][k] + beta C[i][j]   (lower_only: skip tiles above the diagonal)
 *   cholesky: A <- L (lower, in place; strict upper triangle unspecified); inv_blocks [ceil(n/64)][4096] doubles;
 *             d_fail: device int, first non-positive pivot or -1
 *   trsm    : B <- B L^-T with the inv_blocks of cholesky
 * Exported for the parity tests and the tensor-pipe roofline; the filters call the same code internally. */
aslam_status aslam_dense_gemm_nt(void* stream, int32_t M, int32_t N, int32_t K, const double* d_A, size_t lda,
                                 const double* d_B, size_t ldb, double* d_C, size_t ldc, double alpha, double beta,
                                 int32_t lower_only);
aslam_status aslam_dense_cholesky(void* stream, double* d_A, size_t ld, int32_t n, double* d_inv_blocks,
                                  int32_t* d_fail);
aslam_status aslam_dense_trsm(void* stream, double* d_B, size_t ldb, int32_t m, const double* d_L, size_t ldl,
                              int32_t n, const double* d_inv_blocks);
/* Measured throughput of the FP64 tensor pipe (DMMA) of `device`, TFLOP/s. */
aslam_status aslam_probe_dmma_tflops(int32_t device, double* tflops);
/* TF32 variant of the sigma-point products of the large-map UKF (north_star: "FP64/TF32 tensor-core tiles via TMA"):
 * tcgen05.mma kind::tf32 with the accumulator in tensor memory, operands by TMA. mode 0 = FP64 (default, the parity
 * path), 1 = TF32, 3 = error-compensated 3xTF32 (hi/lo split). The sigma-point mean, the w0 rank-1 terms, the
 * factorisations and the solves stay FP64 in every mode. Tolerances are stated separately (tests/util.py). */
aslam_status aslam_large_ukf_set_precision(aslam_large* h, int32_t mode);
/* CUDA-event time (ms) of the three sigma-point products of the last step, in every mode (TF32: split included). */
aslam_status aslam_large_ukf_products_ms(aslam_large* h, float* ms);
/* The TF32 GEMM by itself on FP64 DEVICE matrices (row-major, contraction index contiguous): C = alpha A B' + beta C,
 * the operands are split into TF32 hi / lo parts internally. passes = 1 or 3. Synchronous. */
aslam_status aslam_dense_gemm_nt_tf32(void* stream, int32_t M, int32_t N, int32_t K, const double* d_A, size_t lda,
                                      const double* d_B, size_t ldb, double* d_C, size_t ldc, double alpha, double beta,
                                      int32_t lower_only, int32_t passes);
/* aslam_dense_gemm_nt_tf32 for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it). NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same kernels in the same order, same result; the scratch (FP32 hi / lo copies of both operands + the
 * pipeline's error flag) lives in a file-scope static __device__ array of ASLAM_TF32_STATIC_SCRATCH_BYTES that every
 * GPU loading the library holds until exit, so the call allocates and frees nothing. Runs on the CURRENT device.
 * Limit: (round256(4 M Kp) + round256(4 N Kp)) * (passes == 3 ? 2 : 1) <= ASLAM_TF32_STATIC_SCRATCH_BYTES with
 * Kp = K rounded up to 32 (e.g. M = N = 512, K = 1024, passes 3; or M = N = 1024, K = 1024, passes 1); beyond it the
 * call returns ASLAM_ERR_UNSUPPORTED with the sizes in aslam_last_error() and touches nothing.
 * Every call on a GPU shares the one array: calls that overlap on the same GPU (two streams, two host threads) RACE
 * on it -- serialise them. Synchronous like the allocating form. */
#define ASLAM_TF32_STATIC_SCRATCH_BYTES (8u * 1024u * 1024u)
aslam_status aslam_dense_gemm_nt_tf32_static(void* stream, int32_t M, int32_t N, int32_t K, const double* d_A, size_t lda,
                                             const double* d_B, size_t ldb, double* d_C, size_t ldc, double alpha,
                                             double beta, int32_t lower_only, int32_t passes);
/* Measured kind::tf32 tensor throughput of `device` (MMAs issued back to back from shared memory), TFLOP/s. */
aslam_status aslam_probe_tf32_tflops(int32_t device, double* tflops);

/* ---------------------------------------------------------------- laser front end -------- */
/* The step in front of the filters: the reference's sensor_landmark node,
 * awesome_slam/src/sensor_landmark/sensor_landmark.cpp -- LandMarks::callback (:59-133), bearing2pose
 * (:136-142), circleClassification (:147-186), circleFitting (:192-298) and Circle::toLaserData
 * (include/awesome_slam/structures.h:126-133) -- for a batch of LaserScans: 360 ranges per scan (one ray per
 * degree, +inf = no return, as Gazebo publishes them) in, the (range, bearing) list of the fitted cylinders out,
 * in scan order, exactly what cbSensorLandmark (src/ekf/ekf.cpp:102-114) receives. */
#define ASLAM_LASER_RAYS 360u
#define ASLAM_LASER_MAX_LANDMARKS 32u /* output slots per scan */
typedef struct aslam_laser aslam_laser;
typedef struct aslam_laser_config { /* include/awesome_slam/config.h:40-51 */
    float deg2rad;                  /* DEG2RAD */
    float min_dist_thresh;          /* MIN_DIST_THRESH */
    float std_max;                  /* STD */
    float min_mean;                 /* MIN_MEAN */
    float max_mean;                 /* MAX_MEAN */
    int32_t min_cluster_points;     /* MIN_CLUSTER_POINTS */
} aslam_laser_config;
void aslam_laser_config_default(aslam_laser_config* c);
/* LandMarks::LandMarks + initialize (:38-57): the sin / cos tables are built on the host, as the reference does. */
aslam_status aslam_laser_create(const aslam_laser_config* cfg, uint32_t max_scans, int32_t device, aslam_laser** out);
aslam_status aslam_laser_destroy(aslam_laser* h);
/* LandMarks::callback for nscan scans, host buffers: ranges [nscan][360]; range_out / bearing_out
 * [nscan][ASLAM_LASER_MAX_LANDMARKS]; counts [nscan] = landmarks published for the scan, or -1 where the reference's
 * assert(theta < 359) (:82) fires (first and last ray on the same object). Synchronous. */
aslam_status aslam_laser_extract(aslam_laser* h, const float* ranges, uint32_t nscan, float* range_out,
                                 float* bearing_out, int32_t* counts);
/* Same with device buffers, asynchronous on the handle's stream (aslam_laser_sync / aslam_laser_stream). nscan may
 * exceed max_scans: the batch is then processed in chunks of max_scans (the size of the handle's work list). Two
 * kernel launches per chunk (classification, then the circle fit of the accepted clusters). */
aslam_status aslam_laser_extract_device(aslam_laser* h, const float* d_ranges, uint32_t nscan, float* d_range_out,
                                        float* d_bearing_out, int32_t* d_counts);
/* Synthetic LaserScans, standing in for Gazebo's LDS plugin: 360 rays, one per degree from the
 * sensor's yaw, against cylinders of `radius` at (lm_x, lm_y)[L] (models/awesome_cylinder/model.sdf:18-22: 0.25 m);
 * returns inside 0.12 .. 3.5 m get sigma * noise added (turtlebot3_burger/model.sdf:94-119: sigma 0.01), every other
 * ray is +inf. poses [nscan][3] = x, y, yaw (double); noise [nscan][360] standard-normal floats or NULL; ranges
 * [nscan][360]. Host buffers / device buffers (asynchronous on the handle's stream). */
aslam_status aslam_laser_synthesize(aslam_laser* h, const double* lm_x, const double* lm_y, uint32_t L,
                                    const double* poses, const float* noise, uint32_t nscan, double sigma,
                                    double radius, float* ranges);
aslam_status aslam_laser_synthesize_device(aslam_laser* h, const double* d_lm_x, const double* d_lm_y, uint32_t L,
                                           const double* d_poses, const float* d_noise, uint32_t nscan, double sigma,
                                           double radius, float* d_ranges);
aslam_status aslam_laser_sync(aslam_laser* h);
void* aslam_laser_stream(aslam_laser* h);
uint64_t aslam_laser_launch_count(aslam_laser* h);

/* ---------------------------------------------------------------- synthetic replay ------- */
/* Deterministic replay generator of SURVEY.md sec. 8d (host code, no GPU needed):
 * map (lm_x, lm_y)[L]; `warm` stationary steps then vx, az constant, dt = 1.0f;
 * noise sigma: odom x,y,yaw = 0.01, range = 0.01, bearing = 0.005; per-filter RNG
 * mt19937_64(seed0 + filter) + Box-Muller. Output is step-major:
 *   ox, oy, vx, az: double [steps][batch]; yaw, dt: float [steps][batch];
 *   range, bearing: float [steps][batch][L] listed in increasing true bearing. */
aslam_status aslam_replay_generate(const double* lm_x, const double* lm_y, uint32_t L, uint32_t batch,
                                   uint32_t steps, uint32_t warm, double vx, double az, uint64_t seed0,
                                   uint32_t first_filter, double* ox, double* oy, float* yaw, double* ovx,
                                   double* oaz, float* dt, float* range, float* bearing);
/* The same replay, additionally returning the orientation quaternion (w, 0, 0, z) of every odometry message
 * (qw, qz: double [steps][batch], may be null) -- what nav_msgs/Odometry carries and ekf.cpp:141 feeds to
 * quat2euler; `yaw` is aslam_quat2euler(qw, 0, 0, qz). Both entry points are also built into the host-only
 * library awesomeslam_b200/lib/libaslam_host.so (no CUDA), together with aslam_quat2euler. */
aslam_status aslam_replay_generate_quat(const double* lm_x, const double* lm_y, uint32_t L, uint32_t batch,
                                        uint32_t steps, uint32_t warm, double vx, double az, uint64_t seed0,
                                        uint32_t first_filter, double* ox, double* oy, float* yaw, double* ovx,
                                        double* oaz, float* dt, float* range, float* bearing, double* qw,
                                        double* qz);

#ifdef __cplusplus
}
#endif
#endif /* ASLAM_C_H */
