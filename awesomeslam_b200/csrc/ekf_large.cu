// Large-map EKF-SLAM step (kernels K3-K5): one filter whose covariance
// P (n x n FP64, n = 3 + 2 L) lives in HBM. Same formulas as EKFSlam::slam (src/ekf/ekf.cpp:293-311)
// and updateZandA (ekf.cpp:137-213), with H restricted to the rows in use this step -- the three
// odometry rows and the two rows of every landmark observed in THIS message (row-subset joint
// update; with every landmark observed it is ekf.cpp:300-310 verbatim):
//     S = H_o P H_o' + R_o (q x q),  K = P H_o' S^-1 (n x q),  X += K y,  P -= K (H_o P)
// A = I except A(0,0), A(1,0) (ekf.cpp:210-211) and H_o has <= 5 non-zeros per row, so everything
// but the last product is O(q n); the last one streams P once: read + write 2*8*n^2 bytes.
//
// Kernels of the path cfg4 takes (q <= 20), in launch order
//   large_front_kernel          1 CTA, side stream, under the previous step's update: association (float math,
//                               first-minimum ties), pose predict, A, H_o, y; what depends on the message only
//                               (observation projection, A) before it waits for the previous step
//   large_solve_kernel          gather of the q rows / columns of P (with the predict P = A P A' + Q applied on the fly),
//                               exchange over peer memory when sharded, S, Cholesky + inverse by the whole CTA,
//                               H_o P, K, X += K y. Two CTA shapes: 128 threads (co-resident with the update), 256
//   large_update_stream_kernel  P_out = P_in - K_local (H_o P): the HBM-bound kernel, persistent (1 CTA per SM), predict
//                               fused into the tiles
// Two forms of the step (large_ekf_phase2_fast):
//   overlapped  the update writes OUT of place (P <-> P_alt) and the solve kernel of step k+1 runs on its own stream
//               under the update of step k, computing the entries of P_k it needs from P_{k-1} and the previous K,
//               H_o P, A (PrevUpd / pv_entry): a step costs max(update, front + solve). Shards that stream >= 192 MB.
//   in place    solve and update on one stream, the update launched programmatically behind the solve kernel
//               (NCCL form, small shards, a peer shard on the same GPU, ASLAM_LARGE_NO_OVERLAP)
// 20 < q <= 131 (more than 8 observed landmarks; no BASELINE config) keeps the round-1 chain
//   large_predict_kernel -> large_gather[_fast]_kernel -> large_gain[_fast]_kernel -> large_update_{bulk,generic}_kernel
#include <cuda_runtime.h>

#include "bulk_copy.cuh"
#include "device_math.cuh"
#include "large.hpp"

namespace aslam {

namespace {

__device__ __forceinline__ unsigned long long umin64(unsigned long long a, unsigned long long b) {
    return a < b ? a : b;
}

// ------------------------------------------------------------------------------------------
// Front: ekf.cpp:139-187 (association against the supplied map; growth is outside the
// large-map contract), :206-212 (A), :295-296 (pose predict), :117-134 (H rows), :302-307 (y).
// ------------------------------------------------------------------------------------------
struct FrontArgs {
    double ox, oy, vx, az;
    float yaw, dt;
    uint32_t m;
    uint32_t q_ub;  // 3 + 2 m: rows the exchange / gain buffers are laid out for
    float pi, thresh;
    double kr_pose, kr_lm;
    unsigned long long epoch;  // 1-based index of this step on the handle
    // the message's observations travel in the kernel arguments: no staging copy in front of the launch
    float range[ASLAM_MAX_OBS], bearing[ASLAM_MAX_OBS];
};

// Ordering between the kernels of consecutive steps that run on different streams (the front of step k+1 runs on the
// side stream UNDER the streaming update of step k) is carried by two counters in device memory instead of CUDA
// events: an event recorded between the solve kernel and the update would sit between the two launches and keep the
// update from being launched programmatically behind the solve kernel.
//   sync[SYNC_FRONT]  = k once the front of step k has written the context, A and the predicted pose
//   sync[SYNC_SOLVE]  = k once the solve (or gain) kernel of step k has finished with the context and has written X
//   sync[SYNC_COUNT]  : arrival counter of the solve CTAs of the step in flight
// Every waiter was enqueued after the kernel it waits for, so the spin always terminates; it is bounded anyway.
enum { SYNC_FRONT = 0, SYNC_SOLVE = 1, SYNC_COUNT = 2, SYNC_WORDS = 4 };

__device__ __forceinline__ unsigned long long ld_acquire_gpu(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// One thread: wait until *p >= target. Returns false when the bound expires.
__device__ __forceinline__ bool spin_until(const unsigned long long* p, unsigned long long target) {
    for (int it = 0; it < 8000000; ++it) {
        if (ld_acquire_gpu(p) >= target) return true;
        __nanosleep(64);
    }
    return false;
}

// Optional timeline (ASLAM_LARGE_TRACE=1, diagnostics only): %globaltimer stamps of the kernels of each step,
// slot = (step % TRACE_STEPS) * 8 + {0 front begin, 1 front end, 2 solve first CTA in, 3 solve last CTA out,
// 4 first update CTA past its wait, 5 last update CTA out, 6 first update CTA in}.
constexpr int TRACE_STEPS = 64;
__device__ __forceinline__ unsigned long long globaltimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void trace_min(unsigned long long* tr, unsigned long long epoch, int slot) {
    if (tr) atomicMin(tr + (epoch % TRACE_STEPS) * 16 + slot, globaltimer());
}
// phase stamps of CTA 0 of the solve kernel: slots 8..15
__device__ __forceinline__ void trace_mark(unsigned long long* tr, unsigned long long epoch, int slot) {
    if (tr && blockIdx.x == 0 && threadIdx.x == 0) tr[(epoch % TRACE_STEPS) * 16 + slot] = globaltimer();
}
__device__ __forceinline__ void trace_max(unsigned long long* tr, unsigned long long epoch, int slot) {
    if (tr) atomicMax(tr + (epoch % TRACE_STEPS) * 16 + slot, globaltimer());
}

constexpr int FRONT_THREADS = 256;  // small enough to be co-resident with the streaming update (158 regs x 288 threads)

__global__ void __launch_bounds__(FRONT_THREADS) large_front_kernel(FrontArgs a, uint32_t n, double* X, double* A,
                                                                   LargeCtx* ctx, unsigned long long* sync,
                                                                   unsigned long long* trace) {
    constexpr int NW = FRONT_THREADS / 32;
    if (threadIdx.x == 0 && trace) {
        unsigned long long* sl = trace + (a.epoch % TRACE_STEPS) * 16;
        sl[0] = globaltimer();
        sl[2] = ~0ull; sl[3] = 0; sl[4] = ~0ull; sl[5] = 0; sl[6] = ~0ull;
    }
    __shared__ float s_nb[ASLAM_MAX_OBS];
    __shared__ int s_corr[ASLAM_MAX_OBS];
    __shared__ float s_mind[ASLAM_MAX_OBS];
    __shared__ int s_lm[ASLAM_MAX_OBS], s_ob[ASLAM_MAX_OBS];
    __shared__ int s_cnt;
    __shared__ double s_pose[3];
    __shared__ float2 s_pt[ASLAM_MAX_OBS];
    __shared__ unsigned long long s_red[ASLAM_MAX_OBS][NW];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t L = (n - 3) / 2;
    const double z0 = a.ox, z1 = a.oy, z2 = (double)a.yaw;
    // ---- what depends on the MESSAGE only is done before the wait for the previous step (this kernel is enqueued a
    // whole step ahead): the observations projected with the odometry pose (structures.h:104-111) and the two entries
    // of A (ekf.cpp:206-212; they persist when vx or az is zero: the previous front kernel wrote them, stream order).
    double ja00 = 0.0, ja10 = 0.0;
    if (tid < (int)a.m) {
        const float b = normalize_angle(a.bearing[tid], a.pi);  // ekf.cpp:150
        s_nb[tid] = b;
        s_pt[tid] = to_point(a.range[tid], b, z0, z1, z2);
    } else if (tid == 64) {
        ja00 = A[0];
        ja10 = A[1];
        motion_jacobian(a.vx, a.az, a.dt, z2, ja00, ja10);
    }
    if (sync) {  // X of the previous step must be final and its context no longer in use
        if (threadIdx.x == 0 && !spin_until(sync + SYNC_SOLVE, a.epoch - 1))
            atomicOr(&ctx->status, (uint32_t)ASLAM_ST_STEP_TIMEOUT);
    }
    __syncthreads();
    // (issued now, used by thread 32 after the association: one memory round trip less on the way)
    double px = 0.0, py = 0.0, pth = 0.0;
    if (tid == 32) {
        px = X[0];
        py = X[1];
        pth = X[2];
    }
    // ---- nearest landmark of every observation: every thread keeps a few landmarks in registers and scans all
    // observations against them: one pass over X, one round of warp reductions, one CTA barrier, one final reduction
    // per observation.
    constexpr int LPT = 16;  // landmarks per thread and pass: 4096 landmarks in ONE pass -- every load of X in flight at once, one
                             // round of warp reductions per observation (the kernel is on the critical path of the sharded step)
    if (lane == 0)
        for (uint32_t o = 0; o < a.m; ++o) s_red[o][warp] = ~0ull;
    __syncwarp();
    for (uint32_t l0 = 0; l0 < L; l0 += FRONT_THREADS * LPT) {
        float lx[LPT], ly[LPT];
#pragma unroll
        for (int u = 0; u < LPT; ++u) {
            const uint32_t l = l0 + u * FRONT_THREADS + tid;
            lx[u] = l < L ? (float)X[3 + 2 * l] : 0.0f;
            ly[u] = l < L ? (float)X[4 + 2 * l] : 0.0f;
        }
        for (uint32_t o = 0; o < a.m; ++o) {
            const float2 pt = s_pt[o];
            unsigned long long best = ~0ull;
#pragma unroll
            for (int u = 0; u < LPT; ++u) {
                const uint32_t l = l0 + u * FRONT_THREADS + tid;
                if (l < L) {
                    const float d = euler_distance(pt.x, pt.y, lx[u], ly[u]);
                    // d >= 0: its bit pattern orders like the value; the low word breaks ties toward the
                    // lowest landmark, which is what the strict `<` of ekf.cpp:168 keeps
                    best = umin64(best, ((unsigned long long)__float_as_uint(d) << 32) | l);
                }
            }
            for (int off = 16; off; off >>= 1) best = umin64(best, __shfl_xor_sync(0xffffffffu, best, off));
            if (lane == 0) s_red[o][warp] = umin64(s_red[o][warp], best);
        }
    }
    __syncthreads();
    for (uint32_t o = warp; o < a.m; o += NW) {
        unsigned long long best = lane < NW ? s_red[o][lane] : ~0ull;
        for (int off = 16; off; off >>= 1) best = umin64(best, __shfl_xor_sync(0xffffffffu, best, off));
        if (lane == 0) {
            s_corr[o] = (int)(best & 0xffffffffu) * 2;  // corr_id counts state entries (ekf.cpp:170)
            s_mind[o] = __uint_as_float((unsigned)(best >> 32));
        }
    }
    __syncthreads();
    if (tid == 0) {
        // ---- sequential part: which landmark keeps which observation (later messages overwrite Z)
        int cnt = 0;
        for (uint32_t o = 0; o < a.m; ++o) {
            if (L > 0 && s_mind[o] < a.thresh) {  // ekf.cpp:175
                ctx->assoc[o] = s_corr[o];
                int k = 0;
                while (k < cnt && s_lm[k] != s_corr[o]) ++k;
                if (k == cnt) {
                    s_lm[cnt] = s_corr[o];
                    s_ob[cnt] = (int)o;
                    ++cnt;
                } else {
                    s_ob[k] = (int)o;
                }
            } else {
                ctx->assoc[o] = -1;
            }
        }
        for (int i = 1; i < cnt; ++i) {  // ascending landmark order, as the oracle walks the map
            const int l = s_lm[i], o = s_ob[i];
            int k = i - 1;
            while (k >= 0 && s_lm[k] > l) {
                s_lm[k + 1] = s_lm[k];
                s_ob[k + 1] = s_ob[k];
                --k;
            }
            s_lm[k + 1] = l;
            s_ob[k + 1] = o;
        }
        s_cnt = cnt;
        ctx->q = 3 + 2 * cnt;
    } else if (tid == 64) {
        A[0] = ja00;  // (the solve kernel of the previous step has read the old values: it is done)
        A[1] = ja10;
    } else if (tid == 32) {
        // ---- the pose predict (ekf.cpp:295-296)
        double x = px, y = py, th = pth;
        motion_model(x, y, th, (float)a.vx, (float)a.az, a.dt);
        th = (double)normalize_angle((float)th, a.pi);
        s_pose[0] = x;
        s_pose[1] = y;
        s_pose[2] = th;
    }
    __syncthreads();
    const int cnt = s_cnt, q = 3 + 2 * cnt;
    const double x = s_pose[0], y = s_pose[1], th = s_pose[2];
    // ---- rows of H in use: thread k < 3 the pose rows, thread 3 + c landmark c, the rest the padding rows
    if (tid < 3) {
        const int k = tid;
        X[k] = s_pose[k];
        ctx->rows[k] = k;
        ctx->hnnz[k] = 1;
        ctx->hcol[k][0] = k;
        ctx->hk[k][0] = k;
        ctx->hval[k][0] = 1.0;
        ctx->rdiag[k] = a.kr_pose;
        ctx->y[k] = k == 0 ? a.ox - x : (k == 1 ? a.oy - y : (double)normalize_angle((float)(z2 - th), a.pi));
    } else if (tid < 3 + cnt) {
        const int c = tid - 3;
        const int i = s_lm[c];
        const double lx = X[3 + i], ly = X[4 + i];
        const HEntries h = jacobian_entries(x, y, lx, ly);
        const int kr = 3 + 2 * c, kb = kr + 1;
        ctx->rows[kr] = 3 + i;
        ctx->rows[kb] = 4 + i;
        ctx->hnnz[kr] = 4;
        ctx->hcol[kr][0] = 0;      ctx->hk[kr][0] = 0;   ctx->hval[kr][0] = h.a;
        ctx->hcol[kr][1] = 1;      ctx->hk[kr][1] = 1;   ctx->hval[kr][1] = h.c;
        ctx->hcol[kr][2] = 3 + i;  ctx->hk[kr][2] = kr;  ctx->hval[kr][2] = -h.a;
        ctx->hcol[kr][3] = 4 + i;  ctx->hk[kr][3] = kb;  ctx->hval[kr][3] = -h.c;
        ctx->hnnz[kb] = 5;
        ctx->hcol[kb][0] = 0;      ctx->hk[kb][0] = 0;   ctx->hval[kb][0] = h.b;
        ctx->hcol[kb][1] = 1;      ctx->hk[kb][1] = 1;   ctx->hval[kb][1] = h.d;
        ctx->hcol[kb][2] = 2;      ctx->hk[kb][2] = 2;   ctx->hval[kb][2] = -1.0;
        ctx->hcol[kb][3] = 3 + i;  ctx->hk[kb][3] = kr;  ctx->hval[kb][3] = -h.b;
        ctx->hcol[kb][4] = 4 + i;  ctx->hk[kb][4] = kb;  ctx->hval[kb][4] = -h.d;
        ctx->rdiag[kr] = a.kr_lm;
        ctx->rdiag[kb] = a.kr_lm;
        double rg, br;
        measure_landmark(x, y, th, lx, ly, rg, br);
        ctx->y[kr] = (double)a.range[s_ob[c]] - rg;                                      // Z - h(X)
        ctx->y[kb] = (double)normalize_angle((float)((double)s_nb[s_ob[c]] - br), a.pi);  // ekf.cpp:306
    } else if (tid < (int)a.q_ub + cnt) {  // unused rows q .. q_ub-1: harmless identity padding
        const int k = q + (tid - 3 - cnt);
        if (k < (int)a.q_ub) {
            ctx->rows[k] = -1;
            ctx->hnnz[k] = 0;
            ctx->y[k] = 0.0;
            ctx->rdiag[k] = 1.0;
        }
    }
    if (sync) {  // publish: context, A and the predicted pose of step `epoch` are in place
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            if (trace) trace[(a.epoch % TRACE_STEPS) * 16 + 1] = globaltimer();
            st_release_gpu(sync + SYNC_FRONT, a.epoch);
        }
    }
}

// Legacy chain (q > 20): marks the step's solve stage as done for the front of the next step.
__global__ void large_mark_kernel(unsigned long long* sync, unsigned long long epoch) {
    __threadfence();
    st_release_gpu(sync + SYNC_SOLVE, epoch);
}

// ------------------------------------------------------------------------------------------
// P = A P A' + Q (ekf.cpp:297) with A = I + (a00-1) e0 e0' + a10 e1 e0'. (A P) changes rows 0,1;
// (.) A' then changes columns 0,1. Thread t >= 2 owns P(0..1, t) and P(t, 0..1); thread 0 owns
// the 2x2 corner, where the column pass reads the row pass's results, and the Q diagonal.
// ------------------------------------------------------------------------------------------
__global__ void large_predict_kernel(uint32_t n, uint32_t rb, uint32_t re, size_t ldp, double* P, const double* A,
                                     double kq) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const double a00 = A[0], a10 = A[1];
    const bool own01 = rb == 0;  // the shard with row 0 also has rows 1, 2
    if (t >= 2) {
        if (own01) {  // rows 0,1 of column t
            double* c = P + (size_t)t * ldp;
            const double p0 = c[0], p1 = c[1];
            c[0] = a00 * p0;
            c[1] = a10 * p0 + p1;
        }
        if (t >= rb && t < re) {  // row t of columns 0,1
            const double p0 = P[t - rb], p1 = P[ldp + (t - rb)];
            P[t - rb] = p0 * a00;
            P[ldp + (t - rb)] = p0 * a10 + p1;
        }
        if (t == 2 && own01) P[2 * ldp + 2] += kq;
    } else if (t == 0 && own01) {
        double p00 = P[0], p10 = P[1], p01 = P[ldp], p11 = P[ldp + 1];
        // rows: row0 *= a00; row1 = a10*row0 + row1 (old row0)
        const double r00 = a00 * p00, r01 = a00 * p01;
        const double r10 = a10 * p00 + p10, r11 = a10 * p01 + p11;
        // cols: col0 *= a00; col1 = col0*a10 + col1
        p00 = r00 * a00;
        p01 = r00 * a10 + r01;
        p10 = r10 * a00;
        p11 = r10 * a10 + r11;
        P[0] = p00 + kq;
        P[1] = p10;
        P[ldp] = p01;
        P[ldp + 1] = p11 + kq;
    }
}

// ------------------------------------------------------------------------------------------
// Exchange buffer: E[k][j], k < q_ub: row rows[k] of P when this shard owns it (else 0);
// E[q_ub + k][i]: (P H_o')[i][k] for the owned rows i (else 0). Sums follow the oracle's order.
// ------------------------------------------------------------------------------------------
__global__ void large_gather_kernel(uint32_t n, uint32_t rb, uint32_t re, size_t ldp, const double* P,
                                    const LargeCtx* ctx, uint32_t q_ub, double* E) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int q = ctx->q;
    const bool mine = t >= rb && t < re;
    for (uint32_t k = 0; k < q_ub; ++k) {
        double raw = 0.0, pht = 0.0;
        if ((int)k < q) {
            const uint32_t r = (uint32_t)ctx->rows[k];
            if (r >= rb && r < re) raw = P[(size_t)t * ldp + (r - rb)];
            if (mine) {
                const int nnz = ctx->hnnz[k];
                for (int e = 0; e < nnz; ++e) pht += P[(size_t)ctx->hcol[k][e] * ldp + (t - rb)] * ctx->hval[k][e];
            }
        }
        E[(size_t)k * n + t] = raw;
        E[(size_t)(q_ub + k) * n + t] = pht;
    }
}

// Exchange buffer for q <= QT (QT = 20 or 40): same values as large_gather_kernel, arranged for latency (context in
// shared memory, 2 q independent loads per thread, H_o's fixed row structure with static register indices).
// PEER: instead of the local buffer (zeros where this rank owns nothing, summed by the caller), every value this
// rank owns is stored into the exchange buffer of EVERY rank (peer memory over NVLink), then the CTA publishes its
// stores (system-scope fence) and bumps this rank's arrival counter on every destination.
template <int QT, bool PEER>
__global__ void __launch_bounds__(256) large_gather_fast_kernel(uint32_t n, uint32_t rb, uint32_t re, size_t ldp,
                                                               const double* __restrict__ P,
                                                               const LargeCtx* __restrict__ ctx, uint32_t q_ub,
                                                               double* __restrict__ E, const PeerSet ps) {
    __shared__ double s_hval[QT][5];
    __shared__ int s_rows[QT];
    const int q = ctx->q;
    const int tid = threadIdx.x;
    if (tid < QT) {
        const bool in = tid < q;
        s_rows[tid] = in ? ctx->rows[tid] : 0;
#pragma unroll
        for (int e = 0; e < 5; ++e) s_hval[tid][e] = in ? ctx->hval[tid][e] : 0.0;
    }
    __syncthreads();
    const uint32_t t = blockIdx.x * 256 + tid;
    if (t < n) {
    const bool mine = t >= rb && t < re;
    double raw[QT], colv[QT];
#pragma unroll
    for (int k = 0; k < QT; ++k) {
        const uint32_t r = (uint32_t)s_rows[k];
        raw[k] = (k < q && r >= rb && r < re) ? P[(size_t)t * ldp + (r - rb)] : 0.0;   // P(rows[k], t)
        colv[k] = (k < q && mine) ? P[(size_t)r * ldp + (t - rb)] : 0.0;               // P(t, rows[k])
    }
    double pht[QT];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        double sum = 0.0;
        sum += colv[k] * s_hval[k][0];
        pht[k] = sum;
    }
#pragma unroll
    for (int c = 0; c < (QT - 3) / 2; ++c) {
        const int kr = 3 + 2 * c, kb = kr + 1;
        double sr = 0.0, sb = 0.0;  // entries of the row left to right, as the generic kernel sums them
        sr += colv[0] * s_hval[kr][0];
        sr += colv[1] * s_hval[kr][1];
        sr += colv[kr] * s_hval[kr][2];
        sr += colv[kb] * s_hval[kr][3];
        sb += colv[0] * s_hval[kb][0];
        sb += colv[1] * s_hval[kb][1];
        sb += colv[2] * s_hval[kb][2];
        sb += colv[kr] * s_hval[kb][3];
        sb += colv[kb] * s_hval[kb][4];
        pht[kr] = sr;
        pht[kb] = sb;
    }
    if ((QT - 3) % 2) pht[QT - 1] = 0.0;
    if constexpr (!PEER) {
#pragma unroll
        for (int k = 0; k < QT; ++k) {
            if (k < (int)q_ub) {
                E[(size_t)k * n + t] = raw[k];
                E[(size_t)(q_ub + k) * n + t] = pht[k];
            }
        }
    } else {
#pragma unroll
        for (int k = 0; k < QT; ++k) {
            if (k < q) {
                const uint32_t r = (uint32_t)s_rows[k];
                const bool own_row = r >= rb && r < re;
                for (int p = 0; p < ps.world; ++p) {
                    if (own_row) ps.E[p][(size_t)k * n + t] = raw[k];
                    if (mine) ps.E[p][(size_t)(q_ub + k) * n + t] = pht[k];
                }
            }
        }
    }
    }
    if constexpr (PEER) {
        // One system-scope fence per CTA, by the thread that signals: the CTA barrier orders every thread's stores
        // before it and the fence is cumulative, so the counter bump below publishes all of them (256 per-thread
        // system fences here cost ~20 us per step).
        __syncthreads();
        if (tid == 0) {
            __threadfence_system();
            for (int p = 0; p < ps.world; ++p) atomicAdd_system(ps.flag[p] + ps.rank, 1ULL);
        }
    }
}


// ------------------------------------------------------------------------------------------
// Gain. Every CTA rebuilds S = H_o P H_o' + R_o and inverts it in shared memory (Gauss-Jordan with
// partial pivoting, in place; q <= 131), then produces its slice of H_o P (layout [j][QT], rows
// >= q zero), of K = (P H_o') S^-1 (layout [QT][n]) and of X += K y.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) large_gain_kernel(uint32_t n, const LargeCtx* ctx, uint32_t q_ub,
                                                        const double* E, uint32_t QT, double* hp, double* K,
                                                        double* X) {
    extern __shared__ double sm[];
    const int q = ctx->q;
    double* S = sm;  // q x q, row-major
    __shared__ int piv[LARGE_QMAX];
    const int tid = threadIdx.x, nt = blockDim.x;
    // S(a,b) = sum_e (H_o P)[a][col_b[e]] * v_b[e] + R;  (H_o P)[a][j] = sum_e' v_a[e'] * row_{hk[a][e']}[j]
    for (int idx = tid; idx < q * q; idx += nt) {
        const int a = idx / q, b = idx % q;
        double s = 0.0;
        for (int e = 0; e < ctx->hnnz[b]; ++e) {
            const int j = ctx->hcol[b][e];
            double hpaj = 0.0;
            for (int f = 0; f < ctx->hnnz[a]; ++f) hpaj += ctx->hval[a][f] * E[(size_t)ctx->hk[a][f] * n + j];
            s += hpaj * ctx->hval[b][e];
        }
        S[idx] = s + (a == b ? ctx->rdiag[a] : 0.0);
    }
    __syncthreads();
    // in-place Gauss-Jordan inverse with partial pivoting by ONE warp (q <= 131: a handful of rows per lane);
    // __syncwarp instead of CTA barriers: the 19 x 19 case takes a few microseconds instead of ~50
    if (tid < 32) {
        const int lane = tid;
        for (int k = 0; k < q; ++k) {
            // pivot: largest |S(i,k)|, i >= k; ties to the lowest row (first maximum, as Eigen's PartialPivLU scans)
            double best = -1.0;
            int p = k;
            for (int i = k + lane; i < q; i += 32) {
                const double v = fabs(S[i * q + k]);
                if (v > best) {
                    best = v;
                    p = i;
                }
            }
            for (int off = 16; off; off >>= 1) {
                const double ob = __shfl_xor_sync(0xffffffffu, best, off);
                const int op = __shfl_xor_sync(0xffffffffu, p, off);
                if (ob > best || (ob == best && op < p)) {
                    best = ob;
                    p = op;
                }
            }
            if (lane == 0) piv[k] = p;
            if (p != k)
                for (int j = lane; j < q; j += 32) {
                    const double t = S[k * q + j];
                    S[k * q + j] = S[p * q + j];
                    S[p * q + j] = t;
                }
            __syncwarp();
            const double d = 1.0 / S[k * q + k];
            __syncwarp();
            for (int j = lane; j < q; j += 32) S[k * q + j] = (j == k ? 1.0 : S[k * q + j]) * d;
            __syncwarp();
            for (int i = lane; i < q; i += 32) {  // lane owns whole rows: no conflicts
                if (i == k) continue;
                const double f = S[i * q + k];
                for (int j = 0; j < q; ++j)
                    if (j != k) S[i * q + j] = fma(-f, S[k * q + j], S[i * q + j]);
                S[i * q + k] = -f * S[k * q + k];
            }
            __syncwarp();
        }
        for (int k = q - 1; k >= 0; --k) {  // undo the row swaps as column swaps
            const int p = piv[k];
            if (p != k)
                for (int i = lane; i < q; i += 32) {
                    const double t = S[i * q + k];
                    S[i * q + k] = S[i * q + p];
                    S[i * q + p] = t;
                }
            __syncwarp();
        }
    }
    __syncthreads();
    // ---- slices
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    for (int k = 0; k < (int)QT; ++k) {  // H_o P column t
        double s = 0.0;
        if (k < q)
            for (int e = 0; e < ctx->hnnz[k]; ++e) s += ctx->hval[k][e] * E[(size_t)ctx->hk[k][e] * n + t];
        hp[(size_t)t * QT + k] = s;
    }
    double dx = 0.0;
    for (int k = 0; k < (int)QT; ++k) {  // K row t
        double s = 0.0;
        if (k < q) {
            for (int e = 0; e < q; ++e) s += E[(size_t)(q_ub + e) * n + t] * S[e * q + k];
            dx += s * ctx->y[k];
        }
        K[(size_t)k * n + t] = s;
    }
    X[t] += dx;  // ekf.cpp:309
}

// Gain for q <= QT (QT = 20 or 40), the path cfg4 takes. Same result as large_gain_kernel, arranged for latency: the
// per-step context is copied to shared memory once, S^-1 is done by one warp, and every thread first pulls its 2 q
// values of the exchange buffer with independent (unrolled) loads, then does all arithmetic from registers and shared
// memory. H_o has the fixed structure pose rows = identity, then (range, bearing) pairs -- ekf.cpp:124-132 -- so
// H_o P for landmark pair c needs only raw rows 0, 1, 2, 3+2c, 4+2c: static register indices.
// PEER: the exchange buffer is filled by the gather kernels of all ranks; wait (bounded) until every source's
// arrival counter has reached this step's target, then read it past L1 (ld.cg: the lines came in over NVLink).
template <bool PEER>
__device__ __forceinline__ double ld_exch(const double* p) {
    if constexpr (PEER)
        return __ldcg(p);
    else
        return *p;
}

template <int QT, bool PEER>
__global__ void __launch_bounds__(256) large_gain_fast_kernel(uint32_t n, LargeCtx* __restrict__ ctx, uint32_t q_ub,
                                                             const double* __restrict__ E, double* __restrict__ hp,
                                                             double* __restrict__ K, double* __restrict__ X,
                                                             const unsigned long long* flags, unsigned long long target,
                                                             int world) {
    if constexpr (PEER) {
        if ((int)threadIdx.x < world) {
            unsigned long long v = 0;
            for (int it = 0; it < 4000000; ++it) {
                asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flags + threadIdx.x) : "memory");
                if (v >= target) break;
                __nanosleep(256);
            }
            if (v < target) atomicOr(&ctx->status, (uint32_t)ASLAM_ST_PEER_TIMEOUT);
        }
        __syncthreads();
    }
    __shared__ double S[QT * QT];
    __shared__ double s_hval[QT][5];
    __shared__ int s_hk[QT][5], s_hcol[QT][5], s_hnnz[QT], piv[QT];
    __shared__ double s_y[QT], s_rdiag[QT];
    const int q = ctx->q;
    const int tid = threadIdx.x;
    if (tid < QT) {
        const bool in = tid < q;
        s_hnnz[tid] = in ? ctx->hnnz[tid] : 0;
        s_y[tid] = in ? ctx->y[tid] : 0.0;
        s_rdiag[tid] = in ? ctx->rdiag[tid] : 1.0;
#pragma unroll
        for (int e = 0; e < 5; ++e) {
            s_hval[tid][e] = in ? ctx->hval[tid][e] : 0.0;
            s_hk[tid][e] = in ? ctx->hk[tid][e] : 0;
            s_hcol[tid][e] = in ? ctx->hcol[tid][e] : 0;
        }
    }
    __syncthreads();
    // S = H_s Psub H_s' + R with Psub(k1,k2) = P(rows[k1], rows[k2]) (every column H_o touches is one of its own rows'
    // state indices) and H_s the q x q sparse form of H_o: one round of independent loads, then shared memory only.
    // Zero-padded to QT x QT with the identity on the padding diagonal.
    __shared__ double T[QT * QT];  // Psub, then H_s Psub
    __shared__ int s_rows[QT];
    if (tid < QT) s_rows[tid] = tid < q ? ctx->rows[tid] : 0;
    __syncthreads();
    for (int idx = tid; idx < QT * QT; idx += 256) {
        const int k1 = idx / QT, k2 = idx % QT;
        S[idx] = (k1 < q && k2 < q) ? ld_exch<PEER>(E + (size_t)k1 * n + s_rows[k2]) : 0.0;  // Psub, staged in S
    }
    __syncthreads();
    for (int idx = tid; idx < QT * QT; idx += 256) {
        const int a = idx / QT, j = idx % QT;
        double hpaj = 0.0;
        for (int f = 0; f < s_hnnz[a]; ++f) hpaj += s_hval[a][f] * S[s_hk[a][f] * QT + j];
        T[idx] = hpaj;
    }
    __syncthreads();
    for (int idx = tid; idx < QT * QT; idx += 256) {
        const int a = idx / QT, b = idx % QT;
        double sum = 0.0;
        for (int e = 0; e < s_hnnz[b]; ++e) sum += T[a * QT + s_hk[b][e]] * s_hval[b][e];
        S[idx] = sum + (a == b ? s_rdiag[a] : 0.0);
    }
    __syncthreads();
    // Column t of the exchange buffer and of H_o P: independent of S^-1, so the loads are in flight (and H_o P is
    // stored) while the inversion below runs.
    const uint32_t t = blockIdx.x * 256 + tid;
    const bool valid = t < n;
    const uint32_t tc = valid ? t : n - 1;
    double pht[QT];
    {
        double raw[QT];
#pragma unroll
        for (int k = 0; k < QT; ++k) {
            raw[k] = k < q ? ld_exch<PEER>(E + (size_t)k * n + tc) : 0.0;
            pht[k] = k < q ? ld_exch<PEER>(E + (size_t)(q_ub + k) * n + tc) : 0.0;
        }
        double hpv[QT];
#pragma unroll
        for (int k = 0; k < 3; ++k) hpv[k] = s_hval[k][0] * raw[k];
#pragma unroll
        for (int c = 0; c < (QT - 3) / 2; ++c) {
            const int kr = 3 + 2 * c, kb = kr + 1;
            // same summation order as the generic kernel: entries e = 0.. of the row, left to right
            double sr = 0.0, sb = 0.0;
            sr += s_hval[kr][0] * raw[0];
            sr += s_hval[kr][1] * raw[1];
            sr += s_hval[kr][2] * raw[kr];
            sr += s_hval[kr][3] * raw[kb];
            sb += s_hval[kb][0] * raw[0];
            sb += s_hval[kb][1] * raw[1];
            sb += s_hval[kb][2] * raw[2];
            sb += s_hval[kb][3] * raw[kr];
            sb += s_hval[kb][4] * raw[kb];
            hpv[kr] = sr;
            hpv[kb] = sb;
        }
        if ((QT - 3) % 2) hpv[QT - 1] = 0.0;
        if (valid) {
            double* hpo = hp + (size_t)t * QT;
#pragma unroll
            for (int k = 0; k < QT; k += 2) *reinterpret_cast<double2*>(hpo + k) = make_double2(hpv[k], hpv[k + 1]);
        }
    }
    // Gauss-Jordan inversion with partial pivoting (what `S.inverse()` of ekf.cpp:301 does through PartialPivLU, as
    // one in-place sweep), the whole CTA on it: every thread finds the pivot of column k itself (broadcast reads, no
    // shuffle chain), the eliminated matrix goes to the other buffer (no read/write hazard), ONE barrier per pivot.
    // The arithmetic per element is the one-warp version's, so the result is bit-identical to it.
    double* cur = S;
    double* nxt = T;
    for (int k = 0; k < q; ++k) {
        double best = -1.0;
        int p = k;
        for (int i = k; i < q; ++i) {
            const double v = fabs(cur[i * QT + k]);
            if (v > best) {  // first maximum
                best = v;
                p = i;
            }
        }
        if (tid == 0) piv[k] = p;
        const double d = 1.0 / cur[p * QT + k];
        for (int idx = tid; idx < QT * QT; idx += 256) {
            const int i = idx / QT, j = idx % QT;
            double out = cur[idx];
            if (i < q && j < q) {
                const double pkj = (j == k ? 1.0 : cur[p * QT + j]) * d;  // row k after the swap and the scaling
                if (i == k) {
                    out = pkj;
                } else {
                    const int r = i == p ? k : i;  // row p holds the old row k after the swap
                    const double f = cur[r * QT + k];
                    out = j == k ? -f * pkj : fma(-f, pkj, cur[r * QT + j]);
                }
            }
            nxt[idx] = out;
        }
        __syncthreads();
        double* sw = cur;
        cur = nxt;
        nxt = sw;
    }
    // undo the row swaps on the columns (k = q-1 .. 0), composed into one column permutation
    if (tid == 0) {
        for (int c = 0; c < q; ++c) s_rows[c] = c;  // s_rows is free: Psub has been read
        for (int k = q - 1; k >= 0; --k) {
            const int p = piv[k], a = s_rows[k];
            s_rows[k] = s_rows[p];
            s_rows[p] = a;
        }
    }
    __syncthreads();
    for (int idx = tid; idx < QT * QT; idx += 256) {
        const int i = idx / QT, c = idx % QT;
        nxt[idx] = (i < q && c < q) ? cur[i * QT + s_rows[c]] : 0.0;  // padding must not contribute
    }
    __syncthreads();
    const double* Sinv = nxt;
    if (!valid) return;
    double dx = 0.0;
#pragma unroll
    for (int k = 0; k < QT; ++k) {  // K row t
        double sum = 0.0;
#pragma unroll
        for (int e = 0; e < QT; ++e) sum += pht[e] * Sinv[e * QT + k];
        dx += sum * s_y[k];
        K[(size_t)k * n + t] = sum;
    }
    X[t] += dx;  // ekf.cpp:309
}

// ------------------------------------------------------------------------------------------
// K5: P_local -= K_local (H_o P), the HBM-bound streaming kernel: P is read once and written once, 16 n^2 bytes per
// step against 2 q n^2 flop (HBM-bound for q <~ 40).
// ------------------------------------------------------------------------------------------
// K5, bulk-copy form (launched for 20 < q <= 40; q <= 20 takes the warp-specialised kernel below). The covariance tile never passes through
// registers on its way in or out: one elected thread drives the copy engine (cp.async.bulk, SASS
// UBLKCP) to pull 256*R rows x CS columns of P -- CS contiguous column segments -- plus the matching
// CS x Q block of H_o P into one of NST shared-memory stages, the CTA updates the stage in place
// (thread t: rows R t .. R t + R - 1 of every column, -K slice in registers, H_o P read as a
// broadcast), and the same thread streams the stage back with bulk stores. Loads run NST-2 stages
// ahead, so the bytes in flight per SM are set by shared memory (up to ~190 KB), not by registers.
template <int R, int Q, int CS, int NST>
struct BulkGeo {
    static constexpr int ROWS = 256 * R;
    static constexpr size_t P_BYTES = (size_t)CS * ROWS * sizeof(double);
    static constexpr size_t H_BYTES = (size_t)CS * Q * sizeof(double);
    static constexpr size_t STAGE = P_BYTES + H_BYTES;
    static constexpr size_t SMEM = NST * STAGE + NST * sizeof(uint64_t) + 16;
};

template <int R, int Q, int CS, int NST>
__global__ void __launch_bounds__(256, 1) large_update_bulk_kernel(uint32_t n, uint32_t rb, uint32_t nrows, size_t ldp,
                                                                  double* __restrict__ P,
                                                                  const double* __restrict__ K,
                                                                  const double* __restrict__ hp,
                                                                  uint32_t cols_per_cta) {
    using G = BulkGeo<R, Q, CS, NST>;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + NST * G::STAGE);
    const int tid = threadIdx.x;
    const uint32_t r0 = blockIdx.y * G::ROWS;                          // first local row of the tile
    const uint32_t rows_here = min((uint32_t)G::ROWS, (uint32_t)ldp - r0);  // ldp is a multiple of 16 rows
    const uint32_t seg_bytes = rows_here * sizeof(double);
    const uint32_t c_begin = blockIdx.x * cols_per_cta;
    const uint32_t c_end = min(n, c_begin + cols_per_cta);
    if (c_begin >= c_end) return;
    const uint32_t nchunks = (c_end - c_begin + CS - 1) / CS;

    auto stage_p = [&](int st) { return reinterpret_cast<double*>(smem_raw + st * G::STAGE); };
    auto stage_h = [&](int st) { return reinterpret_cast<double*>(smem_raw + st * G::STAGE + G::P_BYTES); };
    auto issue_load = [&](uint32_t chunk) {  // elected thread only
        const int st = chunk % NST;
        const uint32_t c0 = c_begin + chunk * CS;
        const uint32_t cw = min((uint32_t)CS, c_end - c0);
        bulk::mbar_expect_tx(&full[st], cw * seg_bytes + cw * Q * (uint32_t)sizeof(double));
        double* dp = stage_p(st);
        for (uint32_t c = 0; c < cw; ++c)
            bulk::g2s(dp + c * G::ROWS, P + (size_t)(c0 + c) * ldp + r0, seg_bytes, &full[st]);
        bulk::g2s(stage_h(st), hp + (size_t)c0 * Q, cw * Q * (uint32_t)sizeof(double), &full[st]);
    };

    if (tid == 0) {
        for (int s = 0; s < NST; ++s) bulk::mbar_init(&full[s], 1);
        bulk::fence_barrier_init();
    }
    __syncthreads();
    if (tid == 0)
        for (uint32_t c = 0; c < (uint32_t)(NST - 1) && c < nchunks; ++c) issue_load(c);

    // this thread's R x Q slice of -K
    const uint32_t i0 = r0 + tid * R;
    double nk[Q][R];
#pragma unroll
    for (int k = 0; k < Q; ++k)
#pragma unroll
        for (int r = 0; r < R; ++r) nk[k][r] = (i0 + r < nrows) ? -K[(size_t)k * n + rb + i0 + r] : 0.0;
    const bool active = (uint32_t)(tid * R) < rows_here;

    for (uint32_t chunk = 0; chunk < nchunks; ++chunk) {
        const int st = chunk % NST;
        const uint32_t c0 = c_begin + chunk * CS;
        const uint32_t cw = min((uint32_t)CS, c_end - c0);
        bulk::mbar_wait(&full[st], (chunk / NST) & 1u);
        if (active) {
            double* sp = stage_p(st) + tid * R;
            const double* sh = stage_h(st);
            if (cw == CS) {
#pragma unroll
                for (int c = 0; c < CS; ++c) {
                    double v[R];
                    if constexpr (R == 2) {
                        const double2 t = *reinterpret_cast<const double2*>(sp + c * G::ROWS);
                        v[0] = t.x;
                        v[1] = t.y;
                    } else {
                        v[0] = sp[c * G::ROWS];
                    }
#pragma unroll
                    for (int k = 0; k < Q; k += 2) {
                        const double2 h = *reinterpret_cast<const double2*>(sh + c * Q + k);
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            v[r] = fma(nk[k][r], h.x, v[r]);
                            v[r] = fma(nk[k + 1][r], h.y, v[r]);
                        }
                    }
                    if constexpr (R == 2) {
                        *reinterpret_cast<double2*>(sp + c * G::ROWS) = make_double2(v[0], v[1]);
                    } else {
                        sp[c * G::ROWS] = v[0];
                    }
                }
            } else {
                for (uint32_t c = 0; c < cw; ++c)
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        double v = sp[c * G::ROWS + r];
#pragma unroll
                        for (int k = 0; k < Q; ++k) v = fma(nk[k][r], sh[c * Q + k], v);
                        sp[c * G::ROWS + r] = v;
                    }
            }
            bulk::fence_proxy_async();
        }
        __syncthreads();
        if (tid == 0) {
            const double* sp = stage_p(st);
            for (uint32_t c = 0; c < cw; ++c) bulk::s2g(P + (size_t)(c0 + c) * ldp + r0, sp + c * G::ROWS, seg_bytes);
            bulk::commit_group();
            // refill the stage drained one iteration ago: its store group is the second newest
            const uint32_t nxt = chunk + NST - 1;
            if (nxt < nchunks) {
                bulk::wait_group_read<1>();
                issue_load(nxt);
            }
        }
    }
    if (tid == 0) bulk::wait_group<0>();
}

// K5, persistent warp-specialised form (the kernel cfg4 runs). One CTA per SM for the whole launch; the row blocks x
// column chunks of the local P are laid out as ONE sequence (row-block major) and every CTA takes a contiguous,
// equally heavy piece of it, so its shared-memory ring never drains between tiles (the round-1 kernel gave each CTA
// ~10 chunks and paid the ring fill / drain once per CTA: 84 % of the HBM peak).
//   producer lane : wait done[st] -> bulk-store the stage -> (one chunk later, once that store has been read out of
//                   shared memory) re-arm full[st] and bulk-load the chunk NST-1 ahead (cp.async.bulk, SASS UBLKCP)
//   consumer warp : wait full[st] -> update its 64 rows of the stage in place, 4 columns interleaved (8 independent
//                   DFMA chains per thread) -> fence.proxy.async -> arrive on done[st]
// Predict: P = A P A' + Q (ekf.cpp:297) has NOT been applied to P in memory on this path (see pred_entry below): the
// consumer that owns rows 0, 1 / the chunk that holds columns 0, 1 applies it to the stage in shared memory before the
// rank-q update -- the formulas of large_predict_kernel on the same operands.
// The local rows are cut into equally high blocks of at most ROWS rows (StreamPlan), so every chunk moves about the
// same number of bytes and the chunk sequence can simply be dealt out in equal counts.
// Launched with programmatic stream serialization behind the solve kernel: barrier set-up and the first NST-1 tiles of
// P (which the solve kernel does not write) are under way before griddepcontrol.wait; K, H_o P and predA are touched
// only after it AND after the solve stage's completion counter has been seen (the counter makes correctness
// independent of how the launch was serialized).
__device__ __forceinline__ void mbar_arrive_cta(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bulk::smem_u32(bar)) : "memory");
}

struct StreamPlan {
    uint32_t n, rb, nrows;   // state dimension, first global row and number of rows of this shard
    uint32_t nch;            // column chunks per row block: ceil(n / CS)
    uint32_t nblk;           // row blocks
    uint32_t blk_rows;       // rows per block (multiple of 16, <= ROWS): the local rows are cut into EQUAL blocks, so
    uint32_t last_rows;      // that no short block is left over (its chunks would cost a full chunk's latency each)
};

// Tile geometry of the persistent kernel: CW consumer warps x 32 lanes x R rows per row block, CS columns per chunk.
// CW = 7 (+ the producer warp = 8 warps, two per SM sub-partition) on purpose: the register file is split four ways, a
// ninth warp would put three 168-register warps (16,128 of 16,384 registers) on one sub-partition and leave no room
// there for a warp of any other kernel -- the front of the NEXT step, which is meant to run under this kernel, could
// then only start when a CTA of this (persistent) grid retires, i.e. at the very end (measured: %globaltimer trace).
template <int CW, int R, int Q, int CS, int NST>
struct StreamGeo {
    static constexpr int ROWS = CW * 32 * R;
    static constexpr size_t P_BYTES = (size_t)CS * ROWS * sizeof(double);
    static constexpr size_t H_BYTES = (size_t)CS * Q * sizeof(double);
    static constexpr size_t STAGE = P_BYTES + H_BYTES;
    static constexpr size_t SMEM = NST * STAGE + 2 * NST * sizeof(uint64_t) + 16;
    static constexpr int THREADS = (CW + 1) * 32;
};

template <int CW, int R, int Q, int CS, int NST>
__global__ void __launch_bounds__((CW + 1) * 32, 1) large_update_stream_kernel(
    StreamPlan pl, size_t ldp, const double* Pin, double* Pout, const double* __restrict__ K, const double* __restrict__ hp,
    const double* __restrict__ predA, double kq, const unsigned long long* sync, unsigned long long epoch,
    unsigned long long* trace) {
    using G = StreamGeo<CW, R, Q, CS, NST>;
    static_assert(R == 2 && CS % 4 == 0, "the interleaved inner loop is written for 2 rows x 4 columns");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + NST * G::STAGE);
    uint64_t* done = full + NST;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t n = pl.n, rb = pl.rb, nrows = pl.nrows;
    if (tid == 0) trace_min(trace, epoch, 6);
    const unsigned long long T = (unsigned long long)pl.nblk * pl.nch;  // chunks, row-block major
    const uint32_t g_begin = (uint32_t)(T * blockIdx.x / gridDim.x);
    const uint32_t g_end = (uint32_t)(T * (blockIdx.x + 1) / gridDim.x);
    auto stage_p = [&](int st) { return reinterpret_cast<double*>(smem_raw + st * G::STAGE); };
    auto stage_h = [&](int st) { return reinterpret_cast<double*>(smem_raw + st * G::STAGE + G::P_BYTES); };
    auto rows_of = [&](uint32_t b) { return b + 1 < pl.nblk ? pl.blk_rows : pl.last_rows; };

    if (tid == 0) {
        for (int s = 0; s < NST; ++s) {
            bulk::mbar_init(&full[s], 1);
            bulk::mbar_init(&done[s], CW);
        }
        bulk::fence_barrier_init();
    }
    __syncthreads();

    if (warp == CW) {  // ---------------- producer
        if (lane != 0) return;
        auto issue_p = [&](uint32_t g) {
            const int st = (g - g_begin) % NST;
            const uint32_t b = g / pl.nch, c0 = (g % pl.nch) * CS;
            const uint32_t cw = min((uint32_t)CS, n - c0);
            const uint32_t seg_bytes = rows_of(b) * (uint32_t)sizeof(double);
            bulk::mbar_expect_tx(&full[st], cw * seg_bytes + cw * Q * (uint32_t)sizeof(double));
            double* dp = stage_p(st);
            for (uint32_t c = 0; c < cw; ++c)
                bulk::g2s(dp + c * G::ROWS, Pin + (size_t)(c0 + c) * ldp + (size_t)b * pl.blk_rows, seg_bytes, &full[st]);
        };
        auto issue_h = [&](uint32_t g) {
            const int st = (g - g_begin) % NST;
            const uint32_t c0 = (g % pl.nch) * CS;
            const uint32_t cw = min((uint32_t)CS, n - c0);
            bulk::g2s(stage_h(st), hp + (size_t)c0 * Q, cw * Q * (uint32_t)sizeof(double), &full[st]);
        };
        // P tiles first (independent of the solve kernel), H_o P blocks once it has completed
        for (uint32_t g = g_begin; g < g_begin + (uint32_t)(NST - 1) && g < g_end; ++g) issue_p(g);
        asm volatile("griddepcontrol.wait;" ::: "memory");
        spin_until(sync + SYNC_SOLVE, epoch);
        for (uint32_t g = g_begin; g < g_begin + (uint32_t)(NST - 1) && g < g_end; ++g) issue_h(g);
        for (uint32_t g = g_begin; g < g_end; ++g) {
            const uint32_t li = g - g_begin;
            const int st = li % NST;
            const uint32_t b = g / pl.nch, c0 = (g % pl.nch) * CS;
            const uint32_t cw = min((uint32_t)CS, n - c0);
            const uint32_t seg_bytes = rows_of(b) * (uint32_t)sizeof(double);
            bulk::mbar_wait(&done[st], (li / NST) & 1u);
            const double* sp = stage_p(st);
            for (uint32_t c = 0; c < cw; ++c)
                bulk::s2g(Pout + (size_t)(c0 + c) * ldp + (size_t)b * pl.blk_rows, sp + c * G::ROWS, seg_bytes);
            bulk::commit_group();
            const uint32_t nxt = g + NST - 1;  // goes into the stage drained one chunk ago
            if (nxt < g_end) {
                bulk::wait_group_read<1>();
                issue_p(nxt);
                issue_h(nxt);
            }
        }
        bulk::wait_group<0>();
        trace_max(trace, epoch, 5);
        return;
    }

    // ---------------- consumers
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (lane == 0) spin_until(sync + SYNC_SOLVE, epoch);
    __syncwarp();
    if (tid == 0) trace_min(trace, epoch, 4);
    double a00 = 1.0, a10 = 0.0;
    if (predA) {
        a00 = predA[0];
        a10 = predA[1];
    }
    double nk[Q][R];
    uint32_t cur_b = 0xffffffffu;
    for (uint32_t g = g_begin; g < g_end; ++g) {
        const uint32_t li = g - g_begin;
        const int st = li % NST;
        const uint32_t b = g / pl.nch, c0 = (g % pl.nch) * CS;
        const uint32_t cw = min((uint32_t)CS, n - c0);
        const uint32_t i0 = b * pl.blk_rows + tid * R;  // local index of this thread's first row
        if (b != cur_b) {                           // this thread's R x Q slice of -K
            cur_b = b;
#pragma unroll
            for (int k = 0; k < Q; ++k)
#pragma unroll
                for (int r = 0; r < R; ++r) nk[k][r] = (i0 + r < nrows) ? -K[(size_t)k * n + rb + i0 + r] : 0.0;
        }
        const bool active = (uint32_t)(tid * R) < rows_of(b);
        const uint32_t g0 = rb + i0;  // global index of rows g0, g0 + 1
        bulk::mbar_wait(&full[st], (li / NST) & 1u);
        if (active) {
            double* sp = stage_p(st) + tid * R;
            const double* sh = stage_h(st);
            if (predA) {  // P = A P A' + Q on the part of this thread's two rows that lies in the stage
                if (c0 == 0) {  // columns 0, 1 are the first two columns of this stage (n >= 5 > 2)
                    if (g0 >= 2) {
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            const double p0 = sp[r], p1 = sp[G::ROWS + r];
                            sp[r] = p0 * a00;
                            sp[G::ROWS + r] = p0 * a10 + p1;
                        }
                        if (g0 == 2) sp[2 * G::ROWS] += kq;  // P(2,2) += Q(2,2)
                    } else {  // g0 == 0: the 2 x 2 corner, row pass then column pass
                        const double p00 = sp[0], p10 = sp[1], p01 = sp[G::ROWS], p11 = sp[G::ROWS + 1];
                        const double r00 = a00 * p00, r01 = a00 * p01;
                        const double r10 = a10 * p00 + p10, r11 = a10 * p01 + p11;
                        sp[0] = r00 * a00 + kq;
                        sp[G::ROWS] = r00 * a10 + r01;
                        sp[1] = r10 * a00;
                        sp[G::ROWS + 1] = r10 * a10 + r11 + kq;
                    }
                }
                if (g0 == 0) {  // rows 0, 1 of every other column
                    for (uint32_t c = (c0 == 0 ? 2u : 0u); c < cw; ++c) {
                        const double p0 = sp[c * G::ROWS], p1 = sp[c * G::ROWS + 1];
                        sp[c * G::ROWS] = a00 * p0;
                        sp[c * G::ROWS + 1] = a10 * p0 + p1;
                    }
                }
            }
            if (cw == CS) {
#pragma unroll
                for (int cg = 0; cg < CS; cg += 4) {
                    double2 v[4];
#pragma unroll
                    for (int c = 0; c < 4; ++c) v[c] = *reinterpret_cast<const double2*>(sp + (cg + c) * G::ROWS);
#pragma unroll
                    for (int k = 0; k < Q; k += 2) {
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            const double2 h = *reinterpret_cast<const double2*>(sh + (cg + c) * Q + k);
                            v[c].x = fma(nk[k][0], h.x, v[c].x);
                            v[c].y = fma(nk[k][1], h.x, v[c].y);
                            v[c].x = fma(nk[k + 1][0], h.y, v[c].x);
                            v[c].y = fma(nk[k + 1][1], h.y, v[c].y);
                        }
                    }
#pragma unroll
                    for (int c = 0; c < 4; ++c) *reinterpret_cast<double2*>(sp + (cg + c) * G::ROWS) = v[c];
                }
            } else {
                for (uint32_t c = 0; c < cw; ++c)
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        double v = sp[c * G::ROWS + r];
#pragma unroll
                        for (int k = 0; k < Q; ++k) v = fma(nk[k][r], sh[c * Q + k], v);
                        sp[c * G::ROWS + r] = v;
                    }
            }
            bulk::fence_proxy_async();
        }
        __syncwarp();
        if (lane == 0) mbar_arrive_cta(&done[st]);
    }
}

// Any q (up to LARGE_QMAX): K read from global (L2-resident), FP64-bound for large q anyway.
__global__ void __launch_bounds__(256) large_update_generic_kernel(uint32_t n, uint32_t rb, uint32_t nrows,
                                                                  size_t ldp, double* __restrict__ P,
                                                                  const double* __restrict__ K,
                                                                  const double* __restrict__ hp, uint32_t QT,
                                                                  const LargeCtx* ctx) {
    const int q = ctx->q;
    const uint32_t i = blockIdx.x * 256 + threadIdx.x;
    const uint32_t j0 = blockIdx.y * 32;
    if (i >= nrows) return;
    for (uint32_t j = j0; j < min(n, j0 + 32); ++j) {
        double v = P[(size_t)j * ldp + i];
        for (int k = 0; k < q; ++k) v = fma(-K[(size_t)k * n + rb + i], hp[(size_t)j * QT + k], v);
        P[(size_t)j * ldp + i] = v;
    }
}

// ==========================================================================================
// Fast path, q_ub <= 20 (cfg4: 8 observed landmarks, q = 19): TWO kernels after the front --
//   large_solve_kernel   gather + exchange + S + S^-1 + H_o P + K + X in one launch
//   large_update_ws_kernel (above)   with the predict P = A P A' + Q applied to the tiles as they pass
// P = A P A' + Q (ekf.cpp:297) touches rows / columns 0, 1 and P(2,2) only. Instead of a kernel that rewrites them
// in place before the gather, every reader of P in the solve kernel goes through pred_entry(), which returns the
// PREDICTED entry from the stored one(s), and the streaming update applies the same formulas to the tile rows /
// columns 0, 1 in shared memory before its rank-q update. Same operations on the same operands as
// large_predict_kernel, so the results are bit-identical to the unfused path.
// ==========================================================================================
__device__ __forceinline__ double pred_entry(const double* __restrict__ P, size_t ldp, uint32_t rb, uint32_t i,
                                             uint32_t j, double a00, double a10, double kq) {
    // (i, j) global indices; row i must be owned by this shard (rows 0, 1, 2 live together on the first shard)
    if (i < 2 && j < 2) {  // the 2 x 2 corner: row pass, then column pass (large_predict_kernel, t == 0)
        const double p00 = P[0], p10 = P[1], p01 = P[ldp], p11 = P[ldp + 1];
        const double r00 = a00 * p00, r01 = a00 * p01;
        const double r10 = a10 * p00 + p10, r11 = a10 * p01 + p11;
        if (i == 0) return j == 0 ? r00 * a00 + kq : r00 * a10 + r01;
        return j == 0 ? r10 * a00 : r10 * a10 + r11 + kq;
    }
    const double* c = P + (size_t)j * ldp;
    if (i < 2) {  // j >= 2: rows 0, 1 of column j
        const double p0 = c[0];
        return i == 0 ? a00 * p0 : a10 * p0 + c[1];
    }
    if (j < 2) {  // i >= 2: row i of columns 0, 1
        const double p0 = P[i - rb];
        return j == 0 ? p0 * a00 : p0 * a10 + P[ldp + (i - rb)];
    }
    const double v = c[i - rb];
    return (i == 2 && j == 2) ? v + kq : v;
}

// P H_o' of row t from the predicted column values colv[k] = P(t, rows[k]); entries of each H row left to right, as
// the oracle sums them. H_o has the fixed structure pose rows = identity, then (range, bearing) pairs (ekf.cpp:124-132).
template <int QT>
__device__ __forceinline__ void pht_from_cols(const double (&colv)[QT], const double (*s_hval)[5], double (&pht)[QT]) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        double sum = 0.0;
        sum += colv[k] * s_hval[k][0];
        pht[k] = sum;
    }
#pragma unroll
    for (int c = 0; c < (QT - 3) / 2; ++c) {
        const int kr = 3 + 2 * c, kb = kr + 1;
        double sr = 0.0, sb = 0.0;
        sr += colv[0] * s_hval[kr][0];
        sr += colv[1] * s_hval[kr][1];
        sr += colv[kr] * s_hval[kr][2];
        sr += colv[kb] * s_hval[kr][3];
        sb += colv[0] * s_hval[kb][0];
        sb += colv[1] * s_hval[kb][1];
        sb += colv[2] * s_hval[kb][2];
        sb += colv[kr] * s_hval[kb][3];
        sb += colv[kb] * s_hval[kb][4];
        pht[kr] = sr;
        pht[kb] = sb;
    }
    if ((QT - 3) % 2) pht[QT - 1] = 0.0;
}

// Predicted P(t, rows[k]) for the shard's own row t (column rows[k] of P: consecutive t are consecutive in memory).
// Rows 0, 1 of the whole matrix are the only ones where the predict needs more than the stored entries of row t;
// the two threads with t < 2 take the generic pred_entry(), everybody else the straight loads below.
template <int QT>
__device__ __forceinline__ void gather_cols(const double* __restrict__ P, size_t ldp, uint32_t rb, uint32_t t, bool mine,
                                            int q, const int* s_rows, double a00, double a10, double kq,
                                            double (&colv)[QT]) {
    if (!mine) {
#pragma unroll
        for (int k = 0; k < QT; ++k) colv[k] = 0.0;
    } else if (t >= 2) {
        const double c0 = P[t - rb];
        colv[0] = 0 < q ? c0 * a00 : 0.0;
        colv[1] = 1 < q ? c0 * a10 + P[ldp + (t - rb)] : 0.0;
        const double c2 = P[2 * ldp + (t - rb)];
        colv[2] = 2 < q ? (t == 2 ? c2 + kq : c2) : 0.0;
#pragma unroll
        for (int k = 3; k < QT; ++k) colv[k] = k < q ? P[(size_t)s_rows[k] * ldp + (t - rb)] : 0.0;
    } else {
#pragma unroll 1
        for (int k = 0; k < QT; ++k) {
            const double cv = k < q ? pred_entry(P, ldp, rb, t, (uint32_t)s_rows[k], a00, a10, kq) : 0.0;
#pragma unroll
            for (int kk = 0; kk < QT; ++kk)
                if (kk == k) colv[kk] = cv;
        }
    }
}

// ------------------------------------------------------------------------------------------
// Overlapped steps (large_ekf_phase2_fast): the solve kernel of step k+1 runs WHILE the streaming update of step k
// is still writing P_k -- out of place, into the handle's second buffer. What the solve kernel needs of P_k (q rows /
// columns and the q x q block) it computes itself from the stored P_{k-1} and the previous step's K, H_o P and A:
// exactly what large_update_stream_kernel writes for that entry -- its predict (pred_entry with the previous A) and
// then its rank-Q update, k ascending, one fma per term -- so the value is bit-identical to the one in flight.
// ------------------------------------------------------------------------------------------
struct PrevUpd {
    const double* K;      // [QT][n] gain of the previous step; nullptr: nothing in flight, the stored entries are final
    const double* hp;     // [n][QT] its H_o P
    const double* predA;  // [2] its A(0,0), A(1,0)
    double kq;
};

// The predict of pred_entry() on top of any accessor at(i, j) of the un-predicted matrix (same expressions).
template <class Acc>
__device__ __forceinline__ double pred_entry_acc(Acc at, uint32_t i, uint32_t j, double a00, double a10, double kq) {
    if (i < 2 && j < 2) {
        const double p00 = at(0u, 0u), p10 = at(1u, 0u), p01 = at(0u, 1u), p11 = at(1u, 1u);
        const double r00 = a00 * p00, r01 = a00 * p01;
        const double r10 = a10 * p00 + p10, r11 = a10 * p01 + p11;
        if (i == 0) return j == 0 ? r00 * a00 + kq : r00 * a10 + r01;
        return j == 0 ? r10 * a00 : r10 * a10 + r11 + kq;
    }
    if (i < 2) {
        const double p0 = at(0u, j);
        return i == 0 ? a00 * p0 : a10 * p0 + at(1u, j);
    }
    if (j < 2) {
        const double p0 = at(i, 0u);
        return j == 0 ? p0 * a00 : p0 * a10 + at(i, 1u);
    }
    const double v = at(i, j);
    return (i == 2 && j == 2) ? v + kq : v;
}

// Entry (i, j) of the matrix the previous step's update is producing (row i owned by this shard), from the stored one.
template <int QT>
__device__ __forceinline__ double pv_entry(const double* __restrict__ P, size_t ldp, uint32_t rb, uint32_t n, uint32_t i,
                                           uint32_t j, const PrevUpd& pu) {
    if (!pu.K) return P[(size_t)j * ldp + (i - rb)];
    double v = pred_entry(P, ldp, rb, i, j, pu.predA[0], pu.predA[1], pu.kq);
#pragma unroll
    for (int k = 0; k < QT; ++k) v = fma(-pu.K[(size_t)k * n + i], pu.hp[(size_t)j * QT + k], v);
    return v;
}

// gather_cols() on top of the update in flight: colv[k] = predicted P_k(t, rows[k]) for the shard's own row t.
// s_hpp[k][:] = row rows[k] of the previous H_o P (staged by the caller; rows[0..2] = 0, 1, 2).
template <int QT>
__device__ __forceinline__ void gather_cols_v(const double* __restrict__ P, size_t ldp, uint32_t rb, uint32_t n,
                                              uint32_t t, bool mine, int q, const int* s_rows, double a00, double a10,
                                              double kq, const PrevUpd& pu, const double (*s_hpp)[QT],
                                              double (&colv)[QT]) {
    if (!pu.K) {
        gather_cols<QT>(P, ldp, rb, t, mine, q, s_rows, a00, a10, kq, colv);
        return;
    }
    if (!mine) {
#pragma unroll
        for (int k = 0; k < QT; ++k) colv[k] = 0.0;
    } else if (t >= 2) {
        double nk[QT];
#pragma unroll
        for (int m = 0; m < QT; ++m) nk[m] = -pu.K[(size_t)m * n + t];
        const double pa00 = pu.predA[0], pa10 = pu.predA[1];
        // the previous predict on the stored row t (columns 0, 1 and the (2,2) entry), then the previous rank-Q update
        const double p0 = P[t - rb];
        double e[QT];
        e[0] = p0 * pa00;
        e[1] = p0 * pa10 + P[ldp + (t - rb)];
        const double p2 = P[2 * ldp + (t - rb)];
        e[2] = t == 2 ? p2 + pu.kq : p2;
#pragma unroll
        for (int k = 3; k < QT; ++k) e[k] = k < q ? P[(size_t)s_rows[k] * ldp + (t - rb)] : 0.0;
#pragma unroll
        for (int k = 0; k < QT; ++k) {
            double v = e[k];
#pragma unroll
            for (int m = 0; m < QT; ++m) v = fma(nk[m], s_hpp[k][m], v);
            e[k] = v;
        }
        // this step's predict (gather_cols)
        colv[0] = 0 < q ? e[0] * a00 : 0.0;
        colv[1] = 1 < q ? e[0] * a10 + e[1] : 0.0;
        colv[2] = 2 < q ? (t == 2 ? e[2] + kq : e[2]) : 0.0;
#pragma unroll
        for (int k = 3; k < QT; ++k) colv[k] = k < q ? e[k] : 0.0;
    } else {
        auto at = [&](uint32_t i, uint32_t j) { return pv_entry<QT>(P, ldp, rb, n, i, j, pu); };
#pragma unroll 1
        for (int k = 0; k < QT; ++k) {
            const double cv = k < q ? pred_entry_acc(at, t, (uint32_t)s_rows[k], a00, a10, kq) : 0.0;
#pragma unroll
            for (int kk = 0; kk < QT; ++kk)
                if (kk == k) colv[kk] = cv;
        }
    }
}

// What the shards exchange on the fast path: E[k][t] = (P H_o')(t, k) for every row t (its owner's value), k < q_ub,
// followed by the q x q block Psub(k1, k2) = P(rows[k1], rows[k2]) (owner of row rows[k1]) at E + q_ub n.
__device__ __forceinline__ size_t exch_psub_offset(uint32_t q_ub, uint32_t n) { return (size_t)q_ub * n; }

// NCCL form, first half: this shard's part of the exchange buffer, zeros where it owns nothing (the caller sums the
// buffers of all ranks; every element has exactly one owner, so the sum is exact).
template <int QT>
__global__ void __launch_bounds__(256) large_raw_gather_kernel(uint32_t n, uint32_t rb, uint32_t re, size_t ldp,
                                                              const double* __restrict__ P,
                                                              const double* __restrict__ A, double kq,
                                                              const LargeCtx* __restrict__ ctx, uint32_t q_ub,
                                                              double* __restrict__ E) {
    __shared__ int s_rows[QT];
    __shared__ double s_hval[QT][5];
    const int q = ctx->q, tid = threadIdx.x;
    if (tid < QT) {
        s_rows[tid] = tid < q ? ctx->rows[tid] : 0;
#pragma unroll
        for (int e = 0; e < 5; ++e) s_hval[tid][e] = tid < q ? ctx->hval[tid][e] : 0.0;
    }
    __syncthreads();
    const double a00 = A[0], a10 = A[1];
    if (blockIdx.x == 0) {
        for (int idx = tid; idx < QT * QT; idx += 256) {
            const int k1 = idx / QT, k2 = idx % QT;
            const uint32_t r1 = (uint32_t)s_rows[k1];
            E[exch_psub_offset(q_ub, n) + idx] =
                (k1 < q && k2 < q && r1 >= rb && r1 < re) ? pred_entry(P, ldp, rb, r1, (uint32_t)s_rows[k2], a00, a10, kq) : 0.0;
        }
    }
    const uint32_t t = blockIdx.x * 256 + tid;
    if (t >= n) return;
    const bool mine = t >= rb && t < re;
    double colv[QT], pht[QT];
    gather_cols<QT>(P, ldp, rb, t, mine, q, s_rows, a00, a10, kq, colv);
    pht_from_cols<QT>(colv, s_hval, pht);
#pragma unroll
    for (int k = 0; k < QT; ++k)
        if (k < (int)q_ub) E[(size_t)k * n + t] = pht[k];
}

// Gauss-Jordan inverse with partial pivoting by ONE warp, lane i owns row i of the q x q matrix (q <= QT <= 32) held
// in shared memory with a padded row stride (conflict-free column reads). Per pivot: first-maximum search of column k
// (shuffle butterfly, ties to the lower row), then every lane eliminates its own row into the other buffer (no
// read/write hazard), one __syncwarp per pivot -- no CTA barrier and no 170-cycle barrier round trip per pivot as in
// the CTA-wide version of large_gain_fast_kernel, whose arithmetic this repeats element for element
// (`(j == k ? 1 : p_kj) * d`, `-f * p_kk`, `fma(-f, p_kj, s_ij)`), so the result is bit-identical to it. The row
// swaps are undone as one column permutation when the rows are written to Sinv (QT x QT, zero padded by the caller).
template <int QT>
__device__ __forceinline__ void warp_gauss_jordan(double* bufA, double* bufB, int ld, int q, int lane, int* s_piv,
                                                  int* s_perm, double* Sinv) {
    double* cur = bufA;
    double* nxt = bufB;
    for (int k = 0; k < q; ++k) {
        double best = -1.0;
        if (lane >= k && lane < q) {
            const double v = fabs(cur[lane * ld + k]);
            best = v >= 0.0 ? v : -0.5;  // a NaN never wins against a number, as in `v > best`
        }
        int p = lane;
#pragma unroll
        for (int off = 16; off; off >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, off);
            const int op = __shfl_xor_sync(0xffffffffu, p, off);
            if (ob > best || (ob == best && op < p)) {
                best = ob;
                p = op;
            }
        }
        if (lane == 0) s_piv[k] = p;
        // both rows into registers first (static indices: the loop is fully unrolled), so that the 2 q loads are in
        // flight together instead of one dependent load -> multiply -> store chain per element
        const int r = lane == p ? k : lane;  // row p holds the old row k after the swap
        const int rr = lane < q ? r : 0;
        double prow[QT], srow[QT];
#pragma unroll
        for (int jj = 0; jj < QT; ++jj) {
            prow[jj] = cur[p * ld + jj];
            srow[jj] = cur[rr * ld + jj];
        }
        const double d = 1.0 / cur[p * ld + k];
        const double f = cur[rr * ld + k];
        const bool is_k = lane == k;
        if (lane < q) {
#pragma unroll
            for (int jj = 0; jj < QT; ++jj) {
                if (jj < q) {
                    const double pkj = (jj == k ? 1.0 : prow[jj]) * d;  // row k after the swap and the scaling
                    double out;
                    if (is_k)
                        out = pkj;
                    else
                        out = jj == k ? -f * pkj : fma(-f, pkj, srow[jj]);
                    nxt[lane * ld + jj] = out;
                }
            }
        }
        __syncwarp();
        double* sw = cur;
        cur = nxt;
        nxt = sw;
    }
    // undo the row swaps on the columns (k = q-1 .. 0), composed into one permutation: Sinv(i, c) = cur(i, perm[c])
    if (lane == 0) {
        for (int c = 0; c < q; ++c) s_perm[c] = c;
        for (int k = q - 1; k >= 0; --k) {
            const int pv = s_piv[k], a = s_perm[k];
            s_perm[k] = s_perm[pv];
            s_perm[pv] = a;
        }
    }
    __syncwarp();
    if (lane < q) {
        double o[QT];
#pragma unroll
        for (int c = 0; c < QT; ++c) o[c] = c < q ? cur[lane * ld + s_perm[c]] : 0.0;
#pragma unroll
        for (int c = 0; c < QT; ++c)
            if (c < q) Sinv[lane * QT + c] = o[c];
    }
}

enum { SOLVE_DIRECT = 0, SOLVE_EXCH = 1, SOLVE_PEER = 2 };

// Everything between the front and the streaming update, one launch, SOLVE_NT columns t of P per CTA.
//   DIRECT  one GPU: raw rows and the q x q block of P are read straight from P (no exchange buffer at all)
//   EXCH    NCCL form, second half: raw rows and P H_o' come from the summed local exchange buffer E
//   PEER    fused exchange: every value this rank owns is stored into the exchange buffer of EVERY rank over NVLink peer
//           memory, the CTA publishes its stores (system-scope fence) and bumps its rank's arrival counter on every
//           destination, and only then waits (bounded) for the counters of all sources. On a timeout the step is
//           ABANDONED cleanly: K = 0, X keeps its predicted value, H_o P = 0 -- the update then applies the predict
//           only -- and ASLAM_ST_PEER_TIMEOUT is raised (aslam_large_status returns ASLAM_ERR_TIMEOUT).
// SOLVE_NT threads per CTA (4 warps, at most 168 registers each: one warp per SM sub-partition), so that a CTA of this
// kernel is co-resident with a CTA of the persistent streaming update (8 warps x 168 registers, 180 KB of shared memory)
// on every SM: in the overlapped mode the solve kernel of step k+1 runs under the update of step k, and the update of
// step k+1 -- which may already hold every SM -- spins until this kernel has finished.
constexpr int SOLVE_NT = 128;
// NT = 256 (one CTA per SM, the faster shape when the kernel has the GPU to itself): the in-place form.
template <int QT, int MODE, int NT>
__global__ void __launch_bounds__(NT, NT == SOLVE_NT ? 3 : 1)
large_solve_kernel(uint32_t n, uint32_t rb, uint32_t re, size_t ldp, const double* __restrict__ P,
                   const double* __restrict__ A, double kq, LargeCtx* __restrict__ ctx, uint32_t q_ub,
                   const double* __restrict__ E, double* __restrict__ hp, double* __restrict__ K, double* __restrict__ X,
                   double* __restrict__ predA, const PeerSet ps, unsigned long long target, unsigned long long* sync,
                   unsigned long long epoch, unsigned long long* trace, const PrevUpd pu) {
    // let the streaming update become resident and prefetch its first P tiles while this kernel works
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    static_assert(QT <= 32 && QT + 1 <= 32, "one lane per row");
    constexpr int LD = QT + 1;
    __shared__ double S[QT * LD];
    __shared__ double T[QT * LD];
    __shared__ __align__(16) double Sinv[QT * QT];
    __shared__ double s_hval[QT][5];
    __shared__ int s_hk[QT][5], s_hnnz[QT], s_rows[QT], s_piv[QT], s_perm[QT];
    __shared__ double s_y[QT], s_rdiag[QT], s_invd[QT];
    __shared__ int s_abort, s_chol_ok;
    __shared__ double s_hpp[QT][QT];  // rows rows[k] of the previous step's H_o P (overlapped mode)
    const int tid = threadIdx.x;
    if (tid == 0) {
        s_abort = 0;
        // the front of this step runs on the side stream: wait for its context, A and pose
        if (!spin_until(sync + SYNC_FRONT, epoch)) atomicOr(&ctx->status, (uint32_t)ASLAM_ST_STEP_TIMEOUT);
        trace_min(trace, epoch, 2);
    }
    for (int idx = tid; idx < QT * QT; idx += NT) Sinv[idx] = 0.0;  // padding must not contribute
    __syncthreads();
    const int q = ctx->q;
    const double a00 = A[0], a10 = A[1];
    if (tid < QT) {
        // (loaded unconditionally and masked afterwards: the loads do not wait for q -- every dependent round trip
        // to memory costs several microseconds under a streaming update that keeps the memory system saturated)
        const int r_ = ctx->rows[tid], nz_ = ctx->hnnz[tid];
        const double y_ = ctx->y[tid], rd_ = ctx->rdiag[tid];
        double hv_[5];
        int hk_[5];
#pragma unroll
        for (int e = 0; e < 5; ++e) {
            hv_[e] = ctx->hval[tid][e];
            hk_[e] = ctx->hk[tid][e];
        }
        const bool in = tid < q;
        s_rows[tid] = in ? r_ : 0;
        s_hnnz[tid] = in ? nz_ : 0;
        s_y[tid] = in ? y_ : 0.0;
        s_rdiag[tid] = in ? rd_ : 1.0;
#pragma unroll
        for (int e = 0; e < 5; ++e) {
            s_hval[tid][e] = in ? hv_[e] : 0.0;
            s_hk[tid][e] = in ? hk_[e] : 0;
        }
    }
    if (blockIdx.x == 0 && tid == 0) {  // the streaming update applies the predict with these two entries of A
        predA[0] = a00;
        predA[1] = a10;
    }
    __syncthreads();
    const uint32_t t = blockIdx.x * NT + tid;
    const bool valid = t < n;
    const uint32_t tc = valid ? t : n - 1;
    const bool mine = valid && t >= rb && t < re;
    // ---- (P H_o')(t, :) of the shard's own row t. By the symmetry of P this is also column t of H_o P, which is how
    // the update uses it below: P -= K (P H_o')' instead of P -= K (H_o P) (ekf.cpp:310 read with P = P'). The two
    // differ by the rounding-level asymmetry of P only (1.7e-16 between the covariance forms) and it
    // replaces 19 strided row reads of a column-major matrix per step (155 k scattered sectors, 20 us behind the
    // streaming kernel's prefetch) by 19 contiguous column reads.
    auto at_prev = [&](uint32_t i, uint32_t j) { return pv_entry<QT>(P, ldp, rb, n, i, j, pu); };
    double pht[QT];
    double colv[QT];
    // Rows 0 and 1 of the matrix are the only ones whose predicted entries need more than the stored entry of their own
    // row (pred_entry): one thread per (row, k) computes them side by side -- 2 q entries in one memory round trip
    // instead of the two threads t = 0, 1 walking their q entries one after the other (10 us alone, 50 us under the
    // streaming update: it was the longest chain of the kernel).
    __shared__ double s_c01[2][QT];
    if constexpr (MODE != SOLVE_EXCH) {
        if (blockIdx.x == 0 && rb == 0 && tid < 2 * QT) {
            const int tt = tid / QT, k = tid % QT;
            s_c01[tt][k] = k < q ? pred_entry_acc(at_prev, (uint32_t)tt, (uint32_t)s_rows[k], a00, a10, kq) : 0.0;
        }
    }
    if constexpr (MODE != SOLVE_EXCH) {
        // Overlapped mode: every load this thread needs goes out FIRST (the previous gain of row t, the stored row t at
        // the q columns), then the CTA stages the q rows of the previous H_o P -- one memory round trip, under a
        // streaming update that keeps the memory system saturated, instead of three.
        const bool vfast = pu.K != nullptr && mine && tc >= 2;
        double nk[QT], e[QT];
        double pa00 = 1.0, pa10 = 0.0;
        if (vfast) {
            pa00 = pu.predA[0];
            pa10 = pu.predA[1];
#pragma unroll
            for (int m = 0; m < QT; ++m) nk[m] = pu.K[(size_t)m * n + tc];
            e[0] = P[tc - rb];
            e[1] = P[ldp + (tc - rb)];
            e[2] = P[2 * ldp + (tc - rb)];
#pragma unroll
            for (int k = 3; k < QT; ++k) e[k] = k < q ? P[(size_t)s_rows[k] * ldp + (tc - rb)] : 0.0;
        }
        if (pu.K) {
            for (int idx = tid; idx < QT * QT; idx += NT) {
                const int k1 = idx / QT, m = idx % QT;
                s_hpp[k1][m] = k1 < q ? pu.hp[(size_t)s_rows[k1] * QT + m] : 0.0;
            }
        }
        __syncthreads();  // s_hpp and s_c01 are complete
        if (vfast) {
            // what the update in flight writes for these entries: its predict, then its rank-Q update (k ascending)
            const double p0 = e[0];
            e[0] = p0 * pa00;
            e[1] = p0 * pa10 + e[1];
            if (tc == 2) e[2] = e[2] + pu.kq;
#pragma unroll
            for (int k = 0; k < QT; ++k) {
                double v = e[k];
#pragma unroll
                for (int m = 0; m < QT; ++m) v = fma(-nk[m], s_hpp[k][m], v);
                e[k] = v;
            }
            // ... and this step's predict on top (gather_cols)
            colv[0] = 0 < q ? e[0] * a00 : 0.0;
            colv[1] = 1 < q ? e[0] * a10 + e[1] : 0.0;
            colv[2] = 2 < q ? (tc == 2 ? e[2] + kq : e[2]) : 0.0;
#pragma unroll
            for (int k = 3; k < QT; ++k) colv[k] = k < q ? e[k] : 0.0;
        } else if (mine && tc < 2) {
#pragma unroll
            for (int k = 0; k < QT; ++k) colv[k] = s_c01[tc][k];
        } else {
            gather_cols_v<QT>(P, ldp, rb, n, tc, mine, q, s_rows, a00, a10, kq, pu, s_hpp, colv);
        }
        pht_from_cols<QT>(colv, s_hval, pht);
    } else {
#pragma unroll
        for (int k = 0; k < QT; ++k) pht[k] = colv[k] = 0.0;
    }
    trace_mark(trace, epoch, 8);  // own columns gathered
    if constexpr (MODE == SOLVE_PEER) {
        if (mine) {
            // destination loop outside and ROLLED: with it unrolled inside the 20 rows the kernel grew to 12,000
            // instructions (192 KB) -- more than the instruction cache holds next to the streaming update, and the
            // straight-line factorisation of warp 0 then ran from cold instruction fetches every step (17 us instead of 3)
#pragma unroll 1
            for (int p = 0; p < ps.world; ++p) {
                double* Ep = ps.E[p] + t;
#pragma unroll
                for (int k = 0; k < QT; ++k)
                    if (k < q) Ep[(size_t)k * n] = pht[k];
            }
            // the rows of the q x q block this shard owns: row k1 is what the thread of state row rows[k1] has just
            // gathered -- colv[k2] = predicted P(rows[k1], rows[k2]), the same entry by the same expressions as pred_entry()
            int k1 = -1;
#pragma unroll
            for (int k = 0; k < QT; ++k)
                if (k < q && (uint32_t)s_rows[k] == t) k1 = k;
            if (k1 >= 0) {
#pragma unroll 1
                for (int p = 0; p < ps.world; ++p) {
                    double* Eb = ps.E[p] + exch_psub_offset(q_ub, n) + k1 * QT;
#pragma unroll
                    for (int k2 = 0; k2 < QT; ++k2)
                        if (k2 < q) Eb[k2] = colv[k2];
                }
            }
        }
        // One system-scope fence per CTA, by the thread that signals: the CTA barrier orders every thread's stores
        // before it and the fence is cumulative, so the counter bump below publishes all of them (256 per-thread
        // system fences here cost ~20 us per step).
        __syncthreads();
        if (tid == 0) {
            __threadfence_system();
            // arrivals are counted in groups of SOLVE_NT columns, so that ranks running different CTA shapes (a shard
            // above and one below the overlap threshold) agree on the target: epoch x ceil(n / SOLVE_NT)
            const uint32_t c0 = blockIdx.x * NT;
            const unsigned long long units = (unsigned long long)((min(n, c0 + (uint32_t)NT) - c0 + SOLVE_NT - 1) / SOLVE_NT);
            for (int p = 0; p < ps.world; ++p) atomicAdd_system(ps.flag[p] + ps.rank, units);
        }
        if (tid < ps.world) {  // wait for the other ranks' values
            const unsigned long long* flags = ps.flag[ps.rank];
            unsigned long long v = 0;
            for (int it = 0; it < 4000000; ++it) {
                asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flags + tid) : "memory");
                if (v >= target) break;
                __nanosleep(128);
            }
            if (v < target) {
                s_abort = 1;
                atomicOr(&ctx->status, (uint32_t)ASLAM_ST_PEER_TIMEOUT);
            }
        }
        __syncthreads();
    }
    const bool abort_step = MODE == SOLVE_PEER && s_abort != 0;
    // ---- (P H_o')(t, :) of every row t and the q x q block P(rows, rows)
    if constexpr (MODE == SOLVE_EXCH) {
#pragma unroll
        for (int k = 0; k < QT; ++k) pht[k] = k < q ? *(E + (size_t)k * n + tc) : 0.0;
    }
    if constexpr (MODE == SOLVE_PEER) {  // the shard's own rows are still in registers; the others come from the peers
        if (!mine || abort_step) {
#pragma unroll
            for (int k = 0; k < QT; ++k)
                pht[k] = (k < q && !abort_step) ? ld_exch<true>(E + (size_t)k * n + tc) : 0.0;
        }
    }
    for (int idx = tid; idx < QT * QT; idx += NT) {
        const int k1 = idx / QT, k2 = idx % QT;
        double v = 0.0;
        if (k1 < q && k2 < q && !abort_step) {
            if constexpr (MODE == SOLVE_DIRECT)
                v = pred_entry_acc(at_prev, (uint32_t)s_rows[k1], (uint32_t)s_rows[k2], a00, a10, kq);
            else
                v = ld_exch<MODE == SOLVE_PEER>(E + exch_psub_offset(q_ub, n) + idx);
        }
        S[k1 * LD + k2] = v;  // Psub, staged in S
    }
    __syncthreads();
    trace_mark(trace, epoch, 9);  // exchange waited for, q x q block in shared memory
    // S = H_s Psub H_s' + R (H_s: the q x q sparse form of H_o), zero-padded to QT x QT with the identity on the padding
    for (int idx = tid; idx < QT * QT; idx += NT) {
        const int a = idx / QT, jx = idx % QT;
        double hpaj = 0.0;
        for (int f = 0; f < s_hnnz[a]; ++f) hpaj += s_hval[a][f] * S[s_hk[a][f] * LD + jx];
        T[a * LD + jx] = hpaj;
    }
    __syncthreads();
    for (int idx = tid; idx < QT * QT; idx += NT) {
        const int a = idx / QT, b = idx % QT;
        double sum = 0.0;
        for (int e = 0; e < s_hnnz[b]; ++e) sum += T[a * LD + s_hk[b][e]] * s_hval[b][e];
        S[a * LD + b] = sum + (a == b ? s_rdiag[a] : 0.0);
    }
    __syncthreads();
    trace_mark(trace, epoch, 10);  // S built
    // ---- column t of H_o P (= row t of P H_o', see above): independent of S^-1, on its way before the factorisation
    if (valid) {
        double* hpo = hp + (size_t)t * QT;
#pragma unroll
        for (int k = 0; k < QT; k += 2)
            *reinterpret_cast<double2*>(hpo + k) = abort_step ? make_double2(0.0, 0.0) : make_double2(pht[k], pht[k + 1]);
    }
    // ---- S = L L' and W = L^-1 together, by the whole CTA: thread e < q (q + 1) / 2 owns ONE entry (i, j), i >= j, of
    // the lower triangle -- S(i, j) and the matching entry of X, which starts as the identity and ends as L^-1 (the
    // elimination of [L | I], as in chol_panel_kernel). Right-looking, a rolled loop over the columns with ONE barrier per
    // column: the owners of column k publish l_ik, the owners of row k publish the scaled pivot row of X, and after the
    // barrier every thread applies both rank-1 updates to its own two registers. The next pivot is recomputed by every
    // thread from the published diagonal (same fma as its owner's update), so nobody waits for it. ~50 instructions
    // of code and ~200 cycles per column; the one-warp register version this replaces was 3,500 instructions that each
    // step fetched cold (the kernel runs once per step between two passes of the streaming update) and, rolled into
    // shared memory, paid ~500 cycles per column.
    // If a pivot is not positive, S is not positive definite: S^-1 then comes from the pivoted sweep -- what `S.inverse()`
    // of ekf.cpp:301 does through PartialPivLU -- by warp 0, straight into Sinv.
    {
        double* colb = T;                 // [2][QT]   l_ik of the column in flight
        double* xrb = T + 2 * QT;         // [2][QT]   scaled pivot row of X (zero beyond the pivot)
        double* dgb = T + 4 * QT;         // [2][QT]   diagonal of the trailing matrix
        constexpr int NE = QT * (QT + 1) / 2, EPT = (NE + NT - 1) / NT;  // entries of the triangle per thread
        int ei[EPT], ej[EPT];
        bool own[EPT];
        double sv[EPT], xv[EPT];
#pragma unroll
        for (int u = 0; u < EPT; ++u) {   // e -> (i, j), e = i (i + 1) / 2 + j
            const int e = tid + u * NT;
            int i = (int)((sqrtf(8.0f * (float)e + 1.0f) - 1.0f) * 0.5f);
            while ((i + 1) * (i + 2) / 2 <= e) ++i;
            while (i * (i + 1) / 2 > e) --i;
            ei[u] = i;
            ej[u] = e - i * (i + 1) / 2;
            own[u] = e < NE && i < q;     // (j <= i < q)
            sv[u] = own[u] ? S[ei[u] * LD + ej[u]] : 0.0;
            xv[u] = (own[u] && ei[u] == ej[u]) ? 1.0 : 0.0;
        }
        for (int idx = tid; idx < 6 * QT; idx += NT) T[idx] = 0.0;
        __syncthreads();
#pragma unroll
        for (int u = 0; u < EPT; ++u)
            if (own[u] && ei[u] == ej[u]) dgb[ei[u]] = sv[u];
        __syncthreads();
        double akk = dgb[0];
        bool ok = true;
#pragma unroll 1
        for (int k = 0; k < q; ++k) {
            const int b = k & 1;
            if (!(akk > 0.0)) {  // every thread sees the same pivot: uniform exit
                ok = false;
                break;
            }
            const double inv = rsqrt(akk);
#pragma unroll
            for (int u = 0; u < EPT; ++u) {
                if (own[u] && ej[u] == k) {  // column k of L
                    sv[u] = ei[u] == k ? akk * inv : sv[u] * inv;
                    colb[b * QT + ei[u]] = sv[u];
                }
                if (own[u] && ei[u] == k) {  // pivot row of X
                    xv[u] *= inv;
                    xrb[b * QT + ej[u]] = xv[u];
                }
            }
            __syncthreads();
#pragma unroll
            for (int u = 0; u < EPT; ++u) {
                if (own[u] && ei[u] > k) {
                    const double ci = colb[b * QT + ei[u]];
                    if (ej[u] > k) {
                        sv[u] = fma(-ci, colb[b * QT + ej[u]], sv[u]);
                        if (ei[u] == ej[u]) dgb[(b ^ 1) * QT + ei[u]] = sv[u];
                    }
                    xv[u] = fma(-ci, xrb[b * QT + ej[u]], xv[u]);
                }
            }
            if (k + 1 < q) {  // next pivot: the fma its owner has just done, redone here from the published operands
                const double c1 = colb[b * QT + k + 1];
                akk = fma(-c1, c1, dgb[b * QT + k + 1]);
            }
        }
        if (tid == 0) s_chol_ok = ok ? 1 : 0;
#pragma unroll
        for (int u = 0; u < EPT; ++u)
            if (ok && own[u]) Sinv[ei[u] * QT + ej[u]] = xv[u];  // W = L^-1 (zero above the diagonal)
        trace_mark(trace, epoch, 13);
        if (!ok) {
            __syncthreads();
            if (tid < 32) {
                for (int idx = tid; idx < QT * LD; idx += 32) T[idx] = S[idx];
                for (int idx = tid; idx < QT * QT; idx += 32) Sinv[idx] = 0.0;
                __syncwarp();
                warp_gauss_jordan<QT>(S, T, LD, q, tid, s_piv, s_perm, Sinv);
            }
        }
    }
    __syncthreads();
    trace_mark(trace, epoch, 11);  // factorised, H_o P stored
    if (valid) {
        // K row t = (P H_o')(t, :) S^-1, 20 independent accumulators, rows of the factor read as 16-byte broadcasts.
        //   S = L L', W = L^-1 (in Sinv's storage, zero above the diagonal): S^-1 = W' W, so with u = W p,
        //   K row = W' u -- one pass over the rows of W: u_k = W(k, :) . p, then acc(:) += u_k W(k, :).
        //   (not positive definite: Sinv holds the pivoted-sweep inverse and the pass is the plain product)
        double acc[QT];
#pragma unroll
        for (int k = 0; k < QT; ++k) acc[k] = 0.0;
        if (s_chol_ok) {
#pragma unroll
            for (int k = 0; k < QT; ++k) {
                double w[QT];
#pragma unroll
                for (int a = 0; a <= k; a += 2) {
                    const double2 v = *reinterpret_cast<const double2*>(&Sinv[k * QT + a]);
                    w[a] = v.x;
                    w[a + 1] = v.y;  // (a + 1 == k + 1 reads the zero above the diagonal)
                }
                double u0 = 0.0, u1 = 0.0;
#pragma unroll
                for (int a = 0; a <= k; a += 2) {
                    u0 = fma(w[a], pht[a], u0);
                    if (a + 1 <= k) u1 = fma(w[a + 1], pht[a + 1], u1);
                }
                const double u = u0 + u1;
#pragma unroll
                for (int a = 0; a <= k; ++a) acc[a] = fma(u, w[a], acc[a]);
            }
        } else {
#pragma unroll
            for (int e = 0; e < QT; ++e) {
#pragma unroll
                for (int k = 0; k < QT; k += 2) {
                    const double2 v = *reinterpret_cast<const double2*>(&Sinv[e * QT + k]);
                    acc[k] = fma(pht[e], v.x, acc[k]);
                    acc[k + 1] = fma(pht[e], v.y, acc[k + 1]);
                }
            }
        }
        double dx = 0.0;
#pragma unroll
        for (int k = 0; k < QT; ++k) {
            const double sum = abort_step ? 0.0 : acc[k];
            dx = fma(sum, s_y[k], dx);
            K[(size_t)k * n + t] = sum;
        }
        X[t] += dx;  // ekf.cpp:309
    }
    trace_mark(trace, epoch, 12);  // K and X written
    // ---- the last CTA to get here publishes "X is final, the context is free" for the front of the next step
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        trace_max(trace, epoch, 3);
        const unsigned long long arrived = atomicAdd(sync + SYNC_COUNT, 1ULL) + 1;
        if (arrived == gridDim.x) {
            sync[SYNC_COUNT] = 0;
            __threadfence();
            st_release_gpu(sync + SYNC_SOLVE, epoch);
        }
    }
}

}  // namespace

// ---------------------------------------------------------------------------------- host side
uint32_t large_ekf_qt(uint32_t q_ub) { return q_ub <= 20 ? 20u : (q_ub <= 40 ? 40u : 132u); }

// Per-step view of the mapped peer blocks: the parity-selected exchange buffer and the counters of every rank.
static PeerSet peer_set(const aslam_large* h, uint64_t epoch) {
    PeerSet ps{};
    ps.world = (int32_t)h->peer_world;
    ps.rank = (int32_t)h->peer_rank;
    const size_t ebytes = (size_t)2 * LARGE_QMAX * h->n * sizeof(double);
    for (uint32_t p = 0; p < h->peer_world; ++p) {
        ps.E[p] = reinterpret_cast<double*>(h->peer_base[p] + (epoch & 1) * ebytes);
        ps.flag[p] = reinterpret_cast<unsigned long long*>(h->peer_base[p] + 2 * ebytes);
    }
    return ps;
}

static FrontArgs front_args(const aslam_large* h, double x, double y, float yaw, double vx, double az, float dt,
                            uint32_t m) {
    FrontArgs a;
    a.ox = x; a.oy = y; a.vx = vx; a.az = az; a.yaw = yaw; a.dt = dt;
    a.m = m;
    a.q_ub = 3 + 2 * m;
    a.pi = h->cfg.pi;
    a.thresh = h->cfg.min_dist_thresh;
    a.kr_pose = (double)h->cfg.ekf_kr;
    a.kr_lm = (double)h->cfg.ukf_kr;  // ekf.cpp:278 (UKF constant, as written)
    a.epoch = h->step_epoch + 1;
    for (uint32_t i = 0; i < ASLAM_MAX_OBS; ++i) {
        a.range[i] = i < m ? h->obs_host[i] : 0.0f;
        a.bearing[i] = i < m ? h->obs_host[ASLAM_MAX_OBS + i] : 0.0f;
    }
    return a;
}

// mode: LARGE_STEP_DIRECT (unsharded handle, no exchange buffer), LARGE_STEP_EXCH (the caller sums h->exch between
// phase1 and phase2), LARGE_STEP_PEER (fused exchange over peer memory).
aslam_status large_ekf_phase1(aslam_large* h, double x, double y, float yaw, double vx, double az, float dt,
                              uint32_t m, int mode) {
    const FrontArgs a = front_args(h, x, y, yaw, vx, az, dt, m);
    const uint32_t n = h->n, nb = (n + 255) / 256;
    // every argument / attachment check comes BEFORE the first launch: an error return leaves the handle untouched
    if (mode == LARGE_STEP_PEER && (a.q_ub > 40 || !h->peer_world)) {
        set_error("fused large step: needs peer_attach and at most 18 observations per message");
        return ASLAM_ERR_UNSUPPORTED;
    }
    if (mode == LARGE_STEP_DIRECT && (h->row_begin != 0 || h->row_end != n)) {
        set_error("large step: a sharded handle must use step_phase1 / exchange / step_phase2 or step_fused");
        return ASLAM_ERR_UNSUPPORTED;
    }
    // The front runs on the side stream, so that it overlaps the previous step's streaming update. Ordering against
    // the main stream is carried by the counters in h->sync (see SYNC_*), not by events: the front waits for the
    // previous step's solve stage, this step's solve kernel waits for the front.
    h->step_epoch += 1;
    {
        // The front usually becomes resident BEFORE the update CTA that will share its SM. An SM keeps the shared-memory
        // carve-out of the first kernel that arrives until it is empty again, and the update needs ~180 KB: without
        // this the SM the front sits on cannot take its update CTA, the persistent grid runs that CTA after another one
        // has finished, and the update takes twice as long (measured: 190 -> 461 us).
        static PerDeviceOnce once_;
        int dev_ = -1;
        if (once_.pending(&dev_)) {
            ASLAM_CUDA(cudaFuncSetAttribute(large_front_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                            (int)cudaSharedmemCarveoutMaxShared));
            auto carve = [](auto kernel) {
                return cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                            (int)cudaSharedmemCarveoutMaxShared);
            };
            ASLAM_CUDA(carve(large_solve_kernel<20, SOLVE_DIRECT, SOLVE_NT>));
            ASLAM_CUDA(carve(large_solve_kernel<20, SOLVE_PEER, SOLVE_NT>));
            ASLAM_CUDA(carve(large_solve_kernel<20, SOLVE_DIRECT, 256>));
            ASLAM_CUDA(carve(large_solve_kernel<20, SOLVE_EXCH, 256>));
            ASLAM_CUDA(carve(large_solve_kernel<20, SOLVE_PEER, 256>));
            once_.mark(dev_);
        }
    }
    large_front_kernel<<<1, FRONT_THREADS, 0, h->side>>>(a, n, h->X, h->A, h->ctx, h->sync, h->trace);
    h->launches += 1;
    h->q_host = a.q_ub;
    h->step_mode = mode;
    const double kq = (double)h->cfg.ekf_kq;
    if (a.q_ub <= 20) {  // fast path: no in-place predict, no separate gather (see large_solve_kernel)
        if (mode == LARGE_STEP_PEER) {
            h->peer_epoch += 1;  // the step's epoch (1-based) picks the buffer parity and the arrival target
        } else if (mode == LARGE_STEP_EXCH) {
            ASLAM_CUDA(cudaEventRecord(h->ev_front, h->side));
            ASLAM_CUDA(cudaStreamWaitEvent(h->stream, h->ev_front, 0));
            large_raw_gather_kernel<20><<<nb, 256, 0, h->stream>>>(n, h->row_begin, h->row_end, h->ldp, h->P, h->A, kq,
                                                                   h->ctx, a.q_ub, h->exch);
            h->launches += 1;
        }
        ASLAM_CUDA(cudaGetLastError());
        return ASLAM_OK;
    }
    // round-1 chain for q > 20: stream-ordered behind the front through an event, in place on h->P
    h->pending = false;
    ASLAM_CUDA(cudaEventRecord(h->ev_front, h->side));
    ASLAM_CUDA(cudaStreamWaitEvent(h->stream, h->ev_front, 0));
    large_predict_kernel<<<nb, 256, 0, h->stream>>>(n, h->row_begin, h->row_end, h->ldp, h->P, h->A, kq);
    if (mode == LARGE_STEP_PEER) {
        h->peer_epoch += 1;
        const PeerSet ps = peer_set(h, h->peer_epoch);
        large_gather_fast_kernel<40, true><<<nb, 256, 0, h->stream>>>(n, h->row_begin, h->row_end, h->ldp, h->P,
                                                                      h->ctx, a.q_ub, nullptr, ps);
    } else if (a.q_ub <= 40)
        large_gather_fast_kernel<40, false><<<nb, 256, 0, h->stream>>>(n, h->row_begin, h->row_end, h->ldp, h->P, h->ctx,
                                                                       a.q_ub, h->exch, PeerSet{});
    else
        large_gather_kernel<<<nb, 256, 0, h->stream>>>(n, h->row_begin, h->row_end, h->ldp, h->P, h->ctx, a.q_ub, h->exch);
    ASLAM_CUDA(cudaGetLastError());
    h->launches += 2;
    return ASLAM_OK;
}

// Grid of the streaming update: column ranges x row blocks. The full row blocks fill `waves` whole waves of the
// resident CTAs (one per SM): 12 waves where that still leaves a CTA at least 64 columns to stream (measured on the
// unsharded cfg4 matrix), else 4 (row-sharded handles: short column ranges would spend their time filling the ring).
static dim3 update_grid(const aslam_large* h, uint32_t nrows, uint32_t rows_per_cta, uint32_t cs,
                        uint32_t* cols_per_cta_out) {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
    const uint32_t n = h->n;
    const uint32_t rblocks = (nrows + rows_per_cta - 1) / rows_per_cta;
    const uint32_t full_rblocks = nrows / rows_per_cta ? nrows / rows_per_cta : 1;
    auto ranges = [&](uint32_t w) {
        const uint32_t c = (w * (uint32_t)sms + full_rblocks / 2) / full_rblocks;
        return c < 1 ? 1u : c;
    };
    const uint32_t waves = (n + ranges(12u) - 1) / ranges(12u) >= 64u ? 12u : 4u;
    uint32_t cx = ranges(waves);
    uint32_t cols_per_cta = (n + cx - 1) / cx;
    cols_per_cta = ((cols_per_cta + cs - 1) / cs) * cs;
    cx = (n + cols_per_cta - 1) / cols_per_cta;
    *cols_per_cta_out = cols_per_cta;
    return dim3(cx, rblocks);
}

#ifndef ASLAM_STREAM_NST
#define ASLAM_STREAM_NST 4
#endif
#ifndef ASLAM_STREAM_CS
#define ASLAM_STREAM_CS 12
#endif
// Can a CTA of the solve kernel live on an SM that already holds a CTA of the persistent streaming update? The
// overlapped mode depends on it (the update of step k+1 may hold every SM while it waits for the solve kernel of step
// k+1): checked from the compiled kernels' own resource figures, per SM sub-partition for the registers (allocation
// unit: 256 registers per warp), once per process -- a build whose register counts had grown would otherwise only show
// as step timeouts.
template <class KU, class KS>
static bool solve_fits_beside_update(KU update_kernel, KS solve_kernel, int update_threads, size_t update_dyn_smem) {
    cudaFuncAttributes fu{}, fs{};
    if (cudaFuncGetAttributes(&fu, update_kernel) != cudaSuccess || cudaFuncGetAttributes(&fs, solve_kernel) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    auto per_warp = [](int regs) { return (size_t)((regs * 32 + 255) / 256) * 256; };
    const int wu = (update_threads / 32 + 3) / 4, ws = (SOLVE_NT / 32 + 3) / 4;  // warps per sub-partition
    const bool regs_ok = wu * per_warp(fu.numRegs) + ws * per_warp(fs.numRegs) <= 16384;
    const size_t smem = update_dyn_smem + fu.sharedSizeBytes + fs.sharedSizeBytes + 2 * 1024;  // + 1 KB reserved per CTA
    return regs_ok && smem <= (size_t)228 * 1024 && update_threads + SOLVE_NT <= 2048;
}

static aslam_status large_ekf_phase2_fast(aslam_large* h) {
    const uint32_t n = h->n, nb = (n + SOLVE_NT - 1) / SOLVE_NT, q_ub = h->q_host;
    const uint32_t nrows = h->row_end - h->row_begin;
    const double kq = (double)h->cfg.ekf_kq;
    const unsigned long long epoch = h->step_epoch;
    // ---- Overlapped mode (one GPU per shard, DIRECT and PEER): the solve kernel goes to `solve_stream` and reads the
    // buffer the update IN FLIGHT is reading (P_alt = P_{k-1}) plus that update's K, H_o P and A -- it computes the
    // entries of P_k it needs itself (PrevUpd) -- so it runs UNDER the update of the previous step; this step's update
    // then reads h->P (= P_k, complete by stream order) and writes P_{k+1} into P_alt, and the two pointers swap. The
    // step costs max(update, front + solve) instead of their sum. Not with the NCCL form (the caller's all-reduce sits
    // between the two halves on `stream`) and not with a peer shard on this same GPU (its update could take every SM
    // while this shard's solve kernel still waits for it): those run in place on one stream, as before.
    // ... and only where the update is long enough to hide the small kernels, which run 3-4x slower under it: measured
    // on the cfg4 map (n = 8195) split over N GPUs, overlapped vs in place: N = 1 0.200 vs 0.225 ms, N = 2 0.112 vs 0.139,
    // N = 4 0.077 vs 0.089, N = 8 (1,024-row shards, update 39 us) 0.0614 vs 0.0575 -- so shards that stream less than
    // 192 MB do not share their SMs with the solve kernel (they set some aside for it, below).
    // (h->overlap_min_bytes: 192 MB, or ASLAM_LARGE_OVERLAP_MIN_BYTES from the environment -- the tests set 0 to run the
    // overlapped form on small maps)
    bool overlap = !h->no_overlap && h->step_mode != LARGE_STEP_EXCH &&
                   !(h->step_mode == LARGE_STEP_PEER && h->peer_same_device) &&
                   (size_t)nrows * n * sizeof(double) * 2 >= h->overlap_min_bytes;
    // Short shards (below the threshold, above 32 MB): still overlapped, but with SMs set aside for the solve kernel
    // instead of sharing them. Measured on two GPUs with shards of the N = 8 size (131 MB): in place 0.0548 ms, sharing
    // the SMs 0.0570, 17 / 33 / 48 SMs set aside 0.0492 / 0.0476 / 0.0480; at the N = 4 size (272 MB) 33 SMs aside
    // equal sharing (0.0772 vs 0.0781), at the N = 2 size they lose (0.1387 vs 0.1117: the update needs the SMs).
    int sms_ = 148;
    cudaDeviceGetAttribute(&sms_, cudaDevAttrMultiProcessorCount, h->device);
    const size_t shard_bytes = (size_t)nrows * n * sizeof(double) * 2;
    const int reserve_want = (int)((n + 255) / 256) + 2;
    bool reserve = false;
    if (!overlap && !h->no_overlap && h->reserve_sms != 0 && h->step_mode != LARGE_STEP_EXCH &&
        !(h->step_mode == LARGE_STEP_PEER && h->peer_same_device) && shard_bytes >= h->reserve_min_bytes &&
        reserve_want <= sms_ / 3) {
        overlap = true;
        reserve = true;
    }
    const int reserve_sms = reserve ? (h->reserve_sms > 0 ? h->reserve_sms : reserve_want) : 0;
    if (reserve && reserve_sms < reserve_want - 2) {  // (an override too small for one wave of the solve grid)
        overlap = false;
        reserve = false;
    }
    if (overlap && !reserve) {
        using GU = StreamGeo<7, 2, 20, ASLAM_STREAM_CS, ASLAM_STREAM_NST>;
        static const bool fits =
            solve_fits_beside_update(large_update_stream_kernel<7, 2, 20, ASLAM_STREAM_CS, ASLAM_STREAM_NST>,
                                     large_solve_kernel<20, SOLVE_PEER, SOLVE_NT>, GU::THREADS, GU::SMEM) &&
            solve_fits_beside_update(large_update_stream_kernel<7, 2, 20, ASLAM_STREAM_CS, ASLAM_STREAM_NST>,
                                     large_solve_kernel<20, SOLVE_DIRECT, SOLVE_NT>, GU::THREADS, GU::SMEM);
        if (!fits) overlap = false;
    }
    if (overlap && !h->P_alt) {
        if (cudaMalloc((void**)&h->P_alt, (size_t)n * h->ldp * sizeof(double)) != cudaSuccess) {
            cudaGetLastError();
            h->P_alt = nullptr;
            overlap = false;  // no room for the second buffer: in place
            reserve = false;
        } else {
            // (the padding rows beyond nrows are carried along by the bulk copies: defined values, once)
            ASLAM_CUDA(cudaMemsetAsync(h->P_alt, 0, (size_t)n * h->ldp * sizeof(double), h->stream));
        }
    }
    const int kb = h->kbuf ^ 1;  // the halves of K / hp / predA this step writes (the other ones may still be in use)
    double* Kc = h->K + (size_t)kb * 20 * n;
    double* hpc = h->hp + (size_t)kb * 20 * n;
    double* predAc = h->predA + 2 * kb;
    PrevUpd pu{};
    const double* Pread = h->P;
    cudaStream_t ss = h->stream;
    if (overlap) {
        ss = h->solve_stream;
        if (h->pending) {  // update(k-1) may still be running: wait for everything BEFORE it, read its input
            ASLAM_CUDA(cudaStreamWaitEvent(ss, h->ev_before_upd, 0));
            pu.K = h->K + (size_t)h->kbuf * 20 * n;
            pu.hp = h->hp + (size_t)h->kbuf * 20 * n;
            pu.predA = h->predA + 2 * h->kbuf;
            pu.kq = kq;
            Pread = h->P_alt;
        } else {
            ASLAM_CUDA(cudaEventRecord(h->ev_now, h->stream));
            ASLAM_CUDA(cudaStreamWaitEvent(ss, h->ev_now, 0));
        }
    } else {
        h->pending = false;  // everything on `stream`, in order: h->P is final when the kernels below run
    }
    // CTA shape: SOLVE_NT threads under the update (overlapped), 256 when the kernel has the GPU to itself (in place)
    // `reserve`: the overlapped form for SHORT shards (decided above): the update grid leaves R SMs free and the solve
    // kernel runs in its 256-thread shape, which does not fit beside an update CTA -- so it lands on the free SMs and
    // runs at full speed. R covers the whole solve grid: its CTAs wait for the peers' arrivals, which need ALL of a
    // rank's CTAs to have run, so a second wave could never start.
    auto launch_solve = [&](auto k128, auto k256, const double* E, const PeerSet& ps, unsigned long long target) {
        if (overlap && !reserve)
            k128<<<nb, SOLVE_NT, 0, ss>>>(n, h->row_begin, h->row_end, h->ldp, Pread, h->A, kq, h->ctx, q_ub, E, hpc, Kc, h->X,
                                          predAc, ps, target, h->sync, epoch, h->trace, pu);
        else
            k256<<<(n + 255) / 256, 256, 0, ss>>>(n, h->row_begin, h->row_end, h->ldp, Pread, h->A, kq, h->ctx, q_ub, E, hpc, Kc,
                                                  h->X, predAc, ps, target, h->sync, epoch, h->trace, pu);
    };
    if (h->step_mode == LARGE_STEP_PEER) {
        const PeerSet ps = peer_set(h, h->peer_epoch);
        // arrivals are counted in groups of SOLVE_NT columns whatever the CTA shape (see the kernel)
        const unsigned long long target = (unsigned long long)h->peer_epoch * nb;
        launch_solve(large_solve_kernel<20, SOLVE_PEER, SOLVE_NT>, large_solve_kernel<20, SOLVE_PEER, 256>,
                     ps.E[h->peer_rank], ps, target);
    } else if (h->step_mode == LARGE_STEP_EXCH) {
        launch_solve(large_solve_kernel<20, SOLVE_EXCH, 256>, large_solve_kernel<20, SOLVE_EXCH, 256>, h->exch, PeerSet{},
                     0ull);  // (never overlapped)
    } else {
        launch_solve(large_solve_kernel<20, SOLVE_DIRECT, SOLVE_NT>, large_solve_kernel<20, SOLVE_DIRECT, 256>, nullptr,
                     PeerSet{}, 0ull);
    }
    ASLAM_CUDA(cudaGetLastError());
    const double* Pin = h->P;
    double* Pout = overlap ? h->P_alt : h->P;
    if (overlap) {
        if (h->timed) {  // timing pass: the update's events must not include its wait for the solve kernel
            ASLAM_CUDA(cudaEventRecord(h->ev_solve, ss));
            ASLAM_CUDA(cudaStreamWaitEvent(h->stream, h->ev_solve, 0));
        }
        ASLAM_CUDA(cudaEventRecord(h->ev_before_upd, h->stream));
    }
    if (h->timed) ASLAM_CUDA(cudaEventRecord(h->ev0, h->stream));
    constexpr int NST = ASLAM_STREAM_NST, CW = 7, CS = ASLAM_STREAM_CS;
    using G = StreamGeo<CW, 2, 20, CS, NST>;
    constexpr size_t WS_SMEM = G::SMEM;
    ASLAM_DYN_SMEM((large_update_stream_kernel<CW, 2, 20, CS, NST>), WS_SMEM);
    StreamPlan pl{};
    pl.n = n;
    pl.rb = h->row_begin;
    pl.nrows = nrows;
    pl.nch = (n + CS - 1) / CS;
    pl.nblk = (nrows + (uint32_t)G::ROWS - 1) / (uint32_t)G::ROWS;
    pl.blk_rows = (((nrows + pl.nblk - 1) / pl.nblk + 15) / 16) * 16;
    while ((pl.nblk - 1) * pl.blk_rows >= (uint32_t)h->ldp) pl.nblk -= 1;  // (padding can make the last block empty)
    pl.last_rows = (uint32_t)h->ldp - (pl.nblk - 1) * pl.blk_rows;           // padded to 16 rows, like ldp
    const int sms = sms_;
    cudaLaunchConfig_t lc{};
    lc.gridDim = dim3((unsigned)(reserve ? sms - reserve_sms : sms));  // persistent: one CTA per SM
    lc.blockDim = dim3(G::THREADS);
    lc.dynamicSmemBytes = WS_SMEM;
    lc.stream = h->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at;
    // Early launch needs the two launches to be adjacent in the stream (no timing event between them). With a peer
    // shard on the same GPU the 148 early-resident update CTAs would leave no SM with registers for that peer's solve
    // kernel, which this handle's solve kernel is waiting for: plain stream order there.
    // (only when both kernels are on one stream; overlapped, the update simply follows the previous update)
    const bool early = !overlap && h->pdl && !h->timed && !(h->step_mode == LARGE_STEP_PEER && h->peer_same_device);
    lc.numAttrs = early ? 1 : 0;
    ASLAM_CUDA(cudaLaunchKernelEx(&lc, large_update_stream_kernel<CW, 2, 20, CS, NST>, pl, h->ldp, Pin, Pout,
                                  (const double*)Kc, (const double*)hpc, (const double*)predAc, kq,
                                  (const unsigned long long*)h->sync, epoch, h->trace));
    if (h->timed) ASLAM_CUDA(cudaEventRecord(h->ev1, h->stream));
    ASLAM_CUDA(cudaGetLastError());
    h->launches += 2;
    h->kbuf = kb;
    if (overlap) {
        double* t = h->P;
        h->P = h->P_alt;
        h->P_alt = t;
        h->pending = true;
    }
    return ASLAM_OK;
}

aslam_status large_ekf_phase2(aslam_large* h) {
    if (h->q_host <= 20) return large_ekf_phase2_fast(h);
    const bool peer = h->step_mode == LARGE_STEP_PEER;
    const uint32_t n = h->n, nb = (n + 255) / 256, q_ub = h->q_host;
    const uint32_t QT = large_ekf_qt(q_ub);
    const size_t smem = (size_t)q_ub * q_ub * sizeof(double);
    // generic gain kernel: the opt-in is sized for the largest q it can be asked for, once per device
    if (smem > 48 * 1024) ASLAM_DYN_SMEM(large_gain_kernel, (size_t)LARGE_QMAX * LARGE_QMAX * sizeof(double));
    if (peer) {
        const PeerSet ps = peer_set(h, h->peer_epoch);
        const double* E = ps.E[h->peer_rank];
        const unsigned long long* flags = ps.flag[h->peer_rank];
        const unsigned long long target = (unsigned long long)h->peer_epoch * nb;  // nb gather CTAs per rank and step
        large_gain_fast_kernel<40, true><<<nb, 256, 0, h->stream>>>(n, h->ctx, q_ub, E, h->hp, h->K, h->X, flags, target,
                                                                    (int)h->peer_world);
    } else if (QT == 40)
        large_gain_fast_kernel<40, false><<<nb, 256, 0, h->stream>>>(n, h->ctx, q_ub, h->exch, h->hp, h->K, h->X, nullptr, 0,
                                                                     0);
    else
        large_gain_kernel<<<nb, 256, smem, h->stream>>>(n, h->ctx, q_ub, h->exch, QT, h->hp, h->K, h->X);
    const uint32_t nrows = h->row_end - h->row_begin;
    // X and the context are final for this step once the gain kernel is done (the bulk update kernel reads neither)
    if (QT == 40) large_mark_kernel<<<1, 1, 0, h->stream>>>(h->sync, h->step_epoch);
    if (h->timed) ASLAM_CUDA(cudaEventRecord(h->ev0, h->stream));
    if (QT == 40) {
        using G = BulkGeo<1, 40, 16, 5>;
        ASLAM_DYN_SMEM((large_update_bulk_kernel<1, 40, 16, 5>), G::SMEM);
        uint32_t cols_per_cta = 0;
        const dim3 grid = update_grid(h, nrows, 256u, 16u, &cols_per_cta);
        large_update_bulk_kernel<1, 40, 16, 5><<<grid, 256, G::SMEM, h->stream>>>(n, h->row_begin, nrows, h->ldp, h->P,
                                                                                 h->K, h->hp, cols_per_cta);
    } else {
        dim3 grid((nrows + 255) / 256, (n + 31) / 32);
        large_update_generic_kernel<<<grid, 256, 0, h->stream>>>(n, h->row_begin, nrows, h->ldp, h->P, h->K, h->hp, QT,
                                                               h->ctx);
        large_mark_kernel<<<1, 1, 0, h->stream>>>(h->sync, h->step_epoch);  // this kernel reads the context
    }
    if (h->timed) ASLAM_CUDA(cudaEventRecord(h->ev1, h->stream));
    ASLAM_CUDA(cudaGetLastError());
    h->launches += 2;
    return ASLAM_OK;
}

}  // namespace aslam
