// Large-map EKF-SLAM step (kernels K3-K5): one filter whose covariance
// P (n x n FP64, n = 3 + 2 L) lives in HBM. Same formulas as EKFSlam::slam (src/ekf/ekf.cpp:293-311)
// and updateZandA (ekf.cpp:137-213), with H restricted to the rows in use this step -- the three
// odometry rows and the two rows of every landmark observed in THIS message (row-subset joint
// update; with every landmark observed it is ekf.cpp:300-310 verbatim):
//     S = H_o P H_o' + R_o (q x q),  K = P H_o' S^-1 (n x q),  X += K y,  P -= K (H_o P)
// A = I except A(0,0), A(1,0) (ekf.cpp:210-211) and H_o has <= 5 non-zeros per row, so everything
// but the last product is O(q n); the last one streams P once: read + write 2*8*n^2 bytes.
//
// Kernels, in launch order
//   large_front_kernel    1 CTA   association (float math, first-minimum ties), pose predict, A, H_o, y
//   large_predict_kernel  n thr   P = A P A' + Q on rows/cols 0,1 (ekf.cpp:297)
//   large_gather_kernel   n thr   raw rows rows[k] of P and the local slice of P H_o'  -> exchange buffer
//   (sharded: the caller sums the exchange buffers of all ranks; every element has one owner)
//   large_gain_kernel             S, S^-1 (per CTA, in smem), H_o P, K, X += K y
//   large_update_kernel           P_local -= K_local (H_o P): the HBM-bound streaming kernel
#include <cuda_runtime.h>

#include <cstdlib>

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
};

__global__ void __launch_bounds__(1024) large_front_kernel(FrontArgs a, uint32_t n, double* X, double* A,
                                                          const float* obs, LargeCtx* ctx) {
    __shared__ float s_nb[ASLAM_MAX_OBS];
    __shared__ int s_corr[ASLAM_MAX_OBS];
    __shared__ float s_mind[ASLAM_MAX_OBS];
    __shared__ int s_lm[ASLAM_MAX_OBS], s_ob[ASLAM_MAX_OBS];
    __shared__ int s_cnt;
    __shared__ double s_pose[3];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t L = (n - 3) / 2;
    const double z0 = a.ox, z1 = a.oy, z2 = (double)a.yaw;
    // ---- nearest landmark of every observation. Thread o < m projects observation o (structures.h:104-111); then
    // every thread keeps a few landmarks in registers and scans all observations against them: one pass over X,
    // one round of warp reductions, one CTA barrier, one final reduction per observation.
    __shared__ float2 s_pt[ASLAM_MAX_OBS];
    __shared__ unsigned long long s_red[ASLAM_MAX_OBS][32];
    if (tid < (int)a.m) {
        const float b = normalize_angle(obs[ASLAM_MAX_OBS + tid], a.pi);  // ekf.cpp:150
        s_nb[tid] = b;
        s_pt[tid] = to_point(obs[tid], b, z0, z1, z2);
    }
    __syncthreads();
    constexpr int LPT = 4;  // landmarks per thread and pass
    for (uint32_t o = 0; o < a.m; ++o) s_red[o][warp] = ~0ull;
    for (uint32_t l0 = 0; l0 < L; l0 += 1024 * LPT) {
        float lx[LPT], ly[LPT];
#pragma unroll
        for (int u = 0; u < LPT; ++u) {
            const uint32_t l = l0 + u * 1024 + tid;
            lx[u] = l < L ? (float)X[3 + 2 * l] : 0.0f;
            ly[u] = l < L ? (float)X[4 + 2 * l] : 0.0f;
        }
        for (uint32_t o = 0; o < a.m; ++o) {
            const float2 pt = s_pt[o];
            unsigned long long best = ~0ull;
#pragma unroll
            for (int u = 0; u < LPT; ++u) {
                const uint32_t l = l0 + u * 1024 + tid;
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
    for (uint32_t o = warp; o < a.m; o += 32) {
        unsigned long long best = s_red[o][lane];
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
    } else if (tid == 32) {
        // ---- A (ekf.cpp:206-212) and the pose predict (ekf.cpp:295-296)
        double a00 = A[0], a10 = A[1];
        motion_jacobian(a.vx, a.az, a.dt, z2, a00, a10);
        A[0] = a00;
        A[1] = a10;
        double x = X[0], y = X[1], th = X[2];
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
        ctx->y[kr] = (double)obs[s_ob[c]] - rg;                                          // Z - h(X)
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
        __threadfence_system();
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
// K5: P_local -= K_local (H_o P). Thread owns R consecutive rows; its R x Q slice of -K sits in
// registers for the whole tile. A CTA (256 threads, 256 R rows) walks a range of columns; the
// Q-vector (H_o P)[:, j] of each column is staged in shared memory and read as a broadcast.
// Per column and thread: one 16-byte load, R*Q DFMA, one 16-byte store -- P is read once and
// written once: 16 n^2 bytes per step against 2 q n^2 flop (HBM-bound for q <~ 40).
// ------------------------------------------------------------------------------------------
template <int R, int Q, int U>
__global__ void __launch_bounds__(256) large_update_kernel(uint32_t n, uint32_t rb, uint32_t nrows, size_t ldp,
                                                          double* __restrict__ P, const double* __restrict__ K,
                                                          const double* __restrict__ hp, uint32_t cols_per_cta) {
    constexpr int ROWS = 256 * R;
    constexpr int CW = 64;  // columns staged per smem refill
    __shared__ __align__(16) double s_hp[CW][Q];
    const uint32_t rblk = blockIdx.y;
    const uint32_t i0 = rblk * ROWS + threadIdx.x * R;  // local row
    const uint32_t c_begin = blockIdx.x * cols_per_cta;
    const uint32_t c_end = min(n, c_begin + cols_per_cta);
    double nk[Q][R];
#pragma unroll
    for (int k = 0; k < Q; ++k)
#pragma unroll
        for (int r = 0; r < R; ++r) nk[k][r] = (i0 + r < nrows) ? -K[(size_t)k * n + rb + i0 + r] : 0.0;
    const bool full = i0 + R <= nrows;
    for (uint32_t c0 = c_begin; c0 < c_end; c0 += CW) {
        const uint32_t cw = min((uint32_t)CW, c_end - c0);
        __syncthreads();
        {
            const double* src = hp + (size_t)c0 * Q;
            double* dst = &s_hp[0][0];
            for (uint32_t i = threadIdx.x; i < cw * Q; i += 256) dst[i] = src[i];
        }
        __syncthreads();
        if (i0 >= nrows) continue;
        double* col = P + (size_t)c0 * ldp + i0;
        uint32_t jj = 0;
        if (full) {
            for (; jj + U <= cw; jj += U) {
                double v[U][R];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    if constexpr (R == 2) {
                        const double2 t = __ldcs(reinterpret_cast<const double2*>(col + (size_t)(jj + u) * ldp));
                        v[u][0] = t.x;
                        v[u][1] = t.y;
                    } else {
#pragma unroll
                        for (int r = 0; r < R; ++r) v[u][r] = __ldcs(col + (size_t)(jj + u) * ldp + r);
                    }
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
#pragma unroll
                    for (int k = 0; k < Q; k += 2) {
                        const double2 h = *reinterpret_cast<const double2*>(&s_hp[jj + u][k]);
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            v[u][r] = fma(nk[k][r], h.x, v[u][r]);
                            v[u][r] = fma(nk[k + 1][r], h.y, v[u][r]);
                        }
                    }
                    if constexpr (R == 2) {
                        __stcs(reinterpret_cast<double2*>(col + (size_t)(jj + u) * ldp), make_double2(v[u][0], v[u][1]));
                    } else {
#pragma unroll
                        for (int r = 0; r < R; ++r) __stcs(col + (size_t)(jj + u) * ldp + r, v[u][r]);
                    }
                }
            }
        }
        for (; jj < cw; ++jj) {  // column tail and the ragged last rows
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (i0 + r >= nrows) continue;
                double v = col[(size_t)jj * ldp + r];
#pragma unroll
                for (int k = 0; k < Q; ++k) v = fma(nk[k][r], s_hp[jj][k], v);
                col[(size_t)jj * ldp + r] = v;
            }
        }
    }
}

// K5, bulk-copy form (the one launched for q <= 40). The covariance tile never passes through
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

// K5, warp-specialised form (default). Same tiles and shared-memory ring as the kernel above, but the copy engine is
// driven by a dedicated ninth warp, so the 8 consumer warps never meet in a CTA-wide barrier:
//   producer lane : wait done[st] -> bulk-store the stage -> (one chunk later, once that store has been read out of
//                   shared memory) re-arm full[st] and bulk-load the chunk NST ahead
//   consumer warp : wait full[st] -> update its 64 rows of the stage in place, 4 columns interleaved (8 independent
//                   DFMA chains per thread instead of 2) -> fence.proxy.async -> arrive on done[st]
// ncu on the kernel above (profiles/r01_summary.md): 26 % of the samples sat in __syncthreads behind thread 0's 17 bulk
// instructions per chunk and 37 % in the 20-deep dependent DFMA chains; both are structural and are what this removes.
__device__ __forceinline__ void mbar_arrive_cta(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bulk::smem_u32(bar)) : "memory");
}

template <int R, int Q, int CS, int NST>
__global__ void __launch_bounds__(288, 1) large_update_ws_kernel(uint32_t n, uint32_t rb, uint32_t nrows, size_t ldp,
                                                                double* __restrict__ P, const double* __restrict__ K,
                                                                const double* __restrict__ hp, uint32_t cols_per_cta) {
    using G = BulkGeo<R, Q, CS, NST>;
    static_assert(R == 2 && CS % 4 == 0, "the interleaved inner loop is written for 2 rows x 4 columns");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + NST * G::STAGE);
    uint64_t* done = full + NST;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t r0 = blockIdx.y * G::ROWS;
    const uint32_t rows_here = min((uint32_t)G::ROWS, (uint32_t)ldp - r0);
    const uint32_t seg_bytes = rows_here * sizeof(double);
    const uint32_t c_begin = blockIdx.x * cols_per_cta;
    const uint32_t c_end = min(n, c_begin + cols_per_cta);
    if (c_begin >= c_end) return;
    const uint32_t nchunks = (c_end - c_begin + CS - 1) / CS;
    auto stage_p = [&](int st) { return reinterpret_cast<double*>(smem_raw + st * G::STAGE); };
    auto stage_h = [&](int st) { return reinterpret_cast<double*>(smem_raw + st * G::STAGE + G::P_BYTES); };

    if (tid == 0) {
        for (int s = 0; s < NST; ++s) {
            bulk::mbar_init(&full[s], 1);
            bulk::mbar_init(&done[s], 8);
        }
        bulk::fence_barrier_init();
    }
    __syncthreads();

    if (warp == 8) {  // ---------------- producer
        if (lane != 0) return;
        auto issue_load = [&](uint32_t chunk) {
            const int st = chunk % NST;
            const uint32_t c0 = c_begin + chunk * CS;
            const uint32_t cw = min((uint32_t)CS, c_end - c0);
            bulk::mbar_expect_tx(&full[st], cw * seg_bytes + cw * Q * (uint32_t)sizeof(double));
            double* dp = stage_p(st);
            for (uint32_t c = 0; c < cw; ++c)
                bulk::g2s(dp + c * G::ROWS, P + (size_t)(c0 + c) * ldp + r0, seg_bytes, &full[st]);
            bulk::g2s(stage_h(st), hp + (size_t)c0 * Q, cw * Q * (uint32_t)sizeof(double), &full[st]);
        };
        for (uint32_t c = 0; c < (uint32_t)(NST - 1) && c < nchunks; ++c) issue_load(c);
        for (uint32_t chunk = 0; chunk < nchunks; ++chunk) {
            const int st = chunk % NST;
            const uint32_t c0 = c_begin + chunk * CS;
            const uint32_t cw = min((uint32_t)CS, c_end - c0);
            bulk::mbar_wait(&done[st], (chunk / NST) & 1u);
            const double* sp = stage_p(st);
            for (uint32_t c = 0; c < cw; ++c) bulk::s2g(P + (size_t)(c0 + c) * ldp + r0, sp + c * G::ROWS, seg_bytes);
            bulk::commit_group();
            const uint32_t nxt = chunk + NST - 1;  // goes into the stage drained one chunk ago
            if (nxt < nchunks) {
                bulk::wait_group_read<1>();
                issue_load(nxt);
            }
        }
        bulk::wait_group<0>();
        return;
    }

    // ---------------- consumers
    const uint32_t i0 = r0 + tid * R;
    double nk[Q][R];
#pragma unroll
    for (int k = 0; k < Q; ++k)
#pragma unroll
        for (int r = 0; r < R; ++r) nk[k][r] = (i0 + r < nrows) ? -K[(size_t)k * n + rb + i0 + r] : 0.0;
    const bool active = (uint32_t)(tid * R) < rows_here;

    for (uint32_t chunk = 0; chunk < nchunks; ++chunk) {
        const int st = chunk % NST;
        const uint32_t cw = min((uint32_t)CS, c_end - (c_begin + chunk * CS));
        bulk::mbar_wait(&full[st], (chunk / NST) & 1u);
        if (active) {
            double* sp = stage_p(st) + tid * R;
            const double* sh = stage_h(st);
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

aslam_status large_ekf_phase1(aslam_large* h, double x, double y, float yaw, double vx, double az, float dt,
                              uint32_t m, bool peer) {
    FrontArgs a;
    a.ox = x; a.oy = y; a.vx = vx; a.az = az; a.yaw = yaw; a.dt = dt;
    a.m = m;
    a.q_ub = 3 + 2 * m;
    a.pi = h->cfg.pi;
    a.thresh = h->cfg.min_dist_thresh;
    a.kr_pose = (double)h->cfg.ekf_kr;
    a.kr_lm = (double)h->cfg.ukf_kr;  // ekf.cpp:278 (UKF constant, as written)
    const uint32_t n = h->n, nb = (n + 255) / 256;
    // front on the side stream (stage_obs already made it wait for ev_ctx_free): it overlaps the previous step's
    // streaming update; predict / gather need both the front's context and the updated P
    ASLAM_CUDA(cudaStreamWaitEvent(h->side, h->ev_ctx_free, 0));
    large_front_kernel<<<1, 1024, 0, h->side>>>(a, n, h->X, h->A, h->d_obs, h->ctx);
    ASLAM_CUDA(cudaEventRecord(h->ev_front, h->side));
    ASLAM_CUDA(cudaStreamWaitEvent(h->stream, h->ev_front, 0));
    large_predict_kernel<<<nb, 256, 0, h->stream>>>(n, h->row_begin, h->row_end, h->ldp, h->P, h->A,
                                                    (double)h->cfg.ekf_kq);
    if (peer) {
        // the step's epoch (1-based) picks the buffer parity; phase2 waits for epoch * nb arrivals per source
        if (a.q_ub > 40 || !h->peer_world) {
            set_error("fused large step: needs peer_attach and at most 18 observations per message");
            return ASLAM_ERR_UNSUPPORTED;
        }
        h->peer_epoch += 1;
        const PeerSet ps = peer_set(h, h->peer_epoch);
        if (a.q_ub <= 20)
            large_gather_fast_kernel<20, true><<<nb, 256, 0, h->stream>>>(n, h->row_begin, h->row_end, h->ldp, h->P,
                                                                          h->ctx, a.q_ub, nullptr, ps);
        else
            large_gather_fast_kernel<40, true><<<nb, 256, 0, h->stream>>>(n, h->row_begin, h->row_end, h->ldp, h->P,
                                                                          h->ctx, a.q_ub, nullptr, ps);
    } else if (a.q_ub <= 20)
        large_gather_fast_kernel<20, false><<<nb, 256, 0, h->stream>>>(n, h->row_begin, h->row_end, h->ldp, h->P, h->ctx,
                                                                       a.q_ub, h->exch, PeerSet{});
    else if (a.q_ub <= 40)
        large_gather_fast_kernel<40, false><<<nb, 256, 0, h->stream>>>(n, h->row_begin, h->row_end, h->ldp, h->P, h->ctx,
                                                                       a.q_ub, h->exch, PeerSet{});
    else
        large_gather_kernel<<<nb, 256, 0, h->stream>>>(n, h->row_begin, h->row_end, h->ldp, h->P, h->ctx, a.q_ub, h->exch);
    ASLAM_CUDA(cudaGetLastError());
    h->launches += 3;
    h->q_host = a.q_ub;
    return ASLAM_OK;
}

aslam_status large_ekf_phase2(aslam_large* h, bool peer) {
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
        if (QT == 20)
            large_gain_fast_kernel<20, true><<<nb, 256, 0, h->stream>>>(n, h->ctx, q_ub, E, h->hp, h->K, h->X, flags, target,
                                                                        (int)h->peer_world);
        else
            large_gain_fast_kernel<40, true><<<nb, 256, 0, h->stream>>>(n, h->ctx, q_ub, E, h->hp, h->K, h->X, flags, target,
                                                                        (int)h->peer_world);
    } else if (QT == 20)
        large_gain_fast_kernel<20, false><<<nb, 256, 0, h->stream>>>(n, h->ctx, q_ub, h->exch, h->hp, h->K, h->X, nullptr, 0,
                                                                     0);
    else if (QT == 40)
        large_gain_fast_kernel<40, false><<<nb, 256, 0, h->stream>>>(n, h->ctx, q_ub, h->exch, h->hp, h->K, h->X, nullptr, 0,
                                                                     0);
    else
        large_gain_kernel<<<nb, 256, smem, h->stream>>>(n, h->ctx, q_ub, h->exch, QT, h->hp, h->K, h->X);
    const uint32_t nrows = h->row_end - h->row_begin;
    // X and the context are final for this step once the gain kernel is done (the fast update kernels read neither)
    if (QT == 20 || QT == 40) ASLAM_CUDA(cudaEventRecord(h->ev_ctx_free, h->stream));
    if (h->timed) ASLAM_CUDA(cudaEventRecord(h->ev0, h->stream));
    if (QT == 20 || QT == 40) {
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
        static const char* variant = std::getenv("ASLAM_LARGE_UPDATE");  // "ldg": register-path kernel
        static const char* waves_env = std::getenv("ASLAM_LARGE_WAVES");
        const bool use_ldg = variant && variant[0] == 'l';
        const uint32_t rows_per_cta = QT == 20 ? 512u : 256u;
        const uint32_t rblocks = (nrows + rows_per_cta - 1) / rows_per_cta;
        const uint32_t full_rblocks = nrows / rows_per_cta ? nrows / rows_per_cta : 1;
        // column ranges: the full row blocks fill `waves` whole waves of the resident CTAs
        const uint32_t resident = (uint32_t)sms * (use_ldg ? 2u : 1u);
        const uint32_t cs = QT == 20 ? 8u : 16u;
        // 12 waves where that still leaves a CTA at least 64 columns to stream (measured on the unsharded cfg4 matrix:
        // 201 -> 196 us), else 4 (row-sharded handles: short column ranges would spend their time filling the ring)
        auto ranges = [&](uint32_t w) {
            const uint32_t c = (w * resident + full_rblocks / 2) / full_rblocks;
            return c < 1 ? 1u : c;
        };
        uint32_t waves = (n + ranges(12u) - 1) / ranges(12u) >= 64u ? 12u : 4u;
        if (waves_env) waves = (uint32_t)std::atoi(waves_env);
        uint32_t cx = ranges(waves);
        uint32_t cols_per_cta = (n + cx - 1) / cx;
        cols_per_cta = ((cols_per_cta + cs - 1) / cs) * cs;
        cx = (n + cols_per_cta - 1) / cols_per_cta;
        dim3 grid(cx, rblocks);
        if (use_ldg) {
            if (QT == 20)
                large_update_kernel<2, 20, 4><<<grid, 256, 0, h->stream>>>(n, h->row_begin, nrows, h->ldp, h->P, h->K,
                                                                          h->hp, cols_per_cta);
            else
                large_update_kernel<1, 40, 4><<<grid, 256, 0, h->stream>>>(n, h->row_begin, nrows, h->ldp, h->P, h->K,
                                                                          h->hp, cols_per_cta);
        } else if (QT == 20 && !(variant && variant[0] == 'b')) {  // default; ASLAM_LARGE_UPDATE=bulk: the one-barrier kernel
            using G = BulkGeo<2, 20, 8, 5>;
            constexpr size_t WS_SMEM = G::SMEM + 5 * sizeof(uint64_t);
            ASLAM_DYN_SMEM((large_update_ws_kernel<2, 20, 8, 5>), WS_SMEM);
            large_update_ws_kernel<2, 20, 8, 5><<<grid, 288, WS_SMEM, h->stream>>>(n, h->row_begin, nrows, h->ldp, h->P,
                                                                                  h->K, h->hp, cols_per_cta);
        } else if (QT == 20) {
            using G = BulkGeo<2, 20, 8, 5>;
            ASLAM_DYN_SMEM((large_update_bulk_kernel<2, 20, 8, 5>), G::SMEM);
            large_update_bulk_kernel<2, 20, 8, 5><<<grid, 256, G::SMEM, h->stream>>>(n, h->row_begin, nrows, h->ldp, h->P,
                                                                                    h->K, h->hp, cols_per_cta);
        } else {
            using G = BulkGeo<1, 40, 16, 5>;
            ASLAM_DYN_SMEM((large_update_bulk_kernel<1, 40, 16, 5>), G::SMEM);
            large_update_bulk_kernel<1, 40, 16, 5><<<grid, 256, G::SMEM, h->stream>>>(n, h->row_begin, nrows, h->ldp, h->P,
                                                                                     h->K, h->hp, cols_per_cta);
        }
    } else {
        dim3 grid((nrows + 255) / 256, (n + 31) / 32);
        large_update_generic_kernel<<<grid, 256, 0, h->stream>>>(n, h->row_begin, nrows, h->ldp, h->P, h->K, h->hp, QT,
                                                               h->ctx);
        ASLAM_CUDA(cudaEventRecord(h->ev_ctx_free, h->stream));  // this kernel reads the context
    }
    if (h->timed) ASLAM_CUDA(cudaEventRecord(h->ev1, h->stream));
    ASLAM_CUDA(cudaGetLastError());
    h->launches += 2;
    return ASLAM_OK;
}

}  // namespace aslam
