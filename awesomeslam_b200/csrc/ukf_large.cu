// Large-map UKF-SLAM step (kernels K6-K9): UKFSlam::updateZ + slam (src/ukf/ukf.cpp:113-180,
// 260-392) for ONE filter with thousands of landmarks (cfg5: n = 2051, 4107 sigma points). Same formulas as the
// reference; what changes is how the O(n^3) part is evaluated.
//
//   reference (ukf.cpp)                         here
//   L = Paug.llt().matrixL()           :280     blocked Cholesky of P on the DMMA pipe (chol_blocked.cu); the two
//                                               augmented rows are decoupled (Paug is block diagonal): L(N,N) = std_a
//   XsigAug, XsigPred (f per column)   :283-297 never materialised for rows >= 3 (f only changes rows 0..2): the pose rows
//                                               of all 2N+5 columns are propagated by sigma_pose_kernel
//   X = sum w_i XsigPred_i             :300-304 sigma_mean_kernel (one CTA per state row, FP64)
//   P = sum w_i d_i d_i' + Q           :307-319 D = XsigPred - X (row 2 wrapped) stored [n][NS]; every weight but w_0 is
//                                               the same float (ukf.h:78-79), so  P = wc D D' + (w0 - wc) d_0 d_0' + Q:
//                                               one TMA + DMMA GEMM with the rank-1 term and Q fused into its epilogue
//   Zsig = h(XsigPred_i), Zpred        :322-339 zsig_kernel, zdiff_kernel (wraps as written, float)
//   S, Tc                              :342-375 two more GEMMs of the same kind (S lower tiles only)
//   K = Tc S^-1; X += K dz; P -= K S K':378-391 w_0 = (1-N)/3 << 0 makes S INDEFINITE for large maps (the reference does not
//                                               care: S.inverse() is an LU). S = M + c u u' with M = wc DZ DZ' + R positive
//                                               definite, c = w0 - wc, u = dz_0, so with M = Lm Lm' (Cholesky), W = Tc Lm^-T,
//                                               a = Lm^-1 u, b = Lm^-1 dz, g = W a, den = 1 + c a.a (Sherman-Morrison):
//                                                 K S K' = Tc S^-1 Tc' = W W' - (c/den) g g'
//                                                 K dz   = Tc S^-1 dz  = W b  - (c/den) (a.b) g
//                                               Identical in exact arithmetic; no S^-1, no K, no pivoting.
// All matrices are row-major with the contraction index contiguous. P is kept exactly symmetric (lower triangle
// computed, then mirrored); the reference's P is symmetric to rounding.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <new>

#include "dense_fp64.hpp"
#include "device_math.cuh"
#include "large.hpp"

namespace aslam {

namespace {

struct UkfLarge {
    int n = 0, na = 0, ns = 0;
    size_t ld = 0;     // pitch of the n x n matrices (== h->ldp)
    size_t nsld = 0;   // pitch of the n x NS matrices
    double* Z = nullptr;       // [n] persistent measurement vector (ukf.cpp:154-155)
    double* Lb = nullptr;      // [n][ld] Cholesky factor of P (lower)
    double* S = nullptr;       // [n][ld]
    double* Tc = nullptr;      // [n + 2][ld], later W; rows n, n+1 carry dz_0' and dz' through the triangular solve
    double* D = nullptr;       // [n][nsld]
    double* DZ = nullptr;      // [n][nsld] Zsig, then Zdiff
    double* pose = nullptr;    // [3][nsld] rows 0..2 of XsigPred
    double* mean = nullptr;    // [n]
    double* zpred = nullptr;   // [n]
    double* d0 = nullptr;      // [n] column 0 of D
    double* dz0 = nullptr;     // [n] column 0 of DZ
    double* dzv = nullptr;     // [n] wrap(Z - Zpred), then Lm^-1 of it
    double* gvec = nullptr;    // [n] g = W Lm^-1 dz_0
    double* coef = nullptr;    // [2] Sherman-Morrison scalars
    double* qdiag = nullptr;   // [n] diagonal of Q
    double* rdiag = nullptr;   // [n] diagonal of R
    double* invP = nullptr;    // [nb][64*64]
    double* invS = nullptr;
    double* inv256 = nullptr;  // [ceil(n/256)][256*256] inverses of the 256-wide diagonal blocks of chol(M), and transposed
    double* inv256T = nullptr;
    // TF32 variant of the three sigma-point products (gemm_tf32_tcgen05.cu): 0 = FP64 (default), 1 = TF32, 3 = 3xTF32
    int precision = 0;
    float* Dhi = nullptr;      // [n][tf32_padded_k(ns)] hi / lo parts of D and DZ
    float* Dlo = nullptr;
    float* Zhi = nullptr;
    float* Zlo = nullptr;
    int* tf32_err = nullptr;
    cudaEvent_t evg0 = nullptr, evg1 = nullptr;  // around the three sigma-point products (incl. the TF32 split)
    int* fail = nullptr;       // [2]: first failing pivot of chol(P), chol(S)
    uint32_t* status = nullptr;
    float lambda = 0, wsq = 0;
    double w0 = 0, wc = 0;
};

struct UkfFrontArgs {
    double ox, oy;
    float yaw;
    uint32_t m;
    float pi, thresh;
};

__device__ __forceinline__ unsigned long long umin64(unsigned long long a, unsigned long long b) { return a < b ? a : b; }

// updateZ (ukf.cpp:113-180) against the supplied map: Z(0..2) <- odometry; every observation goes to its nearest
// landmark (float distances, first minimum on ties, strict < MIN_DIST_THRESH) and overwrites that landmark's Z
// entries in message order; unmatched observations would enter the wait-list (map growth is outside the large-map
// contract) and are reported as -1.
__global__ void __launch_bounds__(1024) ukf_front_kernel(UkfFrontArgs a, uint32_t n, const double* X, double* Z,
                                                        const float* obs, LargeCtx* ctx) {
    __shared__ unsigned long long red[32];
    const int tid = threadIdx.x;
    const uint32_t L = (n - 3) / 2;
    const double z0 = a.ox, z1 = a.oy, z2 = (double)a.yaw;
    for (uint32_t o = 0; o < a.m; ++o) {
        const float r = obs[o];
        const float b = normalize_angle(obs[ASLAM_MAX_OBS + o], a.pi);  // ukf.cpp:126
        const float2 pt = to_point(r, b, z0, z1, z2);
        unsigned long long best = ~0ull;
        for (uint32_t l = tid; l < L; l += blockDim.x) {
            const float d = euler_distance(pt.x, pt.y, (float)X[3 + 2 * l], (float)X[4 + 2 * l]);
            best = umin64(best, ((unsigned long long)__float_as_uint(d) << 32) | l);
        }
        for (int off = 16; off; off >>= 1) best = umin64(best, __shfl_xor_sync(0xffffffffu, best, off));
        if ((tid & 31) == 0) red[tid >> 5] = best;
        __syncthreads();
        if (tid < 32) {
            unsigned long long v = tid < (int)(blockDim.x >> 5) ? red[tid] : ~0ull;
            for (int off = 16; off; off >>= 1) v = umin64(v, __shfl_xor_sync(0xffffffffu, v, off));
            if (tid == 0) {
                const int corr = (int)(v & 0xffffffffu) * 2;
                const float mind = __uint_as_float((unsigned)(v >> 32));
                if (L > 0 && mind < a.thresh) {  // ukf.cpp:151-157
                    Z[3 + corr] = (double)r;
                    Z[4 + corr] = (double)b;
                    ctx->assoc[o] = corr;
                } else {
                    ctx->assoc[o] = -1;
                }
            }
        }
        __syncthreads();
    }
    if (tid == 0) {
        Z[0] = z0;
        Z[1] = z1;
        Z[2] = z2;
    }
}

struct SigmaArgs {
    int n, na, ns;
    size_t ld, nsld;
    float vx, az, dt, pi, wsq;
    double la;  // L(N,N) = sqrt(Paug(N,N)) (ukf.cpp:276,280)
    double w0, wc;
};

// entry (k, j) of L as the reference sees it: zero above the diagonal (matrixL()), k < n
__device__ __forceinline__ double lower_at(const double* Lb, size_t ld, int k, int j) {
    return j <= k ? Lb[(size_t)k * ld + j] : 0.0;
}
// sigma column i (0 .. 2N+4) -> (sign, column of L), ukf.cpp:283-289
__device__ __forceinline__ void sigma_col(int i, int na, double& sgn, int& j) {
    if (i == 0) {
        sgn = 0.0;
        j = 0;
    } else if (i <= na) {
        sgn = 1.0;
        j = i - 1;
    } else {
        sgn = -1.0;
        j = i - na - 1;
    }
}
// XsigAug(k, i) for a landmark row k >= 3, which f leaves unchanged: Xaug(k) +- w L(k, j), rounded as written
__device__ __forceinline__ double sigma_entry(const double* X, const double* Lb, size_t ld, int k, double sgn, int j,
                                              double w) {
    const double x = X[k];
    if (sgn == 0.0) return x;
    const double wl = __dmul_rn(w, lower_at(Lb, ld, k, j));
    return sgn > 0.0 ? __dadd_rn(x, wl) : __dsub_rn(x, wl);
}

// rows 0..2 of XsigPred for every sigma column: the augmented point through f (common.h:46-75), theta wrapped
// (ukf.cpp:292-297)
__global__ void sigma_pose_kernel(SigmaArgs a, const double* X, const double* Lb, double* pose) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.ns) return;
    double sgn;
    int j;
    sigma_col(i, a.na, sgn, j);
    const double w = (double)a.wsq;
    double p[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) p[k] = (j < a.n) ? sigma_entry(X, Lb, a.ld, k, sgn, j, w) : X[k];
    double noise = 0.0;  // XsigAug(N, i): Xaug(N) = 0 +- w L(N, N) only in the column of L(:, N)
    if (sgn != 0.0 && j == a.n) noise = sgn > 0.0 ? __dadd_rn(0.0, __dmul_rn(w, a.la)) : __dsub_rn(0.0, __dmul_rn(w, a.la));
    motion_model_aug(p[0], p[1], p[2], noise, a.vx, a.az, a.dt);
    pose[i] = p[0];
    pose[a.nsld + i] = p[1];
    pose[2 * a.nsld + i] = (double)normalize_angle((float)p[2], a.pi);
}

__device__ __forceinline__ double block_sum(double v, double* sh) {
    for (int off = 16; off; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    double t = 0.0;
    const int nw = blockDim.x >> 5;
    for (int k = 0; k < nw; ++k) t += sh[k];
    return t;
}

// X = sum_i w_i XsigPred(:, i) (ukf.cpp:300-304). One CTA per state row.
__global__ void __launch_bounds__(256) sigma_mean_kernel(SigmaArgs a, const double* X, const double* Lb,
                                                        const double* pose, double* mean) {
    __shared__ double sh[8];
    const int k = blockIdx.x;
    const double w = (double)a.wsq;
    double acc = 0.0;
    for (int i = 1 + threadIdx.x; i < a.ns; i += 256) {
        double v;
        if (k < 3) {
            v = pose[(size_t)k * a.nsld + i];
        } else {
            double sgn;
            int j;
            sigma_col(i, a.na, sgn, j);
            v = (j < a.n) ? sigma_entry(X, Lb, a.ld, k, sgn, j, w) : X[k];
        }
        acc += v;
    }
    const double s = block_sum(acc, sh);
    if (threadIdx.x == 0) {
        const double c0 = k < 3 ? pose[(size_t)k * a.nsld] : X[k];
        mean[k] = __dadd_rn(__dmul_rn(a.w0, c0), __dmul_rn(a.wc, s));
    }
}

// D(k, i) = XsigPred(k, i) - X(k), row 2 wrapped (ukf.cpp:310-313); d0 = column 0
__global__ void __launch_bounds__(256) xdiff_kernel(SigmaArgs a, const double* X, const double* Lb, const double* pose,
                                                   const double* mean, double* D, double* d0, float pi) {
    const int k = blockIdx.y;
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= (int)a.nsld) return;
    double d = 0.0;
    if (i < a.ns) {
        double v;
        if (k < 3) {
            v = pose[(size_t)k * a.nsld + i];
        } else {
            double sgn;
            int j;
            sigma_col(i, a.na, sgn, j);
            v = (j < a.n) ? sigma_entry(X, Lb, a.ld, k, sgn, j, (double)a.wsq) : X[k];
        }
        d = __dsub_rn(v, mean[k]);
        if (k == 2) d = (double)normalize_angle((float)d, pi);
        if (i == 0) d0[k] = d;
    }
    D[(size_t)k * a.nsld + i] = d;
}

// Zsig = h(XsigPred(:, i)) (ukf.cpp:322-326, common.h:78-90). One thread per (landmark, sigma column); rows 0..2 are
// the identity part of h.
__global__ void __launch_bounds__(256) zsig_kernel(SigmaArgs a, const double* X, const double* Lb, const double* pose,
                                                  double* ZS) {
    const int l = blockIdx.y;  // landmark; gridDim.y = L + 1, the last row block copies the pose rows
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= a.ns) return;
    const int L = (a.n - 3) / 2;
    if (l == L) {
#pragma unroll
        for (int k = 0; k < 3; ++k) ZS[(size_t)k * a.nsld + i] = pose[(size_t)k * a.nsld + i];
        return;
    }
    double sgn;
    int j;
    sigma_col(i, a.na, sgn, j);
    const double w = (double)a.wsq;
    const int kx = 3 + 2 * l, ky = kx + 1;
    const double lx = (j < a.n) ? sigma_entry(X, Lb, a.ld, kx, sgn, j, w) : X[kx];
    const double ly = (j < a.n) ? sigma_entry(X, Lb, a.ld, ky, sgn, j, w) : X[ky];
    double rg, br;
    measure_landmark(pose[i], pose[a.nsld + i], pose[2 * a.nsld + i], lx, ly, rg, br);
    ZS[(size_t)kx * a.nsld + i] = rg;
    ZS[(size_t)ky * a.nsld + i] = br;
}

// Zpred = sum w_i Zsig_i, bearings wrapped (ukf.cpp:329-339); Zdiff = Zsig - Zpred wrapped per column (:346-351),
// in place; dz0 = column 0; dzv = wrap(Z - Zpred) (:381-386). One CTA per row.
__global__ void __launch_bounds__(256) zdiff_kernel(SigmaArgs a, double* ZS, const double* Z, double* zpred, double* dz0,
                                                   double* dzv) {
    __shared__ double sh[8];
    __shared__ double s_zp;
    const int k = blockIdx.x;
    double* row = ZS + (size_t)k * a.nsld;
    double acc = 0.0;
    for (int i = 1 + threadIdx.x; i < a.ns; i += 256) acc += row[i];
    const double s = block_sum(acc, sh);
    const bool wrap = k >= 2 && (k & 1) == 0;  // indices 2, 4, 6, ... (the `j < N-1; j += 2` loops)
    if (threadIdx.x == 0) {
        double zp = __dadd_rn(__dmul_rn(a.w0, row[0]), __dmul_rn(a.wc, s));
        if (wrap) zp = (double)normalize_angle((float)zp, a.pi);
        s_zp = zp;
        zpred[k] = zp;
        double r = __dsub_rn(Z[k], zp);
        if (wrap) r = (double)normalize_angle((float)r, a.pi);
        dzv[k] = r;
    }
    __syncthreads();
    const double zp = s_zp;
    for (int i = threadIdx.x; i < (int)a.nsld; i += 256) {
        double d = 0.0;
        if (i < a.ns) {
            d = __dsub_rn(row[i], zp);
            if (wrap) d = (double)normalize_angle((float)d, a.pi);
            if (i == 0) dz0[k] = d;
        }
        row[i] = d;
    }
}

// coef[0] = c / den, coef[1] = -(c / den) (a . b), den = 1 + c (a . a)  (Sherman-Morrison scalars). One CTA.
__global__ void __launch_bounds__(256) sm_scalars_kernel(int n, const double* a, const double* b, double c, double* coef,
                                                        uint32_t* status) {
    __shared__ double sh[8];
    double saa = 0.0, sab = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) {
        saa = fma(a[i], a[i], saa);
        sab = fma(a[i], b[i], sab);
    }
    const double taa = block_sum(saa, sh);
    const double tab = block_sum(sab, sh);
    if (threadIdx.x == 0) {
        const double den = 1.0 + c * taa;
        const double k = c / den;
        coef[0] = k;
        coef[1] = -k * tab;
        if (!isfinite(k)) atomicOr(status, (uint32_t)ASLAM_ST_NONFINITE);
    }
}

// g = W a;  X = mean + W b + coef[1] g  (ukf.cpp:389). One warp per row. Flags non-finite results.
__global__ void __launch_bounds__(256) state_update_kernel(int n, size_t ld, const double* W, const double* a,
                                                          const double* b, const double* coef, const double* mean,
                                                          double* g, double* X, uint32_t* status) {
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= n) return;
    const double* wr = W + (size_t)row * ld;
    double ga = 0.0, gb = 0.0;
    for (int j = lane; j < n; j += 32) {
        const double w = wr[j];
        ga = fma(w, a[j], ga);
        gb = fma(w, b[j], gb);
    }
    for (int off = 16; off; off >>= 1) {
        ga += __shfl_xor_sync(0xffffffffu, ga, off);
        gb += __shfl_xor_sync(0xffffffffu, gb, off);
    }
    if (lane == 0) {
        const double x = mean[row] + gb + coef[1] * ga;
        g[row] = ga;
        X[row] = x;
        if (!isfinite(x)) atomicOr(status, (uint32_t)ASLAM_ST_NONFINITE);
    }
}

__global__ void fill_diag_kernel(int n, double* q, double* r, double kq, double kr) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    q[i] = i < 3 ? kq : 0.0;  // ukf.cpp:63-65; grown entries of Q are 0 (ukf.cpp:240)
    r[i] = kr;                // ukf.cpp:59,242
}

__global__ void fail_to_status_kernel(const int* fail, const int* tf32_err, uint32_t* status) {
    if (fail[0] >= 0 || fail[1] >= 0) atomicOr(status, (uint32_t)ASLAM_ST_LLT_FAILED);
    if (tf32_err && *tf32_err) atomicOr(status, (uint32_t)ASLAM_ST_STEP_TIMEOUT);  // a bounded tcgen05 wait expired
}

}  // namespace

// ---------------------------------------------------------------------------------- host side
aslam_status large_ukf_create(aslam_large* h) {
    UkfLarge* u = new (std::nothrow) UkfLarge;
    if (!u) return ASLAM_ERR_ARG;
    h->ukf = u;
    const int n = (int)h->n;
    u->n = n;
    u->na = n + 2;
    u->ns = 2 * n + 5;
    u->ld = h->ldp;
    u->nsld = (((size_t)u->ns + 15) / 16) * 16;
    // ukf.h:73-81, all in float
    u->lambda = (float)(3.0 - (double)(uint32_t)(n + 2));
    const float denom = (u->lambda + (float)(uint32_t)n) + 2.0f;
    const float weight = (float)(0.5 / (double)denom);
    u->wc = (double)weight;
    u->w0 = (double)(u->lambda / denom);
    u->wsq = sqrtf(denom);  // ukf.cpp:284
    const int nb = (n + dense::CHOL_BS - 1) / dense::CHOL_BS;
    aslam_status rc = ASLAM_OK;
    auto A = [&](void** p, size_t bytes) {
        if (rc == ASLAM_OK && !cuda_ok(cudaMalloc(p, bytes), "cudaMalloc (large UKF)")) rc = ASLAM_ERR_CUDA;
    };
    const size_t mat = (size_t)n * u->ld * sizeof(double), sig = (size_t)n * u->nsld * sizeof(double);
    A((void**)&u->Z, n * sizeof(double));
    A((void**)&u->Lb, (size_t)(n + 2) * u->ld * sizeof(double));  // chol(Paug); later W = [Tc; dz_0'; dz'] Lm^-T
    A((void**)&u->S, mat);
    A((void**)&u->Tc, (size_t)(n + 2) * u->ld * sizeof(double));
    A((void**)&u->D, sig);
    A((void**)&u->DZ, sig);
    A((void**)&u->pose, 3 * u->nsld * sizeof(double));
    A((void**)&u->mean, n * sizeof(double));
    A((void**)&u->zpred, n * sizeof(double));
    A((void**)&u->d0, n * sizeof(double));
    A((void**)&u->dz0, n * sizeof(double));
    A((void**)&u->dzv, n * sizeof(double));
    A((void**)&u->gvec, n * sizeof(double));
    A((void**)&u->coef, 2 * sizeof(double));
    A((void**)&u->qdiag, n * sizeof(double));
    A((void**)&u->rdiag, n * sizeof(double));
    A((void**)&u->invP, (size_t)nb * 64 * 64 * sizeof(double));
    A((void**)&u->invS, (size_t)nb * 64 * 64 * sizeof(double));
    const size_t w256 = (size_t)((n + dense::TRSM_WIDE - 1) / dense::TRSM_WIDE) * dense::TRSM_WIDE * dense::TRSM_WIDE * sizeof(double);
    A((void**)&u->inv256, w256);
    A((void**)&u->inv256T, w256);
    A((void**)&u->fail, 2 * sizeof(int));
    A((void**)&u->status, sizeof(uint32_t));
    if (rc != ASLAM_OK) return rc;
    ASLAM_CUDA(cudaEventCreate(&u->evg0));
    ASLAM_CUDA(cudaEventCreate(&u->evg1));
    ASLAM_CUDA(cudaMemsetAsync(u->Z, 0, n * sizeof(double), h->stream));
    ASLAM_CUDA(cudaMemsetAsync(u->inv256, 0, w256, h->stream));   // the blocks above the block diagonal stay zero
    ASLAM_CUDA(cudaMemsetAsync(u->inv256T, 0, w256, h->stream));
    ASLAM_CUDA(cudaMemsetAsync(u->status, 0, sizeof(uint32_t), h->stream));
    ASLAM_CUDA(cudaMemsetAsync(u->fail, 0xff, 2 * sizeof(int), h->stream));
    fill_diag_kernel<<<(n + 255) / 256, 256, 0, h->stream>>>(n, u->qdiag, u->rdiag, (double)h->cfg.ukf_kq,
                                                           (double)h->cfg.ukf_kr);
    ASLAM_CUDA(cudaGetLastError());
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}

void large_ukf_destroy(aslam_large* h) {
    UkfLarge* u = static_cast<UkfLarge*>(h->ukf);
    if (!u) return;
    void* ptrs[] = {u->Z, u->Lb, u->S, u->Tc, u->D, u->DZ, u->pose, u->mean, u->zpred, u->d0, u->dz0, u->dzv, u->gvec, u->coef,
                    u->qdiag, u->rdiag, u->invP, u->invS, u->inv256, u->inv256T, u->fail, u->status, u->Dhi, u->Dlo, u->Zhi, u->Zlo, u->tf32_err};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    if (u->evg0) cudaEventDestroy(u->evg0);
    if (u->evg1) cudaEventDestroy(u->evg1);
    delete u;
    h->ukf = nullptr;
}

// Precision of the three sigma-point products: 0 FP64 (DMMA, the parity path), 1 TF32, 3 error-compensated 3xTF32
// (tcgen05.mma kind::tf32, accumulators in TMEM). The mean, the w0 rank-1 terms, both factorisations and the solves stay
// in FP64 whatever the mode.
aslam_status large_ukf_set_precision(aslam_large* h, int mode) {
    UkfLarge* u = static_cast<UkfLarge*>(h->ukf);
    if (!u || (mode != 0 && mode != 1 && mode != 3)) return ASLAM_ERR_ARG;
    if (mode && !u->Dhi) {
        const size_t bytes = (size_t)u->n * dense::tf32_padded_k(u->ns) * sizeof(float);
        float** bufs[] = {&u->Dhi, &u->Dlo, &u->Zhi, &u->Zlo};
        for (float** b : bufs) ASLAM_CUDA(cudaMalloc((void**)b, bytes));
        ASLAM_CUDA(cudaMalloc((void**)&u->tf32_err, sizeof(int)));
        ASLAM_CUDA(cudaMemset(u->tf32_err, 0, sizeof(int)));
    }
    u->precision = mode;
    return ASLAM_OK;
}

// CUDA-event time of the three sigma-point products of the last step (TF32 modes: the hi/lo split included)
aslam_status large_ukf_products_ms(aslam_large* h, float* ms) {
    UkfLarge* u = static_cast<UkfLarge*>(h->ukf);
    if (!u || !ms || !h->timed) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaEventSynchronize(u->evg1));
    ASLAM_CUDA(cudaEventElapsedTime(ms, u->evg0, u->evg1));
    return ASLAM_OK;
}

// Z of aslam_large_load_state; NULL: Z = h(X) (every landmark "seen" where the state says it is)
namespace {
__global__ void z_from_state_kernel(int n, const double* X, double* Z) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    const int L = (n - 3) / 2;
    if (l == 0) {
        Z[0] = X[0];
        Z[1] = X[1];
        Z[2] = X[2];
    }
    if (l >= L) return;
    double rg, br;
    measure_landmark(X[0], X[1], X[2], X[3 + 2 * l], X[4 + 2 * l], rg, br);
    Z[3 + 2 * l] = rg;
    Z[4 + 2 * l] = br;
}
}  // namespace

aslam_status large_ukf_load_z(aslam_large* h, const double* Z) {
    UkfLarge* u = static_cast<UkfLarge*>(h->ukf);
    if (!u) return ASLAM_ERR_ARG;
    if (Z) {
        ASLAM_CUDA(cudaMemcpyAsync(u->Z, Z, u->n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    } else {
        z_from_state_kernel<<<(u->n / 2 + 255) / 256, 256, 0, h->stream>>>(u->n, h->X, u->Z);
        ASLAM_CUDA(cudaGetLastError());
    }
    ASLAM_CUDA(cudaMemsetAsync(u->status, 0, sizeof(uint32_t), h->stream));
    return ASLAM_OK;
}

aslam_status large_ukf_status(aslam_large* h, uint32_t* status, int32_t* llt_fail) {
    UkfLarge* u = static_cast<UkfLarge*>(h->ukf);
    if (!u) return ASLAM_ERR_ARG;
    int f[2] = {-1, -1};
    uint32_t s = 0;
    ASLAM_CUDA(cudaMemcpyAsync(f, u->fail, sizeof(f), cudaMemcpyDeviceToHost, h->stream));
    ASLAM_CUDA(cudaMemcpyAsync(&s, u->status, sizeof(s), cudaMemcpyDeviceToHost, h->stream));
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    if (status) *status = s;
    if (llt_fail) *llt_fail = f[0];  // Paug.llt() of ukf.cpp:280; a failure of chol(S) only shows in the status bit
    return ASLAM_OK;
}

aslam_status large_ukf_step(aslam_large* h, double x, double y, float yaw, double vx, double az, float dt, uint32_t m) {
    UkfLarge* u = static_cast<UkfLarge*>(h->ukf);
    if (!u) return ASLAM_ERR_ARG;
    cudaStream_t st = h->stream;
    const int n = u->n;
    uint64_t& nl = h->launches;
    // ASLAM_UKF_PHASES=1: CUDA-event time of every phase of the step, printed to stderr (diagnostic; synchronises)
    static const bool phases = std::getenv("ASLAM_UKF_PHASES") != nullptr;
    constexpr int MAXPH = 12;
    cudaEvent_t pev[MAXPH];
    const char* pname[MAXPH];
    int nph = 0;
    auto mark = [&](const char* name) {
        if (!phases || nph >= MAXPH) return;
        if (cudaEventCreate(&pev[nph]) != cudaSuccess) return;
        cudaEventRecord(pev[nph], st);
        pname[nph++] = name;
    };
    // ---- updateZ
    UkfFrontArgs fa;
    fa.ox = x; fa.oy = y; fa.yaw = yaw; fa.m = m;
    fa.pi = h->cfg.pi;
    fa.thresh = h->cfg.min_dist_thresh;
    ukf_front_kernel<<<1, 1024, 0, st>>>(fa, (uint32_t)n, h->X, u->Z, h->d_obs, h->ctx);
    ++nl;
    if (h->timed) ASLAM_CUDA(cudaEventRecord(h->ev0, st));
    mark("start");
    // ---- L = Paug.llt().matrixL()
    ASLAM_CUDA(cudaMemsetAsync(u->fail, 0xff, 2 * sizeof(int), st));
    ASLAM_CUDA(cudaMemcpyAsync(u->Lb, h->P, (size_t)n * u->ld * sizeof(double), cudaMemcpyDeviceToDevice, st));
    aslam_status rc = dense::cholesky_lower(st, u->Lb, u->ld, n, u->invP, u->fail, &nl);
    if (rc != ASLAM_OK) return rc;
    mark("chol(Paug)");
    // ---- sigma points, mean, differences
    SigmaArgs a;
    a.n = n; a.na = u->na; a.ns = u->ns; a.ld = u->ld; a.nsld = u->nsld;
    a.vx = (float)vx; a.az = (float)az; a.dt = dt;  // slam(const float&...) ukf.cpp:90,260
    a.pi = h->cfg.pi;
    a.wsq = u->wsq;
    a.la = std::sqrt((double)(h->cfg.ukf_std_a * h->cfg.ukf_std_a));
    a.w0 = u->w0; a.wc = u->wc;
    const int nsb = (int)((u->nsld + 255) / 256);
    sigma_pose_kernel<<<(u->ns + 255) / 256, 256, 0, st>>>(a, h->X, u->Lb, u->pose);
    sigma_mean_kernel<<<n, 256, 0, st>>>(a, h->X, u->Lb, u->pose, u->mean);
    xdiff_kernel<<<dim3(nsb, n), 256, 0, st>>>(a, h->X, u->Lb, u->pose, u->mean, u->D, u->d0, h->cfg.pi);
    zsig_kernel<<<dim3((u->ns + 255) / 256, (n - 3) / 2 + 1), 256, 0, st>>>(a, h->X, u->Lb, u->pose, u->DZ);
    zdiff_kernel<<<n, 256, 0, st>>>(a, u->DZ, u->Z, u->zpred, u->dz0, u->dzv);
    ASLAM_CUDA(cudaGetLastError());
    nl += 5;
    mark("sigma points");
    // ---- P = sum w d d' + Q;  S = sum w dz dz' + R;  Tc = sum w d dz'
    dense::Epilogue e;
    e.alpha = u->wc;
    e.r1 = u->w0 - u->wc;
    e.lower_only = true;
    e.u = u->d0; e.v = u->d0; e.diag = u->qdiag;
    const int tf = u->precision;
    if (h->timed && u->evg0) ASLAM_CUDA(cudaEventRecord(u->evg0, st));
    if (tf) {  // hi / lo FP32 copies of the centred sigma-point matrices (the only values that are rounded)
        rc = dense::split_tf32(st, u->D, u->nsld, n, u->ns, u->Dhi, tf == 3 ? u->Dlo : nullptr);
        if (rc != ASLAM_OK) return rc;
        rc = dense::split_tf32(st, u->DZ, u->nsld, n, u->ns, u->Zhi, tf == 3 ? u->Zlo : nullptr);
        if (rc != ASLAM_OK) return rc;
        nl += 2;
        mark("tf32 split");
    }
    auto gemm = [&](const double* A, const float* Ahi, const float* Alo, const double* B, const float* Bhi,
                    const float* Blo, double* C) -> aslam_status {
        if (tf) return dense::tf32_gemm_nt(st, n, n, u->ns, Ahi, Alo, Bhi, Blo, C, u->ld, e, tf, u->tf32_err, &nl);
        return dense::dgemm_nt(st, n, n, u->ns, A, u->nsld, B, u->nsld, C, u->ld, e, &nl);
    };
    // M = wc DZ DZ' + R first: the positive definite part of S (its rank-1 term c dz_0 dz_0' is applied analytically
    // below). Its factorisation is a latency chain (one-CTA panels, narrow products: 1.6 ms on an otherwise idle GPU),
    // the P and Tc products are throughput work (1.7 ms) that does not depend on it: in the FP64 mode they run as
    // PERSISTENT grids on all SMs but `chol_sms`, and the factorisation runs beside them on a side stream with the SMs
    // that are left (a product CTA needs 131 KB of shared memory, so nothing else fits on an SM that holds one).
    dense::Epilogue em = e;
    em.u = nullptr; em.v = nullptr; em.r1 = 0.0; em.diag = u->rdiag;
    {
        dense::Epilogue keep = e;
        e = em;
        rc = gemm(u->DZ, u->Zhi, u->Zlo, u->DZ, u->Zhi, u->Zlo, u->S);
        e = keep;
    }
    if (rc != ASLAM_OK) return rc;
    mark("M gemm");
    static cudaEvent_t dbg0 = nullptr, dbg1 = nullptr;  // ASLAM_UKF_PHASES: the factorisation on the side stream
    static const int chol_sms = std::getenv("ASLAM_UKF_CHOL_SMS") ? std::atoi(std::getenv("ASLAM_UKF_CHOL_SMS")) : 16;
    dense::SideStream* cs = nullptr;
    const bool beside = tf == 0 && chol_sms > 0 && n >= 1024;
    if (beside) {
        rc = dense::side_stream(2, &cs);
        if (rc != ASLAM_OK) return rc;
        ASLAM_CUDA(cudaEventRecord(cs->fork, st));
        ASLAM_CUDA(cudaStreamWaitEvent(cs->s, cs->fork, 0));
        if (phases && !dbg0) { cudaEventCreate(&dbg0); cudaEventCreate(&dbg1); }
        if (phases) cudaEventRecord(dbg0, cs->s);
        rc = dense::cholesky_lower(cs->s, u->S, u->ld, n, u->invS, u->fail + 1, &nl);
        if (rc != ASLAM_OK) return rc;
        if (phases) cudaEventRecord(dbg1, cs->s);
        ASLAM_CUDA(cudaEventRecord(cs->join, cs->s));
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
        e.max_ctas = sms - chol_sms;
    }
    rc = gemm(u->D, u->Dhi, u->Dlo, u->D, u->Dhi, u->Dlo, h->P);
    if (rc != ASLAM_OK) return rc;
    mark("P gemm");
    e.lower_only = false;
    e.r1 = u->w0 - u->wc;
    e.u = u->d0; e.v = u->dz0; e.diag = nullptr;
    rc = gemm(u->D, u->Dhi, u->Dlo, u->DZ, u->Zhi, u->Zlo, u->Tc);
    if (rc != ASLAM_OK) return rc;
    if (h->timed && u->evg1) ASLAM_CUDA(cudaEventRecord(u->evg1, st));
    mark("Tc gemm");
    // ---- M = Lm Lm';  [W; a'; b'] = [Tc; dz_0'; dz'] Lm^-T  (a = Lm^-1 dz_0, b = Lm^-1 dz ride along as two extra rows)
    if (beside) {
        ASLAM_CUDA(cudaStreamWaitEvent(st, cs->join, 0));
    } else {
        rc = dense::cholesky_lower(st, u->S, u->ld, n, u->invS, u->fail + 1, &nl);
        if (rc != ASLAM_OK) return rc;
    }
    mark("chol(M)");
    double* arow = u->Tc + (size_t)n * u->ld;
    double* brow = arow + u->ld;
    ASLAM_CUDA(cudaMemcpyAsync(arow, u->dz0, n * sizeof(double), cudaMemcpyDeviceToDevice, st));
    ASLAM_CUDA(cudaMemcpyAsync(brow, u->dzv, n * sizeof(double), cudaMemcpyDeviceToDevice, st));
    // 256-wide out-of-place solve: W lands in the storage of chol(Paug) (dead since the sigma points), the scaled
    // panels L[>j, j] inv256_j in the storage of D (dead since the three products)
    double* W = u->Lb;
    rc = dense::trsm_right_lower_t_wide(st, u->Tc, u->ld, n + 2, u->S, u->ld, n, u->invS, u->inv256, u->inv256T, u->D,
                                        u->ld, W, u->ld, &nl);
    if (rc != ASLAM_OK) return rc;
    mark("trsm");
    arow = W + (size_t)n * u->ld;
    brow = arow + u->ld;
    sm_scalars_kernel<<<1, 256, 0, st>>>(n, arow, brow, u->w0 - u->wc, u->coef, u->status);
    state_update_kernel<<<(n + 7) / 8, 256, 0, st>>>(n, u->ld, W, arow, brow, u->coef, u->mean, u->gvec, h->X,
                                                   u->status);
    ASLAM_CUDA(cudaGetLastError());
    nl += 2;
    mark("state update");
    // ---- P = P_pred - W W' + (c/den) g g'
    dense::Epilogue d;
    d.alpha = -1.0;
    d.beta = 1.0;
    d.lower_only = true;
    d.u = u->gvec; d.v = u->gvec; d.r1_dev = u->coef;
    rc = dense::dgemm_nt(st, n, n, n, W, u->ld, W, u->ld, h->P, u->ld, d, &nl);
    if (rc != ASLAM_OK) return rc;
    rc = dense::symmetrize_from_lower(st, h->P, u->ld, n, &nl);
    if (rc != ASLAM_OK) return rc;
    fail_to_status_kernel<<<1, 1, 0, st>>>(u->fail, tf ? u->tf32_err : nullptr, u->status);
    ASLAM_CUDA(cudaGetLastError());
    ++nl;
    if (h->timed) ASLAM_CUDA(cudaEventRecord(h->ev1, st));
    mark("W W' + symmetrize");
    if (nph > 1) {
        cudaEventSynchronize(pev[nph - 1]);
        if (beside && dbg0) {
            float a = 0, b = 0;
            cudaEventElapsedTime(&a, pev[0], dbg0);
            cudaEventElapsedTime(&b, dbg0, dbg1);
            std::fprintf(stderr, "ukf chol(M) beside the products: starts %.3f ms into the step, takes %.3f ms\n", a, b);
        }
        for (int i = 1; i < nph; ++i) {
            float ms = 0.0f;
            cudaEventElapsedTime(&ms, pev[i - 1], pev[i]);
            std::fprintf(stderr, "ukf phase %-20s %8.3f ms\n", pname[i], ms);
        }
    }
    for (int i = 0; i < nph; ++i) cudaEventDestroy(pev[i]);
    return ASLAM_OK;
}

}  // namespace aslam
