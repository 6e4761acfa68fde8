// K2: batched UKF-SLAM step. Replaces, for every filter of a batch, updateZ +
// slam of the reference (src/ukf/ukf.cpp:113-180, 260-392), the weights of src/ukf/ukf.h:73-81 and the
// models of include/awesome_slam/common.h:46-90.
//
// Mapping: one CTA (128 threads) per filter; the whole working set of a step -- Paug/L, the
// 2N+5 propagated sigma points, their measurement images, P, S, Tc, K -- lives in shared memory
// and P stays there across the `steps` messages of a launch. The kernel is instantiated for the padded state
// dimensions 30 (N <= 29: the reference's own cap, config.h:45; 66 KB, three filters per SM) and 36 (N <= 35: the
// cfg3 shape of 16 landmarks, 75 sigma points; 92 KB, two per SM).
// Barriers: the factorisations are the latency chains of the step. LLT takes ONE barrier per column (the diagonal
// goes to a side array, so nobody waits for thread 0 to patch it in; round 1 took 3), and the pivot search of
// PartialPivLU is a shuffle reduction of warp 0 instead of a serial scan by one thread. (An out-of-place LU pass with
// one barrier per pivot was measured too: it touches N^2 instead of (N-k)^2 entries per pivot and was 15 % slower.)
//
// Numerics: the reference UKF is numerically fragile (w0 = (1-N)/3 < 0: P - K S K' goes indefinite
// and Eigen's LLT silently returns a half-factored L), so every reduction here
// keeps the reference's order -- each output element is one thread's sequential sum over the
// sigma points / inner index, multiply and add rounded separately (the library is built with
// -fmad=false) -- and the LLT, LU and triangular solves follow Eigen's unblocked algorithms step
// for step. Parallelism is across output elements only. What can still differ from the CPU are
// the last bits of sin/cos/atan2.
#include <cuda_runtime.h>

#include "frontend.cuh"
#include "internal.hpp"

namespace aslam {

namespace {

constexpr int UTHREADS = 128;

// e / d for 0 <= e < 2^15 and 1 <= d <= 80 without the integer-division sequence (~20 instructions): the float
// quotient of two small integers is within 1 ulp of the true one, so truncating (e + 0.5) / d is exact.
__device__ __forceinline__ int fdiv_small(int e, float rcp_d) { return (int)(((float)e + 0.5f) * rcp_d); }

// The front end's per-message arrays (FrontSlot's, frontend.cuh): dead once the association / wait-list phase is over,
// so they share the bytes of the sigma-point matrices.
struct UkfFrontArrays {
    float2 opt[ASLAM_MAX_OBS];
    float orng[ASLAM_MAX_OBS];
    float obrg[ASLAM_MAX_OBS];
    float omind[ASLAM_MAX_OBS];
    int32_t ocorr[ASLAM_MAX_OBS];
    float2 wpt[ASLAM_MAX_WAIT];
};
// FrontSlot's member names, the arrays as pointers into UkfFrontArrays (owner / ohit belong to the TRACK mode of phase
// A, which this kernel does not use: null).
struct UkfFront {
    float2* opt;
    float* orng;
    float* obrg;
    float* omind;
    int32_t* ocorr;
    float2* wpt;
    int32_t* owner;
    uint8_t* ohit;
    uint32_t nobs, obs_base, wcount, N, status, flags, grew_from, unassoc;
    int32_t promote;
    double ox, oy, vx, az;
    float yaw, dt;
};

template <int UNP>
struct __align__(16) UkfSlot {
    static constexpr int UNA = UNP + 2;      // augmented
    static constexpr int UNS = 2 * UNP + 5;  // sigma points
    static constexpr int USP = UNS | 1;      // row pitch of the sigma-point matrices (odd: conflict-free row walks)
    // Three generations share these bytes, a barrier between each: the front end's arrays (association, wait-list);
    // the sigma-point matrices; and, once P, S and Tc have been formed, the LU factors of S, its inverse and K.
    union {
        UkfFrontArrays fa;
        struct {
            double XS[UNP][USP];   // XsigPred, later Xdiff       (row = state index, col = sigma point)
            double ZS[UNP][USP];   // Zsig, later Zdiff
        };
        struct {
            double K[UNP][UNP];
            double LU[UNP][UNP];   // PartialPivLU of S, then S^-1 in Si
            double Si[UNP][UNP];
        };
    };
    // Paug -> L (ukf.cpp:263-280), dead after the sigma points; then Tc (written by the covariance sums, dead once K is
    // formed); then K S. UNA^2 >= UNP^2 doubles.
    double L[UNA][UNA];
    double P[UNP][UNP];
    double S[UNP][UNP];
    double X[UNP], Z[UNP], Zp[UNP], Zd[UNP], Ld[UNA];
    int piv[UNP];
    int fail_k;
    UkfFront fr;
    __device__ __forceinline__ double (*Tc())[UNP] { return reinterpret_cast<double(*)[UNP]>(&L[0][0]); }
};

struct UkfParams {
    BatchState st;
    BatchInput in;
    FrontParams fp;
    double kq, kr, kp_lm;
    float std_a, std_yaw;
};

template <int UNP>
__global__ void __launch_bounds__(UTHREADS) ukf_batched_kernel(const UkfParams kp) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    UkfSlot<UNP>& s = *reinterpret_cast<UkfSlot<UNP>*>(smem_raw);
    const int tid = threadIdx.x;
    const uint32_t f = blockIdx.x;
    const float pi = kp.fp.pi;

    // ---- load state
    for (int k = tid; k < UNP * UNP; k += UTHREADS) (&s.P[0][0])[k] = kp.st.P[(size_t)f * UNP * UNP + k];
    for (int k = tid; k < UNP; k += UTHREADS) {
        s.X[k] = kp.st.X[(size_t)f * UNP + k];
        s.Z[k] = kp.st.Z[(size_t)f * UNP + k];
    }
    if (tid == 0) {
        s.fr.N = kp.st.N[f];
        s.fr.status = kp.st.status[f];
        s.fr.flags = kp.st.flags[f];
        s.fr.wcount = kp.st.wcount[f];
        s.fr.promote = 0;
        s.fr.opt = s.fa.opt;
        s.fr.orng = s.fa.orng;
        s.fr.obrg = s.fa.obrg;
        s.fr.omind = s.fa.omind;
        s.fr.ocorr = s.fa.ocorr;
        s.fr.wpt = s.fa.wpt;
        s.fr.owner = nullptr;
        s.fr.ohit = nullptr;
        s.fr.unassoc = 0;
        s.fail_k = kp.st.llt_fail[f];
    }
    __syncthreads();

    for (uint32_t step = 0; step < kp.in.steps; ++step) {
        const size_t si = (size_t)step * kp.st.batch + f;
        if (s.fr.flags & FLAG_INIT_Z) break;  // cbOdom gate (ukf.cpp:72); uniform for the CTA
        // ---- updateZ, ukf.cpp:115-118
        if (tid == 0) {
            s.fr.ox = kp.in.ox[si];
            s.fr.oy = kp.in.oy[si];
            s.fr.yaw = kp.in.yaw[si];
            s.fr.vx = kp.in.vx[si];
            s.fr.az = kp.in.az[si];
            s.fr.dt = kp.in.dt[si];
            s.Z[0] = s.fr.ox;
            s.Z[1] = s.fr.oy;
            s.Z[2] = (double)s.fr.yaw;
            uint32_t base, cnt;
            if (kp.in.offsets) {
                base = kp.in.offsets[f];
                cnt = kp.in.offsets[f + 1] - base;
            } else {
                base = (uint32_t)(si * kp.in.obs_per_filter);
                cnt = kp.in.obs_per_filter;
            }
            if (cnt > ASLAM_MAX_OBS) {
                cnt = ASLAM_MAX_OBS;
                s.fr.status |= ASLAM_ST_OBS_OVERFLOW;
            }
            s.fr.obs_base = base;
            s.fr.nobs = cnt;
        }
        __syncthreads();
        front_phase_a(s.fr, kp.fp, s.X, s.Z, kp.in.range, kp.in.bearing, s.fr.obs_base, f, tid, UTHREADS);
        __syncthreads();
        if (tid == 0) {
            int32_t* assoc = nullptr;
            if (kp.in.assoc && step + 1 == kp.in.steps)
                assoc = kp.in.assoc + (kp.in.offsets ? s.fr.obs_base : f * kp.in.obs_per_filter);
            front_phase_b(s.fr, kp.fp, s.X, s.Z, f, assoc);
            if (s.fr.flags & FLAG_INIT_X) {  // ukf.cpp:83-87
                s.fr.flags &= ~FLAG_INIT_X;
                for (uint32_t j = 0; j < s.fr.N; ++j) s.X[j] = s.Z[j];
            }
        }
        __syncthreads();
        const int N = (int)s.fr.N, NA = N + 2, NS = 2 * N + 5;
        // ---- growth of P (ukf.cpp:241) and the weights (ukf.h:73-81, recomputed: they depend on N only)
        if (s.fr.grew_from) {
            const int g0 = (int)s.fr.grew_from;
            for (int k = tid; k < N * N; k += UTHREADS) {
                const int i = k / N, j = k % N;
                if (i >= g0 || j >= g0) s.P[i][j] = (i == j) ? kp.kp_lm : 0.0;
            }
            __syncthreads();  // (uniform branch) the Paug loop below reads P with a different thread mapping
        }
        const float lambda = (float)(3.0 - (double)(N + 2));
        const float denom = __fadd_rn(__fadd_rn(lambda, (float)N), 2.0f);  // lambda + N + 2 in float
        // weights(0) = lambda / (lambda + N + 2) and the common weight 0.5 / (..), both rounded to float (ukf.h:75-80)
        const double w0 = (double)__fdiv_rn(lambda, denom), wc = (double)(float)(0.5 / (double)denom);
        // ---- Paug (ukf.cpp:262-277) into L
        const float rNA = 1.0f / (float)NA;
        for (int k = tid; k < NA * NA; k += UTHREADS) {
            const int i = fdiv_small(k, rNA), j = k - i * NA;
            double v = 0.0;
            if (i < N && j < N) v = s.P[i][j];
            else if (i == j) v = (double)(i == N ? __fmul_rn(kp.std_a, kp.std_a) : __fmul_rn(kp.std_yaw, kp.std_yaw));
            s.L[i][j] = v;
        }
        if (tid == 0) s.fail_k = -1;
        __syncthreads();
        // ---- L = Paug.llt().matrixL(): Eigen's unblocked left-looking LLT with its failure rule. One barrier per
        // column: every thread forms the pivot x itself (same operands, same order: the same value), the new diagonal
        // goes to Ld (L[k][k] still holds the Paug entry that the other threads may not have read yet) and is
        // patched in with the upper triangle below.
        for (int k = 0; k < NA; ++k) {
            if (s.fail_k >= 0) break;
            // The pivot sum (row k against itself) and this thread's row sum (row i against row k) are independent
            // sequential sums over the same row k: they run in ONE loop, sharing the loads of row k, products formed
            // eight at a time before the dependent adds (the shared-memory latency otherwise sits inside every term
            // of the chain). Same operations in the same order per sum as Eigen's unblocked LLT.
            const int i = k + 1 + tid;  // (NA - k - 1 <= 36 < UTHREADS rows below the pivot: at most one per thread)
            const bool row = i < NA;
            const double* lk = s.L[k];
            const double* li = s.L[row ? i : k];
            double sumx = 0.0, sumi = 0.0;
            int j = 0;
            for (; j + 8 <= k; j += 8) {
                double a[8], b[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    a[u] = lk[j + u];
                    b[u] = li[j + u];
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    sumx += a[u] * a[u];
                    sumi += b[u] * a[u];
                }
            }
            for (; j < k; ++j) {
                const double a = lk[j];
                sumx += a * a;
                sumi += li[j] * a;
            }
            double x = s.L[k][k];
            if (k > 0) x -= sumx;
            if (x <= 0.0) {  // every thread computes the same x: uniform decision
                __syncthreads();
                if (tid == 0) s.fail_k = k;
                __syncthreads();
                break;
            }
            x = sqrt(x);
            if (row) s.L[i][k] = (s.L[i][k] - sumi) / x;
            if (tid == 0) s.Ld[k] = x;
            __syncthreads();
        }
        {
            const int done = s.fail_k >= 0 ? s.fail_k : NA;  // columns that were factorised
            for (int k = tid; k < NA * NA; k += UTHREADS) {
                const int i = fdiv_small(k, rNA), j = k - i * NA;
                if (j > i) s.L[i][j] = 0.0;
                else if (i == j && i < done) s.L[i][i] = s.Ld[i];
            }
        }
        __syncthreads();
        // ---- sigma points (ukf.cpp:283-289) propagated through f (ukf.cpp:292-297)
        {
            const float w = __fsqrt_rn(__fadd_rn(__fadd_rn(lambda, (float)N), 2.0f));
            const float vx = (float)s.fr.vx, az = (float)s.fr.az, dt = s.fr.dt;
            for (int c = tid; c < NS; c += UTHREADS) {
                // column c: 0 -> Xaug; 1..NA -> + w L.col(c-1); NA+1..2NA -> - w L.col(c-1-NA)
                const int lc = c == 0 ? -1 : (c <= NA ? c - 1 : c - 1 - NA);
                const double sg = c <= NA ? (double)w : -(double)w;
                auto aug = [&](int k) -> double {
                    const double xa = k < N ? s.X[k] : 0.0;
                    if (lc < 0) return xa;
                    // Xaug + w*L  and  Xaug - w*L (the product is rounded first in both)
                    return c <= NA ? xa + (double)w * s.L[k][lc] : xa - (double)w * s.L[k][lc];
                };
                (void)sg;
                double x = aug(0), y = aug(1), th = aug(2);
                const double na = aug(N);
                motion_model_aug(x, y, th, na, vx, az, dt);
                s.XS[0][c] = x;
                s.XS[1][c] = y;
                s.XS[2][c] = (double)normalize_angle((float)th, pi);
                for (int k = 3; k < N; ++k) s.XS[k][c] = aug(k);
            }
        }
        __syncthreads();
        // ---- predicted mean (ukf.cpp:300-304) and Zsig (ukf.cpp:322-326)
        for (int k = tid; k < N; k += UTHREADS) {
            double acc = 0.0 + w0 * s.XS[k][0];
#pragma unroll 8
            for (int i = 1; i < NS; ++i) acc += wc * s.XS[k][i];
            s.X[k] = acc;
        }
        {
            const int nl = (N - 3) / 2;
            const float rNSz = 1.0f / (float)NS;
            for (int e = tid; e < NS * (nl + 1); e += UTHREADS) {
                const int l = fdiv_small(e, rNSz), c = e - l * NS;
                if (l == nl) {
                    s.ZS[0][c] = s.XS[0][c];
                    s.ZS[1][c] = s.XS[1][c];
                    s.ZS[2][c] = s.XS[2][c];
                } else {
                    double rg, br;
                    measure_landmark(s.XS[0][c], s.XS[1][c], s.XS[2][c], s.XS[3 + 2 * l][c], s.XS[4 + 2 * l][c], rg, br);
                    s.ZS[3 + 2 * l][c] = rg;
                    s.ZS[4 + 2 * l][c] = br;
                }
            }
        }
        __syncthreads();
        // ---- Zpred (ukf.cpp:329-339); Xdiff in place (ukf.cpp:312-313)
        for (int k = tid; k < N; k += UTHREADS) {
            double acc = 0.0 + w0 * s.ZS[k][0];
#pragma unroll 8
            for (int i = 1; i < NS; ++i) acc += wc * s.ZS[k][i];
            if (k >= 2 && (k & 1) == 0) acc = (double)normalize_angle((float)acc, pi);
            s.Zp[k] = acc;
        }
        const float rNS = 1.0f / (float)NS;
        for (int e = tid; e < N * NS; e += UTHREADS) {
            const int k = fdiv_small(e, rNS), c = e - k * NS;
            double d = s.XS[k][c] - s.X[k];
            if (k == 2) d = (double)normalize_angle((float)d, pi);
            s.XS[k][c] = d;
        }
        __syncthreads();
        // ---- Zdiff in place (ukf.cpp:347-351)
        for (int e = tid; e < N * NS; e += UTHREADS) {
            const int k = fdiv_small(e, rNS), c = e - k * NS;
            double d = s.ZS[k][c] - s.Zp[k];
            if (k >= 2 && (k & 1) == 0) d = (double)normalize_angle((float)d, pi);
            s.ZS[k][c] = d;
        }
        __syncthreads();
        // ---- P (ukf.cpp:307-319), S (ukf.cpp:342-357), Tc (ukf.cpp:360-375): one thread per entry, FOUR entries in flight
        // per thread -- each entry is still one sequential sum over the sigma points in the reference's order (the
        // chains are independent, so interleaving them changes no rounding), but the 8-cycle dependent add of one
        // chain now hides behind the other three.
        {
            const float rNN = 1.0f / (float)(N * N), rN = 1.0f / (float)N;
            auto entry = [&](int e, int& which, int& a, int& b, const double*& pa, const double*& pb) {
                which = fdiv_small(e, rNN);
                const int r = e - which * N * N;
                a = fdiv_small(r, rN);
                b = r - a * N;
                pa = which == 1 ? s.ZS[a] : s.XS[a];
                pb = which == 0 ? s.XS[b] : s.ZS[b];
            };
            auto put = [&](int which, int a, int b, double acc) {
                if (which == 0) s.P[a][b] = acc + ((a == b && a < 3) ? kp.kq : 0.0);
                else if (which == 1) s.S[a][b] = acc + (a == b ? kp.kr : 0.0);
                else s.Tc()[a][b] = acc;
            };
            const int total = 3 * N * N;
            int e0 = tid;
            for (; e0 + 3 * UTHREADS < total; e0 += 4 * UTHREADS) {
                const double* da[4];
                const double* db[4];
                double acc[4];
                int wh[4], ea[4], eb[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    entry(e0 + q * UTHREADS, wh[q], ea[q], eb[q], da[q], db[q]);
                    acc[q] = 0.0 + (w0 * da[q][0]) * db[q][0];
                }
                for (int i = 1; i < NS; ++i) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) acc[q] = acc[q] + (wc * da[q][i]) * db[q][i];
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) put(wh[q], ea[q], eb[q], acc[q]);
            }
            for (; e0 < total; e0 += UTHREADS) {
                const double *da, *db;
                int which, a, b;
                entry(e0, which, a, b, da, db);
                double acc = 0.0 + (w0 * da[0]) * db[0];
#pragma unroll 4
                for (int i = 1; i < NS; ++i) acc = acc + (wc * da[i]) * db[i];
                put(which, a, b, acc);
            }
        }
        __syncthreads();  // last read of the sigma-point matrices: LU / Si / K take over their storage
        const float rNl = 1.0f / (float)N;
        for (int e = tid; e < N * N; e += UTHREADS) {
            const int i = fdiv_small(e, rNl), j = e - i * N;
            s.LU[i][j] = s.S[i][j];
        }
        __syncthreads();
        // ---- S.inverse(): PartialPivLU (unblocked, first largest pivot). Per pivot: warp 0 finds the pivot row with a
        // shuffle reduction (largest |a(i,k)|, i >= k, ties to the lowest row: Eigen's first maximum); the row swap
        // (and its barrier) only happens when the pivot is not on the diagonal.
        for (int k = 0; k < N; ++k) {
            if (tid < 32) {
                double best = -1.0;
                int p = k;
                for (int i = k + tid; i < N; i += 32) {  // (N <= 35: at most two rows per lane, ascending)
                    const double v = fabs(s.LU[i][k]);
                    if (v > best) {
                        best = v;
                        p = i;
                    }
                }
#pragma unroll
                for (int off = 16; off; off >>= 1) {
                    const double ob = __shfl_xor_sync(0xffffffffu, best, off);
                    const int op = __shfl_xor_sync(0xffffffffu, p, off);
                    if (ob > best || (ob == best && op < p)) {
                        best = ob;
                        p = op;
                    }
                }
                if (tid == 0) s.piv[k] = best != 0.0 ? p : -1 - p;  // zero pivot: recorded, no swap/scale
            }
            __syncthreads();
            const int pr = s.piv[k];
            const bool nz = pr >= 0;
            const int p = nz ? pr : -1 - pr;
            if (nz && p != k) {
                for (int j = tid; j < N; j += UTHREADS) {
                    const double t = s.LU[k][j];
                    s.LU[k][j] = s.LU[p][j];
                    s.LU[p][j] = t;
                }
                __syncthreads();  // (uniform branch)
            }
            if (nz) {
                const double pv = s.LU[k][k];
                for (int i = k + 1 + tid; i < N; i += UTHREADS) s.LU[i][k] /= pv;
                __syncthreads();
            }
            const int m = N - k - 1;
            const float rm = 1.0f / (float)(m > 0 ? m : 1);
            for (int e = tid; e < m * m; e += UTHREADS) {
                const int q = fdiv_small(e, rm);
                const int i = k + 1 + q, j = k + 1 + (e - q * m);
                s.LU[i][j] -= s.LU[i][k] * s.LU[k][j];
            }
            __syncthreads();
        }
        double* F = &s.LU[0][0];  // the LU factors
        double* I = &s.Si[0][0];  // S^-1 is built here
        // ... then solve against the permuted identity, one column per thread
        for (int e = tid; e < N * N; e += UTHREADS) {
            const int i = fdiv_small(e, rNl), j = e - i * N;
            I[i * UNP + j] = (i == j) ? 1.0 : 0.0;
        }
        __syncthreads();
        for (int j = tid; j < N; j += UTHREADS) {
            for (int k = 0; k < N; ++k) {
                const int pr = s.piv[k];
                const int p = pr >= 0 ? pr : -1 - pr;
                if (p != k) {
                    const double t = I[k * UNP + j];
                    I[k * UNP + j] = I[p * UNP + j];
                    I[p * UNP + j] = t;
                }
            }
            for (int i = 0; i < N; ++i) {
                double v = I[i * UNP + j];
                for (int k = 0; k < i; ++k) v -= F[i * UNP + k] * I[k * UNP + j];
                I[i * UNP + j] = v;
            }
            for (int i = N - 1; i >= 0; --i) {
                double v = I[i * UNP + j];
                for (int k = i + 1; k < N; ++k) v -= F[i * UNP + k] * I[k * UNP + j];
                I[i * UNP + j] = v / F[i * UNP + i];
            }
        }
        __syncthreads();
        // ---- K = Tc S^-1 (ukf.cpp:378); residual (ukf.cpp:381-386). The three N x N products below keep one
        // sequential k-ascending sum per entry; a thread runs its (up to three) entries side by side.
        const float rN1 = 1.0f / (float)N;
        auto product3 = [&](auto&& lhs, auto&& rhs, auto&& store) {
            for (int e0 = tid; e0 < N * N; e0 += 3 * UTHREADS) {
                int ii[3], jj[3];
                bool in[3];
                double acc[3];
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    const int e = e0 + q * UTHREADS;
                    in[q] = e < N * N;
                    const int ee = in[q] ? e : tid;
                    ii[q] = fdiv_small(ee, rN1);
                    jj[q] = ee - ii[q] * N;
                    acc[q] = 0.0;
                }
                for (int k = 0; k < N; ++k) {
#pragma unroll
                    for (int q = 0; q < 3; ++q) acc[q] += lhs(ii[q], k) * rhs(k, jj[q]);
                }
#pragma unroll
                for (int q = 0; q < 3; ++q)
                    if (in[q]) store(ii[q], jj[q], acc[q]);
            }
        };
        product3([&](int i, int k) { return s.Tc()[i][k]; }, [&](int k, int j) { return I[k * UNP + j]; },
                 [&](int i, int j, double v) { s.K[i][j] = v; });
        for (int k = tid; k < N; k += UTHREADS) {
            double d = s.Z[k] - s.Zp[k];
            if (k >= 2 && (k & 1) == 0) d = (double)normalize_angle((float)d, pi);
            s.Zd[k] = d;
        }
        __syncthreads();
        // ---- X += K Zdiff (ukf.cpp:389); K S into L's storage
        for (int i = tid; i < N; i += UTHREADS) {
            double acc = 0.0;
            for (int k = 0; k < N; ++k) acc += s.K[i][k] * s.Zd[k];
            s.X[i] = s.X[i] + acc;
        }
        double* KS = &s.L[0][0];  // N x N, pitch UNP
        product3([&](int i, int k) { return s.K[i][k]; }, [&](int k, int j) { return s.S[k][j]; },
                 [&](int i, int j, double v) { KS[i * UNP + j] = v; });
        __syncthreads();
        // ---- P -= (K S) K' (ukf.cpp:391)
        product3([&](int i, int k) { return KS[i * UNP + k]; }, [&](int k, int j) { return s.K[j][k]; },
                 [&](int i, int j, double v) { s.P[i][j] = s.P[i][j] - v; });
        if (tid == 0 && s.fail_k >= 0) s.fr.status |= ASLAM_ST_LLT_FAILED;
        __syncthreads();
    }

    // ---- write back
    bool bad = false;
    for (int k = tid; k < UNP * UNP; k += UTHREADS) {
        const double v = (&s.P[0][0])[k];
        kp.st.P[(size_t)f * UNP * UNP + k] = v;
        bad |= !isfinite(v);
    }
    for (int k = tid; k < UNP; k += UTHREADS) {
        kp.st.X[(size_t)f * UNP + k] = s.X[k];
        kp.st.Z[(size_t)f * UNP + k] = s.Z[k];
        bad |= !isfinite(s.X[k]);
    }
    if (bad) atomicOr(&kp.st.status[f], (uint32_t)ASLAM_ST_NONFINITE);
    if (tid == 0) {
        kp.st.N[f] = s.fr.N;
        if (s.fr.status) atomicOr(&kp.st.status[f], s.fr.status);
        kp.st.flags[f] = s.fr.flags;
        kp.st.wcount[f] = s.fr.wcount;
        kp.st.llt_fail[f] = s.fail_k;
    }
}

}  // namespace

template <int UNP>
static aslam_status launch_ukf_t(const UkfParams& kp, cudaStream_t stream) {
    // Resident CTAs per SM are what sets this kernel's throughput (measured by padding the allocation: 3 per SM
    // 5.86 M updates/s at n = 19, 2 per SM 4.14 M, 1 per SM 2.18 M -- every phase is a latency chain), and the
    // footprint is set by the cap the handle was created with, not by the filters' current N.
    const size_t smem = sizeof(UkfSlot<UNP>);
    ASLAM_DYN_SMEM(ukf_batched_kernel<UNP>, smem);
    ukf_batched_kernel<UNP><<<kp.st.batch, UTHREADS, smem, stream>>>(kp);
    ASLAM_CUDA(cudaGetLastError());
    return ASLAM_OK;
}

aslam_status launch_ukf_batched(const aslam_config& cfg, const BatchState& st, const BatchInput& in,
                                cudaStream_t stream) {
    UkfParams kp;
    kp.st = st;
    kp.in = in;
    kp.fp.pi = cfg.pi;
    kp.fp.thresh = cfg.min_dist_thresh;
    kp.fp.occ = cfg.min_landmark_occ;
    kp.fp.cap = cfg.max_state_dim;
    kp.fp.wrange = st.wrange;
    kp.fp.wbearing = st.wbearing;
    kp.fp.whits = st.whits;
    kp.kq = (double)cfg.ukf_kq;
    kp.kr = (double)cfg.ukf_kr;
    kp.kp_lm = (double)cfg.ukf_kp_landmark_pose;
    kp.std_a = cfg.ukf_std_a;
    kp.std_yaw = cfg.ukf_std_yaw;
    switch (st.ld) {
        case 20: return launch_ukf_t<20>(kp, stream);
        case 24: return launch_ukf_t<24>(kp, stream);
        case 30: return launch_ukf_t<30>(kp, stream);
        case 36: return launch_ukf_t<36>(kp, stream);
        default:
            set_error("ukf_batched: built for max_state_dim <= 36 (padded dimensions 30 and 36); larger maps: aslam_large_ukf_*");
            return ASLAM_ERR_UNSUPPORTED;
    }
}

}  // namespace aslam
