// Blocked FP64 Cholesky and triangular solves of the large-map UKF (dense_fp64.hpp): Paug.llt() (ukf.cpp:280) and the
// S^-1 products of ukf.cpp:378-391. Right-looking with 64-wide panels: the diagonal block is factored and inverted by
// one CTA in shared memory; the panel below it and the trailing matrix are DGEMMs on the DMMA pipe (dgemm_tma.cu):
//     L21 = A21 inv(L11)'          A22 -= L21 L21'   (lower tiles only)
// Triangular solves reuse the stored inverses of the diagonal blocks, so they are GEMMs as well.
#include <cuda_runtime.h>

#include "dense_fp64.hpp"

namespace aslam {
namespace dense {

namespace {

constexpr int BS = CHOL_BS;

// One CTA, 256 threads as a 16 x 16 grid; thread (ti, tj) keeps the 4 x 4 sub-blocks A[4ti.., 4tj..] and X[4ti.., 4tj..]
// in registers. A: bs x bs diagonal block (row-major, ld), padded to 64 with the identity. Right-looking Cholesky fused
// with the elimination of [L | I], so that X ends as inv(L): per column k the owners publish column k of A and row k
// of X through (double-buffered) shared memory, one __syncthreads, then every thread applies both rank-1 updates to
// its registers. Writes L over the lower triangle of A and inv(L) (pitch 64, zero above the diagonal and beyond bs) to
// `inv`. k0: global index of the block's first row (for the failure report: Eigen's LLT stops at the first x <= 0).
__global__ void __launch_bounds__(256) chol_panel_kernel(double* A, size_t ld, int bs, double* inv, int* fail, int k0) {
    // One barrier per FOUR columns: the owners publish block column kb of A (64 x 4) and block row kb of X (4 x 64);
    // every working thread then factors the 4 x 4 pivot block itself (D = Ld Ld', M = Ld^-1), forms its own rows of
    // the panel L_i = A_i Ld^-T, the panel rows of its columns L_j, and the eliminated pivot rows of X (M X_kb), and
    // applies both rank-4 updates to its registers. 16 barriers and 16 short rsqrt chains instead of 64 of each.
    __shared__ __align__(16) double cb[2][BS][4];
    __shared__ __align__(16) double xr[2][4][BS];
    const int tid = threadIdx.x, ti = tid >> 4, tj = tid & 15;
    double a[4][4], x[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int i = 4 * ti + r, j = 4 * tj + c;
            a[r][c] = (i < bs && j <= i) ? A[(size_t)i * ld + j] : (i == j ? 1.0 : 0.0);
            x[r][c] = i == j ? 1.0 : 0.0;
        }
    for (int kb = 0; kb < BS / 4; ++kb) {
        const int buf = kb & 1;
        if (tj == kb) {
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                *reinterpret_cast<double2*>(&cb[buf][4 * ti + r][0]) = make_double2(a[r][0], a[r][1]);
                *reinterpret_cast<double2*>(&cb[buf][4 * ti + r][2]) = make_double2(a[r][2], a[r][3]);
            }
        }
        if (ti == kb) {
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                *reinterpret_cast<double2*>(&xr[buf][r][4 * tj]) = make_double2(x[r][0], x[r][1]);
                *reinterpret_cast<double2*>(&xr[buf][r][4 * tj + 2]) = make_double2(x[r][2], x[r][3]);
            }
        }
        __syncthreads();
        if (!(tj <= ti && ti >= kb)) continue;  // upper block triangle and finished rows have nothing to do
        // ---- pivot block: lower triangle of D, Cholesky with Eigen's failure rule reported per column
        double d[4][4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const double2 v01 = *reinterpret_cast<const double2*>(&cb[buf][4 * kb + r][0]);
            const double2 v23 = *reinterpret_cast<const double2*>(&cb[buf][4 * kb + r][2]);
            d[r][0] = v01.x;
            d[r][1] = v01.y;
            d[r][2] = v23.x;
            d[r][3] = v23.y;
        }
        double m[4][4];  // M = Ld^-1 (lower); iv[k] = 1 / Ld[k][k]
        double l[4][4];
        double iv[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            double piv = d[k][k];
#pragma unroll
            for (int q = 0; q < k; ++q) piv = fma(-l[k][q], l[k][q], piv);
            if (!(piv > 0.0)) {
                if (ti == kb && tj == kb && 4 * kb + k < bs) atomicCAS(fail, -1, k0 + 4 * kb + k);
                piv = piv == 0.0 ? 1.0 : fabs(piv);
            }
            iv[k] = rsqrt(piv);  // a few ulp from sqrt/divide: far inside the FP64 tolerance
            l[k][k] = piv * iv[k];
#pragma unroll
            for (int i = k + 1; i < 4; ++i) {
                double v = d[i][k];
#pragma unroll
                for (int q = 0; q < k; ++q) v = fma(-l[i][q], l[k][q], v);
                l[i][k] = v * iv[k];
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            m[i][i] = iv[i];
#pragma unroll
            for (int j = 0; j < i; ++j) {
                double v = 0.0;
#pragma unroll
                for (int q = j; q < i; ++q) v = fma(l[i][q], m[q][j], v);
                m[i][j] = -v * iv[i];
            }
        }
        // ---- eliminated pivot rows of X for this thread's columns: Xk = M X_kb (needed by tj <= kb only)
        double xk[4][4];
        if (tj <= kb) {
            double xo[4][4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const double2 v01 = *reinterpret_cast<const double2*>(&xr[buf][r][4 * tj]);
                const double2 v23 = *reinterpret_cast<const double2*>(&xr[buf][r][4 * tj + 2]);
                xo[r][0] = v01.x;
                xo[r][1] = v01.y;
                xo[r][2] = v23.x;
                xo[r][3] = v23.y;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    double v = m[k][0] * xo[0][c];
#pragma unroll
                    for (int q = 1; q <= k; ++q) v = fma(m[k][q], xo[q][c], v);
                    xk[k][c] = v;
                }
        }
        if (ti == kb) {
            // pivot block row: X rows replaced; the diagonal block of A becomes Ld
            if (tj == kb) {
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int c = 0; c <= r; ++c) a[r][c] = l[r][c];
            }
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) x[r][c] = xk[r][c];
            continue;
        }
        // ---- rows below the pivot block: L_i = A_i Ld^-T
        double li[4][4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const double2 v01 = *reinterpret_cast<const double2*>(&cb[buf][4 * ti + r][0]);
            const double2 v23 = *reinterpret_cast<const double2*>(&cb[buf][4 * ti + r][2]);
            const double ar[4] = {v01.x, v01.y, v23.x, v23.y};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                double v = ar[0] * m[k][0];
#pragma unroll
                for (int q = 1; q <= k; ++q) v = fma(ar[q], m[k][q], v);
                li[r][k] = v;
            }
        }
        if (tj > kb) {  // trailing block of A: needs the panel rows of this thread's columns as well
            double lj[4][4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const double2 v01 = *reinterpret_cast<const double2*>(&cb[buf][4 * tj + c][0]);
                const double2 v23 = *reinterpret_cast<const double2*>(&cb[buf][4 * tj + c][2]);
                const double ar[4] = {v01.x, v01.y, v23.x, v23.y};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    double v = ar[0] * m[k][0];
#pragma unroll
                    for (int q = 1; q <= k; ++q) v = fma(ar[q], m[k][q], v);
                    lj[c][k] = v;
                }
            }
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int k = 0; k < 4; ++k) a[r][c] = fma(-li[r][k], lj[c][k], a[r][c]);
        } else {  // tj <= kb: X, and for tj == kb the panel block itself
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int k = 0; k < 4; ++k) x[r][c] = fma(-li[r][k], xk[k][c], x[r][c]);
            if (tj == kb) {
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int c = 0; c < 4; ++c) a[r][c] = li[r][c];
            }
        }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int i = 4 * ti + r, j = 4 * tj + c;
            if (i < bs && j <= i) A[(size_t)i * ld + j] = a[r][c];
            inv[i * BS + j] = (i < bs && j <= i) ? x[r][c] : 0.0;
        }
}

// v <- L^-1 v: one CTA of 1024 threads walks the 64-row blocks; block b: r = v_b - L[b, 0:b0] v[0:b0], v_b = inv_b r.
__global__ void __launch_bounds__(1024) trsv_lower_kernel(double* v, const double* L, size_t ldl, int n, const double* inv) {
    __shared__ double r[BS];
    __shared__ double part[16][BS];
    const int tid = threadIdx.x;
    for (int b0 = 0; b0 < n; b0 += BS) {
        const int bs = min(BS, n - b0);
        // 16 threads per row, each a strided slice of the already-solved prefix
        const int row = tid >> 4, sl = tid & 15;
        double sum = 0.0;
        if (row < bs) {
            const double* lr = L + (size_t)(b0 + row) * ldl;
            for (int j = sl; j < b0; j += 16) sum = fma(lr[j], v[j], sum);
        }
        part[sl][row] = sum;
        __syncthreads();
        if (tid < bs) {
            double t = v[b0 + tid];
            for (int k = 0; k < 16; ++k) t -= part[k][tid];
            r[tid] = t;
        }
        __syncthreads();
        if (tid < bs) {
            const double* ib = inv + (size_t)(b0 / BS) * BS * BS + (size_t)tid * BS;
            double t = 0.0;
            for (int j = 0; j <= tid; ++j) t = fma(ib[j], r[j], t);
            v[b0 + tid] = t;
        }
        __syncthreads();
    }
}


// ---------------------------------------------------------------------------------- 256-wide triangular solve
// Inverse of the 256 x 256 diagonal blocks of L from the stored inverses of their 64 x 64 diagonal blocks:
// for the block-lower-triangular T (4 x 4 blocks of 64), column block jb of T^-1 is
//     Tinv[jb][jb] = inv64,   Tinv[i][jb] = -inv64_i * sum_{k = jb}^{i-1} T[i][k] Tinv[k][jb]    (i > jb)
// Columns are independent: one CTA per (column block jb, 256-block b), 256 threads, each a 4 x 4 patch of the 64 x 64
// output; operands pass through shared memory in slabs of 32. Writes the inverse (inv256, [r][c]) and its transpose
// (inv256T, [c][r]); everything above the block diagonal stays zero (the buffers are zero-filled once).
constexpr int TW = TRSM_WIDE;

// acc += A (64 x 64, row-major lda, rows >= a_rows read as 0) * B (64 x 64, row-major ldb, rows >= b_rows read as 0)
__device__ __forceinline__ void tri_prod64(double (&acc)[4][4], const double* __restrict__ A, size_t lda, int a_rows,
                                           const double* __restrict__ B, size_t ldb, int b_rows, double (*As)[33],
                                           double (*Bs)[65]) {
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    for (int k0 = 0; k0 < 64; k0 += 32) {
        __syncthreads();  // the previous slab (or the caller's last use of the tiles) is done
        for (int e = tid; e < 64 * 32; e += 256) {
            const int r = e >> 5, k = e & 31;
            As[r][k] = r < a_rows ? A[(size_t)r * lda + k0 + k] : 0.0;
        }
        for (int e = tid; e < 32 * 64; e += 256) {
            const int k = e >> 6, c = e & 63;
            Bs[k][c] = (k0 + k) < b_rows ? B[(size_t)(k0 + k) * ldb + c] : 0.0;
        }
        __syncthreads();
#pragma unroll 8
        for (int k = 0; k < 32; ++k) {
            double a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[4 * ty + i][k];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = Bs[k][4 * tx + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
        }
    }
}

__global__ void __launch_bounds__(256) tri_inv256_kernel(const double* __restrict__ L, size_t ldl, int n,
                                                         const double* __restrict__ inv64, double* __restrict__ inv256,
                                                         double* __restrict__ inv256T) {
    __shared__ double As[64][33];
    __shared__ double Bs[32][65];
    const int jb = blockIdx.x, b = blockIdx.y;
    const int base = b * TW;
    const int bs256 = min(TW, n - base);
    const int nsub = (bs256 + 63) / 64;
    if (jb >= nsub) return;
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    double* out = inv256 + (size_t)b * TW * TW;
    double* outT = inv256T + (size_t)b * TW * TW;
    auto store = [&](int i, const double (&v)[4][4]) {  // block (i, jb) of the inverse, and its transpose
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int rr = 64 * i + 4 * ty + r, cc = 64 * jb + 4 * tx + c;
                out[(size_t)rr * TW + cc] = v[r][c];
                outT[(size_t)cc * TW + rr] = v[r][c];
            }
    };
    {
        const double* d = inv64 + (size_t)(4 * b + jb) * BS * BS;
        double v[4][4];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) v[r][c] = d[(4 * ty + r) * BS + 4 * tx + c];
        store(jb, v);
    }
    for (int i = jb + 1; i < nsub; ++i) {
        double acc[4][4];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[r][c] = 0.0;
        const int rows_i = min(64, n - (base + 64 * i));  // valid rows of block row i of L
        for (int k = jb; k < i; ++k)  // T[i][k] * Tinv[k][jb]  (Tinv[k][jb] was written by this CTA: same column)
            tri_prod64(acc, L + (size_t)(base + 64 * i) * ldl + base + 64 * k, ldl, rows_i, out + (size_t)(64 * k) * TW + 64 * jb,
                       TW, 64, As, Bs);
        store(i, acc);  // parked in the output block: the second product reads it back as its right operand
        double res[4][4];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) res[r][c] = 0.0;
        tri_prod64(res, inv64 + (size_t)(4 * b + i) * BS * BS, BS, 64, out + (size_t)(64 * i) * TW + 64 * jb, TW, 64, As, Bs);
        __syncthreads();  // every thread has read the parked block
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) res[r][c] = -res[r][c];
        store(i, res);
    }
}

__global__ void symmetrize_kernel(double* A, size_t ld, int n) {
    __shared__ double tile[32][33];
    const int bi = blockIdx.y, bj = blockIdx.x;
    if (bj > bi) return;
    const int tx = threadIdx.x, ty = threadIdx.y;  // 32 x 8
    for (int k = ty; k < 32; k += 8) {
        const int i = bi * 32 + k, j = bj * 32 + tx;
        tile[k][tx] = (i < n && j < n) ? A[(size_t)i * ld + j] : 0.0;
    }
    __syncthreads();
    for (int k = ty; k < 32; k += 8) {
        const int j = bj * 32 + k, i = bi * 32 + tx;  // destination A[j][i] = source A[i][j]
        if (i < n && j < n && i > j) A[(size_t)j * ld + i] = tile[tx][k];
    }
}

}  // namespace

// Right-looking with one block column of lookahead: after L21 of panel k, the NEXT block column of the trailing matrix
// is updated first (a narrow GEMM), the one-CTA panel kernel of block k+1 then runs on a side stream while the rest
// of the trailing update keeps the other SMs busy on the caller's stream; the two meet again before L21 of k+1.
aslam_status cholesky_lower(cudaStream_t st, double* A, size_t ld, int n, double* inv_blocks, int* d_fail,
                            uint64_t* launches) {
    SideStream* ss = nullptr;  // per thread and device
    {
        const aslam_status rc = side_stream(0, &ss);
        if (rc != ASLAM_OK) return rc;
    }
    const cudaStream_t side = ss->s;
    const cudaEvent_t ev_col = ss->fork, ev_panel = ss->join;
    chol_panel_kernel<<<1, 256, 0, st>>>(A, ld, n < BS ? n : BS, inv_blocks, d_fail, 0);
    ASLAM_CUDA(cudaGetLastError());
    if (launches) ++*launches;
    for (int k0 = 0; k0 < n; k0 += BS) {
        const int bs = n - k0 < BS ? n - k0 : BS;
        const int rs = n - k0 - bs;
        if (rs <= 0) break;
        double* inv = inv_blocks + (size_t)(k0 / BS) * BS * BS;
        double* A21 = A + (size_t)(k0 + bs) * ld + k0;
        double* A22 = A + (size_t)(k0 + bs) * ld + (k0 + bs);
        Epilogue e;  // L21 = A21 inv(L11)'  (in place: one CTA owns all 64 columns of its row block, K = 64 <= 4 slabs)
        aslam_status rc = dgemm_nt(st, rs, bs, bs, A21, ld, inv, BS, A21, ld, e, launches);
        if (rc != ASLAM_OK) return rc;
        const int nb = rs < BS ? rs : BS;  // width of the next block column
        Epilogue t;  // A22 -= L21 L21'
        t.alpha = -1.0;
        t.beta = 1.0;
        rc = dgemm_nt(st, rs, nb, bs, A21, ld, A21, ld, A22, ld, t, launches);  // next block column first
        if (rc != ASLAM_OK) return rc;
        ASLAM_CUDA(cudaEventRecord(ev_col, st));
        ASLAM_CUDA(cudaStreamWaitEvent(side, ev_col, 0));
        chol_panel_kernel<<<1, 256, 0, side>>>(A22, ld, nb, inv + (size_t)BS * BS, d_fail, k0 + bs);
        ASLAM_CUDA(cudaGetLastError());
        if (launches) ++*launches;
        ASLAM_CUDA(cudaEventRecord(ev_panel, side));
        if (rs > nb) {  // the rest of the trailing matrix, lower tiles only, concurrently with the panel
            t.lower_only = true;
            double* L2 = A21 + (size_t)nb * ld;
            rc = dgemm_nt(st, rs - nb, rs - nb, bs, L2, ld, L2, ld, A22 + (size_t)nb * ld + nb, ld, t, launches);
            if (rc != ASLAM_OK) return rc;
        }
        ASLAM_CUDA(cudaStreamWaitEvent(st, ev_panel, 0));
    }
    return ASLAM_OK;
}

aslam_status trsm_right_lower_t(cudaStream_t st, double* B, size_t ldb, int m, const double* L, size_t ldl, int n,
                                const double* inv_blocks, uint64_t* launches) {
    // X L' = B, right-looking by 64-column blocks: X_j = B_j inv(L_jj)' (in place), then the columns to the right lose
    // X_j L[>j, j]'. Every update is a wide GEMM (all SMs busy); the left-looking form leaves 17 CTAs walking a long K.
    for (int j0 = 0; j0 < n; j0 += BS) {
        const int bs = n - j0 < BS ? n - j0 : BS;
        Epilogue f;
        aslam_status rc = dgemm_nt(st, m, bs, bs, B + j0, ldb, inv_blocks + (size_t)(j0 / BS) * BS * BS, BS, B + j0, ldb, f,
                                   launches);
        if (rc != ASLAM_OK) return rc;
        const int rest = n - j0 - bs;
        if (rest > 0) {
            Epilogue e;
            e.alpha = -1.0;
            e.beta = 1.0;
            rc = dgemm_nt(st, m, rest, bs, B + j0, ldb, L + (size_t)(j0 + bs) * ldl + j0, ldl, B + j0 + bs, ldb, e, launches);
            if (rc != ASLAM_OK) return rc;
        }
    }
    return ASLAM_OK;
}


aslam_status trsm_right_lower_t_wide(cudaStream_t st, double* B, size_t ldb, int m, const double* L, size_t ldl, int n,
                                     const double* inv_blocks, double* inv256, double* inv256T, double* Lt, size_t ldt,
                                     double* X, size_t ldx, uint64_t* launches) {
    const int nb = (n + TW - 1) / TW;
    tri_inv256_kernel<<<dim3(4, nb), 256, 0, st>>>(L, ldl, n, inv_blocks, inv256, inv256T);
    ASLAM_CUDA(cudaGetLastError());
    if (launches) ++*launches;
    aslam_status rc;
    if (nb > 1) {  // Lt[r][c] = sum_{k in blk(c)} L[r][k] inv256[k][c]: rows below the block diagonal are what is used
        Epilogue e;
        e.lower_only = true;
        e.kblock = TW;
        e.a_cols = n;
        rc = dgemm_nt(st, n, n, TW, L, ldl, inv256T, TW, Lt, ldt, e, launches);
        if (rc != ASLAM_OK) return rc;
    }
    for (int j0 = 0; j0 + TW < n; j0 += TW) {  // B[:, > j] -= B[:, j] Lt[> j, j]'
        const int rest = n - j0 - TW;
        Epilogue e;
        e.alpha = -1.0;
        e.beta = 1.0;
        rc = dgemm_nt(st, m, rest, TW, B + j0, ldb, Lt + (size_t)(j0 + TW) * ldt + j0, ldt, B + j0 + TW, ldb, e, launches);
        if (rc != ASLAM_OK) return rc;
    }
    Epilogue x;  // X[:, j] = B[:, j] inv256_j'
    x.kblock = TW;
    x.a_cols = n;
    return dgemm_nt(st, m, n, TW, B, ldb, inv256, TW, X, ldx, x, launches);
}

aslam_status trsv_lower(cudaStream_t st, double* v, const double* L, size_t ldl, int n, const double* inv_blocks,
                        uint64_t* launches) {
    trsv_lower_kernel<<<1, 1024, 0, st>>>(v, L, ldl, n, inv_blocks);
    ASLAM_CUDA(cudaGetLastError());
    if (launches) ++*launches;
    return ASLAM_OK;
}

aslam_status symmetrize_from_lower(cudaStream_t st, double* A, size_t ld, int n, uint64_t* launches) {
    const int nb = (n + 31) / 32;
    symmetrize_kernel<<<dim3(nb, nb), dim3(32, 8), 0, st>>>(A, ld, n);
    ASLAM_CUDA(cudaGetLastError());
    if (launches) ++*launches;
    return ASLAM_OK;
}

}  // namespace dense
}  // namespace aslam
