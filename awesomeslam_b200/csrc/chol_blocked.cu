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
    __shared__ double col[2][BS];
    __shared__ double xrow[2][BS];
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
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            const int k = 4 * kb + kk, buf = kk & 1;
            if (tj == kb) {
#pragma unroll
                for (int r = 0; r < 4; ++r) col[buf][4 * ti + r] = a[r][kk];
            }
            if (ti == kb) {
#pragma unroll
                for (int c = 0; c < 4; ++c) xrow[buf][4 * tj + c] = x[kk][c];
            }
            __syncthreads();
            double akk = col[buf][k];
            if (!(akk > 0.0)) {
                if (tid == 0 && k < bs) atomicCAS(fail, -1, k0 + k);
                akk = akk == 0.0 ? 1.0 : fabs(akk);
            }
            const double rd = rsqrt(akk), d = akk * rd;  // a few ulp from sqrt/divide: far inside the FP64 tolerance
            double li[4], lj[4], xk[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                li[r] = col[buf][4 * ti + r] * rd;
                lj[r] = col[buf][4 * tj + r] * rd;
                xk[r] = xrow[buf][4 * tj + r] * rd;
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int i = 4 * ti + r;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int j = 4 * tj + c;
                    if (i > k && j > k) a[r][c] = fma(-li[r], lj[c], a[r][c]);
                    if (i > k) x[r][c] = fma(-li[r], xk[c], x[r][c]);
                }
            }
            if (tj == kb) {  // column k of L
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const int i = 4 * ti + r;
                    if (i > k) a[r][kk] = li[r];
                    if (i == k) a[r][kk] = d;
                }
            }
            if (ti == kb) {  // row k of inv(L)
#pragma unroll
                for (int c = 0; c < 4; ++c) x[kk][c] = xk[c];
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

aslam_status cholesky_lower(cudaStream_t st, double* A, size_t ld, int n, double* inv_blocks, int* d_fail,
                            uint64_t* launches) {
    for (int k0 = 0; k0 < n; k0 += BS) {
        const int bs = n - k0 < BS ? n - k0 : BS;
        double* A11 = A + (size_t)k0 * ld + k0;
        double* inv = inv_blocks + (size_t)(k0 / BS) * BS * BS;
        chol_panel_kernel<<<1, 256, 0, st>>>(A11, ld, bs, inv, d_fail, k0);
        ASLAM_CUDA(cudaGetLastError());
        if (launches) ++*launches;
        const int rs = n - k0 - bs;
        if (rs <= 0) break;
        double* A21 = A + (size_t)(k0 + bs) * ld + k0;
        double* A22 = A + (size_t)(k0 + bs) * ld + (k0 + bs);
        Epilogue e;  // L21 = A21 inv(L11)'  (in place: one CTA owns all 64 columns of its row block, K = 64 <= 4 slabs)
        aslam_status rc = dgemm_nt(st, rs, bs, bs, A21, ld, inv, BS, A21, ld, e, launches);
        if (rc != ASLAM_OK) return rc;
        Epilogue t;  // A22 -= L21 L21'
        t.alpha = -1.0;
        t.beta = 1.0;
        t.lower_only = true;
        rc = dgemm_nt(st, rs, rs, bs, A21, ld, A21, ld, A22, ld, t, launches);
        if (rc != ASLAM_OK) return rc;
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
