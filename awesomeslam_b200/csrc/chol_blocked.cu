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
constexpr int PITCH = BS + 1;

// One CTA, 256 threads. A: bs x bs diagonal block (row-major, ld). Factor in place (lower), write inv(L) (pitch 64,
// zero above the diagonal and beyond bs). k0: global index of the block's first row (for the failure report).
__global__ void __launch_bounds__(256) chol_panel_kernel(double* A, size_t ld, int bs, double* inv, int* fail, int k0) {
    extern __shared__ __align__(16) unsigned char panel_smem[];
    double(*s)[PITCH] = reinterpret_cast<double(*)[PITCH]>(panel_smem);
    double(*si)[PITCH] = s + BS;
    const int tid = threadIdx.x;
    for (int idx = tid; idx < BS * BS; idx += 256) {
        const int i = idx / BS, j = idx % BS;
        s[i][j] = (i < bs && j <= i) ? A[(size_t)i * ld + j] : (i == j ? 1.0 : 0.0);
        si[i][j] = 0.0;
    }
    __syncthreads();
    for (int k = 0; k < bs; ++k) {
        if (tid == 0) {
            double x = s[k][k];
            if (!(x > 0.0)) {  // Eigen: `if (x <= 0) return k;` -- recorded, then carried on with |x| so nothing traps
                atomicCAS(fail, -1, k0 + k);
                x = x == 0.0 ? 1.0 : fabs(x);
            }
            s[k][k] = sqrt(x);
        }
        __syncthreads();
        const double d = s[k][k];
        for (int i = k + 1 + tid; i < bs; i += 256) s[i][k] = s[i][k] / d;
        __syncthreads();
        const int rem = bs - k - 1;
        for (int idx = tid; idx < rem * rem; idx += 256) {
            const int i = k + 1 + idx / rem, j = k + 1 + idx % rem;
            if (j <= i) s[i][j] = fma(-s[i][k], s[j][k], s[i][j]);
        }
        __syncthreads();
    }
    for (int idx = tid; idx < bs * bs; idx += 256) {
        const int i = idx / bs, j = idx % bs;
        if (j <= i) A[(size_t)i * ld + j] = s[i][j];
    }
    // inverse: column c of inv(L) by forward substitution, 4 lanes per column splitting the inner sum
    // (every lane runs the same trip count: the shuffles are full-warp)
    const int c = tid >> 2, part = tid & 3;
    if (c < bs && part == 0) si[c][c] = 1.0 / s[c][c];
    __syncwarp();
    for (int i = 1; i < bs; ++i) {
        const bool act = c < bs && i > c;
        double sum = 0.0;
        if (act)
            for (int j = c + part; j < i; j += 4) sum = fma(s[i][j], si[j][c], sum);
        sum += __shfl_xor_sync(0xffffffffu, sum, 1);
        sum += __shfl_xor_sync(0xffffffffu, sum, 2);
        if (act && part == 0) si[i][c] = -sum / s[i][i];
        __syncwarp();
    }
    __syncthreads();
    for (int idx = tid; idx < BS * BS; idx += 256) {
        const int i = idx / BS, j = idx % BS;
        inv[idx] = (i < bs && j <= i) ? si[i][j] : 0.0;
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
    constexpr int PANEL_SMEM = 2 * BS * PITCH * (int)sizeof(double);
    static bool attr = false;
    if (!attr) {
        ASLAM_CUDA(cudaFuncSetAttribute(chol_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PANEL_SMEM));
        attr = true;
    }
    for (int k0 = 0; k0 < n; k0 += BS) {
        const int bs = n - k0 < BS ? n - k0 : BS;
        double* A11 = A + (size_t)k0 * ld + k0;
        double* inv = inv_blocks + (size_t)(k0 / BS) * BS * BS;
        chol_panel_kernel<<<1, 256, PANEL_SMEM, st>>>(A11, ld, bs, inv, d_fail, k0);
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
    // X L' = B, column block j: X_j = (B_j - X_{<j} L[j, <j]') inv(L_jj)'
    for (int j0 = 0; j0 < n; j0 += BS) {
        const int bs = n - j0 < BS ? n - j0 : BS;
        if (j0 > 0) {
            Epilogue e;
            e.alpha = -1.0;
            e.beta = 1.0;
            aslam_status rc = dgemm_nt(st, m, bs, j0, B, ldb, L + (size_t)j0 * ldl, ldl, B + j0, ldb, e, launches);
            if (rc != ASLAM_OK) return rc;
        }
        Epilogue f;
        aslam_status rc = dgemm_nt(st, m, bs, bs, B + j0, ldb, inv_blocks + (size_t)(j0 / BS) * BS * BS, BS, B + j0, ldb, f,
                                   launches);
        if (rc != ASLAM_OK) return rc;
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
