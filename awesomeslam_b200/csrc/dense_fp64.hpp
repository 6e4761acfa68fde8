// FP64 dense building blocks of the large-map UKF (kernels K6, K8, K9): the contractions
// the reference runs as Eigen products / rank-1 loops / LLT / inverse (src/ukf/ukf.cpp:280, 307-391).
// All matrices are ROW-major with the contraction index contiguous ("NT" form): every product is
//     C[i][j] = alpha * sum_k A[i][k] * B[j][k]  (+ beta * C[i][j])  (+ r1 * u[i] * v[j])  (+ diag[i] if i == j)
// Tiles are moved by TMA (cp.async.bulk.tensor.2d, SASS UTMALDG) into a 4-stage shared-memory ring by a dedicated
// producer warp and contracted on the FP64 tensor pipe (mma.sync m8n8k4 f64, SASS DMMA.8x8x4) -- tcgen05 has no
// FP64 kind, DMMA is the FP64 tensor path of sm_100a.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "internal.hpp"

namespace aslam {
namespace dense {

constexpr int CHOL_BS = 64;  // block size of the blocked Cholesky / triangular solves

struct Epilogue {
    double alpha = 1.0, beta = 0.0;
    const double* u = nullptr;     // rank-1 term r1 * u[i] * v[j] (u indexed by the row of C, v by its column)
    const double* v = nullptr;
    double r1 = 0.0;
    const double* r1_dev = nullptr;  // when set, the coefficient is read from device memory instead of r1
    const double* diag = nullptr;  // added to C[i][i]
    bool lower_only = false;       // skip tiles that lie entirely above the diagonal
    // Block-diagonal second operand: when kblock > 0, B is [N][kblock] (the kblock x kblock diagonal blocks stacked)
    // and column tile c contracts A's columns [blk, blk + kblock) with blk = (c / kblock) * kblock:
    //     C[i][c] = alpha * sum_{k < kblock} A[i][blk + k] * B[c][k] + ...      (K must equal kblock)
    // a_cols = number of valid columns of A (the tensor map zero-fills beyond it).
    int kblock = 0;
    int a_cols = 0;
    // > 0: persistent form on at most max_ctas CTAs (one per SM), each walking a list of 128 x 128 tiles. Leaves the
    // other SMs to whatever runs on other streams (the large-map UKF factors M while the P and Tc products run).
    int max_ctas = 0;
};

// C (M x N, ldc) = alpha * A (M x K, lda) * B (N x K, ldb)' + ...   Base pointers 16-byte aligned, ld even.
// Returns the number of kernels launched through *launches (may be null).
aslam_status dgemm_nt(cudaStream_t st, int M, int N, int K, const double* A, size_t lda, const double* B, size_t ldb,
                      double* C, size_t ldc, const Epilogue& e, uint64_t* launches);

// In-place blocked Cholesky of the lower triangle of A (n x n, row-major, ld): A <- L with L L' = A. The strict
// upper triangle is left with unspecified values. inv_blocks [ceil(n/64)][64*64] receives the inverse of every
// 64 x 64 diagonal block of L (row-major, pitch 64). *d_fail (device int, must hold -1 on entry) receives the
// index of the first non-positive pivot (Eigen's LLT would stop there, ukf.cpp:280); the factorisation carries on
// with unspecified values after it.
aslam_status cholesky_lower(cudaStream_t st, double* A, size_t ld, int n, double* inv_blocks, int* d_fail,
                            uint64_t* launches);

// B (m x n, ldb) <- B * L^-T for the lower-triangular L (n x n, ldl) whose diagonal-block inverses are inv_blocks.
aslam_status trsm_right_lower_t(cudaStream_t st, double* B, size_t ldb, int m, const double* L, size_t ldl, int n,
                                const double* inv_blocks, uint64_t* launches);

// The same solve OUT of place with 256-wide blocks (the form the large-map UKF uses): X (m x n, ldx) = B L^-T; B is
// destroyed. inv256 / inv256T: scratch [ceil(n/256)][256*256] each (zero-filled once by the caller), Lt: scratch
// n x ldt. Critical chain = ceil(n/256) wide GEMMs with K = 256 instead of 2 n/64 dependent ones with K = 64:
//   inv256_j = inverse of the j-th 256 x 256 diagonal block of L (tri_inv256_kernel, from the 64-block inverses)
//   Lt[>j, j] = L[>j, j] inv256_j      (one launch for all j: block-diagonal operand)
//   for j: B[:, >j] -= B[:, j] Lt[>j, j]'
//   X = B blockdiag(inv256_j)'          (one launch)
aslam_status trsm_right_lower_t_wide(cudaStream_t st, double* B, size_t ldb, int m, const double* L, size_t ldl, int n,
                                     const double* inv_blocks, double* inv256, double* inv256T, double* Lt, size_t ldt,
                                     double* X, size_t ldx, uint64_t* launches);
constexpr int TRSM_WIDE = 256;

// v (n) <- L^-1 v (forward substitution with the same blocked data), one CTA.
aslam_status trsv_lower(cudaStream_t st, double* v, const double* L, size_t ldl, int n, const double* inv_blocks,
                        uint64_t* launches);

// A[j][i] <- A[i][j] for i > j.
aslam_status symmetrize_from_lower(cudaStream_t st, double* A, size_t ld, int n, uint64_t* launches);

// A side stream with a fork and a join event, owned by the calling thread, one per (slot, device): small kernels that
// are independent of the caller's stream for a while (the Cholesky's lookahead panel, a GEMM's ragged-edge rows) run on
// it. Created on first use for the CURRENT device. Slots: 0 the Cholesky, 1 dgemm_nt, 2 the large-map UKF (chol(M) beside the products).
struct SideStream {
    cudaStream_t s = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
};
aslam_status side_stream(int slot, SideStream** out);

// Measured throughput of the DMMA pipe of the current device, TFLOP/s (roofline denominator for the GEMMs).
aslam_status probe_dmma_tflops(double* tflops);

// ---- TF32 variant (gemm_tf32_tcgen05.cu): tcgen05.mma kind::tf32, accumulator in TMEM, operands by TMA.
// hi = tf32(x), lo = tf32(x - hi) of an FP64 matrix [rows][ld], stored FP32 [rows][tf32_padded_k(K)] (lo may be null).
size_t tf32_padded_k(int K);
aslam_status split_tf32(cudaStream_t st, const double* src, size_t ld, int rows, int K, float* hi, float* lo);
// C (FP64, M x N, ldc) = alpha * A B' + the Epilogue terms. passes = 1: A_hi B_hi'; passes = 3: + A_hi B_lo' + A_lo B_hi'.
// *d_err (device int, 0 on entry) becomes non-zero when a bounded wait inside the kernel expired.
aslam_status tf32_gemm_nt(cudaStream_t st, int M, int N, int K, const float* Ahi, const float* Alo, const float* Bhi,
                          const float* Blo, double* C, size_t ldc, const Epilogue& e, int passes, int* d_err,
                          uint64_t* launches = nullptr);
// Measured kind::tf32 throughput of the current device (MMAs issued back to back from shared memory), TFLOP/s.
aslam_status probe_tf32_tflops(double* tflops);

}  // namespace dense
}  // namespace aslam
