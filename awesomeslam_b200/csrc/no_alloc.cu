// aslam_dense_gemm_nt_tf32_static: the TF32 (tcgen05) GEMM of include/aslam_c.h for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it). No counterpart in the reference.
// Same arithmetic as aslam_dense_gemm_nt_tf32 (the same split and GEMM kernels of
// gemm_tf32_tcgen05.cu in the same order); the one difference is where the scratch lives.
//
// aslam_dense_gemm_nt_tf32 takes its scratch -- the FP32 hi / lo copies of both operands and the pipeline's error flag --
// from cudaMalloc on every call. Here it is a file-scope `static __device__` array of ASLAM_TF32_STATIC_SCRATCH_BYTES
// (8 MiB) + 256 bytes: every GPU that loads the library holds it until the process exits. In HOST code the array's name is
// a host-side stand-in, not a device address, so the address is looked up with cudaGetSymbolAddress -- in this file, once
// per device, with that device current -- and only that pointer is handed to the kernels.
//
// Every call on a device shares the array and finds the last call's contents in it: the split kernels rewrite every float
// the GEMM reads (padding columns included) and the error flag is cleared on the caller's stream before the GEMM.
// Two calls that overlap on one GPU (different streams, or different host threads) RACE on it: serialise them.
#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "dense_fp64.hpp"

namespace {

constexpr size_t SCRATCH_BYTES = ASLAM_TF32_STATIC_SCRATCH_BYTES;
constexpr size_t FLAG_BYTES = 256;  // the error flag of the tcgen05 pipeline lives behind the operand copies

static __device__ __align__(256) unsigned char g_tf32_scratch[SCRATCH_BYTES + FLAG_BYTES];

// Device address of g_tf32_scratch on the CURRENT device, looked up once per device.
aslam_status scratch_address(unsigned char** out) {
    constexpr int MAXDEV = 64;
    static std::mutex mu;
    static unsigned char* addr[MAXDEV] = {};
    int dev = -1;
    ASLAM_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAXDEV) {
        aslam::set_error("aslam_dense_gemm_nt_tf32_static: device index out of range");
        return ASLAM_ERR_ARG;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void* p = nullptr;
        ASLAM_CUDA(cudaGetSymbolAddress(&p, g_tf32_scratch));
        if (!p) {
            aslam::set_error("aslam_dense_gemm_nt_tf32_static: cudaGetSymbolAddress returned a null address");
            return ASLAM_ERR_CUDA;
        }
        addr[dev] = static_cast<unsigned char*>(p);
    }
    *out = addr[dev];
    return ASLAM_OK;
}

size_t round256(size_t b) { return (b + 255) / 256 * 256; }

}  // namespace

extern "C" aslam_status aslam_dense_gemm_nt_tf32_static(void* stream, int32_t M, int32_t N, int32_t K, const double* d_A,
                                                        size_t lda, const double* d_B, size_t ldb, double* d_C,
                                                        size_t ldc, double alpha, double beta, int32_t lower_only,
                                                        int32_t passes) {
    using namespace aslam;
    if (!d_A || !d_B || !d_C || M <= 0 || N <= 0 || K <= 0 || (passes != 1 && passes != 3)) return ASLAM_ERR_ARG;
    // scratch layout: A_hi | B_hi | (passes == 3: A_lo | B_lo), each [rows][tf32_padded_k(K)] floats, 256-byte aligned
    const size_t ldf = dense::tf32_padded_k(K);
    const size_t a_bytes = round256((size_t)M * ldf * sizeof(float)), b_bytes = round256((size_t)N * ldf * sizeof(float));
    const size_t need = (a_bytes + b_bytes) * (passes == 3 ? 2 : 1);
    if (need > SCRATCH_BYTES) {
        set_error("aslam_dense_gemm_nt_tf32_static: M = " + std::to_string(M) + ", N = " + std::to_string(N) + ", K = " +
                  std::to_string(K) + ", passes = " + std::to_string(passes) + " needs " + std::to_string(need) +
                  " bytes of scratch, the static scratch holds " + std::to_string(SCRATCH_BYTES) +
                  " (ASLAM_TF32_STATIC_SCRATCH_BYTES); use aslam_dense_gemm_nt_tf32");
        return ASLAM_ERR_UNSUPPORTED;
    }
    unsigned char* base = nullptr;
    aslam_status rc = scratch_address(&base);
    if (rc != ASLAM_OK) return rc;
    float* ahi = reinterpret_cast<float*>(base);
    float* bhi = reinterpret_cast<float*>(base + a_bytes);
    float* alo = passes == 3 ? reinterpret_cast<float*>(base + a_bytes + b_bytes) : nullptr;
    float* blo = passes == 3 ? reinterpret_cast<float*>(base + 2 * a_bytes + b_bytes) : nullptr;
    int* err = reinterpret_cast<int*>(base + SCRATCH_BYTES);
    cudaStream_t st = (cudaStream_t)stream;
    ASLAM_CUDA(cudaMemsetAsync(err, 0, sizeof(int), st));  // the one thing that must start at zero
    rc = dense::split_tf32(st, d_A, lda, M, K, ahi, alo);
    if (rc == ASLAM_OK) rc = dense::split_tf32(st, d_B, ldb, N, K, bhi, blo);
    dense::Epilogue e;
    e.alpha = alpha;
    e.beta = beta;
    e.lower_only = lower_only != 0;
    if (rc == ASLAM_OK) rc = dense::tf32_gemm_nt(st, M, N, K, ahi, alo, bhi, blo, d_C, ldc, e, passes, err);
    int herr = 0;
    if (rc == ASLAM_OK && !cuda_ok(cudaMemcpyAsync(&herr, err, sizeof(int), cudaMemcpyDeviceToHost, st), "tf32 err")) rc = ASLAM_ERR_CUDA;
    if (!cuda_ok(cudaStreamSynchronize(st), "tf32 gemm (static scratch)")) rc = ASLAM_ERR_CUDA;
    if (rc == ASLAM_OK && herr) {
        set_error("tf32 gemm: a wait inside the tcgen05 pipeline expired (code " + std::to_string(herr) + ")");
        return ASLAM_ERR_CUDA;
    }
    return rc;
}
