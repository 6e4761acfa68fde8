// TF32 variant of the sigma-point contractions of the large-map UKF (north_star: "FP64/TF32 tensor-core tiles via
// TMA"; SURVEY.md 8d cfg5: "TF32 variant: separate row, tolerance stated separately"). The three uniform-weight products
//     P ~ wc D D',   M ~ wc DZ DZ',   Tc ~ wc D DZ'          (src/ukf/ukf.cpp:307-319, 342-357, 360-375)
// are the one place on this path where Blackwell's own tensor path applies: tcgen05.mma kind::tf32 with the
// accumulator in tensor memory (TMEM). The sigma-point MEAN and the w0 rank-1 term stay in FP64 (FP32
// accumulation of the w0-weighted sums is off by 1e-2 .. 28 in P) -- only the already centred D, DZ are rounded.
//
//   split_tf32_kernel     FP64 matrix -> hi = tf32(x), lo = tf32(x - hi), both stored as FP32 (K padded to 32, zeros)
//   tf32_gemm_nt_kernel   C(FP64) = alpha * A B' (+ beta C + r1 u v' + diag), 128 x 128 tile per CTA:
//       warp 0      one lane: TMA producer (cp.async.bulk.tensor.2d, 128-byte swizzle) into an S-stage ring
//       warp 1      allocates 128 TMEM columns; one lane issues tcgen05.mma.cta_group::1.kind::tf32 (M 128, N 128, K 8,
//                   operands from shared-memory descriptors), tcgen05.commit frees the stage / publishes the accumulator
//       warps 2..5  epilogue: tcgen05.ld 32x32b (each warp its own 32 TMEM lanes = 32 rows), FP32 -> FP64, fused
//                   epilogue, 256-byte row segments to C
//     PASSES = 1: A_hi B_hi'.   PASSES = 3: A_hi B_hi' + A_hi B_lo' + A_lo B_hi' (error-compensated split: the
//     product of the two lo parts, ~2^-22 of the result, is dropped; what remains is the FP32 accumulation in TMEM).
// Every wait in the kernel is bounded: a broken pipeline raises an error flag and the CTA leaves instead of hanging.
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <mutex>
#include <string>

#include "bulk_copy.cuh"
#include "dense_fp64.hpp"

namespace aslam {
namespace dense {

namespace {

constexpr int TBM = 128, TBN = 128;   // CTA tile
constexpr int TBK = 32;               // floats per k-block: 128 bytes = one swizzle row
constexpr int UMMA_K = 8;             // K of one kind::tf32 instruction
constexpr uint32_t TMEM_COLS = 128;   // FP32 accumulator: 128 lanes x 128 columns
constexpr size_t OP_BYTES = (size_t)TBM * TBK * sizeof(float);  // one operand tile of one k-block: 16 KB

struct Tf32Params {
    double* C;
    size_t ldc;
    int M, N, KB;  // KB = k-blocks of TBK floats
    double alpha, beta, r1;
    const double* u;
    const double* v;
    const double* diag;
    const double* r1_dev;
    int lower_only;
    int* err;  // device flag: a bounded wait expired
};

__device__ __forceinline__ void tma_load_2d_f32(void* dst_smem, const CUtensorMap* map, int c_inner, int c_outer,
                                                uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
            bulk::smem_u32(dst_smem)),
        "l"(map), "r"(c_inner), "r"(c_outer), "r"(bulk::smem_u32(bar))
        : "memory");
}

// Bounded mbarrier wait: false when the phase did not complete within ~1 s.
__device__ __forceinline__ bool mbar_wait_bounded(uint64_t* bar, uint32_t parity) {
    for (uint32_t it = 0; it < (1u << 24); ++it) {
        uint32_t done;
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(bulk::smem_u32(bar)), "r"(parity)
            : "memory");
        if (done) return true;
    }
    return false;
}

// Shared-memory matrix descriptor of a K-major operand tile stored as rows of 128 bytes with the 128-byte swizzle
// (what TMA writes with CU_TENSOR_MAP_SWIZZLE_128B and a 32-float inner box): start address, leading byte offset
// (unused for swizzled K-major, 1), stride byte offset = 1024 (one 8-row swizzle atom), descriptor version 1
// (Blackwell), layout type 2 = SWIZZLE_128B. All offsets in units of 16 bytes.
__device__ __forceinline__ uint64_t umma_desc_k_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFFu);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

// Instruction descriptor: D = F32 (bits 4-5 = 1), A and B = TF32 (bits 7-9, 10-12 = 2), both K-major (bits 15, 16 = 0),
// N >> 3 at bits 17-22, M >> 4 at bits 24-28.
__host__ __device__ constexpr uint32_t umma_idesc_tf32(int m, int n) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bulk::smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

}  // namespace

namespace {

// FP64 [rows][ld] -> hi, lo FP32 [rows][ldf] (ldf = K rounded up to TBK; the padding columns are zero)
__global__ void __launch_bounds__(256) split_tf32_kernel(const double* __restrict__ src, size_t ld, int rows, int K,
                                                        float* __restrict__ hi, float* __restrict__ lo, size_t ldf) {
    const size_t total = (size_t)rows * ldf;
    for (size_t idx = (size_t)blockIdx.x * 256 + threadIdx.x; idx < total; idx += (size_t)gridDim.x * 256) {
        const size_t r = idx / ldf, k = idx % ldf;
        float h = 0.0f, l = 0.0f;
        if (k < (size_t)K) {
            const double x = src[r * ld + k];
            uint32_t hb, lb;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hb) : "f"((float)x));
            h = __uint_as_float(hb);
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lb) : "f"((float)(x - (double)h)));
            l = __uint_as_float(lb);
        }
        hi[idx] = h;
        if (lo) lo[idx] = l;
    }
}

template <int PASSES>
struct Tf32Geo {
    static constexpr int NOPS = PASSES == 1 ? 2 : 4;               // operand tiles per stage: A_hi B_hi (A_lo B_lo)
    static constexpr size_t STAGE = NOPS * OP_BYTES;
    static constexpr int NSTG = PASSES == 1 ? 6 : 3;
    static constexpr size_t SMEM = NSTG * STAGE + (2 * NSTG + 1) * sizeof(uint64_t) + 16 + 1024;  // + alignment slack
};

template <int PASSES>
__global__ void __launch_bounds__(192, 1)
tf32_gemm_nt_kernel(const __grid_constant__ CUtensorMap tmAhi, const __grid_constant__ CUtensorMap tmAlo,
                    const __grid_constant__ CUtensorMap tmBhi, const __grid_constant__ CUtensorMap tmBlo,
                    const Tf32Params p) {
    using G = Tf32Geo<PASSES>;
    extern __shared__ unsigned char smem_dyn[];
    // SWIZZLE_128B tiles need 1024-byte alignment
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + G::NSTG * G::STAGE);
    uint64_t* empty = full + G::NSTG;
    uint64_t* accum = empty + G::NSTG;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row0 = blockIdx.y * TBM, col0 = blockIdx.x * TBN;
    const bool skip = p.lower_only && row0 + TBM - 1 < col0;  // tile entirely above the diagonal (CTA-uniform)
    if (skip) return;

    if (threadIdx.x == 0) {
        for (int s = 0; s < G::NSTG; ++s) {
            bulk::mbar_init(&full[s], 1);
            bulk::mbar_init(&empty[s], 1);
        }
        bulk::mbar_init(accum, 1);
        bulk::fence_barrier_init();
    }
    if (warp == 1) {  // TMEM allocation: one warp, the same warp frees it at the end
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(bulk::smem_u32(tmem_slot)),
                     "r"(TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {  // ---------------- TMA producer
        if (lane == 0) {
            for (int kb = 0; kb < p.KB; ++kb) {
                const int st = kb % G::NSTG;
                if (kb >= G::NSTG && !mbar_wait_bounded(&empty[st], ((kb / G::NSTG) - 1) & 1)) {
                    atomicExch(p.err, 1);
                    break;
                }
                bulk::mbar_expect_tx(&full[st], (uint32_t)G::STAGE);
                unsigned char* s0 = smem + st * G::STAGE;
                tma_load_2d_f32(s0, &tmAhi, kb * TBK, row0, &full[st]);
                tma_load_2d_f32(s0 + OP_BYTES, &tmBhi, kb * TBK, col0, &full[st]);
                if constexpr (PASSES == 3) {
                    tma_load_2d_f32(s0 + 2 * OP_BYTES, &tmAlo, kb * TBK, row0, &full[st]);
                    tma_load_2d_f32(s0 + 3 * OP_BYTES, &tmBlo, kb * TBK, col0, &full[st]);
                }
            }
        }
    } else if (warp == 1) {  // ---------------- MMA issuer
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_tf32(TBM, TBN);
            bool ok = true;
            for (int kb = 0; kb < p.KB && ok; ++kb) {
                const int st = kb % G::NSTG;
                if (!mbar_wait_bounded(&full[st], (kb / G::NSTG) & 1)) {
                    atomicExch(p.err, 2);
                    ok = false;
                    break;
                }
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t a_hi = bulk::smem_u32(smem + st * G::STAGE);
                const uint32_t b_hi = a_hi + (uint32_t)OP_BYTES;
#pragma unroll
                for (int k4 = 0; k4 < TBK / UMMA_K; ++k4) {
                    const uint32_t off = k4 * UMMA_K * (uint32_t)sizeof(float);  // 32 bytes along K inside the swizzle row
                    umma_tf32(tmem_base, umma_desc_k_sw128(a_hi + off), umma_desc_k_sw128(b_hi + off), idesc,
                              (kb > 0 || k4 > 0) ? 1u : 0u);
                    if constexpr (PASSES == 3) {
                        const uint32_t a_lo = a_hi + 2 * (uint32_t)OP_BYTES, b_lo = a_hi + 3 * (uint32_t)OP_BYTES;
                        umma_tf32(tmem_base, umma_desc_k_sw128(a_hi + off), umma_desc_k_sw128(b_lo + off), idesc, 1u);
                        umma_tf32(tmem_base, umma_desc_k_sw128(a_lo + off), umma_desc_k_sw128(b_hi + off), idesc, 1u);
                    }
                }
                umma_commit(&empty[st]);  // the stage is free once these MMAs have read it
            }
            umma_commit(accum);  // accumulator complete
        }
    } else {  // ---------------- epilogue warps 2..5: TMEM lanes (warp % 4) * 32 .. + 31
        const int quarter = warp & 3;
        const int r = row0 + quarter * 32 + lane;
        const bool ok = mbar_wait_bounded(accum, 0);
        if (!ok && lane == 0) atomicExch(p.err, 3);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const double r1 = p.r1_dev ? *p.r1_dev : p.r1;
        const double ur = (p.u && r < p.M) ? r1 * p.u[r] : 0.0;
#pragma unroll 1
        for (int c0 = 0; c0 < TBN; c0 += 32) {
            uint32_t acc[32];
            tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)c0, acc);  // (all lanes: .aligned)
            if (r < p.M && ok) {
                double* crow = p.C + (size_t)r * p.ldc;
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int c = col0 + c0 + j;
                    if (c < p.N && !(p.lower_only && c > r)) {
                        double o = p.alpha * (double)__uint_as_float(acc[j]);
                        if (p.beta != 0.0) o = fma(p.beta, crow[c], o);
                        if (p.u) o = fma(ur, p.v[c], o);
                        if (p.diag && c == r) o += p.diag[r];
                        crow[c] = o;
                    }
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

// Tensor-pipe peak probe: every CTA issues `iters` x 4 kind::tf32 MMAs (M 128, N 256, K 8) on operands that sit in
// shared memory -- no loads in the loop -- and waits for the last commit.
__global__ void __launch_bounds__(64, 1) tf32_probe_kernel(int iters, int* err) {
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 3 * OP_BYTES);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (size_t i = threadIdx.x; i < 3 * OP_BYTES / 4; i += 64) reinterpret_cast<float*>(smem)[i] = 0.0f;
    if (threadIdx.x == 0) {
        bulk::mbar_init(bar, 1);
        bulk::fence_barrier_init();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(bulk::smem_u32(tmem_slot)),
                     "r"(256u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;
    if (warp == 1 && lane == 0) {
        constexpr uint32_t idesc = umma_idesc_tf32(128, 256);
        const uint32_t a = bulk::smem_u32(smem), b = a + (uint32_t)OP_BYTES;  // B: 256 rows = two operand tiles
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4)
                umma_tf32(tmem_base, umma_desc_k_sw128(a + k4 * 32), umma_desc_k_sw128(b + k4 * 32), idesc, 1u);
        }
        umma_commit(bar);
        if (!mbar_wait_bounded(bar, 0)) atomicExch(err, 4);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256u) : "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_fn32() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

// rows x kcols window of an FP32 matrix with pitch ldf (both multiples of TBK), box = 128 rows x 32 floats, 128-byte
// swizzle, zero fill out of range
bool make_map_f32(CUtensorMap* m, const float* base, int rows, size_t ldf, size_t kcols) {
    EncodeTiledFn fn = encode_fn32();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled is not available from the driver");
        return false;
    }
    const cuuint64_t gdim[2] = {(cuuint64_t)kcols, (cuuint64_t)rows};
    const cuuint64_t gstride[1] = {(cuuint64_t)ldf * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)TBK, (cuuint32_t)TBM};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled (FP32, 128B swizzle) failed (" + std::to_string((int)r) + ")");
        return false;
    }
    return true;
}

}  // namespace

size_t tf32_padded_k(int K) { return ((size_t)K + TBK - 1) / TBK * TBK; }

aslam_status split_tf32(cudaStream_t st, const double* src, size_t ld, int rows, int K, float* hi, float* lo) {
    const size_t ldf = tf32_padded_k(K);
    const size_t total = (size_t)rows * ldf;
    const unsigned grid = (unsigned)std::min<size_t>((total + 255) / 256, 148 * 16);
    split_tf32_kernel<<<grid, 256, 0, st>>>(src, ld, rows, K, hi, lo, ldf);
    ASLAM_CUDA(cudaGetLastError());
    return ASLAM_OK;
}

// The accumulator in TMEM is FP32 and the tensor core's accumulation is what limits the 3-pass form once K is long
// (measured: 5e-7 at K = 32, 3e-5 at K = 4107). The 3-pass product is therefore cut into slabs of TF32X3_KCHUNK columns
// that are added up in FP64 by the epilogue (beta = 1 after the first slab): one launch per slab.
constexpr int TF32X3_KCHUNK = 512;

aslam_status tf32_gemm_nt(cudaStream_t st, int M, int N, int K, const float* Ahi, const float* Alo, const float* Bhi,
                          const float* Blo, double* C, size_t ldc, const Epilogue& e, int passes, int* d_err,
                          uint64_t* launches) {
    if (passes != 1 && passes != 3) return ASLAM_ERR_ARG;
    const size_t ldf = tf32_padded_k(K);
    const int kchunk = passes == 3 ? TF32X3_KCHUNK : (int)ldf;
    dim3 grid((N + TBN - 1) / TBN, (M + TBM - 1) / TBM);
    for (size_t k0 = 0; k0 < ldf; k0 += kchunk) {
        const size_t kw = std::min<size_t>(kchunk, ldf - k0);
        // the maps describe the slab [k0, k0 + kw) of each operand (same pitch)
        CUtensorMap ta, tal, tb, tbl;
        if (!make_map_f32(&ta, Ahi + k0, M, ldf, kw) || !make_map_f32(&tb, Bhi + k0, N, ldf, kw)) return ASLAM_ERR_CUDA;
        if (!make_map_f32(&tal, (passes == 3 ? Alo : Ahi) + k0, M, ldf, kw) ||
            !make_map_f32(&tbl, (passes == 3 ? Blo : Bhi) + k0, N, ldf, kw))
            return ASLAM_ERR_CUDA;
        Tf32Params p{};
        p.C = C;
        p.ldc = ldc;
        p.M = M;
        p.N = N;
        p.KB = (int)(kw / TBK);
        p.alpha = e.alpha;
        p.lower_only = e.lower_only ? 1 : 0;
        p.err = d_err;
        if (k0 == 0) {
            p.beta = e.beta;
            p.r1 = e.r1;
            p.u = e.u;
            p.v = e.v;
            p.diag = e.diag;
            p.r1_dev = e.r1_dev;
        } else {
            p.beta = 1.0;
        }
        if (passes == 1) {
            ASLAM_DYN_SMEM(tf32_gemm_nt_kernel<1>, Tf32Geo<1>::SMEM);
            tf32_gemm_nt_kernel<1><<<grid, 192, Tf32Geo<1>::SMEM, st>>>(ta, tal, tb, tbl, p);
        } else {
            ASLAM_DYN_SMEM(tf32_gemm_nt_kernel<3>, Tf32Geo<3>::SMEM);
            tf32_gemm_nt_kernel<3><<<grid, 192, Tf32Geo<3>::SMEM, st>>>(ta, tal, tb, tbl, p);
        }
        ASLAM_CUDA(cudaGetLastError());
        if (launches) ++*launches;
    }
    return ASLAM_OK;
}

aslam_status probe_tf32_tflops(double* tflops) {
    int sms = 148;
    int dev = 0;
    ASLAM_CUDA(cudaGetDevice(&dev));
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const size_t smem = 3 * OP_BYTES + 64 + 1024;
    ASLAM_DYN_SMEM(tf32_probe_kernel, smem);
    int* d_err = nullptr;
    ASLAM_CUDA(cudaMalloc((void**)&d_err, sizeof(int)));
    ASLAM_CUDA(cudaMemset(d_err, 0, sizeof(int)));
    cudaEvent_t e0, e1;
    ASLAM_CUDA(cudaEventCreate(&e0));
    ASLAM_CUDA(cudaEventCreate(&e1));
    const int iters = 4096;
    tf32_probe_kernel<<<sms, 64, smem>>>(64, d_err);
    ASLAM_CUDA(cudaEventRecord(e0));
    tf32_probe_kernel<<<sms, 64, smem>>>(iters, d_err);
    ASLAM_CUDA(cudaEventRecord(e1));
    ASLAM_CUDA(cudaEventSynchronize(e1));
    float ms = 0.0f;
    ASLAM_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    int herr = 0;
    ASLAM_CUDA(cudaMemcpy(&herr, d_err, sizeof(int), cudaMemcpyDeviceToHost));
    cudaFree(d_err);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (herr) {
        set_error("tf32 probe: the tcgen05 pipeline timed out");
        return ASLAM_ERR_CUDA;
    }
    *tflops = (double)sms * iters * 4 * (2.0 * 128 * 256 * 8) / (ms * 1e-3) / 1e12;
    return ASLAM_OK;
}

}  // namespace dense
}  // namespace aslam
