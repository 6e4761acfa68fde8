// FP64 "NT" GEMM for the large-map UKF (dense_fp64.hpp): TMA tile loads + DMMA.
//
// CTA tile BM x BN, K walked in slabs of BK = 16 doubles (128 bytes per row: one TMA box per operand per slab).
// Warp NW (the last one) is the producer: one lane waits for a free stage (empty mbarrier), arms the full
// mbarrier with the slab's byte count and issues the two cp.async.bulk.tensor.2d loads; out-of-range rows and the K
// tail are zero-filled by the TMA unit, so the main loop has no edge cases. Warps 0..NW-1 own 32 x WN sub-tiles:
// per slab each lane fetches its A and B fragments with 16-byte shared loads (two consecutive k per load: the
// contraction index is permuted identically for A and B, which a dot product does not care about) and issues
// mma.sync.m8n8k4.f64; a warp releases a stage (arrives on its empty mbarrier) once it has moved on to the next slab. No CTA-wide barrier in
// the loop. The accumulators (WM/8 x WN/8 x 2 doubles per lane) are written once, with the fused epilogue
//     C = alpha * acc + beta * C + r1 * u[i] v[j] + (i == j) diag[i].
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <mutex>

#include "bulk_copy.cuh"
#include "dense_fp64.hpp"

namespace aslam {
namespace dense {

namespace {

constexpr int BK = 16;
constexpr int NST = 4;

struct GemmParams {
    double* C;
    size_t ldc;
    int M, N, K;
    double alpha, beta, r1;
    const double* u;
    const double* v;
    const double* diag;
    const double* r1_dev;
    int lower_only;
    int kblock;  // > 0: block-diagonal second operand (Epilogue::kblock)
    // persistent form (Epilogue::max_ctas): the grid is 1-D, CTA b walks tiles b, b + gridDim.x, ... of the list
    // (row-major over the tile grid; lower_only: the tiles on or below the block diagonal, row by row)
    int persist, tiles, tcols;
};

__device__ __forceinline__ void tma_load_2d(void* dst_smem, const CUtensorMap* map, int c_inner, int c_outer,
                                            uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
            bulk::smem_u32(dst_smem)),
        "l"(map), "r"(c_inner), "r"(c_outer), "r"(bulk::smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bulk::smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

template <int BM, int BN, int WARPS_M, int WARPS_N>
struct Geo {
    static constexpr int NW = WARPS_M * WARPS_N;
    static constexpr int THREADS = (NW + 1) * 32;
    static constexpr int WM = BM / WARPS_M, WN = BN / WARPS_N;
    static constexpr int MI = WM / 8, NI = WN / 8;
    static constexpr size_t A_BYTES = (size_t)BM * BK * sizeof(double);
    static constexpr size_t B_BYTES = (size_t)BN * BK * sizeof(double);
    static constexpr size_t STAGE = A_BYTES + B_BYTES;
    static constexpr size_t SMEM = NST * STAGE + 2 * NST * sizeof(uint64_t) + 128;
    static_assert(WM % 8 == 0 && WN % 8 == 0, "warp tile must be a multiple of the 8 x 8 mma tile");
};

template <int BM, int BN, int WARPS_M, int WARPS_N>
__global__ void __launch_bounds__(Geo<BM, BN, WARPS_M, WARPS_N>::THREADS, 1)
dgemm_nt_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
    using G = Geo<BM, BN, WARPS_M, WARPS_N>;
    // TMA destinations need 128-byte alignment; the kernel has no static shared memory, so the dynamic window starts
    // at the (1024-byte aligned) base of the CTA's shared memory. No pointer arithmetic through integers here: the
    // fragment loads must stay in the shared address space (LDS), not become generic loads.
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + NST * G::STAGE);
    uint64_t* empty = full + NST;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KT = (p.K + BK - 1) / BK;
    if (!p.persist && p.lower_only && (int)blockIdx.y * BM + BM - 1 < (int)blockIdx.x * BN) return;  // above the diagonal

    if (threadIdx.x == 0) {
        for (int s = 0; s < NST; ++s) {
            bulk::mbar_init(&full[s], 1);
            bulk::mbar_init(&empty[s], G::NW);
        }
        bulk::fence_barrier_init();
    }
    __syncthreads();

    // tile list of this CTA: one tile (bm, bn) = blockIdx in the plain form; tiles first, first + step, ... in the
    // persistent one. The slab counter `it` runs on across tiles, so the mbarrier ring never restarts.
    const int first = p.persist ? (int)blockIdx.x : 0, step = p.persist ? (int)gridDim.x : 1;
    const int ntiles = p.persist ? p.tiles : 1;
    auto tile_of = [&](int id, int& bm, int& bn) {
        if (!p.persist) {
            bm = blockIdx.y;
            bn = blockIdx.x;
        } else if (p.lower_only) {  // (BM == BN) id = bm (bm + 1) / 2 + bn, bn <= bm
            bm = (int)((sqrtf(8.0f * (float)id + 1.0f) - 1.0f) * 0.5f);
            while ((bm + 1) * (bm + 2) / 2 <= id) ++bm;
            while (bm * (bm + 1) / 2 > id) --bm;
            bn = id - bm * (bm + 1) / 2;
        } else {
            bm = id / p.tcols;
            bn = id % p.tcols;
        }
    };

    if (warp == G::NW) {  // ---------------- producer
        if (lane == 0) {
            int it = 0;
            for (int id = first; id < ntiles; id += step) {
                int bm, bn;
                tile_of(id, bm, bn);
                const int row0 = bm * BM, col0 = bn * BN;
                const int ka0 = p.kblock ? (col0 / p.kblock) * p.kblock : 0;  // block-diagonal operand: A's column window
                for (int kt = 0; kt < KT; ++kt, ++it) {
                    const int st = it % NST;
                    if (it >= NST) bulk::mbar_wait(&empty[st], ((it / NST) - 1) & 1);
                    bulk::mbar_expect_tx(&full[st], (uint32_t)G::STAGE);
                    unsigned char* sa = smem + st * G::STAGE;
                    tma_load_2d(sa, &tmA, ka0 + kt * BK, row0, &full[st]);
                    tma_load_2d(sa + G::A_BYTES, &tmB, kt * BK, col0, &full[st]);
                }
            }
        }
        return;
    }

    // ---------------- consumers
    const int wm = (warp / WARPS_N) * G::WM, wn = (warp % WARPS_N) * G::WN;
    const int g = lane >> 2, t = lane & 3;
    int it = 0;
    for (int id = first; id < ntiles; id += step) {
    int bm, bn;
    tile_of(id, bm, bn);
    const int row0 = bm * BM, col0 = bn * BN;
    double acc[G::MI][G::NI][2];
#pragma unroll
    for (int mi = 0; mi < G::MI; ++mi)
#pragma unroll
        for (int ni = 0; ni < G::NI; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;

    for (int kt = 0; kt < KT; ++kt, ++it) {
        const int st = it % NST;
        bulk::mbar_wait(&full[st], (it / NST) & 1);
        // Release the PREVIOUS stage here, behind the spin loop: every mma of the previous slab was issued before the
        // loop, and an mma cannot issue until the shared loads feeding it have returned, so the stage is no longer
        // being read. (Arriving at the end of the slab's own iteration is wrong: ptxas hoists the arrive above the DMMAs
        // -- there is no register dependence -- and the TMA refill then races the loads still in flight.)
        if (it > 0 && lane == 0) mbar_arrive(&empty[(it - 1) % NST]);
        const double* As = reinterpret_cast<const double*>(smem + st * G::STAGE);
        const double* Bs = reinterpret_cast<const double*>(smem + st * G::STAGE + G::A_BYTES);
#pragma unroll
        for (int kg = 0; kg < BK / 8; ++kg) {
            double2 a[G::MI], b[G::NI];
#pragma unroll
            for (int mi = 0; mi < G::MI; ++mi)
                a[mi] = *reinterpret_cast<const double2*>(As + (wm + mi * 8 + g) * BK + kg * 8 + 2 * t);
#pragma unroll
            for (int ni = 0; ni < G::NI; ++ni)
                b[ni] = *reinterpret_cast<const double2*>(Bs + (wn + ni * 8 + g) * BK + kg * 8 + 2 * t);
#pragma unroll
            for (int mi = 0; mi < G::MI; ++mi)
#pragma unroll
                for (int ni = 0; ni < G::NI; ++ni) {
                    dmma(acc[mi][ni][0], acc[mi][ni][1], a[mi].x, b[ni].x);
                    dmma(acc[mi][ni][0], acc[mi][ni][1], a[mi].y, b[ni].y);
                }
        }
    }

    // ---------------- epilogue
    const bool has_beta = p.beta != 0.0;
    const double r1 = p.r1_dev ? *p.r1_dev : p.r1;
#pragma unroll
    for (int mi = 0; mi < G::MI; ++mi) {
        const int r = row0 + wm + mi * 8 + g;
        if (r >= p.M) continue;
        const double ur = p.u ? r1 * p.u[r] : 0.0;
        double* crow = p.C + (size_t)r * p.ldc;
#pragma unroll
        for (int ni = 0; ni < G::NI; ++ni) {
            const int c = col0 + wn + ni * 8 + 2 * t;
            if (c >= p.N) continue;
            const bool two = c + 1 < p.N;
            double o0 = p.alpha * acc[mi][ni][0], o1 = p.alpha * acc[mi][ni][1];
            if (has_beta) {
                if (two) {
                    const double2 old = *reinterpret_cast<const double2*>(crow + c);
                    o0 = fma(p.beta, old.x, o0);
                    o1 = fma(p.beta, old.y, o1);
                } else {
                    o0 = fma(p.beta, crow[c], o0);
                }
            }
            if (p.u) {
                o0 = fma(ur, p.v[c], o0);
                if (two) o1 = fma(ur, p.v[c + 1], o1);
            }
            if (p.diag) {
                if (c == r) o0 += p.diag[r];
                if (c + 1 == r) o1 += p.diag[r];
            }
            if (two)
                *reinterpret_cast<double2*>(crow + c) = make_double2(o0, o1);
            else
                crow[c] = o0;
        }
    }
    }  // tiles
}

// ---------------------------------------------------------------------------------- tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
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

// rows x K matrix, row pitch ld doubles, box = box_rows x BK
bool make_map(CUtensorMap* m, const double* base, int rows, int K, size_t ld, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled is not available from the driver");
        return false;
    }
    const cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    const cuuint64_t gstride[1] = {(cuuint64_t)ld * sizeof(double)};
    const cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double*>(base), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + "): rows " + std::to_string(rows) + " K " +
                  std::to_string(K) + " ld " + std::to_string(ld));
        return false;
    }
    return true;
}

template <int BM, int BN, int WARPS_M, int WARPS_N>
aslam_status launch(cudaStream_t st, const double* A, size_t lda, const double* B, size_t ldb, const GemmParams& p,
                    int a_cols = 0) {
    using G = Geo<BM, BN, WARPS_M, WARPS_N>;
    ASLAM_DYN_SMEM((dgemm_nt_kernel<BM, BN, WARPS_M, WARPS_N>), G::SMEM);
    CUtensorMap ta, tb;
    if (!make_map(&ta, A, p.M, a_cols ? a_cols : p.K, lda, BM) || !make_map(&tb, B, p.N, p.K, ldb, BN))
        return ASLAM_ERR_CUDA;
    dim3 grid((p.N + BN - 1) / BN, (p.M + BM - 1) / BM);
    if (p.persist) grid = dim3((unsigned)p.persist);
    dgemm_nt_kernel<BM, BN, WARPS_M, WARPS_N><<<grid, G::THREADS, G::SMEM, st>>>(ta, tb, p);
    ASLAM_CUDA(cudaGetLastError());
    return ASLAM_OK;
}

// Ragged edge. A last tile row (or column) holding only a few rows would still cost a full pass over K for every
// tile in it -- at cfg5 (2051 = 16 * 128 + 3) that is a whole extra wave. Those few rows / columns are computed here
// instead: one warp per index of the long side, the <= 16 short-side vectors (L2-resident) against it.
constexpr int SKINNY_MAX = 16;
__global__ void __launch_bounds__(256) dgemm_skinny_kernel(GemmParams p, const double* __restrict__ A, size_t lda,
                                                          const double* __restrict__ B, size_t ldb, int i0, int i1, int j0,
                                                          int j1, int long_is_col) {
    const int warp = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    const int nlong = long_is_col ? j1 - j0 : i1 - i0;
    const int nshort = long_is_col ? i1 - i0 : j1 - j0;
    if (warp >= nlong) return;
    const double* lv = long_is_col ? B + (size_t)(j0 + warp) * ldb : A + (size_t)(i0 + warp) * lda;
    const double* sv = long_is_col ? A + (size_t)i0 * lda : B + (size_t)j0 * ldb;
    const size_t sld = long_is_col ? lda : ldb;
    double acc[SKINNY_MAX];
#pragma unroll
    for (int s = 0; s < SKINNY_MAX; ++s) acc[s] = 0.0;
    // 4 k per lane and trip: independent loads in flight (the short vectors come from L2)
    for (int k0 = 0; k0 < p.K; k0 += 128) {
        double x[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int k = k0 + q * 32 + lane;
            x[q] = k < p.K ? lv[k] : 0.0;
        }
#pragma unroll
        for (int s = 0; s < SKINNY_MAX; ++s) {
            if (s < nshort) {
                const double* sp = sv + (size_t)s * sld;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int k = k0 + q * 32 + lane;
                    acc[s] = fma(x[q], k < p.K ? sp[k] : 0.0, acc[s]);
                }
            }
        }
    }
    const double r1 = p.r1_dev ? *p.r1_dev : p.r1;
#pragma unroll
    for (int s = 0; s < SKINNY_MAX; ++s) {
        if (s >= nshort) break;
        double v = acc[s];
        for (int off = 16; off; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
        if (lane == 0) {
            const int i = long_is_col ? i0 + s : i0 + warp;
            const int j = long_is_col ? j0 + warp : j0 + s;
            if (p.lower_only && j > i) continue;
            double* c = p.C + (size_t)i * p.ldc + j;
            double o = p.alpha * v;
            if (p.beta != 0.0) o = fma(p.beta, *c, o);
            if (p.u) o = fma(r1 * p.u[i], p.v[j], o);
            if (p.diag && i == j) o += p.diag[i];
            *c = o;
        }
    }
}

}  // namespace

aslam_status side_stream(int slot, SideStream** out) {
    constexpr int MAXDEV = 64, SLOTS = 3;
    static thread_local SideStream ctx[SLOTS][MAXDEV];
    int dev = 0;
    ASLAM_CUDA(cudaGetDevice(&dev));
    if (slot < 0 || slot >= SLOTS || dev < 0 || dev >= MAXDEV) return ASLAM_ERR_ARG;
    SideStream& c = ctx[slot][dev];
    if (!c.s) {
        ASLAM_CUDA(cudaStreamCreateWithFlags(&c.s, cudaStreamNonBlocking));
        ASLAM_CUDA(cudaEventCreateWithFlags(&c.fork, cudaEventDisableTiming));
        ASLAM_CUDA(cudaEventCreateWithFlags(&c.join, cudaEventDisableTiming));
    }
    *out = &c;
    return ASLAM_OK;
}

aslam_status dgemm_nt(cudaStream_t st, int M, int N, int K, const double* A, size_t lda, const double* B, size_t ldb,
                      double* C, size_t ldc, const Epilogue& e, uint64_t* launches) {
    if (M <= 0 || N <= 0 || K <= 0) return ASLAM_OK;
    if ((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B) | reinterpret_cast<uintptr_t>(C)) & 15u ||
        (lda & 1u) || (ldb & 1u) || (ldc & 1u)) {
        set_error("dgemm_nt: operands must be 16-byte aligned with even leading dimensions");
        return ASLAM_ERR_ARG;
    }
    GemmParams p;
    p.C = C; p.ldc = ldc; p.M = M; p.N = N; p.K = K;
    p.alpha = e.alpha; p.beta = e.beta; p.r1 = e.r1; p.u = e.u; p.v = e.v; p.diag = e.diag; p.r1_dev = e.r1_dev;
    p.lower_only = e.lower_only ? 1 : 0;
    p.kblock = e.kblock;
    p.persist = 0;
    p.tiles = 0;
    p.tcols = 0;
    if (e.kblock && (K != e.kblock || e.kblock % 128 != 0 || e.a_cols <= 0)) {
        set_error("dgemm_nt: block-diagonal operand needs K == kblock (a multiple of 128) and a_cols");
        return ASLAM_ERR_ARG;
    }
    // ragged edge (see dgemm_skinny_kernel); not when C aliases an operand (the in-place panel products)
    const bool alias = C == A || C == B || K < 512 || e.kblock;  // short K: the ragged tile is cheap, an extra launch is not
    int Mm = M, Nm = N;
    if (!alias && M > 128 && M % 128 != 0 && M % 128 <= SKINNY_MAX) Mm = M - M % 128;
    if (!alias && N > 128 && N % 128 != 0 && N % 128 <= SKINNY_MAX) Nm = N - N % 128;
    if (e.lower_only && Nm > Mm) Nm = Mm;  // columns beyond the main rows only matter for the tail rows
    // The ragged-edge kernels write rows / columns of C the tile kernel does not touch and are latency-bound (one warp
    // walks all of K): they run on the side stream, under the tile kernel, and join before anything that follows.
    const bool tail_rows = Mm < M, tail_cols = Nm < N && !e.lower_only;
    SideStream* ss = nullptr;
    if (tail_rows || tail_cols) {
        const aslam_status rc = side_stream(1, &ss);
        if (rc != ASLAM_OK) return rc;
        ASLAM_CUDA(cudaEventRecord(ss->fork, st));
        ASLAM_CUDA(cudaStreamWaitEvent(ss->s, ss->fork, 0));
    }
    p.M = Mm;
    p.N = Nm;
    if (launches) ++*launches;
    // tile shape: 128 x 128 when that fills the 148 SMs in whole waves about as well as 128 x 64 would, narrow N -> 128 x 64
    const long tr = (Mm + 127) / 128, tc128 = (Nm + 127) / 128, tc64 = (Nm + 63) / 64;
    long t128 = 0, t64 = 0;
    for (long i = 0; i < tr; ++i) {
        const long lim128 = e.lower_only ? std::min(tc128, i + 1) : tc128;            // tiles with col0 <= last row
        const long lim64 = e.lower_only ? std::min(tc64, 2 * (i + 1)) : tc64;
        t128 += lim128;
        t64 += lim64;
    }
    const long sms = 148;
    const double cost128 = 2.0 * (double)((t128 + sms - 1) / sms), cost64 = 1.08 * (double)((t64 + sms - 1) / sms);
    const int ac = e.kblock ? e.a_cols : 0;
    const bool use128 = (Nm > 64 && cost128 <= cost64) || e.max_ctas > 0;
    if (e.max_ctas > 0 && !(e.lower_only && tc128 < tr)) {  // persistent: 128 x 128 tiles, row-major (lower_only: on or below the block diagonal)
        p.tcols = (int)tc128;
        p.tiles = (int)t128;
        p.persist = (int)std::min<long>(e.max_ctas, t128);
    }
    const aslam_status rc = use128 ? launch<128, 128, 4, 2>(st, A, lda, B, ldb, p, ac)
                                   : launch<128, 64, 4, 2>(st, A, lda, B, ldb, p, ac);
    p.persist = 0;
    if (rc != ASLAM_OK) return rc;
    // the tile kernel goes to the GPU first (one CTA per SM); the ragged-edge CTAs then fit beside it
    p.M = M;
    p.N = N;
    if (tail_rows) {  // tail rows, all columns
        dgemm_skinny_kernel<<<(N + 7) / 8, 256, 0, ss->s>>>(p, A, lda, B, ldb, Mm, M, 0, N, 1);
        ASLAM_CUDA(cudaGetLastError());
        if (launches) ++*launches;
    }
    if (tail_cols) {  // tail columns, main rows
        dgemm_skinny_kernel<<<(Mm + 7) / 8, 256, 0, ss->s>>>(p, A, lda, B, ldb, 0, Mm, Nm, N, 0);
        ASLAM_CUDA(cudaGetLastError());
        if (launches) ++*launches;
    }
    if (ss) ASLAM_CUDA(cudaEventRecord(ss->join, ss->s));
    if (ss) ASLAM_CUDA(cudaStreamWaitEvent(st, ss->join, 0));
    return ASLAM_OK;
}

// ---------------------------------------------------------------------------------- DMMA pipe probe
namespace {
__global__ void dmma_probe_kernel(double* out, int iters) {
    double c[8][2];
#pragma unroll
    for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = 0.0;
    const double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) dmma(c[i][0], c[i][1], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
}  // namespace

aslam_status probe_dmma_tflops(double* tflops) {
    int dev = 0, sms = 0;
    ASLAM_CUDA(cudaGetDevice(&dev));
    ASLAM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int threads = 256, blocks = sms * 4, iters = 4096;
    double* d = nullptr;
    ASLAM_CUDA(cudaMalloc(&d, (size_t)threads * blocks * sizeof(double)));
    cudaEvent_t e0, e1;
    ASLAM_CUDA(cudaEventCreate(&e0));
    ASLAM_CUDA(cudaEventCreate(&e1));
    dmma_probe_kernel<<<blocks, threads>>>(d, 64);
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        ASLAM_CUDA(cudaEventRecord(e0));
        dmma_probe_kernel<<<blocks, threads>>>(d, iters);
        ASLAM_CUDA(cudaEventRecord(e1));
        ASLAM_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        ASLAM_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    const double flops = (double)blocks * (threads / 32) * (double)iters * 8.0 * 512.0;  // 8x8x4 mma = 512 flop
    *tflops = flops / (best * 1e-3) / 1e12;
    return ASLAM_OK;
}

}  // namespace dense
}  // namespace aslam
