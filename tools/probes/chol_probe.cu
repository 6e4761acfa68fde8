#include <cstdio>
#include <cuda_runtime.h>
constexpr int BS = 64;
template <int DBG>
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
            if (!(DBG & 2)) __syncthreads();
            double akk = col[buf][k];
            if (!(akk > 0.0)) {
                if (tid == 0 && k < bs) atomicCAS(fail, -1, k0 + k);
                akk = akk == 0.0 ? 1.0 : fabs(akk);
            }
            const double rd = (DBG & 1) ? akk * 1e-3 : rsqrt(akk), d = akk * rd;  // a few ulp from sqrt/divide: far inside the FP64 tolerance
            // Only the lower block triangle works (tj <= ti), and only rows at or below the pivot block (ti >= kb).
            // Which of a thread's 16 elements a column touches is decided by comparing BLOCK indices (uniform over
            // the unrolled 4 x 4 loops) except inside the pivot block row / column, where the bounds are the
            // compile-time kk: no per-element runtime predicate is left in the update loops.
            if (!(DBG & 4) && tj <= ti && ti >= kb) {
                double li[4], lj[4], xk[4];
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    li[r] = col[buf][4 * ti + r] * rd;
                    lj[r] = col[buf][4 * tj + r] * rd;
                    xk[r] = xrow[buf][4 * tj + r] * rd;
                }
                if (ti > kb) {
                    if (tj > kb) {  // trailing block of A
#pragma unroll
                        for (int r = 0; r < 4; ++r)
#pragma unroll
                            for (int c = 0; c < 4; ++c) if (!(DBG & 8)) a[r][c] = fma(-li[r], lj[c], a[r][c]);
                    } else {  // tj <= kb: X columns at or left of the pivot block (row k of X is zero right of k)
#pragma unroll
                        for (int r = 0; r < 4; ++r)
#pragma unroll
                            for (int c = 0; c < 4; ++c) if (!(DBG & 16)) x[r][c] = fma(-li[r], xk[c], x[r][c]);
                        if (tj == kb) {  // pivot block column: columns right of k, then column k of L
#pragma unroll
                            for (int r = 0; r < 4; ++r) {
#pragma unroll
                                for (int c = 0; c < 4; ++c)
                                    if (c > kk) if (!(DBG & 8)) a[r][c] = fma(-li[r], lj[c], a[r][c]);
                                a[r][kk] = li[r];
                            }
                        }
                    }
                } else {  // ti == kb: rows of the pivot block, below k only
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        if (r > kk) {
#pragma unroll
                            for (int c = 0; c < 4; ++c) if (!(DBG & 16)) x[r][c] = fma(-li[r], xk[c], x[r][c]);
                        }
                    }
                    if (tj == kb) {
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            if (r > kk) {
#pragma unroll
                                for (int c = 0; c < 4; ++c)
                                    if (c > kk) if (!(DBG & 8)) a[r][c] = fma(-li[r], lj[c], a[r][c]);
                                a[r][kk] = li[r];
                            }
                        }
                        a[kk][kk] = d;
                    }
#pragma unroll
                    for (int c = 0; c < 4; ++c) x[kk][c] = xk[c];  // row k of inv(L)
                }
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


template <int DBG> float run(double* A, double* inv, int* fail) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0);
        chol_panel_kernel<DBG><<<1, 256>>>(A, 64, 64, inv, fail, 0);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    return best * 1e3f;
}
int main() {
    double *A, *inv; int* fail;
    cudaMalloc(&A, 64 * 64 * 8); cudaMalloc(&inv, 64 * 64 * 8); cudaMalloc(&fail, 4);
    double h[64 * 64];
    for (int i = 0; i < 64; ++i) for (int j = 0; j < 64; ++j) h[i * 64 + j] = (i == j ? 100.0 : 0.0) + 1.0 / (1 + i + j);
    cudaMemcpy(A, h, sizeof(h), cudaMemcpyHostToDevice);
    printf("noA %.1f noX %.1f noAX %.1f | full %.1f us | no rsqrt %.1f | no barrier %.1f | no update %.1f | none of them %.1f\n", run<8>(A, inv, fail), run<16>(A, inv, fail), run<24>(A, inv, fail), run<0>(A, inv, fail),
           run<1>(A, inv, fail), run<2>(A, inv, fail), run<4>(A, inv, fail), run<7>(A, inv, fail));
    return 0;
}
