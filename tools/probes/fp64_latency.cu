// Dependent-issue latency of FP64 instructions on the device (one warp, one chain): cycles per DFMA / DMUL / rsqrt /
// MUFU.RSQ64H-seeded Newton. Build: nvcc -gencode arch=compute_100a,code=sm_100a -o fp64_latency fp64_latency.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double* out, long long* cyc, int n) {
    double a = out[0], b = out[1], c = out[2];
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) a = fma(a, b, c);
    long long t1 = clock64();
    double r = out[3];
    for (int i = 0; i < n; ++i) r = rsqrt(r + 1.5);
    long long t2 = clock64();
    float f = (float)out[4];
    for (int i = 0; i < n; ++i) f = fmaf(f, 0.999f, 0.5f);
    long long t3 = clock64();
    double s = out[5];
    for (int i = 0; i < n; ++i) s = sqrt(s + 2.0);
    long long t4 = clock64();
    double d = out[6];
    for (int i = 0; i < n; ++i) d = 1.0 / (d + 1.25);
    long long t5 = clock64();
    if (threadIdx.x == 0) {
        cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; cyc[3] = t4 - t3; cyc[4] = t5 - t4;
    }
    out[7] = a + r + f + s + d;
}
int main() {
    double h[8] = {0.5, 0.999, 0.25, 1.0, 0.3, 2.0, 1.0, 0};
    double* d; long long* c;
    cudaMalloc(&d, 64); cudaMalloc(&c, 64);
    cudaMemcpy(d, h, 64, cudaMemcpyHostToDevice);
    const int n = 4096;
    for (int rep = 0; rep < 2; ++rep) k<<<1, 32>>>(d, c, n);
    long long hc[5];
    cudaMemcpy(hc, c, 40, cudaMemcpyDeviceToHost);
    printf("cycles per dependent op: DFMA %.1f  rsqrt(double)+DADD %.1f  FFMA %.1f  sqrt(double)+DADD %.1f  1/x(double)+DADD %.1f\n",
           hc[0] / (double)n, hc[1] / (double)n, hc[2] / (double)n, hc[3] / (double)n, hc[4] / (double)n);
    return 0;
}
