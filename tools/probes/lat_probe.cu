// Developer probe: dependent-issue latency of DADD / DMUL / DFMA and of a shared-memory load chain on this GPU (one warp).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/tools/lat_probe tools/probes/lat_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double *out, long long *cyc, double a, double b, int n) {
    __shared__ unsigned sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = (i * 37 + 11) & 1023;
    __syncthreads();
    double x = a + threadIdx.x;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < n; ++i) x = x + b;
    long long t1 = clock64();
#pragma unroll 16
    for (int i = 0; i < n; ++i) x = x * b;
    long long t2 = clock64();
#pragma unroll 16
    for (int i = 0; i < n; ++i) x = fma(x, b, a);
    long long t3 = clock64();
    unsigned p = threadIdx.x;
#pragma unroll 16
    for (int i = 0; i < n; ++i) p = sm[p];
    long long t4 = clock64();
    float f = (float)a;
#pragma unroll 16
    for (int i = 0; i < n; ++i) f = fmaf(f, (float)b, 1.0f);
    long long t5 = clock64();
    out[threadIdx.x] = x + p + f;
    if (threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; cyc[3] = t4 - t3; cyc[4] = t5 - t4; }
}
int main() {
    double *o; long long *c, h[5];
    cudaMalloc(&o, 8 * 1024); cudaMalloc(&c, 64);
    const int n = 4096;
    for (int threads : {32, 128, 512}) {
        k<<<1, threads>>>(o, c, 1.0, 1.0000001, n);
        cudaMemcpy(h, c, 40, cudaMemcpyDeviceToHost);
        printf("threads %4d: DADD %.1f  DMUL %.1f  DFMA %.1f  LDS chain %.1f  FFMA %.1f cycles per dependent op\n", threads, (double)h[0] / n, (double)h[1] / n, (double)h[2] / n, (double)h[3] / n, (double)h[4] / n);
    }
    return 0;
}
