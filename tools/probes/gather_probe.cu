// Developer probe: what bandwidth do RANDOM 512-byte row gathers reach on this GPU? (the access pattern of the fused / level
// kernels: a warp reads one 512-byte message row per gather, rows scattered over a multi-GB scratch). Every warp reads `rows`
// pseudo-random rows of a buffer much larger than L2, `depth` independent row loads in flight per warp (registers), grid =
// SMs x occupancy. Prints GB/s per (warps per SM, depth).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/tools/gather_probe tools/probes/gather_probe.cu && build/tools/gather_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int DEPTH>
__global__ void gather_kernel(const double2 *__restrict__ buf, uint64_t n_rows, int rows_per_warp, double *out) {
    const int lane = threadIdx.x & 31;
    const uint64_t warp = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    uint64_t s = warp * 0x9E3779B97F4A7C15ull + 12345;
    double acc = 0.0;
    for (int i = 0; i < rows_per_warp; i += DEPTH) {
        double2 v[DEPTH];
#pragma unroll
        for (int d = 0; d < DEPTH; ++d) {
            s = s * 6364136223846793005ull + 1442695040888963407ull;   // warp-uniform LCG
            const uint64_t row = (s >> 20) & (n_rows - 1);   // n_rows is a power of two
            v[d] = buf[row * 32 + lane];   // LDG.E.128: 32 lanes x 16 B = one 512-byte row
        }
#pragma unroll
        for (int d = 0; d < DEPTH; ++d) acc += v[d].x * v[d].y;
    }
    if (acc == 12345.678) out[0] = acc;
}

template <int DEPTH>
static void run(const double2 *buf, uint64_t n_rows, int sms, int warps_per_sm, double *out) {
    const int threads = 256, blocks = sms * warps_per_sm / (threads / 32), rows_per_warp = 4096;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    gather_kernel<DEPTH><<<blocks, threads>>>(buf, n_rows, rows_per_warp, out);
    cudaEventRecord(a);
    gather_kernel<DEPTH><<<blocks, threads>>>(buf, n_rows, rows_per_warp, out);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    const double bytes = (double)blocks * (threads / 32) * rows_per_warp * 512.0;
    printf("warps/SM %2d depth %2d: %7.1f GB/s  (%.2f ms)\n", warps_per_sm, DEPTH, bytes / ms / 1e6, ms);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const uint64_t bytes = 8ull << 30, n_rows = bytes / 512;
    double2 *buf; double *out;
    if (cudaMalloc(&buf, bytes) != cudaSuccess || cudaMalloc(&out, 8) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    cudaMemset(buf, 0, bytes);
    printf("%s, %d SMs, random 512-byte rows over %llu MB\n", p.name, p.multiProcessorCount, (unsigned long long)(bytes >> 20));
    for (int w : {8, 16, 32, 64}) {
        run<2>(buf, n_rows, p.multiProcessorCount, w, out);
        run<8>(buf, n_rows, p.multiProcessorCount, w, out);
        run<16>(buf, n_rows, p.multiProcessorCount, w, out);
    }
    return 0;
}
