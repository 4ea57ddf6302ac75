// fbn_score_device_static: the on-device scoring of fbn_score_device (kernels.cu: score_cases_kernel + score_reduce_kernel)
// for callers that must not allocate — code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// No counterpart in the reference.
//
// fbn_score_device keeps its per-case scratch (three doubles per case) in a cudaMalloc'ed buffer that grows with the batch.
// Here that scratch is a file-scope __device__ array of a fixed, modest size (FBN_STATIC_SCORE_MAX_CASES cases = 1.5 MiB);
// every GPU that loads the library holds it until the process exits. The entry point calls no cudaMalloc / cudaFree / allocating
// library routine, does not synchronise, and writes its result to DEVICE memory the caller passes.
//
// In host code the array's NAME is a host-side stand-in, not a device address: the address the kernels get is looked up with
// cudaGetSymbolAddress, once per device, with that device current (below). All calls share the array and it keeps the last
// call's contents: every entry a call reads is written by the same call first (nothing has to start at zero), and calls that
// overlap on the GPU (two streams, two host threads) race on it — serialise them.
#include <cstdint>
#include <mutex>
#include <string>

#include <cuda_runtime.h>

#include "../../include/fastbn_b200.h"

int fbn_set_error(int code, const std::string &msg);  // capi.cu

namespace {

constexpr int64_t kMaxCases = FBN_STATIC_SCORE_MAX_CASES;
constexpr int kMaxDevices = 64;

// per case {mse, hd, correct}
static __device__ __align__(256) double g_score_scratch[3 * kMaxCases];

// Round(number, 7), src/Inference.cpp:195-206
__device__ __forceinline__ double round7_static(double number) {
    long long ip = (long long)number;
    number -= (double)ip;
    for (int i = 0; i < 7; ++i) number *= 10;
    number = (double)(long long)(number + 0.5);
    for (int i = 0; i < 7; ++i) number /= 10;
    return (double)ip + number;
}

// One warp per case, same arithmetic and summation order as score_cases_kernel (kernels.cu): Inference::CalculateMSE /
// CalculateHellingerDistance / Accuracy, src/Inference.cpp:46-61,153-192. var_off = prefix sums of the variables' dims.
__global__ void score_cases_static_kernel(int64_t n_cases, int n_vars, const int32_t *__restrict__ var_off,
                                          const double *__restrict__ post, const double *__restrict__ golden,
                                          const int32_t *__restrict__ labels, const int32_t *__restrict__ truths,
                                          double *__restrict__ out_case) {
    const int lane = threadIdx.x & 31;
    const int64_t c = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (c >= n_cases) return;
    const int sum_dims = var_off[n_vars];
    const double *p = post + c * sum_dims, *g = golden + c * sum_dims;
    double e_mse = 0.0, e_hd = 0.0;
    int num = 0;
    for (int v = lane; v < n_vars; v += 32) {
        const int off = var_off[v], dim = var_off[v + 1] - off;
        if (g[off] > 0) {
            num += dim;
            for (int d = 0; d < dim; ++d) {
                const double a = round7_static(p[off + d]), b = g[off + d];
                e_mse += (a - b) * (a - b);
                const double h = sqrt(a) - sqrt(b);
                e_hd += h * h;
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        e_mse += __shfl_xor_sync(0xffffffffu, e_mse, o);
        e_hd += __shfl_xor_sync(0xffffffffu, e_hd, o);
        num += __shfl_xor_sync(0xffffffffu, num, o);
    }
    if (lane == 0) {
        out_case[3 * c] = sqrt(e_mse / num);
        out_case[3 * c + 1] = sqrt(e_hd / num);
        out_case[3 * c + 2] = (labels[c] == truths[c]) ? 1.0 : 0.0;
    }
}

// deterministic single-block reduction (same order as score_reduce_kernel): out4 = {n_correct, n_total, sum_mse, sum_hd}
__global__ void score_reduce_static_kernel(int64_t n_cases, const double *__restrict__ out_case, double *__restrict__ out4) {
    __shared__ double s[3][1024];
    double a = 0, b = 0, c = 0;
    for (int64_t i = threadIdx.x; i < n_cases; i += blockDim.x) {
        a += out_case[3 * i];
        b += out_case[3 * i + 1];
        c += out_case[3 * i + 2];
    }
    s[0][threadIdx.x] = a; s[1][threadIdx.x] = b; s[2][threadIdx.x] = c;
    __syncthreads();
    for (int o = blockDim.x / 2; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o)
            for (int k = 0; k < 3; ++k) s[k][threadIdx.x] += s[k][threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) { out4[0] = s[2][0]; out4[1] = (double)n_cases; out4[2] = s[0][0]; out4[3] = s[1][0]; }
}

// device address of g_score_scratch on every device that asked for it (set-up: once per device, that device current)
std::mutex g_mu;
double *g_scratch_ptr[kMaxDevices] = {};

struct DeviceRestore {  // the caller's current device is put back on every exit
    int prev = -1;
    DeviceRestore() { if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; } }
    ~DeviceRestore() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace

extern "C" int fbn_score_device_static(int32_t device, int64_t n_cases, int32_t n_vars, const int32_t *var_offsets,
                                       const double *posteriors, const int32_t *labels, const double *golden,
                                       const int32_t *truths, double *out4_device, void *stream) {
    if (n_cases < 0 || n_vars < 1 || !var_offsets || !out4_device) return fbn_set_error(FBN_ERR_INVALID, "fbn_score_device_static: bad arguments");
    if (n_cases > kMaxCases)
        return fbn_set_error(FBN_ERR_INVALID, "fbn_score_device_static: n_cases " + std::to_string(n_cases) + " exceeds the static scratch (" +
                                                  std::to_string(kMaxCases) + " cases, FBN_STATIC_SCORE_MAX_CASES): split the batch or use fbn_score_device");
    if (n_cases > 0 && (!posteriors || !labels || !golden || !truths)) return fbn_set_error(FBN_ERR_INVALID, "fbn_score_device_static: null device array");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        return fbn_set_error(FBN_ERR_NO_DEVICE, "no CUDA device: fastbn_b200 has no CPU fallback");
    }
    if (device < 0 || device >= ndev || device >= kMaxDevices) return fbn_set_error(FBN_ERR_INVALID, "fbn_score_device_static: device ordinal out of range");
    DeviceRestore restore;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fbn_set_error(FBN_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    double *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_mu);
        if (!g_scratch_ptr[device]) {  // set-up of this device: the array's device address
            double *ptr = nullptr;
            e = cudaGetSymbolAddress((void **)&ptr, g_score_scratch);
            if (e != cudaSuccess || !ptr) return fbn_set_error(FBN_ERR_CUDA, std::string("cudaGetSymbolAddress(g_score_scratch): ") + cudaGetErrorString(e));
            g_scratch_ptr[device] = ptr;
        }
        scratch = g_scratch_ptr[device];
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (n_cases > 0) {
        const int wpb = 8;
        score_cases_static_kernel<<<(unsigned)((n_cases + wpb - 1) / wpb), wpb * 32, 0, st>>>(n_cases, n_vars, var_offsets, posteriors, golden, labels,
                                                                                             truths, scratch);
    }
    score_reduce_static_kernel<<<1, 1024, 0, st>>>(n_cases, scratch, out4_device);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fbn_set_error(FBN_ERR_CUDA, std::string("fbn_score_device_static launch: ") + cudaGetErrorString(e));
    return FBN_OK;
}
