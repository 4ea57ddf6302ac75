// sm_100a kernels of the junction-tree hot path. One persistent kernel runs whole test cases:
//   evidence -> Collect (deepest level first) -> Distribute -> per-variable marginals -> normalise + argmax.
// Thread mapping: a lane holds kW test cases of a 32*kW-case tile, a warp = one block of targets of one SumOp (plan.h). All
// table indices and loop bounds are warp-uniform; all scratch accesses are coalesced rows of 32 x (8*kW)-byte vectors and
// lane-private, so no shuffles or atomics are needed on the data path. Clique tables are never written: every op recomputes
// the entries it sums, and evidence costs no per-entry test (pattern-folded potentials, ops.cuh).
#include <cstdint>
#include <cuda_runtime.h>

#include "ops.cuh"
#include "plan.h"

namespace fbn {

// One SumOp: all units (target block x row chunk) of the op that this warp owns.
template <int TEAM, int CB>
__device__ __forceinline__ void sum_op(const DevPlan &P, const SumOp &__restrict__ op, double *S, const uint8_t *pat, int lane, int warp,
                                       PotStage &st) {
    const V nacc = sum_units<CB, false>(P, op, S, pat, lane, (warp - op.rot) & (TEAM - 1), TEAM, st);
    // this warp's share of sum(clique): one row per (collect op of the phase, warp), summed by the phase's FIN_NORM item
    if (op.kind == SUM_COLLECT) stv(row_ptr(S, P.normpart_row0 + op.normslot * P.nb_max + warp), nacc);
}


template <int TEAM, int CB>
__global__ void __launch_bounds__(TEAM * 32, (FBN_OCC / TEAM) > 0 ? (FBN_OCC / TEAM) : 1) jt_fused_kernel(const DevPlan P, double *__restrict__ scratch, int64_t n_cases,
                                                            const int64_t *__restrict__ ev_off, int64_t ev_base,
                                                            const int32_t *__restrict__ ev_vars,
                                                            const int32_t *__restrict__ ev_vals, int32_t *__restrict__ labels,
                                                            double *__restrict__ post, int64_t n_tiles) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint8_t *pat = smem_raw;  // [n_vars][kTileCases] evidence pattern bytes of the tile
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    PotStage st = PotStage();  // potentials come in by coalesced loads here (shared-memory staging is a level-mode path, ops.cuh)
    double *S = scratch + (int64_t)blockIdx.x * P.tile_rows * kRowDoubles + lane * kW;  // lane-private columns of this CTA's tile
    auto sync = [] { if (TEAM > 1) __syncthreads(); else __syncwarp(); };

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        sync();  // previous tile's readers of the patterns are done (lanes may diverge)
        for (int i = threadIdx.x; i < P.pat_bytes / 4; i += TEAM * 32) reinterpret_cast<uint32_t *>(pat)[i] = 0u;
        sync();
        if (warp == 0) load_patterns(P, pat, tile, n_cases, ev_off, ev_base, ev_vars, ev_vals, lane);
        sync();
        write_lambda_rows(P, S, pat, lane, warp, TEAM);
        sync();

        for (int ph = 0; ph < P.n_phases; ++ph) {
            const Phase p = P.phases[ph];
            for (int oi = p.op_begin; oi < p.op_end; ++oi) sum_op<TEAM, CB>(P, P.ops[P.oplist[oi]], S, pat, lane, warp, st);
            if (p.fin_end > p.fin_begin) {
                sync();
                for (int f = p.fin_begin + warp; f < p.fin_end; f += TEAM) fin_item(P, P.fins[f], S, TEAM);
            }
            sync();  // the next phase overwrites the partial slots and reads this phase's rows
        }

        for (int v = warp; v < P.n_vars; v += TEAM) output_var(P, S, pat, v, tile, n_cases, lane, labels, post);
    }
}

// ---- scoring against golden posteriors (Inference::CalculateMSE / CalculateHellingerDistance / Round / Accuracy,
//      src/Inference.cpp:46-61,153-206). One warp per case; out_case[c] = {mse, hd}.
__device__ __forceinline__ double round7(double number) {  // Round(number, 7), src/Inference.cpp:195-206
    long long ip = (long long)number;
    number -= (double)ip;
    for (int i = 0; i < 7; ++i) number *= 10;
    number = (double)(long long)(number + 0.5);
    for (int i = 0; i < 7; ++i) number /= 10;
    return (double)ip + number;
}

__global__ void score_cases_kernel(int64_t n_cases, int n_vars, int sum_dims, const VarInfo *__restrict__ vars,
                                   const double *__restrict__ post, const double *__restrict__ golden,
                                   const int32_t *__restrict__ labels, const int32_t *__restrict__ truths,
                                   double *__restrict__ out_case) {
    const int lane = threadIdx.x & 31;
    const int64_t c = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (c >= n_cases) return;
    const double *p = post + c * sum_dims, *g = golden + c * sum_dims;
    double e_mse = 0.0, e_hd = 0.0;
    int num = 0;
    for (int v = lane; v < n_vars; v += 32) {
        const VarInfo vi = vars[v];
        if (g[vi.off] > 0) {
            num += vi.dim;
            for (int d = 0; d < vi.dim; ++d) {
                const double a = round7(p[vi.off + d]), b = g[vi.off + d];
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
        out_case[3 * c + 2] = (labels[c] == truths[c]) ? 1.0 : 0.0;  // Inference::Accuracy, src/Inference.cpp:46-61
    }
}

// deterministic single-block reduction: out4 = {n_correct, n_total, sum_mse, sum_hd}
__global__ void score_reduce_kernel(int64_t n_cases, const double *__restrict__ out_case, double *__restrict__ out4) {
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

// ---- launch helpers (called from capi.cu) ------------------------------------------------------------------------
size_t fused_smem_bytes(const DevPlan &P) { return ((size_t)P.pat_bytes + 15) & ~(size_t)15; }

template <int TEAM, int CB>
static cudaError_t prep(const DevPlan &P, int smem_optin, int *blocks_per_sm) {
    size_t smem = fused_smem_bytes(P);
    // per function and per context, not per plan: raised to the device's opt-in maximum so that two live plans of the same
    // instantiation never lower it for each other (a launch still asks only for what its plan needs)
    cudaError_t e = cudaFuncSetAttribute(jt_fused_kernel<TEAM, CB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, jt_fused_kernel<TEAM, CB>, TEAM * 32, smem);
}

// dispatch on (team, targets per unit): team 1 runs cb 1, larger teams cb 2 or 4
#define FBN_FUSED_DISPATCH(F)                                   \
    switch (P.team * 8 + P.cb) {                                \
        case 1 * 8 + 1: F(1, 1); break;                         \
        case 2 * 8 + 2: F(2, 2); break;                         \
        case 2 * 8 + 4: F(2, 4); break;                         \
        case 4 * 8 + 2: F(4, 2); break;                         \
        case 4 * 8 + 4: F(4, 4); break;                         \
        case 8 * 8 + 2: F(8, 2); break;                         \
        case 8 * 8 + 4: F(8, 4); break;                         \
        case 16 * 8 + 2: F(16, 2); break;                       \
        case 16 * 8 + 4: F(16, 4); break;                       \
        case 32 * 8 + 2: F(32, 2); break;                       \
        case 32 * 8 + 4: F(32, 4); break;                       \
        default: return cudaErrorInvalidValue;                  \
    }

cudaError_t fused_prepare(const DevPlan &P, int smem_optin, int *blocks_per_sm) {
#define FBN_PREP(T, C) return prep<T, C>(P, smem_optin, blocks_per_sm)
    FBN_FUSED_DISPATCH(FBN_PREP)
#undef FBN_PREP
    return cudaErrorInvalidValue;
}

cudaError_t fused_launch(const DevPlan &P, int grid, double *scratch, int64_t n_cases, const int64_t *ev_off, int64_t ev_base,
                         const int32_t *ev_vars, const int32_t *ev_vals, int32_t *labels, double *post, cudaStream_t st) {
    const int64_t n_tiles = (n_cases + kTileCases - 1) / kTileCases;
    if (n_tiles == 0) return cudaSuccess;
    if (grid > n_tiles) grid = (int)n_tiles;
    size_t smem = fused_smem_bytes(P);
#define FBN_LAUNCH(T, C) jt_fused_kernel<T, C><<<grid, T * 32, smem, st>>>(P, scratch, n_cases, ev_off, ev_base, ev_vars, ev_vals, labels, post, n_tiles)
    FBN_FUSED_DISPATCH(FBN_LAUNCH)
#undef FBN_LAUNCH
    return cudaGetLastError();
}

// per-case scores of n_cases cases into out_case[3*i..] (i relative to the arrays passed)
cudaError_t score_cases_launch(int64_t n_cases, int n_vars, int sum_dims, const VarInfo *vars, const double *post, const double *golden,
                               const int32_t *labels, const int32_t *truths, double *out_case, cudaStream_t st) {
    if (n_cases > 0) {
        const int wpb = 8;
        score_cases_kernel<<<(unsigned)((n_cases + wpb - 1) / wpb), wpb * 32, 0, st>>>(n_cases, n_vars, sum_dims, vars, post, golden,
                                                                                      labels, truths, out_case);
    }
    return cudaGetLastError();
}

cudaError_t score_reduce_launch(int64_t n_cases, const double *out_case, double *out4, cudaStream_t st) {
    score_reduce_kernel<<<1, 1024, 0, st>>>(n_cases, out_case, out4);
    return cudaGetLastError();
}

}  // namespace fbn
