// sm_100a kernels of the junction-tree hot path. One persistent kernel runs whole test cases:
//   evidence -> Collect (deepest level first) -> Distribute -> per-variable marginals -> normalise + argmax.
// Thread mapping: a lane holds kW test cases of a 32*kW-case tile, a warp = one block of targets of one SumOp (plan.h). All
// table indices, evidence masks (32-bit lambda words, one bit per case) and loop bounds are warp-uniform; all scratch
// accesses are coalesced rows of 32 x (8*kW)-byte vectors and lane-private, so no shuffles or atomics are needed on the
// data path. Clique tables are never written: every op recomputes the entries it sums (ops.cuh).
#include <cstdint>
#include <cuda_runtime.h>

#include "ops.cuh"
#include "plan.h"

namespace fbn {

// One SumOp: all units (target block x row chunk) of the op that this warp owns.
template <int TEAM, int CB>
__device__ __forceinline__ void sum_op(const DevPlan &P, const SumOp &__restrict__ op, double *S, const uint32_t *lam, const uint32_t *mask,
                                       int lane, int warp) {
    const V scale = op_scale(P, op, S);
    const int units = (op.cols + CB - 1) / CB * op.chunks;
    const bool tables = op.rowmask_off >= 0;
    V nacc = vset(0.0);
    for (int u = (warp - op.rot) & (TEAM - 1); u < units; u += TEAM) {
        int j0, nc, ck, r0, r1;
        sum_unit<CB>(P, op, u, j0, nc, ck, r0, r1);
        M cm[CB];
        V acc[CB];
#pragma unroll
        for (int c = 0; c < CB; ++c) acc[c] = vset(0.0);
        if (tables) {
#pragma unroll
            for (int c = 0; c < CB; ++c) {
                const uint32_t *mw = mask + (op.colmask_off + min(j0 + c, op.cols - 1)) * kW;
#pragma unroll
                for (int i = 0; i < kW; ++i) cm[c].w[i] = mw[i];
            }
            RowMaskTable rm{mask + op.rowmask_off * kW};
            sum_block_any<CB>(P, op, rm, cm, S, lane, j0, r0, r1, scale, acc);
        } else {
#pragma unroll
            for (int c = 0; c < CB; ++c) cm[c] = col_mask_from_lam(P, op, lam, min(j0 + c, op.cols - 1));
            RowMaskLam rm{lam, P.lam16 + op.rowlam_off, op.nrest, M()};
            sum_block_any<CB>(P, op, rm, cm, S, lane, j0, r0, r1, scale, acc);
        }
        vadd(nacc, store_unit<CB>(P, op, S, j0, nc, ck, acc));
    }
    // this warp's share of sum(clique): one row per (collect op of the phase, warp), summed by the phase's FIN_NORM item
    if (op.kind == SUM_COLLECT) stv(row_ptr(S, P.normpart_row0 + op.normslot * P.nb_max + warp), nacc);
}

template <int TEAM, int CB>
__global__ void __launch_bounds__(TEAM * 32, (FBN_OCC / TEAM) > 0 ? (FBN_OCC / TEAM) : 1) jt_fused_kernel(const DevPlan P, double *__restrict__ scratch, int64_t n_cases,
                                                            const int64_t *__restrict__ ev_off, int64_t ev_base,
                                                            const int32_t *__restrict__ ev_vars,
                                                            const int32_t *__restrict__ ev_vals, int32_t *__restrict__ labels,
                                                            double *__restrict__ post, int64_t n_tiles) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t *lam = reinterpret_cast<uint32_t *>(smem_raw);  // [lam_words]: allowed-state words, then observed words
    uint32_t *mask = reinterpret_cast<uint32_t *>(smem_raw + (((size_t)P.lam_words * 4 + 15) & ~(size_t)15));  // per-phase mask tables

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *S = scratch + (int64_t)blockIdx.x * P.tile_rows * kRowDoubles + lane * kW;  // lane-private columns of this CTA's tile
    auto sync = [] { if (TEAM > 1) __syncthreads(); else __syncwarp(); };

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        sync();  // previous tile's readers of lam are done (lanes may diverge)
        for (int i = threadIdx.x; i < P.lam_words; i += TEAM * 32) lam[i] = i < P.sum_dims * kW ? 0xffffffffu : 0u;
        sync();
        load_evidence(P, lam, tile, n_cases, ev_off, ev_base, ev_vars, ev_vals, lane, warp, TEAM);
        sync();

        for (int ph = 0; ph < P.n_phases; ++ph) {
            const Phase p = P.phases[ph];
            // evidence mask tables of every op of this phase: threads work on different rows (warp-uniform data)
            for (int oi = p.op_begin; oi < p.op_end; ++oi) {
                const SumOp *__restrict__ op = P.ops + P.oplist[oi];
                if (op->rowmask_off < 0) continue;
                const int nrest = op->nrest, ncolv = op->ncolv, rows = op->rows, cols = op->cols;
                const uint16_t *__restrict__ rl16 = P.lam16 + op->rowlam_off;
                for (int r = threadIdx.x; r < rows; r += TEAM * 32) {
                    uint32_t m[kW];
#pragma unroll
                    for (int i = 0; i < kW; ++i) m[i] = 0xffffffffu;
                    for (int k = 0; k < nrest; ++k) {
                        const uint32_t *lw = lam + (int)__ldg(rl16 + (int64_t)r * nrest + k) * kW;
#pragma unroll
                        for (int i = 0; i < kW; ++i) m[i] &= lw[i];
                    }
#pragma unroll
                    for (int i = 0; i < kW; ++i) mask[(op->rowmask_off + r) * kW + i] = m[i];
                }
                const uint16_t *__restrict__ cl16 = P.lam16 + op->collam_off;
                for (int j = threadIdx.x; j < cols; j += TEAM * 32) {
                    uint32_t m[kW];
#pragma unroll
                    for (int i = 0; i < kW; ++i) m[i] = 0xffffffffu;
                    for (int k = 0; k < ncolv; ++k) {
                        const uint32_t *lw = lam + (int)__ldg(cl16 + (int64_t)j * ncolv + k) * kW;
#pragma unroll
                        for (int i = 0; i < kW; ++i) m[i] &= lw[i];
                    }
#pragma unroll
                    for (int i = 0; i < kW; ++i) mask[(op->colmask_off + j) * kW + i] = m[i];
                }
            }
            sync();
            for (int oi = p.op_begin; oi < p.op_end; ++oi) sum_op<TEAM, CB>(P, P.ops[P.oplist[oi]], S, lam, mask, lane, warp);
            if (p.fin_end > p.fin_begin) {
                sync();
                for (int f = p.fin_begin + warp; f < p.fin_end; f += TEAM) fin_item(P, P.fins[f], S, TEAM);
            }
            sync();  // the next phase overwrites the mask tables / partial slots and reads this phase's rows
        }

        for (int v = warp; v < P.n_vars; v += TEAM) output_var(P, S, lam, v, tile, n_cases, lane, labels, post);
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
size_t fused_smem_bytes(const DevPlan &P) {
    size_t o = ((size_t)P.lam_words * 4 + 15) & ~(size_t)15;
    return o + (((size_t)P.max_mask_words * kW * 4 + 15) & ~(size_t)15);
}

template <int TEAM>
static cudaError_t prep(const DevPlan &P, int smem_optin, int *blocks_per_sm) {
    size_t smem = fused_smem_bytes(P);
    constexpr int CB = TEAM == 1 ? 1 : 2;
    if (P.cb != CB) return cudaErrorInvalidValue;
    // per function and per context, not per plan: raised to the device's opt-in maximum so that two live plans of the same
    // instantiation never lower it for each other (a launch still asks only for what its plan needs)
    cudaError_t e = cudaFuncSetAttribute(jt_fused_kernel<TEAM, CB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, jt_fused_kernel<TEAM, CB>, TEAM * 32, smem);
}

cudaError_t fused_prepare(const DevPlan &P, int smem_optin, int *blocks_per_sm) {
    switch (P.team) {
        case 1: return prep<1>(P, smem_optin, blocks_per_sm);
        case 2: return prep<2>(P, smem_optin, blocks_per_sm);
        case 4: return prep<4>(P, smem_optin, blocks_per_sm);
        case 8: return prep<8>(P, smem_optin, blocks_per_sm);
        case 16: return prep<16>(P, smem_optin, blocks_per_sm);
        case 32: return prep<32>(P, smem_optin, blocks_per_sm);
    }
    return cudaErrorInvalidValue;
}

cudaError_t fused_launch(const DevPlan &P, int grid, double *scratch, int64_t n_cases, const int64_t *ev_off, int64_t ev_base,
                         const int32_t *ev_vars, const int32_t *ev_vals, int32_t *labels, double *post, cudaStream_t st) {
    const int64_t n_tiles = (n_cases + kTileCases - 1) / kTileCases;
    if (n_tiles == 0) return cudaSuccess;
    if (grid > n_tiles) grid = (int)n_tiles;
    size_t smem = fused_smem_bytes(P);
#define FBN_LAUNCH(T) jt_fused_kernel<T, (T == 1 ? 1 : 2)><<<grid, T * 32, smem, st>>>(P, scratch, n_cases, ev_off, ev_base, ev_vars, ev_vals, labels, post, n_tiles)
    switch (P.team) {
        case 1: FBN_LAUNCH(1); break;
        case 2: FBN_LAUNCH(2); break;
        case 4: FBN_LAUNCH(4); break;
        case 8: FBN_LAUNCH(8); break;
        case 16: FBN_LAUNCH(16); break;
        case 32: FBN_LAUNCH(32); break;
        default: return cudaErrorInvalidValue;
    }
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
