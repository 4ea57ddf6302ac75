// sm_100a kernels of the junction-tree hot path. One persistent kernel runs whole test cases:
//   evidence -> Collect (deepest level first) -> Distribute -> per-variable marginals -> normalise + argmax.
// Thread mapping: lane = test case inside a 32-case tile, warp = one column of one table. All table indices, evidence
// masks (32-bit lambda words, one bit per case) and loop bounds are warp-uniform; all scratch accesses are coalesced
// 256-byte rows and lane-private, so no shuffles or atomics are needed on the data path.
#include <cstdint>
#include <cuda_runtime.h>

#include "ops.cuh"
#include "plan.h"

namespace fbn {

#ifndef FBN_FUSED_STREAM
#define FBN_FUSED_STREAM 1
#endif
constexpr bool kFusedStream = FBN_FUSED_STREAM != 0;
// tuning experiments: ops as separate functions (own register allocation) instead of inlined into the phase loop
#ifdef FBN_OPS_NOINLINE
#define FBN_OP_INLINE __noinline__
#else
#define FBN_OP_INLINE __forceinline__
#endif

struct Smem {
    uint32_t *lam;    // [sum_dims]  bit l set <=> state allowed for case l of the tile
    uint32_t *obs;    // [n_vars]    bit l set <=> variable observed in case l
    uint32_t *mask;   // [max_mask_words] per-phase evidence mask tables: one word per clique row / column
    double *normacc;  // TEAM == 1: [max_normslots][32]; TEAM > 1: unused (norm shares live in scratch rows, normpart_row0)
    double *part;     // [max_pslots][32]
};

// COLLECT op: all units (column block x row chunk) of one clique that this warp owns.
template <int TEAM, int kColBlock>
__device__ FBN_OP_INLINE void collect_op(const DevPlan &P, const CollectOp &__restrict__ op, double *S, const Smem &sm, int lane, int warp) {
    double scale = 1.0;
    const int32_t *__restrict__ invrows = P.idx32 + op.invrow_off;
    for (int k = 0; k < op.nchild; ++k) scale *= *row_ptr(S, __ldg(invrows + k));
    const int units = (op.cols + kColBlock - 1) / kColBlock * op.chunks;
    const bool tables = op.rowmask_off >= 0;
    double nacc = 0.0;
    for (int u = (warp - op.rot) & (TEAM - 1); u < units; u += TEAM) {
        int j0, nc, ck, r0, r1;
        collect_unit<kColBlock>(P, op, u, j0, nc, ck, r0, r1);
        uint32_t cm[kColBlock];
        double acc[kColBlock];
#pragma unroll
        for (int c = 0; c < kColBlock; ++c) acc[c] = 0.0;
        if (tables) {
#pragma unroll
            for (int c = 0; c < kColBlock; ++c) cm[c] = sm.mask[op.colmask_off + min(j0 + c, op.cols - 1)];
            RowMaskTable rm{sm.mask + op.rowmask_off};
            collect_block_any<kColBlock, kFusedStream>(P, op, rm, cm, S, lane, j0, nc, r0, r1, scale, acc);
        } else {
#pragma unroll
            for (int c = 0; c < kColBlock; ++c) {
                const uint16_t *__restrict__ cl16 = P.lam16 + op.collam_off + (int64_t)min(j0 + c, op.cols - 1) * op.ncolv;
                uint32_t m = 0xffffffffu;
                for (int i = 0; i < op.ncolv; ++i) m &= sm.lam[__ldg(cl16 + i)];
                cm[c] = m;
            }
            RowMaskLam rm{sm.lam, P.lam16 + op.rowlam_off, op.nrest, 0u};
            collect_block_any<kColBlock, kFusedStream>(P, op, rm, cm, S, lane, j0, nc, r0, r1, scale, acc);
        }
#pragma unroll
        for (int c = 0; c < kColBlock; ++c)
            if (c < nc) {
                if (op.chunks > 1) sm.part[(op.pslot0 + (j0 + c) * op.chunks + ck) * kLanes + lane] = acc[c];
                else *row_ptr(S, op.mup_row + j0 + c) = acc[c];
                nacc += acc[c];
            }
    }
    if (TEAM == 1) sm.normacc[op.normslot * kLanes + lane] = nacc;
    else *row_ptr(S, P.normpart_row0 + op.normslot * TEAM + warp) = nacc;
}

// PULL op: out[q] = sum_r src[base_q + rest_r] * fac[basecol_q + restcol_r]
template <int TEAM, int kColBlock>
__device__ FBN_OP_INLINE void pull_op(const DevPlan &P, const PullOp &__restrict__ op, double *S, const Smem &sm, int lane, int warp) {
    const int units = (op.nT + kColBlock - 1) / kColBlock * op.chunks;
    for (int u = (warp - op.rot) & (TEAM - 1); u < units; u += TEAM) {
        int q0, nq, ck, r0, r1;
        pull_unit<kColBlock>(op, u, q0, nq, ck, r0, r1);
        double acc[kColBlock], mup[kColBlock];
        const bool divide = op.chunks == 1 && op.kind == 1;
#pragma unroll
        for (int c = 0; c < kColBlock; ++c) {
            acc[c] = 0.0;
            // the divisor is requested before the sum is formed: one memory latency less on the unit's critical path
            mup[c] = divide ? *row_ptr(S, op.mup_row + min(q0 + c, op.nT - 1)) : 1.0;
        }
        pull_block<kColBlock, kFusedStream>(P, op, S, q0, r0, r1, acc);
#pragma unroll
        for (int c = 0; c < kColBlock; ++c)
            if (c < nq) {
                const int q = q0 + c;
                if (op.chunks > 1) {
                    sm.part[(op.pslot0 + q * op.chunks + ck) * kLanes + lane] = acc[c];
                } else if (op.kind == 1) {
                    const double m = mup[c];
                    *row_ptr(S, op.dst_row + q) = (m == 0.0) ? 0.0 : acc[c] / m;  // TableDivision 0-guard, PotentialTable.cpp:718-734
                } else {
                    *row_ptr(S, op.dst_row + q) = acc[c];
                }
            }
    }
}

template <int TEAM>
__device__ __forceinline__ void fin_item(const DevPlan &P, const FinItem &F, double *S, const Smem &sm, int lane) {
    double s = 0.0;
    if (F.type == FIN_NORM) {
        if (TEAM == 1) {
            s = sm.normacc[F.slot0 * kLanes + lane];
        } else {
            const double *np = row_ptr(S, P.normpart_row0 + F.slot0 * TEAM);
            constexpr int G = TEAM < 8 ? TEAM : 8;  // independent loads in flight; summed in warp order
#pragma unroll 1
            for (int w0 = 0; w0 < TEAM; w0 += G) {
                double v[G];
#pragma unroll
                for (int w = 0; w < G; ++w) v[w] = np[(w0 + w) * kLanes];
#pragma unroll
                for (int w = 0; w < G; ++w) s += v[w];
            }
        }
        *row_ptr(S, F.dst_row) = 1.0 / s;
        return;
    }
    for (int i = 0; i < F.nslots; ++i) s += sm.part[(F.slot0 + i) * kLanes + lane];
    if (F.type == FIN_D) {
        const double m = *row_ptr(S, F.aux_row);
        s = (m == 0.0) ? 0.0 : s / m;
    }
    *row_ptr(S, F.dst_row) = s;
}

template <int TEAM, int CB>
__global__ void __launch_bounds__(TEAM * 32, (FBN_OCC / TEAM) > 0 ? (FBN_OCC / TEAM) : 1) jt_fused_kernel(const DevPlan P, double *__restrict__ scratch, int64_t n_cases,
                                                            const int64_t *__restrict__ ev_off, int64_t ev_base,
                                                            const int32_t *__restrict__ ev_vars,
                                                            const int32_t *__restrict__ ev_vals, int32_t *__restrict__ labels,
                                                            double *__restrict__ post, int64_t n_tiles) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem sm;
    sm.lam = reinterpret_cast<uint32_t *>(smem_raw);
    sm.obs = sm.lam + P.sum_dims;
    size_t o = ((size_t)(P.sum_dims + P.n_vars) * 4 + 15) & ~(size_t)15;
    sm.mask = reinterpret_cast<uint32_t *>(smem_raw + o);
    o += ((size_t)P.max_mask_words * 4 + 15) & ~(size_t)15;
    sm.normacc = reinterpret_cast<double *>(smem_raw + o);
    sm.part = sm.normacc + (size_t)P.max_normslots * (TEAM == 1 ? kLanes : 0);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *S = scratch + (int64_t)blockIdx.x * P.tile_rows * kLanes + lane;  // lane-private column of this CTA's tile

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        // ---- evidence -> lambda bit words (LoadDiscreteEvidence, src/JunctionTree.cpp:316-383, mask form :386-457)
        if (TEAM > 1) __syncthreads(); else __syncwarp();  // previous tile's readers of lam/obs are done (lanes may diverge)
        for (int i = threadIdx.x; i < P.sum_dims + P.n_vars; i += TEAM * 32) sm.lam[i] = i < P.sum_dims ? 0xffffffffu : 0u;
        if (TEAM > 1) __syncthreads(); else __syncwarp();
        const int64_t cs = tile * kLanes + lane;
        const bool valid = cs < n_cases;
        if (valid) {
            const int64_t k1 = ev_off[cs + 1] - ev_base;
            for (int64_t k = ev_off[cs] - ev_base + warp; k < k1; k += TEAM) {
                const int v = ev_vars[k], val = ev_vals[k];
                if (v < 0 || v >= P.n_vars) continue;  // src/JunctionTree.cpp:326-331
                const VarInfo vi = P.vars[v];
                atomicOr(&sm.obs[v], 1u << lane);
                for (int d = 0; d < vi.dim; ++d)
                    if (d != val) atomicAnd(&sm.lam[vi.off + d], ~(1u << lane));
            }
        }
        if (TEAM > 1) __syncthreads(); else __syncwarp();

        for (int ph = 0; ph < P.n_phases; ++ph) {
            const Phase p = P.phases[ph];
            if (p.kind == 0) {
                // evidence mask tables of every clique of this phase: lanes work on different rows (warp-uniform data)
                for (int oi = p.op_begin; oi < p.op_end; ++oi) {
                    const CollectOp *__restrict__ op = P.cops + P.oplist[oi];
                    if (op->rowmask_off < 0) continue;
                    const int nrest = op->nrest, ncolv = op->ncolv, rows = op->rows, cols = op->cols;
                    const uint16_t *__restrict__ rl16 = P.lam16 + op->rowlam_off;
                    for (int r = threadIdx.x; r < rows; r += TEAM * 32) {
                        uint32_t m = 0xffffffffu;
                        for (int i = 0; i < nrest; ++i) m &= sm.lam[__ldg(rl16 + (int64_t)r * nrest + i)];
                        sm.mask[op->rowmask_off + r] = m;
                    }
                    const uint16_t *__restrict__ cl16 = P.lam16 + op->collam_off;
                    for (int j = threadIdx.x; j < cols; j += TEAM * 32) {
                        uint32_t m = 0xffffffffu;
                        for (int i = 0; i < ncolv; ++i) m &= sm.lam[__ldg(cl16 + (int64_t)j * ncolv + i)];
                        sm.mask[op->colmask_off + j] = m;
                    }
                }
                if (TEAM > 1) __syncthreads(); else __syncwarp();
                for (int oi = p.op_begin; oi < p.op_end; ++oi) {
                    collect_op<TEAM, CB>(P, P.cops[P.oplist[oi]], S, sm, lane, warp);
                }
            } else {
                for (int oi = p.op_begin; oi < p.op_end; ++oi) {
                    pull_op<TEAM, CB>(P, P.pops[P.oplist[oi]], S, sm, lane, warp);
                }
            }
            if (p.fin_end > p.fin_begin) {
                if (TEAM > 1) __syncthreads();
                for (int f = p.fin_begin + warp; f < p.fin_end; f += TEAM) fin_item<TEAM>(P, P.fins[f], S, sm, lane);
            }
            if (TEAM > 1) __syncthreads(); else __syncwarp();  // the next phase overwrites the mask tables / partial slots
        }

        // ---- posteriors + label (GetProbabilitiesOneNode :1392-1454, Inference::ArgMax Inference.cpp:92-102)
        for (int v = warp; v < P.n_vars; v += TEAM) {
            const VarInfo vi = P.vars[v];
            const bool observed = (sm.obs[v] >> lane) & 1u;
            const double *mg = row_ptr(S, P.mg_row0 + vi.off);
            double sum = 0.0;
            for (int d = 0; d < vi.dim; ++d) sum += mg[d * kLanes];
            double best = 0.0;
            int arg = -1;
            for (int d = 0; d < vi.dim; ++d) {
                const double pr = observed ? 0.0 : mg[d * kLanes] / sum;
                if (valid && post) post[cs * P.sum_dims + vi.off + d] = pr;
                if (pr > best) { best = pr; arg = d; }
            }
            if (v == P.query_var && valid && labels) labels[cs] = arg;
        }
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
    size_t o = ((size_t)(P.sum_dims + P.n_vars) * 4 + 15) & ~(size_t)15;
    o += ((size_t)P.max_mask_words * 4 + 15) & ~(size_t)15;
    return o + ((size_t)(P.team == 1 ? P.max_normslots : 0) + (size_t)P.max_pslots) * kLanes * sizeof(double);
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
    const int64_t n_tiles = (n_cases + kLanes - 1) / kLanes;
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
