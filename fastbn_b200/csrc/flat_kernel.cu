// jt_flat_kernel: the fused junction-tree kernel for SMALL trees (ALARM-size: one warp runs a whole 32*kW-case tile).
// Same SumOps, scratch layout and phase order as jt_fused_kernel<1,1> (kernels.cu), but the per-table index arithmetic is
// taken out of the inner loop: the planner writes one RECORD per view entry (pristine potential + the absolute scratch rows
// of the message entries it multiplies with), target-major — plan.h, SumOp::flat_off. The loop is then: load a batch of
// records (warp-uniform 16-byte loads), issue all their loads (each case's potential under its own evidence pattern, one
// 128-bit load per lane and gather for kW = 2), do the arithmetic; the end of a target's sum is a warp-uniform counter.
// Nothing but message rows is written: the entries are recomputed by every op that needs them, so a case costs its messages
// in HBM traffic, not its cliques; evidence costs no per-entry test (pattern-folded potentials, plan.h).
// Measured and rejected on B200 (round 2): a software pipeline over the batches (two register sets, the loads of batch k+1
// in flight while batch k is summed): 131-149 M ALARM cases/s at 16-32 warps per SM against 158 M without (spills).
#include <cstdint>
#include <type_traits>

#include <cuda_runtime.h>

#include "ops.cuh"
#include "plan.h"

namespace fbn {

namespace {

#ifndef FBN_FLAT_BATCH
#define FBN_FLAT_BATCH 2   // measured on B200 (ALARM, 64 registers x 32 warps per SM): batch 2 158 M cases/s, batch 4 (spills) 125 M
#endif
// The 0 guard of TableDivision is not applied by any kernel (plan.h): the down ratios of targets the evidence rules out are zeroed
// by the per-target evidence check of distribute views (finish_target). Without the guard's dependent load per target: ALARM
// 151 M -> 172 M cases/s (B200, round 2).
// records per batch: their loads are all issued before the first use. Ops with many gathers per entry take smaller batches
// (the gathered vectors of a batch live in registers: batch x gathers x 2*kW registers).
template <int NG>
struct FlatBatch {
    static constexpr int value = NG <= 2 ? FBN_FLAT_BATCH : (NG <= 4 ? (FBN_FLAT_BATCH + 1) / 2 : 1);
};

// one SumOp of a flat plan, NG = gathers per entry (0..4)
template <int NG>
__device__ __forceinline__ void flat_sum(const DevPlan &P, const SumOp &__restrict__ op, double *S, const uint8_t *pat, int lane) {
    constexpr int kBatch = FlatBatch<NG>::value;
    const V scale = op_scale(P, op, S);
    const Fx fx = op_fx(P, op, pat, lane);
    constexpr int kStride = NG > 3 ? 2 : 1;
    const int4 *__restrict__ rec = reinterpret_cast<const int4 *>(P.flatc + op.flat_off);
    const double *__restrict__ potl[kW];   // this lane's pattern slots of the potentials
#pragma unroll
    for (int i = 0; i < kW; ++i) potl[i] = P.pot + fx.x[i];
    const int rows = op.rows, cols = op.cols, n = rows * cols;
    double *dst = row_ptr(S, op.dst_row);
    int r = 0, j = 0;
    V acc = vset(0.0), nacc = vset(0.0);
    auto batch = [&](int e0, auto full) {
        constexpr bool kFull = decltype(full)::value;  // full batch: no bounds checks
        int4 A[kBatch], B[kBatch];
#pragma unroll
        for (int b = 0; b < kBatch; ++b) {
            const int e = kFull ? e0 + b : min(e0 + b, n - 1);
            A[b] = __ldg(rec + (int64_t)e * kStride);
            if (NG > 3) B[b] = __ldg(rec + (int64_t)e * kStride + 1);
        }
        double pv[kBatch][kW];
        V m[kBatch][NG > 0 ? NG : 1];
#pragma unroll
        for (int b = 0; b < kBatch; ++b) {
#pragma unroll
            for (int i = 0; i < kW; ++i) pv[b][i] = __ldg(potl[i] + A[b].x);
            if (NG > 0) m[b][0] = ldv(row_ptr(S, A[b].y));
            if (NG > 1) m[b][1] = ldv(row_ptr(S, A[b].z));
            if (NG > 2) m[b][2] = ldv(row_ptr(S, A[b].w));
            if (NG > 3) m[b][3] = ldv(row_ptr(S, B[b].x));
        }
#pragma unroll
        for (int b = 0; b < kBatch; ++b) {
            if (kFull || e0 + b < n) {
#pragma unroll
                for (int i = 0; i < kW; ++i) {
                    double v = pv[b][i];
#pragma unroll
                    for (int g = 0; g < NG; ++g) v *= m[b][g].x[i];
                    acc.x[i] += v;
                }
                if (++r == rows) {  // end of target j
                    const V a = finish_target(P, op, pat, lane, j, acc, scale);
                    stv(dst + (int64_t)j * kRowDoubles, a);
                    vadd(nacc, a);
                    acc = vset(0.0);
                    r = 0;
                    ++j;
                }
            }
        }
    };
    int e0 = 0;
    for (; e0 + kBatch <= n; e0 += kBatch) batch(e0, std::true_type());
    if (e0 < n) batch(e0, std::false_type());
    if (op.kind == SUM_COLLECT) {  // PotentialTableBase::Normalize as a scalar applied by the consumers
#pragma unroll
        for (int i = 0; i < kW; ++i) nacc.x[i] = 1.0 / nacc.x[i];
        stv(row_ptr(S, op.inv_row), nacc);
    }
}

// ops with more than four gathers per entry (cliques with more than three children: rare): plain loop, its own register
// allocation, so that the common instantiations above keep theirs
__device__ __noinline__ void flat_sum_generic(const DevPlan &P, const SumOp &__restrict__ op, double *S, const uint8_t *pat, int lane) {
    const V scale = op_scale(P, op, S);
    const Fx fx = op_fx(P, op, pat, lane);
    const int ng = op.ng, stride = ng > 3 ? 8 : 4;
    const int32_t *__restrict__ rec = P.flatc + op.flat_off;
    V nacc = vset(0.0);
    for (int j = 0; j < op.cols; ++j) {
        V acc = vset(0.0);
        for (int r = 0; r < op.rows; ++r) {
            const int32_t *q = rec + ((int64_t)j * op.rows + r) * stride;
            V v;
#pragma unroll
            for (int i = 0; i < kW; ++i) v.x[i] = __ldg(P.pot + __ldg(q) + fx.x[i]);
            for (int g = 0; g < ng; ++g) vmul(v, ldv(row_ptr(S, __ldg(q + 1 + g))));
            vadd(acc, v);
        }
        V a = finish_target(P, op, pat, lane, j, acc, scale);
        vadd(nacc, a);
        stv(row_ptr(S, op.dst_row + j), a);
    }
    if (op.kind == SUM_COLLECT) {
#pragma unroll
        for (int i = 0; i < kW; ++i) nacc.x[i] = 1.0 / nacc.x[i];
        stv(row_ptr(S, op.inv_row), nacc);
    }
}

}  // namespace

#ifndef FBN_FLAT_OCC
#define FBN_FLAT_OCC 32   // resident warps per SM the register allocation is bounded for (measured on B200, batch 2: 32 -> 158 M, 24 -> 151 M ALARM cases/s)
#endif
#ifndef FBN_FLAT_WPC
#define FBN_FLAT_WPC 1    // warps per CTA; every warp runs its own tiles (an SM holds at most 32 CTAs, so > 32 warps need WPC > 1)
#endif
constexpr int kWpc = FBN_FLAT_WPC;
__global__ void __launch_bounds__(32 * kWpc, FBN_FLAT_OCC / kWpc) jt_flat_kernel(const DevPlan P, double *__restrict__ scratch, int64_t n_cases,
                                                         const int64_t *__restrict__ ev_off, int64_t ev_base,
                                                         const int32_t *__restrict__ ev_vars, const int32_t *__restrict__ ev_vals,
                                                         int32_t *__restrict__ labels, double *__restrict__ post, int64_t n_tiles) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t warp_bytes = ((size_t)P.pat_bytes + 15) & ~(size_t)15;
    uint8_t *pat = smem_raw + warp * warp_bytes;   // [n_vars][kTileCases] evidence pattern bytes of this warp's tile
    const int64_t slot = (int64_t)blockIdx.x * kWpc + warp;  // this warp's resident tile
    double *S = scratch + slot * P.tile_rows * kRowDoubles + lane * kW;  // lane-private columns of this warp's tile

    for (int64_t tile = slot; tile < n_tiles; tile += (int64_t)gridDim.x * kWpc) {
        __syncwarp();
        for (int i = lane; i < P.pat_bytes / 4; i += 32) reinterpret_cast<uint32_t *>(pat)[i] = 0u;
        __syncwarp();
        load_patterns(P, pat, tile, n_cases, ev_off, ev_base, ev_vars, ev_vals, lane);
        __syncwarp();
        write_lambda_rows(P, S, pat, lane, 0, 1);

        for (int ph = 0; ph < P.n_phases; ++ph) {
            const Phase p = P.phases[ph];
            for (int oi = p.op_begin; oi < p.op_end; ++oi) {
                const SumOp &op = P.ops[P.oplist[oi]];
                switch (op.ng) {
                    case 0: flat_sum<0>(P, op, S, pat, lane); break;
                    case 1: flat_sum<1>(P, op, S, pat, lane); break;
                    case 2: flat_sum<2>(P, op, S, pat, lane); break;
                    case 3: flat_sum<3>(P, op, S, pat, lane); break;
                    case 4: flat_sum<4>(P, op, S, pat, lane); break;
                    default: flat_sum_generic(P, op, S, pat, lane); break;
                }
            }
        }

        for (int v = 0; v < P.n_vars; ++v) output_var(P, S, pat, v, tile, n_cases, lane, labels, post);
    }
}

// ---- launch helpers (called from capi.cu) ------------------------------------------------------------------------
size_t flat_smem_bytes(const DevPlan &P) { return ((size_t)P.pat_bytes + 15) & ~(size_t)15; }

// *tiles_per_sm = resident warps (tiles) per SM.
// The attribute is per function and per context, not per plan: it is raised to the device's opt-in maximum so that plans
// with different needs never lower it for each other (a launch still asks only for what its plan needs).
cudaError_t flat_prepare(const DevPlan &P, int smem_optin, int *tiles_per_sm) {
    const size_t smem = flat_smem_bytes(P) * kWpc;
    cudaError_t e = cudaFuncSetAttribute(jt_flat_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin);
    if (e != cudaSuccess) return e;
    int blocks = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks, jt_flat_kernel, 32 * kWpc, smem);
    *tiles_per_sm = blocks * kWpc;
    return e;
}

cudaError_t flat_launch(const DevPlan &P, int grid, double *scratch, int64_t n_cases, const int64_t *ev_off, int64_t ev_base,
                        const int32_t *ev_vars, const int32_t *ev_vals, int32_t *labels, double *post, cudaStream_t st) {
    const int64_t n_tiles = (n_cases + kTileCases - 1) / kTileCases;
    if (n_tiles == 0) return cudaSuccess;
    if (grid > n_tiles) grid = (int)n_tiles;
    grid = (grid + kWpc - 1) / kWpc;  // `grid` counts resident tiles (warps)
    jt_flat_kernel<<<grid, 32 * kWpc, flat_smem_bytes(P) * kWpc, st>>>(P, scratch, n_cases, ev_off, ev_base, ev_vars, ev_vals, labels, post, n_tiles);
    return cudaGetLastError();
}

}  // namespace fbn
