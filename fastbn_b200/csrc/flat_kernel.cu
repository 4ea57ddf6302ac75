// jt_flat_kernel: the fused junction-tree kernel for SMALL trees (ALARM-size: one warp runs a whole 32*kW-case tile).
// Same SumOps, scratch layout and phase order as jt_fused_kernel<1,1> (kernels.cu), but the per-table index arithmetic is
// taken out of the inner loop: the planner writes one RECORD per view entry (pristine potential + the absolute scratch rows
// of the message entries it multiplies with), target-major — plan.h, SumOp::flat_off. The loop is then: load a batch of
// records (warp-uniform 16-byte loads), issue all their gathers (one 128-bit load per lane and gather for kW = 2), do the
// arithmetic; the end of a target's sum is a warp-uniform counter. Nothing but message rows is written: the entries are
// recomputed by every op that needs them, so a case costs its messages in HBM traffic, not its cliques.
#include <cstdint>
#include <type_traits>

#include <cuda_runtime.h>

#include "ops.cuh"
#include "plan.h"

namespace fbn {

namespace {

#ifndef FBN_FLAT_BATCH
#define FBN_FLAT_BATCH 4
#endif
constexpr int kBatch = FBN_FLAT_BATCH;  // records per batch: their loads are all issued before the first use

// one SumOp of a flat plan, NG = gathers per entry (0..6)
template <int NG>
__device__ __forceinline__ void flat_sum(const DevPlan &P, const SumOp &__restrict__ op, double *S, const uint32_t *mask, int lane) {
    const V scale = op_scale(P, op, S);
    constexpr int kStride = NG > 2 ? 2 : 1;
    const int4 *__restrict__ rec = reinterpret_cast<const int4 *>(P.flatc + op.flat_off);
    const uint32_t *rmt = mask + op.rowmask_off * kW, *cmt = mask + op.colmask_off * kW;
    const uint32_t lanebit = 1u << lane;
    const int rows = op.rows, cols = op.cols, n = rows * cols;
    const bool guard = op.kind == SUM_DISTRIBUTE;
    double *dst = row_ptr(S, op.dst_row);
    const double *aux = row_ptr(S, guard ? op.aux_row : op.dst_row);
    int r = 0, j = 0;
    bool con[kW];
#pragma unroll
    for (int i = 0; i < kW; ++i) con[i] = (cmt[i] & lanebit) != 0;
    V up = vset(1.0);  // the child's up message entry of the current target, requested when the target begins (0 guard)
    if (guard) up = ldv(aux);
    V acc = vset(0.0), nacc = vset(0.0);
    auto batch = [&](int e0, auto full) {
        constexpr bool kFull = decltype(full)::value;  // full batch: no bounds checks
        int4 A[kBatch], B[kBatch];
#pragma unroll
        for (int b = 0; b < kBatch; ++b) {
            const int e = kFull ? e0 + b : min(e0 + b, n - 1);
            A[b] = __ldg(rec + (int64_t)e * kStride);
            if (NG > 2) B[b] = __ldg(rec + (int64_t)e * kStride + 1);
        }
        V m[kBatch][NG > 0 ? NG : 1];
#pragma unroll
        for (int b = 0; b < kBatch; ++b) {
            if (NG > 0) m[b][0] = ldv(row_ptr(S, A[b].z));
            if (NG > 1) m[b][1] = ldv(row_ptr(S, A[b].w));
            if (NG > 2) m[b][2] = ldv(row_ptr(S, B[b].x));
            if (NG > 3) m[b][3] = ldv(row_ptr(S, B[b].y));
            if (NG > 4) m[b][4] = ldv(row_ptr(S, B[b].z));
            if (NG > 5) m[b][5] = ldv(row_ptr(S, B[b].w));
        }
#pragma unroll
        for (int b = 0; b < kBatch; ++b) {
            if (kFull || e0 + b < n) {
                const double pv = __hiloint2double(A[b].y, A[b].x);
#pragma unroll
                for (int i = 0; i < kW; ++i) {
                    double v = pv * scale.x[i];
#pragma unroll
                    for (int g = 0; g < NG; ++g) v *= m[b][g].x[i];
                    acc.x[i] += (con[i] && (rmt[r * kW + i] & lanebit)) ? v : 0.0;
                }
                if (++r == rows) {  // end of target j
                    V d = acc;
                    if (guard) {
#pragma unroll
                        for (int i = 0; i < kW; ++i) d.x[i] = (up.x[i] == 0.0) ? 0.0 : acc.x[i];  // TableDivision 0 guard, PotentialTable.cpp:718-734
                    }
                    stv(dst + (int64_t)j * kRowDoubles, d);
                    vadd(nacc, acc);
                    acc = vset(0.0);
                    r = 0;
                    ++j;
                    const int jn = min(j, cols - 1);
#pragma unroll
                    for (int i = 0; i < kW; ++i) con[i] = j < cols && (cmt[jn * kW + i] & lanebit) != 0;
                    if (guard) up = ldv(aux + (int64_t)jn * kRowDoubles);
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

}  // namespace

#ifndef FBN_FLAT_OCC
#define FBN_FLAT_OCC 16   // resident warps per SM the register allocation is bounded for
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
    const size_t lam_bytes = ((size_t)P.lam_words * 4 + 15) & ~(size_t)15;
    const size_t warp_bytes = lam_bytes + (((size_t)P.max_mask_words * kW * 4 + 15) & ~(size_t)15);
    uint32_t *lam = reinterpret_cast<uint32_t *>(smem_raw + warp * warp_bytes);
    uint32_t *mask = reinterpret_cast<uint32_t *>(smem_raw + warp * warp_bytes + lam_bytes);
    const int64_t slot = (int64_t)blockIdx.x * kWpc + warp;  // this warp's resident tile
    double *S = scratch + slot * P.tile_rows * kRowDoubles + lane * kW;  // lane-private columns of this warp's tile

    for (int64_t tile = slot; tile < n_tiles; tile += (int64_t)gridDim.x * kWpc) {
        __syncwarp();
        for (int i = lane; i < P.lam_words; i += 32) lam[i] = i < P.sum_dims * kW ? 0xffffffffu : 0u;
        __syncwarp();
        load_evidence(P, lam, tile, n_cases, ev_off, ev_base, ev_vars, ev_vals, lane, 0, 1);
        __syncwarp();

        for (int ph = 0; ph < P.n_phases; ++ph) {
            const Phase p = P.phases[ph];
            // evidence mask tables of the ops of this phase: kW words per view row / target, lanes build different entries
            for (int oi = p.op_begin; oi < p.op_end; ++oi) {
                const SumOp *__restrict__ op = P.ops + P.oplist[oi];
                const int nrest = op->nrest, ncolv = op->ncolv, rows = op->rows, cols = op->cols;
                const uint16_t *__restrict__ rl16 = P.lam16 + op->rowlam_off;
                for (int r = lane; r < rows; r += 32) {
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
                for (int j = lane; j < cols; j += 32) {
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
            __syncwarp();
            for (int oi = p.op_begin; oi < p.op_end; ++oi) {
                const SumOp &op = P.ops[P.oplist[oi]];
                switch (op.ng) {
                    case 0: flat_sum<0>(P, op, S, mask, lane); break;
                    case 1: flat_sum<1>(P, op, S, mask, lane); break;
                    case 2: flat_sum<2>(P, op, S, mask, lane); break;
                    case 3: flat_sum<3>(P, op, S, mask, lane); break;
                    case 4: flat_sum<4>(P, op, S, mask, lane); break;
                    case 5: flat_sum<5>(P, op, S, mask, lane); break;
                    default: flat_sum<6>(P, op, S, mask, lane); break;
                }
            }
            __syncwarp();  // the next phase overwrites the mask tables
        }

        for (int v = 0; v < P.n_vars; ++v) output_var(P, S, lam, v, tile, n_cases, lane, labels, post);
    }
}

// ---- launch helpers (called from capi.cu) ------------------------------------------------------------------------
size_t flat_smem_bytes(const DevPlan &P) {
    size_t o = ((size_t)P.lam_words * 4 + 15) & ~(size_t)15;
    return o + (((size_t)P.max_mask_words * kW * 4 + 15) & ~(size_t)15);
}

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
