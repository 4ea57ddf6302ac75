// jt_flat_kernel: the fused junction-tree kernel for SMALL trees (ALARM-size: one warp runs a whole 32-case tile).
// Same algorithm, scratch layout and phase order as jt_fused_kernel<1,1> (kernels.cu), but the per-table index arithmetic is
// taken out of the inner loops: the planner writes one RECORD per clique entry (pristine potential + the absolute scratch
// rows of the child-message entries it multiplies with) and one per (target, rest) pair of every marginalisation (absolute
// source row, absolute factor row) -- plan.h, CollectOp::flat_off / PullOp::flat_off. The loops below are then: load 4
// records (warp-uniform 16 / 8 byte loads), issue all their data loads, do the arithmetic; table boundaries (end of a clique
// column, end of a target's sum) are warp-uniform counters. ~2.5x fewer instructions per case than the generic kernel, and
// four independent entries in flight per warp instead of one dependent chain per column.
#include <cstdint>
#include <type_traits>

#include <cuda_runtime.h>

#include "ops.cuh"
#include "plan.h"

namespace fbn {

namespace {

#ifndef FBN_FLAT_BATCH_C
#define FBN_FLAT_BATCH_C 4
#endif
#ifndef FBN_FLAT_BATCH_P
#define FBN_FLAT_BATCH_P 8   // measured on B200 (ALARM, tools/sweep_flat.sh): (4, 8) 121 M cases/s, (4, 4) 110 M, (8, 8) 107 M, (4, 16) 84 M
#endif
constexpr int kBatchC = FBN_FLAT_BATCH_C;  // COLLECT records per batch: their loads are all issued before the first use
constexpr int kBatchP = FBN_FLAT_BATCH_P;  // PULL records per batch

// COLLECT of one clique (reset + evidence mask + child messages + column sums + norm), NC = number of children (0..4)
template <int NC>
__device__ __forceinline__ void flat_collect(const DevPlan &P, const CollectOp &__restrict__ op, double *S, const uint32_t *mask, int lane) {
    double scale = 1.0;
    const int32_t *__restrict__ invrows = P.idx32 + op.invrow_off;
#pragma unroll
    for (int k = 0; k < NC; ++k) scale *= *row_ptr(S, __ldg(invrows + k));
    constexpr int kStride = NC > 2 ? 2 : 1;
    const int4 *__restrict__ rec = reinterpret_cast<const int4 *>(P.flatc + op.flat_off);
    const uint32_t *rmt = mask + op.rowmask_off, *cmt = mask + op.colmask_off;
    const uint32_t lanebit = 1u << lane;
    const int rows = op.rows, cols = op.cols, n = rows * cols;
    const int64_t cstride = (int64_t)cols * kLanes;
    double *colp = row_ptr(S, op.cl_row);
    double *mup = row_ptr(S, op.mup_row);
    double *p = colp;
    int r = 0, j = 0;
    bool con = (cmt[0] & lanebit) != 0;
    double acc = 0.0, nacc = 0.0;
    auto batch = [&](int e0, auto full) {
        constexpr bool kFull = decltype(full)::value;  // full batch: no bounds checks
        int4 A[kBatchC], B[kBatchC];
#pragma unroll
        for (int i = 0; i < kBatchC; ++i) {
            const int e = kFull ? e0 + i : min(e0 + i, n - 1);
            A[i] = __ldg(rec + (int64_t)e * kStride);
            if (NC > 2) B[i] = __ldg(rec + (int64_t)e * kStride + 1);
        }
        double m[kBatchC][NC > 0 ? NC : 1];
#pragma unroll
        for (int i = 0; i < kBatchC; ++i) {
            if (NC > 0) m[i][0] = *row_ptr(S, A[i].z);
            if (NC > 1) m[i][1] = *row_ptr(S, A[i].w);
            if (NC > 2) m[i][2] = *row_ptr(S, B[i].x);
            if (NC > 3) m[i][3] = *row_ptr(S, B[i].y);
        }
#pragma unroll
        for (int i = 0; i < kBatchC; ++i) {
            if (kFull || e0 + i < n) {
                double v = __hiloint2double(A[i].y, A[i].x) * scale;
#pragma unroll
                for (int k = 0; k < NC; ++k) v *= m[i][k];
                v = (con && (rmt[r] & lanebit)) ? v : 0.0;
                store_clique<true>(p, v);
                acc += v;
                p += cstride;
                if (++r == rows) {  // end of column j: its sum is the upward message entry j
                    mup[(int64_t)j * kLanes] = acc;
                    nacc += acc;
                    acc = 0.0;
                    r = 0;
                    ++j;
                    p = colp + (int64_t)j * kLanes;
                    con = j < cols && (cmt[j] & lanebit) != 0;
                }
            }
        }
    };
    int e0 = 0;
    for (; e0 + kBatchC <= n; e0 += kBatchC) batch(e0, std::true_type());
    if (e0 < n) batch(e0, std::false_type());
    *row_ptr(S, op.inv_row) = 1.0 / nacc;  // PotentialTableBase::Normalize as a scalar applied by the consumers
}

// PULL of one op (distribute message or marginal of one variable): dst[q] = sum_r src * fac, then the 0-guarded division
__device__ __forceinline__ void flat_pull(const DevPlan &P, const PullOp &__restrict__ op, double *S) {
    const int2 *__restrict__ rec = reinterpret_cast<const int2 *>(P.flatp + op.flat_off);
    const int nR = op.nR, nT = op.nT, n = nT * nR;
    double *dst = row_ptr(S, op.dst_row);
    int r = 0, q = 0;
    double acc = 0.0;
    auto batch = [&](int e0, auto full) {
        constexpr bool kFull = decltype(full)::value;  // full batch: no bounds checks
        int2 t[kBatchP];
#pragma unroll
        for (int i = 0; i < kBatchP; ++i) t[i] = __ldg(rec + (kFull ? e0 + i : min(e0 + i, n - 1)));
        double s[kBatchP], f[kBatchP];
#pragma unroll
        for (int i = 0; i < kBatchP; ++i) {
            s[i] = load_clique<true>(row_ptr(S, t[i].x));
            f[i] = *row_ptr(S, t[i].y);
        }
#pragma unroll
        for (int i = 0; i < kBatchP; ++i) {
            if (kFull || e0 + i < n) {
                acc += s[i] * f[i];
                if (++r == nR) {
                    dst[(int64_t)q * kLanes] = acc;
                    acc = 0.0;
                    r = 0;
                    ++q;
                }
            }
        }
    };
    int e0 = 0;
    for (; e0 + kBatchP <= n; e0 += kBatchP) batch(e0, std::true_type());
    if (e0 < n) batch(e0, std::false_type());
    if (op.kind == 1) {  // m_down = sum / m_up with the reference's 0 guard (TableDivision, PotentialTable.cpp:718-734)
        const double *mup = row_ptr(S, op.mup_row);
        for (int q0 = 0; q0 < nT; q0 += kBatchP) {
            double a[kBatchP], m[kBatchP];
#pragma unroll
            for (int i = 0; i < kBatchP; ++i) {
                const int qq = min(q0 + i, nT - 1);
                a[i] = dst[(int64_t)qq * kLanes];
                m[i] = mup[(int64_t)qq * kLanes];
            }
#pragma unroll
            for (int i = 0; i < kBatchP; ++i)
                if (q0 + i < nT) dst[(int64_t)(q0 + i) * kLanes] = (m[i] == 0.0) ? 0.0 : a[i] / m[i];
        }
    }
}

}  // namespace

#ifndef FBN_FLAT_OCC
#define FBN_FLAT_OCC 32   // resident warps per SM the register allocation is bounded for
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
    const size_t lam_bytes = ((size_t)(P.sum_dims + P.n_vars) * 4 + 15) & ~(size_t)15;
    const size_t warp_bytes = lam_bytes + (((size_t)P.max_mask_words * 4 + 15) & ~(size_t)15);
    uint32_t *lam = reinterpret_cast<uint32_t *>(smem_raw + warp * warp_bytes);  // [sum_dims] bit l: state allowed for case l; then [n_vars] observed
    uint32_t *obs = lam + P.sum_dims;
    uint32_t *mask = reinterpret_cast<uint32_t *>(smem_raw + warp * warp_bytes + lam_bytes);
    const int64_t slot = (int64_t)blockIdx.x * kWpc + warp;  // this warp's resident tile
    double *S = scratch + slot * P.tile_rows * kLanes + lane;  // lane-private column of this warp's tile

    for (int64_t tile = slot; tile < n_tiles; tile += (int64_t)gridDim.x * kWpc) {
        // ---- evidence -> lambda bit words (LoadDiscreteEvidence, src/JunctionTree.cpp:316-383, mask form :386-457)
        __syncwarp();
        for (int i = lane; i < P.sum_dims + P.n_vars; i += 32) lam[i] = i < P.sum_dims ? 0xffffffffu : 0u;
        __syncwarp();
        const int64_t cs = tile * kLanes + lane;
        const bool valid = cs < n_cases;
        if (valid) {
            const int64_t k1 = ev_off[cs + 1] - ev_base;
            for (int64_t k = ev_off[cs] - ev_base; k < k1; ++k) {
                const int v = ev_vars[k], val = ev_vals[k];
                if (v < 0 || v >= P.n_vars) continue;  // src/JunctionTree.cpp:326-331
                const VarInfo vi = P.vars[v];
                atomicOr(&obs[v], 1u << lane);
                for (int d = 0; d < vi.dim; ++d)
                    if (d != val) atomicAnd(&lam[vi.off + d], ~(1u << lane));
            }
        }
        __syncwarp();

        for (int ph = 0; ph < P.n_phases; ++ph) {
            const Phase p = P.phases[ph];
            if (p.kind == 0) {
                // evidence mask tables of the cliques of this phase: one word per clique row / column, lanes build different words
                for (int oi = p.op_begin; oi < p.op_end; ++oi) {
                    const CollectOp *__restrict__ op = P.cops + P.oplist[oi];
                    const int nrest = op->nrest, ncolv = op->ncolv, rows = op->rows, cols = op->cols;
                    const uint16_t *__restrict__ rl16 = P.lam16 + op->rowlam_off;
                    for (int r = lane; r < rows; r += 32) {
                        uint32_t m = 0xffffffffu;
                        for (int i = 0; i < nrest; ++i) m &= lam[__ldg(rl16 + (int64_t)r * nrest + i)];
                        mask[op->rowmask_off + r] = m;
                    }
                    const uint16_t *__restrict__ cl16 = P.lam16 + op->collam_off;
                    for (int j = lane; j < cols; j += 32) {
                        uint32_t m = 0xffffffffu;
                        for (int i = 0; i < ncolv; ++i) m &= lam[__ldg(cl16 + (int64_t)j * ncolv + i)];
                        mask[op->colmask_off + j] = m;
                    }
                }
                __syncwarp();
                for (int oi = p.op_begin; oi < p.op_end; ++oi) {
                    const CollectOp &op = P.cops[P.oplist[oi]];
                    switch (op.nchild) {
                        case 0: flat_collect<0>(P, op, S, mask, lane); break;
                        case 1: flat_collect<1>(P, op, S, mask, lane); break;
                        case 2: flat_collect<2>(P, op, S, mask, lane); break;
                        case 3: flat_collect<3>(P, op, S, mask, lane); break;
                        default: flat_collect<4>(P, op, S, mask, lane); break;
                    }
                }
            } else {
                for (int oi = p.op_begin; oi < p.op_end; ++oi) flat_pull(P, P.pops[P.oplist[oi]], S);
            }
            __syncwarp();  // the next phase overwrites the mask tables
        }

        // ---- posteriors + label (GetProbabilitiesOneNode :1392-1454, Inference::ArgMax Inference.cpp:92-102)
        for (int v = 0; v < P.n_vars; ++v) {
            const VarInfo vi = P.vars[v];
            const bool observed = (obs[v] >> lane) & 1u;
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

// ---- launch helpers (called from capi.cu) ------------------------------------------------------------------------
size_t flat_smem_bytes(const DevPlan &P) {
    size_t o = ((size_t)(P.sum_dims + P.n_vars) * 4 + 15) & ~(size_t)15;
    return o + (((size_t)P.max_mask_words * 4 + 15) & ~(size_t)15);
}

// *tiles_per_sm = resident warps (tiles) per SM
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
    const int64_t n_tiles = (n_cases + kLanes - 1) / kLanes;
    if (n_tiles == 0) return cudaSuccess;
    if (grid > n_tiles) grid = (int)n_tiles;
    grid = (grid + kWpc - 1) / kWpc;  // `grid` counts resident tiles (warps)
    jt_flat_kernel<<<grid, 32 * kWpc, flat_smem_bytes(P) * kWpc, st>>>(P, scratch, n_cases, ev_off, ev_base, ev_vars, ev_vals, labels, post, n_tiles);
    return cudaGetLastError();
}

}  // namespace fbn
