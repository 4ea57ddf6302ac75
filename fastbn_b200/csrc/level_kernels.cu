// Level-mode kernels: the same COLLECT / PULL / FIN ops as jt_fused_kernel (kernels.cu), but ONE LAUNCH PER PHASE over all
// resident tiles with a 2-D grid (CTAs per tile, tiles). This is the path for trees whose tables are too large for one
// CTA per tile (PIGS-/MUNIN-shape): a 10^7-entry clique is split over up to nb_max CTAs of kLevelWarps warps each, columns
// (and row chunks) are dealt round-robin to the warps of all those CTAs, and partial sums meet in scratch rows:
//   split columns  -> part rows (two-stage, deterministic: partials written by the op kernel, summed by lvl_fin_kernel)
//   clique norms   -> one partial row per (clique, CTA), summed by lvl_fin_kernel
// Kernel boundaries are the only synchronisation (no inter-CTA spinning, nothing can hang). Layout and index maps are the
// planner's (plan.h): lane = test case, one 256-byte row per table entry, every access coalesced.
#include <algorithm>
#include <cstdint>
#include <vector>

#include <cuda_runtime.h>

#include "ops.cuh"
#include "plan.h"

namespace fbn {

#ifndef FBN_LVL_OCC4
#define FBN_LVL_OCC4 3   // resident CTAs per SM the cb = 4 collect kernel is register-bounded for
#endif

namespace {

// evidence -> per-tile lambda bit words in global memory (LoadDiscreteEvidence, src/JunctionTree.cpp:316-383, mask form
// :386-457). One CTA per tile; words [0, sum_dims): bit l set <=> state allowed for case l; [sum_dims, +n_vars): observed.
__global__ void __launch_bounds__(128) lvl_evidence_kernel(const DevPlan P, uint32_t *__restrict__ lamG, int64_t tile0,
                                                           int64_t n_cases, const int64_t *__restrict__ ev_off, int64_t ev_base,
                                                           const int32_t *__restrict__ ev_vars, const int32_t *__restrict__ ev_vals) {
    uint32_t *lam = lamG + (int64_t)blockIdx.x * P.lam_words;
    for (int i = threadIdx.x; i < P.lam_words; i += blockDim.x) lam[i] = i < P.sum_dims ? 0xffffffffu : 0u;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int64_t cs = (tile0 + blockIdx.x) * kLanes + lane;
    if (cs < n_cases) {
        const int64_t k1 = ev_off[cs + 1] - ev_base;
        for (int64_t k = ev_off[cs] - ev_base + warp; k < k1; k += nw) {
            const int v = ev_vars[k], val = ev_vals[k];
            if (v < 0 || v >= P.n_vars) continue;  // src/JunctionTree.cpp:326-331
            const VarInfo vi = P.vars[v];
            atomicOr(&lam[P.sum_dims + v], 1u << lane);
            for (int d = 0; d < vi.dim; ++d)
                if (d != val) atomicAnd(&lam[vi.off + d], ~(1u << lane));
        }
    }
}

// COLLECT phase: grid (CTAs per tile, tiles). Reference steps fused here: reset from backup, evidence masking, clique *=
// child messages, column sums into the upstream separator (see CollectOp in plan.h). Row masks are evaluated
// lane-parallel from the tile's lambda words (RowMaskLam, ops.cuh).
template <int kColBlock>
__global__ void __launch_bounds__(kLevelWarps * 32, kColBlock >= 8 ? 2 : FBN_LVL_OCC4) lvl_collect_kernel(const DevPlan P, double *__restrict__ scratch,
                                                                       const uint32_t *__restrict__ lamG, int phase_idx) {
    __shared__ double red[kLevelWarps][kLanes];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int gw = blockIdx.x * kLevelWarps + warp, GW = gridDim.x * kLevelWarps;
    double *S = scratch + (int64_t)blockIdx.y * P.tile_rows * kLanes + lane;
    const uint32_t *lam = lamG + (int64_t)blockIdx.y * P.lam_words;
    const Phase ph = P.phases[phase_idx];
    for (int oi = ph.op_begin; oi < ph.op_end; ++oi) {
        const CollectOp &__restrict__ op = P.cops[P.oplist[oi]];
        double scale = 1.0;
        const int32_t *__restrict__ invrows = P.idx32 + op.invrow_off;
        for (int k = 0; k < op.nchild; ++k) scale *= *row_ptr(S, __ldg(invrows + k));
        const int units = (op.cols + kColBlock - 1) / kColBlock * op.chunks;
        int start = (gw - op.rot % GW) % GW;
        if (start < 0) start += GW;
        double nacc = 0.0;
        for (int u = start; u < units; u += GW) {
            int j0, nc, ck, r0, r1;
            collect_unit<kColBlock>(P, op, u, j0, nc, ck, r0, r1);
            uint32_t cm[kColBlock];
            double acc[kColBlock];
#pragma unroll
            for (int c = 0; c < kColBlock; ++c) {
                acc[c] = 0.0;
                const uint16_t *__restrict__ cl16 = P.lam16 + op.collam_off + (int64_t)min(j0 + c, op.cols - 1) * op.ncolv;
                uint32_t m = 0xffffffffu;
                for (int i = 0; i < op.ncolv; ++i) m &= __ldg(lam + __ldg(cl16 + i));
                cm[c] = m;
            }
            RowMaskLam rm{lam, P.lam16 + op.rowlam_off, op.nrest, 0u};
            collect_block_any<kColBlock, true>(P, op, rm, cm, S, lane, j0, nc, r0, r1, scale, acc);
#pragma unroll
            for (int c = 0; c < kColBlock; ++c)
                if (c < nc) {
                    if (op.chunks > 1) *row_ptr(S, P.part_row0 + op.pslot0 + (j0 + c) * op.chunks + ck) = acc[c];
                    else *row_ptr(S, op.mup_row + j0 + c) = acc[c];
                    nacc += acc[c];
                }
        }
        // this CTA's share of sum(clique): one partial row per (clique, CTA)
        red[warp][lane] = nacc;
        __syncthreads();
        if (warp == 0) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < kLevelWarps; ++w) s += red[w][lane];
            *row_ptr(S, P.normpart_row0 + op.normslot * P.nb_max + blockIdx.x) = s;
        }
        __syncthreads();
    }
}

// PULL phase (distribute messages and per-variable marginals): out[q] = sum_r src[base_q + rest_r] * fac[basecol_q + restcol_r]
template <int kColBlock>
__global__ void __launch_bounds__(kLevelWarps * 32, kColBlock >= 8 ? 2 : 3) lvl_pull_kernel(const DevPlan P, double *__restrict__ scratch, int phase_idx) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int gw = blockIdx.x * kLevelWarps + warp, GW = gridDim.x * kLevelWarps;
    double *S = scratch + (int64_t)blockIdx.y * P.tile_rows * kLanes + lane;
    const Phase ph = P.phases[phase_idx];
    for (int oi = ph.op_begin; oi < ph.op_end; ++oi) {
        const int oidx = P.oplist[oi];
        if (oidx < 0) {  // two children of one parent from one pass (Pull2Op, plan.h)
            const Pull2Op &__restrict__ op2 = P.p2ops[-1 - oidx];
            const int units2 = op2.n12 * op2.nblk1;
            int start2 = (gw - op2.rot % GW) % GW;
            if (start2 < 0) start2 += GW;
            for (int u = start2; u < units2; u += GW) pull2_unit<kColBlock>(P, op2, S, u);
            continue;
        }
        const PullOp &__restrict__ op = P.pops[oidx];
        const int units = (op.nT + kColBlock - 1) / kColBlock * op.chunks;
        int start = (gw - op.rot % GW) % GW;
        if (start < 0) start += GW;
        for (int u = start; u < units; u += GW) {
            int q0, nq, ck, r0, r1;
            pull_unit<kColBlock>(op, u, q0, nq, ck, r0, r1);
            double acc[kColBlock];
#pragma unroll
            for (int c = 0; c < kColBlock; ++c) acc[c] = 0.0;
            pull_block<kColBlock, true>(P, op, S, q0, r0, r1, acc);
#pragma unroll
            for (int c = 0; c < kColBlock; ++c)
                if (c < nq) {
                    const int q = q0 + c;
                    if (op.chunks > 1) {
                        *row_ptr(S, P.part_row0 + op.pslot0 + q * op.chunks + ck) = acc[c];
                    } else if (op.kind == 1) {
                        const double m = *row_ptr(S, op.mup_row + q);
                        *row_ptr(S, op.dst_row + q) = (m == 0.0) ? 0.0 : acc[c] / m;  // TableDivision 0-guard, PotentialTable.cpp:718-734
                    } else {
                        *row_ptr(S, op.dst_row + q) = acc[c];
                    }
                }
        }
    }
}

// finalisers of a phase: one warp per item. nb = CTAs per tile the phase's op kernel was launched with.
__global__ void __launch_bounds__(kLevelWarps * 32) lvl_fin_kernel(const DevPlan P, double *__restrict__ scratch, int phase_idx, int nb) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *S = scratch + (int64_t)blockIdx.y * P.tile_rows * kLanes + lane;
    const Phase ph = P.phases[phase_idx];
    const int f = ph.fin_begin + blockIdx.x * kLevelWarps + warp;
    if (f >= ph.fin_end) return;
    const FinItem F = P.fins[f];
    double s = 0.0;
    if (F.type == FIN_NORM) {
        const double *np = row_ptr(S, P.normpart_row0 + F.slot0 * P.nb_max);
        for (int b = 0; b < nb; ++b) s += np[(int64_t)b * kLanes];
        *row_ptr(S, F.dst_row) = 1.0 / s;
        return;
    }
    const double *pp = row_ptr(S, P.part_row0 + F.slot0);
    for (int i = 0; i < F.nslots; ++i) s += pp[(int64_t)i * kLanes];
    if (F.type == FIN_D) {
        const double m = *row_ptr(S, F.aux_row);
        s = (m == 0.0) ? 0.0 : s / m;
    }
    *row_ptr(S, F.dst_row) = s;
}

// posteriors + label (GetProbabilitiesOneNode src/JunctionTree.cpp:1392-1454, Inference::ArgMax src/Inference.cpp:92-102)
__global__ void __launch_bounds__(kLevelWarps * 32) lvl_output_kernel(const DevPlan P, double *__restrict__ scratch,
                                                                      const uint32_t *__restrict__ lamG, int64_t tile0, int64_t n_cases,
                                                                      int32_t *__restrict__ labels, double *__restrict__ post) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int v = blockIdx.x * kLevelWarps + warp;
    if (v >= P.n_vars) return;
    double *S = scratch + (int64_t)blockIdx.y * P.tile_rows * kLanes + lane;
    const uint32_t *lam = lamG + (int64_t)blockIdx.y * P.lam_words;
    const int64_t cs = (tile0 + blockIdx.y) * kLanes + lane;
    const bool valid = cs < n_cases;
    const VarInfo vi = P.vars[v];
    const bool observed = (lam[P.sum_dims + v] >> lane) & 1u;
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

}  // namespace

// Enqueues all launches for n_cases cases (several waves of `resident` tiles). Returns the number of launches via *n_launch.
cudaError_t level_launch(const DevPlan &P, const std::vector<Phase> &phases, int resident, double *scratch, uint32_t *lamG,
                         int64_t n_cases, const int64_t *ev_off, int64_t ev_base, const int32_t *ev_vars, const int32_t *ev_vals,
                         int32_t *labels, double *post, cudaStream_t st, int64_t *n_launch) {
    if (P.cb != 4 && P.cb != 8) return cudaErrorInvalidValue;
    const int64_t n_tiles = (n_cases + kLanes - 1) / kLanes;
    int64_t launches = 0;
    for (int64_t tile0 = 0; tile0 < n_tiles; tile0 += resident) {
        const unsigned wt = (unsigned)std::min<int64_t>(resident, n_tiles - tile0);
        lvl_evidence_kernel<<<wt, 128, 0, st>>>(P, lamG, tile0, n_cases, ev_off, ev_base, ev_vars, ev_vals);
        ++launches;
        for (size_t pi = 0; pi < phases.size(); ++pi) {
            const Phase &ph = phases[pi];
            // enough CTAs per tile to give every warp ~2 units, and enough CTAs overall to fill the GPU
            int nb = (ph.n_units + 2 * kLevelWarps - 1) / (2 * kLevelWarps);
            nb = std::max(1, std::min(nb, P.nb_max));
            dim3 grid((unsigned)nb, wt);
            if (P.cb == 8) {
                if (ph.kind == 0) lvl_collect_kernel<8><<<grid, kLevelWarps * 32, 0, st>>>(P, scratch, lamG, (int)pi);
                else lvl_pull_kernel<8><<<grid, kLevelWarps * 32, 0, st>>>(P, scratch, (int)pi);
            } else {
                if (ph.kind == 0) lvl_collect_kernel<4><<<grid, kLevelWarps * 32, 0, st>>>(P, scratch, lamG, (int)pi);
                else lvl_pull_kernel<4><<<grid, kLevelWarps * 32, 0, st>>>(P, scratch, (int)pi);
            }
            ++launches;
            const int nfin = ph.fin_end - ph.fin_begin;
            if (nfin > 0) {
                dim3 fg((unsigned)((nfin + kLevelWarps - 1) / kLevelWarps), wt);
                lvl_fin_kernel<<<fg, kLevelWarps * 32, 0, st>>>(P, scratch, (int)pi, nb);
                ++launches;
            }
        }
        dim3 og((unsigned)((P.n_vars + kLevelWarps - 1) / kLevelWarps), wt);
        lvl_output_kernel<<<og, kLevelWarps * 32, 0, st>>>(P, scratch, lamG, tile0, n_cases, labels, post);
        ++launches;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    if (n_launch) *n_launch = launches;
    return cudaSuccess;
}

}  // namespace fbn
