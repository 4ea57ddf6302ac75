// Level-mode kernels: the same SumOps / finalisers as jt_fused_kernel (kernels.cu), but ONE LAUNCH PER PHASE over all
// resident tiles with a 2-D grid (CTAs per tile, tiles). This is the path for trees whose tables are too large for one
// CTA per tile (MUNIN-shape): a 10^7-entry view is split over up to nb_max CTAs of kLevelWarps warps each, target blocks
// (and row chunks) are dealt round-robin to the warps of all those CTAs, and partial sums meet in scratch rows:
//   split targets  -> part rows (two-stage, deterministic: partials written by the op kernel, summed by lvl_fin_kernel)
//   clique norms   -> one share row per (collect op, CTA), summed by lvl_fin_kernel
// Kernel boundaries are the only synchronisation (no inter-CTA spinning, nothing can hang). Layout and index maps are the
// planner's (plan.h): a lane holds kW test cases, one row per message entry, every access coalesced; clique entries are
// recomputed in registers (ops.cuh), never stored.
#include <algorithm>
#include <cstdint>
#include <vector>

#include <cuda_runtime.h>

#include "ops.cuh"
#include "plan.h"

#ifndef FBN_LEVEL_TMA
#define FBN_LEVEL_TMA 1   // potentials of a unit through shared memory (cp.async.bulk + mbarrier); 0: broadcast loads like the fused kernels
#endif

namespace fbn {

#ifndef FBN_LVL_OCC
#define FBN_LVL_OCC 2   // resident CTAs (of kLevelWarps warps) per SM the op kernel is register-bounded for
#endif

namespace {

// evidence -> per-tile pattern bytes in global memory (+ lambda rows of the variables some op gathers). One CTA per tile.
__global__ void __launch_bounds__(128) lvl_evidence_kernel(const DevPlan P, double *__restrict__ scratch, uint8_t *__restrict__ patG, int64_t tile0,
                                                           int64_t n_cases, const int64_t *__restrict__ ev_off, int64_t ev_base,
                                                           const int32_t *__restrict__ ev_vars, const int32_t *__restrict__ ev_vals) {
    uint8_t *pat = patG + (int64_t)blockIdx.x * P.pat_bytes;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < P.pat_bytes / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(pat)[i] = 0u;
    __syncthreads();
    if (warp == 0) load_patterns(P, pat, tile0 + blockIdx.x, n_cases, ev_off, ev_base, ev_vars, ev_vals, lane);
    __syncthreads();
    double *S = scratch + (int64_t)blockIdx.x * P.tile_rows * kRowDoubles + lane * kW;
    write_lambda_rows(P, S, pat, lane, warp, blockDim.x >> 5);
}

// One phase (tree level) of SumOps: grid (CTAs per tile, tiles). Dynamic shared memory: two potential buffers per warp, two
// mbarriers per warp, the norm reduction rows.
constexpr uint32_t kLvlPotBuf = 4096;
constexpr size_t kLvlSmem = (size_t)kLevelWarps * (2 * kLvlPotBuf + 16) + (size_t)kLevelWarps * kRowDoubles * sizeof(double);
template <int CB>
__global__ void __launch_bounds__(kLevelWarps * 32, FBN_LVL_OCC) lvl_sum_kernel(const DevPlan P, double *__restrict__ scratch,
                                                                                const uint8_t *__restrict__ patG, int phase_idx) {
    extern __shared__ __align__(128) unsigned char lsm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint64_t *bars = reinterpret_cast<uint64_t *>(lsm + (size_t)kLevelWarps * 2 * kLvlPotBuf);
    double(*red)[kRowDoubles] = reinterpret_cast<double(*)[kRowDoubles]>(lsm + (size_t)kLevelWarps * (2 * kLvlPotBuf + 16));
    PotStage st;
    pot_stage_init(st, lsm + (size_t)warp * 2 * kLvlPotBuf, bars + 2 * warp, kLvlPotBuf, lane);
    const int gw = blockIdx.x * kLevelWarps + warp, GW = gridDim.x * kLevelWarps;
    double *S = scratch + (int64_t)blockIdx.y * P.tile_rows * kRowDoubles + lane * kW;
    const uint8_t *pat = patG + (int64_t)blockIdx.y * P.pat_bytes;
    const Phase ph = P.phases[phase_idx];
    for (int oi = ph.op_begin; oi < ph.op_end; ++oi) {
        const SumOp &__restrict__ op = P.ops[P.oplist[oi]];
        int start = (gw - op.rot % GW) % GW;
        if (start < 0) start += GW;
        const V nacc = sum_units<CB, FBN_LEVEL_TMA != 0>(P, op, S, pat, lane, start, GW, st);
        if (op.kind == SUM_COLLECT) {  // this CTA's share of sum(clique): one share row per (collect op, CTA)
#pragma unroll
            for (int i = 0; i < kW; ++i) red[warp][lane * kW + i] = nacc.x[i];
            __syncthreads();
            if (warp == 0) {
                V s = vset(0.0);
#pragma unroll
                for (int w = 0; w < kLevelWarps; ++w)
#pragma unroll
                    for (int i = 0; i < kW; ++i) s.x[i] += red[w][lane * kW + i];
                stv(row_ptr(S, P.normpart_row0 + op.normslot * P.nb_max + blockIdx.x), s);
            }
            __syncthreads();
        }
    }
}

// finalisers of a phase: one warp per item. nb = CTAs per tile the phase's op kernel was launched with.
__global__ void __launch_bounds__(kLevelWarps * 32) lvl_fin_kernel(const DevPlan P, double *__restrict__ scratch, int phase_idx, int nb) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *S = scratch + (int64_t)blockIdx.y * P.tile_rows * kRowDoubles + lane * kW;
    const Phase ph = P.phases[phase_idx];
    const int f = ph.fin_begin + blockIdx.x * kLevelWarps + warp;
    if (f >= ph.fin_end) return;
    fin_item(P, P.fins[f], S, nb);
}

// posteriors + label
__global__ void __launch_bounds__(kLevelWarps * 32) lvl_output_kernel(const DevPlan P, double *__restrict__ scratch,
                                                                      const uint8_t *__restrict__ patG, int64_t tile0, int64_t n_cases,
                                                                      int32_t *__restrict__ labels, double *__restrict__ post) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int v = blockIdx.x * kLevelWarps + warp;
    if (v >= P.n_vars) return;
    const double *S = scratch + (int64_t)blockIdx.y * P.tile_rows * kRowDoubles + lane * kW;
    output_var(P, S, patG + (int64_t)blockIdx.y * P.pat_bytes, v, tile0 + blockIdx.y, n_cases, lane, labels, post);
}

}  // namespace

// Enqueues all launches for n_cases cases (several waves of `resident` tiles). Returns the number of launches via *n_launch.
cudaError_t level_launch(const DevPlan &P, const std::vector<Phase> &phases, int resident, double *scratch, uint8_t *patG,
                         int64_t n_cases, const int64_t *ev_off, int64_t ev_base, const int32_t *ev_vars, const int32_t *ev_vals,
                         int32_t *labels, double *post, cudaStream_t st, int64_t *n_launch) {
    if (P.cb != 4 && P.cb != 8) return cudaErrorInvalidValue;
    static bool attr_done = false;  // more than 48 KB of dynamic shared memory needs the opt-in, once per context is enough
    if (!attr_done) {
        cudaError_t ea = cudaFuncSetAttribute(lvl_sum_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kLvlSmem);
        if (ea == cudaSuccess) ea = cudaFuncSetAttribute(lvl_sum_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kLvlSmem);
        if (ea != cudaSuccess) return ea;
    }
    const int64_t n_tiles = (n_cases + kTileCases - 1) / kTileCases;
    int64_t launches = 0;
    for (int64_t tile0 = 0; tile0 < n_tiles; tile0 += resident) {
        const unsigned wt = (unsigned)std::min<int64_t>(resident, n_tiles - tile0);
        lvl_evidence_kernel<<<wt, 128, 0, st>>>(P, scratch, patG, tile0, n_cases, ev_off, ev_base, ev_vars, ev_vals);
        ++launches;
        for (size_t pi = 0; pi < phases.size(); ++pi) {
            const Phase &ph = phases[pi];
            // enough CTAs per tile to give every warp ~2 units, and enough CTAs overall to fill the GPU
            int nb = (ph.n_units + 2 * kLevelWarps - 1) / (2 * kLevelWarps);
            nb = std::max(1, std::min(nb, P.nb_max));
            dim3 grid((unsigned)nb, wt);
            if (P.cb == 8) lvl_sum_kernel<8><<<grid, kLevelWarps * 32, kLvlSmem, st>>>(P, scratch, patG, (int)pi);
            else lvl_sum_kernel<4><<<grid, kLevelWarps * 32, kLvlSmem, st>>>(P, scratch, patG, (int)pi);
            ++launches;
            const int nfin = ph.fin_end - ph.fin_begin;
            if (nfin > 0) {
                dim3 fg((unsigned)((nfin + kLevelWarps - 1) / kLevelWarps), wt);
                lvl_fin_kernel<<<fg, kLevelWarps * 32, 0, st>>>(P, scratch, (int)pi, nb);
                ++launches;
            }
        }
        dim3 og((unsigned)((P.n_vars + kLevelWarps - 1) / kLevelWarps), wt);
        lvl_output_kernel<<<og, kLevelWarps * 32, 0, st>>>(P, scratch, patG, tile0, n_cases, labels, post);
        ++launches;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    if (n_launch) *n_launch = launches;
    return cudaSuccess;
}

}  // namespace fbn
