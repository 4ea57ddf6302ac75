// Device-side building blocks shared by the fused kernel (kernels.cu) and the level kernels (level_kernels.cu).
// Thread mapping everywhere: a lane holds kW test cases of a 32*kW-case tile (one 8*kW-byte vector per scratch row: 128-bit
// loads / stores for kW = 2); a warp owns a UNIT = a block of up to CB adjacent targets of one SumOp (plan.h) times a
// range of rows of its view. Register blocking over the targets gives CB independent row gathers per index load
// (memory-level parallelism) and amortises the warp-uniform index loads; all scratch accesses are coalesced rows.
// Clique entries are recomputed in registers, summed and dropped: the only scratch traffic is message rows.
#pragma once
#include <cstdint>
#include <type_traits>
#include <cuda_runtime.h>

#include "plan.h"

namespace fbn {

// unroll factor of the single-register-set row loop of the CB = 2 kernels
#ifndef FBN_SUM_UNROLL_CB2
#define FBN_SUM_UNROLL_CB2 2
#endif
constexpr int kSumUnrollCB2 = FBN_SUM_UNROLL_CB2;

// ---- per-lane vectors: the kW cases a lane owns -------------------------------------------------------------------
struct V {
    double x[kW];
};
struct M {   // evidence mask words: bit l of w[i] <=> case (lane l, slot i) of the tile
    uint32_t w[kW];
};

// S points at this lane's first double of row 0 of the tile (tile base + lane * kW)
__device__ __forceinline__ double *row_ptr(double *S, int32_t row) { return S + (int64_t)row * kRowDoubles; }
__device__ __forceinline__ const double *row_ptr(const double *S, int32_t row) { return S + (int64_t)row * kRowDoubles; }

__device__ __forceinline__ V ldv(const double *p) {
    V v;
    if constexpr (kW == 1) {
        v.x[0] = *p;
    } else {
#pragma unroll
        for (int i = 0; i < kW; i += 2) {
            const double2 t = *reinterpret_cast<const double2 *>(p + i);   // LDG.E.128
            v.x[i] = t.x;
            v.x[i + 1] = t.y;
        }
    }
    return v;
}
__device__ __forceinline__ void stv(double *p, const V &v) {
    if constexpr (kW == 1) {
        *p = v.x[0];
    } else {
#pragma unroll
        for (int i = 0; i < kW; i += 2) *reinterpret_cast<double2 *>(p + i) = make_double2(v.x[i], v.x[i + 1]);   // STG.E.128
    }
}
__device__ __forceinline__ V vset(double a) {
    V v;
#pragma unroll
    for (int i = 0; i < kW; ++i) v.x[i] = a;
    return v;
}
__device__ __forceinline__ void vmul(V &a, const V &b) {
#pragma unroll
    for (int i = 0; i < kW; ++i) a.x[i] *= b.x[i];
}
__device__ __forceinline__ void vadd(V &a, const V &b) {
#pragma unroll
    for (int i = 0; i < kW; ++i) a.x[i] += b.x[i];
}

// ---- row evidence masks -----------------------------------------------------------------------------------------
// mask of view row r: bit l of word i set <=> case (l, i) of the tile allows that row (AND of the lambda words of the row's digits)
struct RowMaskTable {  // fused mode, tables built once per phase in shared memory
    static constexpr int kSpan = 1 << 30;  // rows per begin(): the whole range
    const uint32_t *tab;
    __device__ __forceinline__ void begin(int, int, int) {}
    __device__ __forceinline__ M get(int r, int) const {
        M m;
#pragma unroll
        for (int i = 0; i < kW; ++i) m.w[i] = tab[r * kW + i];
        return m;
    }
};
struct RowMaskLam {  // evaluated on the fly, lane-parallel: lane i builds the words of row rb+i, get() broadcasts them
    static constexpr int kSpan = 32;
    const uint32_t *lam;       // lambda words of the tile (shared or global memory), kW per (variable, state)
    const uint16_t *rowlam;    // op's rows x nrest lambda indices
    int nrest;
    M mine;
    __device__ __forceinline__ void begin(int rb, int r1, int lane) {
#pragma unroll
        for (int i = 0; i < kW; ++i) mine.w[i] = 0xffffffffu;
        if (rb + lane < r1) {
            const uint16_t *__restrict__ rl = rowlam + (int64_t)(rb + lane) * nrest;
            for (int k = 0; k < nrest; ++k) {
                const uint32_t *lw = lam + (int)__ldg(rl + k) * kW;
#pragma unroll
                for (int i = 0; i < kW; ++i) mine.w[i] &= lw[i];
            }
        }
    }
    __device__ __forceinline__ M get(int r, int rb) const {
        M m;
#pragma unroll
        for (int i = 0; i < kW; ++i) m.w[i] = __shfl_sync(0xffffffffu, mine.w[i], r - rb);
        return m;
    }
};

// mask of target j of an op, from the lambda words
__device__ __forceinline__ M col_mask_from_lam(const DevPlan &P, const SumOp &op, const uint32_t *lam, int j) {
    M m;
#pragma unroll
    for (int i = 0; i < kW; ++i) m.w[i] = 0xffffffffu;
    const uint16_t *__restrict__ cl16 = P.lam16 + op.collam_off + (int64_t)j * op.ncolv;
    for (int k = 0; k < op.ncolv; ++k) {
        const uint32_t *lw = lam + (int)__ldg(cl16 + k) * kW;
#pragma unroll
        for (int i = 0; i < kW; ++i) m.w[i] &= lw[i];
    }
    return m;
}

// ---- SUM ---------------------------------------------------------------------------------------------------------
// Targets [j0, j0+nc) (nc <= CB), rows [r0, r1) of one view:
//   e = pristine[r, j] * scale * mask(r, j) * prod_g message_g[rowpart_g(r) + colpart_g(j)];  acc[c] += e(r, j0+c)
// NG = number of gathers per entry (compile-time 0..4, -1 = generic). Loads of targets beyond nc are clamped to the last one
// (valid addresses, results discarded), so the inner loop has no divergent control flow.
// UMASK (compile time): gathers whose row is the same for all CB targets of the block (one gather per row, folded into the
// row's scalar factor).
// Software pipeline: the gathers of row r+1 are issued BEFORE row r is multiplied, masked and summed (two register sets,
// used alternately); the row parts of the index maps are fetched two rows ahead, the coalesced pristine-potential load one
// group ahead. Rows that need many gathers keep one register set (gathers issued together at the top of the row).
template <int NG, int UMASK, int CB, class RowMask>
__device__ __forceinline__ void sum_rows(const DevPlan &P, const SumOp &op, RowMask &rmask, const M (&cm)[CB], const double *S,
                                         int lane, int j0, int r0, int r1, const V &scale, const int32_t (&cp)[NG > 0 ? NG : 1][CB],
                                         V (&acc)[CB]) {
    constexpr int NGC = NG > 0 ? NG : 1;
    constexpr int NU = (UMASK & 1) + ((UMASK >> 1) & 1) + ((UMASK >> 2) & 1) + ((UMASK >> 3) & 1);
    constexpr int kGathers = NG > 0 ? (NG - NU) * CB + NU : 0;  // per row
    constexpr bool kTwoSets = NG > 0 && CB >= 4 && kGathers * kW <= 12;
    const int ng = NG >= 0 ? NG : op.ng, cols = op.cols;
    const bool has_pot = op.pot_off >= 0;
    const double *__restrict__ pot = P.pot + (has_pot ? op.pot_off : 0);
    const int32_t *__restrict__ rowpart = P.idx32 + op.rowpart_off;
    const int32_t *__restrict__ colpart = P.idx32 + op.colpart_off;
    const uint32_t lanebit = 1u << lane;
    uint32_t colon[kW];  // bit c of word i: case (lane, i) allows target j0+c
#pragma unroll
    for (int i = 0; i < kW; ++i) {
        colon[i] = 0;
#pragma unroll
        for (int c = 0; c < CB; ++c)
            if (cm[c].w[i] & lanebit) colon[i] |= 1u << c;
    }
    // pristine potentials: ONE coalesced load serves RPL rows x CB targets (lane = row-in-group * CB + target), the row
    // loop picks its values by shuffle; a warp-uniform 8-byte load per entry would cost a load instruction (and a
    // scoreboard stall) per entry
    constexpr int RPL = 32 / CB;
    const int lane_c = lane % CB, lane_r = lane / CB;
    const double *potl = pot + min(j0 + lane_c, cols - 1);
    double potv = 1.0, potv_next = has_pot ? __ldg(potl + (int64_t)min(r0 + lane_r, r1 - 1) * cols) : 1.0;
    int rb = r0;

    auto load_parts = [&](int rr, int32_t (&rp)[NGC]) {
        if (NG > 0) {
            const int32_t *q = rowpart + (int64_t)min(rr, r1 - 1) * NG;
#pragma unroll
            for (int g = 0; g < NG; ++g) rp[g] = __ldg(q + g);
        }
    };
    auto gather = [&](const int32_t (&rp)[NGC], V (&m)[NGC][CB]) {
        if (NG > 0) {
#pragma unroll
            for (int g = 0; g < NG; ++g) {
                const double *mb = S + (int64_t)rp[g] * kRowDoubles;
                if ((UMASK >> g) & 1) {
                    m[g][0] = ldv(mb + (int64_t)cp[g][0] * kRowDoubles);
                } else {
#pragma unroll
                    for (int c = 0; c < CB; ++c) m[g][c] = ldv(mb + (int64_t)cp[g][c] * kRowDoubles);
                }
            }
        }
    };
    auto finish = [&](int r, const V (&m)[NGC][CB]) {
        if (RowMask::kSpan <= 32 && ((r - r0) & 31) == 0) {
            rb = r;
            rmask.begin(rb, r1, lane);
        }
        const int rg = (r - r0) & (RPL - 1);  // row inside its group of RPL
        if (rg == 0) {
            potv = potv_next;
            if (has_pot) potv_next = __ldg(potl + (int64_t)min(r + RPL + lane_r, r1 - 1) * cols);
        }
        const M rm = rmask.get(r, rb);
        V u = scale;
        if (NG > 0) {
#pragma unroll
            for (int g = 0; g < NG; ++g)
                if ((UMASK >> g) & 1) vmul(u, m[g][0]);
        }
        double pc[CB];
#pragma unroll
        for (int c = 0; c < CB; ++c) pc[c] = __shfl_sync(0xffffffffu, potv, rg * CB + c);
#pragma unroll
        for (int i = 0; i < kW; ++i) {
            const uint32_t on = (rm.w[i] & lanebit) ? colon[i] : 0u;
#pragma unroll
            for (int c = 0; c < CB; ++c) {
                double v = pc[c] * u.x[i];
                if (NG > 0) {
#pragma unroll
                    for (int g = 0; g < NG; ++g)
                        if (!((UMASK >> g) & 1)) v *= m[g][c].x[i];
                } else if (NG < 0) {
                    for (int g = 0; g < ng; ++g) {
                        const int32_t row = __ldg(rowpart + (int64_t)r * ng + g) + __ldg(colpart + (int64_t)min(j0 + c, cols - 1) * ng + g);
                        v *= row_ptr(S, row)[i];
                    }
                }
                acc[c].x[i] += ((on >> c) & 1u) ? v : 0.0;
            }
        }
    };

    if (RowMask::kSpan > 32) rmask.begin(r0, r1, lane);
    if constexpr (kTwoSets) {
        int32_t rp0[NGC], rp1[NGC];
        V m0[NGC][CB], m1[NGC][CB];
        load_parts(r0, rp0);
        load_parts(r0 + 1, rp1);
        gather(rp0, m0);          // row r0
        load_parts(r0 + 2, rp0);  // parts of row r0 + 2
        for (int r = r0; r < r1; r += 2) {
            // here: m0 = row r (in flight), rp1 = parts of row r+1, rp0 = parts of row r+2 (in flight)
            gather(rp1, m1);
            load_parts(r + 3, rp1);
            finish(r, m0);
            if (r + 1 < r1) {
                gather(rp0, m0);
                load_parts(r + 4, rp0);
                finish(r + 1, m1);
            }
        }
    } else {
        int32_t rp[NGC], rpn[NGC];
        V m[NGC][CB];
        load_parts(r0, rpn);
#pragma unroll(CB >= 4 ? 1 : kSumUnrollCB2)
        for (int r = r0; r < r1; ++r) {
#pragma unroll
            for (int g = 0; g < NGC; ++g) rp[g] = rpn[g];
            gather(rp, m);
            load_parts(r + 1, rpn);
            finish(r, m);
        }
    }
}

template <int NG, int CB, class RowMask>
__device__ __forceinline__ void sum_block(const DevPlan &P, const SumOp &op, RowMask &rmask, const M (&cm)[CB], const double *S,
                                          int lane, int j0, int r0, int r1, const V &scale, V (&acc)[CB]) {
    const int cols = op.cols;
    int jc[CB];
#pragma unroll
    for (int c = 0; c < CB; ++c) jc[c] = min(j0 + c, cols - 1);
    int32_t cp[NG > 0 ? NG : 1][CB];
    const int32_t *__restrict__ colpart = P.idx32 + op.colpart_off;
    // umask bit g: all CB targets of the block read the SAME row of gather g (the block's target variables are outside
    // that source table) -> one gather per row instead of CB, folded into the row's scalar factor
    int umask = 0;
    if (NG > 0) {
#pragma unroll
        for (int g = 0; g < NG; ++g) {
            bool u = true;
#pragma unroll
            for (int c = 0; c < CB; ++c) {
                cp[g][c] = __ldg(colpart + (int64_t)jc[c] * NG + g);
                u = u && cp[g][c] == cp[g][0];
            }
            if (u) umask |= 1 << g;
        }
    }
    if constexpr (NG > 0 && NG <= 2 && CB >= 2) {
        switch (umask) {
            case 0: sum_rows<NG, 0, CB>(P, op, rmask, cm, S, lane, j0, r0, r1, scale, cp, acc); break;
            case 1: sum_rows<NG, 1, CB>(P, op, rmask, cm, S, lane, j0, r0, r1, scale, cp, acc); break;
            case 2: sum_rows<NG, 2 % (1 << NG), CB>(P, op, rmask, cm, S, lane, j0, r0, r1, scale, cp, acc); break;
            default: sum_rows<NG, (1 << NG) - 1, CB>(P, op, rmask, cm, S, lane, j0, r0, r1, scale, cp, acc); break;
        }
    } else if constexpr (NG >= 3 && CB >= 2) {
        // three or four gathers: specialise the all-uniform block and the common "only the last gather (the clique's own down
        // ratio, indexed by its trailing variables) differs per target" block
        constexpr int kAll = (1 << NG) - 1, kButLast = kAll >> 1;
        if (umask == kAll) sum_rows<NG, kAll, CB>(P, op, rmask, cm, S, lane, j0, r0, r1, scale, cp, acc);
        else if ((umask & kButLast) == kButLast) sum_rows<NG, kButLast, CB>(P, op, rmask, cm, S, lane, j0, r0, r1, scale, cp, acc);
        else sum_rows<NG, 0, CB>(P, op, rmask, cm, S, lane, j0, r0, r1, scale, cp, acc);
    } else {
        sum_rows<NG, 0, CB>(P, op, rmask, cm, S, lane, j0, r0, r1, scale, cp, acc);
    }
}

template <int CB, class RowMask>
__device__ __forceinline__ void sum_block_any(const DevPlan &P, const SumOp &op, RowMask &rmask, const M (&cm)[CB], const double *S,
                                              int lane, int j0, int r0, int r1, const V &scale, V (&acc)[CB]) {
    switch (op.ng) {
        case 0: sum_block<0, CB>(P, op, rmask, cm, S, lane, j0, r0, r1, scale, acc); break;
        case 1: sum_block<1, CB>(P, op, rmask, cm, S, lane, j0, r0, r1, scale, acc); break;
        case 2: sum_block<2, CB>(P, op, rmask, cm, S, lane, j0, r0, r1, scale, acc); break;
        case 3: sum_block<3, CB>(P, op, rmask, cm, S, lane, j0, r0, r1, scale, acc); break;
        case 4: sum_block<4, CB>(P, op, rmask, cm, S, lane, j0, r0, r1, scale, acc); break;
        default: sum_block<-1, CB>(P, op, rmask, cm, S, lane, j0, r0, r1, scale, acc); break;
    }
}

// unit u of an op -> target block + row range
template <int CB>
__device__ __forceinline__ void sum_unit(const DevPlan &P, const SumOp &op, int u, int &j0, int &nc, int &ck, int &r0, int &r1) {
    int jb = u;
    ck = 0; r0 = 0; r1 = op.rows;
    if (op.chunks > 1) {
        jb = u / op.chunks;
        ck = u - jb * op.chunks;
        r0 = ck * op.rows_per_chunk;
        r1 = min(op.rows, r0 + op.rows_per_chunk);
    }
    if (op.perm_off >= 0) jb = __ldg(P.idx32 + op.perm_off + jb);  // planner's visiting order of the target blocks
    j0 = jb * CB;
    nc = min(CB, op.cols - j0);
}

// product of the 1/norm rows that scale every entry of an op (PotentialTableBase::Normalize as scalars applied by the consumers)
__device__ __forceinline__ V op_scale(const DevPlan &P, const SumOp &op, const double *S) {
    V scale = vset(1.0);
    const int32_t *__restrict__ invrows = P.idx32 + op.invrow_off;
    for (int k = 0; k < op.ninv; ++k) vmul(scale, ldv(row_ptr(S, __ldg(invrows + k))));
    return scale;
}

// what happens to the finished sums of a unit: partial slot, or the op's own finaliser. Returns the sum of the unit's targets
// (a collect op's share of the clique norm).
template <int CB>
__device__ __forceinline__ V store_unit(const DevPlan &P, const SumOp &op, double *S, int j0, int nc, int ck, const V (&acc)[CB]) {
    V nacc = vset(0.0);
#pragma unroll
    for (int c = 0; c < CB; ++c)
        if (c < nc) {
            const int j = j0 + c;
            if (op.chunks > 1) {
                stv(row_ptr(S, P.part_row0 + op.pslot0 + j * op.chunks + ck), acc[c]);
            } else if (op.kind == SUM_DISTRIBUTE) {
                const V up = ldv(row_ptr(S, op.aux_row + j));
                V d;
#pragma unroll
                for (int i = 0; i < kW; ++i) d.x[i] = (up.x[i] == 0.0) ? 0.0 : acc[c].x[i];  // TableDivision 0 guard, PotentialTable.cpp:718-734
                stv(row_ptr(S, op.dst_row + j), d);
            } else {
                stv(row_ptr(S, op.dst_row + j), acc[c]);
            }
            vadd(nacc, acc[c]);
        }
    return nacc;
}

// finaliser of a phase: sums of partial slots / norm shares. `shares` = norm shares per collect op actually written.
__device__ __forceinline__ void fin_item(const DevPlan &P, const FinItem &F, double *S, int shares) {
    V s = vset(0.0);
    if (F.type == FIN_NORM) {
        const double *np = row_ptr(S, P.normpart_row0 + F.slot0 * P.nb_max);
        for (int b0 = 0; b0 < shares; b0 += 4) {   // four independent loads in flight; summed in share order
            V v[4];
#pragma unroll
            for (int b = 0; b < 4; ++b) v[b] = ldv(np + (int64_t)min(b0 + b, shares - 1) * kRowDoubles);
#pragma unroll
            for (int b = 0; b < 4; ++b)
                if (b0 + b < shares) vadd(s, v[b]);
        }
#pragma unroll
        for (int i = 0; i < kW; ++i) s.x[i] = 1.0 / s.x[i];
        stv(row_ptr(S, F.dst_row), s);
        return;
    }
    const double *pp = row_ptr(S, P.part_row0 + F.slot0);
    for (int i = 0; i < F.nslots; ++i) vadd(s, ldv(pp + (int64_t)i * kRowDoubles));
    if (F.type == FIN_D) {
        const V up = ldv(row_ptr(S, F.aux_row));
#pragma unroll
        for (int i = 0; i < kW; ++i)
            if (up.x[i] == 0.0) s.x[i] = 0.0;
    }
    stv(row_ptr(S, F.dst_row), s);
}

// ---- evidence -> lambda words (LoadDiscreteEvidence, src/JunctionTree.cpp:316-383, mask form :386-457) ------------------
// lam[(off_v + d) * kW + i] bit l: state d of variable v allowed for case (l, i); lam[(sum_dims + v) * kW + i] bit l: v observed.
// Called by `nw` warps that share the tile (warp `warp` takes every nw-th evidence pair); lam must be initialised.
__device__ __forceinline__ void load_evidence(const DevPlan &P, uint32_t *lam, int64_t tile, int64_t n_cases, const int64_t *ev_off,
                                              int64_t ev_base, const int32_t *ev_vars, const int32_t *ev_vals, int lane, int warp, int nw) {
#pragma unroll
    for (int i = 0; i < kW; ++i) {
        const int64_t cs = tile * kTileCases + lane * kW + i;
        if (cs >= n_cases) continue;
        const int64_t k1 = ev_off[cs + 1] - ev_base;
        for (int64_t k = ev_off[cs] - ev_base + warp; k < k1; k += nw) {
            const int v = ev_vars[k], val = ev_vals[k];
            if (v < 0 || v >= P.n_vars) continue;  // src/JunctionTree.cpp:326-331
            const VarInfo vi = P.vars[v];
            atomicOr(&lam[(P.sum_dims + v) * kW + i], 1u << lane);
            for (int d = 0; d < vi.dim; ++d)
                if (d != val) atomicAnd(&lam[(vi.off + d) * kW + i], ~(1u << lane));
        }
    }
}

// ---- posteriors + label of one variable (GetProbabilitiesOneNode src/JunctionTree.cpp:1392-1454, Inference::ArgMax
//      src/Inference.cpp:92-102): normalise the unnormalised marginal rows, zeros for an observed variable
__device__ __forceinline__ void output_var(const DevPlan &P, const double *S, const uint32_t *lam, int v, int64_t tile, int64_t n_cases,
                                           int lane, int32_t *labels, double *post) {
    const VarInfo vi = P.vars[v];
    const double *mg = row_ptr(S, P.mg_row0 + vi.off);
#pragma unroll
    for (int i = 0; i < kW; ++i) {
        const int64_t cs = tile * kTileCases + lane * kW + i;
        const bool valid = cs < n_cases;
        const bool observed = (lam[(P.sum_dims + v) * kW + i] >> lane) & 1u;
        double sum = 0.0;
        for (int d = 0; d < vi.dim; ++d) sum += mg[(int64_t)d * kRowDoubles + i];
        double best = 0.0;
        int arg = -1;
        for (int d = 0; d < vi.dim; ++d) {
            const double pr = observed ? 0.0 : mg[(int64_t)d * kRowDoubles + i] / sum;
            if (valid && post) post[cs * P.sum_dims + vi.off + d] = pr;
            if (pr > best) { best = pr; arg = d; }
        }
        if (v == P.query_var && valid && labels) labels[cs] = arg;
    }
}

}  // namespace fbn
