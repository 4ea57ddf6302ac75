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

// ---- evidence patterns ------------------------------------------------------------------------------------------
// pat[v * kTileCases + lane * kW + i]: pattern byte of variable v for case (lane, i) of the tile: 0 = not observed, 1 + d =
// observed in state d, dim + 1 = observed in no valid state (or in two different ones): LoadDiscreteEvidence,
// src/JunctionTree.cpp:316-383. Filled by ONE warp (a lane walks the evidence pairs of its own cases, so every byte has a
// single writer); pat must be zeroed before.
__device__ __forceinline__ void load_patterns(const DevPlan &P, uint8_t *pat, int64_t tile, int64_t n_cases, const int64_t *ev_off,
                                              int64_t ev_base, const int32_t *ev_vars, const int32_t *ev_vals, int lane) {
#pragma unroll
    for (int i = 0; i < kW; ++i) {
        const int64_t cs = tile * kTileCases + lane * kW + i;
        if (cs >= n_cases) continue;
        const int64_t k1 = ev_off[cs + 1] - ev_base;
        for (int64_t k = ev_off[cs] - ev_base; k < k1; ++k) {
            const int v = ev_vars[k], val = ev_vals[k];
            if (v < 0 || v >= P.n_vars) continue;  // src/JunctionTree.cpp:326-331
            const int dim = P.vars[v].dim;
            const int np = (val < 0 || val >= dim) ? dim + 1 : 1 + val;
            uint8_t *b = pat + v * kTileCases + lane * kW + i;
            const int old = *b;
            *b = (uint8_t)((old == 0 || old == np) ? np : dim + 1);
        }
    }
}

// lambda rows (evidence indicators 1.0 / 0.0 per state) of lamvars[k], k = first, first + step, ...
__device__ __forceinline__ void write_lambda_rows(const DevPlan &P, double *S, const uint8_t *pat, int lane, int first, int step) {
    for (int k = first; k < P.n_lamvars; k += step) {
        const int v = P.lamvars[k];
        const VarInfo vi = P.vars[v];
        for (int d = 0; d < vi.dim; ++d) {
            V l;
#pragma unroll
            for (int i = 0; i < kW; ++i) {
                const int pv = pat[v * kTileCases + lane * kW + i];
                l.x[i] = (pv == 0 || pv == 1 + d) ? 1.0 : 0.0;
            }
            stv(row_ptr(S, P.lam_row0 + vi.off + d), l);
        }
    }
}

// this lane's evidence pattern index of an op's folded variables (one per case)
struct Fx {
    uint32_t x[kW];
};
__device__ __forceinline__ Fx op_fx(const DevPlan &P, const SumOp &op, const uint8_t *pat, int lane) {
    Fx f;
#pragma unroll
    for (int i = 0; i < kW; ++i) f.x[i] = 0;
    const int32_t *__restrict__ fl = P.idx32 + op.fold_off;
    for (int k = 0; k < op.nfold; ++k) {
        const int v = __ldg(fl + 2 * k), st = __ldg(fl + 2 * k + 1);
#pragma unroll
        for (int i = 0; i < kW; ++i) f.x[i] += (uint32_t)pat[v * kTileCases + lane * kW + i] * (uint32_t)st;
    }
    return f;
}

// ---- SUM ---------------------------------------------------------------------------------------------------------
// Targets [j0, j0+nc) (nc <= CB), rows [r0, r1) of one view:
//   e = potX[r, j][fx] * prod_g message_g[rowpart_g(r) + colpart_g(j)];  acc[c] += e(r, j0+c)
// NG = number of gathers per entry (compile-time 0..4, -1 = generic). Loads of targets beyond nc are clamped to the last one
// (valid addresses, results discarded), so the inner loop has no divergent control flow.
// UMASK (compile time): gathers whose row is the same for all CB targets of the block (one gather per row, folded into the
// row's scalar factor). FOLD: the view has folded variables: every case reads its own pattern's potential (per-lane 8-byte
// loads from nfx adjacent values); otherwise ONE coalesced load serves 32/CB rows x CB targets and the row loop picks its
// values by shuffle.
// Software pipeline: the loads of row r+1 are issued BEFORE row r is multiplied and summed (two register sets, used
// alternately); the row parts of the index maps are fetched two rows ahead. Rows that need many gathers keep one register
// set (loads issued together at the top of the row).
template <int NG, int UMASK, int CB, bool FOLD>
__device__ __forceinline__ void sum_rows(const DevPlan &P, const SumOp &op, const Fx &fx, const double *S, int lane, int j0, int r0, int r1,
                                         const int32_t (&cp)[NG > 0 ? NG : 1][CB], V (&acc)[CB]) {
    constexpr int NGC = NG > 0 ? NG : 1;
    constexpr int NU = (UMASK & 1) + ((UMASK >> 1) & 1) + ((UMASK >> 2) & 1) + ((UMASK >> 3) & 1);
    constexpr int kLoads = (NG > 0 ? (NG - NU) * CB + NU : 0) * kW + (FOLD ? CB * kW / 2 : 0);  // 16-byte units per row
    constexpr bool kTwoSets = NG > 0 && CB >= 4 && kLoads <= 12;
    const int ng = NG >= 0 ? NG : op.ng, cols = op.cols;
    const bool has_pot = op.pot_off >= 0;
    const double *__restrict__ pot = P.pot + (has_pot ? op.pot_off : 0);
    const int32_t *__restrict__ rowpart = P.idx32 + op.rowpart_off;
    const int32_t *__restrict__ colpart = P.idx32 + op.colpart_off;
    constexpr int RPL = 32 / CB;
    const int lane_c = lane % CB, lane_r = lane / CB;
    const double *potl = pot + min(j0 + lane_c, cols - 1);
    double potv = 1.0, potv_next = (has_pot && !FOLD) ? __ldg(potl + (int64_t)min(r0 + lane_r, r1 - 1) * cols) : 1.0;
    int jc[CB];
#pragma unroll
    for (int c = 0; c < CB; ++c) jc[c] = min(j0 + c, cols - 1);
    const int64_t nfx = op.nfx;

    struct Set {
        V m[NGC][CB];
        double pv[FOLD ? CB : 1][kW];
    };
    auto load_parts = [&](int rr, int32_t (&rp)[NGC]) {
        if (NG > 0) {
            const int32_t *q = rowpart + (int64_t)min(rr, r1 - 1) * NG;
#pragma unroll
            for (int g = 0; g < NG; ++g) rp[g] = __ldg(q + g);
        }
    };
    auto gather = [&](int rr, const int32_t (&rp)[NGC], Set &s) {
        if (FOLD) {
            const int64_t rbase = (int64_t)min(rr, r1 - 1) * cols;
#pragma unroll
            for (int c = 0; c < CB; ++c)
#pragma unroll
                for (int i = 0; i < kW; ++i) s.pv[FOLD ? c : 0][i] = __ldg(pot + (rbase + jc[c]) * nfx + fx.x[i]);
        }
        if (NG > 0) {
#pragma unroll
            for (int g = 0; g < NG; ++g) {
                const double *mb = S + (int64_t)rp[g] * kRowDoubles;
                if ((UMASK >> g) & 1) {
                    s.m[g][0] = ldv(mb + (int64_t)cp[g][0] * kRowDoubles);
                } else {
#pragma unroll
                    for (int c = 0; c < CB; ++c) s.m[g][c] = ldv(mb + (int64_t)cp[g][c] * kRowDoubles);
                }
            }
        }
    };
    auto finish = [&](int r, const Set &s) {
        double pc[CB];
        if (!FOLD) {
            const int rg = (r - r0) & (RPL - 1);  // row inside its group of RPL
            if (rg == 0) {
                potv = potv_next;
                if (has_pot) potv_next = __ldg(potl + (int64_t)min(r + RPL + lane_r, r1 - 1) * cols);
            }
#pragma unroll
            for (int c = 0; c < CB; ++c) pc[c] = __shfl_sync(0xffffffffu, potv, rg * CB + c);
        }
        V u = vset(1.0);
        if (NG > 0) {
#pragma unroll
            for (int g = 0; g < NG; ++g)
                if ((UMASK >> g) & 1) vmul(u, s.m[g][0]);
        }
#pragma unroll
        for (int i = 0; i < kW; ++i) {
#pragma unroll
            for (int c = 0; c < CB; ++c) {
                double v = (FOLD ? s.pv[FOLD ? c : 0][i] : pc[c]) * u.x[i];
                if (NG > 0) {
#pragma unroll
                    for (int g = 0; g < NG; ++g)
                        if (!((UMASK >> g) & 1)) v *= s.m[g][c].x[i];
                } else if (NG < 0) {
                    for (int g = 0; g < ng; ++g) {
                        const int32_t row = __ldg(rowpart + (int64_t)r * ng + g) + __ldg(colpart + (int64_t)jc[c] * ng + g);
                        v *= row_ptr(S, row)[i];
                    }
                }
                acc[c].x[i] += v;
            }
        }
    };

    if constexpr (kTwoSets) {
        int32_t rp0[NGC], rp1[NGC];
        Set s0, s1;
        load_parts(r0, rp0);
        load_parts(r0 + 1, rp1);
        gather(r0, rp0, s0);      // row r0
        load_parts(r0 + 2, rp0);  // parts of row r0 + 2
        for (int r = r0; r < r1; r += 2) {
            // here: s0 = row r (in flight), rp1 = parts of row r+1, rp0 = parts of row r+2 (in flight)
            gather(r + 1, rp1, s1);
            load_parts(r + 3, rp1);
            finish(r, s0);
            if (r + 1 < r1) {
                gather(r + 2, rp0, s0);
                load_parts(r + 4, rp0);
                finish(r + 1, s1);
            }
        }
    } else {
        int32_t rp[NGC], rpn[NGC];
        Set s1;
        load_parts(r0, rpn);
#pragma unroll(CB >= 4 ? 1 : kSumUnrollCB2)
        for (int r = r0; r < r1; ++r) {
#pragma unroll
            for (int g = 0; g < NGC; ++g) rp[g] = rpn[g];
            gather(r, rp, s1);
            load_parts(r + 1, rpn);
            finish(r, s1);
        }
    }
}

template <int NG, int CB>
__device__ __forceinline__ void sum_block(const DevPlan &P, const SumOp &op, const Fx &fx, const double *S, int lane, int j0, int r0, int r1,
                                          V (&acc)[CB]) {
    const int cols = op.cols;
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
                cp[g][c] = __ldg(colpart + (int64_t)min(j0 + c, cols - 1) * NG + g);
                u = u && cp[g][c] == cp[g][0];
            }
            if (u) umask |= 1 << g;
        }
    }
    auto run = [&](auto um) {
        constexpr int UM = decltype(um)::value;
        if (op.nfx > 1) sum_rows<NG, UM, CB, true>(P, op, fx, S, lane, j0, r0, r1, cp, acc);
        else sum_rows<NG, UM, CB, false>(P, op, fx, S, lane, j0, r0, r1, cp, acc);
    };
    if constexpr (NG > 0 && NG <= 2 && CB >= 2) {
        switch (umask) {
            case 0: run(std::integral_constant<int, 0>{}); break;
            case 1: run(std::integral_constant<int, 1>{}); break;
            case 2: run(std::integral_constant<int, 2 % (1 << NG)>{}); break;
            default: run(std::integral_constant<int, (1 << NG) - 1>{}); break;
        }
    } else if constexpr (NG >= 3 && CB >= 2) {
        // three or four gathers: specialise the all-uniform block and the common "only the last gather (the clique's own down
        // ratio, indexed by its trailing variables) differs per target" block
        constexpr int kAll = (1 << NG) - 1, kButLast = kAll >> 1;
        if (umask == kAll) run(std::integral_constant<int, kAll>{});
        else if ((umask & kButLast) == kButLast) run(std::integral_constant<int, kButLast>{});
        else run(std::integral_constant<int, 0>{});
    } else {
        run(std::integral_constant<int, 0>{});
    }
}

template <int CB>
__device__ __forceinline__ void sum_block_any(const DevPlan &P, const SumOp &op, const Fx &fx, const double *S, int lane, int j0, int r0,
                                              int r1, V (&acc)[CB]) {
    switch (op.ng) {
        case 0: sum_block<0, CB>(P, op, fx, S, lane, j0, r0, r1, acc); break;
        case 1: sum_block<1, CB>(P, op, fx, S, lane, j0, r0, r1, acc); break;
        case 2: sum_block<2, CB>(P, op, fx, S, lane, j0, r0, r1, acc); break;
        case 3: sum_block<3, CB>(P, op, fx, S, lane, j0, r0, r1, acc); break;
        case 4: sum_block<4, CB>(P, op, fx, S, lane, j0, r0, r1, acc); break;
        default: sum_block<-1, CB>(P, op, fx, S, lane, j0, r0, r1, acc); break;
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

// product of the 1/norm rows that scale the sums of an op (PotentialTableBase::Normalize as scalars applied by the consumers)
__device__ __forceinline__ V op_scale(const DevPlan &P, const SumOp &op, const double *S) {
    V scale = vset(1.0);
    const int32_t *__restrict__ invrows = P.idx32 + op.invrow_off;
    for (int k = 0; k < op.ninv; ++k) vmul(scale, ldv(row_ptr(S, __ldg(invrows + k))));
    return scale;
}

// a finished sum of target j: the op's scale, and zero where the case observed another state of an uncovered target variable
__device__ __forceinline__ V finish_target(const DevPlan &P, const SumOp &op, const uint8_t *pat, int lane, int j, V acc, const V &scale) {
    vmul(acc, scale);
    const int32_t *__restrict__ cmk = P.idx32 + op.cmask_off;
    for (int k = 0; k < op.ncmask; ++k) {
        const int v = __ldg(cmk + 3 * k), d = (j / __ldg(cmk + 3 * k + 1)) % __ldg(cmk + 3 * k + 2);
#pragma unroll
        for (int i = 0; i < kW; ++i) {
            const int pv = pat[v * kTileCases + lane * kW + i];
            if (pv != 0 && pv != 1 + d) acc.x[i] = 0.0;
        }
    }
    return acc;
}

// what happens to the finished sums of a unit: partial slot, or the op's own finaliser. Returns the sum of the unit's targets
// (a collect op's share of the clique norm).
template <int CB>
__device__ __forceinline__ V store_unit(const DevPlan &P, const SumOp &op, double *S, const uint8_t *pat, int lane, int j0, int nc, int ck,
                                        const V &scale, const V (&acc)[CB]) {
    V nacc = vset(0.0);
#pragma unroll
    for (int c = 0; c < CB; ++c)
        if (c < nc) {
            const int j = j0 + c;
            const V a = finish_target(P, op, pat, lane, j, acc[c], scale);
            if (op.chunks > 1) {
                stv(row_ptr(S, P.part_row0 + op.pslot0 + j * op.chunks + ck), a);
            } else if (op.kind == SUM_DISTRIBUTE) {
                const V up = ldv(row_ptr(S, op.aux_row + j));
                V d;
#pragma unroll
                for (int i = 0; i < kW; ++i) d.x[i] = (up.x[i] == 0.0) ? 0.0 : a.x[i];  // TableDivision 0 guard, PotentialTable.cpp:718-734
                stv(row_ptr(S, op.dst_row + j), d);
            } else {
                stv(row_ptr(S, op.dst_row + j), a);
            }
            vadd(nacc, a);
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

// ---- posteriors + label of one variable (GetProbabilitiesOneNode src/JunctionTree.cpp:1392-1454, Inference::ArgMax
//      src/Inference.cpp:92-102): normalise the unnormalised marginal rows, zeros for an observed variable
__device__ __forceinline__ void output_var(const DevPlan &P, const double *S, const uint8_t *pat, int v, int64_t tile, int64_t n_cases,
                                           int lane, int32_t *labels, double *post) {
    const VarInfo vi = P.vars[v];
    const double *mg = row_ptr(S, P.mg_row0 + vi.off);
#pragma unroll
    for (int i = 0; i < kW; ++i) {
        const int64_t cs = tile * kTileCases + lane * kW + i;
        const bool valid = cs < n_cases;
        const bool observed = pat[v * kTileCases + lane * kW + i] != 0;
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
