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

#ifndef FBN_TWOSET_MAX
#define FBN_TWOSET_MAX 16   // 16-byte loads per lane and row up to which the row loop keeps two register sets (B200: 12 -> 16 HEPAR2 shape +5 %, MUNIN shape +4 %, PIGS shape -1 %; 24: spills, -3..-6 %)
#endif
#ifndef FBN_POT_BCAST
#define FBN_POT_BCAST 1
#endif
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

// ---- potentials through shared memory: cp.async.bulk (TMA) + mbarrier ---------------------------------------------
// The potentials of one unit of warp work are one contiguous stream (unit-major potX, plan.h). Every warp owns two
// shared-memory buffers and two mbarriers: lane 0 starts the bulk copy of the NEXT block of rows (or of the next unit's
// first block) while the warp sums the current one; the row loop then reads its potentials from shared memory (same
// address in all lanes: one broadcast LDS.128 per two targets) instead of waiting on a global load and passing values
// around by shuffles. Copies are issued and waited for in the same order, buffer = count & 1, mbarrier parity = (count >> 1) & 1.
struct PotStage {
    unsigned char *base;   // this warp's two buffers (generic address), buf_bytes each
    uint32_t base_s;       // the same as a shared-memory address
    uint32_t bar_s;        // shared-memory address of this warp's two mbarriers (8 bytes each)
    uint32_t buf_bytes;
    uint32_t issued, waited;
};
__device__ __forceinline__ void pot_stage_init(PotStage &st, unsigned char *bufs, uint64_t *bars, uint32_t buf_bytes, int lane) {
    st.base = bufs;
    st.base_s = (uint32_t)__cvta_generic_to_shared(bufs);
    st.bar_s = (uint32_t)__cvta_generic_to_shared(bars);
    st.buf_bytes = buf_bytes;
    st.issued = st.waited = 0;
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(st.bar_s));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(st.bar_s + 8));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
}
// start the copy of `bytes` (multiple of 16, <= buf_bytes) from global `src` (16-byte aligned) into the next buffer
__device__ __forceinline__ void pot_issue(PotStage &st, const double *src, uint32_t bytes, int lane) {
    __syncwarp();  // every lane has finished reading the buffer this copy overwrites
    if (lane == 0) {
        const uint32_t slot = st.issued & 1u, bar = st.bar_s + 8 * slot;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the warp's reads of the buffer (generic proxy) before the copy engine's write
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(st.base_s + slot * st.buf_bytes),
                     "l"(src), "r"(bytes), "r"(bar)
                     : "memory");
    }
    ++st.issued;
}
// wait for the oldest copy in flight; returns its buffer
__device__ __forceinline__ const double *pot_wait(PotStage &st) {
    const uint32_t slot = st.waited & 1u, parity = (st.waited >> 1) & 1u, bar = st.bar_s + 8 * slot;
    uint32_t done = 0;
    while (!done) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(bar), "r"(parity)
                     : "memory");
    }
    ++st.waited;
    return reinterpret_cast<const double *>(st.base + slot * st.buf_bytes);
}
// can this op's potentials go through the staged path? (plain potentials; folded views read per-case pattern slots)
__device__ __forceinline__ bool pot_staged(const SumOp &op) { return op.pot_off >= 0 && op.nfx == 1; }
// first block of rows of unit (j0, r0, r1): source address and bytes
template <int CB>
__device__ __forceinline__ void pot_unit_block(const DevPlan &P, const SumOp &op, const PotStage &st, int j0, int r0, int r1, const double *&src,
                                               uint32_t &bytes) {
    src = P.pot + op.pot_off + ((int64_t)(j0 / CB) * op.rows + r0) * CB;
    bytes = (uint32_t)min(r1 - r0, (int)(st.buf_bytes / (CB * 8))) * (CB * 8);
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
template <int NG, int UMASK, int CB, int POT, bool TMA>   // POT: 0 = all ones, 1 = plain (TMA: staged through shared memory, else coalesced load + shuffle), 2 = folded
__device__ __forceinline__ void sum_rows(const DevPlan &P, const SumOp &op, const Fx &fx, const double *S, int lane, int j0, int r0, int r1,
                                         const int32_t (&cp)[NG > 0 ? NG : 1][CB], V (&acc)[CB], PotStage &st, const double *next_src,
                                         uint32_t next_bytes) {
    constexpr bool FOLD = POT == 2;
    constexpr bool STAGE = POT == 1 && CB >= 2 && TMA;
    // plain potentials without staging: every lane loads the row's CB values itself (same address in all lanes: one L1 broadcast
    // per 16 bytes, requested with the row's gathers) instead of one coalesced load per 32 / CB rows + two shuffles per value
    constexpr bool BCAST = POT == 1 && !STAGE && CB >= 2 && FBN_POT_BCAST;
    constexpr int NGC = NG > 0 ? NG : 1;
    constexpr int NU = (UMASK & 1) + ((UMASK >> 1) & 1) + ((UMASK >> 2) & 1) + ((UMASK >> 3) & 1);
    constexpr int kLoads = (NG > 0 ? (NG - NU) * CB + NU : 0) * kW + (POT == 2 ? CB * kW / 2 : 0);  // 16-byte units per row
    constexpr bool kTwoSets = NG > 0 && CB >= 4 && kLoads <= FBN_TWOSET_MAX;
    const int ng = NG >= 0 ? NG : op.ng, cols = op.cols;
    constexpr bool has_pot = POT != 0;
    const double *__restrict__ pot = P.pot + (has_pot ? op.pot_off : 0);
    const int32_t *__restrict__ rowpart = P.idx32 + op.rowpart_off;
    const int32_t *__restrict__ colpart = P.idx32 + op.colpart_off;
    // POT 1, cb 1: ONE coalesced load serves 32 rows, the row loop picks its value by shuffle
    constexpr int RPL = 32 / CB;
    const double *potl = pot + (int64_t)(j0 / CB) * op.rows * CB * (POT == 2 ? op.nfx : 1);   // the unit's block of potX (unit-major)
    double potv = 1.0, potv_next = (POT == 1 && !STAGE && !BCAST) ? __ldg(potl + min((int64_t)r0 * CB + lane, (int64_t)r1 * CB - 1)) : 1.0;
    // POT 1, cb >= 2: blocks of `sb` rows come in through shared memory (PotStage above); the copy of rows [r0, r0 + sb) was
    // started by the caller
    const int sb = STAGE ? (int)(st.buf_bytes / (CB * 8)) : 0;
    int sb_begin = r0, sb_end = STAGE ? min(r1, r0 + sb) : r1;
    const double *pb = nullptr;
    auto next_block = [&]() {  // rows [sb_begin, sb_end) are about to be summed: start the copy after them, wait for theirs
        if (sb_end < r1) pot_issue(st, potl + (int64_t)sb_end * CB, (uint32_t)min(r1 - sb_end, sb) * (CB * 8), lane);
        else if (next_src) pot_issue(st, next_src, next_bytes, lane);
        pb = pot_wait(st);
    };
    if (STAGE) next_block();
    int jc[CB];
#pragma unroll
    for (int c = 0; c < CB; ++c) jc[c] = min(j0 + c, cols - 1);
    const int64_t nfx = op.nfx;

    struct Set {
        V m[NGC][CB];
        double pv[POT == 2 ? CB : 1][kW];
        double pc[BCAST ? CB : 1];
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
            const double *q = potl + (int64_t)min(rr, r1 - 1) * CB * nfx;
#pragma unroll
            for (int c = 0; c < CB; ++c)
#pragma unroll
                for (int i = 0; i < kW; ++i) s.pv[FOLD ? c : 0][i] = __ldg(q + c * nfx + fx.x[i]);
        }
        if (BCAST) {
            const double2 *q = reinterpret_cast<const double2 *>(potl + (int64_t)min(rr, r1 - 1) * CB);
#pragma unroll
            for (int c = 0; c < CB; c += 2) {
                const double2 t = __ldg(q + c / 2);
                s.pc[BCAST ? c : 0] = t.x;
                s.pc[BCAST ? c + 1 : 0] = t.y;
            }
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
        if (STAGE) {
            if (r == sb_end) {
                sb_begin = sb_end;
                sb_end = min(r1, sb_begin + sb);
                next_block();
            }
            const double *q = pb + (r - sb_begin) * CB;
#pragma unroll
            for (int c = 0; c < CB; c += 2) {
                const double2 t = *reinterpret_cast<const double2 *>(q + c);  // LDS.128, same address in all lanes
                pc[c] = t.x;
                pc[c + (CB > 1 ? 1 : 0)] = t.y;
            }
        } else if (BCAST) {
#pragma unroll
            for (int c = 0; c < CB; ++c) pc[c] = s.pc[BCAST ? c : 0];
        } else if (POT == 1) {
            const int rg = (r - r0) & (RPL - 1);  // row inside its group of RPL
            if (rg == 0) {
                potv = potv_next;
                potv_next = __ldg(potl + min((int64_t)(r + RPL) * CB + lane, (int64_t)r1 * CB - 1));
            }
#pragma unroll
            for (int c = 0; c < CB; ++c) pc[c] = __shfl_sync(0xffffffffu, potv, rg * CB + c);
        } else if (POT == 0) {
#pragma unroll
            for (int c = 0; c < CB; ++c) pc[c] = 1.0;
        }
        V u = vset(1.0);
        if (NG > 0) {
#pragma unroll
            for (int g = 0; g < NG; ++g)
                if ((UMASK >> g) & 1) vmul(u, s.m[g][0]);
        }
        if constexpr (NG < 0) {
            // more than four gathers per entry (cliques with more than three or four neighbours): gather by gather, one index load
            // per row and one per target, 128-bit row loads (both cases of a lane at once), four loads in flight
            V w[CB];
#pragma unroll
            for (int c = 0; c < CB; ++c)
#pragma unroll
                for (int i = 0; i < kW; ++i) w[c].x[i] = (FOLD ? s.pv[FOLD ? c : 0][i] : pc[c]);
            for (int g = 0; g < ng; ++g) {
                const int32_t rpg = __ldg(rowpart + (int64_t)r * ng + g);
                V m[CB];
#pragma unroll
                for (int c = 0; c < CB; ++c) m[c] = ldv(row_ptr(S, rpg + __ldg(colpart + (int64_t)jc[c] * ng + g)));
#pragma unroll
                for (int c = 0; c < CB; ++c) vmul(w[c], m[c]);
            }
#pragma unroll
            for (int c = 0; c < CB; ++c) vadd(acc[c], w[c]);
        } else {
#pragma unroll
            for (int i = 0; i < kW; ++i) {
#pragma unroll
                for (int c = 0; c < CB; ++c) {
                    double v = (FOLD ? s.pv[FOLD ? c : 0][i] : pc[c]) * u.x[i];
                    if (NG > 0) {
#pragma unroll
                        for (int g = 0; g < NG; ++g)
                            if (!((UMASK >> g) & 1)) v *= s.m[g][c].x[i];
                    }
                    acc[c].x[i] += v;
                }
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

template <int NG, int CB, bool TMA>
__device__ __forceinline__ void sum_block(const DevPlan &P, const SumOp &op, const Fx &fx, const double *S, int lane, int j0, int r0, int r1,
                                          V (&acc)[CB], PotStage &st, const double *next_src, uint32_t next_bytes) {
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
        if (op.pot_off < 0) sum_rows<NG, UM, CB, 0, TMA>(P, op, fx, S, lane, j0, r0, r1, cp, acc, st, next_src, next_bytes);
        else if (op.nfx > 1) sum_rows<NG, UM, CB, 2, TMA>(P, op, fx, S, lane, j0, r0, r1, cp, acc, st, next_src, next_bytes);
        else sum_rows<NG, UM, CB, 1, TMA>(P, op, fx, S, lane, j0, r0, r1, cp, acc, st, next_src, next_bytes);
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

template <int CB, bool TMA>
__device__ __forceinline__ void sum_block_any(const DevPlan &P, const SumOp &op, const Fx &fx, const double *S, int lane, int j0, int r0,
                                              int r1, V (&acc)[CB], PotStage &st, const double *next_src, uint32_t next_bytes) {
    switch (op.ng) {
        case 0: sum_block<0, CB, TMA>(P, op, fx, S, lane, j0, r0, r1, acc, st, next_src, next_bytes); break;
        case 1: sum_block<1, CB, TMA>(P, op, fx, S, lane, j0, r0, r1, acc, st, next_src, next_bytes); break;
        case 2: sum_block<2, CB, TMA>(P, op, fx, S, lane, j0, r0, r1, acc, st, next_src, next_bytes); break;
        case 3: sum_block<3, CB, TMA>(P, op, fx, S, lane, j0, r0, r1, acc, st, next_src, next_bytes); break;
        case 4: sum_block<4, CB, TMA>(P, op, fx, S, lane, j0, r0, r1, acc, st, next_src, next_bytes); break;
        default: sum_block<-1, CB, TMA>(P, op, fx, S, lane, j0, r0, r1, acc, st, next_src, next_bytes); break;
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
        const int v = __ldg(cmk + k), want = __ldg(cmk + op.ncmask + j * op.ncmask + k);
#pragma unroll
        for (int i = 0; i < kW; ++i) {
            const int pv = pat[v * kTileCases + lane * kW + i];
            if (pv != 0 && pv != want) acc.x[i] = 0.0;
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
    // the op's scale, then the evidence checks of uncovered target variables (finish_target), variable by variable: a case's
    // pattern bytes of a checked variable are read once per unit, not once per target
    V a[CB];
#pragma unroll
    for (int c = 0; c < CB; ++c) {
        a[c] = acc[c];
        vmul(a[c], scale);
    }
    const int32_t *__restrict__ cmk = P.idx32 + op.cmask_off;
    for (int k = 0; k < op.ncmask; ++k) {
        const int v = __ldg(cmk + k);
        int pv[kW];
#pragma unroll
        for (int i = 0; i < kW; ++i) pv[i] = pat[v * kTileCases + lane * kW + i];
#pragma unroll
        for (int c = 0; c < CB; ++c) {
            const int want = __ldg(cmk + op.ncmask + min(j0 + c, op.cols - 1) * op.ncmask + k);
#pragma unroll
            for (int i = 0; i < kW; ++i)
                if (pv[i] != 0 && pv[i] != want) a[c].x[i] = 0.0;
        }
    }
#pragma unroll
    for (int c = 0; c < CB; ++c)
        if (c < nc) {
            const int j = j0 + c;
            if (op.chunks > 1) {
                stv(row_ptr(S, P.part_row0 + op.pslot0 + j * op.chunks + ck), a[c]);
            } else {
                stv(row_ptr(S, op.dst_row + j), a[c]);
            }
            vadd(nacc, a[c]);
        }
    return nacc;
}

// All units u = first, first + stride, ... < units of one op for one warp: the potentials of unit k+1 are on their way into
// shared memory while unit k is summed. Returns this warp's share of the op's total (a collect op's norm).
template <int CB, bool TMA>
__device__ __forceinline__ V sum_units(const DevPlan &P, const SumOp &op, double *S, const uint8_t *pat, int lane, int first, int stride,
                                       PotStage &st) {
    const V scale = op_scale(P, op, S);
    const Fx fx = op_fx(P, op, pat, lane);
    const int units = (op.cols + CB - 1) / CB * op.chunks;
    const bool staged = TMA && CB >= 2 && pot_staged(op);
    V nacc = vset(0.0);
    int j0, nc, ck, r0, r1;
    const double *src = nullptr;
    uint32_t bytes = 0;
    if (first < units) {
        sum_unit<CB>(P, op, first, j0, nc, ck, r0, r1);
        if (staged) {
            pot_unit_block<CB>(P, op, st, j0, r0, r1, src, bytes);
            pot_issue(st, src, bytes, lane);
        }
    }
    for (int u = first; u < units; u += stride) {
        int j0n = 0, ncn = 0, ckn = 0, r0n = 0, r1n = 0;
        src = nullptr;
        if (u + stride < units) {
            sum_unit<CB>(P, op, u + stride, j0n, ncn, ckn, r0n, r1n);
            if (staged) pot_unit_block<CB>(P, op, st, j0n, r0n, r1n, src, bytes);
        }
        V acc[CB];
#pragma unroll
        for (int c = 0; c < CB; ++c) acc[c] = vset(0.0);
        sum_block_any<CB, TMA>(P, op, fx, S, lane, j0, r0, r1, acc, st, src, bytes);
        vadd(nacc, store_unit<CB>(P, op, S, pat, lane, j0, nc, ck, scale, acc));
        j0 = j0n; nc = ncn; ck = ckn; r0 = r0n; r1 = r1n;
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
