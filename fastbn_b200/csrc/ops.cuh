// Device-side building blocks shared by the fused kernel (kernels.cu) and the level kernels (level_kernels.cu).
// Thread mapping everywhere: lane = test case inside a 32-case tile; a warp owns a UNIT = a block of up to CB
// adjacent columns (collect) / target entries (pull) of one table times a range of rows. Register blocking over the
// columns gives CB independent 256-byte row loads per index load (memory-level parallelism) and amortises the
// warp-uniform index loads; all scratch accesses are coalesced 256-byte rows.
#pragma once
#include <cstdint>
#include <type_traits>
#include <cuda_runtime.h>

#include "plan.h"

namespace fbn {

// rest entries per batch of the QB = 2 PULL loop (fused kernel, >= 2 warps per tile); 1 = plain loop
#ifndef FBN_PULL_BATCH
#define FBN_PULL_BATCH 8
#endif
// two register sets in the COLLECT row loop of the CB = 2 (fused, >= 2 warps per tile) kernels too
#ifndef FBN_TWOSETS_CB2
#define FBN_TWOSETS_CB2 0
#endif
// unroll factor of the single-register-set COLLECT row loop of the CB = 2 kernels
#ifndef FBN_COLLECT_UNROLL_CB2
#define FBN_COLLECT_UNROLL_CB2 2
#endif
constexpr int kCollectUnrollCB2 = FBN_COLLECT_UNROLL_CB2;

__device__ __forceinline__ double *row_ptr(double *S, int32_t row) { return S + (int64_t)row * kLanes; }

// Cache policy of the two big streams (clique entries written by COLLECT, read once per child by PULL). FBN_STREAM_HINTS
// marks them evict-first so that the small, heavily re-used rows (messages, ratios, index maps, pristine potentials)
// keep the L2.
#ifndef FBN_STREAM_HINTS
#define FBN_STREAM_HINTS 1
#endif
template <bool STREAM>
__device__ __forceinline__ void store_clique(double *p, double v) {
    if (STREAM && FBN_STREAM_HINTS) __stcs(p, v);
    else *p = v;
}
template <bool STREAM>
__device__ __forceinline__ double load_clique(const double *p) {
    if (STREAM && FBN_STREAM_HINTS) return __ldcs(p);
    return *p;
}

// ---- row evidence masks -----------------------------------------------------------------------------------------
// mask word of clique row r: bit l set <=> case l of the tile allows that row (AND of the lambda words of the row's digits)
struct RowMaskTable {  // fused mode, tables built once per phase in shared memory
    static constexpr int kSpan = 1 << 30;  // rows per begin(): the whole range
    const uint32_t *tab;
    __device__ __forceinline__ void begin(int, int, int) {}
    __device__ __forceinline__ uint32_t get(int r, int) const { return tab[r]; }
};
struct RowMaskLam {  // evaluated on the fly, lane-parallel: lane i builds the word of row rb+i, get() broadcasts it
    static constexpr int kSpan = 32;
    const uint32_t *lam;       // lambda words of the tile (shared or global memory)
    const uint16_t *rowlam;    // op's rows x nrest lambda indices
    int nrest;
    uint32_t mine;
    __device__ __forceinline__ void begin(int rb, int r1, int lane) {
        mine = 0xffffffffu;
        if (rb + lane < r1) {
            const uint16_t *__restrict__ rl = rowlam + (int64_t)(rb + lane) * nrest;
            for (int i = 0; i < nrest; ++i) mine &= lam[__ldg(rl + i)];
        }
    }
    __device__ __forceinline__ uint32_t get(int r, int rb) const { return __shfl_sync(0xffffffffu, mine, r - rb); }
};

// ---- COLLECT -----------------------------------------------------------------------------------------------------
// Columns [j0, j0+nc) (nc <= CB), rows [r0, r1) of one clique:
//   C[e] = pristine[e] * scale * mask(e) * prod_k message_k[rowpart_k(r) + colpart_k(j)];  acc[c] += C[r, j0+c]
// NC = number of children (compile-time 0..3, -1 = generic). Loads of columns beyond nc are clamped to the last column
// (valid addresses, results discarded), so the inner loop has no divergent control flow.
// Row loop of collect_block. UMASK (compile time): children whose message row is the same for all CB columns of the block
// (one gather per row, folded into the row's scalar factor).
// Software pipeline: the gathers of row r+1 are issued BEFORE row r is multiplied, masked and stored (two register sets,
// used alternately), so the memory latency of a row overlaps the arithmetic and the stores of the previous one; the row parts
// of the child maps are fetched two rows ahead, the coalesced pristine-potential load one group ahead. Rows that need more
// than 9 gathers keep one register set (gathers issued together at the top of the row). The clique row pointer advances by
// a 64-bit add, stores go to immediate offsets c * 256 B.
template <int NC, int UMASK, int CB, bool STREAM, class RowMask>
__device__ __forceinline__ void collect_rows(const DevPlan &P, const CollectOp &op, RowMask &rmask, const uint32_t (&cm)[CB], double *S,
                                             int lane, int j0, int nc, int r0, int r1, double scale, const int32_t (&cp)[NC > 0 ? NC : 1][CB],
                                             double (&acc)[CB]) {
    constexpr int NCC = NC > 0 ? NC : 1;
    constexpr int NU = (UMASK & 1) + ((UMASK >> 1) & 1) + ((UMASK >> 2) & 1);
    constexpr int kGathers = NC > 0 ? (NC - NU) * CB + NU : 0;  // per row
    constexpr bool kTwoSets = NC > 0 && ((CB >= 4 && kGathers <= 9) || (FBN_TWOSETS_CB2 != 0 && CB == 2 && kGathers <= 6));
    const int nchild = NC >= 0 ? NC : op.nchild, cols = op.cols;
    const double *__restrict__ pot = P.pot + op.pot_off;
    const int32_t *__restrict__ rowpart = P.idx32 + op.rowpart_off;
    const int32_t *__restrict__ colpart = P.idx32 + op.colpart_off;
    const int64_t rstride = (int64_t)cols * kLanes;
    double *clp = row_ptr(S, op.cl_row) + ((int64_t)r0 * cols + j0) * kLanes;
    const uint32_t lanebit = 1u << lane;
    uint32_t colon = 0;  // bit c: this lane's case allows column j0+c
#pragma unroll
    for (int c = 0; c < CB; ++c)
        if (cm[c] & lanebit) colon |= 1u << c;
    // pristine potentials: ONE coalesced load serves RPL rows x CB columns (lane = row-in-group * CB + column), the row
    // loop picks its values by shuffle; a warp-uniform 8-byte load per entry would cost a load instruction (and a
    // scoreboard stall) per entry
    constexpr int RPL = 32 / CB;
    const int lane_c = lane % CB, lane_r = lane / CB;
    const double *potl = pot + min(j0 + lane_c, cols - 1);
    double potv = 0.0, potv_next = __ldg(potl + (int64_t)min(r0 + lane_r, r1 - 1) * cols);
    int rb = r0;

    auto load_parts = [&](int rr, int32_t (&rp)[NCC]) {
        if (NC > 0) {
            const int32_t *q = rowpart + (int64_t)min(rr, r1 - 1) * NC;
#pragma unroll
            for (int k = 0; k < NC; ++k) rp[k] = __ldg(q + k);
        }
    };
    auto gather = [&](const int32_t (&rp)[NCC], double (&m)[NCC][CB]) {
        if (NC > 0) {
#pragma unroll
            for (int k = 0; k < NC; ++k) {
                const double *mb = S + (int64_t)rp[k] * kLanes;
                if ((UMASK >> k) & 1) {
                    m[k][0] = mb[(int64_t)cp[k][0] * kLanes];
                } else {
#pragma unroll
                    for (int c = 0; c < CB; ++c) m[k][c] = mb[(int64_t)cp[k][c] * kLanes];
                }
            }
        }
    };
    auto finish = [&](int r, const double (&m)[NCC][CB]) {
        if (RowMask::kSpan <= 32 && ((r - r0) & 31) == 0) {
            rb = r;
            rmask.begin(rb, r1, lane);
        }
        const int rg = (r - r0) & (RPL - 1);  // row inside its group of RPL
        if (rg == 0) {
            potv = potv_next;
            potv_next = __ldg(potl + (int64_t)min(r + RPL + lane_r, r1 - 1) * cols);
        }
        const uint32_t on = (rmask.get(r, rb) & lanebit) ? colon : 0u;
        double u = scale;
        if (NC > 0) {
#pragma unroll
            for (int k = 0; k < NC; ++k)
                if ((UMASK >> k) & 1) u *= m[k][0];
        }
        double v[CB];
#pragma unroll
        for (int c = 0; c < CB; ++c) v[c] = __shfl_sync(0xffffffffu, potv, rg * CB + c) * u;
        if (NC > 0) {
#pragma unroll
            for (int k = 0; k < NC; ++k)
                if (!((UMASK >> k) & 1)) {
#pragma unroll
                    for (int c = 0; c < CB; ++c) v[c] *= m[k][c];
                }
        } else if (NC < 0) {
            for (int k = 0; k < nchild; ++k) {
                const int32_t rpk = __ldg(rowpart + (int64_t)r * nchild + k);
#pragma unroll
                for (int c = 0; c < CB; ++c) v[c] *= *row_ptr(S, rpk + __ldg(colpart + (int64_t)min(j0 + c, cols - 1) * nchild + k));
            }
        }
#pragma unroll
        for (int c = 0; c < CB; ++c) {
            const double w = ((on >> c) & 1u) ? v[c] : 0.0;
            if (c < nc) store_clique<STREAM>(clp + c * kLanes, w);
            acc[c] += w;
        }
        clp += rstride;
    };

    if (RowMask::kSpan > 32) rmask.begin(r0, r1, lane);
    if constexpr (kTwoSets) {
        int32_t rp0[NCC], rp1[NCC];
        double m0[NCC][CB], m1[NCC][CB];
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
        int32_t rp[NCC], rpn[NCC];
        double m[NCC][CB];
        load_parts(r0, rpn);
#pragma unroll(CB >= 4 ? 1 : kCollectUnrollCB2)
        for (int r = r0; r < r1; ++r) {
#pragma unroll
            for (int k = 0; k < NCC; ++k) rp[k] = rpn[k];
            gather(rp, m);
            load_parts(r + 1, rpn);
            finish(r, m);
        }
    }
}

template <int NC, int CB, bool STREAM, class RowMask>
__device__ __forceinline__ void collect_block(const DevPlan &P, const CollectOp &op, RowMask &rmask, const uint32_t (&cm)[CB], double *S,
                                              int lane, int j0, int nc, int r0, int r1, double scale, double (&acc)[CB]) {
    const int nchild = NC >= 0 ? NC : op.nchild, cols = op.cols;
    if constexpr (CB == 1 && RowMask::kSpan > 32) {
        // one column, mask tables in shared memory: the small-tree path (ALARM-size, one warp per tile). Kept as the plain
        // loop — with 2-4 rows per column the blocking / shuffle / prefetch machinery below costs more than it saves.
        const int32_t *__restrict__ cpart = P.idx32 + op.colpart_off + (int64_t)j0 * nchild;
        int32_t cp1[NC > 0 ? NC : 1];
        if (NC > 0) {
#pragma unroll
            for (int k = 0; k < NC; ++k) cp1[k] = __ldg(cpart + k);
        }
        const int32_t *__restrict__ rp = P.idx32 + op.rowpart_off + (int64_t)r0 * nchild;
        const double *__restrict__ pot1 = P.pot + op.pot_off + j0;
        double *cl1 = row_ptr(S, op.cl_row + j0);
        double a = 0.0;
#pragma unroll 2
        for (int r = r0; r < r1; ++r, rp += nchild) {
            const uint32_t m = cm[0] & rmask.get(r, r0);
            const int64_t e = (int64_t)r * cols;
            double v = __ldg(pot1 + e) * scale;
            if (NC > 0) {
#pragma unroll
                for (int k = 0; k < NC; ++k) v *= *row_ptr(S, __ldg(rp + k) + cp1[k]);
            } else if (NC < 0) {
                for (int k = 0; k < nchild; ++k) v *= *row_ptr(S, __ldg(rp + k) + __ldg(cpart + k));
            }
            v = ((m >> lane) & 1u) ? v : 0.0;
            store_clique<STREAM>(cl1 + e * kLanes, v);
            a += v;
        }
        acc[0] += a;
        return;
    }
    int jc[CB];
#pragma unroll
    for (int c = 0; c < CB; ++c) jc[c] = min(j0 + c, cols - 1);
    int32_t cp[NC > 0 ? NC : 1][CB];
    const int32_t *__restrict__ colpart = P.idx32 + op.colpart_off;
    // umask bit k: all CB columns of the block read the SAME message row of child k (the block's column variables are
    // outside separator k) -> one gather per row instead of CB, folded into the row's scalar factor
    int umask = 0;
    if (NC > 0) {
#pragma unroll
        for (int k = 0; k < NC; ++k) {
            bool u = true;
#pragma unroll
            for (int c = 0; c < CB; ++c) {
                cp[k][c] = __ldg(colpart + (int64_t)jc[c] * NC + k);
                u = u && cp[k][c] == cp[k][0];
            }
            if (u) umask |= 1 << k;
        }
    }
    if constexpr (NC > 0 && NC <= 2 && CB >= 4) {
        switch (umask) {
            case 0: collect_rows<NC, 0, CB, STREAM>(P, op, rmask, cm, S, lane, j0, nc, r0, r1, scale, cp, acc); break;
            case 1: collect_rows<NC, 1, CB, STREAM>(P, op, rmask, cm, S, lane, j0, nc, r0, r1, scale, cp, acc); break;
            case 2: collect_rows<NC, 2 % (1 << NC), CB, STREAM>(P, op, rmask, cm, S, lane, j0, nc, r0, r1, scale, cp, acc); break;
            default: collect_rows<NC, (1 << NC) - 1, CB, STREAM>(P, op, rmask, cm, S, lane, j0, nc, r0, r1, scale, cp, acc); break;
        }
    } else if constexpr (NC == 3 && CB >= 4) {
        if (umask == 7) collect_rows<NC, 7, CB, STREAM>(P, op, rmask, cm, S, lane, j0, nc, r0, r1, scale, cp, acc);
        else collect_rows<NC, 0, CB, STREAM>(P, op, rmask, cm, S, lane, j0, nc, r0, r1, scale, cp, acc);
    } else {
        collect_rows<NC, 0, CB, STREAM>(P, op, rmask, cm, S, lane, j0, nc, r0, r1, scale, cp, acc);
    }
}

template <int CB, bool STREAM, class RowMask>
__device__ __forceinline__ void collect_block_any(const DevPlan &P, const CollectOp &op, RowMask &rmask, const uint32_t (&cm)[CB],
                                                  double *S, int lane, int j0, int nc, int r0, int r1, double scale, double (&acc)[CB]) {
    switch (op.nchild) {
        case 0: collect_block<0, CB, STREAM>(P, op, rmask, cm, S, lane, j0, nc, r0, r1, scale, acc); break;
        case 1: collect_block<1, CB, STREAM>(P, op, rmask, cm, S, lane, j0, nc, r0, r1, scale, acc); break;
        case 2: collect_block<2, CB, STREAM>(P, op, rmask, cm, S, lane, j0, nc, r0, r1, scale, acc); break;
        case 3: collect_block<3, CB, STREAM>(P, op, rmask, cm, S, lane, j0, nc, r0, r1, scale, acc); break;
        default: collect_block<-1, CB, STREAM>(P, op, rmask, cm, S, lane, j0, nc, r0, r1, scale, acc); break;
    }
}

// unit u of a collect op -> column block + row range
template <int CB>
__device__ __forceinline__ void collect_unit(const DevPlan &P, const CollectOp &op, int u, int &j0, int &nc, int &ck, int &r0, int &r1) {
    int jb = u;
    ck = 0; r0 = 0; r1 = op.rows;
    if (op.chunks > 1) {
        jb = u / op.chunks;
        ck = u - jb * op.chunks;
        r0 = ck * op.rows_per_chunk;
        r1 = min(op.rows, r0 + op.rows_per_chunk);
    }
    if (op.perm_off >= 0) jb = __ldg(P.idx32 + op.perm_off + jb);  // planner's visiting order of the column blocks
    j0 = jb * CB;
    nc = min(CB, op.cols - j0);
}

// ---- PULL --------------------------------------------------------------------------------------------------------
// targets [q0, q0+nq) (nq <= QB), rest entries [r0, r1):  acc[q] += src[tbase[q].x + trest[r].x] * fac[tbase[q].y + trest[r].y]
template <int QB, bool STREAM>
__device__ __forceinline__ void pull_block(const DevPlan &P, const PullOp &op, double *S, int q0, int r0, int r1, double (&acc)[QB]) {
    const int2 *__restrict__ tb = reinterpret_cast<const int2 *>(P.idx32 + op.tbase_off);
    const int2 *__restrict__ tr = reinterpret_cast<const int2 *>(P.idx32 + op.trest_off);
    if constexpr (QB == 1) {  // small-tree path: plain loop, four rest entries in flight
        const int2 b = __ldg(tb + q0);
        const double *src1 = row_ptr(S, op.src_row + b.x), *fac1 = row_ptr(S, op.fac_row + b.y);
        double a = 0.0;
#pragma unroll 4
        for (int r = r0; r < r1; ++r) {
            const int2 t = __ldg(tr + r);
            a += load_clique<STREAM>(src1 + (int64_t)t.x * kLanes) * fac1[(int64_t)t.y * kLanes];
        }
        acc[0] += a;
        return;
    }
    const double *src[QB], *fac[QB];
    bool uni = true;  // all QB targets share the factor row (their variables are outside the source's column variables)
#pragma unroll
    for (int q = 0; q < QB; ++q) {
        const int2 b = __ldg(tb + min(q0 + q, op.nT - 1));
        src[q] = row_ptr(S, op.src_row + b.x);
        fac[q] = row_ptr(S, op.fac_row + b.y);
        uni = uni && fac[q] == fac[0];
    }
    if constexpr (QB == 2 && FBN_PULL_BATCH > 1) {
        // fused kernel, several warps per tile: batches of RB rest entries. The RB index entries come first (L1-resident,
        // warp-uniform), then ALL data loads of the batch (RB x QB source rows + RB or RB x QB factor rows) are issued before
        // the first multiply-add; the sums run in rest order (same rounding as the plain loop). One DRAM latency per batch
        // instead of one per dependent index -> address -> data chain (measured on B200: RB 4 / 6 / 8 / 12 / 16 -> PIGS-shape
        // 423 / 428 / 440 / 434 / 435 k cases/s; fetching the next batch's indices ahead costs 8 registers and loses).
        constexpr int RB = FBN_PULL_BATCH;
        int r = r0;
        auto batch = [&](auto uni_c, auto n_c) {  // N rest entries starting at r
            constexpr bool UNI = decltype(uni_c)::value;
            constexpr int N = decltype(n_c)::value, NF = UNI ? 1 : QB;
            int2 t[N];
#pragma unroll
            for (int i = 0; i < N; ++i) t[i] = __ldg(tr + r + i);
            double s[N][QB], f[N][NF];
#pragma unroll
            for (int i = 0; i < N; ++i) {
#pragma unroll
                for (int q = 0; q < QB; ++q) s[i][q] = load_clique<STREAM>(src[q] + (int64_t)t[i].x * kLanes);
#pragma unroll
                for (int q = 0; q < NF; ++q) f[i][q] = fac[q][(int64_t)t[i].y * kLanes];
            }
#pragma unroll
            for (int i = 0; i < N; ++i) {
#pragma unroll
                for (int q = 0; q < QB; ++q) acc[q] += s[i][q] * f[i][UNI ? 0 : q];
            }
            r += N;
        };
        auto batches = [&](auto uni_c) {
            while (r + RB <= r1) batch(uni_c, std::integral_constant<int, RB>{});
            // the last < RB rest entries in batches of 4, 2, 1 (table sizes are rarely multiples of RB)
            if constexpr (RB > 4) {
                while (r + 4 <= r1) batch(uni_c, std::integral_constant<int, 4>{});
            }
            if (r + 2 <= r1) batch(uni_c, std::integral_constant<int, 2>{});
            if (r + 2 <= r1) batch(uni_c, std::integral_constant<int, 2>{});
            if (r < r1) batch(uni_c, std::integral_constant<int, 1>{});
        };
        if (uni) batches(std::true_type{});
        else batches(std::false_type{});
        return;
    }
    int2 t_next = __ldg(tr + min(r0, r1 - 1));  // index prefetch, as in collect_block
    if (uni) {
#pragma unroll(QB >= 4 ? 2 : 4)
        for (int r = r0; r < r1; ++r) {
            const int2 t = t_next;
            t_next = __ldg(tr + min(r + 1, r1 - 1));
            const double f = fac[0][(int64_t)t.y * kLanes];
#pragma unroll
            for (int q = 0; q < QB; ++q) acc[q] += load_clique<STREAM>(src[q] + (int64_t)t.x * kLanes) * f;
        }
        return;
    }
#pragma unroll(QB >= 4 ? 2 : 4)
    for (int r = r0; r < r1; ++r) {
        const int2 t = t_next;
        t_next = __ldg(tr + min(r + 1, r1 - 1));
#pragma unroll
        for (int q = 0; q < QB; ++q) acc[q] += load_clique<STREAM>(src[q] + (int64_t)t.x * kLanes) * fac[q][(int64_t)t.y * kLanes];
    }
}

// ---- PULL2 (level mode) ----------------------------------------------------------------------------------------------
// Unit u = (G12 configuration a, block b of QB configurations of G1) of a Pull2Op (plan.h): one pass over the G2
// configurations gives QB complete entries of separator A and one partial row per G2 configuration for separator B.
template <int QB>
__device__ __forceinline__ void pull2_unit(const DevPlan &P, const Pull2Op &op, double *S, int u) {
    const int4 *__restrict__ ta = reinterpret_cast<const int4 *>(P.idx32 + op.a_off);
    const int4 *__restrict__ t1 = reinterpret_cast<const int4 *>(P.idx32 + op.g1_off);
    const int4 *__restrict__ t2 = reinterpret_cast<const int4 *>(P.idx32 + op.g2_off);
    const int a = u / op.nblk1, b = u - a * op.nblk1;
    const int4 A = __ldg(ta + a);
    const int g0 = b * QB, nq = min(QB, op.n1 - g0);
    const double *src[QB], *fac[QB];
    int32_t qa[QB];
    double acc[QB];
#pragma unroll
    for (int c = 0; c < QB; ++c) {
        const int4 G = __ldg(t1 + min(g0 + c, op.n1 - 1));
        src[c] = row_ptr(S, op.src_row + A.x + G.x);
        fac[c] = row_ptr(S, op.fac_row + A.y + G.y);
        qa[c] = A.z + G.z;
        acc[c] = 0.0;
    }
    double *part = row_ptr(S, P.part_row0 + op.pslot0) + ((int64_t)A.w * op.nblk1 + b) * kLanes;
    const int64_t pstride = (int64_t)op.nblk1 * kLanes;
    // software pipeline, two register sets: the 2*QB loads of configuration r+1 are issued before configuration r is
    // multiplied, accumulated and its partial row stored; the index entries are fetched two configurations ahead
    auto loads = [&](const int4 &R, double (&sv)[QB], double (&fv)[QB]) {
#pragma unroll
        for (int c = 0; c < QB; ++c) {
            sv[c] = load_clique<true>(src[c] + (int64_t)R.x * kLanes);
            fv[c] = fac[c][(int64_t)R.y * kLanes];
        }
    };
    auto finish = [&](const int4 &R, const double (&sv)[QB], const double (&fv)[QB]) {
        double prod[QB];
#pragma unroll
        for (int c = 0; c < QB; ++c) {
            prod[c] = sv[c] * fv[c];
            acc[c] += prod[c];
            if (c >= nq) prod[c] = 0.0;  // clamped duplicates of a ragged last block do not belong to the partial
        }
        // fixed pairwise order (deterministic)
#pragma unroll
        for (int w = 1; w < QB; w *= 2)
#pragma unroll
            for (int c = 0; c + w < QB; c += 2 * w) prod[c] += prod[c + w];
        part[(int64_t)R.z * pstride] = prod[0];
    };
    double s0[QB], f0[QB], s1[QB], f1[QB];
    int4 Ra = __ldg(t2), Rb = __ldg(t2 + min(1, op.n2 - 1));
    loads(Ra, s0, f0);
    for (int r = 0; r < op.n2; r += 2) {
        // here: (s0, f0) = configuration r (in flight), Ra its index entry, Rb the entry of r+1
        loads(Rb, s1, f1);
        const int4 Rc = __ldg(t2 + min(r + 2, op.n2 - 1));
        finish(Ra, s0, f0);
        if (r + 1 < op.n2) {
            loads(Rc, s0, f0);
            const int4 Rd = __ldg(t2 + min(r + 3, op.n2 - 1));
            finish(Rb, s1, f1);
            Rb = Rd;
        }
        Ra = Rc;
    }
#pragma unroll
    for (int c = 0; c < QB; ++c)
        if (c < nq) {
            const double m = *row_ptr(S, op.mupA_row + qa[c]);
            *row_ptr(S, op.dstA_row + qa[c]) = (m == 0.0) ? 0.0 : acc[c] / m;  // TableDivision 0-guard, PotentialTable.cpp:718-734
        }
}

template <int CB>
__device__ __forceinline__ void pull_unit(const PullOp &op, int u, int &q0, int &nq, int &ck, int &r0, int &r1) {
    int qb = u;
    ck = 0; r0 = 0; r1 = op.nR;
    if (op.chunks > 1) {
        qb = u / op.chunks;
        ck = u - qb * op.chunks;
        r0 = ck * op.rows_per_chunk;
        r1 = min(op.nR, r0 + op.rows_per_chunk);
    }
    q0 = qb * CB;
    nq = min(CB, op.nT - q0);
}

}  // namespace fbn
