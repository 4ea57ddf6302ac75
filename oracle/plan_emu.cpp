// TEST INFRASTRUCTURE ONLY — scalar CPU interpreter of the device plan (fastbn_b200/csrc/plan.h) so that the
// planner's index maps, phases, split columns and finalisers can be unit-tested in the no-GPU container
// (`pytest -m "not gpu"`). It mirrors jt_fused_kernel one lane at a time. Never shipped, never timed, never
// reachable from fastbn_b200/: built into oracle/_build/libplan_emu.so by tests only.
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "../fastbn_b200/csrc/plan.h"
#include "../include/fastbn_b200.h"

using namespace fbn;

static std::string g_err;

extern "C" const char *emu_last_error() { return g_err.c_str(); }

extern "C" int emu_run(const fbn_tree_desc *tree, int team, int64_t n_cases, const int64_t *ev_off, const int32_t *ev_vars,
                       const int32_t *ev_vals, int32_t *labels, double *post, int64_t *info /* 8 */) {
    HostPlan P;
    const bool level = team < 0;  // negative: level mode planned for -team gang warps (fastbn_b200/csrc/level_kernels.cu)
    if (level) team = -team;
    g_err = build_plan(*tree, team, P, level);
    if (!g_err.empty()) return 1;
    if (info) {
        info[0] = (int64_t)P.phases.size(); info[1] = P.n_units; info[2] = (int64_t)P.fins.size();
        info[3] = P.max_pslots; info[4] = P.max_normslots; info[5] = P.tile_rows; info[6] = P.n_levels; info[7] = P.parent_entries;
    }
    const int kColBlock = P.cb;
    const bool use_flat = P.flat;  // team 1 fused plans of small trees carry flat records: interpret those
    std::vector<double> S((size_t)P.tile_rows);
    std::vector<char> lam((size_t)P.sum_dims), obs((size_t)P.n_vars);
    std::vector<double> normacc((size_t)std::max(1, P.max_normslots) * team), part((size_t)std::max(1, P.max_pslots));
    for (int64_t cs = 0; cs < n_cases; ++cs) {
        std::fill(S.begin(), S.end(), -12345.0);  // poison: reading a row before writing it shows up in the results
        std::fill(lam.begin(), lam.end(), 1);
        std::fill(obs.begin(), obs.end(), 0);
        for (int64_t k = ev_off[cs]; k < ev_off[cs + 1]; ++k) {
            int v = ev_vars[k], val = ev_vals[k];
            if (v < 0 || v >= P.n_vars) continue;
            obs[v] = 1;
            for (int d = 0; d < P.vars[v].dim; ++d)
                if (d != val) lam[P.vars[v].off + d] = 0;
        }
        for (const Phase &ph : P.phases) {
            if (ph.kind == 0) std::fill(normacc.begin(), normacc.end(), 0.0);
            for (int oi = ph.op_begin; oi < ph.op_end; ++oi) {
                const int o = P.oplist[oi];
                if (ph.kind == 1 && o < 0) {  // PULL2: two children of one parent from one pass (plan.h)
                    const Pull2Op &op = P.p2ops[-1 - o];
                    const int32_t *A = &P.idx32[op.a_off], *G1 = &P.idx32[op.g1_off], *G2 = &P.idx32[op.g2_off];
                    for (int a = 0; a < op.n12; ++a)
                        for (int b = 0; b < op.nblk1; ++b) {
                            std::vector<double> acc(kColBlock, 0.0);
                            for (int r = 0; r < op.n2; ++r) {
                                double pb = 0.0;
                                for (int c = 0; c < kColBlock; ++c) {
                                    const int g = b * kColBlock + c;
                                    if (g >= op.n1) break;
                                    const double prod = S[op.src_row + A[4 * a] + G1[4 * g] + G2[4 * r]] *
                                                        S[op.fac_row + A[4 * a + 1] + G1[4 * g + 1] + G2[4 * r + 1]];
                                    acc[c] += prod;
                                    pb += prod;
                                }
                                part[op.pslot0 + (int64_t)(A[4 * a + 3] + G2[4 * r + 2]) * op.nblk1 + b] = pb;
                            }
                            for (int c = 0; c < kColBlock; ++c) {
                                const int g = b * kColBlock + c;
                                if (g >= op.n1) break;
                                const int q = A[4 * a + 2] + G1[4 * g + 2];
                                const double m = S[op.mupA_row + q];
                                S[op.dstA_row + q] = m == 0.0 ? 0.0 : acc[c] / m;
                            }
                        }
                    continue;
                }
                const int ncolsop = ph.kind == 0 ? P.cops[o].cols : P.pops[o].nT;
                const int chunks = ph.kind == 0 ? P.cops[o].chunks : P.pops[o].chunks;
                const int units = (ncolsop + kColBlock - 1) / kColBlock * chunks;
                for (int u = 0; u < units; ++u)
                  for (int cc = 0; cc < kColBlock; ++cc) {  // the columns / targets of the unit's block
                    int jb = u / chunks;
                    const int ck = u - jb * chunks;
                    if (ph.kind == 0 && P.cops[o].perm_off >= 0) jb = P.idx32[P.cops[o].perm_off + jb];  // visiting order
                    if (jb * kColBlock + cc >= ncolsop) break;
                    if (ph.kind == 0) {
                        const CollectOp &op = P.cops[o];
                        const int warp = (op.rot + u) % team;
                        const int j = jb * kColBlock + cc;
                        int r0 = 0, r1 = op.rows;
                        if (op.chunks > 1) { r0 = ck * op.rows_per_chunk; r1 = std::min(op.rows, r0 + op.rows_per_chunk); }
                        bool cm = true;
                        for (int i = 0; i < op.ncolv; ++i) cm = cm && lam[P.lam16[op.collam_off + (int64_t)j * op.ncolv + i]];
                        double scale = 1.0;
                        for (int k = 0; k < op.nchild; ++k) scale *= S[P.idx32[op.invrow_off + k]];
                        double acc = 0.0;
                        for (int r = r0; r < r1; ++r) {
                            bool m = cm;
                            for (int i = 0; i < op.nrest; ++i) m = m && lam[P.lam16[op.rowlam_off + (int64_t)r * op.nrest + i]];
                            const int64_t e = (int64_t)r * op.cols + j;
                            double v;
                            if (use_flat) {  // the flat records of jt_flat_kernel (column-major, absolute message rows)
                                const int64_t x = op.flat_off + ((int64_t)j * op.rows + r) * (op.nchild > 2 ? 8 : 4);
                                double pv;
                                memcpy(&pv, &P.flatc[x], 8);
                                v = pv * scale;
                                for (int k = 0; k < op.nchild; ++k) v *= S[P.flatc[x + 2 + k]];  // {pot, m0, m1} {m2, m3, 0, 0}
                            } else {
                                v = P.pot[op.pot_off + e] * scale;
                                for (int k = 0; k < op.nchild; ++k)
                                    v *= S[P.idx32[op.rowpart_off + (int64_t)r * op.nchild + k] + P.idx32[op.colpart_off + (int64_t)j * op.nchild + k]];
                            }
                            if (!m) v = 0.0;
                            S[op.cl_row + e] = v;
                            acc += v;
                        }
                        if (op.chunks > 1) part[op.pslot0 + j * op.chunks + ck] = acc;
                        else S[op.mup_row + j] = acc;
                        normacc[op.normslot * team + warp] += acc;
                    } else {
                        const PullOp &op = P.pops[o];
                        const int q = jb * kColBlock + cc;
                        int r0 = 0, r1 = op.nR;
                        if (op.chunks > 1) { r0 = ck * op.rows_per_chunk; r1 = std::min(op.nR, r0 + op.rows_per_chunk); }
                        const int32_t bx = P.idx32[op.tbase_off + 2 * q], by = P.idx32[op.tbase_off + 2 * q + 1];
                        double acc = 0.0;
                        for (int r = r0; r < r1; ++r) {
                            if (use_flat) {
                                const int64_t x = op.flat_off + 2 * ((int64_t)q * op.nR + r);
                                acc += S[P.flatp[x]] * S[P.flatp[x + 1]];
                            } else {
                                acc += S[op.src_row + bx + P.idx32[op.trest_off + 2 * r]] * S[op.fac_row + by + P.idx32[op.trest_off + 2 * r + 1]];
                            }
                        }
                        if (op.chunks > 1) part[op.pslot0 + q * op.chunks + ck] = acc;
                        else if (op.kind == 1) { double m = S[op.mup_row + q]; S[op.dst_row + q] = m == 0.0 ? 0.0 : acc / m; }
                        else S[op.dst_row + q] = acc;
                    }
                }
            }
            for (int f = ph.fin_begin; f < ph.fin_end; ++f) {
                const FinItem &F = P.fins[f];
                double s = 0.0;
                if (F.type == FIN_NORM) {
                    for (int w = 0; w < team; ++w) s += normacc[F.slot0 * team + w];
                    S[F.dst_row] = 1.0 / s;
                    continue;
                }
                for (int i = 0; i < F.nslots; ++i) s += part[F.slot0 + i];
                if (F.type == FIN_D) { double m = S[F.aux_row]; s = m == 0.0 ? 0.0 : s / m; }
                S[F.dst_row] = s;
            }
        }
        for (int v = 0; v < P.n_vars; ++v) {
            const VarInfo vi = P.vars[v];
            double sum = 0.0;
            for (int d = 0; d < vi.dim; ++d) sum += S[P.mg_row0 + vi.off + d];
            double best = 0.0;
            int arg = -1;
            for (int d = 0; d < vi.dim; ++d) {
                double pr = obs[v] ? 0.0 : S[P.mg_row0 + vi.off + d] / sum;
                if (post) post[cs * P.sum_dims + vi.off + d] = pr;
                if (pr > best) { best = pr; arg = d; }
            }
            if (v == P.query_var && labels) labels[cs] = arg;
        }
    }
    return 0;
}
