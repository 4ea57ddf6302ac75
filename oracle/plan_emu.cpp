// TEST INFRASTRUCTURE ONLY — scalar CPU interpreter of the device plan (fastbn_b200/csrc/plan.h) so that the
// planner's views, index maps, phases, split columns and finalisers can be unit-tested in the no-GPU container
// (`pytest -m "not gpu"`). It mirrors the kernels one lane at a time: every op is a SumOp over a view of a clique,
// entries are recomputed from pristine potentials and message rows, nothing else is stored. Never shipped, never timed,
// never reachable from fastbn_b200/: built into oracle/_build/libplan_emu.so by tests only.
#include <algorithm>
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
                       const int32_t *ev_vals, int32_t *labels, double *post, int64_t *info /* 16 */) {
    HostPlan P;
    const bool level = team < 0;  // negative: level mode planned for -team gang warps (fastbn_b200/csrc/level_kernels.cu)
    if (level) team = -team;
    g_err = build_plan(*tree, team, P, level);
    if (!g_err.empty()) return 1;
    if (info) {
        info[0] = (int64_t)P.phases.size(); info[1] = P.n_units; info[2] = (int64_t)P.fins.size();
        info[3] = P.max_pslots; info[4] = P.max_normslots; info[5] = P.tile_rows; info[6] = P.n_levels; info[7] = P.parent_entries;
        info[8] = P.view_entries; info[9] = P.gather_rows; info[10] = P.flat ? 1 : 0; info[11] = (int64_t)P.pot.size();
        info[12] = P.folded_vars; info[13] = P.lambda_gathers;
    }
    const int kColBlock = P.cb;
    const int shares = level ? P.nb_max : team;   // norm shares per collect op
    const bool use_flat = P.flat;  // team 1 fused plans of small trees carry flat records: interpret those
    std::vector<double> S((size_t)P.tile_rows);
    std::vector<int> pat((size_t)P.n_vars);   // evidence pattern of every variable: 0 not observed, 1 + state, dim + 1 = no valid state
    for (int64_t cs = 0; cs < n_cases; ++cs) {
        std::fill(S.begin(), S.end(), -12345.0);  // poison: reading a row before writing it shows up in the results
        std::fill(pat.begin(), pat.end(), 0);
        for (int64_t k = ev_off[cs]; k < ev_off[cs + 1]; ++k) {
            int v = ev_vars[k], val = ev_vals[k];
            if (v < 0 || v >= P.n_vars) continue;
            const int dim = P.vars[v].dim;
            const int np = (val < 0 || val >= dim) ? dim + 1 : 1 + val;
            pat[v] = (pat[v] == 0 || pat[v] == np) ? np : dim + 1;  // two different states observed: nothing is consistent
        }
        for (int v : P.lamvars)  // lambda rows of the variables some op gathers
            for (int d = 0; d < P.vars[v].dim; ++d) S[P.lam_row0 + P.vars[v].off + d] = (pat[v] == 0 || pat[v] == 1 + d) ? 1.0 : 0.0;
        for (const Phase &ph : P.phases) {
            std::vector<double> normshare((size_t)std::max(1, ph.n_normslots) * shares, 0.0);
            std::vector<double> part((size_t)std::max(1, P.max_pslots), -777.0);
            for (int oi = ph.op_begin; oi < ph.op_end; ++oi) {
                const SumOp &op = P.ops[P.oplist[oi]];
                const int units = (op.cols + kColBlock - 1) / kColBlock * op.chunks;
                double scale = 1.0;
                for (int k = 0; k < op.ninv; ++k) scale *= S[P.idx32[op.invrow_off + k]];
                int64_t fx = 0;  // this case's evidence pattern of the op's folded variables
                for (int k = 0; k < op.nfold; ++k) fx += (int64_t)pat[P.idx32[op.fold_off + 2 * k]] * P.idx32[op.fold_off + 2 * k + 1];
                const int rstride = 1 + (op.ng > 3 ? 1 : 0);  // int4 per flat record
                for (int u = 0; u < units; ++u)
                    for (int cc = 0; cc < kColBlock; ++cc) {  // the targets of the unit's block
                        int jb = u / op.chunks;
                        const int ck = u - jb * op.chunks;
                        if (op.perm_off >= 0) jb = P.idx32[op.perm_off + jb];  // visiting order
                        const int j = jb * kColBlock + cc;
                        if (j >= op.cols) break;
                        const int share = (op.rot + u) % shares;
                        int r0 = 0, r1 = op.rows;
                        if (op.chunks > 1) { r0 = ck * op.rows_per_chunk; r1 = std::min(op.rows, r0 + op.rows_per_chunk); }
                        double acc = 0.0;
                        for (int r = r0; r < r1; ++r) {
                            double v;
                            if (use_flat) {  // the flat records of jt_flat_kernel (target-major, absolute gather rows)
                                const int64_t x = op.flat_off + ((int64_t)j * op.rows + r) * 4 * rstride;
                                v = P.pot[P.flatc[x] + fx];
                                for (int g = 0; g < op.ng; ++g) v *= S[P.flatc[x + 1 + g]];  // {pot index, g0, g1, g2} {g3, g4, g5, g6}
                            } else {
                                v = op.pot_off < 0 ? 1.0 : P.pot[op.pot_off + (((int64_t)(j / kColBlock) * op.rows + r) * kColBlock + j % kColBlock) * op.nfx + fx];
                                for (int g = 0; g < op.ng; ++g)
                                    v *= S[P.idx32[op.rowpart_off + (int64_t)r * op.ng + g] + P.idx32[op.colpart_off + (int64_t)j * op.ng + g]];
                            }
                            acc += v;
                        }
                        acc *= scale;
                        for (int k = 0; k < op.ncmask; ++k) {  // uncovered target variables of a collect / distribute view
                            const int32_t *cmk = &P.idx32[op.cmask_off];
                            const int pv = pat[cmk[k]], want = cmk[op.ncmask + (int64_t)j * op.ncmask + k];
                            if (pv != 0 && pv != want) acc = 0.0;
                        }
                        if (op.chunks > 1) {
                            part[op.pslot0 + j * op.chunks + ck] = acc;
                        } else {
                            S[op.dst_row + j] = acc;
                        }
                        if (op.kind == SUM_COLLECT) normshare[(size_t)op.normslot * shares + share] += acc;
                    }
            }
            for (int f = ph.fin_begin; f < ph.fin_end; ++f) {
                const FinItem &F = P.fins[f];
                double s = 0.0;
                if (F.type == FIN_NORM) {
                    for (int w = 0; w < shares; ++w) s += normshare[(size_t)F.slot0 * shares + w];
                    S[F.dst_row] = 1.0 / s;
                    continue;
                }
                for (int i = 0; i < F.nslots; ++i) s += part[F.slot0 + i];
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
                double pr = pat[v] ? 0.0 : S[P.mg_row0 + vi.off + d] / sum;
                if (post) post[cs * P.sum_dims + vi.off + d] = pr;
                if (pr > best) { best = pr; arg = d; }
            }
            if (v == P.query_var && labels) labels[cs] = arg;
        }
    }
    return 0;
}

// ---- resident-tile plan (fastbn_b200/csrc/resident.h): interprets the per-phase constant blocks exactly as
//      jt_resident_kernel does (items warp by warp), one case at a time. The state is poisoned before every case and sized to
//      the plan's live-row high-water mark, so a wrong live range shows up in the results.
#include "../fastbn_b200/csrc/resident.h"

extern "C" int emu_run_resident(const fbn_tree_desc *tree, int warps, int64_t smem_limit, int64_t n_cases, const int64_t *ev_off,
                                const int32_t *ev_vars, const int32_t *ev_vals, int32_t *labels, double *post, int64_t *info /* 16 */) {
    HostPlan P;
    g_err = build_plan(*tree, 1, P, false);
    if (!g_err.empty()) return 1;
    ResidentPlan R;
    build_resident(P, warps, smem_limit, R);
    if (info) {
        info[0] = R.ok ? 1 : 0; info[1] = R.state_rows; info[2] = R.max_block_bytes; info[3] = R.smem_bytes; info[4] = R.n_blocks;
        info[5] = R.n_items; info[6] = (int64_t)(R.balance * 1000); info[7] = P.tile_rows; info[8] = (int64_t)R.blob.size() * 4;
        info[9] = (int64_t)P.pot.size();
    }
    if (!R.ok) { g_err = R.why; return 2; }
    const int rowb = kResRowBytes;
    std::vector<double> S((size_t)R.state_rows);
    std::vector<int> pat((size_t)P.n_vars);
    int64_t n_evid_items = 0;
    for (int64_t cs = 0; cs < n_cases; ++cs) {
        std::fill(S.begin(), S.end(), -12345.0);
        std::fill(pat.begin(), pat.end(), 0);
        for (int64_t k = ev_off[cs]; k < ev_off[cs + 1]; ++k) {
            int v = ev_vars[k], val = ev_vals[k];
            if (v < 0 || v >= P.n_vars) continue;
            const int dim = P.vars[v].dim;
            const int np = (val < 0 || val >= dim) ? dim + 1 : 1 + val;
            pat[v] = (pat[v] == 0 || pat[v] == np) ? np : dim + 1;
        }
        std::vector<char> seen((size_t)P.n_vars, 0);
        for (const ResLam &l : R.lams)
            for (uint32_t d = 0; d < l.dim; ++d) S[l.state_off / rowb + d] = (pat[l.var] == 0 || pat[l.var] == (int)(1 + d)) ? 1.0 : 0.0;
        for (int b = 0; b < R.n_blocks; ++b) {
            const uint32_t *blk = R.blob.data() + R.block_off[(size_t)b] / 4;
            const unsigned char *bytes = reinterpret_cast<const unsigned char *>(blk);
            if (blk[3] != R.block_off[(size_t)b + 1] - R.block_off[(size_t)b]) { g_err = "block size mismatch"; return 3; }
            const ResItem *items = reinterpret_cast<const ResItem *>(bytes + blk[2]);
            // a block's items run concurrently on the GPU: none may read what another one of the same block writes. Writes go
            // to a copy that replaces the state at the block's barrier.
            std::vector<double> W(S);
            for (int w = 0; w < warps; ++w)
                for (uint32_t it = blk[4 + w]; it < blk[4 + w + 1]; ++it) {
                    const ResItem &I = items[it];
                    const uint32_t kind = I.kind_ng & 255u, ng = (I.kind_ng >> 8) & 255u;
                    if (kind == RES_EVID) { if (cs == 0) ++n_evid_items; continue; }
                    if (kind != RES_OUT && kind != RES_SUM && kind != RES_SUMC) { g_err = "unknown item kind"; return 6; }
                    if (kind != RES_OUT && (I.kind_ng >> 16) != (1u << (ng < 5 ? ng : 5))) { g_err = "item without its body bit"; return 6; }
                    if (kind == RES_OUT) {
                        const int v = (int)I.rec_off, dim = (int)I.n_targets, off = (int)I.rows;
                        const double *mg = &S[I.dst_off / rowb];
                        double sum = 0.0;
                        for (int d = 0; d < dim; ++d) sum += mg[d];
                        double best = 0.0;
                        int arg = -1;
                        for (int d = 0; d < dim; ++d) {
                            double pr = pat[v] ? 0.0 : mg[d] / sum;
                            if (post) post[cs * P.sum_dims + off + d] = pr;
                            if (pr > best) { best = pr; arg = d; }
                        }
                        if (v == P.query_var && labels) labels[cs] = arg;
                        seen[(size_t)v] = 1;
                        continue;
                    }
                    const uint32_t *fd = reinterpret_cast<const uint32_t *>(bytes + I.desc_off);
                    // the item repeats what the kernel needs of the descriptor (resident.h): both must say the same, and the
                    // emulator reads what the kernel reads (the item, unless the op folds more than two variables)
                    const uint32_t ncm = fd[1], *cm = fd + 4 + 2 * fd[0];
                    if ((I.nfold_ncm & 255u) != fd[0] || (I.nfold_ncm >> 8) != ncm) { g_err = "item and descriptor disagree on the counts"; return 8; }
                    const uint32_t ifv[2] = {I.fold_var0, I.fold_var1}, ifs[2] = {I.fold_stride0, I.fold_stride1};
                    for (uint32_t k = 0; k < 2; ++k) {
                        if (k < fd[0] ? (ifv[k] != fd[4 + 2 * k] || ifs[k] != fd[5 + 2 * k]) : (ifs[k] != 0 || ifv[k] >= (uint32_t)P.n_vars)) { g_err = "item and descriptor disagree on the folded variables"; return 8; }
                    }
                    for (uint32_t k = 0; k < (uint32_t)kResMaxChecks; ++k)
                        if (k < ncm ? I.chk[k] != cm[k] : I.chk[k] >= (uint32_t)P.n_vars) { g_err = "item and descriptor disagree on the checked variables"; return 8; }
                    if (I.want_off != (ncm ? fd[2] + I.j0 * ncm * 4u : 0u)) { g_err = "item and descriptor disagree on the want words"; return 8; }
                    uint32_t fx = 0;
                    if (fd[0] <= 2) fx = (uint32_t)pat[I.fold_var0] * I.fold_stride0 + (uint32_t)pat[I.fold_var1] * I.fold_stride1;
                    else for (uint32_t k = 0; k < fd[0]; ++k) fx += (uint32_t)pat[fd[4 + 2 * k]] * fd[5 + 2 * k];
                    const uint32_t *want = reinterpret_cast<const uint32_t *>(bytes + I.want_off) - (size_t)I.j0 * ncm;   // indexed by the op's target below
                    if ((ncm != 0) != (kind == RES_SUMC)) { g_err = "item kind does not match the op's checks"; return 7; }
                    const uint32_t stride = ng > 3 ? 8 : 4;
                    const uint32_t *rec = reinterpret_cast<const uint32_t *>(bytes + I.rec_off);
                    for (uint32_t j = 0; j < I.n_targets; ++j) {
                        double acc = 0.0;
                        for (uint32_t r = 0; r < I.rows; ++r) {
                            const uint32_t *q = rec + ((size_t)j * I.rows + r) * stride;
                            double v;
                            memcpy(&v, bytes + q[0] + fx, 8);
                            for (uint32_t g = 0; g < ng; ++g) v *= S[q[1 + g] / rowb];
                            acc += v;
                        }
                        for (uint32_t k = 0; k < ncm; ++k) {   // checked target variables: zero where the case observed another state
                            const int pv = pat[I.chk[k]];
                            if (pv != 0 && pv != (int)want[(size_t)(I.j0 + j) * ncm + k]) acc = 0.0;
                        }
                        W[I.dst_off / rowb + j] = acc;
                    }
                }
            S.swap(W);
        }
        for (int v = 0; v < P.n_vars; ++v)
            if (!seen[(size_t)v]) { g_err = "variable without an OUT item"; return 4; }
    }
    if (n_cases > 0 && n_evid_items != 1) { g_err = "schedule must hold exactly one EVID item"; return 5; }
    return 0;
}
