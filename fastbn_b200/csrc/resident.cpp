// Host-side planner of the resident-tile kernel: flat plan -> live-range row allocation + static per-warp schedule + one
// constant block per phase (resident.h). Replaces, like plan.cpp, what the reference recomputes per case and per level
// (src/JunctionTree.cpp:586-639, :834-880); the ops themselves are the SumOps of plan.h.
#include "resident.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <numeric>

namespace fbn {
namespace {

struct Obj {
    int first_row = 0, rows = 0;   // rows of the flat plan
    int birth = 0, death = 0;      // phases
    int new_row = -1;
};

struct Piece {
    int kind;      // ResItemKind
    int op;        // SUM: op index; OUT: variable
    int j0, j1;    // SUM: targets [j0, j1)
    double cost;
};

// instruction-count model of the kernel's loops (two cases per lane): only the ratios matter
double sum_piece_cost(const SumOp &op, int targets) {
    // measured on B200 (ncu, ALARM): ~100 instructions per item, ~20 per record with two gathers, ~10 per checked target variable
    const double setup = 100.0 + 12.0 * op.nfold + 6.0 * op.ncmask;
    const double per_target = 8.0 + 10.0 * op.ncmask;
    const double per_entry = 10.0 + 5.0 * op.ng + (op.ng > 3 ? 1.0 : 0.0);
    return setup + targets * (per_target + per_entry * op.rows);
}
double out_cost(int dim) { return 180.0 + 20.0 * dim; }
const double kEvidCost = getenv("FBN_RES_EVID_COST") ? atof(getenv("FBN_RES_EVID_COST")) : 1200.0;   // developer knob for sweeps

}  // namespace

void build_resident(const HostPlan &P, int warps, int64_t smem_limit, ResidentPlan &R) {
    R = ResidentPlan();
    R.warps = warps;
    auto fail = [&](const std::string &w) { R.ok = false; R.why = w; };
    if (!P.flat) return fail("not a flat plan");
    if (warps < 1 || warps > 32) return fail("warps out of range");
    for (int v = 0; v < P.n_vars; ++v)
        if (P.vars[v].dim > 254) return fail("variable with more than 254 states");
    // ---- can Normalize wait for the final marginals? Every message / marginal entry is a sum of products of one entry per
    //      clique of a subtree: when it is not zero it lies between prod(min positive entry) and prod(sum of entries) ----
    {
        if ((int)P.clique_off.size() != P.n_cliques + 1) return fail("plan without clique offsets");
        double lo = 0.0, hi = 0.0;
        for (int c = 0; c < P.n_cliques; ++c) {
            double mn = 0.0, sum = 0.0;
            for (int64_t e = P.clique_off[(size_t)c]; e < P.clique_off[(size_t)c + 1]; ++e) {
                const double x = P.pot[(size_t)e];
                if (!(x >= 0.0) || !std::isfinite(x)) return fail("negative or non-finite potential");
                if (x > 0.0 && (mn == 0.0 || x < mn)) mn = x;
                sum += x;
            }
            if (mn > 0.0) lo += std::min(0.0, std::log10(mn));
            if (sum > 0.0) hi += std::max(0.0, std::log10(sum));
        }
        if (lo < -280.0 || hi > 280.0) return fail("unnormalised messages could leave the range of a double");
    }
    const int n_ph = (int)P.phases.size();
    std::vector<int> op_phase(P.ops.size(), -1);
    for (int p = 0; p < n_ph; ++p)
        for (int oi = P.phases[p].op_begin; oi < P.phases[p].op_end; ++oi) op_phase[P.oplist[oi]] = p;

    // ---- objects: every op's targets ----
    std::vector<Obj> objs;
    std::vector<int> row2obj((size_t)P.tile_rows, -1), obj_op;
    // lambda rows (evidence indicators of the variables some view gathers): written when a tile's patterns are known, i.e.
    // during the previous tile's last block; they keep their rows for good
    for (int v : P.lamvars) {
        Obj ob;
        ob.first_row = P.lam_row0 + P.vars[(size_t)v].off; ob.rows = P.vars[(size_t)v].dim; ob.birth = -1; ob.death = 1 << 30;
        for (int r = ob.first_row; r < ob.first_row + ob.rows; ++r) row2obj[(size_t)r] = (int)objs.size();
        ob.new_row = R.state_rows;
        R.state_rows += ob.rows;
        R.lams.push_back({(uint32_t)v, (uint32_t)ob.rows, (uint32_t)ob.new_row * (uint32_t)kResRowBytes, 0u});
        objs.push_back(ob);
        obj_op.push_back(-1);
    }
    for (size_t o = 0; o < P.ops.size(); ++o) {
        const SumOp &op = P.ops[o];
        if (op_phase[o] < 0) return fail("op outside every phase");
        if (op.chunks != 1 || op.ng > 7 || op.flat_off < 0) return fail("op without flat records");
        if (op.ncmask > kResMaxChecks) return fail("view checks more than " + std::to_string(kResMaxChecks) + " target variables against the evidence");
        for (int r = op.dst_row; r < op.dst_row + op.cols; ++r) {
            if (r < 0 || r >= P.tile_rows || row2obj[(size_t)r] != -1) return fail("two ops write one row");
            row2obj[(size_t)r] = (int)objs.size();
        }
        Obj ob;
        ob.first_row = op.dst_row; ob.rows = op.cols; ob.birth = ob.death = op_phase[o];
        if (op.kind == SUM_MARGINAL) ob.death = op_phase[o] + 1;   // read by the OUT item of the next block
        objs.push_back(ob);
        obj_op.push_back((int)o);
    }
    // ---- needed ops (marginals and everything they read, transitively) and liveness; readers run in later phases than writers ----
    std::vector<char> needed(P.ops.size(), 0);
    for (size_t o = 0; o < P.ops.size(); ++o) needed[o] = P.ops[o].kind == SUM_MARGINAL;
    for (int p = n_ph - 1; p >= 0; --p)
        for (int oi = P.phases[p].op_begin; oi < P.phases[p].op_end; ++oi) {
            const int o = P.oplist[oi];
            if (!needed[(size_t)o]) continue;
            const SumOp &op = P.ops[(size_t)o];
            const int stride = op.ng > 3 ? 8 : 4;
            for (int64_t e = 0; e < (int64_t)op.rows * op.cols; ++e)
                for (int g = 0; g < op.ng; ++g) {
                    const int row = P.flatc[(size_t)(op.flat_off + e * stride + 1 + g)];
                    if (row < 0 || row >= P.tile_rows || row2obj[(size_t)row] < 0) return fail("op reads a row nobody writes");
                    Obj &ob = objs[(size_t)row2obj[(size_t)row]];
                    if (obj_op[(size_t)row2obj[(size_t)row]] < 0) continue;   // a lambda row
                    if (ob.birth >= p) return fail("op reads a row no earlier phase wrote");
                    ob.death = std::max(ob.death, p);
                    needed[(size_t)obj_op[(size_t)row2obj[(size_t)row]]] = 1;
                }
        }
    // ---- best-fit allocation, phase by phase: births of a phase first, then the deaths of the phase are released ----
    {
        std::map<int, int> free_list;   // first row -> rows
        int high = R.state_rows;   // above the lambda rows
        auto release = [&](int first, int rows) {
            auto it = free_list.emplace(first, rows).first;
            auto nx = std::next(it);
            if (nx != free_list.end() && it->first + it->second == nx->first) { it->second += nx->second; free_list.erase(nx); }
            if (it != free_list.begin()) {
                auto pv = std::prev(it);
                if (pv->first + pv->second == it->first) { pv->second += it->second; free_list.erase(it); }
            }
        };
        for (int p = 0; p <= n_ph; ++p) {
            std::vector<int> born;
            for (size_t i = 0; i < objs.size(); ++i)
                if (objs[i].birth == p && obj_op[i] >= 0 && needed[(size_t)obj_op[i]]) born.push_back((int)i);
            std::stable_sort(born.begin(), born.end(), [&](int a, int b) { return objs[(size_t)a].rows > objs[(size_t)b].rows; });
            for (int i : born) {
                Obj &ob = objs[(size_t)i];
                auto best = free_list.end();
                for (auto it = free_list.begin(); it != free_list.end(); ++it)
                    if (it->second >= ob.rows && (best == free_list.end() || it->second < best->second)) best = it;
                if (best != free_list.end()) {
                    ob.new_row = best->first;
                    const int rest = best->second - ob.rows, at = best->first + ob.rows;
                    free_list.erase(best);
                    if (rest > 0) free_list.emplace(at, rest);
                } else {
                    ob.new_row = high;
                    high += ob.rows;
                }
            }
            for (size_t i = 0; i < objs.size(); ++i)
                if (objs[i].death == p && obj_op[i] >= 0 && objs[i].new_row >= 0) release(objs[i].new_row, objs[i].rows);
        }
        R.state_rows = high;
        if (getenv("FBN_RES_DEBUG"))
            for (int p = 0; p <= n_ph; ++p) {
                int up = 0, dn = 0, mg = 0;
                for (size_t i = 0; i < objs.size(); ++i) {
                    if (obj_op[i] < 0 || !needed[(size_t)obj_op[i]] || objs[i].birth > p || objs[i].death < p) continue;
                    const int k = P.ops[(size_t)obj_op[i]].kind;
                    (k == SUM_COLLECT ? up : k == SUM_DISTRIBUTE ? dn : mg) += objs[i].rows;
                }
                fprintf(stderr, "phase %2d live rows: up %3d dn %3d mg %3d total %3d\n", p, up, dn, mg, up + dn + mg);
            }
    }
    auto state_off = [&](int old_row) {
        const Obj &ob = objs[(size_t)row2obj[(size_t)old_row]];
        return (uint32_t)(ob.new_row + (old_row - ob.first_row)) * (uint32_t)kResRowBytes;
    };

    // ---- blocks = phases with needed ops (+ one last block for the OUT items of the last phase) ----
    std::vector<std::vector<int>> blk_ops, blk_outs;   // per block: ops summed in it; marginal ops written out by it
    {
        std::vector<int> pending;   // marginal ops of the previous block
        for (int p = 0; p < n_ph; ++p) {
            std::vector<int> ops;
            for (int oi = P.phases[p].op_begin; oi < P.phases[p].op_end; ++oi)
                if (needed[(size_t)P.oplist[oi]]) ops.push_back(P.oplist[oi]);
            if (ops.empty() && pending.empty()) continue;
            // an empty phase between two phases cannot hold marginals that are still being summed: OUT items read rows of the
            // previous BLOCK, whose barrier has passed
            blk_ops.push_back(ops);
            blk_outs.push_back(pending);
            pending.clear();
            for (int o : ops)
                if (P.ops[(size_t)o].kind == SUM_MARGINAL) pending.push_back(o);
        }
        blk_ops.push_back({});
        blk_outs.push_back(pending);
    }
    // the live ranges above are in PHASES; blocks follow the phase order and an OUT item runs in the block after its marginal
    // op's, i.e. not later than phase + 1 of the range. (A skipped empty phase only shortens the distance.)
    R.n_blocks = (int)blk_ops.size();
    // a CTA barrier per phase: deep trees with little to sum per level (a chain of small cliques) are better off with one warp
    // per tile (jt_flat_kernel). ALARM: 166 view entries per phase
    const int64_t kMinWork = getenv("FBN_RES_MIN_WORK") ? atoll(getenv("FBN_RES_MIN_WORK")) : 48;   // tests force small nets through with 0
    if (P.view_entries < kMinWork * R.n_blocks) return fail("too little work per phase for a CTA: " + std::to_string(P.view_entries) + " view entries in " + std::to_string(R.n_blocks) + " phases");
    std::vector<std::vector<Piece>> pieces((size_t)R.n_blocks);
    std::vector<double> work((size_t)R.n_blocks, 0.0);
    for (int b = 0; b < R.n_blocks; ++b) {
        for (int o : blk_ops[(size_t)b]) work[(size_t)b] += sum_piece_cost(P.ops[(size_t)o], P.ops[(size_t)o].cols);
        for (int o : blk_outs[(size_t)b]) work[(size_t)b] += out_cost(P.ops[(size_t)o].cols);
    }
    // not the last block: the next tile's lambda rows are written there, from the patterns the EVID item has loaded
    const int evid_block = (int)(std::max_element(work.begin(), work.end() - 1) - work.begin());
    work[(size_t)evid_block] += kEvidCost;
    const double kGrain = getenv("FBN_RES_GRAIN") ? atof(getenv("FBN_RES_GRAIN")) : 200.0;   // measured on B200 (ALARM, 24 warps): 60 -> 264, 120 -> 264, 200 -> 268 M cases/s
    for (int b = 0; b < R.n_blocks; ++b) {
        const double grain = std::max(kGrain, work[(size_t)b] / warps / 2.0);   // aimed cost of a piece
        for (int o : blk_ops[(size_t)b]) {
            const SumOp &op = P.ops[(size_t)o];
            int k = (int)std::min<double>(op.cols, std::max(1.0, sum_piece_cost(op, op.cols) / grain + 0.5));
            for (int i = 0; i < k; ++i) {
                const int j0 = (int)((int64_t)op.cols * i / k), j1 = (int)((int64_t)op.cols * (i + 1) / k);
                if (j1 > j0) pieces[(size_t)b].push_back({RES_SUM, o, j0, j1, sum_piece_cost(op, j1 - j0)});
            }
        }
        for (int o : blk_outs[(size_t)b]) {
            const SumOp &op = P.ops[(size_t)o];
            int v = -1;
            for (int x = 0; x < P.n_vars; ++x)
                if (P.mg_row0 + P.vars[(size_t)x].off == op.dst_row) v = x;
            if (v < 0 || P.vars[(size_t)v].dim != op.cols) return fail("marginal op without a variable");
            pieces[(size_t)b].push_back({RES_OUT, v, 0, 0, out_cost(op.cols)});
        }
        if (b == evid_block) pieces[(size_t)b].push_back({RES_EVID, 0, 0, 0, kEvidCost});
    }

    // ---- the blocks' bytes ----
    double total_cost = 0, total_max = 0;
    R.block_off.assign(1, 0);
    for (int b = 0; b < R.n_blocks; ++b) {
        std::vector<Piece> &pc = pieces[(size_t)b];
        std::stable_sort(pc.begin(), pc.end(), [](const Piece &a, const Piece &c) { return a.cost > c.cost; });
        std::vector<double> load((size_t)warps, 0.0);
        std::vector<std::vector<int>> mine((size_t)warps);
        for (size_t i = 0; i < pc.size(); ++i) {   // longest processing time first
            const int w = (int)(std::min_element(load.begin(), load.end()) - load.begin());
            load[(size_t)w] += pc[i].cost;
            mine[(size_t)w].push_back((int)i);
            total_cost += pc[i].cost;
        }
        total_max += *std::max_element(load.begin(), load.end());

        const std::vector<int> &bops = blk_ops[(size_t)b];
        const uint32_t hdr_words = (uint32_t)((4 + warps + 1 + 3) & ~3);
        const uint32_t items_off = hdr_words * 4;
        const uint32_t n_items = (uint32_t)pc.size();
        uint32_t cur = items_off + n_items * (uint32_t)sizeof(ResItem);
        std::map<int, uint32_t> desc_off, rec_off, pot_off, tab_off;
        for (int o : bops) {
            desc_off[o] = cur;
            cur += ((4u + 2u * (uint32_t)P.ops[(size_t)o].nfold + (uint32_t)P.ops[(size_t)o].ncmask + 3u) & ~3u) * 4u;
        }
        for (int o : bops) {
            const SumOp &op = P.ops[(size_t)o];
            rec_off[o] = cur;
            cur += (uint32_t)op.rows * (uint32_t)op.cols * (op.ng > 3 ? 32u : 16u);
        }
        const uint32_t one_off = cur;   // a 1.0 for the views without potentials
        cur += 16;
        for (int o : bops) {
            const SumOp &op = P.ops[(size_t)o];
            if (op.pot_off < 0) { pot_off[o] = one_off; continue; }
            pot_off[o] = cur;
            cur += (((uint32_t)op.rows * (uint32_t)op.cols * (uint32_t)op.nfx * 8u) + 15u) & ~15u;
        }
        for (int o : bops) {   // per target: the pattern byte it wants of every checked variable
            const SumOp &op = P.ops[(size_t)o];
            tab_off[o] = cur;
            cur += (((uint32_t)op.cols * (uint32_t)op.ncmask * 4u) + 15u) & ~15u;
        }
        const uint32_t block_bytes = cur;
        if (block_bytes >= (1u << 20)) return fail("constant block of a phase exceeds the mbarrier transaction range");
        std::vector<uint32_t> blk(block_bytes / 4, 0u);
        blk[0] = n_items; blk[1] = 0; blk[2] = items_off; blk[3] = block_bytes;
        {
            uint32_t at = 0;
            ResItem *items = reinterpret_cast<ResItem *>(blk.data() + items_off / 4);
            for (int w = 0; w < warps; ++w) {
                blk[(size_t)(4 + w)] = at;
                for (int i : mine[(size_t)w]) {
                    const Piece &q = pc[(size_t)i];
                    ResItem it = ResItem();
                    if (q.kind == RES_SUM) {
                        const SumOp &op = P.ops[(size_t)q.op];
                        it.kind_ng = (op.ncmask ? RES_SUMC : RES_SUM) | ((uint32_t)op.ng << 8) | (1u << (16 + std::min(op.ng, 5)));   // bits 16..21: the kernel body (by gathers per entry)
                        it.rec_off = rec_off[q.op] + (uint32_t)q.j0 * (uint32_t)op.rows * (op.ng > 3 ? 32u : 16u);
                        it.n_targets = (uint32_t)(q.j1 - q.j0);
                        it.rows = (uint32_t)op.rows;
                        it.dst_off = state_off(op.dst_row) + (uint32_t)q.j0 * (uint32_t)kResRowBytes;
                        it.desc_off = desc_off[q.op];
                        it.j0 = (uint32_t)q.j0;
                        it.want_off = op.ncmask ? tab_off[q.op] + (uint32_t)q.j0 * (uint32_t)op.ncmask * 4u : 0u;
                        it.nfold_ncm = (uint32_t)op.nfold | ((uint32_t)op.ncmask << 8);
                        if (op.nfold > 0) { it.fold_var0 = (uint32_t)P.idx32[(size_t)op.fold_off]; it.fold_stride0 = (uint32_t)P.idx32[(size_t)op.fold_off + 1] * 8u; }
                        if (op.nfold > 1) { it.fold_var1 = (uint32_t)P.idx32[(size_t)op.fold_off + 2]; it.fold_stride1 = (uint32_t)P.idx32[(size_t)op.fold_off + 3] * 8u; }
                        for (int k = 0; k < kResMaxChecks; ++k) it.chk[k] = op.ncmask ? (uint32_t)P.idx32[(size_t)(op.cmask_off + std::min(k, op.ncmask - 1))] : 0u;
                    } else if (q.kind == RES_OUT) {
                        const VarInfo vi = P.vars[(size_t)q.op];
                        it.kind_ng = RES_OUT;
                        it.rec_off = (uint32_t)q.op;
                        it.n_targets = (uint32_t)vi.dim;
                        it.rows = (uint32_t)vi.off;
                        it.dst_off = state_off(P.mg_row0 + vi.off);
                    } else {
                        it.kind_ng = RES_EVID;
                    }
                    items[at++] = it;
                }
            }
            blk[(size_t)(4 + warps)] = at;
        }
        for (int o : bops) {
            const SumOp &op = P.ops[(size_t)o];
            uint32_t *d = blk.data() + desc_off[o] / 4;
            d[0] = (uint32_t)op.nfold; d[1] = (uint32_t)op.ncmask;
            for (int i = 0; i < op.nfold; ++i) {
                d[4 + 2 * i] = (uint32_t)P.idx32[(size_t)(op.fold_off + 2 * i)];
                d[5 + 2 * i] = (uint32_t)P.idx32[(size_t)(op.fold_off + 2 * i + 1)] * 8u;
            }
            d[2] = tab_off[o];
            for (int k = 0; k < op.ncmask; ++k) d[4 + 2 * op.nfold + k] = (uint32_t)P.idx32[(size_t)(op.cmask_off + k)];
            for (int64_t i = 0; i < (int64_t)op.cols * op.ncmask; ++i)
                blk[(size_t)(tab_off[o] / 4 + i)] = (uint32_t)P.idx32[(size_t)(op.cmask_off + op.ncmask + i)];
            uint32_t *rec = blk.data() + rec_off[o] / 4;
            const int stride = op.ng > 3 ? 8 : 4;
            for (int64_t e = 0; e < (int64_t)op.rows * op.cols; ++e) {
                const int32_t *src = &P.flatc[(size_t)(op.flat_off + e * stride)];
                uint32_t *q = rec + e * stride;
                q[0] = op.pot_off < 0 ? one_off : pot_off[o] + (uint32_t)(src[0] - op.pot_off) * 8u;
                for (int g = 0; g < stride - 1; ++g) q[1 + g] = g < op.ng ? state_off(src[1 + g]) : 0u;
            }
            if (op.pot_off >= 0)
                memcpy(blk.data() + pot_off[o] / 4, P.pot.data() + op.pot_off, (size_t)op.rows * op.cols * op.nfx * 8);
        }
        {
            const double one = 1.0;
            memcpy(blk.data() + one_off / 4, &one, 8);
        }
        R.blob.insert(R.blob.end(), blk.begin(), blk.end());
        R.block_off.push_back((uint32_t)(R.blob.size() * 4));
        R.max_block_bytes = std::max(R.max_block_bytes, (int)block_bytes);
        R.n_items += (int)n_items;
    }
    R.balance = total_max > 0 ? total_cost / (total_max * warps) : 1.0;
    const ResSmem L = res_smem_layout(R.state_rows, P.n_vars, R.max_block_bytes, R.n_blocks, (int)R.lams.size());
    R.smem_bytes = L.total;
    if (R.smem_bytes > smem_limit)
        return fail("live messages of a tile need " + std::to_string(R.smem_bytes) + " bytes of shared memory, the device has " + std::to_string(smem_limit));
    R.ok = true;
}

}  // namespace fbn
