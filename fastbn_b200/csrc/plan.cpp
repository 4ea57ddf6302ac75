// Host-side planner: flat reference tree (fbn_tree_desc) -> per-tree constant arrays + warp work items.
// Replaces, once per tree, what the reference recomputes for every level of every case in the "pre" phases
// of src/JunctionTree.cpp:586-639 and :834-880 (index maps by GetConfigValueByTableIndex /
// GetTableIndexByConfigValue, src/PotentialTableBase.cpp:547-566).
#include "plan.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <string>

#include "../../include/fastbn_b200.h"

namespace fbn {
namespace {

struct Table {
    std::vector<int> vars, dims, stride;
    int64_t size = 1;
    int find(int v) const {
        for (size_t i = 0; i < vars.size(); ++i)
            if (vars[i] == v) return (int)i;
        return -1;
    }
};

// cum_levels of the reference: stride[n-1] = 1, stride[i] = stride[i+1]*dims[i+1] (PotentialTableBase.cpp:599-607)
bool finish_table(Table &t, const int32_t *var_dims, int n_vars, std::string &err) {
    size_t n = t.vars.size();
    t.dims.resize(n);
    t.stride.resize(n);
    t.size = 1;
    for (size_t i = 0; i < n; ++i) {
        if (t.vars[i] < 0 || t.vars[i] >= n_vars) {
            err = "table variable out of range";
            return false;
        }
        for (size_t k = 0; k < i; ++k)
            if (t.vars[k] == t.vars[i]) {
                err = "duplicate variable in a table";
                return false;
            }
        t.dims[i] = var_dims[t.vars[i]];
    }
    for (size_t i = n; i-- > 0;) {
        t.stride[i] = (int)t.size;
        t.size *= t.dims[i];
        if (t.size > (int64_t)1 << 30) {
            err = "table exceeds 2^30 entries (the reference itself is limited to int32 table sizes)";
            return false;
        }
    }
    return true;
}

// For every configuration of `sub` (a list of positions into tab.vars, row-major in the given order) the partial
// entry index sum digit*stride. Enumerated with an odometer, no div/mod.
std::vector<int32_t> enumerate_offsets(const Table &tab, const std::vector<int> &pos) {
    int64_t n = 1;
    for (int p : pos) n *= tab.dims[p];
    std::vector<int32_t> out((size_t)n);
    std::vector<int> dig(pos.size(), 0);
    int32_t cur = 0;
    for (int64_t i = 0; i < n; ++i) {
        out[(size_t)i] = cur;
        for (size_t k = pos.size(); k-- > 0;) {
            int p = pos[k];
            if (++dig[k] < tab.dims[p]) {
                cur += tab.stride[p];
                break;
            }
            cur -= (tab.dims[p] - 1) * tab.stride[p];
            dig[k] = 0;
        }
    }
    return out;
}

// digits of every configuration of `pos` (row-major), flattened n x pos.size()
std::vector<int> enumerate_digits(const Table &tab, const std::vector<int> &pos) {
    int64_t n = 1;
    for (int p : pos) n *= tab.dims[p];
    std::vector<int> out((size_t)n * pos.size());
    std::vector<int> dig(pos.size(), 0);
    for (int64_t i = 0; i < n; ++i) {
        for (size_t k = 0; k < pos.size(); ++k) out[(size_t)i * pos.size() + k] = dig[k];
        for (size_t k = pos.size(); k-- > 0;) {
            if (++dig[k] < tab.dims[pos[k]]) break;
            dig[k] = 0;
        }
    }
    return out;
}

}  // namespace

std::string build_plan(const fbn_tree_desc &t, int team, HostPlan &P, bool level) {
    std::string err;
    const int nv = t.n_vars, nc = t.n_cliques, ns = t.n_separators;
    if (nv <= 0 || nc <= 0 || ns != nc - 1) return "need n_vars > 0, n_cliques > 0 and n_separators == n_cliques - 1";
    if (!t.var_dims || !t.table_nvars || !t.table_vars || !t.clique_potentials || (ns > 0 && (!t.sep_parent || !t.sep_child)))
        return "null array in fbn_tree_desc";
    if (t.root < 0 || t.root >= nc) return "root out of range";
    if (t.query_var < 0 || t.query_var >= nv) return "query_var out of range";
    if (!level && (team < 1 || team > 32 || (team & (team - 1)))) return "team_warps must be a power of two in [1,32]";
    if (level && (team < 1 || team > 65536)) return "level mode: gang warps must be in [1,65536]";
    P = HostPlan();
    P.level = level;
    {
        int64_t max_table = 0, pos = 0;
        for (int i = 0; i < nc; ++i) {
            int64_t sz = 1;
            for (int k = 0; k < t.table_nvars[i]; ++k) sz = std::min<int64_t>((int64_t)1 << 40, sz * std::max(1, t.var_dims[t.table_vars[pos + k]]));
            pos += t.table_nvars[i];
            max_table = std::max(max_table, sz);
        }
        P.cb = level ? (max_table >= ((int64_t)1 << 20) ? 8 : 4) : (team == 1 ? 1 : 2);
        if (const char *e = getenv("FBN_CB")) {
            int v = atoi(e);
            if (level ? (v == 4 || v == 8) : (v == 1 || v == 2)) P.cb = v;
        }
    }
    const int64_t kColBlock = P.cb;
    P.n_vars = nv; P.n_cliques = nc; P.n_seps = ns; P.query_var = t.query_var; P.team = team;
    P.vars.resize(nv);
    int64_t sd = 0;
    for (int v = 0; v < nv; ++v) {
        if (t.var_dims[v] <= 0) return "variable with non-positive domain size";
        P.vars[v].off = (int32_t)sd;
        P.vars[v].dim = t.var_dims[v];
        sd += t.var_dims[v];
    }
    if (sd > 65535) return "sum of domain sizes exceeds 65535 (lambda index is 16 bit)";
    P.sum_dims = (int32_t)sd;

    // ---- tables ----
    std::vector<Table> clq(nc), sep(ns);
    {
        int64_t pos = 0;
        for (int i = 0; i < nc + ns; ++i) {
            Table &tb = i < nc ? clq[i] : sep[i - nc];
            if (t.table_nvars[i] <= 0) return "empty table";
            tb.vars.assign(t.table_vars + pos, t.table_vars + pos + t.table_nvars[i]);
            pos += t.table_nvars[i];
            if (!finish_table(tb, t.var_dims, nv, err)) return err;
        }
    }
    std::vector<int64_t> pot_off(nc + 1, 0);
    for (int c = 0; c < nc; ++c) pot_off[c + 1] = pot_off[c] + clq[c].size;
    P.clique_entries = pot_off[nc];

    // ---- topology (MarkLevel, src/JunctionTree.cpp:137-181) ----
    std::vector<int> up_sep(nc, -1), depth(nc, -1);
    std::vector<std::vector<int>> kids(nc);
    for (int s = 0; s < ns; ++s) {
        int p = t.sep_parent[s], c = t.sep_child[s];
        if (p < 0 || p >= nc || c < 0 || c >= nc || p == c) return "separator endpoints out of range";
        if (up_sep[c] != -1) return "clique with two upstream separators";
        up_sep[c] = s;
        kids[p].push_back(s);
    }
    if (up_sep[t.root] != -1) return "root has an upstream separator";
    int maxdepth = 0;
    {
        std::vector<int> q{t.root};
        depth[t.root] = 0;
        for (size_t h = 0; h < q.size(); ++h)
            for (int s : kids[q[h]]) {
                int c = t.sep_child[s];
                if (depth[c] != -1) return "cycle in the tree";
                depth[c] = depth[q[h]] + 1;
                maxdepth = std::max(maxdepth, depth[c]);
                q.push_back(c);
            }
        if ((int)q.size() != nc) return "tree is not connected";
    }
    P.n_levels = 2 * maxdepth + 1;
    // separator scope checks + ReorganizeTableStorage invariant (src/JunctionTree.cpp:235-281)
    for (int s = 0; s < ns; ++s) {
        const Table &S = sep[s], &C = clq[t.sep_child[s]], &Pa = clq[t.sep_parent[s]];
        if (S.vars.size() > C.vars.size()) return "separator larger than its child clique";
        size_t base = C.vars.size() - S.vars.size();
        for (size_t i = 0; i < S.vars.size(); ++i) {
            if (C.vars[base + i] != S.vars[i]) return "child clique does not end with its upstream separator's variables";
            if (Pa.find(S.vars[i]) < 0) return "separator variable missing from parent clique";
        }
    }

    // ---- scratch row regions ----
    std::vector<int64_t> sep_off(ns + 1, 0);
    for (int s = 0; s < ns; ++s) sep_off[s + 1] = sep_off[s] + sep[s].size;
    P.sep_entries = sep_off[ns];
    int64_t rows_total = P.clique_entries + 2 * (P.sep_entries + 1) + nc + sd;
    if (rows_total > ((int64_t)1 << 30)) return "tree too large: more than 2^30 scratch rows per tile";
    P.cl_row0 = 0;
    P.mup_row0 = (int32_t)P.clique_entries;
    P.dn_row0 = P.mup_row0 + (int32_t)P.sep_entries + 1;
    P.inv_row0 = P.dn_row0 + (int32_t)P.sep_entries + 1;
    P.mg_row0 = P.inv_row0 + nc;
    P.tile_rows = (int32_t)rows_total;
    auto cl_row = [&](int c) { return (int32_t)pot_off[c]; };
    auto mup_row = [&](int c) { return c == t.root ? P.mup_row0 + (int32_t)P.sep_entries : P.mup_row0 + (int32_t)sep_off[up_sep[c]]; };
    auto fac_row = [&](int c) { return c == t.root ? P.inv_row0 + c : P.dn_row0 + (int32_t)sep_off[up_sep[c]]; };
    auto ncols = [&](int c) { return c == t.root ? (int64_t)1 : sep[up_sep[c]].size; };

    P.pot.assign(t.clique_potentials, t.clique_potentials + pot_off[nc]);

    // ---- collect ops ----
    P.cops.resize(nc);
    for (int c = 0; c < nc; ++c) {
        const Table &C = clq[c];
        CollectOp &op = P.cops[c];
        op = CollectOp();
        op.pot_off = pot_off[c];
        op.ncolv = c == t.root ? 0 : (int32_t)sep[up_sep[c]].vars.size();
        op.nrest = (int32_t)C.vars.size() - op.ncolv;
        op.cols = (int32_t)ncols(c);
        op.rows = (int32_t)(C.size / op.cols);
        op.cl_row = cl_row(c);
        op.mup_row = mup_row(c);
        op.inv_row = P.inv_row0 + c;
        op.nchild = (int32_t)kids[c].size();
        std::vector<int> rpos(op.nrest), cpos(op.ncolv);
        std::iota(rpos.begin(), rpos.end(), 0);
        std::iota(cpos.begin(), cpos.end(), op.nrest);
        // lambda-bit indices (evidence mask): var offset + digit
        op.rowlam_off = (int64_t)P.lam16.size();
        {
            std::vector<int> dg = enumerate_digits(C, rpos);
            for (int64_t r = 0; r < op.rows; ++r)
                for (int i = 0; i < op.nrest; ++i)
                    P.lam16.push_back((uint16_t)(P.vars[C.vars[rpos[i]]].off + dg[(size_t)r * op.nrest + i]));
        }
        op.collam_off = (int64_t)P.lam16.size();
        {
            std::vector<int> dg = enumerate_digits(C, cpos);
            for (int64_t j = 0; j < op.cols; ++j)
                for (int i = 0; i < op.ncolv; ++i)
                    P.lam16.push_back((uint16_t)(P.vars[C.vars[cpos[i]]].off + dg[(size_t)j * op.ncolv + i]));
        }
        // child message index maps, split into row part + column part (TableExtensionMain, PotentialTable.cpp:603-613)
        std::vector<std::vector<int32_t>> rp(op.nchild), cp(op.nchild);
        for (int k = 0; k < op.nchild; ++k) {
            int s = kids[c][k];
            Table V = C;  // same shape, strides of the separator (0 for variables outside it)
            for (size_t i = 0; i < V.vars.size(); ++i) {
                int ps = sep[s].find(V.vars[i]);
                V.stride[i] = ps < 0 ? 0 : sep[s].stride[ps];
            }
            rp[k] = enumerate_offsets(V, rpos);
            cp[k] = enumerate_offsets(V, cpos);
            int32_t base = P.mup_row0 + (int32_t)sep_off[s];
            for (auto &x : rp[k]) x += base;
        }
        op.rowpart_off = (int64_t)P.idx32.size();
        for (int64_t r = 0; r < op.rows; ++r)
            for (int k = 0; k < op.nchild; ++k) P.idx32.push_back(rp[k][(size_t)r]);
        op.colpart_off = (int64_t)P.idx32.size();
        for (int64_t j = 0; j < op.cols; ++j)
            for (int k = 0; k < op.nchild; ++k) P.idx32.push_back(cp[k][(size_t)j]);
        op.invrow_off = (int64_t)P.idx32.size();
        for (int k = 0; k < op.nchild; ++k) P.idx32.push_back(P.inv_row0 + t.sep_child[kids[c][k]]);
        // Visiting order of the column blocks. Child k's message row of entry (r, j) depends only on the digits of the
        // variables in separator k. Consecutive units run on the warps of one CTA at the same time, so the order decides
        // which gathers they share: column variables that belong only to children whose row is the SAME for a whole block
        // (one gather per row) vary fastest, variables of children that need one gather per column vary slower (the warps
        // of a CTA then read the same rows -> L1 hits), variables shared by several children slowest.
        op.perm_off = -1;
        const int64_t nblk = (op.cols + kColBlock - 1) / kColBlock;
        if (op.nchild > 0 && nblk > 1 && C.size >= 32768 && !getenv("FBN_NO_PERM")) {
            std::vector<int> order(op.ncolv);  // column variables, slowest first
            std::iota(order.begin(), order.end(), 0);
            std::vector<int> cnt(op.ncolv, 0), first(op.ncolv, op.nchild), tier(op.ncolv, 3);
            std::vector<char> in_block(op.ncolv, 0), child_uniform(op.nchild, 1);
            {
                int64_t span = 1;  // the fastest column variables, up to a product >= the block width, vary inside a block
                for (int i = op.ncolv; i-- > 0 && span < kColBlock;) { in_block[i] = 1; span *= C.dims[cpos[i]]; }
            }
            for (int i = 0; i < op.ncolv; ++i)
                for (int k = op.nchild; k-- > 0;)
                    if (sep[kids[c][k]].find(C.vars[cpos[i]]) >= 0) {
                        ++cnt[i]; first[i] = k;
                        if (in_block[i]) child_uniform[k] = 0;
                    }
            const bool old_rule = getenv("FBN_PERM_OLD") != nullptr;
            for (int i = 0; i < op.ncolv; ++i)
                tier[i] = cnt[i] >= 2 ? 0 : cnt[i] == 0 ? 3 : (old_rule || !child_uniform[first[i]]) ? 1 : 2;
            std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
                if (tier[a] != tier[b]) return tier[a] < tier[b];
                return first[a] < first[b];
            });
            bool identity = true;
            for (int i = 0; i < op.ncolv; ++i) identity = identity && order[i] == i;
            if (!identity) {
                std::vector<int64_t> w(op.ncolv);  // weight of column variable i in the visiting key
                int64_t acc = 1;
                for (int i = op.ncolv; i-- > 0;) { w[order[i]] = acc; acc *= C.dims[cpos[order[i]]]; }
                std::vector<std::pair<int64_t, int32_t>> key((size_t)nblk);
                for (int64_t b = 0; b < nblk; ++b) {
                    int64_t j = b * kColBlock, kk = 0;
                    for (int i = op.ncolv; i-- > 0;) { kk += (j % C.dims[cpos[i]]) * w[i]; j /= C.dims[cpos[i]]; }
                    key[(size_t)b] = {kk, (int32_t)b};
                }
                std::sort(key.begin(), key.end());
                op.perm_off = (int64_t)P.idx32.size();
                for (auto &kb : key) P.idx32.push_back(kb.second);
            }
        }
    }

    // ---- pull ops: distribute (one per non-root clique) + marginals (one per variable) ----
    auto make_pull = [&](const Table &X, const std::vector<int> &tpos, int64_t xcols, bool col_is_idx) {
        PullOp op = PullOp();
        std::vector<int> rest;
        for (int i = 0; i < (int)X.vars.size(); ++i)
            if (std::find(tpos.begin(), tpos.end(), i) == tpos.end()) rest.push_back(i);
        std::vector<int32_t> tb = enumerate_offsets(X, tpos), tr = enumerate_offsets(X, rest);
        op.nT = (int32_t)tb.size();
        op.nR = (int32_t)tr.size();
        if (P.idx32.size() & 1) P.idx32.push_back(0);  // int2 alignment
        op.tbase_off = (int64_t)P.idx32.size();
        for (int32_t x : tb) { P.idx32.push_back(x); P.idx32.push_back(col_is_idx ? x : (int32_t)(x % xcols)); }
        op.trest_off = (int64_t)P.idx32.size();
        for (int32_t x : tr) { P.idx32.push_back(x); P.idx32.push_back(col_is_idx ? x : (int32_t)(x % xcols)); }
        return op;
    };
    std::vector<int> dist_op(nc, -1);
    for (int c = 0; c < nc; ++c) {
        if (c == t.root) continue;
        int s = up_sep[c], p = t.sep_parent[s];
        std::vector<int> tpos;
        for (int v : sep[s].vars) tpos.push_back(clq[p].find(v));
        PullOp op = make_pull(clq[p], tpos, ncols(p), false);
        op.kind = 1;
        op.src_row = cl_row(p);
        op.fac_row = fac_row(p);
        op.mup_row = mup_row(c);
        op.dst_row = P.dn_row0 + (int32_t)sep_off[s];
        dist_op[c] = (int)P.pops.size();
        P.pops.push_back(op);
        P.parent_entries += clq[p].size;
    }
    // marginal source: the smallest table (separator preferred on ties) containing the variable; any calibrated
    // table gives the same marginal (the reference takes the smallest clique, src/JunctionTree.cpp:1417-1434)
    std::vector<int> marg_op(nv, -1), marg_avail(nv, 0);
    P.marg_source.assign(nv, -1);
    double ref_marg_entries = 0;
    for (int v = 0; v < nv; ++v) {
        int best = -1;
        int64_t best_size = INT64_MAX;
        for (int s = 0; s < ns; ++s)
            if (sep[s].find(v) >= 0 && sep[s].size < best_size) { best = nc + s; best_size = sep[s].size; }
        for (int c = 0; c < nc; ++c)
            if (clq[c].find(v) >= 0 && clq[c].size < best_size) { best = c; best_size = clq[c].size; }
        if (best < 0) return "variable " + std::to_string(v) + " does not appear in any clique";
        {   // reference rule, for the algorithmic-bytes figure only
            size_t mn = SIZE_MAX; int64_t sz = 0;
            for (int c = 0; c < nc; ++c)
                if (clq[c].vars.size() < mn && clq[c].find(v) >= 0) { mn = clq[c].vars.size(); sz = clq[c].size; }
            ref_marg_entries += (double)sz;
        }
        P.marg_source[v] = best;
        PullOp op;
        if (best >= nc) {
            int s = best - nc, c = t.sep_child[s];
            op = make_pull(sep[s], {sep[s].find(v)}, 1, true);
            op.src_row = P.mup_row0 + (int32_t)sep_off[s];
            op.fac_row = P.dn_row0 + (int32_t)sep_off[s];
            marg_avail[v] = depth[c];
        } else {
            int c = best;
            op = make_pull(clq[c], {clq[c].find(v)}, ncols(c), false);
            op.src_row = cl_row(c);
            op.fac_row = fac_row(c);
            marg_avail[v] = depth[c];
        }
        op.kind = 2;
        op.mup_row = 0;
        op.dst_row = P.mg_row0 + P.vars[v].off;
        marg_op[v] = (int)P.pops.size();
        P.pops.push_back(op);
        P.marginal_entries += best_size;
    }

    // ---- phases, warp work units, split-column partial slots ----
    // A block of cb adjacent columns (collect) / target entries (pull) is one unit of warp work unless its row range
    // is long and the phase has too few column blocks to keep `team` warps busy; then the rows are cut into chunks whose
    // partial sums meet in shared memory (fused mode) or scratch rows (level mode). Evidence-mask tables (one 32-bit word per row and per column of every clique of the phase) are built in
    // shared memory at the start of a collect phase when they fit.
    // fused mode keeps partial sums and mask tables in shared memory; level mode keeps partials in scratch rows
    const int64_t kMaxPslots = level ? std::max<int64_t>(192, std::min<int64_t>(1 << 16, P.clique_entries / 8)) : 192;
    const int kMaxMaskWords = (level || getenv("FBN_NO_TABLES")) ? 0 : 6144;
    struct P2Host { int32_t dstB_row, mupB_row; int64_t sizeB; };  // host-only facts of a Pull2Op: separator B's rows
    std::vector<P2Host> p2host;
    auto emit_phase = [&](int kind, const std::vector<int> &ops, const std::vector<int> &p2 = std::vector<int>()) {
        Phase ph = Phase();
        ph.kind = kind;
        ph.op_begin = (int32_t)P.oplist.size();
        ph.fin_begin = (int32_t)P.fins.size();
        int64_t work = 0;
        for (int o : ops) work += kind == 0 ? (int64_t)P.cops[o].rows * P.cops[o].cols : (int64_t)P.pops[o].nT * P.pops[o].nR;
        static const int kSplit = getenv("FBN_SPLIT") ? atoi(getenv("FBN_SPLIT")) : 4;
        int64_t target = std::max<int64_t>(128, work / (kSplit * (int64_t)team));  // table entries per unit
        auto chunks_of = [&](int64_t rows, int64_t cols, int64_t tgt) {
            if (team == 1) return (int64_t)1;
            tgt = std::max<int64_t>(8, tgt / std::min<int64_t>(cols, kColBlock));  // rows per unit
            int64_t ch = std::min<int64_t>(rows, (rows + tgt - 1) / tgt);
            int64_t per = (rows + ch - 1) / ch;
            return (rows + per - 1) / per;
        };
        for (int pass = 0; pass < 12; ++pass) {  // coarsen the split until the partial slots fit
            int64_t pslots = 0;
            for (int o : ops) {
                int64_t rows = kind == 0 ? P.cops[o].rows : P.pops[o].nR, cols = kind == 0 ? P.cops[o].cols : P.pops[o].nT;
                int64_t ch = chunks_of(rows, cols, target);
                if (ch > 1) pslots += ch * cols;
            }
            if (pslots <= kMaxPslots) break;
            target *= 4;
        }
        // fused mode deals the units of a phase round-robin to the warps of the tile's CTA: ops with the most expensive units
        // first, so that the small units at the end even out the load (longest-processing-time order)
        std::vector<int> ordered(ops);
        if (!level && team > 1 && !getenv("FBN_NO_LPT")) {
            auto unit_cost = [&](int o) {
                int64_t rows = kind == 0 ? P.cops[o].rows : P.pops[o].nR, cols = kind == 0 ? P.cops[o].cols : P.pops[o].nT;
                int64_t ch = chunks_of(rows, cols, target);
                int64_t per = (rows + ch - 1) / ch;
                return per * std::min<int64_t>(cols, kColBlock) * (kind == 0 ? 1 + P.cops[o].nchild : 2);
            };
            std::stable_sort(ordered.begin(), ordered.end(), [&](int a, int b) { return unit_cost(a) > unit_cost(b); });
        }
        int pslots = 0, normslots = 0, mask_words = 0;
        int64_t units = 0;
        for (int o : ordered) {
            int64_t rows = kind == 0 ? P.cops[o].rows : P.pops[o].nR, cols = kind == 0 ? P.cops[o].cols : P.pops[o].nT;
            int64_t chunks = chunks_of(rows, cols, target);
            if (chunks > 1 && pslots + chunks * cols > kMaxPslots) chunks = 1;
            int64_t per = (rows + chunks - 1) / chunks;
            int32_t pslot0 = chunks > 1 ? pslots : -1;
            if (chunks > 1) pslots += (int)(chunks * cols);
            if (kind == 0) {
                CollectOp &op = P.cops[o];
                op.normslot = normslots++;
                op.chunks = (int32_t)chunks; op.rows_per_chunk = (int32_t)per; op.pslot0 = pslot0; op.rot = (int32_t)(level ? units : units % team);
                if (mask_words + rows + cols <= kMaxMaskWords) {
                    op.rowmask_off = mask_words; op.colmask_off = mask_words + (int32_t)rows;
                    mask_words += (int32_t)(rows + cols);
                } else {
                    op.rowmask_off = op.colmask_off = -1;
                }
            } else {
                PullOp &op = P.pops[o];
                op.chunks = (int32_t)chunks; op.rows_per_chunk = (int32_t)per; op.pslot0 = pslot0; op.rot = (int32_t)(level ? units : units % team);
            }
            units += chunks * ((cols + kColBlock - 1) / kColBlock);
            P.oplist.push_back(o);
            if (chunks > 1)
                for (int64_t j = 0; j < cols; ++j) {
                    FinItem f = FinItem();
                    if (kind == 0) { f.type = FIN_MUP; f.dst_row = P.cops[o].mup_row + (int32_t)j; }
                    else if (P.pops[o].kind == 1) { f.type = FIN_D; f.dst_row = P.pops[o].dst_row + (int32_t)j; f.aux_row = P.pops[o].mup_row + (int32_t)j; }
                    else { f.type = FIN_MARG; f.dst_row = P.pops[o].dst_row + (int32_t)j; }
                    f.slot0 = pslot0 + (int32_t)(j * chunks);
                    f.nslots = (int32_t)chunks;
                    P.fins.push_back(f);
                }
            if (kind == 0) {
                FinItem f = FinItem();
                f.type = FIN_NORM;
                f.dst_row = P.cops[o].inv_row;
                f.slot0 = P.cops[o].normslot;
                f.nslots = 1;
                P.fins.push_back(f);
            }
        }
        for (int x : p2) {  // fused two-children pulls: their own partial slots (one per entry of separator B and block)
            Pull2Op &op = P.p2ops[x];
            op.pslot0 = pslots;
            op.rot = (int32_t)units;
            units += (int64_t)op.n12 * op.nblk1;
            P.oplist.push_back(-1 - x);
            for (int64_t tB = 0; tB < p2host[x].sizeB; ++tB) {
                FinItem f = FinItem();
                f.type = FIN_D;
                f.dst_row = p2host[x].dstB_row + (int32_t)tB;
                f.aux_row = p2host[x].mupB_row + (int32_t)tB;
                f.slot0 = op.pslot0 + (int32_t)(tB * op.nblk1);
                f.nslots = op.nblk1;
                P.fins.push_back(f);
            }
            pslots += (int)(p2host[x].sizeB * op.nblk1);
        }
        P.n_units += units;
        ph.op_end = (int32_t)P.oplist.size();
        ph.fin_end = (int32_t)P.fins.size();
        ph.n_normslots = normslots;
        ph.mask_words = mask_words;
        ph.n_units = (int32_t)std::min<int64_t>(units, INT32_MAX);
        P.max_normslots = std::max(P.max_normslots, normslots);
        P.max_pslots = std::max(P.max_pslots, pslots);
        P.max_mask_words = std::max(P.max_mask_words, mask_words);
        P.phases.push_back(ph);
    };
    int64_t pull2_saved_entries = 0, pull2_partial_rows = 0;
    for (int d = maxdepth; d >= 0; --d) {  // Collect: deepest cliques first (src/JunctionTree.cpp:1240-1306)
        std::vector<int> ops;
        for (int c = 0; c < nc; ++c)
            if (depth[c] == d) ops.push_back(c);
        emit_phase(0, ops);
    }
    for (int d = 1; d <= maxdepth + 1; ++d) {  // Distribute (src/JunctionTree.cpp:1308-1333) + marginals as soon as ready
        std::vector<int> ops, p2;
        std::vector<char> fused(nc, 0);
        if (level && !getenv("FBN_NO_PULL2")) {
            // two children A, B of one big parent whose separators cover every parent variable: one pass over the parent
            for (int p = 0; p < nc; ++p) {
                if (depth[p] != d - 1 || kids[p].size() < 2 || clq[p].size < 32768) continue;
                const Table &X = clq[p];
                int bestA = -1, bestB = -1;
                int64_t best_slots = INT64_MAX;
                for (int sa : kids[p])
                    for (int sb : kids[p]) {
                        if (sa == sb || fused[t.sep_child[sa]] || fused[t.sep_child[sb]]) continue;
                        bool cover = true;
                        int64_t n1 = 1, n2 = 1;
                        for (size_t i = 0; i < X.vars.size(); ++i) {
                            const bool a = sep[sa].find(X.vars[i]) >= 0, b = sep[sb].find(X.vars[i]) >= 0;
                            if (!a && !b) cover = false;
                            if (a && !b) n1 *= X.dims[i];
                            if (b && !a) n2 *= X.dims[i];
                        }
                        if (!cover || n1 < 2 || n2 < 8) continue;
                        const int64_t slots = sep[sb].size * ((n1 + kColBlock - 1) / kColBlock);
                        if (slots < best_slots) { best_slots = slots; bestA = sa; bestB = sb; }
                    }
                if (bestA < 0 || best_slots > X.size / 3 || best_slots > ((int64_t)1 << 26)) continue;  // partials cost 2 rows each
                std::vector<int> g12, g1, g2;
                for (int i = 0; i < (int)X.vars.size(); ++i) {
                    const bool a = sep[bestA].find(X.vars[i]) >= 0, b = sep[bestB].find(X.vars[i]) >= 0;
                    (a && b ? g12 : a ? g1 : g2).push_back(i);
                }
                Table VA = X, VB = X;  // parent shape with the strides of separator A / B (0 outside)
                for (size_t i = 0; i < X.vars.size(); ++i) {
                    int pa = sep[bestA].find(X.vars[i]), pb = sep[bestB].find(X.vars[i]);
                    VA.stride[i] = pa < 0 ? 0 : sep[bestA].stride[pa];
                    VB.stride[i] = pb < 0 ? 0 : sep[bestB].stride[pb];
                }
                const int64_t xcols = ncols(p);
                auto emit_tab = [&](const std::vector<int> &pos, const Table *T3, const Table *T4) {
                    while (P.idx32.size() & 3) P.idx32.push_back(0);  // int4 alignment
                    const int64_t off = (int64_t)P.idx32.size();
                    std::vector<int32_t> x = enumerate_offsets(X, pos);
                    std::vector<int32_t> e3 = T3 ? enumerate_offsets(*T3, pos) : std::vector<int32_t>(x.size(), 0);
                    std::vector<int32_t> e4 = T4 ? enumerate_offsets(*T4, pos) : std::vector<int32_t>(x.size(), 0);
                    for (size_t i = 0; i < x.size(); ++i) {
                        P.idx32.push_back(x[i]);
                        P.idx32.push_back((int32_t)(x[i] % xcols));
                        P.idx32.push_back(e3[i]);
                        P.idx32.push_back(e4[i]);
                    }
                    return off;
                };
                Pull2Op op = Pull2Op();
                op.a_off = emit_tab(g12, &VA, &VB);
                op.g1_off = emit_tab(g1, &VA, nullptr);
                op.g2_off = emit_tab(g2, &VB, nullptr);
                op.n12 = 1; op.n1 = 1; op.n2 = 1;
                for (int i : g12) op.n12 *= X.dims[i];
                for (int i : g1) op.n1 *= X.dims[i];
                for (int i : g2) op.n2 *= X.dims[i];
                op.nblk1 = (int32_t)((op.n1 + kColBlock - 1) / kColBlock);
                op.src_row = cl_row(p);
                op.fac_row = fac_row(p);
                const int cA = t.sep_child[bestA], cB = t.sep_child[bestB];
                op.mupA_row = mup_row(cA);
                op.dstA_row = P.dn_row0 + (int32_t)sep_off[bestA];
                p2host.push_back({P.dn_row0 + (int32_t)sep_off[bestB], mup_row(cB), sep[bestB].size});
                p2.push_back((int)P.p2ops.size());
                P.p2ops.push_back(op);
                fused[cA] = fused[cB] = 1;
                pull2_saved_entries += X.size;
                pull2_partial_rows += best_slots;
            }
        }
        for (int c = 0; c < nc; ++c)
            if (depth[c] == d && !fused[c]) ops.push_back(dist_op[c]);
        for (int v = 0; v < nv; ++v)
            if (marg_avail[v] == d - 1) ops.push_back(marg_op[v]);
        if (!ops.empty() || !p2.empty()) emit_phase(1, ops, p2);
    }

    if (level) {  // partial-sum rows and per-CTA norm rows are appended to the tile
        P.nb_max = (team + kLevelWarps - 1) / kLevelWarps;
        P.part_row0 = P.tile_rows;
        P.normpart_row0 = P.part_row0 + P.max_pslots;
        int64_t rows = (int64_t)P.normpart_row0 + (int64_t)P.max_normslots * P.nb_max;
        if (rows > ((int64_t)1 << 30)) return "tree too large: more than 2^30 scratch rows per tile";
        P.tile_rows = (int32_t)rows;
    } else if (team > 1) {
        // fused mode, several warps per tile: the per-warp shares of every clique norm of a phase go to scratch rows (written
        // once, read once by the phase's FIN_NORM items, L2-resident) instead of max_normslots x team x 256 B of shared
        // memory, which was what kept big trees (PIGS shape: 15 cliques per phase) at one CTA per SM
        P.nb_max = team;
        P.part_row0 = 0;
        P.normpart_row0 = P.tile_rows;
        int64_t rows = (int64_t)P.normpart_row0 + (int64_t)P.max_normslots * team;
        if (rows > ((int64_t)1 << 30)) return "tree too large: more than 2^30 scratch rows per tile";
        P.tile_rows = (int32_t)rows;
    }

    // ---- flat records (small trees run by one warp per tile, jt_flat_kernel): one record per clique entry / per (target, rest)
    //      pair with ABSOLUTE scratch rows, so the kernel's inner loops are a record load, the data loads and the arithmetic
    for (auto &op : P.cops) op.flat_off = -1;
    for (auto &op : P.pops) op.flat_off = -1;
    if (!level && team == 1 && !getenv("FBN_NO_FLAT")) {
        bool ok = true;
        for (const auto &op : P.cops) ok = ok && op.nchild <= 4 && op.rowmask_off >= 0 && op.chunks == 1;
        for (const auto &op : P.pops) ok = ok && op.chunks == 1;
        if (ok) {
            for (auto &op : P.cops) {
                while (P.flatc.size() & 3) P.flatc.push_back(0);
                op.flat_off = (int64_t)P.flatc.size();
                const int nk = op.nchild;
                for (int64_t j = 0; j < op.cols; ++j)
                    for (int64_t r = 0; r < op.rows; ++r) {
                        int32_t m[4] = {0, 0, 0, 0};
                        for (int k = 0; k < nk; ++k)
                            m[k] = P.idx32[(size_t)(op.rowpart_off + r * nk + k)] + P.idx32[(size_t)(op.colpart_off + j * nk + k)];
                        const double pv = P.pot[(size_t)(op.pot_off + r * op.cols + j)];
                        int32_t bits[2];
                        memcpy(bits, &pv, 8);
                        P.flatc.push_back(bits[0]); P.flatc.push_back(bits[1]); P.flatc.push_back(m[0]); P.flatc.push_back(m[1]);
                        if (nk > 2) { P.flatc.push_back(m[2]); P.flatc.push_back(m[3]); P.flatc.push_back(0); P.flatc.push_back(0); }
                    }
            }
            for (auto &op : P.pops) {
                if (P.flatp.size() & 1) P.flatp.push_back(0);
                op.flat_off = (int64_t)P.flatp.size();
                for (int64_t q = 0; q < op.nT; ++q)
                    for (int64_t r = 0; r < op.nR; ++r) {
                        P.flatp.push_back(op.src_row + P.idx32[(size_t)(op.tbase_off + 2 * q)] + P.idx32[(size_t)(op.trest_off + 2 * r)]);
                        P.flatp.push_back(op.fac_row + P.idx32[(size_t)(op.tbase_off + 2 * q + 1)] + P.idx32[(size_t)(op.trest_off + 2 * r + 1)]);
                    }
            }
            P.flat = true;
        }
    }

    // ---- byte accounting ----
    // B = 8(Tc+Ts) + sum_edges(24|child| + 24|parent| + 32|sep|) + sum_vars 8|marginal clique| + 8 sum_dims
    double child_entries = (double)(P.clique_entries - clq[t.root].size);
    P.alg_bytes_per_case = 8.0 * (double)(P.clique_entries + P.sep_entries) + 24.0 * child_entries +
                           24.0 * (double)P.parent_entries + 32.0 * (double)P.sep_entries + 8.0 * ref_marg_entries + 8.0 * (double)sd;
    // what the fused kernel must move per case: clique write (collect) + parent reads (distribute) + marginal reads
    // + separator up/down rows written once and read once + posterior/label output
    P.kernel_bytes_per_case = 8.0 * (double)(P.clique_entries + P.parent_entries - pull2_saved_entries + 2 * pull2_partial_rows + P.marginal_entries) +
                              32.0 * (double)P.sep_entries + 8.0 * (double)sd + 4.0;
    return "";
}

}  // namespace fbn
