// Host-side planner: flat reference tree (fbn_tree_desc) -> per-tree constant arrays + warp work items.
// Replaces, once per tree, what the reference recomputes for every level of every case in the "pre" phases
// of src/JunctionTree.cpp:586-639 and :834-880 (index maps by GetConfigValueByTableIndex /
// GetTableIndexByConfigValue, src/PotentialTableBase.cpp:547-566).
// Every step of the per-case pipeline becomes a SumOp over a VIEW of a clique (plan.h): the clique tables themselves are
// never written to memory.
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

// For every configuration of `pos` (a list of positions into tab.vars, row-major in the given order) the partial
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

struct GatherSrc {
    const Table *tab;   // source table (its variables are a subset of the viewed table's)
    int32_t base_row;   // scratch row of the source's entry 0
};

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
    P.cb = level ? 4 : (team == 1 ? 1 : 4);  // measured on B200 (fused, HEPAR2- / PIGS-shape): 4 targets per unit +31 % / +26 % over 2
    if (const char *e = getenv("FBN_CB")) {
        int v = atoi(e);
        if (level ? (v == 4 || v == 8) : (team == 1 ? v == 1 : (v == 2 || v == 4))) P.cb = v;
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
    P.clique_off = pot_off;

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

    // ---- scratch row regions (no clique rows: cliques are recomputed, never stored) ----
    std::vector<int64_t> sep_off(ns + 1, 0);
    for (int s = 0; s < ns; ++s) sep_off[s + 1] = sep_off[s] + sep[s].size;
    P.sep_entries = sep_off[ns];
    int64_t rows_total = 2 * (P.sep_entries + 1) + nc + 2 * sd;
    if (rows_total > ((int64_t)1 << 30)) return "tree too large: more than 2^30 scratch rows per tile";
    P.mup_row0 = 0;
    P.dn_row0 = P.mup_row0 + (int32_t)P.sep_entries + 1;
    P.inv_row0 = P.dn_row0 + (int32_t)P.sep_entries + 1;
    P.mg_row0 = P.inv_row0 + nc;
    P.lam_row0 = P.mg_row0 + (int32_t)sd;
    P.tile_rows = (int32_t)rows_total;
    auto up_row = [&](int s) { return P.mup_row0 + (int32_t)sep_off[s]; };
    auto dn_row = [&](int s) { return P.dn_row0 + (int32_t)sep_off[s]; };
    const int32_t dummy_row = P.mup_row0 + (int32_t)P.sep_entries;  // the root's "up message" (one entry, never read)

    P.pot.assign(t.clique_potentials, t.clique_potentials + pot_off[nc]);
    const int64_t one_idx = (int64_t)P.pot.size();  // a 1.0 for the views without potentials (flat records point at it)
    P.pot.push_back(1.0);
    // unary "tables" of the evidence indicators (lambda rows), gathered like messages by views whose pattern table would get too big
    std::vector<Table> lamtab(nv);
    for (int v = 0; v < nv; ++v) {
        lamtab[v].vars = {v};
        if (!finish_table(lamtab[v], t.var_dims, nv, err)) return err;
    }
    std::vector<char> lam_needed(nv, 0);
    const int64_t kMaxFx = getenv("FBN_MAX_FX") ? std::max(1, atoi(getenv("FBN_MAX_FX"))) : 64;  // evidence patterns per entry at most

    // ---- views -------------------------------------------------------------------------------------------------------
    // A view of table X (clique c >= 0, or a separator: c < 0) with the variables at positions `tpos` as targets.
    // `native` = the targets are X's trailing variables in storage order: the original potentials serve as they are.
    auto make_view = [&](const Table &X, int c, const std::vector<int> &tpos, std::vector<GatherSrc> gs,
                         const std::vector<int32_t> &invrows, int kind) {
        SumOp op = SumOp();
        op.kind = kind;
        auto is_target = [&](int i) { return std::find(tpos.begin(), tpos.end(), i) != tpos.end(); };
        // uncovered variables: in none of the gathered tables. Target variables of marginal views need nothing (an observed
        // variable is reported as zeros); those of collect AND distribute views are checked once per finished sum, so that a
        // down ratio is zero wherever the evidence rules its target out: no kernel applies the 0 guard of TableDivision (plan.h).
        std::vector<int> fold_pos, cmask_pos;
        int64_t nfx = 1;
        if (c >= 0) {
            std::vector<int> unc;
            for (int i = 0; i < (int)X.vars.size(); ++i) {
                bool covered = false;
                for (const GatherSrc &g : gs) covered = covered || g.tab->find(X.vars[i]) >= 0;
                if (covered) continue;
                if (!is_target(i)) unc.push_back(i);
                else if (kind != SUM_MARGINAL) cmask_pos.push_back(i);  // checked once per finished sum, not per entry
            }
            std::stable_sort(unc.begin(), unc.end(), [&](int a, int b) { return X.dims[a] < X.dims[b]; });
            for (int i : unc) {
                if (nfx * (X.dims[i] + 2) <= kMaxFx) {
                    fold_pos.push_back(i);
                    nfx *= X.dims[i] + 2;
                    ++P.folded_vars;
                } else {
                    gs.push_back({&lamtab[X.vars[i]], P.lam_row0 + P.vars[X.vars[i]].off});
                    lam_needed[X.vars[i]] = 1;
                    ++P.lambda_gathers;
                }
            }
        }
        std::vector<int> rpos, frpos;  // row variables of the view; folded row variables (summed out)
        for (int i = 0; i < (int)X.vars.size(); ++i) {
            if (is_target(i)) continue;
            if (std::find(fold_pos.begin(), fold_pos.end(), i) != fold_pos.end()) frpos.push_back(i);
            else rpos.push_back(i);
        }
        const std::vector<int32_t> roff = enumerate_offsets(X, rpos), coff = enumerate_offsets(X, tpos);
        op.rows = (int32_t)roff.size();
        op.cols = (int32_t)coff.size();
        op.nfold = (int32_t)fold_pos.size();
        op.nfx = (int32_t)nfx;
        op.fold_off = (int64_t)P.idx32.size();
        {
            int64_t st = 1;
            for (int i : fold_pos) { P.idx32.push_back(X.vars[i]); P.idx32.push_back((int32_t)st); st *= X.dims[i] + 2; }
        }
        op.cmask_off = (int64_t)P.idx32.size();
        op.ncmask = (int32_t)cmask_pos.size();
        if (!cmask_pos.empty()) {  // the variables, then for every target its pattern byte "1 + digit" of each of them
            for (int i : cmask_pos) P.idx32.push_back(X.vars[i]);
            std::vector<int64_t> cst;  // stride of the variable's digit in the target index (targets are enumerated row-major over tpos)
            for (int i : cmask_pos) {
                int64_t s = 1;
                for (size_t k = tpos.size(); k-- > 0 && tpos[k] != i;) s *= X.dims[tpos[k]];
                cst.push_back(s);
            }
            for (int64_t j = 0; j < op.cols; ++j)
                for (size_t k = 0; k < cmask_pos.size(); ++k) P.idx32.push_back(1 + (int32_t)((j / cst[k]) % X.dims[cmask_pos[k]]));
        }
        if (c < 0) {
            op.pot_off = -1;  // separators carry no potentials (all ones in the reference)
        } else {
            // potX in UNIT-MAJOR order: block b of cb adjacent targets, then row, then target inside the block, then pattern:
            //   potX[(((j / cb) * rows + r) * cb + j % cb) * nfx + fx]
            // so that the potentials of one unit of warp work (target block x row range) are ONE contiguous stream (fetched
            // by cp.async.bulk into shared memory, ops.cuh) and a flat plan (cb = 1) reads them in record order. The last
            // block is padded with zeros. Entry (r, j, fx): the folded row variables summed out under evidence pattern fx.
            while (P.pot.size() & 1) P.pot.push_back(0.0);  // 16-byte alignment of every view (bulk copies)
            op.pot_off = (int64_t)P.pot.size();
            const int64_t nblk = (op.cols + kColBlock - 1) / kColBlock;
            P.pot.resize(P.pot.size() + (size_t)(nblk * op.rows * kColBlock * nfx), 0.0);
            double *dst = P.pot.data() + op.pot_off;
            const double *src = t.clique_potentials + pot_off[c];
            const std::vector<int32_t> froff = enumerate_offsets(X, frpos);
            const std::vector<int> frdig = enumerate_digits(X, frpos);
            const int nfr = (int)frpos.size(), nfold = (int)fold_pos.size();
            std::vector<int> fr_slot(nfr);  // index in fold_pos of every folded row variable
            for (int a = 0; a < nfr; ++a) fr_slot[a] = (int)(std::find(fold_pos.begin(), fold_pos.end(), frpos[a]) - fold_pos.begin());
            std::vector<int> pat(nfold);
            std::vector<int32_t> combo;  // offsets of the folded-variable configurations pattern fx allows
            for (int64_t fx = 0; fx < nfx; ++fx) {
                int64_t x = fx;
                bool dead = false;  // a variable observed in no valid state: everything is zero
                for (int k = 0; k < nfold; ++k) {
                    const int d = X.dims[fold_pos[k]];
                    pat[k] = (int)(x % (d + 2));
                    x /= d + 2;
                    dead = dead || pat[k] == d + 1;
                }
                combo.clear();
                for (size_t q = 0; q < froff.size() && !dead; ++q) {
                    bool ok = true;
                    for (int a = 0; a < nfr && ok; ++a) {
                        const int pk = pat[fr_slot[a]];
                        ok = pk == 0 || pk - 1 == frdig[q * nfr + a];
                    }
                    if (ok) combo.push_back(froff[q]);
                }
                for (int64_t j = 0; j < op.cols; ++j) {
                    double *d2 = dst + ((j / kColBlock) * op.rows * kColBlock + j % kColBlock) * nfx + fx;
                    for (int64_t r = 0; r < op.rows; ++r) {
                        const int64_t e0 = (int64_t)roff[(size_t)r] + coff[(size_t)j];
                        double sum = 0.0;
                        for (int32_t o : combo) sum += src[e0 + o];
                        d2[r * kColBlock * nfx] = sum;
                    }
                }
            }
        }
        // index maps of the gathered tables, split into row part + column part (TableExtensionMain, PotentialTable.cpp:603-613)
        op.ng = (int32_t)gs.size();
        std::vector<std::vector<int32_t>> rp(op.ng), cp(op.ng);
        for (int g = 0; g < op.ng; ++g) {
            Table V = X;  // same shape, strides of the source (0 for variables outside it)
            for (size_t i = 0; i < V.vars.size(); ++i) {
                int ps = gs[g].tab->find(V.vars[i]);
                V.stride[i] = ps < 0 ? 0 : gs[g].tab->stride[ps];
            }
            rp[g] = enumerate_offsets(V, rpos);
            cp[g] = enumerate_offsets(V, tpos);
            for (auto &x : rp[g]) x += gs[g].base_row;
        }
        op.rowpart_off = (int64_t)P.idx32.size();
        for (int64_t r = 0; r < op.rows; ++r)
            for (int g = 0; g < op.ng; ++g) P.idx32.push_back(rp[g][(size_t)r]);
        op.colpart_off = (int64_t)P.idx32.size();
        for (int64_t j = 0; j < op.cols; ++j)
            for (int g = 0; g < op.ng; ++g) P.idx32.push_back(cp[g][(size_t)j]);
        op.invrow_off = (int64_t)P.idx32.size();
        op.ninv = (int32_t)invrows.size();
        for (int32_t r : invrows) P.idx32.push_back(r);
        op.flat_off = -1;
        // Visiting order of the column blocks. Gather g's row of entry (r, j) depends only on the digits of the variables
        // of its source table. Consecutive units run on the warps of one CTA at the same time, so the order decides which
        // gathers they share: target variables that belong only to sources whose row is the SAME for a whole block (one
        // gather per row) vary fastest, variables of sources that need one gather per column vary slower (the warps of a
        // CTA then read the same rows -> L1 hits), variables shared by several sources slowest.
        op.perm_off = -1;
        const int64_t nblk = (op.cols + kColBlock - 1) / kColBlock;
        if (op.ng > 0 && nblk > 1 && (int64_t)op.rows * op.cols >= 32768 && !getenv("FBN_NO_PERM")) {
            const int ncv = (int)tpos.size();
            std::vector<int> order(ncv);  // target variables, slowest first
            std::iota(order.begin(), order.end(), 0);
            std::vector<int> cnt(ncv, 0), first(ncv, op.ng), tier(ncv, 3);
            std::vector<char> in_block(ncv, 0), src_uniform(op.ng, 1);
            {
                int64_t span = 1;  // the fastest target variables, up to a product >= the block width, vary inside a block
                for (int i = ncv; i-- > 0 && span < kColBlock;) { in_block[i] = 1; span *= X.dims[tpos[i]]; }
            }
            for (int i = 0; i < ncv; ++i)
                for (int g = op.ng; g-- > 0;)
                    if (gs[g].tab->find(X.vars[tpos[i]]) >= 0) {
                        ++cnt[i]; first[i] = g;
                        if (in_block[i]) src_uniform[g] = 0;
                    }
            for (int i = 0; i < ncv; ++i) tier[i] = cnt[i] >= 2 ? 0 : cnt[i] == 0 ? 3 : !src_uniform[first[i]] ? 1 : 2;
            std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
                if (tier[a] != tier[b]) return tier[a] < tier[b];
                return first[a] < first[b];
            });
            bool identity = true;
            for (int i = 0; i < ncv; ++i) identity = identity && order[i] == i;
            if (!identity) {
                std::vector<int64_t> w(ncv);  // weight of target variable i in the visiting key
                int64_t acc = 1;
                for (int i = ncv; i-- > 0;) { w[order[i]] = acc; acc *= X.dims[tpos[order[i]]]; }
                std::vector<std::pair<int64_t, int32_t>> key((size_t)nblk);
                for (int64_t b = 0; b < nblk; ++b) {
                    int64_t j = b * kColBlock, kk = 0;
                    for (int i = ncv; i-- > 0;) { kk += (j % X.dims[tpos[i]]) * w[i]; j /= X.dims[tpos[i]]; }
                    key[(size_t)b] = {kk, (int32_t)b};
                }
                std::sort(key.begin(), key.end());
                op.perm_off = (int64_t)P.idx32.size();
                for (auto &kb : key) P.idx32.push_back(kb.second);
            }
        }
        P.view_entries += (int64_t)op.rows * op.cols;
        P.gather_rows += (int64_t)op.rows * op.cols * op.ng;
        return op;
    };
    auto inv_rows_of = [&](int c) {  // 1/norm rows that scale every entry of clique c: its children's, and its own at the root
        std::vector<int32_t> r;
        for (int s : kids[c]) r.push_back(P.inv_row0 + t.sep_child[s]);
        if (c == t.root) r.push_back(P.inv_row0 + c);
        return r;
    };
    auto trailing = [&](const Table &X, size_t n) {
        std::vector<int> p(n);
        std::iota(p.begin(), p.end(), (int)(X.vars.size() - n));
        return p;
    };

    // ---- collect ops (one per clique) ----
    std::vector<int> collect_op(nc, -1);
    for (int c = 0; c < nc; ++c) {
        std::vector<GatherSrc> gs;
        for (int s : kids[c]) gs.push_back({&sep[s], up_row(s)});
        std::vector<int32_t> inv;
        for (int s : kids[c]) inv.push_back(P.inv_row0 + t.sep_child[s]);
        SumOp op = make_view(clq[c], c, c == t.root ? std::vector<int>() : trailing(clq[c], sep[up_sep[c]].vars.size()), gs, inv, SUM_COLLECT);
        op.dst_row = c == t.root ? dummy_row : up_row(up_sep[c]);
        op.inv_row = P.inv_row0 + c;
        collect_op[c] = (int)P.ops.size();
        P.ops.push_back(op);
    }
    // ---- distribute ops (one per non-root clique): message of parent p into separator s = up_sep[c] ----
    std::vector<int> dist_op(nc, -1);
    for (int c = 0; c < nc; ++c) {
        if (c == t.root) continue;
        const int s = up_sep[c], p = t.sep_parent[s];
        std::vector<int> tpos;
        for (int v : sep[s].vars) tpos.push_back(clq[p].find(v));
        std::vector<GatherSrc> gs;
        for (int s2 : kids[p])
            if (s2 != s) gs.push_back({&sep[s2], up_row(s2)});
        if (p != t.root) gs.push_back({&sep[up_sep[p]], dn_row(up_sep[p])});
        SumOp op = make_view(clq[p], p, tpos, gs, inv_rows_of(p), SUM_DISTRIBUTE);
        op.dst_row = dn_row(s);
        op.aux_row = up_row(s);   // informational: no kernel applies the 0 guard (plan.h)
        dist_op[c] = (int)P.ops.size();
        P.ops.push_back(op);
        P.parent_entries += clq[p].size;
    }
    // ---- marginal ops (one per variable). Source: the smallest table (separator preferred on ties) containing the
    //      variable; any calibrated table gives the same marginal (the reference takes the smallest clique,
    //      src/JunctionTree.cpp:1417-1434)
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
        SumOp op;
        if (best >= nc) {
            const int s = best - nc;
            op = make_view(sep[s], -1, {sep[s].find(v)}, {{&sep[s], up_row(s)}, {&sep[s], dn_row(s)}}, {}, SUM_MARGINAL);
            marg_avail[v] = depth[t.sep_child[s]];
        } else {
            const int c = best;
            std::vector<GatherSrc> gs;
            for (int s : kids[c]) gs.push_back({&sep[s], up_row(s)});
            if (c != t.root) gs.push_back({&sep[up_sep[c]], dn_row(up_sep[c])});
            op = make_view(clq[c], c, {clq[c].find(v)}, gs, inv_rows_of(c), SUM_MARGINAL);
            marg_avail[v] = depth[c];
        }
        op.dst_row = P.mg_row0 + P.vars[v].off;
        marg_op[v] = (int)P.ops.size();
        P.ops.push_back(op);
        P.marginal_entries += best_size;
    }

    // ---- phases, warp work units, split-column partial slots ----
    // A block of cb adjacent targets is one unit of warp work unless its row range is long and the phase has too few column
    // blocks to keep `team` warps busy; then the rows are cut into chunks whose partial sums meet in scratch rows.
    const int64_t kMaxPslots = level ? std::max<int64_t>(192, std::min<int64_t>(1 << 16, P.clique_entries / 8)) : (team == 1 ? 0 : 512);
    auto emit_phase = [&](int kind, const std::vector<int> &ops) {
        Phase ph = Phase();
        ph.kind = kind;
        ph.op_begin = (int32_t)P.oplist.size();
        ph.fin_begin = (int32_t)P.fins.size();
        int64_t work = 0;
        for (int o : ops) work += (int64_t)P.ops[o].rows * P.ops[o].cols;
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
                int64_t ch = chunks_of(P.ops[o].rows, P.ops[o].cols, target);
                if (ch > 1) pslots += ch * P.ops[o].cols;
            }
            if (pslots <= kMaxPslots) break;
            target *= 4;
        }
        // fused mode deals the units of a phase round-robin to the warps of the tile's CTA: ops with the most expensive units
        // first, so that the small units at the end even out the load (longest-processing-time order)
        std::vector<int> ordered(ops);
        if (!level && team > 1 && !getenv("FBN_NO_LPT")) {
            auto unit_cost = [&](int o) {
                int64_t rows = P.ops[o].rows, cols = P.ops[o].cols;
                int64_t ch = chunks_of(rows, cols, target);
                int64_t per = (rows + ch - 1) / ch;
                return per * std::min<int64_t>(cols, kColBlock) * (1 + P.ops[o].ng);
            };
            std::stable_sort(ordered.begin(), ordered.end(), [&](int a, int b) { return unit_cost(a) > unit_cost(b); });
        }
        int pslots = 0, normslots = 0;
        int64_t units = 0;
        for (int o : ordered) {
            SumOp &op = P.ops[o];
            const int64_t rows = op.rows, cols = op.cols;
            int64_t chunks = chunks_of(rows, cols, target);
            if (chunks > 1 && pslots + chunks * cols > kMaxPslots) chunks = 1;
            const int64_t per = (rows + chunks - 1) / chunks;
            const int32_t pslot0 = chunks > 1 ? pslots : -1;
            if (chunks > 1) pslots += (int)(chunks * cols);
            op.normslot = op.kind == SUM_COLLECT ? normslots++ : -1;
            op.chunks = (int32_t)chunks; op.rows_per_chunk = (int32_t)per; op.pslot0 = pslot0; op.rot = (int32_t)(level ? units : units % team);
            units += chunks * ((cols + kColBlock - 1) / kColBlock);
            P.oplist.push_back(o);
            if (chunks > 1)
                for (int64_t j = 0; j < cols; ++j) {
                    FinItem f = FinItem();
                    f.type = FIN_SUM;
                    f.dst_row = op.dst_row + (int32_t)j;
                    f.aux_row = 0;
                    f.slot0 = pslot0 + (int32_t)(j * chunks);
                    f.nslots = (int32_t)chunks;
                    P.fins.push_back(f);
                }
            if (op.kind == SUM_COLLECT) {
                FinItem f = FinItem();
                f.type = FIN_NORM;
                f.dst_row = op.inv_row;
                f.slot0 = op.normslot;
                f.nslots = 1;
                P.fins.push_back(f);
            }
        }
        P.n_units += units;
        ph.op_end = (int32_t)P.oplist.size();
        ph.fin_end = (int32_t)P.fins.size();
        ph.n_normslots = normslots;
        ph.n_units = (int32_t)std::min<int64_t>(units, INT32_MAX);
        P.max_normslots = std::max(P.max_normslots, normslots);
        P.max_pslots = std::max(P.max_pslots, pslots);
        P.phases.push_back(ph);
    };
    for (int d = maxdepth; d >= 0; --d) {  // Collect: deepest cliques first (src/JunctionTree.cpp:1240-1306)
        std::vector<int> ops;
        for (int c = 0; c < nc; ++c)
            if (depth[c] == d) ops.push_back(collect_op[c]);
        emit_phase(0, ops);
    }
    for (int d = 1; d <= maxdepth + 1; ++d) {  // Distribute (src/JunctionTree.cpp:1308-1333) + marginals as soon as ready
        std::vector<int> ops;
        for (int c = 0; c < nc; ++c)
            if (depth[c] == d) ops.push_back(dist_op[c]);
        for (int v = 0; v < nv; ++v)
            if (marg_avail[v] == d - 1) ops.push_back(marg_op[v]);
        if (!ops.empty()) emit_phase(1, ops);
    }

    // partial-sum rows and norm-share rows are appended to the tile (several warps / CTAs per tile)
    if (level) P.nb_max = (team + kLevelWarps - 1) / kLevelWarps;
    else P.nb_max = team;
    {
        P.part_row0 = P.tile_rows;
        P.normpart_row0 = P.part_row0 + P.max_pslots;
        int64_t rows = (int64_t)P.normpart_row0 + (int64_t)P.max_normslots * P.nb_max;
        if (rows > ((int64_t)1 << 30)) return "tree too large: more than 2^30 scratch rows per tile";
        P.tile_rows = (int32_t)rows;
    }

    // ---- flat records (small trees run by one warp per tile, jt_flat_kernel): one record per view entry with ABSOLUTE
    //      scratch rows, so the kernel's inner loop is a record load, the data loads and the arithmetic
    if (!level && team == 1 && !getenv("FBN_NO_FLAT")) {
        bool ok = P.pot.size() < ((size_t)1 << 31);
        for (const auto &op : P.ops) ok = ok && op.ng <= 7 && op.chunks == 1;
        if (ok) {
            for (auto &op : P.ops) {
                while (P.flatc.size() & 3) P.flatc.push_back(0);
                op.flat_off = (int64_t)P.flatc.size();
                const int ng = op.ng;
                for (int64_t j = 0; j < op.cols; ++j)
                    for (int64_t r = 0; r < op.rows; ++r) {
                        int32_t m[7] = {0, 0, 0, 0, 0, 0, 0};
                        for (int g = 0; g < ng; ++g)
                            m[g] = P.idx32[(size_t)(op.rowpart_off + r * ng + g)] + P.idx32[(size_t)(op.colpart_off + j * ng + g)];
                        const int64_t pi = op.pot_off < 0 ? one_idx : op.pot_off + (j * op.rows + r) * op.nfx;  // cb = 1: target-major
                        P.flatc.push_back((int32_t)pi); P.flatc.push_back(m[0]); P.flatc.push_back(m[1]); P.flatc.push_back(m[2]);
                        if (ng > 3) { P.flatc.push_back(m[3]); P.flatc.push_back(m[4]); P.flatc.push_back(m[5]); P.flatc.push_back(m[6]); }
                    }
            }
            P.flat = true;
        }
    }

    for (int v = 0; v < nv; ++v)
        if (lam_needed[v]) P.lamvars.push_back(v);

    // ---- byte accounting ----
    // B = 8(Tc+Ts) + sum_edges(24|child| + 24|parent| + 32|sep|) + sum_vars 8|marginal clique| + 8 sum_dims
    double child_entries = (double)(P.clique_entries - clq[t.root].size);
    P.alg_bytes_per_case = 8.0 * (double)(P.clique_entries + P.sep_entries) + 24.0 * child_entries +
                           24.0 * (double)P.parent_entries + 32.0 * (double)P.sep_entries + 8.0 * ref_marg_entries + 8.0 * (double)sd;
    // what the kernels must move per case when every scratch row misses the caches once: each message / ratio / norm /
    // marginal row written once and read once + posterior/label output (gathers beyond the first hit L1/L2)
    P.kernel_bytes_per_case = 16.0 * (double)(2 * P.sep_entries + nc + sd) + 8.0 * (double)sd + 4.0;
    (void)one_idx;
    return "";
}

}  // namespace fbn
