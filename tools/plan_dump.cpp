// Developer tool: prints the structure of a plan (row regions, constants, ops per phase, entries, live message rows).
// Build: g++ -O2 -std=c++17 -I include -I fastbn_b200/csrc tools/plan_dump.cpp fastbn_b200/csrc/plan.cpp -shared -fPIC -o build/tools/libplan_dump.so
// Use:   python tools/plan_dump.py <workload> [team] [level]
#include <algorithm>
#include <cstdio>
#include <vector>

#include "fastbn_b200.h"
#include "plan.h"

extern "C" int plan_dump(const fbn_tree_desc *tree, int team, int level, int verbose) {
    fbn::HostPlan h;
    std::string err = fbn::build_plan(*tree, team, h, level != 0);
    if (!err.empty()) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
    printf("team %d cb %d flat %d phases %zu ops %zu tile_rows %d (mup %d dn %d inv %d mg %d lam %d part %d normpart %d)\n", h.team, h.cb, (int)h.flat,
           h.phases.size(), h.ops.size(), h.tile_rows, h.mup_row0, h.dn_row0, h.inv_row0, h.mg_row0, h.lam_row0, h.part_row0, h.normpart_row0);
    printf("cliques %d seps %d vars %d sum_dims %d clique_entries %lld sep_entries %lld view_entries %lld gather_rows %lld\n", h.n_cliques, h.n_seps,
           h.n_vars, h.sum_dims, (long long)h.clique_entries, (long long)h.sep_entries, (long long)h.view_entries, (long long)h.gather_rows);
    printf("constants: pot %zu doubles (%.1f KB) idx32 %zu (%.1f KB) flatc %zu (%.1f KB) lamvars %zu\n", h.pot.size(), h.pot.size() * 8 / 1024.0,
           h.idx32.size(), h.idx32.size() * 4 / 1024.0, h.flatc.size(), h.flatc.size() * 4 / 1024.0, h.lamvars.size());
    for (size_t pi = 0; pi < h.phases.size(); ++pi) {
        const fbn::Phase &p = h.phases[pi];
        long long entries = 0, maxop = 0, targets = 0;
        int nk[3] = {0, 0, 0};
        for (int oi = p.op_begin; oi < p.op_end; ++oi) {
            const fbn::SumOp &op = h.ops[h.oplist[oi]];
            entries += (long long)op.rows * op.cols;
            maxop = std::max(maxop, (long long)op.rows * op.cols);
            targets += op.cols;
            nk[op.kind]++;
            if (verbose > 1)
                printf("    op %3d kind %d rows %5d cols %4d ng %d nfx %3d nfold %d ncmask %d ninv %d chunks %d dst %d\n", h.oplist[oi], op.kind, op.rows, op.cols,
                       op.ng, op.nfx, op.nfold, op.ncmask, op.ninv, op.chunks, op.dst_row);
        }
        if (verbose)
            printf("ph %3zu kind %d ops %3d (collect %d distribute %d marginal %d) targets %4lld entries %7lld maxop %6lld units %d fins %d\n", pi, p.kind,
                   p.op_end - p.op_begin, nk[0], nk[1], nk[2], targets, entries, maxop, p.n_units, p.fin_end - p.fin_begin);
    }
    return 0;
}

#include "resident.h"
// resident-tile schedule of the tree: per block the work of every warp in model instructions
extern "C" int resident_dump(const fbn_tree_desc *tree, int warps, int verbose) {
    fbn::HostPlan h;
    std::string err = fbn::build_plan(*tree, 1, h, false);
    if (!err.empty()) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
    if (verbose > 1)
        for (size_t o = 0; o < h.ops.size(); ++o) {
            const fbn::SumOp &op = h.ops[o];
            printf("    op %3zu kind %d rows %5d cols %4d ng %d nfx %3d nfold %d ncmask %d\n", o, op.kind, op.rows, op.cols, op.ng, op.nfx, op.nfold, op.ncmask);
        }
    fbn::ResidentPlan R;
    fbn::build_resident(h, warps, 232448, R);
    printf("resident: ok %d (%s) warps %d blocks %d state_rows %d (%.1f KB) max_block %d B smem %lld B items %d balance %.3f constants %.1f KB pot %.1f KB\n", (int)R.ok,
           R.why.c_str(), R.warps, R.n_blocks, R.state_rows, R.state_rows * 0.5, R.max_block_bytes, (long long)R.smem_bytes, R.n_items, R.balance,
           R.blob.size() * 4 / 1024.0, h.pot.size() * 8 / 1024.0);
    if (!R.ok) return 2;
    for (int b = 0; b < R.n_blocks && verbose; ++b) {
        const uint32_t *blk = R.blob.data() + R.block_off[b] / 4;
        const fbn::ResItem *items = reinterpret_cast<const fbn::ResItem *>(reinterpret_cast<const unsigned char *>(blk) + blk[2]);
        printf("block %2d bytes %6u items %3u | entries per warp:", b, blk[3], blk[0]);
        for (int w = 0; w < warps; ++w) {
            unsigned e = 0, n = 0;
            for (uint32_t i = blk[4 + w]; i < blk[4 + w + 1]; ++i) {
                if ((items[i].kind_ng & 255) == fbn::RES_SUM || (items[i].kind_ng & 255) == fbn::RES_SUMC) e += items[i].n_targets * items[i].rows;
                else if ((items[i].kind_ng & 255) == fbn::RES_OUT) e += 1000;
                else e += 50000;
                ++n;
            }
            printf(" %u/%u", e, n);
        }
        printf("\n");
        if (verbose > 2) {   // the schedule model's cost of every warp (resident.cpp: sum_piece_cost / out_cost / kEvidCost)
            printf("         model cost per warp:");
            for (int w = 0; w < warps; ++w) {
                double c = 0;
                for (uint32_t i = blk[4 + w]; i < blk[4 + w + 1]; ++i) {
                    const uint32_t kind = items[i].kind_ng & 255, ng = (items[i].kind_ng >> 8) & 255;
                    if (kind == fbn::RES_SUM || kind == fbn::RES_SUMC) {
                        const uint32_t *d = blk + items[i].desc_off / 4;
                        c += 100.0 + 12.0 * d[0] + 6.0 * d[1] + items[i].n_targets * (8.0 + 10.0 * d[1] + (10.0 + 5.0 * ng + (ng > 3 ? 1.0 : 0.0)) * items[i].rows);
                    } else if (kind == fbn::RES_OUT) c += 180.0 + 20.0 * items[i].n_targets;
                    else c += 1200.0;
                }
                printf(" %5.0f", c);
            }
            printf("\n");
        }
    }
    return 0;
}

// how the gathers of a flat plan depend on rows / targets: entries x gathers by category, and what hoisting would leave
extern "C" int gather_stats(const fbn_tree_desc *tree) {
    fbn::HostPlan h;
    std::string err = fbn::build_plan(*tree, 1, h, false);
    if (!err.empty()) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
    long long both = 0, rowonly = 0, colonly = 0, none = 0, entries = 0, hoisted = 0, potless = 0, nofold = 0;
    for (const fbn::SumOp &op : h.ops) {
        entries += (long long)op.rows * op.cols;
        if (op.pot_off < 0) potless += (long long)op.rows * op.cols;
        else if (op.nfold == 0) nofold += (long long)op.rows * op.cols;
        for (int g = 0; g < op.ng; ++g) {
            bool rv = false, cv = false;
            for (int r = 1; r < op.rows; ++r) rv = rv || h.idx32[op.rowpart_off + (long long)r * op.ng + g] != h.idx32[op.rowpart_off + g];
            for (int j = 1; j < op.cols; ++j) cv = cv || h.idx32[op.colpart_off + (long long)j * op.ng + g] != h.idx32[op.colpart_off + g];
            const long long e = (long long)op.rows * op.cols;
            if (rv && cv) { both += e; hoisted += e; }
            else if (rv) { rowonly += e; hoisted += e; }   // target-independent: reloaded per target unless targets are blocked
            else if (cv) { colonly += e; hoisted += op.cols; }   // row-independent: once per target
            else { none += e; hoisted += 1; }
        }
    }
    printf("entries %lld (no potential %lld, unfolded potential %lld)  gathers: row+target %lld, row only %lld, target only %lld, constant %lld -> %lld loads after hoisting the row-independent ones out of the row loop\n",
           entries, potless, nofold, both, rowonly, colonly, none, hoisted);
    return 0;
}

// view entries of a fused plan by gathers per entry
extern "C" int ng_hist(const fbn_tree_desc *tree, int team) {
    fbn::HostPlan h;
    std::string err = fbn::build_plan(*tree, team, h, false);
    if (!err.empty()) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
    long long e[10] = {0}, tot = 0;
    for (const fbn::SumOp &op : h.ops) { e[std::min(op.ng, 9)] += (long long)op.rows * op.cols; tot += (long long)op.rows * op.cols; }
    printf("view entries by gathers per entry:");
    for (int g = 0; g < 10; ++g) if (e[g]) printf("  ng %d: %.1f%%", g, 100.0 * e[g] / tot);
    printf("\n");
    return 0;
}
