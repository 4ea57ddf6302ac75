// Developer tool: prints the phase structure of a fused / level plan (ops, units, entries per phase, per-warp balance).
// Build: g++ -O2 -std=c++17 -I include -I fastbn_b200/csrc tools/plan_dump.cpp fastbn_b200/csrc/plan.cpp -shared -fPIC -o build/tools/libplan_dump.so
#include <algorithm>
#include <cstdio>
#include <vector>

#include "fastbn_b200.h"
#include "plan.h"

extern "C" int plan_dump(const fbn_tree_desc *tree, int team, int level, int verbose) {
    fbn::HostPlan h;
    std::string err = fbn::build_plan(*tree, team, h, level != 0);
    if (!err.empty()) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
    printf("team %d cb %d phases %zu tile_rows %d clique_entries %lld smem: mask_words %d normslots %d pslots %d\n", h.team, h.cb,
           h.phases.size(), h.tile_rows, (long long)h.clique_entries, h.max_mask_words, h.max_normslots, h.max_pslots);
    double tot = 0, totmax = 0;
    for (size_t pi = 0; pi < h.phases.size(); ++pi) {
        const fbn::Phase &p = h.phases[pi];
        std::vector<double> w(h.team, 0.0);
        double entries = 0;
        int units = 0, maxop = 0;
        for (int oi = p.op_begin; oi < p.op_end; ++oi) {
            int id = h.oplist[oi];
            if (p.kind == 0) {
                const fbn::CollectOp &op = h.cops[id];
                int nu = (op.cols + h.cb - 1) / h.cb * op.chunks;
                for (int u = 0; u < nu; ++u) {
                    int ck = u % op.chunks;
                    int r0 = ck * op.rows_per_chunk, r1 = std::min(op.rows, r0 + op.rows_per_chunk);
                    double e = (double)(r1 - r0) * std::min(h.cb, op.cols) * (1 + op.nchild);
                    w[(u + op.rot) % h.team] += e;
                    entries += e;
                }
                units += nu;
                maxop = std::max(maxop, op.rows * op.cols);
            } else if (id >= 0) {
                const fbn::PullOp &op = h.pops[id];
                int nu = (op.nT + h.cb - 1) / h.cb * op.chunks;
                for (int u = 0; u < nu; ++u) {
                    int ck = u % op.chunks;
                    int r0 = ck * op.rows_per_chunk, r1 = std::min(op.nR, r0 + op.rows_per_chunk);
                    double e = (double)(r1 - r0) * std::min(h.cb, op.nT) * 2;
                    w[(u + op.rot) % h.team] += e;
                    entries += e;
                }
                units += nu;
                maxop = std::max(maxop, op.nT * op.nR);
            }
        }
        double mx = *std::max_element(w.begin(), w.end());
        tot += entries; totmax += mx * h.team;
        if (verbose) printf("ph %3zu kind %d ops %3d units %5d fins %3d loads %9.0f max/warp %8.0f eff %.2f maxop %d\n", pi, p.kind, p.op_end - p.op_begin, units,
               p.fin_end - p.fin_begin, entries, mx, entries / (mx * h.team + 1e-9), maxop);
    }
    printf("total loads %.0f  sum(max*team) %.0f  balance efficiency %.3f\n", tot, totmax, tot / totmax);
    return 0;
}
