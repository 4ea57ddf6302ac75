// Flattened junction-tree plan shared by the host-side planner (plan.cpp) and the sm_100a kernels.
//
// Round-2 design: clique tables are NEVER materialised. Every step of the reference's per-case pipeline is one kind of
// op, a SUM over a VIEW of a clique (SumOp below): the view lists the clique's entries as rows x cols, cols = the
// configurations of the op's TARGET variables, and every entry is recomputed where it is needed from per-tree constants
// (pristine potential, index maps) and small message rows:
//     entry(r, j) = pristine[r, j] * evidence_mask(r, j) * prod_i 1/norm_i * prod_g message_g[rowpart_g(r) + colpart_g(j)]
//     out[j]      = sum_r entry(r, j)
//   COLLECT of clique C   targets = C's upstream separator, gathers = up messages of C's children
//                         (reference: reset + LoadEvidenceToNodes + CliqueLevelCollection + SeparatorLevelCollectionOptimized)
//   DISTRIBUTE P -> K     targets = separator K, gathers = up messages of P's OTHER children + P's own down ratio
//                         (reference: SeparatorLevelDistribution + CliqueLevelDistributionOptimized; the division by K's up
//                         message is carried out algebraically: K's message is simply left out of the product. The
//                         reference's 0 guard of TableDivision (PotentialTable.cpp:718-734: ratio = 0 where K's up message
//                         is 0) is NOT applied: a down ratio is only ever multiplied into entries of K's clique and into
//                         K's up message itself, and where the up message entry is 0 every one of those is already 0 (a sum
//                         of non-negative terms), so the guarded and the unguarded ratio give the same finite products;
//                         targets the evidence itself rules out are zeroed by the per-target check (cmask). The guard cost
//                         one dependent load per target at the end of every unit: without it ALARM +14 % (flat kernel),
//                         HEPAR2 shape +12 %, PIGS shape +9 %, MUNIN shape +6 % (B200, round 2))
//   MARGINAL of variable v targets = {v}, source = the smallest table containing v (a clique: all child messages + its down
//                         ratio; a separator: up message * down ratio)
// HBM traffic per case is then the messages (sum of separator sizes) instead of the cliques (round 1: every clique entry
// written once and read back 1-2 times).
//
// Data layout in HBM ("scratch"): one TILE = kTileCases = 32 * kW test cases, interleaved. A tile owns `tile_rows` ROWS of
// 32 * kW doubles (lane l holds cases l*kW .. l*kW+kW-1 of the tile: one 8*kW-byte vector per lane, 128-bit accesses for
// kW = 2), so every access a warp makes is one fully coalesced row no matter how irregular the table index map is, and all
// index arithmetic is warp-uniform. Row regions of a tile:
//   [mup_row0 ..)  one row per separator entry, up message   (reference: Separator::p_table after Collect)
//   [dn_row0 ..)   one row per separator entry, down ratio   (reference: Separator::p_table after Distribute / old)
//   [inv_row0 ..)  one row per clique: 1 / (sum of the clique after Collect)  (reference: Normalize(), PotentialTableBase.cpp:433-445)
//   [mg_row0 ..)   one row per (variable, state): unnormalised marginal
//   [lam_row0 ..)  one row per (variable, state): evidence indicator 1.0 / 0.0 (written only for the variables some op gathers)
//   [part_row0 ..) / [normpart_row0 ..)  partial sums of split ops (several warps per tile)
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace fbn {

constexpr int kLanes = 32;     // lanes per warp
#ifndef FBN_W
#define FBN_W 2                // test cases per lane: 2 -> 16-byte (128-bit) row accesses
#endif
constexpr int kW = FBN_W;
constexpr int kTileCases = kLanes * kW;   // test cases per tile
constexpr int kRowDoubles = kLanes * kW;  // doubles per scratch row
#ifndef FBN_OCC
#define FBN_OCC 16   // fused kernel: resident warps per SM the register allocation is bounded for
#endif
// Column blocking: adjacent targets a warp processes together (register blocking: cb independent row loads per index
// load). Chosen per plan: fused, 1 warp per tile: 1; fused, >= 2 warps per tile: 2; level: 4.
constexpr int kMaxColBlock = 8;

// --- device-visible op descriptors -------------------------------------------------------------------------------
enum SumKind : int32_t { SUM_COLLECT = 0, SUM_DISTRIBUTE = 1, SUM_MARGINAL = 2 };

// One sum over a rows x cols view of a table (see the header comment).
//
// Evidence never costs a per-entry mask test. A variable of the view that none of the gathered messages contains
// ("uncovered": the messages of covered variables are already zero where the evidence rules a state out) is either
//   FOLDED into the potentials: the planner sums the variable out of the view for every EVIDENCE PATTERN (0 = not observed,
//     1 + d = observed in state d, dim + 1 = observed in no valid state) and stores nfx = prod (dim + 2) values per entry; a
//     case looks its pattern index fx up once per op (fold list below, per-tile pattern bytes) and reads potX[entry*nfx + fx];
//   (row variables only; an uncovered TARGET variable of a collect / distribute view is checked once per finished sum, cmask list below)
//   or GATHERED as a unary message: the tile's lambda rows (lam_row0 + var offset + state: 1.0 allowed / 0.0 ruled out) are one
//     more gather of the entry (used when the pattern table would get too big).
struct SumOp {
    int64_t pot_off;      // into pot[]: potX, unit-major: (((j / cb) * rows + r) * cb + j % cb) * nfx + fx (16-byte aligned, last
                          //   target block zero-padded), -1 = all ones
    int64_t fold_off;     // into idx32[]: nfold pairs {variable, pattern stride}
    int64_t cmask_off;    // into idx32[]: ncmask variables, then cols x ncmask pattern bytes (1 + the target's digit of the variable):
                          //   uncovered TARGET variables of a collect / distribute view; a finished sum is zeroed when the case
                          //   observed another state
    int64_t rowpart_off;  // into idx32[]: rows*ng  (absolute scratch row of gather g's source, row part)
    int64_t colpart_off;  // into idx32[]: cols*ng  (column part)
    int64_t invrow_off;   // into idx32[]: ninv absolute scratch rows whose product scales the sums (1/norm of cliques)
    int64_t flat_off;     // flat plans (small trees, one warp per tile): into flatc[] (int4 records), -1 = none. One record
                          //   per entry, TARGET-major (all rows of target 0, then target 1, ...): first int4 = {index of the
                          //   entry's first potX value, absolute scratch row of gather 0, of gather 1, of gather 2}; ops with
                          //   more than three gathers append one int4 of further rows
    int64_t perm_off;     // into idx32[]: visiting order of the column blocks, -1 = storage order
    int32_t kind;         // SumKind
    int32_t dst_row;      // scratch row of target 0: up message (root: dummy row) / down ratio / unnormalised marginal
    int32_t aux_row;      // SUM_DISTRIBUTE: row of entry 0 of the child's up message (informational: the 0 guard is not applied)
    int32_t inv_row;      // SUM_COLLECT: scratch row receiving 1/sum(clique)
    int32_t rows, cols;
    int32_t nfold, nfx;   // folded variables, evidence patterns per entry (1 = plain potentials)
    int32_t ncmask;
    int32_t ng;           // gathers per entry
    int32_t ninv;
    int32_t normslot;     // SUM_COLLECT: index among the collect ops of its phase
    // work decomposition: unit u in [0, ceil(cols/cb)*chunks): column block b = u / chunks, rows [(u % chunks)*rows_per_chunk, ...)
    int32_t chunks, rows_per_chunk;
    int32_t pslot0;       // chunks > 1: partial slot of (col, chunk) = pslot0 + col*chunks + chunk
    int32_t rot;          // units of earlier ops in the phase (fused mode: mod team): keeps the round-robin over warps going
    int32_t pad0;
};

enum FinType : int32_t { FIN_SUM = 0, FIN_NORM = 1 };
struct FinItem {
    int32_t type;      // FIN_SUM: dst = sum of partials; FIN_NORM: dst = 1 / sum of the norm shares
    int32_t dst_row;
    int32_t aux_row;
    int32_t slot0;     // first partial slot (FIN_NORM: normslot)
    int32_t nslots;
    int32_t pad0, pad1, pad2;
};

struct Phase {
    int32_t kind;  // 0 = collect level, 1 = distribute level (+ marginals); informational, the ops are all SumOps
    int32_t op_begin, op_end;  // into oplist[] (indices into ops[])
    int32_t fin_begin, fin_end;
    int32_t n_normslots;
    int32_t pad;
    int32_t n_units;           // warp work units of the phase (level mode: sizes the grid)
};

struct VarInfo {
    int32_t off;  // prefix sum of dims
    int32_t dim;
};

// Kernel parameter block (device pointers)
struct DevPlan {
    const double *pot;
    const int32_t *idx32;
    const int32_t *lamvars;  // variables whose lambda rows some op gathers
    int32_t n_lamvars;
    int32_t lam_row0;        // scratch row of (variable 0, state 0)'s lambda row
    const int32_t *flatc;   // flat plans: entry records (int4 aligned)
    const SumOp *ops;
    const int32_t *oplist;
    const FinItem *fins;
    const Phase *phases;
    const VarInfo *vars;
    int32_t n_phases;
    int32_t n_vars, sum_dims, query_var;
    int32_t tile_rows;   // rows per tile
    int32_t mg_row0;
    int32_t max_normslots, max_pslots;
    int32_t team;        // warps per tile (level mode: nominal gang warps the split was planned for)
    int32_t cb;          // targets per unit the plan was made for
    int32_t part_row0;      // first of max_pslots rows (partial sums of split ops)
    int32_t normpart_row0;  // first of max_normslots * nb_max rows: row (normslot * nb_max + share)
    int32_t nb_max;         // norm shares per collect op (fused: team warps; level: CTAs per tile a phase may be launched with)
    int32_t pat_bytes;      // n_vars * kTileCases: evidence pattern byte of (variable, case of the tile)
    int32_t flat;           // 1: every op has flat records (jt_flat_kernel)
};

// --- host-side plan ----------------------------------------------------------------------------------------------
struct HostPlan {
    std::vector<double> pot;
    std::vector<int32_t> idx32;
    std::vector<int32_t> flatc;
    std::vector<int32_t> lamvars;
    bool flat = false;
    std::vector<SumOp> ops;
    std::vector<int32_t> oplist;
    std::vector<FinItem> fins;
    std::vector<Phase> phases;
    std::vector<VarInfo> vars;
    int32_t n_vars = 0, n_cliques = 0, n_seps = 0, sum_dims = 0, query_var = 0;
    int32_t n_levels = 0;
    int32_t tile_rows = 0, mup_row0 = 0, dn_row0 = 0, inv_row0 = 0, mg_row0 = 0, lam_row0 = 0;
    int32_t max_normslots = 0, max_pslots = 0;
    int32_t team = 1, cb = 1;
    bool level = false;          // level mode (one launch per phase, many CTAs per tile)
    int32_t part_row0 = 0, normpart_row0 = 0, nb_max = 1;
    int64_t n_units = 0;  // warp work units per tile (for tests / info)
    int64_t clique_entries = 0, sep_entries = 0, parent_entries = 0, marginal_entries = 0;
    int64_t view_entries = 0;    // sum over all ops of rows*cols (entries recomputed per case)
    int64_t gather_rows = 0;     // sum over all ops of rows*cols*ng (message rows gathered per case, before register blocking)
    int64_t folded_vars = 0, lambda_gathers = 0;  // uncovered (view, variable) pairs folded into pattern tables / gathered as lambda rows
    // per-variable marginal source, for tests: table index (clique c, or n_cliques + s)
    std::vector<int32_t> marg_source;
    std::vector<int64_t> clique_off;   // [n_cliques + 1] offsets of the cliques' pristine potentials in pot[]
    double alg_bytes_per_case = 0, kernel_bytes_per_case = 0;
};

}  // namespace fbn

struct fbn_tree_desc;
namespace fbn {
constexpr int kLevelWarps = 8;  // warps per CTA of the level-mode kernels
// Validates the tree and builds the host plan. Fused mode (level = false): `team` warps (power of two <= 32) of ONE CTA
// work on a tile. Level mode: `team` is the nominal number of warps (any positive count, over several CTAs) that work on
// a tile in one phase launch. Returns empty string on success.
std::string build_plan(const fbn_tree_desc &t, int team, HostPlan &out, bool level = false);
}
