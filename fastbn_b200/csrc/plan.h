// Flattened junction-tree plan shared by the host-side planner (plan.cpp) and the sm_100a kernels (kernels.cu).
//
// Data layout in HBM ("scratch"): one TILE = 32 test cases, interleaved. A tile owns `tile_rows` ROWS of
// 32 doubles (256 B, lane = case), so every access a warp makes is one fully coalesced 256-byte row no matter how
// irregular the table index map is, and all index arithmetic is warp-uniform. Row regions of a tile:
//   [cl_row0 ..)   one row per clique-table entry           (reference: Clique::p_table.potentials)
//   [mup_row0 ..)  one row per separator entry, up message   (reference: Separator::p_table after Collect)
//   [dn_row0 ..)   one row per separator entry, down ratio   (reference: Separator::p_table after Distribute / old)
//   [inv_row0 ..)  one row per clique: 1 / (sum of the clique after Collect)  (reference: Normalize(), PotentialTableBase.cpp:433-445)
//   [mg_row0 ..)   one row per (variable, state): unnormalised marginal
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace fbn {

constexpr int kLanes = 32;     // cases per tile == warp width
#ifndef FBN_OCC
#define FBN_OCC 32   // fused kernel: resident warps per SM the register allocation is bounded for
#endif
// Column blocking: adjacent columns (collect) / target entries (pull) a warp processes together (register blocking:
// cb independent row loads per index load). Chosen per plan from measurements on B200 (profiles/):
//   fused, 1 warp per tile (ALARM-size)  cb 1      fused, >= 2 warps per tile  cb 2
//   level, largest table < 2^20 entries  cb 4      level, larger (MUNIN-shape) cb 8
constexpr int kMaxColBlock = 8;

// --- device-visible op descriptors -------------------------------------------------------------------------------
// COLLECT of clique C (fuses reference steps: reset from backup [JunctionTree.cpp:302-311], evidence masking
// [LoadEvidenceToNodes :386-457], clique *= child messages [CliqueLevelCollection :829-941], column sums into the
// upstream separator [SeparatorLevelCollectionOptimized :1056-1148]).
// The clique table is viewed as rows x cols, cols = |upstream separator| (trailing variables), root: cols = 1.
struct CollectOp {
    int64_t pot_off;      // into pot[]: pristine potentials (rows*cols doubles, per-tree constant)
    int64_t rowlam_off;   // into lam16[]: rows*nrest lambda-bit indices of the leading (non-separator) variables
    int64_t collam_off;   // into lam16[]: cols*ncolv lambda-bit indices of the trailing (separator) variables
    int64_t rowpart_off;  // into idx32[]: rows*nchild  (absolute scratch row of child k's message, row part)
    int64_t colpart_off;  // into idx32[]: cols*nchild  (column part)
    int64_t invrow_off;   // into idx32[]: nchild absolute scratch rows holding 1/norm of each child clique
    int64_t flat_off;     // flat plans (small trees, one warp per tile): into flatc[] (int4 records), -1 = none. One record per
                          //   entry, COLUMN-major (all rows of column 0, then column 1, ...): {pristine potential (2 x 32 bit),
                          //   absolute scratch row of child 0's message entry, of child 1's}; ops with 3-4 children use two
                          //   int4 per entry, the second = {child 2's row, child 3's row, 0, 0}
    int64_t perm_off;     // into idx32[]: visiting order of the column blocks (unit u works on block perm[u / chunks]), -1 =
                          //   storage order. The planner orders blocks so that blocks visited together gather the same child
                          //   message rows (column variables outside the child separators vary fastest): keeps the gathers in L2.
    int32_t cl_row;       // scratch row of entry 0 of this clique
    int32_t mup_row;      // scratch row of entry 0 of this clique's upward message (root: dummy row)
    int32_t inv_row;      // scratch row receiving 1/sum(clique)
    int32_t rows, cols;
    int32_t nrest, ncolv;
    int32_t nchild;
    int32_t normslot;     // index of this op among the collect ops of its phase
    // work decomposition: unit u in [0, ceil(cols/cb)*chunks): column block b = u / chunks (columns [cb*b, cb*b+cb)),
    // rows [(u % chunks)*rows_per_chunk, ...)
    int32_t chunks, rows_per_chunk;
    int32_t pslot0;       // chunks > 1: partial slot of (col, chunk) = pslot0 + col*chunks + chunk
    int32_t rot;          // units of earlier ops in the phase (fused mode: mod team): keeps the round-robin over warps going
    int32_t rowmask_off;  // shared-memory word offsets of this op's per-tile evidence mask tables (rows / cols words),
    int32_t colmask_off;  //   -1: tables do not fit, fall back to per-entry mask evaluation
    int32_t pad0, pad1;
};

// PULL: out[q] = sum_r src[tbase[q].x + trest[r].x] * fac[tbase[q].y + trest[r].y]
//  kind 1 (distribute, reference SeparatorLevelDistribution :700-816 + CliqueLevelDistributionOptimized :1150-1238):
//          src = parent clique, fac = parent's down ratio; dn[q] = mup[q]==0 ? 0 : out[q]/mup[q]
//  kind 2 (marginal of one variable, reference GetProbabilitiesOneNode :1392-1454): mg[q] = out[q]
struct PullOp {
    int64_t tbase_off;  // into idx32[] (int2 pairs): nT pairs
    int64_t trest_off;  // into idx32[] (int2 pairs): nR pairs
    int64_t flat_off;   // flat plans: into flatp[] (int2 records), -1 = none: nT x nR pairs {absolute src row, absolute fac row},
                        //   target-major
    int32_t src_row;
    int32_t fac_row;
    int32_t mup_row;    // kind 1 only
    int32_t dst_row;
    int32_t nT, nR;
    int32_t kind;
    int32_t chunks, rows_per_chunk, pslot0, rot;  // as in CollectOp (cols = nT, rows = nR): blocks of cb targets
    int32_t pad;
};


// PULL2 (level mode): the distribute messages of TWO children A and B of one parent clique from ONE pass over the parent, when
// every parent variable belongs to separator A or B. Parent variables split into G12 (in both), G1 (A only), G2 (B only).
// Unit (a, b): G12 configuration a, block b of QB configurations of G1. The warp loops over the G2 configurations r:
//   prod = src[x12[a] + x1[g] + x2[r]] * fac[col...]      (QB values g of the block)
//   accA[g] += prod                      -> complete sums of QB entries of separator A (divided by mup, written to dn)
//   partB(a, r, b) = sum_g prod          -> ONE partial row per r; separator B's entry (a, r) = sum over the blocks b, by a
//                                           FIN_D item (reads nblk1 adjacent partial rows)
// The parent is read once instead of twice; the partials add 2/QB of a read.
struct Pull2Op {
    int64_t a_off;    // idx32 (int4 per G12 configuration): {src row offset, fac row offset, A entry offset, B entry offset}
    int64_t g1_off;   // idx32 (int4 per G1 configuration):  {src row offset, fac row offset, A entry offset, 0}
    int64_t g2_off;   // idx32 (int4 per G2 configuration):  {src row offset, fac row offset, B entry offset, 0}
    int32_t src_row, fac_row;
    int32_t mupA_row, dstA_row;
    int32_t n12, n1, n2, nblk1;
    int32_t pslot0;   // partial slot of (B entry t, block b) = pslot0 + t * nblk1 + b
    int32_t rot;
    int32_t pad0, pad1;
};

enum FinType : int32_t { FIN_MUP = 0, FIN_NORM = 1, FIN_D = 2, FIN_MARG = 3 };
struct FinItem {
    int32_t type;
    int32_t dst_row;   // FIN_MUP: mup row; FIN_NORM: inv row; FIN_D: dn row; FIN_MARG: mg row
    int32_t aux_row;   // FIN_D: mup row (divisor)
    int32_t slot0;     // first partial slot (FIN_NORM: normslot)
    int32_t nslots;    // number of partial slots
    int32_t pad0, pad1, pad2;
};

struct Phase {
    int32_t kind;  // 0 = collect ops, 1 = pull ops
    int32_t op_begin, op_end;  // into oplist[] (indices into cops[] / pops[]; < 0 in a pull phase: p2ops[-1 - entry])
    int32_t fin_begin, fin_end;
    int32_t n_normslots;
    int32_t mask_words;        // shared-memory words of evidence mask tables this phase builds
    int32_t n_units;           // warp work units of the phase (level mode: sizes the grid)
};

struct VarInfo {
    int32_t off;  // prefix sum of dims
    int32_t dim;
};

// Kernel parameter block (device pointers)
struct DevPlan {
    const double *pot;
    const uint16_t *lam16;
    const int32_t *idx32;
    const int32_t *flatc;   // flat plans: COLLECT entry records (int4 aligned)
    const int32_t *flatp;   // flat plans: PULL pair records (int2 aligned)
    const CollectOp *cops;
    const PullOp *pops;
    const Pull2Op *p2ops;   // oplist entries < 0 of a pull phase: p2ops[-1 - entry]
    const int32_t *oplist;
    const FinItem *fins;
    const Phase *phases;
    const VarInfo *vars;
    int32_t n_phases;
    int32_t n_vars, sum_dims, query_var;
    int32_t tile_rows;   // rows per tile
    int32_t mg_row0;
    int32_t max_normslots, max_pslots, max_mask_words;
    int32_t team;        // warps per tile (level mode: nominal gang warps the split was planned for)
    int32_t cb;          // columns / targets per unit the plan was made for
    // level mode only (one launch per phase, many CTAs per tile): split-column partials and per-CTA norm partials live in
    // the tile's scratch rows instead of shared memory; evidence bit words live in a per-tile global array
    int32_t part_row0;      // first of max_pslots rows
    int32_t normpart_row0;  // first of max_normslots * nb_max rows: row (normslot * nb_max + cta)
    int32_t nb_max;         // CTAs per tile a phase may be launched with
    int32_t lam_words;      // sum_dims + n_vars
    int32_t flat;           // 1: every op has flat records (jt_flat_kernel)
};

// --- host-side plan ----------------------------------------------------------------------------------------------
struct HostPlan {
    std::vector<double> pot;
    std::vector<uint16_t> lam16;
    std::vector<int32_t> idx32;
    std::vector<int32_t> flatc, flatp;
    bool flat = false;
    std::vector<CollectOp> cops;
    std::vector<PullOp> pops;
    std::vector<Pull2Op> p2ops;
    std::vector<int32_t> oplist;
    std::vector<FinItem> fins;
    std::vector<Phase> phases;
    std::vector<VarInfo> vars;
    int32_t n_vars = 0, n_cliques = 0, n_seps = 0, sum_dims = 0, query_var = 0;
    int32_t n_levels = 0;
    int32_t tile_rows = 0, cl_row0 = 0, mup_row0 = 0, dn_row0 = 0, inv_row0 = 0, mg_row0 = 0;
    int32_t max_normslots = 0, max_pslots = 0, max_mask_words = 0;
    int32_t team = 1, cb = 1;
    bool level = false;          // level mode (see DevPlan)
    int32_t part_row0 = 0, normpart_row0 = 0, nb_max = 1;
    int64_t n_units = 0;  // warp work units per tile (for tests / info)
    int64_t clique_entries = 0, sep_entries = 0, parent_entries = 0, marginal_entries = 0;
    // per-variable marginal source, for tests: table index (clique c, or n_cliques + s)
    std::vector<int32_t> marg_source;
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
