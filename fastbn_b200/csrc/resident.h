// Resident-tile plan: small trees whose LIVE messages of a 64-case tile fit into one SM's shared memory (ALARM: 174 KB).
// jt_resident_kernel (resident_kernel.cu) then runs a tile with one CTA per SM and never touches HBM between the evidence
// it reads and the posteriors it writes: message rows live in shared memory, and the per-tree constants of a phase (work
// items, per-entry records, potentials) are streamed into a double-buffered shared-memory block by cp.async.bulk (TMA) while
// the previous phase is computed.
//
// Built from a FLAT plan (plan.h, team 1: one record per view entry) by build_resident():
//   * the per-target evidence checks of a view (cmask, plan.h: at most kResMaxChecks variables) read the case's pattern bytes
//     once per item, not per target;
//   * Normalize (PotentialTableBase.cpp:433-445) is left to the final marginals: a bound over the cliques' potentials shows
//     that no message can leave the range of a double (else the tree is not eligible and the flat kernel runs it), so there
//     are no 1/norm rows, no scaling and no second barrier per collect level; ops nobody reads (the root's collect) are dropped;
//   * liveness: every message (an op's targets) gets shared-memory rows only from the phase that writes it to the last phase
//     that reads it (best-fit over phases): ALARM's 775 scratch rows become ~350;
//   * a STATIC schedule: the ops of a phase are cut into pieces (ranges of targets = contiguous records), the pieces are
//     dealt to the CTA's warps by longest-processing-time-first; marginals are normalised and written out by OUT items of
//     the following phase, the next tile's evidence patterns are loaded by an EVID item;
//   * one contiguous constant block per phase: header, items, op descriptors, records, potentials; every offset in it is a
//     byte offset (into the block, or into the tile's state), so the kernel's inner loop is LDS + multiply-add only.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "plan.h"

namespace fbn {

enum ResItemKind : uint32_t { RES_SUM = 0, RES_OUT = 1, RES_EVID = 2, RES_SUMC = 3 /* SUM whose op checks target variables against the evidence */ };

// Block of one phase (uint32 words; all *_off are BYTE offsets from the block start unless they say "state"):
//   [0] n_items  [1] 0  [2] items_off  [3] block_bytes
//   [4 .. 4 + warps]  warp_begin[w]: first item of warp w (items of a warp are adjacent), warp_begin[warps] = n_items
//   items    : ResItem (64 B each)
//   descs    : per op {nfold, ncm, want_off, 0} + nfold x {variable, pattern stride in BYTES} + ncm checked target variables,
//              padded to 16 bytes; want_off -> cols x ncm words: the pattern byte (1 + state) target j has of checked variable k
//   records  : per entry, target-major: int4 {pot_off, state offsets of gathers 0..2} (+ int4 {gathers 3..6} when the op has
//              more than 3 gathers)
//   pots     : doubles (nfx per entry), 16-byte aligned per op
struct ResItem {
    uint32_t kind_ng;    // kind | ng << 8 | 1 << (16 + min(ng, 5)) (SUM items: the bit names the kernel body)
    uint32_t rec_off;    // SUM: first record.                     OUT: variable
    uint32_t n_targets;  // SUM: targets of the piece.             OUT: dim
    uint32_t rows;       // SUM: records per target.               OUT: posterior column of state 0 (VarInfo::off)
    uint32_t dst_off;    // SUM: state offset of the first target. OUT: state offset of the unnormalised marginal of state 0
    uint32_t desc_off;   // SUM: the op's descriptor
    uint32_t j0;         // SUM: first target of the piece
    uint32_t want_off;   // SUM with checks: the want words of target j0 (ncm per target), else 0
    // what the kernel needs of the descriptor before its first record, so that an item's set-up is ONE round of shared-memory
    // loads (item, then pattern bytes) instead of a chain of five (item -> descriptor -> fold pair -> pattern byte, twice):
    uint32_t nfold_ncm;     // folded variables | checked variables << 8
    uint32_t fold_var0;     // the first two folded variables and their strides in BYTES (stride 0: none); ops that fold more
    uint32_t fold_stride0;  // than two go through the descriptor
    uint32_t fold_var1;
    uint32_t fold_stride1;
    uint32_t chk[3];        // checked target variables; unused slots repeat a valid variable (their loads are unconditional)
};
static_assert(sizeof(ResItem) == 64, "ResItem is four 16-byte words");

constexpr int kResMaxChecks = 3;   // checked target variables per op the kernel keeps in registers
constexpr int kResRowBytes = kRowDoubles * 8;   // bytes of one state row (64 cases)

// lambda rows of one variable (evidence indicators 1.0 / 0.0 per state, gathered like a message by the views that could not
// fold the variable into their pattern tables)
struct ResLam {
    uint32_t var, dim, state_off, pad;
};

struct ResidentPlan {
    bool ok = false;
    std::string why;                 // why the tree is not eligible (ok == false)
    int warps = 0;                   // warps per CTA the schedule was made for
    int n_blocks = 0;                // phases of the flat plan with needed ops + 1 (the last block holds the OUT items of the last phase)
    int state_rows = 0;              // shared-memory rows of the tile state (high-water mark of the live rows)
    int max_block_bytes = 0;
    std::vector<uint32_t> blob;      // all blocks, each 16-byte aligned
    std::vector<uint32_t> block_off; // [n_blocks + 1] byte offsets into blob
    int64_t smem_bytes = 0;          // state + 2 pattern buffers + 2 block buffers + mbarriers + dims
    double balance = 0;              // sum of the item costs / (warps x sum over phases of the most loaded warp): schedule quality
    int n_items = 0;
    std::vector<ResLam> lams;
};

// smem_limit: bytes of dynamic shared memory one CTA may use (sharedMemPerBlockOptin)
void build_resident(const HostPlan &flat, int warps, int64_t smem_limit, ResidentPlan &out);

// layout of the CTA's dynamic shared memory (shared by the kernel and the host)
struct ResSmem {
    uint32_t state, pat0, pat1, buf0, buf1, bars, dims, boff, lams, total;
};
inline ResSmem res_smem_layout(int state_rows, int n_vars, int max_block_bytes, int n_blocks, int n_lams) {
    ResSmem L;
    uint32_t o = 0;
    L.state = o; o += (uint32_t)state_rows * kResRowBytes;
    const uint32_t pat = ((uint32_t)n_vars * kTileCases + 15u) & ~15u;
    L.pat0 = o; o += pat;
    L.pat1 = o; o += pat;
    L.buf0 = o; o += ((uint32_t)max_block_bytes + 15u) & ~15u;
    L.buf1 = o; o += ((uint32_t)max_block_bytes + 15u) & ~15u;
    L.bars = o; o += 16;
    L.dims = o; o += ((uint32_t)n_vars + 15u) & ~15u;   // one byte per variable: its number of states
    L.boff = o; o += (((uint32_t)n_blocks + 1u) * 4u + 15u) & ~15u;   // byte offsets of the blocks in the constant blob
    L.lams = o; o += (uint32_t)n_lams * 16u;   // ResLam list
    L.total = o;
    return L;
}


// kernel parameter block of jt_resident_kernel (device pointers)
struct ResDev {
    const unsigned char *blob;     // constant blocks (global memory, 16-byte aligned)
    const uint32_t *block_off;     // [n_blocks + 1]
    const VarInfo *vars;
    const ResLam *lams;
    long long *clk;                // developer builds (-DFBN_RES_CLOCK_BUILD=1) with FBN_RES_CLOCK=1: cycle stamps of CTA 0, else NULL
    int32_t n_blocks, n_vars, sum_dims, query_var, n_lams;
    ResSmem L;
};

}  // namespace fbn
