/*
 * fastbn_b200 — C ABI of the B200-native junction-tree inference hot path.
 *
 * The reference (Tonanguyxiro/FastBN) has no FFI layer: its boundary for this path is the C++ class API
 *   JunctionTree(Network*, Dataset*, bool)            include/JunctionTree.h:40
 *   virtual double EvaluateAccuracy(string, int)      include/Inference.h:44, src/JunctionTree.cpp:57-129
 * and everything below replaces what happens INSIDE JunctionTree::PredictUseJTInfer (src/JunctionTree.cpp:1473-1534)
 * for a whole batch of test cases at once. A maintainer of the reference binds these entry points from
 * JunctionTree.cpp; our own host-side mirror (fastbn_b200/host/) does exactly that.
 *
 * Plain C types only; every pointer is caller-owned; nothing is retained after a call returns except the plan.
 * All functions return 0 on success, non-zero on error; fbn_last_error() gives the message (thread-local).
 * There is no CPU fallback: fbn_plan_create fails with FBN_ERR_NO_DEVICE when no sm_100 device is usable.
 */
#ifndef FASTBN_B200_H
#define FASTBN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FBN_OK 0
#define FBN_ERR_INVALID 1   /* malformed tree / arguments */
#define FBN_ERR_NO_DEVICE 2 /* no CUDA device (the product path never falls back to the CPU) */
#define FBN_ERR_CUDA 3      /* a CUDA runtime call failed */
#define FBN_ERR_NOMEM 4
#define FBN_ERR_COMM 5      /* NCCL could not be loaded or an NCCL call failed (fbn_comm_*, fbn_reduce) */

/* Execution modes. FUSED: one persistent kernel, one CTA per tile (fbn_plan_info.tile_cases test cases), whole tree per
 * launch (small / medium trees). LEVEL: one launch per tree level over all resident tiles, many CTAs per tile, big tables
 * split across CTAs with a two-stage reduction (MUNIN-shape). AUTO picks by table sizes. */
#define FBN_MODE_AUTO 0
#define FBN_MODE_FUSED 1
#define FBN_MODE_LEVEL 2

/*
 * Flat, immutable description of a calibrated-to-be junction tree, i.e. of the reference's
 * JunctionTreeStructure + JunctionTree state right after the JunctionTree ctor (src/JunctionTree.cpp:3-46):
 *  - tables [0, n_cliques) are tree->vector_clique_ptr_container[i]->p_table, tables
 *    [n_cliques, n_cliques+n_separators) are tree->vector_separator_ptr_container[i]->p_table;
 *  - table_vars is PotentialTableBase::vec_related_variables in STORAGE order (first = most significant digit,
 *    src/PotentialTableBase.cpp:599-607) — for a non-root clique the variables of its upstream separator are the
 *    trailing ones, in separator order (JunctionTree::ReorganizeTableStorage, src/JunctionTree.cpp:235-281);
 *  - clique_potentials are PotentialTableBase::potentials of every clique, concatenated in clique order
 *    (separator tables are all-ones in the reference and are not passed);
 *  - sep_parent / sep_child are Separator::ptr_upstream_clique / ptr_downstream_cliques[0] as clique indices;
 *  - root is jt_root; query_var is Inference::query_index.
 */
typedef struct fbn_tree_desc {
    int32_t n_vars;
    const int32_t *var_dims;          /* [n_vars] DiscreteNode::GetDomainSize */
    int32_t n_cliques;
    int32_t n_separators;             /* == n_cliques - 1 for a tree, fewer for a forest is NOT supported */
    const int32_t *table_nvars;       /* [n_cliques + n_separators] */
    const int32_t *table_vars;        /* concatenated, sum(table_nvars) entries */
    const double *clique_potentials;  /* concatenated, sum over cliques of prod(dims) entries */
    const int32_t *sep_parent;        /* [n_separators] */
    const int32_t *sep_child;         /* [n_separators] */
    int32_t root;
    int32_t query_var;
} fbn_tree_desc;

typedef struct fbn_plan fbn_plan; /* opaque: device-resident flattened tree + schedule */

typedef struct fbn_plan_options {
    int32_t device;        /* CUDA device ordinal; -1 = current device */
    int32_t team_warps;    /* fused mode: warps (power of two <= 32) of the one CTA that owns a tile;
                              level mode: nominal warps per tile over all CTAs of a phase launch; 0 = choose */
    int32_t max_resident;  /* cap on resident tiles (scratch = max_resident x tile bytes); 0 = choose */
    int32_t mode;          /* FBN_MODE_AUTO / FBN_MODE_FUSED / FBN_MODE_LEVEL */
    int32_t cases_hint;    /* > 0: the caller will run about this many cases per call; scratch is sized for no more tiles */
    int32_t reserved[3];
} fbn_plan_options;

typedef struct fbn_plan_info {
    int32_t n_vars, n_cliques, n_separators, sum_dims;
    int32_t n_levels;            /* == reference max_level (cliques and separators alternate) */
    int32_t team_warps;          /* chosen warps per tile (see fbn_plan_options) */
    int32_t resident_tiles;      /* tiles in flight (grid size) */
    int32_t n_phases;
    int32_t mode;                /* FBN_MODE_FUSED or FBN_MODE_LEVEL */
    int32_t launches_per_wave;   /* kernels per wave of resident tiles (1 in fused mode) */
    int64_t clique_entries;      /* sum |clique| */
    int64_t separator_entries;   /* sum |separator| */
    int64_t parent_entries;      /* sum over edges of |parent clique| */
    int64_t marginal_entries;    /* sum over vars of |table the marginal is read from| */
    int64_t scratch_bytes;       /* device scratch for all resident tiles */
    int64_t plan_bytes;          /* device bytes of per-tree constants */
    double alg_bytes_per_case;   /* SURVEY.md 8(d) algorithmic bytes per case (fixed-shape fp64 tables), evidence excluded */
    double kernel_bytes_per_case;/* bytes the fused kernel itself moves per case (scratch r/w + output) */
    int32_t flat_records;        /* small trees, per-entry records: 1 = jt_flat_kernel (one warp per tile, tile state in HBM / L2),
                                    2 = jt_resident_kernel (one CTA per tile, the tile's live messages in shared memory); else 0 */
    int32_t kernel_rev;          /* revision of the kernels' memory behaviour: ncu traffic figures are valid for one revision only */
    int32_t tile_cases;          /* test cases per tile (32 lanes x cases per lane) */
    int32_t reserved;
    int64_t view_entries;        /* clique entries recomputed per case, summed over all ops (collect + distribute + marginals) */
    int64_t gather_rows;         /* message rows gathered per case, before register blocking */
} fbn_plan_info;

typedef struct fbn_stats {
    double kernel_ms;      /* device time of the inference kernel(s), CUDA events on the launching stream */
    double total_ms;       /* fbn_run only: wall time of the whole call incl. copies */
    int64_t n_launches;    /* kernels launched by this call */
    int64_t h2d_bytes, d2h_bytes;
} fbn_stats;

const char *fbn_last_error(void);
int fbn_device_count(void);
/* Creates the CUDA context of `device` (about a second on a B200 box) so that a caller can overlap it with its own file
 * parsing from another thread; fbn_plan_create does the same implicitly. */
int fbn_device_warmup(int32_t device);

/* Build the device plan once per tree (replaces BackUpJunctionTree + the per-level "pre" phases,
 * src/JunctionTree.cpp:289-300, :586-639, :834-880). opt may be NULL. */
int fbn_plan_create(const fbn_tree_desc *tree, const fbn_plan_options *opt, fbn_plan **out);
void fbn_plan_destroy(fbn_plan *plan);
int fbn_plan_get_info(const fbn_plan *plan, fbn_plan_info *info);

/*
 * Run n_cases test cases (replaces the loop of src/JunctionTree.cpp:1517-1528 and everything it calls).
 * Evidence is CSR: case i observes (ev_vars[k], ev_vals[k]) for k in [ev_offsets[i], ev_offsets[i+1]).
 * Pairs with var >= n_vars or var < 0 are ignored (src/JunctionTree.cpp:326-331).
 * labels_out[i]      = Inference::ArgMax of the query variable's posterior (src/Inference.cpp:92-102), -1 if none > 0
 * posteriors_out     = n_cases x sum_dims doubles, row i = probs_one_sample of case i flattened in variable order,
 *                      zeros for evidence variables (src/JunctionTree.cpp:1392-1406). May be NULL (labels only).
 * HOST buffers (pageable or pinned); copies are chunked and overlapped with compute. stats may be NULL.
 */
int fbn_run(fbn_plan *plan, int64_t n_cases, const int64_t *ev_offsets, const int32_t *ev_vars,
            const int32_t *ev_vals, int32_t *labels_out, double *posteriors_out, fbn_stats *stats);

/* Same, but every buffer is DEVICE memory on the plan's device and the work is enqueued on `stream`
 * (a cudaStream_t, NULL = default stream); returns after enqueueing unless stats != NULL (then it synchronises
 * to fill kernel_ms). */
int fbn_run_device(fbn_plan *plan, int64_t n_cases, const int64_t *ev_offsets, const int32_t *ev_vars,
                   const int32_t *ev_vals, int32_t *labels_out, double *posteriors_out, void *stream,
                   fbn_stats *stats);

/*
 * fbn_run + scoring in one pipelined call: what JunctionTree::EvaluateAccuracy needs (src/JunctionTree.cpp:57-129:
 * PredictUseJTInfer for every case, CalculateMSE / CalculateHellingerDistance per case src/Inference.cpp:153-192, Accuracy
 * src/Inference.cpp:46-61). HOST buffers. golden = n_cases x sum_dims, flattened ground_truth_probability_tables
 * (src/Inference.cpp:108-146: first entry of an evidence node is -1); truths = ground_truths. posteriors_out may be NULL
 * (posteriors then never leave the device). out4 = {n_correct, n_total, sum_mse, sum_hd}.
 */
int fbn_evaluate(fbn_plan *plan, int64_t n_cases, const int64_t *ev_offsets, const int32_t *ev_vars,
                 const int32_t *ev_vals, const double *golden, const int32_t *truths, int32_t *labels_out,
                 double *posteriors_out, double *out4, fbn_stats *stats);

/*
 * On-device scoring against golden posteriors (replaces Inference::CalculateMSE / CalculateHellingerDistance /
 * Accuracy, src/Inference.cpp:46-61,153-206). golden is n_cases x sum_dims with golden[i][off_v] <= 0 meaning
 * "skip node v" (src/Inference.cpp:160). out4 = {n_correct, n_total, sum_mse, sum_hd} as doubles, host memory.
 * All array arguments are device pointers.
 */
int fbn_score_device(fbn_plan *plan, int64_t n_cases, const double *posteriors, const int32_t *labels,
                     const double *golden, const int32_t *truths, double *out4_host, void *stream);

/*
 * The scoring of fbn_score_device for callers that cannot allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc). No counterpart in the reference.
 * Needs no plan. Calls no cudaMalloc / cudaFree / allocating library routine and does not synchronise: the two kernels are
 * enqueued on `stream` and the result {n_correct, n_total, sum_mse, sum_hd} is written to out4_device (DEVICE memory, 4 doubles).
 * The per-case scratch that fbn_score_device allocates is a static __device__ array of FBN_STATIC_SCORE_MAX_CASES x 3 doubles
 * (1.5 MiB, held by every GPU that loads the library until the process exits); n_cases beyond that is FBN_ERR_INVALID with a
 * message. All calls share that array: calls that overlap on the GPU (different streams or host threads) race — serialise
 * them. The first call on a device looks the array's device address up (cudaGetSymbolAddress): make it outside a capture.
 * var_offsets: DEVICE array of n_vars + 1 int32 prefix sums of the variables' dims (var_offsets[n_vars] = sum_dims); the
 * other arrays as in fbn_score_device, all device pointers on `device`.
 */
#define FBN_STATIC_SCORE_MAX_CASES 65536
int fbn_score_device_static(int32_t device, int64_t n_cases, int32_t n_vars, const int32_t *var_offsets,
                            const double *posteriors, const int32_t *labels, const double *golden, const int32_t *truths,
                            double *out4_device, void *stream);

/*
 * Multi-GPU: test cases are independent, so every GPU runs its own contiguous slice with a replicated plan and NO data-path
 * collective (reference loop src/JunctionTree.cpp:1517-1528 carries only the mse / hd accumulators between cases). The one
 * exchange step is the sum of out4 = {n_correct, n_total, sum_mse, sum_hd} over the ranks: ONE ncclAllReduce of four doubles
 * (counts are exact below 2^53). The reference has no counterpart (its USE_MPI flag is a dead stub, CMakeLists.txt:10-16).
 *
 *   one process per GPU : rank 0 calls fbn_comm_unique_id and hands the 128 bytes to the other ranks by any means (file,
 *                         socket, torch.distributed store); every rank calls fbn_comm_create, later fbn_reduce.
 *   one process, n GPUs : fbn_comm_create_all gives one communicator per device; fbn_reduce must then be called
 *                         concurrently, one host thread per communicator (our C++ JunctionTree does that).
 * NCCL is loaded at run time (libnccl.so.2; a copy the process already loaded is preferred); FBN_ERR_COMM if absent.
 */
#define FBN_COMM_ID_BYTES 128
typedef struct fbn_comm fbn_comm;
int fbn_comm_unique_id(void *out128);
int fbn_comm_create(int32_t device, int32_t rank, int32_t world, const void *id128, fbn_comm **out);
int fbn_comm_create_all(int32_t n_devices, const int32_t *devices, fbn_comm **out /* [n_devices] */);
/* in-place sum over all ranks of four host doubles; collective: every rank must call it; returns when the result is there */
int fbn_reduce(fbn_comm *comm, double inout4[4]);
/* same on four DEVICE doubles, enqueued on `stream` (no synchronisation) */
int fbn_reduce_device(fbn_comm *comm, double *dev4, void *stream);
int fbn_comm_info(const fbn_comm *comm, int32_t *rank, int32_t *world, int32_t *nccl_version);
void fbn_comm_destroy(fbn_comm *comm);

#ifdef __cplusplus
}
#endif
#endif /* FASTBN_B200_H */
