// C ABI (include/fastbn_b200.h) over the planner + kernels. No CPU fallback anywhere: without a usable CUDA device
// fbn_plan_create fails with FBN_ERR_NO_DEVICE.
#include <chrono>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/fastbn_b200.h"
#include "plan.h"
#include "resident.h"

namespace fbn {
size_t fused_smem_bytes(const DevPlan &P);
cudaError_t fused_prepare(const DevPlan &P, int smem_optin, int *blocks_per_sm);
cudaError_t fused_launch(const DevPlan &P, int grid, double *scratch, int64_t n_cases, const int64_t *ev_off, int64_t ev_base,
                         const int32_t *ev_vars, const int32_t *ev_vals, int32_t *labels, double *post, cudaStream_t st);
size_t flat_smem_bytes(const DevPlan &P);
cudaError_t flat_prepare(const DevPlan &P, int smem_optin, int *blocks_per_sm);
cudaError_t flat_launch(const DevPlan &P, int grid, double *scratch, int64_t n_cases, const int64_t *ev_off, int64_t ev_base,
                        const int32_t *ev_vars, const int32_t *ev_vals, int32_t *labels, double *post, cudaStream_t st);
cudaError_t resident_prepare(int warps, int smem_optin);
cudaError_t resident_launch(const ResDev &R, int warps, int grid, int64_t n_cases, const int64_t *ev_off, int64_t ev_base, const int32_t *ev_vars,
                            const int32_t *ev_vals, int32_t *labels, double *post, cudaStream_t st);
cudaError_t level_launch(const DevPlan &P, const std::vector<Phase> &phases, int resident, double *scratch, uint8_t *patG,
                         int64_t n_cases, const int64_t *ev_off, int64_t ev_base, const int32_t *ev_vars, const int32_t *ev_vals,
                         int32_t *labels, double *post, cudaStream_t st, int64_t *n_launch);
cudaError_t score_cases_launch(int64_t n_cases, int n_vars, int sum_dims, const VarInfo *vars, const double *post, const double *golden,
                               const int32_t *labels, const int32_t *truths, double *out_case, cudaStream_t st);
cudaError_t score_reduce_launch(int64_t n_cases, const double *out_case, double *out4, cudaStream_t st);
}  // namespace fbn

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) {
    g_err = msg;
    return code;
}
int fbn_set_error(int code, const std::string &msg) { return fail(code, msg); }  // for comm.cu

// Revision of the kernels' memory behaviour. bench.py attaches an ncu traffic figure (profiles/traffic.json) to a run only
// when it was captured at the same revision: bump on every change that alters what the kernels read or write.
static const int kKernelRev = 8;
#define CK(call)                                                                                       \
    do {                                                                                               \
        cudaError_t e__ = (call);                                                                      \
        if (e__ != cudaSuccess)                                                                        \
            return fail(FBN_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__));            \
    } while (0)

namespace {
// every entry point switches to the plan's device and puts the caller's current device back on return
struct DeviceGuard {
    int prev = -1;
    DeviceGuard() { if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; cudaGetLastError(); } }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
constexpr int kNBuf = 3;            // pipeline depth of fbn_run
struct ChunkBuf {                   // device + pinned staging for one in-flight chunk
    int64_t *d_off = nullptr;
    int32_t *d_vars = nullptr, *d_vals = nullptr, *d_labels = nullptr;
    double *d_post = nullptr;
    double *d_gold = nullptr, *h_gold = nullptr;   // fbn_evaluate only
    int32_t *d_truth = nullptr, *h_truth = nullptr;
    int64_t *h_off = nullptr;       // pinned staging (rebased copies / pageable callers)
    int32_t *h_vars = nullptr, *h_vals = nullptr, *h_labels = nullptr;
    double *h_post = nullptr;
    int64_t cap_cases = 0, cap_ev = 0;
    bool staged_out = false;
    cudaEvent_t in_done = nullptr, k_done = nullptr, out_done = nullptr;
};
}  // namespace

struct fbn_plan {
    fbn::HostPlan hp;
    fbn::DevPlan dp;
    int device = 0;
    int n_sms = 0, blocks_per_sm = 0, grid = 0;
    std::vector<void *> dev_allocs;
    double *scratch = nullptr;
    uint8_t *patG = nullptr;    // level mode: evidence pattern bytes of the resident tiles
    bool level = false;
    bool resident = false;      // small tree run by jt_resident_kernel (tile state in shared memory, resident.h)
    fbn::ResidentPlan rp;
    fbn::ResDev rd;
    int launches_per_wave = 1;
    int64_t scratch_bytes = 0, plan_bytes = 0;
    cudaStream_t s_in = nullptr, s_k = nullptr, s_out = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    ChunkBuf buf[kNBuf];
    double *score_case = nullptr;
    int64_t score_cap = 0;
    double *score_out4 = nullptr;
    std::mutex mu;              // serialises the calls that use the plan's scratch and buffers
};

extern "C" const char *fbn_last_error(void) { return g_err.c_str(); }

extern "C" int fbn_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

extern "C" int fbn_device_warmup(int32_t device) {
    int ndev = fbn_device_count();
    if (ndev <= 0) return fail(FBN_ERR_NO_DEVICE, "no CUDA device: fastbn_b200 has no CPU fallback");
    if (device < 0 || device >= ndev) return fail(FBN_ERR_INVALID, "device ordinal out of range");
    DeviceGuard guard;
    CK(cudaSetDevice(device));
    CK(cudaFree(nullptr));
    return FBN_OK;
}

template <class T>
static int upload(fbn_plan *p, const std::vector<T> &v, const T **out) {
    void *d = nullptr;
    size_t bytes = std::max<size_t>(v.size(), 1) * sizeof(T);
    CK(cudaMalloc(&d, bytes));
    p->dev_allocs.push_back(d);
    if (!v.empty()) CK(cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    p->plan_bytes += (int64_t)bytes;
    *out = (const T *)d;
    return FBN_OK;
}

static double tree_entries(const fbn_tree_desc *t) {  // sum of clique table sizes
    double entries = 0;
    int64_t pos = 0;
    for (int c = 0; c < t->n_cliques; ++c) {
        double sz = 1;
        for (int i = 0; i < t->table_nvars[c]; ++i) sz *= t->var_dims[t->table_vars[pos + i]];
        pos += t->table_nvars[c];
        entries += sz;
    }
    return entries;
}

// above this many clique entries per case AUTO uses level mode (tile >= ~0.5 GB: too few tiles fit for one CTA per tile)
static const double kLevelAutoEntries = 2.0e6;
static const int64_t kLevelMaxResident = 4096;
// Level mode: tiles in flight by default (FBN_LEVEL_RESIDENT overrides). The CTAs of a phase launch are scheduled tile by tile
// (blockIdx.x = CTA of the tile varies fastest), so a big view keeps the whole GPU on one or two tiles at a time and the
// message rows it gathers over and over (each is re-read |view| / |message| times) stay in the 126 MB L2; more tiles in
// flight only add parallelism to the small phases.
static const int64_t kLevelResident = 32;

static int pick_team(double entries) {
    // fused mode: more warps per tile for bigger tables; 16 resident warps per SM at 128 registers (two cases per lane), so a
    // team of T warps runs 16 / T tiles per SM
    if (entries <= 8192) return 1;
    if (entries <= 131072) return 8;   // HEPAR2 shape: 8 warps 2.62 M cases/s, 4 warps 2.47 M
    return 16;                         // PIGS shape: 16 warps 573 k cases/s, 8 warps 537 k
}

// one inference pass over n_cases cases whose inputs are on the device: fused kernel or the per-level launches
static cudaError_t launch_inference(fbn_plan *p, int64_t n_cases, const int64_t *ev_off, int64_t ev_base, const int32_t *ev_vars,
                                    const int32_t *ev_vals, int32_t *labels, double *post, cudaStream_t st, int64_t *n_launch) {
    if (n_cases <= 0) { *n_launch = 0; return cudaSuccess; }
    if (p->level)
        return fbn::level_launch(p->dp, p->hp.phases, p->grid, p->scratch, p->patG, n_cases, ev_off, ev_base, ev_vars, ev_vals,
                                 labels, post, st, n_launch);
    *n_launch = 1;
    if (p->resident)
        return fbn::resident_launch(p->rd, p->rp.warps, p->grid, n_cases, ev_off, ev_base, ev_vars, ev_vals, labels, post, st);
    if (p->dp.flat)
        return fbn::flat_launch(p->dp, p->grid, p->scratch, n_cases, ev_off, ev_base, ev_vars, ev_vals, labels, post, st);
    return fbn::fused_launch(p->dp, p->grid, p->scratch, n_cases, ev_off, ev_base, ev_vars, ev_vals, labels, post, st);
}

extern "C" int fbn_plan_create(const fbn_tree_desc *tree, const fbn_plan_options *opt, fbn_plan **out) {
    if (!tree || !out) return fail(FBN_ERR_INVALID, "null argument");
    *out = nullptr;
    int ndev = fbn_device_count();
    if (ndev <= 0) return fail(FBN_ERR_NO_DEVICE, "no CUDA device: fastbn_b200 has no CPU fallback");
    int dev = opt && opt->device >= 0 ? opt->device : -1;
    if (dev < 0) CK(cudaGetDevice(&dev));
    if (dev >= ndev) return fail(FBN_ERR_INVALID, "device ordinal out of range");
    DeviceGuard guard;
    CK(cudaSetDevice(dev));
    if (!tree->table_nvars || !tree->table_vars || !tree->var_dims) return fail(FBN_ERR_INVALID, "null array in fbn_tree_desc");
    for (int v = 0; v < tree->n_vars; ++v)
        if (tree->var_dims[v] <= 0) return fail(FBN_ERR_INVALID, "variable with non-positive domain size");
    {
        int64_t pos = 0;
        for (int i = 0; i < tree->n_cliques + tree->n_separators; ++i) {
            for (int k = 0; k < tree->table_nvars[i]; ++k)
                if (tree->table_vars[pos + k] < 0 || tree->table_vars[pos + k] >= tree->n_vars)
                    return fail(FBN_ERR_INVALID, "table variable out of range");
            pos += tree->table_nvars[i];
        }
    }
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, dev));
    size_t free_b = 0, total_b = 0;
    CK(cudaMemGetInfo(&free_b, &total_b));
    double frac = 0.45;  // of the free device memory for the tile scratch: leaves room for the caller's evidence / posterior buffers
    if (const char *e = getenv("FBN_SCRATCH_FRAC")) frac = std::min(0.9, std::max(0.05, atof(e)));
    const int64_t budget = (int64_t)(free_b * frac);
    const double entries = tree_entries(tree);
    int mode = opt ? opt->mode : FBN_MODE_AUTO;
    if (const char *e = getenv("FBN_MODE")) mode = atoi(e);
    if (mode != FBN_MODE_FUSED && mode != FBN_MODE_LEVEL) mode = entries > kLevelAutoEntries ? FBN_MODE_LEVEL : FBN_MODE_FUSED;
    const bool level = mode == FBN_MODE_LEVEL;
    int team = opt && opt->team_warps > 0 ? opt->team_warps : 0;
    if (!team && !level) team = pick_team(entries);
    if (!team && level) {
        // nominal warps per tile: enough CTAs over all resident tiles to fill the GPU several times over. A tile holds the
        // messages only (cliques are recomputed), roughly 2 rows per separator entry; separators are not known here, so the
        // estimate is refined after planning (resident is computed from the plan's real tile size below)
        int64_t res = kLevelResident;
        if (opt && opt->max_resident > 0) res = std::min<int64_t>(res, opt->max_resident);
        int64_t nb = (8 * (int64_t)prop.multiProcessorCount + res - 1) / res;
        nb = std::max<int64_t>(2, std::min<int64_t>(nb, 512));
        team = (int)nb * fbn::kLevelWarps;
    }

    fbn_plan *p = new fbn_plan();
    p->device = dev;
    p->level = level;
    std::string err = fbn::build_plan(*tree, team, p->hp, level);
    if (!err.empty()) {
        delete p;
        return fail(FBN_ERR_INVALID, err);
    }
    const fbn::HostPlan &h = p->hp;
    fbn::DevPlan &d = p->dp;
    int rc;
#define UP(field) if ((rc = upload(p, h.field, &d.field)) != FBN_OK) { fbn_plan_destroy(p); return rc; }
    UP(pot) UP(idx32) UP(flatc) UP(lamvars) UP(ops) UP(oplist) UP(fins) UP(phases) UP(vars)
#undef UP
    d.n_phases = (int32_t)h.phases.size();
    d.n_vars = h.n_vars; d.sum_dims = h.sum_dims; d.query_var = h.query_var;
    d.tile_rows = h.tile_rows; d.mg_row0 = h.mg_row0;
    d.max_normslots = h.max_normslots; d.max_pslots = h.max_pslots; d.team = h.team; d.cb = h.cb;
    d.n_lamvars = (int32_t)h.lamvars.size(); d.lam_row0 = h.lam_row0;
    d.flat = h.flat ? 1 : 0;
    d.part_row0 = h.part_row0; d.normpart_row0 = h.normpart_row0; d.nb_max = h.nb_max; d.pat_bytes = h.n_vars * fbn::kTileCases;

    p->n_sms = prop.multiProcessorCount;
    cudaError_t ce;
    int64_t tile_bytes = (int64_t)h.tile_rows * fbn::kRowDoubles * 8;
    int64_t resident;
    // Small trees (flat plans): when the LIVE message rows of a tile fit into one SM's shared memory the tile never touches HBM
    // between its evidence and its posteriors (resident.h). Measured on B200 (ALARM): see DESIGN.md. FBN_NO_RESIDENT=1 keeps
    // the one-warp-per-tile kernel with the tile state in HBM / L2.
    if (!level && h.flat && fbn::kW == 2 && !getenv("FBN_NO_RESIDENT")) {   // (the resident kernel is written for two cases per lane)
        int warps = 24;   // measured on B200 (ALARM, M cases/s): 8 warps 199, 16: 258, 24: 268, 32: 265
        if (const char *e = getenv("FBN_RES_WARPS")) warps = std::max(1, std::min(32, atoi(e)));
        fbn::build_resident(h, warps, (int64_t)prop.sharedMemPerBlockOptin, p->rp);
        if (p->rp.ok) {
            const unsigned char *blob = nullptr;
            std::vector<unsigned char> bytes((const unsigned char *)p->rp.blob.data(), (const unsigned char *)p->rp.blob.data() + p->rp.blob.size() * 4);
            if ((rc = upload(p, bytes, &blob)) != FBN_OK) { fbn_plan_destroy(p); return rc; }
            if ((rc = upload(p, p->rp.block_off, &p->rd.block_off)) != FBN_OK) { fbn_plan_destroy(p); return rc; }
            if ((rc = upload(p, p->rp.lams, &p->rd.lams)) != FBN_OK) { fbn_plan_destroy(p); return rc; }
            p->rd.n_lams = (int32_t)p->rp.lams.size();
            p->rd.clk = nullptr;
            if (getenv("FBN_RES_CLOCK")) {   // developer knob: cycles per block of CTA 0, printed when the plan is destroyed
                std::vector<long long> z(8300 + 32 * 32, 0);
                const long long *dz = nullptr;
                if ((rc = upload(p, z, &dz)) != FBN_OK) { fbn_plan_destroy(p); return rc; }
                p->rd.clk = const_cast<long long *>(dz);
            }
            p->rd.blob = blob; p->rd.vars = d.vars;
            p->rd.n_blocks = p->rp.n_blocks; p->rd.n_vars = h.n_vars; p->rd.sum_dims = h.sum_dims; p->rd.query_var = h.query_var;
            p->rd.L = fbn::res_smem_layout(p->rp.state_rows, h.n_vars, p->rp.max_block_bytes, p->rp.n_blocks, (int)p->rp.lams.size());
            ce = fbn::resident_prepare(warps, (int)prop.sharedMemPerBlockOptin);
            if (ce != cudaSuccess) {
                fbn_plan_destroy(p);
                return fail(FBN_ERR_CUDA, std::string("kernel setup failed (is this an sm_100 device?): ") + cudaGetErrorString(ce));
            }
            p->resident = true;
        }
    }
    if (p->resident) {
        tile_bytes = 512;   // no tile state in HBM
        p->blocks_per_sm = 1;
        resident = p->n_sms;
    } else if (level) {
        int64_t want = kLevelResident;
        if (const char *e = getenv("FBN_LEVEL_RESIDENT")) want = std::max(1, atoi(e));
        resident = std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(want, kLevelMaxResident), budget / tile_bytes));
        p->launches_per_wave = 2;
        for (const fbn::Phase &ph : h.phases) p->launches_per_wave += 1 + (ph.fin_end > ph.fin_begin ? 1 : 0);
    } else {
        if ((d.flat ? fbn::flat_smem_bytes(d) : fbn::fused_smem_bytes(d)) > (size_t)prop.sharedMemPerBlockOptin) {
            fbn_plan_destroy(p);
            return fail(FBN_ERR_INVALID, "tree needs more shared memory per tile than the device has");
        }
        ce = d.flat ? fbn::flat_prepare(d, (int)prop.sharedMemPerBlockOptin, &p->blocks_per_sm)
                    : fbn::fused_prepare(d, (int)prop.sharedMemPerBlockOptin, &p->blocks_per_sm);
        if (ce != cudaSuccess) {
            fbn_plan_destroy(p);
            return fail(FBN_ERR_CUDA, std::string("kernel setup failed (is this an sm_100 device?): ") + cudaGetErrorString(ce));
        }
        if (p->blocks_per_sm < 1) p->blocks_per_sm = 1;
        resident = (int64_t)p->n_sms * p->blocks_per_sm;
        if (resident * tile_bytes > budget) resident = std::max<int64_t>(1, budget / tile_bytes);
    }
    if (opt && opt->max_resident > 0) resident = std::min<int64_t>(resident, opt->max_resident);
    if (opt && opt->cases_hint > 0) resident = std::min<int64_t>(resident, ((int64_t)opt->cases_hint + fbn::kTileCases - 1) / fbn::kTileCases);
    if (!level && resident >= p->n_sms) resident = resident / p->n_sms * p->n_sms;  // whole waves over the SMs
    p->grid = (int)resident;
    p->scratch_bytes = resident * tile_bytes;
    ce = cudaMalloc((void **)&p->scratch, (size_t)p->scratch_bytes);
    if (ce != cudaSuccess) {
        fbn_plan_destroy(p);
        return fail(FBN_ERR_NOMEM, "cannot allocate tile scratch");
    }
    // from here on a failing CUDA call must release the half-built plan
#define CKP(call)                                                                                      \
    do {                                                                                               \
        cudaError_t e__ = (call);                                                                      \
        if (e__ != cudaSuccess) {                                                                      \
            fbn_plan_destroy(p);                                                                       \
            return fail(FBN_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__));            \
        }                                                                                              \
    } while (0)
    CKP(cudaMemset(p->scratch, 0, (size_t)p->scratch_bytes));
    if (level) {
        CKP(cudaMalloc((void **)&p->patG, (size_t)resident * d.pat_bytes));
        p->plan_bytes += resident * (int64_t)d.pat_bytes;
    }
    CKP(cudaStreamCreateWithFlags(&p->s_in, cudaStreamNonBlocking));
    CKP(cudaStreamCreateWithFlags(&p->s_k, cudaStreamNonBlocking));
    CKP(cudaStreamCreateWithFlags(&p->s_out, cudaStreamNonBlocking));
    CKP(cudaEventCreate(&p->ev0));
    CKP(cudaEventCreate(&p->ev1));
    CKP(cudaMalloc((void **)&p->score_out4, 4 * sizeof(double)));
#undef CKP
    *out = p;
    return FBN_OK;
}

static void free_chunk(ChunkBuf &b) {
    cudaFree(b.d_off); cudaFree(b.d_vars); cudaFree(b.d_vals); cudaFree(b.d_labels); cudaFree(b.d_post);
    cudaFree(b.d_gold); cudaFree(b.d_truth); cudaFreeHost(b.h_gold); cudaFreeHost(b.h_truth);
    cudaFreeHost(b.h_off); cudaFreeHost(b.h_vars); cudaFreeHost(b.h_vals); cudaFreeHost(b.h_labels); cudaFreeHost(b.h_post);
    if (b.in_done) cudaEventDestroy(b.in_done);
    if (b.k_done) cudaEventDestroy(b.k_done);
    if (b.out_done) cudaEventDestroy(b.out_done);
    b = ChunkBuf();
}

extern "C" void fbn_plan_destroy(fbn_plan *p) {
    if (!p) return;
    DeviceGuard guard;
    cudaSetDevice(p->device);
    cudaDeviceSynchronize();
    if (p->resident && p->rd.clk) {
        std::vector<long long> c(4 * 2048);
        const int nb = p->rp.n_blocks;
        {   // -DFBN_RES_CLOCK_BUILD=2: mean cycles per warp and block of CTA 0
            std::vector<long long> w(1 + 32 * 32);
            if (cudaMemcpy(w.data(), p->rd.clk + 8299, w.size() * 8, cudaMemcpyDeviceToHost) == cudaSuccess && w[0] > 0) {
                fprintf(stderr, "FBN_RES_CLOCK per warp (CTA 0, mean over %lld tiles): cycles from the block's constants to the end of the warp's items\n", w[0]);
                for (int b = 0; b < nb && b < 32; ++b) {
                    fprintf(stderr, "  block %2d:", b);
                    for (int k = 0; k < p->rp.warps; ++k) fprintf(stderr, " %5.0f", (double)w[1 + (size_t)b * 32 + k] / (double)w[0]);
                    fprintf(stderr, "\n");
                }
            }
        }
        if (cudaMemcpy(c.data(), p->rd.clk, c.size() * 8, cudaMemcpyDeviceToHost) == cudaSuccess && c[4 * (size_t)nb] != 0) {
            // tiles 1.. of CTA 0 (the first one includes the start-up), thread 0: cycles for requesting the next block + waiting
            // for this one, for warp 0's items, at the CTA barrier
            const int tiles = std::min(2048 / nb - 1, 40);
            fprintf(stderr, "FBN_RES_CLOCK (CTA 0, warp 0, mean over tiles 1..%d): block: issue+wait / items / barrier cycles\n", tiles - 1);
            double tot = 0;
            for (int b = 0; b < nb; ++b) {
                double w = 0, it = 0, br = 0;
                int n = 0;
                for (int k = 1; k < tiles; ++k) {
                    const size_t g = (size_t)k * nb + b;
                    if (c[4 * g + 4] == 0) break;
                    w += (double)(c[4 * g + 1] - c[4 * g]); it += (double)(c[4 * g + 2] - c[4 * g + 1]); br += (double)(c[4 * g + 4] - c[4 * g + 2]); ++n;
                }
                if (n) { fprintf(stderr, "  block %2d: %6.0f / %6.0f / %6.0f\n", b, w / n, it / n, br / n); tot += (w + it + br) / n; }
            }
            fprintf(stderr, "  tile: %.0f cycles\n", tot);
            static const char *seg[] = {"barrier->before wait", "wait for block", "header + list", "item load + dispatch", "fold", "check setup", "record loop",
                                        "items done", "CTA barrier", "OUT: item load + dispatch", "OUT: pattern + sum of the states", "OUT: reciprocal, posteriors, argmax"};
            std::vector<long long> sc(40);
            if (cudaMemcpy(sc.data(), p->rd.clk + 8100, 40 * 8, cudaMemcpyDeviceToHost) == cudaSuccess)
                for (int k = 0; k < 12; ++k)
                    if (sc[20 + k]) fprintf(stderr, "  segment %-22s: %8.0f cycles mean x %lld\n", seg[k], (double)sc[k] / sc[20 + k], sc[20 + k]);
        }
    }
    for (void *d : p->dev_allocs) cudaFree(d);
    cudaFree(p->scratch);
    cudaFree(p->patG);
    cudaFree(p->score_case);
    cudaFree(p->score_out4);
    for (auto &b : p->buf) free_chunk(b);
    if (p->s_in) cudaStreamDestroy(p->s_in);
    if (p->s_k) cudaStreamDestroy(p->s_k);
    if (p->s_out) cudaStreamDestroy(p->s_out);
    if (p->ev0) cudaEventDestroy(p->ev0);
    if (p->ev1) cudaEventDestroy(p->ev1);
    cudaGetLastError();
    delete p;
}

extern "C" int fbn_plan_get_info(const fbn_plan *p, fbn_plan_info *o) {
    if (!p || !o) return fail(FBN_ERR_INVALID, "null argument");
    const fbn::HostPlan &h = p->hp;
    memset(o, 0, sizeof(*o));
    o->n_vars = h.n_vars; o->n_cliques = h.n_cliques; o->n_separators = h.n_seps; o->sum_dims = h.sum_dims;
    o->n_levels = h.n_levels; o->team_warps = h.team; o->resident_tiles = p->grid; o->n_phases = (int32_t)h.phases.size();
    o->mode = p->level ? FBN_MODE_LEVEL : FBN_MODE_FUSED; o->launches_per_wave = p->launches_per_wave;
    o->clique_entries = h.clique_entries; o->separator_entries = h.sep_entries; o->parent_entries = h.parent_entries;
    o->marginal_entries = h.marginal_entries; o->scratch_bytes = p->scratch_bytes; o->plan_bytes = p->plan_bytes;
    o->alg_bytes_per_case = h.alg_bytes_per_case; o->kernel_bytes_per_case = h.kernel_bytes_per_case;
    o->flat_records = p->resident ? 2 : p->dp.flat;
    if (p->resident) {
        o->team_warps = p->rp.warps;
        o->kernel_bytes_per_case = 8.0 * h.sum_dims + 4.0;   // the tile state never leaves shared memory: posteriors + label out (+ the evidence in)
    }
    o->tile_cases = fbn::kTileCases; o->view_entries = h.view_entries; o->gather_rows = h.gather_rows;
    o->kernel_rev = kKernelRev;
    return FBN_OK;
}

extern "C" int fbn_run_device(fbn_plan *p, int64_t n_cases, const int64_t *ev_offsets, const int32_t *ev_vars,
                              const int32_t *ev_vals, int32_t *labels_out, double *posteriors_out, void *stream,
                              fbn_stats *stats) {
    if (!p) return fail(FBN_ERR_INVALID, "null plan");
    if (n_cases < 0 || (n_cases > 0 && !ev_offsets)) return fail(FBN_ERR_INVALID, "bad n_cases / ev_offsets");
    std::lock_guard<std::mutex> lock(p->mu);
    DeviceGuard guard;
    CK(cudaSetDevice(p->device));
    cudaStream_t st = (cudaStream_t)stream;
    if (stats) CK(cudaEventRecord(p->ev0, st));
    int64_t nl = 0;
    CK(launch_inference(p, n_cases, ev_offsets, 0, ev_vars, ev_vals, labels_out, posteriors_out, st, &nl));
    if (stats) {
        CK(cudaEventRecord(p->ev1, st));
        CK(cudaEventSynchronize(p->ev1));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, p->ev0, p->ev1));
        memset(stats, 0, sizeof(*stats));
        stats->kernel_ms = ms;
        stats->total_ms = ms;
        stats->n_launches = nl;
    }
    return FBN_OK;
}

// staging copies between pageable caller memory and the pinned chunk buffers: large ones are split over a few host threads
// (one thread moves ~8 GB/s, the PCIe link ~55 GB/s)
static void host_copy(void *dst, const void *src, size_t bytes) {
    const size_t kMin = (size_t)8 << 20;
    if (bytes < 2 * kMin) { memcpy(dst, src, bytes); return; }
    const int parts = (int)std::min<size_t>(8, bytes / kMin);
    const size_t per = (bytes / parts + 63) & ~(size_t)63;
#pragma omp parallel for num_threads(parts) schedule(static, 1)
    for (int i = 0; i < parts; ++i) {
        const size_t o = (size_t)i * per;
        if (o < bytes) memcpy((char *)dst + o, (const char *)src + o, std::min(per, bytes - o));
    }
}

static bool is_pinned(const void *ptr) {
    if (!ptr) return true;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, ptr) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

static int ensure_chunk(ChunkBuf &b, int64_t cases, int64_t ev, int sum_dims, bool want_post, bool want_post_host, bool stage_in,
                        bool stage_out, bool want_gold, bool stage_gold) {
    if (!b.in_done) {
        CK(cudaEventCreateWithFlags(&b.in_done, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&b.k_done, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&b.out_done, cudaEventDisableTiming));
    }
    if (cases > b.cap_cases) {
        cudaFree(b.d_off); cudaFree(b.d_labels); cudaFree(b.d_post); cudaFree(b.d_gold); cudaFree(b.d_truth);
        cudaFreeHost(b.h_off); cudaFreeHost(b.h_labels); cudaFreeHost(b.h_post); cudaFreeHost(b.h_gold); cudaFreeHost(b.h_truth);
        b.h_off = nullptr; b.h_labels = nullptr; b.h_post = nullptr; b.d_post = nullptr;
        b.d_gold = nullptr; b.h_gold = nullptr; b.d_truth = nullptr; b.h_truth = nullptr;
        b.d_off = nullptr; b.d_labels = nullptr;
        b.cap_cases = 0;   // a failing allocation below must not leave dangling pointers behind a stale capacity
        CK(cudaMalloc((void **)&b.d_off, (size_t)(cases + 1) * 8));
        CK(cudaMalloc((void **)&b.d_labels, (size_t)cases * 4));
        b.cap_cases = cases;
    }
    if (want_post && !b.d_post) CK(cudaMalloc((void **)&b.d_post, (size_t)b.cap_cases * sum_dims * 8));
    if (want_gold && !b.d_gold) {
        CK(cudaMalloc((void **)&b.d_gold, (size_t)b.cap_cases * sum_dims * 8));
        CK(cudaMalloc((void **)&b.d_truth, (size_t)b.cap_cases * 4));
    }
    if (want_gold && stage_gold && !b.h_gold) {
        CK(cudaMallocHost((void **)&b.h_gold, (size_t)b.cap_cases * sum_dims * 8));
        CK(cudaMallocHost((void **)&b.h_truth, (size_t)b.cap_cases * 4));
    }
    if (ev > b.cap_ev) {
        cudaFree(b.d_vars); cudaFree(b.d_vals); cudaFreeHost(b.h_vars); cudaFreeHost(b.h_vals);
        b.h_vars = nullptr; b.h_vals = nullptr; b.d_vars = nullptr; b.d_vals = nullptr;
        b.cap_ev = 0;
        CK(cudaMalloc((void **)&b.d_vars, (size_t)ev * 4));
        CK(cudaMalloc((void **)&b.d_vals, (size_t)ev * 4));
        b.cap_ev = ev;
    }
    if (stage_in) {
        if (!b.h_off) CK(cudaMallocHost((void **)&b.h_off, (size_t)(b.cap_cases + 1) * 8));
        if (!b.h_vars) { CK(cudaMallocHost((void **)&b.h_vars, (size_t)b.cap_ev * 4)); CK(cudaMallocHost((void **)&b.h_vals, (size_t)b.cap_ev * 4)); }
    }
    if (stage_out) {
        if (!b.h_labels) CK(cudaMallocHost((void **)&b.h_labels, (size_t)b.cap_cases * 4));
        if (want_post_host && !b.h_post) CK(cudaMallocHost((void **)&b.h_post, (size_t)b.cap_cases * sum_dims * 8));
    }
    return FBN_OK;
}

static int ensure_score(fbn_plan *p, int64_t n_cases) {
    if (n_cases > p->score_cap) {
        cudaFree(p->score_case);
        p->score_case = nullptr;
        p->score_cap = 0;
        CK(cudaMalloc((void **)&p->score_case, (size_t)n_cases * 24));
        p->score_cap = n_cases;
    }
    return FBN_OK;
}

// Chunked, triple-buffered pipeline H2D -> fused kernel (-> scoring kernel) -> D2H shared by fbn_run and fbn_evaluate.
static int run_pipeline(fbn_plan *p, int64_t n_cases, const int64_t *ev_offsets, const int32_t *ev_vars, const int32_t *ev_vals,
                        const double *golden, const int32_t *truths, int32_t *labels_out, double *posteriors_out, double *out4,
                        fbn_stats *stats);

// A plan owns ONE tile scratch and one set of chunk buffers: calls on the same plan are serialised here. When the pipeline
// fails half-way, earlier chunks may still have copies in flight that target the caller's buffers and the pinned staging
// areas, so the streams are drained before the error is returned.
static int run_impl(fbn_plan *p, int64_t n_cases, const int64_t *ev_offsets, const int32_t *ev_vars, const int32_t *ev_vals,
                    const double *golden, const int32_t *truths, int32_t *labels_out, double *posteriors_out, double *out4,
                    fbn_stats *stats) {
    if (!p) return fail(FBN_ERR_INVALID, "null plan");
    std::lock_guard<std::mutex> lock(p->mu);
    DeviceGuard guard;
    const int rc = run_pipeline(p, n_cases, ev_offsets, ev_vars, ev_vals, golden, truths, labels_out, posteriors_out, out4, stats);
    if (rc != FBN_OK) {
        const std::string msg = g_err;
        cudaSetDevice(p->device);
        cudaStreamSynchronize(p->s_in);
        cudaStreamSynchronize(p->s_k);
        cudaStreamSynchronize(p->s_out);
        cudaGetLastError();
        g_err = msg;
    }
    return rc;
}

static int run_pipeline(fbn_plan *p, int64_t n_cases, const int64_t *ev_offsets, const int32_t *ev_vars, const int32_t *ev_vals,
                        const double *golden, const int32_t *truths, int32_t *labels_out, double *posteriors_out, double *out4,
                        fbn_stats *stats) {
    if (n_cases < 0 || (n_cases > 0 && (!ev_offsets || !labels_out))) return fail(FBN_ERR_INVALID, "bad arguments");
    const bool want_gold = golden != nullptr;
    if (want_gold && (!truths || !out4)) return fail(FBN_ERR_INVALID, "fbn_evaluate needs truths and out4");
    auto t0 = std::chrono::steady_clock::now();
    CK(cudaSetDevice(p->device));
    if (stats) memset(stats, 0, sizeof(*stats));
    if (out4) { out4[0] = out4[1] = out4[2] = out4[3] = 0.0; }
    if (n_cases == 0) return FBN_OK;
    if (ev_offsets[n_cases] < ev_offsets[0]) return fail(FBN_ERR_INVALID, "ev_offsets must be non-decreasing");
    if (ev_offsets[n_cases] > ev_offsets[0] && (!ev_vars || !ev_vals)) return fail(FBN_ERR_INVALID, "null evidence arrays");
    // the per-case monotonicity check runs chunk by chunk inside the pipeline (a 10^6-entry scan up front would delay the
    // first copy by ~1 ms); chunk c is checked before anything of it is enqueued
    auto offsets_ok = [&](int64_t c0, int64_t c1) {
        for (int64_t i = c0; i < c1; ++i)
            if (ev_offsets[i + 1] < ev_offsets[i]) return false;
        return true;
    };
    const int sd = p->hp.sum_dims;
    const bool want_post_host = posteriors_out != nullptr;
    const bool want_post = want_post_host || want_gold;  // scoring reads the posteriors on the device
    int rc;
    if (want_gold && (rc = ensure_score(p, n_cases)) != FBN_OK) return rc;
    const bool stage_gold = want_gold && !(is_pinned(golden) && is_pinned(truths));
    // chunking: ~96 MB of posteriors per chunk, at least half a wave of resident tiles
    double chunk_mb = 96.0;  // measured on B200 (tools/e2e_sweep.py): 96 MB chunks give the shortest pipeline for 1M ALARM cases
    if (const char *e = getenv("FBN_CHUNK_MB")) chunk_mb = std::max(1.0, atof(e));
    int64_t chunk = (int64_t)(chunk_mb * 1048576.0 / (8.0 * sd));
    if (!getenv("FBN_CHUNK_MB")) chunk = std::max<int64_t>(chunk, (int64_t)p->grid * fbn::kTileCases / 2);  // at least half a wave of tiles
    chunk = std::min<int64_t>(chunk, std::max<int64_t>(fbn::kTileCases, (int64_t)(1024.0 * 1048576.0 / (8.0 * sd))));
    {   // whole waves of resident tiles per chunk when a chunk holds more than one (a ragged last wave runs nearly empty)
        const int64_t wave = (int64_t)p->grid * fbn::kTileCases;
        if (chunk > wave) chunk = chunk / wave * wave;
    }
    chunk = (chunk + fbn::kTileCases - 1) / fbn::kTileCases * fbn::kTileCases;
    chunk = std::min<int64_t>(chunk, (n_cases + fbn::kTileCases - 1) / fbn::kTileCases * fbn::kTileCases);
    std::vector<int64_t> bounds{0};  // chunk boundaries (a ramp of smaller first chunks was measured: no gain)
    while (bounds.back() < n_cases) bounds.push_back(std::min(n_cases, bounds.back() + chunk));
    const int64_t n_chunks = (int64_t)bounds.size() - 1;
    int64_t max_ev = 1;
    if (!offsets_ok(0, bounds[1])) return fail(FBN_ERR_INVALID, "ev_offsets must be non-decreasing");
    for (int64_t c = 0; c < n_chunks; ++c) max_ev = std::max(max_ev, std::max<int64_t>(0, ev_offsets[bounds[c + 1]] - ev_offsets[bounds[c]]));
    // later chunks are validated only when their turn comes: a corrupt offset there must not size the buffers
    max_ev = std::min(max_ev, std::max<int64_t>(1, ev_offsets[n_cases] - ev_offsets[0]));
    const bool stage_in = !(is_pinned(ev_offsets) && is_pinned(ev_vars) && is_pinned(ev_vals));
    const bool stage_out = !(is_pinned(labels_out) && is_pinned(posteriors_out));
    struct Pending { int64_t c0 = 0, n = 0; bool active = false; } pend[kNBuf];
    auto drain = [&](int bi) -> int {  // wait for the chunk that last used buffer bi and hand its results to the caller
        if (!pend[bi].active) return FBN_OK;
        ChunkBuf &b = p->buf[bi];
        CK(cudaEventSynchronize(b.out_done));
        if (stage_out) {
            memcpy(labels_out + pend[bi].c0, b.h_labels, (size_t)pend[bi].n * 4);
            if (want_post_host) host_copy(posteriors_out + pend[bi].c0 * sd, b.h_post, (size_t)pend[bi].n * sd * 8);
        }
        pend[bi].active = false;
        return FBN_OK;
    };
    CK(cudaEventRecord(p->ev0, p->s_k));
    for (int64_t c = 0; c < n_chunks; ++c) {
        const int bi = (int)(c % kNBuf);
        ChunkBuf &b = p->buf[bi];
        if ((rc = drain(bi)) != FBN_OK) return rc;
        const int64_t c0 = bounds[c], n = bounds[c + 1] - c0;
        if (c > 0 && !offsets_ok(c0, c0 + n)) {  // earlier chunks are in flight: let them finish, then report
            for (int b2 = 0; b2 < kNBuf; ++b2) drain(b2);
            cudaStreamSynchronize(p->s_k);
            return fail(FBN_ERR_INVALID, "ev_offsets must be non-decreasing");
        }
        const int64_t e0 = ev_offsets[c0], ne = ev_offsets[c0 + n] - e0;
        if ((rc = ensure_chunk(b, chunk, max_ev, sd, want_post, want_post_host, stage_in, stage_out, want_gold, stage_gold)) != FBN_OK) return rc;
        const int64_t *src_off = ev_offsets + c0;
        const int32_t *src_vars = ev_vars + e0, *src_vals = ev_vals + e0;
        if (stage_in) {
            memcpy(b.h_off, src_off, (size_t)(n + 1) * 8);
            if (ne) { memcpy(b.h_vars, src_vars, (size_t)ne * 4); memcpy(b.h_vals, src_vals, (size_t)ne * 4); }
            src_off = b.h_off; src_vars = b.h_vars; src_vals = b.h_vals;
        }
        CK(cudaMemcpyAsync(b.d_off, src_off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, p->s_in));
        if (ne) {
            CK(cudaMemcpyAsync(b.d_vars, src_vars, (size_t)ne * 4, cudaMemcpyHostToDevice, p->s_in));
            CK(cudaMemcpyAsync(b.d_vals, src_vals, (size_t)ne * 4, cudaMemcpyHostToDevice, p->s_in));
        }
        if (want_gold) {
            const double *src_g = golden + c0 * sd;
            const int32_t *src_t = truths + c0;
            if (stage_gold) {
                host_copy(b.h_gold, src_g, (size_t)n * sd * 8);
                memcpy(b.h_truth, src_t, (size_t)n * 4);
                src_g = b.h_gold; src_t = b.h_truth;
            }
            CK(cudaMemcpyAsync(b.d_gold, src_g, (size_t)n * sd * 8, cudaMemcpyHostToDevice, p->s_in));
            CK(cudaMemcpyAsync(b.d_truth, src_t, (size_t)n * 4, cudaMemcpyHostToDevice, p->s_in));
        }
        CK(cudaEventRecord(b.in_done, p->s_in));
        CK(cudaStreamWaitEvent(p->s_k, b.in_done, 0));
        int64_t nl = 0;
        CK(launch_inference(p, n, b.d_off, e0, b.d_vars, b.d_vals, b.d_labels, want_post ? b.d_post : nullptr, p->s_k, &nl));
        if (want_gold)
            CK(fbn::score_cases_launch(n, p->hp.n_vars, sd, p->dp.vars, b.d_post, b.d_gold, b.d_labels, b.d_truth,
                                       p->score_case + 3 * c0, p->s_k));
        CK(cudaEventRecord(b.k_done, p->s_k));
        CK(cudaStreamWaitEvent(p->s_out, b.k_done, 0));
        CK(cudaMemcpyAsync(stage_out ? b.h_labels : labels_out + c0, b.d_labels, (size_t)n * 4, cudaMemcpyDeviceToHost, p->s_out));
        if (want_post_host)
            CK(cudaMemcpyAsync(stage_out ? b.h_post : posteriors_out + c0 * sd, b.d_post, (size_t)n * sd * 8, cudaMemcpyDeviceToHost, p->s_out));
        CK(cudaEventRecord(b.out_done, p->s_out));
        pend[bi].c0 = c0; pend[bi].n = n; pend[bi].active = true;
        if (stats) {
            stats->n_launches += nl + (want_gold ? 1 : 0);
            stats->h2d_bytes += (n + 1) * 8 + ne * 8 + (want_gold ? n * ((int64_t)sd * 8 + 4) : 0);
            stats->d2h_bytes += n * 4 + (want_post_host ? n * (int64_t)sd * 8 : 0);
        }
    }
    if (want_gold) {
        CK(fbn::score_reduce_launch(n_cases, p->score_case, p->score_out4, p->s_k));
        CK(cudaMemcpyAsync(out4, p->score_out4, 32, cudaMemcpyDeviceToHost, p->s_k));
        if (stats) { stats->n_launches += 1; stats->d2h_bytes += 32; }
    }
    CK(cudaEventRecord(p->ev1, p->s_k));
    for (int bi = 0; bi < kNBuf; ++bi)
        if ((rc = drain(bi)) != FBN_OK) return rc;
    CK(cudaStreamSynchronize(p->s_k));
    CK(cudaStreamSynchronize(p->s_in));
    if (stats) {
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, p->ev0, p->ev1));
        stats->kernel_ms = ms;  // span of the compute stream (kernels + waits on their inputs)
        stats->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return FBN_OK;
}

extern "C" int fbn_run(fbn_plan *p, int64_t n_cases, const int64_t *ev_offsets, const int32_t *ev_vars,
                       const int32_t *ev_vals, int32_t *labels_out, double *posteriors_out, fbn_stats *stats) {
    return run_impl(p, n_cases, ev_offsets, ev_vars, ev_vals, nullptr, nullptr, labels_out, posteriors_out, nullptr, stats);
}

extern "C" int fbn_evaluate(fbn_plan *p, int64_t n_cases, const int64_t *ev_offsets, const int32_t *ev_vars,
                            const int32_t *ev_vals, const double *golden, const int32_t *truths, int32_t *labels_out,
                            double *posteriors_out, double *out4, fbn_stats *stats) {
    if (!golden || !truths || !out4) return fail(FBN_ERR_INVALID, "fbn_evaluate: golden, truths and out4 are required");
    return run_impl(p, n_cases, ev_offsets, ev_vars, ev_vals, golden, truths, labels_out, posteriors_out, out4, stats);
}

extern "C" int fbn_score_device(fbn_plan *p, int64_t n_cases, const double *posteriors, const int32_t *labels,
                                const double *golden, const int32_t *truths, double *out4_host, void *stream) {
    if (!p || !out4_host || n_cases < 0) return fail(FBN_ERR_INVALID, "bad arguments");
    if (n_cases > 0 && (!posteriors || !labels || !golden || !truths)) return fail(FBN_ERR_INVALID, "null device array");
    std::lock_guard<std::mutex> lock(p->mu);
    DeviceGuard guard;
    CK(cudaSetDevice(p->device));
    cudaStream_t st = (cudaStream_t)stream;
    int rc;
    if ((rc = ensure_score(p, n_cases)) != FBN_OK) return rc;
    CK(fbn::score_cases_launch(n_cases, p->hp.n_vars, p->hp.sum_dims, p->dp.vars, posteriors, golden, labels, truths,
                               p->score_case, st));
    CK(fbn::score_reduce_launch(n_cases, p->score_case, p->score_out4, st));
    CK(cudaMemcpyAsync(out4_host, p->score_out4, 32, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return FBN_OK;
}
