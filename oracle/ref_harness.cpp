// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// Thin extern "C" shim around the UNMODIFIED reference (Tonanguyxiro/FastBN) compiled from its own
// sources where they lie under /root/reference (see oracle/build_ref.sh). Output: oracle/_ref/libfastbn_ref.so.
// It is used (a) to pin the C restatement in oracle/jt_oracle.c, (b) to generate the golden vectors committed
// under tests/golden/, (c) as the "reference" CPU baseline in bench.py. Nothing in fastbn_b200/ links or loads it.
//
// The per-case sequence below is the one the reference runs in JunctionTree::PredictUseJTInfer
// (src/JunctionTree.cpp:1473-1502); protected members are reached with the include trick of SURVEY.md App. C.
#include <cstdint>
#include <climits>
#include <cstdio>
#include <cstring>
#include <unistd.h>
#include <fcntl.h>

#include <bits/stdc++.h>  // pull every std header in BEFORE the access trick so libstdc++ is untouched
#include <omp.h>

#define protected public
#include "JunctionTree.h"
#undef protected
#include "CustomNetwork.h"
#include "Dataset.h"

namespace {
struct RefCtx {
    Dataset *tester = nullptr;
    CustomNetwork *network = nullptr;
    JunctionTree *jt = nullptr;
    Timer *timer = nullptr;
    std::vector<int> var_dims;
    std::vector<int> var_off;
    int sum_dims = 0;
};

struct Quiet {  // silence the reference's cout/printf banners while a call runs
    int saved = -1;
    explicit Quiet(bool on) {
        if (!on) return;
        fflush(stdout);
        std::cout.flush();
        saved = dup(1);
        int nul = open("/dev/null", O_WRONLY);
        dup2(nul, 1);
        close(nul);
    }
    ~Quiet() {
        if (saved < 0) return;
        fflush(stdout);
        std::cout.flush();
        dup2(saved, 1);
        close(saved);
    }
};
bool g_quiet = true;
}  // namespace

extern "C" {

void fbnref_set_quiet(int q) { g_quiet = q != 0; }

// main.cpp:71-76 (ALGJT branch) without the Parameter "../dataset/" prefixing.
void *fbnref_create(const char *xml_path, const char *libsvm_path) {
    Quiet q(g_quiet);
    RefCtx *c = new RefCtx();
    c->tester = new Dataset();
    c->network = new CustomNetwork(true);
    c->network->GetNetFromXMLBIFFile(xml_path);
    c->tester->LoadLIBSVMDataKnownNetwork(libsvm_path, c->network->num_nodes);
    c->jt = new JunctionTree(c->network, c->tester, false);
    c->timer = new Timer();
    int n = c->network->num_nodes;
    c->var_dims.resize(n);
    c->var_off.resize(n + 1);
    c->var_off[0] = 0;
    for (int i = 0; i < n; ++i) {
        c->var_dims[i] = dynamic_cast<DiscreteNode *>(c->network->FindNodePtrByIndex(i))->GetDomainSize();
        c->var_off[i + 1] = c->var_off[i] + c->var_dims[i];
    }
    c->sum_dims = c->var_off[n];
    return c;
}

void fbnref_destroy(void *h) {
    RefCtx *c = (RefCtx *)h;
    delete c->timer;
    delete c->jt;
    delete c->network;
    delete c->tester;
    delete c;
}

int fbnref_num_cases(void *h) { return ((RefCtx *)h)->jt->num_instances; }
int fbnref_num_vars(void *h) { return ((RefCtx *)h)->network->num_nodes; }
int fbnref_sum_dims(void *h) { return ((RefCtx *)h)->sum_dims; }
int fbnref_max_level(void *h) { return ((RefCtx *)h)->jt->max_level; }

// out[0..7] = n_vars, n_cliques, n_separators, total table vars (cliques then separators),
//             total clique entries, root clique index, query var, max_level
void fbnref_tree_sizes(void *h, int64_t *out) {
    RefCtx *c = (RefCtx *)h;
    auto &cl = c->jt->tree->vector_clique_ptr_container;
    auto &sp = c->jt->tree->vector_separator_ptr_container;
    int64_t tv = 0, ce = 0;
    int root = -1;
    for (size_t i = 0; i < cl.size(); ++i) {
        tv += cl[i]->p_table.num_variables;
        ce += cl[i]->p_table.table_size;
        if (cl[i] == c->jt->jt_root) root = (int)i;
    }
    for (size_t i = 0; i < sp.size(); ++i) tv += sp[i]->p_table.num_variables;
    out[0] = c->network->num_nodes;
    out[1] = (int64_t)cl.size();
    out[2] = (int64_t)sp.size();
    out[3] = tv;
    out[4] = ce;
    out[5] = root;
    out[6] = c->jt->query_index;
    out[7] = c->jt->max_level;
}

// Walks the live reference objects after the JunctionTree ctor (i.e. after ReorganizeTableStorage),
// exactly as SURVEY.md App. C describes. Tables: cliques [0,nc) then separators [nc, nc+ns).
void fbnref_tree_fill(void *h, int32_t *var_dims, int32_t *table_nvars, int32_t *table_vars, double *clique_pots,
                      int32_t *sep_parent, int32_t *sep_child, int32_t *clique_level, int32_t *child_order) {
    RefCtx *c = (RefCtx *)h;
    auto &cl = c->jt->tree->vector_clique_ptr_container;
    auto &sp = c->jt->tree->vector_separator_ptr_container;
    for (int i = 0; i < c->network->num_nodes; ++i) var_dims[i] = c->var_dims[i];
    std::map<Clique *, int> cidx;
    std::map<Clique *, int> sidx;
    for (size_t i = 0; i < cl.size(); ++i) cidx[cl[i]] = (int)i;
    for (size_t i = 0; i < sp.size(); ++i) sidx[sp[i]] = (int)i;
    int64_t tv = 0, ce = 0;
    for (size_t i = 0; i < cl.size(); ++i) {
        PotentialTable &t = cl[i]->p_table;
        table_nvars[i] = t.num_variables;
        for (int k = 0; k < t.num_variables; ++k) table_vars[tv++] = t.vec_related_variables[k];
        for (int k = 0; k < t.table_size; ++k) clique_pots[ce++] = t.potentials[k];
    }
    for (size_t i = 0; i < sp.size(); ++i) {
        PotentialTable &t = sp[i]->p_table;
        table_nvars[cl.size() + i] = t.num_variables;
        for (int k = 0; k < t.num_variables; ++k) table_vars[tv++] = t.vec_related_variables[k];
        sep_parent[i] = cidx.at(sp[i]->ptr_upstream_clique);
        sep_child[i] = cidx.at(sp[i]->ptr_downstream_cliques.at(0));
    }
    // level of each clique (nodes_by_level even levels) and rank of each separator among its parent's children
    for (size_t l = 0; l < c->jt->nodes_by_level.size(); l += 2)
        for (auto *q : c->jt->nodes_by_level[l]) clique_level[cidx.at(q)] = (int)l;
    for (size_t i = 0; i < cl.size(); ++i) {
        auto &down = cl[i]->ptr_downstream_cliques;
        for (size_t k = 0; k < down.size(); ++k) child_order[sidx.at(down[k])] = (int)k;
    }
}

// One case through the reference's own per-case sequence (src/JunctionTree.cpp:1473-1502).
static int run_evidence(RefCtx *c, const DiscreteConfig &E, double *post, int32_t *label) {
    JunctionTree *jt = c->jt;
    jt->LoadDiscreteEvidence(E, 1, c->timer);
    for (size_t i = 0; i < jt->tree->vector_clique_ptr_container.size(); ++i)
        jt->tree->vector_clique_ptr_container[i]->p_table.Normalize();
    jt->MessagePassingUpdateJT(1, c->timer);
    int lab = jt->InferenceUsingJT();
    jt->GetProbabilitiesAllNodes(E);
    for (int v = 0; v < c->network->num_nodes; ++v)
        for (int d = 0; d < c->var_dims[v]; ++d) post[c->var_off[v] + d] = jt->probs_one_sample[v][d];
    jt->ResetJunctionTree();
    *label = lab;
    return 0;
}

int fbnref_run_case(void *h, int i, double *post, int32_t *label) {
    RefCtx *c = (RefCtx *)h;
    Quiet q(g_quiet);
    return run_evidence(c, c->jt->evidences.at(i), post, label);
}

int fbnref_run_custom(void *h, int n_ev, const int32_t *vars, const int32_t *vals, double *post, int32_t *label) {
    RefCtx *c = (RefCtx *)h;
    Quiet q(g_quiet);
    DiscreteConfig E;
    for (int i = 0; i < n_ev; ++i) E.insert(std::pair<int, int>(vars[i], vals[i]));
    return run_evidence(c, E, post, label);
}

// Batch form used for timing the reference (bench.py --impl reference / cpu_baseline):
// cases given as CSR; returns seconds spent inside the per-case sequence (omp_get_wtime).
double fbnref_run_batch(void *h, int64_t n_cases, const int64_t *ev_off, const int32_t *vars, const int32_t *vals,
                        double *post /* may be null */, int32_t *labels /* may be null */) {
    RefCtx *c = (RefCtx *)h;
    Quiet q(g_quiet);
    std::vector<double> tmp(c->sum_dims);
    std::vector<DiscreteConfig> Es(n_cases);
    for (int64_t i = 0; i < n_cases; ++i)
        for (int64_t k = ev_off[i]; k < ev_off[i + 1]; ++k) Es[i].insert(std::pair<int, int>(vars[k], vals[k]));
    double t0 = omp_get_wtime();
    for (int64_t i = 0; i < n_cases; ++i) {
        int32_t lab;
        run_evidence(c, Es[i], post ? post + i * c->sum_dims : tmp.data(), &lab);
        if (labels) labels[i] = lab;
    }
    return omp_get_wtime() - t0;
}

// Batch form with the reference's intra-case OpenMP parallelism (`-t N`): the same per-case sequence with num_threads
// passed to LoadDiscreteEvidence / MessagePassingUpdateJT (src/JunctionTree.cpp:1476,1487). Works for any number of cases
// (the stock loop of PredictUseJTInfer divides by n/20).
double fbnref_run_batch_threads(void *h, int64_t n_cases, const int64_t *ev_off, const int32_t *vars, const int32_t *vals,
                                double *post /* may be null */, int32_t *labels /* may be null */, int num_threads) {
    RefCtx *c = (RefCtx *)h;
    Quiet q(g_quiet);
    JunctionTree *jt = c->jt;
    std::vector<DiscreteConfig> Es(n_cases);
    for (int64_t i = 0; i < n_cases; ++i)
        for (int64_t k = ev_off[i]; k < ev_off[i + 1]; ++k) Es[i].insert(std::pair<int, int>(vars[k], vals[k]));
    double t0 = omp_get_wtime();
    for (int64_t i = 0; i < n_cases; ++i) {
        jt->LoadDiscreteEvidence(Es[i], num_threads, c->timer);
        for (size_t k = 0; k < jt->tree->vector_clique_ptr_container.size(); ++k)
            jt->tree->vector_clique_ptr_container[k]->p_table.Normalize();
        jt->MessagePassingUpdateJT(num_threads, c->timer);
        int lab = jt->InferenceUsingJT();
        jt->GetProbabilitiesAllNodes(Es[i]);
        if (post)
            for (int v = 0; v < c->network->num_nodes; ++v)
                for (int d = 0; d < c->var_dims[v]; ++d) post[i * c->sum_dims + c->var_off[v] + d] = jt->probs_one_sample[v][d];
        jt->ResetJunctionTree();
        if (labels) labels[i] = lab;
    }
    double secs = omp_get_wtime() - t0;
    omp_set_num_threads(1);
    return secs;
}

// The stock driver loop itself, timed like the reference's own `jt` timer (src/JunctionTree.cpp:66-74): golden table load
// outside, then PredictUseJTInfer(num_threads, timer) incl. CalculateMSE / CalculateHellingerDistance per case and the
// progress lines, then Accuracy. Needs >= 20 loaded cases (quirk 5). Returns the accuracy, *secs = the timed span.
double fbnref_predict_timed(void *h, const char *pt_path, int num_threads, double *secs) {
    RefCtx *c = (RefCtx *)h;
    Quiet q(g_quiet);
    c->jt->LoadGroundTruthProbabilityTable(pt_path);
    Timer *timer = new Timer();
    double t0 = omp_get_wtime();
    std::vector<int> predictions = c->jt->PredictUseJTInfer(num_threads, timer);
    double acc = c->jt->Accuracy(predictions);
    *secs = omp_get_wtime() - t0;
    delete timer;
    omp_set_num_threads(1);
    return acc;
}

// The reference's own public entry point (include/Inference.h:44), banners and timers included.
double fbnref_evaluate_accuracy(void *h, const char *pt_path, int num_threads, int quiet) {
    RefCtx *c = (RefCtx *)h;
    Quiet q(quiet != 0);
    return c->jt->EvaluateAccuracy(pt_path, num_threads);
}

// evidence of the i-th loaded test case (for fixtures): returns count, fills up to cap pairs
int fbnref_case_evidence(void *h, int i, int cap, int32_t *vars, int32_t *vals, int32_t *truth) {
    RefCtx *c = (RefCtx *)h;
    const DiscreteConfig &E = c->jt->evidences.at(i);
    int k = 0;
    for (auto &e : E) {
        if (k < cap) {
            vars[k] = e.first;
            vals[k] = e.second;
        }
        ++k;
    }
    *truth = c->jt->ground_truths.at(i);
    return k;
}

}  // extern "C"
