// extern "C" access to the host mirror for the Python harness (tests, bench.py): build the junction tree of an XMLBIF
// network natively and hand it out in the flat fbn_tree_desc layout; load LIBSVM test cases; run EvaluateAccuracy.
// Declared in include/fastbn_b200_host.h.
#include <cstring>
#include <iostream>
#include <sstream>

#include "../../include/fastbn_b200_host.h"
#include "fastbn_host.h"

struct fbn_host_model {
    CustomNetwork *net = nullptr;
    Dataset *data = nullptr;
    JunctionTree *jt = nullptr;
    JunctionTree::FlatTree flat;
};

namespace {
// the mirror prints the reference's progress lines to stdout; the C entry points keep them out of the caller's stdout
struct CoutMute {
    std::streambuf *old;
    std::ostringstream sink;
    explicit CoutMute(bool on) : old(nullptr) { if (on) old = std::cout.rdbuf(sink.rdbuf()); }
    ~CoutMute() { if (old) std::cout.rdbuf(old); }
};
}  // namespace

extern "C" int fbn_host_model_create(const char *xmlbif_path, const char *libsvm_path, int quiet, fbn_host_model **out) {
    if (!xmlbif_path || !out) return FBN_ERR_INVALID;
    CoutMute mute(quiet != 0);
    fbn_host_model *m = new fbn_host_model();
    m->net = new CustomNetwork(true);
    m->net->GetNetFromXMLBIFFile(xmlbif_path);
    m->data = new Dataset();
    if (libsvm_path && *libsvm_path) m->data->LoadLIBSVMDataKnownNetwork(libsvm_path, m->net->num_nodes);
    else m->data->ev_offsets.assign(1, 0);
    m->jt = new JunctionTree(m->net, m->data, false);
    m->jt->quiet = quiet != 0;
    m->jt->Flatten(m->flat);
    *out = m;
    return FBN_OK;
}

extern "C" void fbn_host_model_destroy(fbn_host_model *m) {
    if (!m) return;
    delete m->jt;
    delete m->net;
    delete m->data;
    delete m;
}

extern "C" const fbn_tree_desc *fbn_host_model_tree(const fbn_host_model *m) { return m ? &m->flat.desc : nullptr; }

extern "C" int fbn_host_model_info(const fbn_host_model *m, int64_t *out8) {
    if (!m || !out8) return FBN_ERR_INVALID;
    int64_t tv = 0, ce = 0;
    for (int32_t x : m->flat.table_nvars) tv += x;
    ce = (int64_t)m->flat.clique_potentials.size();
    int max_clique = 0;
    for (Clique *c : m->jt->tree->vector_clique_ptr_container) max_clique = std::max(max_clique, c->p_table.table_size);
    out8[0] = m->net->num_nodes; out8[1] = m->flat.desc.n_cliques; out8[2] = m->flat.desc.n_separators; out8[3] = tv;
    out8[4] = ce; out8[5] = m->jt->max_level; out8[6] = m->jt->num_instances; out8[7] = max_clique;
    return FBN_OK;
}

extern "C" int fbn_host_model_cases(const fbn_host_model *m, const int64_t **ev_offsets, const int32_t **ev_vars,
                                    const int32_t **ev_vals, const int32_t **truths) {
    if (!m) return FBN_ERR_INVALID;
    if (ev_offsets) *ev_offsets = m->jt->ev_offsets.data();
    if (ev_vars) *ev_vars = m->jt->ev_vars.data();
    if (ev_vals) *ev_vals = m->jt->ev_vals.data();
    if (truths) *truths = m->jt->ground_truths.data();
    return FBN_OK;
}

extern "C" int fbn_host_model_evaluate(fbn_host_model *m, const char *pt_path, const int32_t *devices, int n_devices,
                                       int32_t *labels_out, double *posteriors_out, double *out5) {
    if (!m || !pt_path || !out5) return FBN_ERR_INVALID;
    CoutMute mute(m->jt->quiet);
    if (devices && n_devices > 0) m->jt->devices.assign(devices, devices + n_devices);
    m->jt->keep_posteriors = posteriors_out != nullptr;
    double acc = m->jt->EvaluateAccuracy(pt_path, 1);
    out5[0] = acc; out5[1] = m->jt->average_mse; out5[2] = m->jt->average_hd; out5[3] = m->jt->jt_seconds; out5[4] = m->jt->kernel_ms;
    if (labels_out) memcpy(labels_out, m->jt->predictions.data(), sizeof(int32_t) * m->jt->predictions.size());
    if (posteriors_out) memcpy(posteriors_out, m->jt->posteriors.data(), sizeof(double) * m->jt->posteriors.size());
    return FBN_OK;
}

extern "C" int fbn_host_model_load_golden(fbn_host_model *m, const char *pt_path, double *golden_out) {
    if (!m || !pt_path || !golden_out) return FBN_ERR_INVALID;
    CoutMute mute(m->jt->quiet);
    m->jt->LoadGroundTruthProbabilityTable(pt_path);
    memcpy(golden_out, m->jt->ground_truth_probability_tables.data(), sizeof(double) * m->jt->ground_truth_probability_tables.size());
    return FBN_OK;
}

extern "C" int fbn_host_model_write_pt(fbn_host_model *m, const char *path) {
    if (!m || !path) return FBN_ERR_INVALID;
    m->jt->WritePotentialTableFile(path);
    return FBN_OK;
}
