// Inference + JunctionTree: the reference's driver classes for `-a 2` (include/Inference.h, include/JunctionTree.h),
// with the per-test-case loop of src/JunctionTree.cpp:1508-1534 replaced by ONE batched call into the CUDA library
// (fbn_evaluate, include/fastbn_b200.h) per device. There is no host implementation of message passing here: without a
// CUDA device EvaluateAccuracy prints the library's error and exits(1), the reference's error convention.
#include <algorithm>
#include <array>
#include <chrono>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <memory>
#include <thread>
#include <unistd.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#include "fastbn_host.h"

namespace {
double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
}  // namespace

// ---- Inference -------------------------------------------------------------------------------------------------
// Copies evidences and ground truths out of the Dataset (src/Inference.cpp:13-42): everything except the class variable
// is evidence. is_dense (fill missing variables with value 0, :66-86) is honoured.
Inference::Inference(Network *net, Dataset *dts, bool is_dense)
    : network(net), num_instances(dts->num_instance), query_index(dts->class_var_index) {
    var_offsets.assign(net->num_nodes + 1, 0);
    for (int v = 0; v < net->num_nodes; ++v) var_offsets[v + 1] = var_offsets[v] + net->FindNodePtrByIndex(v)->GetDomainSize();
    ground_truths.assign(dts->labels.begin(), dts->labels.end());
    {   // a value outside the variable's domain silently wipes every table (NaN posteriors, label -1) exactly as in the
        // reference; a malformed test file should at least be visible
        int64_t bad = 0;
        for (size_t k = 0; k < dts->ev_vars.size(); ++k) {
            const int v = dts->ev_vars[k];
            if (v >= 0 && v < net->num_nodes && (dts->ev_vals[k] < 0 || dts->ev_vals[k] >= var_offsets[v + 1] - var_offsets[v])) ++bad;
        }
        if (bad) fprintf(stderr, "Warning: %lld evidence value(s) outside the variable's domain: those cases have no consistent state (nan posteriors)\n", (long long)bad);
    }
    if (!is_dense) {
        // DiscreteConfig is a std::set of (variable, value): ascending and without duplicates (src/Inference.cpp:21-33)
        ev_offsets.assign(1, 0);
        ev_vars.reserve(dts->ev_vars.size());
        ev_vals.reserve(dts->ev_vals.size());
        vector<pair<int, int>> tmp;
        for (int i = 0; i < num_instances; ++i) {
            const int64_t k0 = dts->ev_offsets[i], k1 = dts->ev_offsets[i + 1];
            bool sorted = true;
            for (int64_t k = k0 + 1; k < k1 && sorted; ++k)
                sorted = dts->ev_vars[k - 1] < dts->ev_vars[k] ||
                         (dts->ev_vars[k - 1] == dts->ev_vars[k] && dts->ev_vals[k - 1] < dts->ev_vals[k]);
            if (sorted) {
                ev_vars.insert(ev_vars.end(), dts->ev_vars.begin() + k0, dts->ev_vars.begin() + k1);
                ev_vals.insert(ev_vals.end(), dts->ev_vals.begin() + k0, dts->ev_vals.begin() + k1);
            } else {
                tmp.clear();
                for (int64_t k = k0; k < k1; ++k) tmp.emplace_back(dts->ev_vars[k], dts->ev_vals[k]);
                std::sort(tmp.begin(), tmp.end());
                tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
                for (auto &vv : tmp) { ev_vars.push_back(vv.first); ev_vals.push_back(vv.second); }
            }
            ev_offsets.push_back((int64_t)ev_vars.size());
        }
        return;
    }
    ev_offsets.assign(1, 0);
    vector<char> seen(net->num_nodes);
    for (int i = 0; i < num_instances; ++i) {
        std::fill(seen.begin(), seen.end(), 0);
        for (int64_t k = dts->ev_offsets[i]; k < dts->ev_offsets[i + 1]; ++k) {
            ev_vars.push_back(dts->ev_vars[k]);
            ev_vals.push_back(dts->ev_vals[k]);
            if (dts->ev_vars[k] >= 0 && dts->ev_vars[k] < net->num_nodes) seen[dts->ev_vars[k]] = 1;
        }
        for (int v = 0; v < net->num_nodes; ++v)
            if (v != query_index && !seen[v]) { ev_vars.push_back(v); ev_vals.push_back(0); }
        ev_offsets.push_back((int64_t)ev_vars.size());
    }
}

DiscreteConfig Inference::Evidence(int i) const {
    DiscreteConfig e;
    for (int64_t k = ev_offsets[i]; k < ev_offsets[i + 1]; ++k) e.insert(pair<int, int>(ev_vars[k], ev_vals[k]));
    return e;
}

double Inference::Accuracy(vector<int> predictions) {  // src/Inference.cpp:46-61
    int correct = 0, wrong = 0;
    for (size_t i = 0; i < ground_truths.size(); ++i) (ground_truths[i] == predictions.at(i) ? correct : wrong)++;
    return correct / (double)(correct + wrong);
}

int Inference::ArgMax(const vector<double> &array) {  // first value strictly above the running max, which starts at 0
    double max_prob = 0;
    int max_index = -1;
    for (size_t i = 0; i < array.size(); ++i)
        if (array[i] > max_prob) { max_prob = array[i]; max_index = (int)i; }
    return max_index;
}

// One text line per (case, node): dim numbers "%.7f " each, or an empty line for an evidence node, which is stored as
// first entry -1 (src/Inference.cpp:108-146). Lines missing at the end of a short file count as empty.
void Inference::LoadGroundTruthProbabilityTable(string file_path) {
    FILE *f = fopen(file_path.c_str(), "rb");
    if (!f) {
        fprintf(stderr, "Error in function %s!", __FUNCTION__);
        fprintf(stderr, "Unable to open file %s!", file_path.c_str());
        exit(1);
    }
    std::cout << "Data file opened. Begin to load ground truth probability tables computed by JT. " << std::endl;
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    // uninitialised buffer (a std::string would zero-fill ~0.9 GB first), NUL-terminated for the strtod fallback
    struct RawBuf {
        std::unique_ptr<char[]> p;
        size_t n = 0;
        size_t size() const { return n; }
        bool empty() const { return n == 0; }
        char &operator[](size_t i) { return p[i]; }
        const char *data() const { return p.get(); }
        char back() const { return p[n - 1]; }
    } buf;
    buf.n = sz > 0 ? (size_t)sz : 0;
    buf.p.reset(new char[buf.n + 1]);
    buf.p[buf.n] = 0;
    {   // the file is read and scanned in parallel pieces (887 MB for 10^6 ALARM cases)
        const int fd = fileno(f);
        int nth = 1;
#ifdef _OPENMP
        nth = omp_get_max_threads();
#endif
        const int64_t piece = ((int64_t)buf.size() + nth - 1) / std::max(nth, 1);
        bool ok = true;
#pragma omp parallel for schedule(static, 1)
        for (int t = 0; t < nth; ++t) {
            int64_t o = (int64_t)t * piece, end = std::min<int64_t>((int64_t)buf.size(), o + piece);
            while (o < end) {
                ssize_t g = pread(fd, &buf[(size_t)o], (size_t)(end - o), (off_t)o);
                if (g <= 0) { ok = false; break; }
                o += g;
            }
        }
        if (!ok) {
            fprintf(stderr, "Error in function %s!Unable to read file %s!", __FUNCTION__, file_path.c_str());
            exit(1);
        }
    }
    fclose(f);
    const int n_nodes = network->num_nodes, sd = var_offsets[n_nodes];
    ground_truth_probability_tables.assign((size_t)num_instances * sd, 0.0);
    // line starts (newlines counted per piece, then placed), then parallel numeric parsing
    const int64_t want = (int64_t)num_instances * n_nodes;
    const char *base = buf.data();
    vector<int64_t> line_at;
    {
        int nth = 1;
#ifdef _OPENMP
        nth = omp_get_max_threads();
#endif
        const int64_t piece = ((int64_t)buf.size() + nth - 1) / std::max(nth, 1);
        vector<int64_t> cnt((size_t)nth + 1, 0);
#pragma omp parallel for schedule(static, 1)
        for (int t = 0; t < nth; ++t) {
            const char *p = base + std::min<int64_t>((int64_t)buf.size(), (int64_t)t * piece);
            const char *pe = base + std::min<int64_t>((int64_t)buf.size(), (int64_t)(t + 1) * piece);
            int64_t c = 0;
            while (p < pe) {
                const char *nl = (const char *)memchr(p, '\n', (size_t)(pe - p));
                if (!nl) break;
                ++c;
                p = nl + 1;
            }
            cnt[(size_t)t + 1] = c;
        }
        for (int t = 0; t < nth; ++t) cnt[(size_t)t + 1] += cnt[(size_t)t];
        // line k+1 starts after the k-th newline; line 0 starts at 0; a last piece without '\n' is a line too
        const int64_t n_nl = cnt[(size_t)nth];
        const bool tail = !buf.empty() && buf.back() != '\n';
        const int64_t n_lines = buf.empty() ? 0 : n_nl + (tail ? 1 : 0);
        line_at.assign((size_t)std::min(want, n_lines) + 1, 0);
        const int64_t keep = (int64_t)line_at.size() - 1;  // lines used
#pragma omp parallel for schedule(static, 1)
        for (int t = 0; t < nth; ++t) {
            const char *p = base + std::min<int64_t>((int64_t)buf.size(), (int64_t)t * piece);
            const char *pe = base + std::min<int64_t>((int64_t)buf.size(), (int64_t)(t + 1) * piece);
            int64_t k = cnt[(size_t)t];  // newlines before this piece
            while (p < pe) {
                const char *nl = (const char *)memchr(p, '\n', (size_t)(pe - p));
                if (!nl) break;
                ++k;
                if (k <= keep) line_at[(size_t)k] = nl + 1 - base;  // start of line k == end of line k-1
                p = nl + 1;
            }
        }
        if (keep > n_nl) line_at[(size_t)keep] = (int64_t)buf.size();  // the last used line is the unterminated tail
    }
    const int64_t have = (int64_t)line_at.size() - 1;
#pragma omp parallel for schedule(static)
    for (int64_t ln = 0; ln < want; ++ln) {
        const int64_t i = ln / n_nodes;
        const int j = (int)(ln % n_nodes);
        double *dst = &ground_truth_probability_tables[(size_t)i * sd + var_offsets[j]];
        const int dim = var_offsets[j + 1] - var_offsets[j];
        if (ln >= have) { dst[0] = -1; continue; }
        const char *p = base + line_at[ln];
        const char *e = base + line_at[ln + 1];
        while (e > p && (unsigned char)e[-1] < 33) --e;
        if (p == e) { dst[0] = -1; continue; }
        for (int k = 0; k < dim && p < e; ++k) {
            const char *q = textio::ParseDouble(p, e, &dst[k]);
            if (q == p) break;  // not a number: the rest of the block stays 0, like a short line
            p = q;
            while (p < e && *p == ' ') ++p;
        }
    }
}

// ---- JunctionTree ----------------------------------------------------------------------------------------------
JunctionTree::JunctionTree(Network *net, Dataset *dts, bool is_dense) : Inference(net, dts, is_dense) {
    double t0 = now_s();
    tree = new JunctionTreeStructure(net);
    // root = the clique that minimises the depth of the tree, first one on ties (src/JunctionTree.cpp:14-24)
    int best = GetMaxLevel(tree->vector_clique_ptr_container[0]);
    jt_root = tree->vector_clique_ptr_container[0];
    for (size_t i = 1; i < tree->vector_clique_ptr_container.size(); ++i) {
        int lv = GetMaxLevel(tree->vector_clique_ptr_container[i]);
        if (lv < best) { best = lv; jt_root = tree->vector_clique_ptr_container[i]; }
    }
    MarkLevel();
    std::cout << "Finish MarkLevel, maximum level of the junction tree = " << max_level << std::endl;
    ReorganizeTableStorage(num_instances);
    probs_one_sample.resize(network->num_nodes);
    if (const char *env = getenv("FASTBN_DEVICES")) {
        for (const char *p = env; *p;) {
            char *e = nullptr;
            long d = strtol(p, &e, 10);
            if (e == p) break;
            devices.push_back((int)d);
            p = (*e == ',') ? e + 1 : e;
        }
    }
    if (devices.empty()) devices.push_back(0);
    // the CUDA contexts come up (about a second each) while the caller goes on to read the potential-table file
    {
        vector<int> devs = devices;
        warmup = std::thread([devs] { for (int d : devs) fbn_device_warmup(d); });
    }
    std::cout << "==================================================" << std::endl;
    std::cout << "construct jt: " << now_s() - t0 << " s " << std::endl;
}

JunctionTree::~JunctionTree() {
    if (warmup.joinable()) warmup.join();
    delete tree;
}

// number of alternating clique / separator levels when `root` is the root: 2 * eccentricity + 1
int JunctionTree::GetMaxLevel(Clique *root) {
    vector<pair<Clique *, Clique *>> frontier{{root, nullptr}}, next;
    int levels = 0;
    while (!frontier.empty()) {
        ++levels;
        next.clear();
        for (auto &cf : frontier)
            for (Clique *nb : cf.first->set_neighbours_ptr)
                if (nb != cf.second) next.push_back({nb, cf.first});
        frontier.swap(next);
    }
    return levels;
}

// level order from the root; fills nodes_by_level (cliques on even levels, separators on odd ones),
// separators_by_level, ptr_upstream_clique and ptr_downstream_cliques (src/JunctionTree.cpp:137-181)
void JunctionTree::MarkLevel() {
    nodes_by_level.clear();
    separators_by_level.clear();
    vector<Clique *> level{jt_root};
    while (!level.empty()) {
        nodes_by_level.push_back(level);
        const bool children_are_separators = nodes_by_level.size() % 2 == 1;
        vector<Clique *> below;
        vector<Separator *> below_sep;
        for (Clique *c : level)
            for (Clique *nb : c->set_neighbours_ptr) {
                if (nb == c->ptr_upstream_clique) continue;
                c->ptr_downstream_cliques.push_back(nb);
                nb->ptr_upstream_clique = c;
                below.push_back(nb);
                if (children_are_separators) below_sep.push_back(static_cast<Separator *>(nb));
            }
        if (children_are_separators && !below_sep.empty()) separators_by_level.push_back(below_sep);
        level.swap(below);
    }
    max_level = (int)nodes_by_level.size();
}

// every non-root clique stores the variables of its upstream separator LAST, in separator order, the others first in
// their previous relative order (src/JunctionTree.cpp:235-281, src/PotentialTable.cpp:215-256)
void JunctionTree::ReorganizeTableStorage(int) {
    for (Clique *c : tree->vector_clique_ptr_container) {
        if (!c->ptr_upstream_clique) continue;
        const vector<int> &sv = c->ptr_upstream_clique->p_table.vec_related_variables;
        const vector<int> &cur = c->p_table.vec_related_variables;
        vector<int> order;
        for (int v : cur)
            if (std::find(sv.begin(), sv.end(), v) == sv.end()) order.push_back(v);
        order.insert(order.end(), sv.begin(), sv.end());
        if (order != cur) c->p_table.TableReorganization(order);
    }
}

void JunctionTree::Flatten(FlatTree &o) const {
    const int nv = network->num_nodes;
    map<const Clique *, int> cidx;
    o = FlatTree();
    for (int v = 0; v < nv; ++v) o.var_dims.push_back(var_offsets[v + 1] - var_offsets[v]);
    size_t total = 0;
    for (Clique *c : tree->vector_clique_ptr_container) total += (size_t)c->p_table.table_size;
    o.clique_potentials.reserve(total);
    for (size_t i = 0; i < tree->vector_clique_ptr_container.size(); ++i) {
        const Clique *c = tree->vector_clique_ptr_container[i];
        cidx[c] = (int)i;
        o.table_nvars.push_back(c->p_table.num_variables);
        o.table_vars.insert(o.table_vars.end(), c->p_table.vec_related_variables.begin(), c->p_table.vec_related_variables.end());
        o.clique_potentials.insert(o.clique_potentials.end(), c->p_table.potentials.begin(), c->p_table.potentials.end());
    }
    for (const Separator *s : tree->vector_separator_ptr_container) {
        o.table_nvars.push_back(s->p_table.num_variables);
        o.table_vars.insert(o.table_vars.end(), s->p_table.vec_related_variables.begin(), s->p_table.vec_related_variables.end());
        o.sep_parent.push_back(cidx[s->ptr_upstream_clique]);
        o.sep_child.push_back(cidx[s->ptr_downstream_cliques.at(0)]);
    }
    fbn_tree_desc &d = o.desc;
    memset(&d, 0, sizeof(d));
    d.n_vars = nv;
    d.var_dims = o.var_dims.data();
    d.n_cliques = (int32_t)tree->vector_clique_ptr_container.size();
    d.n_separators = (int32_t)tree->vector_separator_ptr_container.size();
    d.table_nvars = o.table_nvars.data();
    d.table_vars = o.table_vars.data();
    d.clique_potentials = o.clique_potentials.data();
    d.sep_parent = o.sep_parent.data();
    d.sep_child = o.sep_child.data();
    d.root = cidx[jt_root];
    d.query_var = query_index;
}

// All test cases in one batched call per device (replaces the loop of src/JunctionTree.cpp:1517-1528). Cases are cut
// into contiguous slices, one per entry of `devices`; the only cross-device step is the sum of
// {n_correct, n_total, sum_mse, sum_hd}.
vector<int> JunctionTree::PredictUseJTInfer(int) {
    if (!quiet) std::cout << "Progress indicator: " << std::endl;
    if (warmup.joinable()) warmup.join();
    FlatTree flat;
    Flatten(flat);
    const int sd = var_offsets[network->num_nodes];
    const int n = num_instances;
    predictions.assign(n, 0);
    if (keep_posteriors) posteriors.assign((size_t)n * sd, 0.0);
    const int nd = (int)devices.size();
    // several devices: the four scores of the slices meet in ONE ncclAllReduce (fbn_reduce), one communicator per device
    vector<fbn_comm *> comms(nd, nullptr);
    if (nd > 1 && fbn_comm_create_all(nd, devices.data(), comms.data()) != FBN_OK) {
        fprintf(stderr, "Error in function %s! %s\n", __FUNCTION__, fbn_last_error());
        exit(1);
    }
    vector<std::array<double, 4>> part(nd);
    vector<string> errors(nd);
    vector<double> kms(nd, 0.0);
    vector<double> last_post(sd, 0.0);
    auto worker = [&](int di) {
        const int64_t c0 = (int64_t)n * di / nd, c1 = (int64_t)n * (di + 1) / nd;
        part[di] = {0, 0, 0, 0};
        fbn_plan_options opt;
        memset(&opt, 0, sizeof(opt));
        opt.device = devices[di];
        opt.cases_hint = (int32_t)std::min<int64_t>(c1 - c0, INT32_MAX);
        fbn_plan *plan = nullptr;
        const double tw0 = now_s();
        if (fbn_plan_create(&flat.desc, &opt, &plan) != FBN_OK) { errors[di] = fbn_last_error(); return; }
        const double tw1 = now_s();
        // slice-local CSR offsets are absolute into ev_vars/ev_vals (fbn_run rebases them per chunk)
        fbn_stats st;
        memset(&st, 0, sizeof(st));
        const bool need_last = (di == nd - 1) && c1 > c0;
        double *post_out = keep_posteriors ? posteriors.data() + (size_t)c0 * sd : nullptr;
        int rc = fbn_evaluate(plan, c1 - c0, ev_offsets.data() + c0, ev_vars.data(), ev_vals.data(),
                              ground_truth_probability_tables.data() + (size_t)c0 * sd, ground_truths.data() + c0,
                              predictions.data() + c0, post_out, part[di].data(), &st);
        if (rc != FBN_OK) errors[di] = fbn_last_error();
        kms[di] = st.kernel_ms;
        if (getenv("FASTBN_TIMING"))
            fprintf(stderr, "[timing] device %d: plan %.3f s, fbn_evaluate %.3f s (compute stream %.3f s) for %lld cases\n", devices[di],
                    tw1 - tw0, now_s() - tw1, st.kernel_ms * 1e-3, (long long)(c1 - c0));
        if (rc == FBN_OK && need_last) {  // probs_one_sample holds the last case, as after the reference's loop
            if (keep_posteriors) {
                memcpy(last_post.data(), posteriors.data() + (size_t)(n - 1) * sd, sizeof(double) * sd);
            } else {
                int32_t lab;
                rc = fbn_run(plan, 1, ev_offsets.data() + (n - 1), ev_vars.data(), ev_vals.data(), &lab, last_post.data(), nullptr);
                if (rc != FBN_OK) errors[di] = fbn_last_error();
            }
        }
        fbn_plan_destroy(plan);
    };
    auto reduce_worker = [&](int di) {
        if (fbn_reduce(comms[di], part[di].data()) != FBN_OK) errors[di] = fbn_last_error();
    };
    if (nd == 1) {
        worker(0);
    } else {
        vector<std::thread> th;
        for (int di = 0; di < nd; ++di) th.emplace_back(worker, di);
        for (auto &t : th) t.join();
    }
    auto check_errors = [&]() {
        for (int di = 0; di < nd; ++di)
            if (!errors[di].empty()) {
                fprintf(stderr, "Error in function %s! device %d: %s\n", __FUNCTION__, devices[di], errors[di].c_str());
                exit(1);
            }
    };
    check_errors();  // a failed slice must not enter the collective (the other ranks would wait for it)
    if (nd > 1) {    // every device thread contributes its slice's scores; afterwards each holds the sum over all slices
        vector<std::thread> th;
        for (int di = 0; di < nd; ++di) th.emplace_back(reduce_worker, di);
        for (auto &t : th) t.join();
        check_errors();
        for (int di = 0; di < nd; ++di) fbn_comm_destroy(comms[di]);
    }
    const double mse = part[0][2], hd = part[0][3];
    n_correct_reduced = (int64_t)(part[0][0] + 0.5);
    kernel_ms = 0;
    for (int di = 0; di < nd; ++di) kernel_ms = std::max(kernel_ms, kms[di]);
    for (int v = 0; v < network->num_nodes; ++v)
        probs_one_sample[v].assign(last_post.begin() + var_offsets[v], last_post.begin() + var_offsets[v + 1]);
    average_mse = mse / n;
    average_hd = hd / n;
    if (!quiet && n >= 20) {
        // the reference prints one line per n/20 cases while it walks them (src/JunctionTree.cpp:1509-1524); here the whole
        // batch is back at once, the same lines follow each other
        const int every = n / 20;
        for (int progress = every; progress <= n; progress += every)
            fprintf(stdout, "%s\n", (std::to_string((double)progress / n * 100) + "%...").c_str());
        fflush(stdout);
    }
    std::cout << "average MSE = " << average_mse << std::endl;
    std::cout << "average HD = " << average_hd << std::endl;
    return predictions;
}

double JunctionTree::EvaluateAccuracy(string path, int num_threads) {
    LoadGroundTruthProbabilityTable(path);
    std::cout << "==================================================" << '\n'
              << "Begin testing the trained network." << std::endl;
    double t0 = now_s();
    vector<int> pred = PredictUseJTInfer(num_threads);
    double accuracy = Accuracy(pred);
    jt_seconds = now_s() - t0;
    std::cout << "==================================================" << std::endl;
    // the reference's timer block (src/JunctionTree.cpp:76-124), same names. What the names mean here: the whole batch is ONE
    // pipelined call, evidence loading, message passing and the marginals + ArgMax + MSE/HD run inside the same kernels, so
    // "msg passing" is the device time of those kernels (max over the devices), "load evidence" the rest of the call (staging
    // and host<->device copies that the pipeline could not hide), "predict" and "reset" have no separate pass (0).
    {
        const double dev_s = std::min(jt_seconds, kernel_ms * 1e-3), rest = std::max(0.0, jt_seconds - dev_s);
        auto pct = [&](double x) { return jt_seconds > 0 ? x / jt_seconds * 100 : 0.0; };
        std::cout << "jt: " << jt_seconds << " s " << "removing reset: " << jt_seconds << " s " << "removing norm: " << jt_seconds << " s ";
        std::cout << "load evidence: " << rest << " s (" << pct(rest) << "%) ";
        std::cout << "msg passing: " << dev_s << " s (" << pct(dev_s) << "%) ";
        std::cout << "reset: " << 0.0 << " s (" << 0.0 << "%) ";
        std::cout << "predict: " << 0.0 << " s (" << 0.0 << "%)" << std::endl;
        std::cout << "device kernels: " << kernel_ms * 1e-3 << " s (" << devices.size() << " GPU" << (devices.size() > 1 ? "s" : "") << ", "
                  << (jt_seconds > 0 ? num_instances / jt_seconds : 0.0) << " cases/s)" << std::endl << std::endl;
    }
    return accuracy;
}

// The file format of dataset/alarm/alarm_1k_pt: one line per (case, node), "%.7f " per state, an empty line for an
// evidence node. Needs keep_posteriors = true before EvaluateAccuracy.
void JunctionTree::WritePotentialTableFile(const string &path) const {
    const int n_nodes = network->num_nodes, sd = var_offsets[n_nodes];
    if (posteriors.size() != (size_t)num_instances * sd) {
        fprintf(stderr, "Error in function %s! set keep_posteriors before EvaluateAccuracy\n", __FUNCTION__);
        exit(1);
    }
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) {
        fprintf(stderr, "Error in function %s!Unable to open file %s!", __FUNCTION__, path.c_str());
        exit(1);
    }
    // cases are formatted in parallel (snprintf "%.7f " is the whole cost at 10^6 cases x 105 states), block by block in
    // file order
    const int kBlock = 2048;
    int nthreads = 1;
#ifdef _OPENMP
    nthreads = omp_get_max_threads();
#endif
    const int64_t super = (int64_t)kBlock * nthreads;
    vector<string> out((size_t)nthreads);
    for (int64_t s0 = 0; s0 < num_instances; s0 += super) {
#pragma omp parallel for schedule(static, 1)
        for (int t = 0; t < nthreads; ++t) {
            string &line = out[(size_t)t];
            line.clear();
            vector<char> is_ev(n_nodes);
            char tmp[360];
            const int64_t i0 = s0 + (int64_t)t * kBlock, i1 = std::min<int64_t>(num_instances, i0 + kBlock);
            for (int64_t i = i0; i < i1; ++i) {
                std::fill(is_ev.begin(), is_ev.end(), 0);
                for (int64_t k = ev_offsets[i]; k < ev_offsets[i + 1]; ++k)
                    if (ev_vars[k] >= 0 && ev_vars[k] < n_nodes) is_ev[ev_vars[k]] = 1;
                for (int v = 0; v < n_nodes; ++v) {
                    if (!is_ev[v])
                        for (int k = var_offsets[v]; k < var_offsets[v + 1]; ++k) {
                            int len = textio::FormatFixed7(posteriors[(size_t)i * sd + k], tmp);
                            tmp[len++] = ' ';
                            line.append(tmp, len);
                        }
                    line.push_back('\n');
                }
            }
        }
        for (int t = 0; t < nthreads; ++t) fwrite(out[(size_t)t].data(), 1, out[(size_t)t].size(), f);
    }
    fclose(f);
}
