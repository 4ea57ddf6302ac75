// Dataset::LoadLIBSVMDataKnownNetwork — test cases "label idx:val idx:val ... \n" into CSR arrays.
// Contract of the reference loader (src/Dataset.cpp:162-262) that is kept:
//   * `getline` + `while(!eof)`: a last line WITHOUT a trailing '\n' is dropped (:179-227)
//   * every line is right-trimmed of bytes < 33 before it is split on single spaces (:183-185)
//   * the first token is the value of the class variable (index 0), every other token is "index:value"
// What differs: one pass over the whole file with a hand-written integer scanner instead of
// getline/Split/stoi/vector<VarVal> per token — at 10^6 cases the reference-style ingestion takes longer than the GPU work.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>

#include "fastbn_host.h"

namespace {

[[noreturn]] void bad_line(int64_t line_no, const char *why) {
    fprintf(stderr, "Error in function LoadLIBSVMDataKnownNetwork! line %lld: %s\n", (long long)(line_no + 1), why);
    exit(1);
}

// scans an optionally signed decimal integer (stoi semantics: leading blanks were already consumed); false if none
inline bool scan_int(const char *&p, const char *e, int &out) {
    const char *s = p;
    bool neg = false;
    if (s < e && (*s == '-' || *s == '+')) neg = *s++ == '-';
    if (s >= e || *s < '0' || *s > '9') return false;
    long long v = 0;
    while (s < e && *s >= '0' && *s <= '9') {
        v = v * 10 + (*s++ - '0');
        if (v > 2147483648LL) return false;
    }
    out = (int)(neg ? -v : v);
    p = s;
    return true;
}

}  // namespace

void Dataset::LoadLIBSVMDataKnownNetwork(string data_file_path, int num_nodes, set<int> cont_vars) {
    FILE *f = fopen(data_file_path.c_str(), "rb");
    if (!f) {
        fprintf(stderr, "Error in function %s!", __FUNCTION__);
        fprintf(stderr, "Unable to open file %s!", data_file_path.c_str());
        exit(1);
    }
    if (!cont_vars.empty()) {
        fprintf(stderr, "Error in function %s! continuous variables are not supported on the junction-tree path\n", __FUNCTION__);
        exit(1);
    }
    string buf;
    {
        fseek(f, 0, SEEK_END);
        long sz = ftell(f);
        fseek(f, 0, SEEK_SET);
        buf.resize(sz > 0 ? (size_t)sz : 0);
        size_t got = buf.empty() ? 0 : fread(&buf[0], 1, buf.size(), f);
        buf.resize(got);
        fclose(f);
    }
    class_var_index = 0;
    ev_offsets.assign(1, 0);
    ev_vars.clear();
    ev_vals.clear();
    labels.clear();
    const char *p = buf.data(), *end = p + buf.size();
    int64_t line_no = 0;
    while (p < end) {
        const char *nl = (const char *)memchr(p, '\n', (size_t)(end - p));
        if (!nl) break;  // unterminated last line: dropped, like the reference's eof loop
        const char *e = nl;
        while (e > p && (unsigned char)e[-1] < 33) --e;  // TrimRight
        const char *q = p;
        int label;
        if (!scan_int(q, e, label)) bad_line(line_no, "no class label");
        labels.push_back(label);
        while (q < e) {
            if (*q != ' ') bad_line(line_no, "expected a single space between tokens");
            ++q;
            int idx, val;
            if (!scan_int(q, e, idx)) bad_line(line_no, "bad feature index");
            if (q >= e || *q != ':') bad_line(line_no, "expected index:value");
            ++q;
            if (!scan_int(q, e, val)) bad_line(line_no, "bad feature value");
            ev_vars.push_back(idx);
            ev_vals.push_back(val);
        }
        ev_offsets.push_back((int64_t)ev_vars.size());
        p = nl + 1;
        ++line_no;
    }
    num_instance = (int)labels.size();
    num_vars = num_nodes;
    std::cout << "Finish loading data. "
              << "Number of instances: " << num_instance << ". "
              << "Number of features: " << num_vars << ". " << std::endl;
}

vector<pair<int, int>> Dataset::Instance(int i) const {
    vector<pair<int, int>> out;
    out.emplace_back(class_var_index, labels[i]);
    for (int64_t k = ev_offsets[i]; k < ev_offsets[i + 1]; ++k) out.emplace_back(ev_vars[k], ev_vals[k]);
    return out;
}
