// Dataset::LoadLIBSVMDataKnownNetwork — test cases "label idx:val idx:val ... \n" into CSR arrays.
// Contract of the reference loader (src/Dataset.cpp:162-262) that is kept:
//   * `getline` + `while(!eof)`: a last line WITHOUT a trailing '\n' is dropped (:179-227)
//   * every line is right-trimmed of bytes < 33 before it is split on single spaces (:183-185)
//   * the first token is the value of the class variable (index 0), every other token is "index:value"
// What differs: the file is mmap'ed and parsed in one pass by several threads (pieces of whole lines, a hand-written integer
// scanner) instead of getline/Split/stoi/vector<VarVal> per token — at 10^6 cases the reference-style ingestion takes longer
// than the GPU work — and the parsed arrays can be kept in a binary cache beside the file (FASTBN_EVIDENCE_CACHE=1 / -C).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#ifdef _OPENMP
#include <omp.h>
#endif

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

namespace {

// one piece of the file (whole lines): parsed independently, concatenated in file order afterwards
struct Piece {
    vector<int> labels, vars, vals;
    vector<int64_t> ends;   // evidence pairs of the piece up to and including line i
    int64_t first_line = 0;
};

void parse_piece(const char *p, const char *end, Piece &out) {
    int64_t line_no = out.first_line;
    while (p < end) {
        const char *nl = (const char *)memchr(p, '\n', (size_t)(end - p));
        if (!nl) break;  // unterminated last line: dropped, like the reference's eof loop
        const char *e = nl;
        while (e > p && (unsigned char)e[-1] < 33) --e;  // TrimRight
        const char *q = p;
        int label;
        if (!scan_int(q, e, label)) bad_line(line_no, "no class label");
        out.labels.push_back(label);
        while (q < e) {
            if (*q != ' ') bad_line(line_no, "expected a single space between tokens");
            ++q;
            int idx, val;
            if (!scan_int(q, e, idx)) bad_line(line_no, "bad feature index");
            if (q >= e || *q != ':') bad_line(line_no, "expected index:value");
            ++q;
            if (!scan_int(q, e, val)) bad_line(line_no, "bad feature value");
            out.vars.push_back(idx);
            out.vals.push_back(val);
        }
        out.ends.push_back((int64_t)out.vars.size());
        p = nl + 1;
        ++line_no;
    }
}

// Binary evidence cache "<file>.fbncache" (opt-in: FASTBN_EVIDENCE_CACHE=1 or the CLI's -C): the parsed CSR arrays behind a
// header that pins the text file's size and modification time; anything that does not match is ignored and rewritten.
struct CacheHeader {
    char magic[8];        // "FBNEVC1\0"
    int64_t src_size, src_mtime_ns, n_cases, n_pairs;
};
const char kCacheMagic[8] = {'F', 'B', 'N', 'E', 'V', 'C', '1', 0};

bool cache_enabled() {
    const char *e = getenv("FASTBN_EVIDENCE_CACHE");
    return e && *e && *e != '0';
}

bool read_cache(const string &path, const struct stat &st, Dataset &d) {
    FILE *f = fopen((path + ".fbncache").c_str(), "rb");
    if (!f) return false;
    CacheHeader h;
    bool ok = fread(&h, sizeof(h), 1, f) == 1 && memcmp(h.magic, kCacheMagic, 8) == 0 && h.src_size == (int64_t)st.st_size &&
              h.src_mtime_ns == (int64_t)st.st_mtim.tv_sec * 1000000000LL + st.st_mtim.tv_nsec && h.n_cases >= 0 && h.n_pairs >= 0 &&
              h.n_cases < ((int64_t)1 << 31);
    if (ok) {
        d.labels.resize((size_t)h.n_cases);
        d.ev_offsets.resize((size_t)h.n_cases + 1);
        d.ev_vars.resize((size_t)h.n_pairs);
        d.ev_vals.resize((size_t)h.n_pairs);
        ok = fread(d.labels.data(), 4, (size_t)h.n_cases, f) == (size_t)h.n_cases &&
             fread(d.ev_offsets.data(), 8, (size_t)h.n_cases + 1, f) == (size_t)h.n_cases + 1 &&
             fread(d.ev_vars.data(), 4, (size_t)h.n_pairs, f) == (size_t)h.n_pairs &&
             fread(d.ev_vals.data(), 4, (size_t)h.n_pairs, f) == (size_t)h.n_pairs && d.ev_offsets[0] == 0 &&
             d.ev_offsets[(size_t)h.n_cases] == h.n_pairs;
        for (int64_t i = 0; ok && i < h.n_cases; ++i) ok = d.ev_offsets[(size_t)i] <= d.ev_offsets[(size_t)i + 1];
    }
    fclose(f);
    return ok;
}

void write_cache(const string &path, const struct stat &st, const Dataset &d) {
    const string tmp = path + ".fbncache.tmp" + std::to_string((long long)getpid());
    FILE *f = fopen(tmp.c_str(), "wb");
    if (!f) return;  // read-only data directory: no cache, no complaint
    CacheHeader h;
    memcpy(h.magic, kCacheMagic, 8);
    h.src_size = (int64_t)st.st_size;
    h.src_mtime_ns = (int64_t)st.st_mtim.tv_sec * 1000000000LL + st.st_mtim.tv_nsec;
    h.n_cases = (int64_t)d.labels.size();
    h.n_pairs = (int64_t)d.ev_vars.size();
    bool ok = fwrite(&h, sizeof(h), 1, f) == 1 && fwrite(d.labels.data(), 4, d.labels.size(), f) == d.labels.size() &&
              fwrite(d.ev_offsets.data(), 8, d.ev_offsets.size(), f) == d.ev_offsets.size() &&
              fwrite(d.ev_vars.data(), 4, d.ev_vars.size(), f) == d.ev_vars.size() &&
              fwrite(d.ev_vals.data(), 4, d.ev_vals.size(), f) == d.ev_vals.size();
    ok = fclose(f) == 0 && ok;
    if (!ok || rename(tmp.c_str(), (path + ".fbncache").c_str()) != 0) remove(tmp.c_str());
}

}  // namespace

void Dataset::LoadLIBSVMDataKnownNetwork(string data_file_path, int num_nodes, set<int> cont_vars) {
    const int fd = open(data_file_path.c_str(), O_RDONLY);
    struct stat st;
    if (fd < 0 || fstat(fd, &st) != 0) {
        fprintf(stderr, "Error in function %s!", __FUNCTION__);
        fprintf(stderr, "Unable to open file %s!", data_file_path.c_str());
        exit(1);
    }
    if (!cont_vars.empty()) {
        fprintf(stderr, "Error in function %s! continuous variables are not supported on the junction-tree path\n", __FUNCTION__);
        exit(1);
    }
    class_var_index = 0;
    const bool use_cache = cache_enabled();
    if (use_cache && read_cache(data_file_path, st, *this)) {
        close(fd);
    } else {
        // the file is mapped, not copied; pieces of whole lines are parsed by several threads (one pass, hand-written integer
        // scanner) and concatenated in file order
        const size_t size = (size_t)st.st_size;
        const char *base = size ? (const char *)mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0) : nullptr;
        if (size && base == (const char *)MAP_FAILED) {
            fprintf(stderr, "Error in function %s! cannot map %s\n", __FUNCTION__, data_file_path.c_str());
            exit(1);
        }
        if (size) madvise((void *)base, size, MADV_SEQUENTIAL);
        int nthreads = 1;
#ifdef _OPENMP
        nthreads = omp_get_max_threads();
#endif
        const int npieces = (int)std::max<size_t>(1, std::min<size_t>((size_t)nthreads, size / ((size_t)4 << 20)));
        vector<size_t> cut((size_t)npieces + 1, size);
        cut[0] = 0;
        for (int k = 1; k < npieces; ++k) {  // piece boundaries just after a newline
            size_t c = size / (size_t)npieces * (size_t)k;
            const char *nl = c < size ? (const char *)memchr(base + c, '\n', size - c) : nullptr;
            cut[(size_t)k] = nl ? (size_t)(nl - base) + 1 : size;
        }
        vector<Piece> pieces((size_t)npieces);
        if (npieces > 1) {  // line numbers (for error messages) need the newline counts of the earlier pieces
            vector<int64_t> nls((size_t)npieces, 0);
#pragma omp parallel for schedule(static, 1)
            for (int k = 0; k < npieces; ++k) {
                int64_t c = 0;
                for (const char *q = base + cut[(size_t)k], *e = base + cut[(size_t)k + 1]; q < e;) {
                    const char *nl = (const char *)memchr(q, '\n', (size_t)(e - q));
                    if (!nl) break;
                    ++c;
                    q = nl + 1;
                }
                nls[(size_t)k] = c;
            }
            for (int k = 1; k < npieces; ++k) pieces[(size_t)k].first_line = pieces[(size_t)k - 1].first_line + nls[(size_t)k - 1];
        }
#pragma omp parallel for schedule(static, 1)
        for (int k = 0; k < npieces; ++k) parse_piece(base + cut[(size_t)k], base + cut[(size_t)k + 1], pieces[(size_t)k]);
        size_t n_cases = 0, n_pairs = 0;
        for (const Piece &pc : pieces) { n_cases += pc.labels.size(); n_pairs += pc.vars.size(); }
        labels.resize(n_cases);
        ev_offsets.resize(n_cases + 1);
        ev_vars.resize(n_pairs);
        ev_vals.resize(n_pairs);
        ev_offsets[0] = 0;
        size_t c0 = 0, e0 = 0;
        for (const Piece &pc : pieces) {
            if (!pc.labels.empty()) memcpy(&labels[c0], pc.labels.data(), pc.labels.size() * 4);
            if (!pc.vars.empty()) {
                memcpy(&ev_vars[e0], pc.vars.data(), pc.vars.size() * 4);
                memcpy(&ev_vals[e0], pc.vals.data(), pc.vals.size() * 4);
            }
            for (size_t i = 0; i < pc.ends.size(); ++i) ev_offsets[c0 + i + 1] = (int64_t)e0 + pc.ends[i];
            c0 += pc.labels.size();
            e0 += pc.vars.size();
        }
        if (size) munmap((void *)base, size);
        close(fd);
        if (use_cache) write_cache(data_file_path, st, *this);
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
