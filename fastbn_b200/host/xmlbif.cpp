// Network / DiscreteNode / CustomNetwork::GetNetFromXMLBIFFile: XMLBIF -> nodes with integer-count CPTs.
// Behaviour follows the reference loader (src/CustomNetwork.cpp:20-41, src/XMLBIFParser.cpp:3-181):
//   * node index = order of the <VARIABLE> elements; only TYPE "discrete" is kept (XMLBIFParser.cpp:45-69)
//   * <TABLE> digit order: the FOR variable is the most significant digit, then the <GIVEN>s in file order (:134-146)
//   * every entry becomes an INTEGER count (int)(p*10000) (:176, include/DiscreteNode.h:38) and probabilities are read
//     back with Laplace smoothing (count+1)/(total+dim) (src/DiscreteNode.cpp:152-161), so no entry is ever zero
// The XML reader below is a small purpose-built element-tree parser (the reference vendors tinyxml2 for this).
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>

#include "fastbn_host.h"

namespace {

struct XmlElem {
    string name, text;
    vector<std::unique_ptr<XmlElem>> kids;
    const XmlElem *first(const char *n) const {
        for (auto &k : kids)
            if (k->name == n) return k.get();
        return nullptr;
    }
    vector<const XmlElem *> all(const char *n) const {
        vector<const XmlElem *> out;
        for (auto &k : kids)
            if (k->name == n) out.push_back(k.get());
        return out;
    }
};

[[noreturn]] void die(const char *func, const string &msg) {  // the reference's error convention: print + exit(1)
    fprintf(stderr, "Error in function %s! %s\n", func, msg.c_str());
    exit(1);
}

string decode_entities(const string &s) {
    if (s.find('&') == string::npos) return s;
    static const pair<const char *, char> ent[] = {{"&lt;", '<'}, {"&gt;", '>'}, {"&amp;", '&'}, {"&quot;", '"'}, {"&apos;", '\''}};
    string out;
    for (size_t i = 0; i < s.size();) {
        bool hit = false;
        if (s[i] == '&')
            for (auto &e : ent) {
                size_t n = strlen(e.first);
                if (s.compare(i, n, e.first) == 0) {
                    out.push_back(e.second);
                    i += n;
                    hit = true;
                    break;
                }
            }
        if (!hit) out.push_back(s[i++]);
    }
    return out;
}

string trim(const string &s) {
    size_t a = 0, b = s.size();
    while (a < b && (unsigned char)s[a] < 33) ++a;
    while (b > a && (unsigned char)s[b - 1] < 33) --b;
    return s.substr(a, b - a);
}

// Recursive-descent over the byte buffer; attributes, comments, processing instructions and DOCTYPE are skipped.
class XmlReader {
public:
    XmlReader(const char *p, size_t n) : p_(p), end_(p + n) {}
    std::unique_ptr<XmlElem> parse_document() {
        std::unique_ptr<XmlElem> root;
        while (skip_misc()) {
            if (root) die("GetNetFromXMLBIFFile", "more than one root element in the XML file");
            root = parse_element();
        }
        return root;
    }

private:
    const char *p_, *end_;
    bool starts(const char *lit) const { size_t n = strlen(lit); return (size_t)(end_ - p_) >= n && memcmp(p_, lit, n) == 0; }
    void skip_until(const char *lit) {
        size_t n = strlen(lit);
        while (p_ < end_ && !starts(lit)) ++p_;
        if (p_ < end_) p_ += n;
    }
    // positions p_ at the next '<' that opens an element; false at end of input
    bool skip_misc() {
        while (p_ < end_) {
            if (*p_ != '<') { ++p_; continue; }
            if (starts("<!--")) skip_until("-->");
            else if (starts("<?")) skip_until("?>");
            else if (starts("<!")) skip_until(">");
            else return true;
        }
        return false;
    }
    std::unique_ptr<XmlElem> parse_element() {
        ++p_;  // '<'
        std::unique_ptr<XmlElem> e(new XmlElem());
        while (p_ < end_ && !isspace((unsigned char)*p_) && *p_ != '>' && *p_ != '/') e->name.push_back(*p_++);
        bool self_closing = false;
        char quote = 0;
        while (p_ < end_) {  // skip attributes
            char c = *p_++;
            if (quote) { if (c == quote) quote = 0; continue; }
            if (c == '"' || c == '\'') quote = c;
            else if (c == '/' && p_ < end_ && *p_ == '>') { ++p_; self_closing = true; break; }
            else if (c == '>') break;
        }
        if (self_closing) return e;
        string text;
        while (p_ < end_) {
            if (*p_ != '<') { text.push_back(*p_++); continue; }
            if (starts("<!--")) { skip_until("-->"); continue; }
            if (starts("<![CDATA[")) {
                p_ += 9;
                while (p_ < end_ && !starts("]]>")) text.push_back(*p_++);
                if (p_ < end_) p_ += 3;
                continue;
            }
            if (starts("<?")) { skip_until("?>"); continue; }
            if (starts("</")) { skip_until(">"); break; }
            e->kids.push_back(parse_element());
        }
        e->text = decode_entities(text);
        return e;
    }
};

string read_file(const string &path, const char *func) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) {
        fprintf(stderr, "Error in function %s!", func);
        fprintf(stderr, "Unable to open file %s!", path.c_str());
        exit(1);
    }
    string buf;
    char tmp[1 << 16];
    size_t n;
    while ((n = fread(tmp, 1, sizeof(tmp), f)) > 0) buf.append(tmp, n);
    fclose(f);
    return buf;
}

}  // namespace

// ---- DiscreteNode ----------------------------------------------------------------------------------------------
void DiscreteNode::SetCounts(vector<int> counts, vector<int> gdims) {
    given_dims = std::move(gdims);
    size_t ncfg = 1;
    for (int d : given_dims) ncfg *= (size_t)d;
    cpt_counts = std::move(counts);
    cfg_totals.assign(ncfg, 0);
    for (int q = 0; q < num_potential_vals; ++q)
        for (size_t c = 0; c < ncfg; ++c) cfg_totals[c] += cpt_counts[(size_t)q * ncfg + c];
}

void DiscreteNode::AddCount(int query_val, int cfg, int count) {  // src/DiscreteNode.cpp:139-145
    cpt_counts[(size_t)query_val * cfg_totals.size() + cfg] += count;
    cfg_totals[cfg] += count;
}

double DiscreteNode::GetProbabilityFlat(int query_val, int cfg) const {  // src/DiscreteNode.cpp:152-161
    int frequency_count = cpt_counts[(size_t)query_val * cfg_totals.size() + cfg];
    int total = cfg_totals[cfg];
    return (frequency_count + laplace_smooth) / (total + laplace_smooth * GetDomainSize());
}

int DiscreteNode::ParentConfigIndex(const DiscreteConfig &parents_config) const {
    int idx = 0;
    for (size_t i = 0; i < vec_given_order.size(); ++i) {
        int val = -1;
        for (auto &vv : parents_config)
            if (vv.first == vec_given_order[i]) { val = vv.second; break; }
        if (val < 0 || val >= given_dims[i]) die(__FUNCTION__, "parent configuration is not full");
        idx = idx * given_dims[i] + val;
    }
    return idx;
}

double DiscreteNode::GetProbability(int query_val, const DiscreteConfig &parents_config) const {
    return GetProbabilityFlat(query_val, ParentConfigIndex(parents_config));
}

// ---- Network ---------------------------------------------------------------------------------------------------
Network::~Network() {
    for (auto &kv : map_idx_node_ptr) delete kv.second;
}

Node *Network::FindNodePtrByIndex(const int &index) const {
    auto it = map_idx_node_ptr.find(index);
    if (it == map_idx_node_ptr.end()) die(__FUNCTION__, "Invalid index " + std::to_string(index));
    return it->second;
}

Node *Network::FindNodePtrByName(const string &name) const {
    for (auto &kv : map_idx_node_ptr)
        if (kv.second->node_name == name) return kv.second;
    return nullptr;
}

int **Network::ConvertDAGNetworkToAdjacencyMatrix() const {
    int **m = new int *[num_nodes];
    for (int i = 0; i < num_nodes; ++i) m[i] = new int[num_nodes]();
    for (auto &kv : map_idx_node_ptr)
        for (int p : kv.second->set_parent_indexes) m[p][kv.first] = 1;
    return m;
}

vector<int> Network::GetTopoOrd() const {  // Kahn, smallest index first
    vector<int> indeg(num_nodes, 0), order;
    vector<vector<int>> kids(num_nodes);
    for (auto &kv : map_idx_node_ptr)
        for (int p : kv.second->set_parent_indexes) { kids[p].push_back(kv.first); ++indeg[kv.first]; }
    set<int> ready;
    for (int i = 0; i < num_nodes; ++i)
        if (!indeg[i]) ready.insert(i);
    while (!ready.empty()) {
        int v = *ready.begin();
        ready.erase(ready.begin());
        order.push_back(v);
        for (int c : kids[v])
            if (--indeg[c] == 0) ready.insert(c);
    }
    return order;
}

// ---- CustomNetwork ---------------------------------------------------------------------------------------------
void CustomNetwork::GetNetFromXMLBIFFile(string file_path) {
    string buf = read_file(file_path, __FUNCTION__);
    XmlReader reader(buf.data(), buf.size());
    std::unique_ptr<XmlElem> doc = reader.parse_document();
    if (!doc || doc->name != "BIF") die(__FUNCTION__, "no <BIF> root element");
    const XmlElem *net = doc->first("NETWORK");
    if (!net) die(__FUNCTION__, "no <NETWORK> element");
    if (const XmlElem *nm = net->first("NAME")) network_name = trim(nm->text);
    vector<const XmlElem *> xvars = net->all("VARIABLE"), xprobs = net->all("PROBABILITY");
    if (xvars.size() != xprobs.size()) die("LoadFile", "Numbers of variables and probabilities not match!");
    if (xvars.empty()) die("GetUnconnectedNodes", "No variables!");

    map<string, DiscreteNode *> by_name;
    for (const XmlElem *xv : xvars) {
        const XmlElem *ty = xv->first("TYPE");
        if (!ty || trim(ty->text) != "discrete") continue;  // continuous variables are silently dropped by the reference too
        DiscreteNode *n = new DiscreteNode((int)map_idx_node_ptr.size());
        if (const XmlElem *nm = xv->first("NAME")) n->node_name = nm->text;
        for (const XmlElem *val : xv->all("VALUE")) n->vec_str_potential_vals.push_back(val->text);
        n->SetDomainSize((int)n->vec_str_potential_vals.size());
        map_idx_node_ptr[n->GetNodeIndex()] = n;
        if (!by_name.count(n->node_name)) by_name[n->node_name] = n;  // the reference matches the FIRST node of a name
    }
    num_nodes = (int)map_idx_node_ptr.size();

    for (const XmlElem *xp : xprobs) {
        const XmlElem *xfor = xp->first("FOR");
        auto it = xfor ? by_name.find(xfor->text) : by_name.end();
        if (it == by_name.end()) die("AssignProbsToNodes", "Probability does not match any variable!");
        DiscreteNode *node = it->second;
        vector<int> given, gdims;
        for (const XmlElem *xg : xp->all("GIVEN")) {
            auto ig = by_name.find(xg->text);
            if (ig == by_name.end()) die("AssignProbsToNodes", "\"GIVEN " + xg->text + "\" does not match any variable!");
            given.push_back(ig->second->GetNodeIndex());
            gdims.push_back(ig->second->GetDomainSize());
            node->set_parent_indexes.insert(ig->second->GetNodeIndex());
        }
        node->vec_given_order = given;
        const XmlElem *xt = xp->first("TABLE");
        if (!xt) die("AssignProbsToNodes", "PROBABILITY without TABLE");
        size_t expect = (size_t)node->GetDomainSize();
        for (int d : gdims) expect *= (size_t)d;
        vector<int> counts;
        counts.reserve(expect);
        const char *s = xt->text.c_str();
        while (true) {
            while (*s && (unsigned char)*s < 33) ++s;
            if (!*s) break;
            char *e = nullptr;
            double p = strtod(s, &e);
            if (e == s) die("AssignProbsToNodes", "bad number in TABLE");
            counts.push_back((int)(p * 10000));  // int count: truncation, e.g. 0.69 -> 6899 (src/XMLBIFParser.cpp:176)
            s = e;
        }
        if (counts.size() != expect) die("AssignProbsToNodes", "Two vectors have different sizes!");
        node->SetCounts(std::move(counts), gdims);
    }
}
