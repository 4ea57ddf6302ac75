// Host-side mirror of the reference's front end for the `-a 2` junction-tree path: same class names, constructor /
// virtual signatures, public members and CLI as Tonanguyxiro/FastBN, written from scratch. The tree is built here on
// the host (once per run); every per-test-case computation happens in the CUDA library behind include/fastbn_b200.h.
//
//   reference                                   here
//   include/Network.h, CustomNetwork.h          Network, CustomNetwork::GetNetFromXMLBIFFile      (xmlbif.cpp)
//   include/DiscreteNode.h                      DiscreteNode (flat integer count table, same Laplace semantics)
//   include/Dataset.h                           Dataset::LoadLIBSVMDataKnownNetwork (CSR storage)  (dataset.cpp)
//   include/PotentialTableBase.h, PotentialTable.h   PotentialTable                                 (potential_table.cpp)
//   include/Clique.h, Separator.h               Clique, Separator
//   include/JunctionTreeStructure.h             JunctionTreeStructure                              (jt_structure.cpp)
//   include/Inference.h                         Inference
//   include/JunctionTree.h                      JunctionTree                                       (junction_tree.cpp)
//   src/main.cpp, src/Parameter.cpp             main.cpp -> fastbn_b200/bin/BayesianNetwork
#pragma once
#include <cstdint>
#include <map>
#include <set>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/fastbn_b200.h"

using std::map;
using std::pair;
using std::set;
using std::string;
using std::vector;

typedef set<pair<int, int>> DiscreteConfig;  // include/common.h:22

// exact fast paths of "%.7f" formatting and decimal parsing for the potential-table files (textio.cpp)
namespace textio {
int FormatFixed7(double x, char *out352);
const char *ParseDouble(const char *p, const char *e, double *out);
}  // namespace textio

// ---------------------------------------------------------------------------------------------------------------
// A discrete variable with its CPT kept as the reference keeps it: INTEGER counts (int)(p*10000) per (value, parent
// configuration) and Laplace smoothing at read time (src/XMLBIFParser.cpp:176, src/DiscreteNode.cpp:129-161).
class DiscreteNode {
public:
    string node_name;
    vector<string> vec_str_potential_vals;
    set<int> set_parent_indexes;
    vector<int> vec_given_order;  // parents in <GIVEN> order (the digit order of <TABLE>, src/XMLBIFParser.cpp:134-146)
    bool is_discrete = true;

    explicit DiscreteNode(int index = -1) : node_index(index) {}
    int GetNodeIndex() const { return node_index; }
    int GetDomainSize() const { return num_potential_vals; }
    void SetDomainSize(int size) { num_potential_vals = size; }
    bool HasParents() const { return !set_parent_indexes.empty(); }
    int GetNumParents() const { return (int)set_parent_indexes.size(); }
    double GetLaplaceSmooth() const { return laplace_smooth; }
    // counts are FOR-major: counts[query * n_parent_configs + cfg], cfg row-major over vec_given_order
    void SetCounts(vector<int> counts, vector<int> given_dims);
    void AddCount(int query_val, int parent_config_index, int count);
    double GetProbabilityFlat(int query_val, int parent_config_index) const;
    // reference signature (parents_config must be a full parent configuration)
    double GetProbability(int query_val, const DiscreteConfig &parents_config) const;
    int ParentConfigIndex(const DiscreteConfig &parents_config) const;
    const vector<int> &GivenDims() const { return given_dims; }

private:
    int node_index;
    int num_potential_vals = -1;
    double laplace_smooth = 1;  // include/DiscreteNode.h:60
    vector<int> cpt_counts, cfg_totals, given_dims;
};
typedef DiscreteNode Node;

class Network {
public:
    string network_name;
    int num_nodes = 0;
    bool pure_discrete;
    map<int, Node *> map_idx_node_ptr;  // owns the nodes (include/Network.h)

    explicit Network(bool pure_disc = true) : pure_discrete(pure_disc) {}
    virtual ~Network();
    Node *FindNodePtrByIndex(const int &index) const;
    Node *FindNodePtrByName(const string &name) const;
    int **ConvertDAGNetworkToAdjacencyMatrix() const;  // [parent][child] = 1; caller frees (src/Network.cpp)
    vector<int> GetTopoOrd() const;
};

class CustomNetwork : public Network {
public:
    explicit CustomNetwork(bool pure_disc = true) : Network(pure_disc) {}
    void GetNetFromXMLBIFFile(string file_path);  // src/CustomNetwork.cpp:20-41 + src/XMLBIFParser.cpp
};

// ---------------------------------------------------------------------------------------------------------------
// Test cases in LIBSVM format "label idx:val idx:val ...". Same loader contract as src/Dataset.cpp:162-262 (a last
// line without '\n' is dropped, lines are right-trimmed), but stored as CSR, the layout fbn_run takes.
class Dataset {
public:
    int num_instance = 0;
    int num_vars = 0;
    int class_var_index = 0;
    vector<int64_t> ev_offsets;  // [num_instance + 1]
    vector<int32_t> ev_vars, ev_vals;
    vector<int32_t> labels;      // value of the class variable per case

    void LoadLIBSVMDataKnownNetwork(string data_file_path, int num_nodes, set<int> cont_vars = set<int>());
    // reference view of one case: (class_var_index, label) first, then the features (vector_dataset_all_vars[i])
    vector<pair<int, int>> Instance(int i) const;
};

// ---------------------------------------------------------------------------------------------------------------
// Dense factor, row-major with vec_related_variables[0] most significant (include/PotentialTableBase.h:23-38).
class PotentialTable {
public:
    vector<int> vec_related_variables;
    vector<double> potentials;
    vector<int> var_dims;
    vector<int> cum_levels;
    int num_variables = 0;
    int table_size = 0;

    PotentialTable() = default;
    PotentialTable(DiscreteNode *disc_node, Network *net);      // CPT as a factor, variables ascending
    void ConstructEmptyPotentialTable(const set<int> &set_node_index, Network *net);  // all ones
    void ConstructVarDimsAndCumLevels(Network *net);
    void ConstructCumLevels();
    int GetVariableIndex(const int &variable) const;            // position, num_variables when absent
    void GetConfigValueByTableIndex(const int &table_index, int *config_value) const;
    int GetTableIndexByConfigValue(const int *config_value) const;
    void TableMultiplication(const PotentialTable &second_table);  // this *= second (scope(second) within scope(this))
    void TableReorganization(const vector<int> &new_order);        // same variables, new storage order
    void TableMarginalization(const vector<int> &keep_variables);  // sum out everything else
    void TableDivision(const PotentialTable &second_table);        // 0-guarded (src/PotentialTable.cpp:718-734)
    void Normalize();
};

class Clique {
public:
    set<int> clique_variables;
    PotentialTable p_table;
    vector<Clique *> set_neighbours_ptr;  // insertion order (the reference's set<Clique*> iterates in pointer order)
    Clique *ptr_upstream_clique = nullptr;
    vector<Clique *> ptr_downstream_cliques;

    Clique() = default;
    Clique(set<int> set_node_index, Network *net);
    virtual ~Clique() {}
};

class Separator : public Clique {
public:
    bool is_in_jt = false;
    Separator() = default;
    Separator(set<int> set_node_index, Network *net, bool with_table);
};

class JunctionTreeStructure {
public:
    Network *network = nullptr;
    vector<Clique *> vector_clique_ptr_container;
    vector<Separator *> vector_separator_ptr_container;

    JunctionTreeStructure() = default;
    explicit JunctionTreeStructure(Network *net);
    ~JunctionTreeStructure();

protected:
    static void Moralize(vector<vector<char>> &adj, const Network *net);
    void Triangulate(vector<vector<char>> &adj);
    void FormJunctionTree();
    void AssignPotentials();
    float GetAveAndMaxCliqueSize(int &max_size) const;
};

// ---------------------------------------------------------------------------------------------------------------
class Inference {
public:
    Network *network = nullptr;
    int num_instances = 0;
    int query_index = 0;
    // evidences of every test case, CSR (the reference keeps vector<DiscreteConfig>; Evidence(i) gives that view)
    vector<int64_t> ev_offsets;
    vector<int32_t> ev_vars, ev_vals;
    vector<int> ground_truths;
    // flattened ground_truth_probability_tables: num_instances x sum_dims, first entry of an evidence node = -1
    vector<double> ground_truth_probability_tables;
    vector<int> var_offsets;  // prefix sums of domain sizes, [num_nodes + 1]

    Inference(Network *net, Dataset *dts, bool is_dense);
    virtual ~Inference() {}
    virtual double EvaluateAccuracy(string path, int num_threads) = 0;
    double Accuracy(vector<int> predictions);
    int ArgMax(const vector<double> &array);
    void LoadGroundTruthProbabilityTable(string file_path);
    DiscreteConfig Evidence(int i) const;
};

class JunctionTree : public Inference {
public:
    JunctionTreeStructure *tree = nullptr;
    Clique *jt_root = nullptr;
    vector<vector<Clique *>> nodes_by_level;
    vector<vector<Separator *>> separators_by_level;
    int max_level = 0;
    vector<vector<double>> probs_one_sample;  // posteriors of the LAST test case, as after the reference's loop

    // extensions (not in the reference)
    vector<int> devices;             // CUDA devices to shard the cases over; default {0} or $FASTBN_DEVICES
    bool keep_posteriors = false;    // keep all posteriors on the host after EvaluateAccuracy
    vector<double> posteriors;       // num_instances x sum_dims when keep_posteriors
    vector<int> predictions;         // labels of the last EvaluateAccuracy
    double average_mse = 0, average_hd = 0, jt_seconds = 0, kernel_ms = 0;
    int64_t n_correct_reduced = 0;  // correct predictions summed over the device slices (fbn_reduce when several devices)
    bool quiet = false;
    std::thread warmup;              // brings the CUDA contexts of `devices` up in the background (joined before the first plan)

    JunctionTree(Network *net, Dataset *dts, bool is_dense);
    ~JunctionTree();
    virtual double EvaluateAccuracy(string path, int num_threads);
    void ResetJunctionTree() {}  // host tables are never mutated: the device resets from its pristine copy per case
    // flat description of the tree in the layout of fbn_tree_desc (arrays owned by the returned object)
    struct FlatTree {
        vector<int32_t> var_dims, table_nvars, table_vars, sep_parent, sep_child;
        vector<double> clique_potentials;
        fbn_tree_desc desc;
    };
    void Flatten(FlatTree &out) const;
    void WritePotentialTableFile(const string &path) const;  // "%.7f " per state, empty line per evidence node

protected:
    int GetMaxLevel(Clique *root);
    void MarkLevel();
    void ReorganizeTableStorage(int num_threads);
    vector<int> PredictUseJTInfer(int num_threads);
};
