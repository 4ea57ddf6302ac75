// PotentialTable / Clique / Separator: the dense factor type of the junction tree (host side, used to BUILD the tree;
// per-test-case table arithmetic runs on the GPU). Layout and operations follow the reference:
//   row-major, vec_related_variables[0] most significant; cum_levels[n-1] = 1, cum_levels[i] = cum_levels[i+1]*var_dims[i+1]
//   (src/PotentialTableBase.cpp:599-607); index <-> digits (:547-566); CPT-as-factor with ascending variables
//   (src/PotentialTable.cpp:16-78); multiplication into the larger table (:649-667); 0-guarded division (:718-734).
// Index maps are generated with an odometer over strides instead of per-entry div/mod + heap arrays.
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "fastbn_host.h"

namespace {
// For a table with dimensions `dims`, walks all entries in storage order and yields sum(digit[i] * stride[i]) for an
// arbitrary stride vector (0 for variables a smaller table does not contain).
template <class F>
void for_each_offset(const vector<int> &dims, const vector<int64_t> &stride, F f) {
    const int n = (int)dims.size();
    int64_t total = 1;
    for (int d : dims) total *= d;
    vector<int> dig(n, 0);
    int64_t cur = 0;
    for (int64_t e = 0; e < total; ++e) {
        f(e, cur);
        for (int i = n - 1; i >= 0; --i) {
            if (++dig[i] < dims[i]) { cur += stride[i]; break; }
            cur -= (int64_t)(dims[i] - 1) * stride[i];
            dig[i] = 0;
        }
    }
}
}  // namespace

void PotentialTable::ConstructCumLevels() {
    cum_levels.assign(num_variables, 1);
    for (int i = num_variables - 2; i >= 0; --i) cum_levels[i] = cum_levels[i + 1] * var_dims[i + 1];
}

void PotentialTable::ConstructVarDimsAndCumLevels(Network *net) {
    var_dims.resize(num_variables);
    int64_t size = 1;
    for (int i = 0; i < num_variables; ++i) {
        var_dims[i] = net->FindNodePtrByIndex(vec_related_variables[i])->GetDomainSize();
        size *= var_dims[i];
        if (size > (int64_t)1 << 30) {  // the reference overflows int32 silently here
            fprintf(stderr, "Error in function %s! a table over %d variables exceeds 2^30 entries\n", __FUNCTION__, num_variables);
            exit(1);
        }
    }
    ConstructCumLevels();
    table_size = (int)size;
}

PotentialTable::PotentialTable(DiscreteNode *node, Network *net) {
    set<int> scope(node->set_parent_indexes);
    scope.insert(node->GetNodeIndex());
    vec_related_variables.assign(scope.begin(), scope.end());
    num_variables = (int)vec_related_variables.size();
    ConstructVarDimsAndCumLevels(net);
    // stride of every table variable inside the node's flat CPT (value-major, then parents in <GIVEN> order)
    const vector<int> &given = node->vec_given_order, &gdims = node->GivenDims();
    int64_t ncfg = 1;
    for (int d : gdims) ncfg *= d;
    vector<int64_t> stride(num_variables, 0);
    for (int i = 0; i < num_variables; ++i) {
        int v = vec_related_variables[i];
        if (v == node->GetNodeIndex()) { stride[i] = ncfg; continue; }
        int64_t s = 1;
        for (int k = (int)given.size() - 1; k >= 0; --k) {
            if (given[k] == v) { stride[i] = s; break; }  // a parent listed twice keeps its last (least significant) digit
            s *= gdims[k];
        }
    }
    potentials.resize(table_size);
    for_each_offset(var_dims, stride, [&](int64_t e, int64_t off) {
        potentials[e] = node->GetProbabilityFlat((int)(off / ncfg), (int)(off % ncfg));
    });
}

void PotentialTable::ConstructEmptyPotentialTable(const set<int> &set_node_index, Network *net) {
    vec_related_variables.assign(set_node_index.begin(), set_node_index.end());
    num_variables = (int)vec_related_variables.size();
    ConstructVarDimsAndCumLevels(net);
    potentials.assign(table_size, 1.0);
}

int PotentialTable::GetVariableIndex(const int &variable) const {
    for (int i = 0; i < num_variables; ++i)
        if (vec_related_variables[i] == variable) return i;
    return num_variables;
}

void PotentialTable::GetConfigValueByTableIndex(const int &table_index, int *config_value) const {
    int rest = table_index;
    for (int i = 0; i < num_variables; ++i) {
        config_value[i] = rest / cum_levels[i];
        rest -= config_value[i] * cum_levels[i];
    }
}

int PotentialTable::GetTableIndexByConfigValue(const int *config_value) const {
    int idx = 0;
    for (int i = 0; i < num_variables; ++i) idx += config_value[i] * cum_levels[i];
    return idx;
}

void PotentialTable::TableMultiplication(const PotentialTable &second) {
    vector<int64_t> stride(num_variables, 0);
    int found = 0;
    for (int i = 0; i < num_variables; ++i) {
        int loc = second.GetVariableIndex(vec_related_variables[i]);
        if (loc < second.num_variables) { stride[i] = second.cum_levels[loc]; ++found; }
    }
    if (found != second.num_variables) {
        fprintf(stderr, "Error in function %s! the second table has variables outside this table\n", __FUNCTION__);
        exit(1);
    }
    for_each_offset(var_dims, stride, [&](int64_t e, int64_t off) { potentials[e] *= second.potentials[off]; });
}

void PotentialTable::TableReorganization(const vector<int> &new_order) {
    if ((int)new_order.size() != num_variables) {
        fprintf(stderr, "Error in function %s! the new order must be a permutation of the table's variables\n", __FUNCTION__);
        exit(1);
    }
    PotentialTable nt;
    nt.vec_related_variables = new_order;
    nt.num_variables = num_variables;
    nt.var_dims.resize(num_variables);
    vector<int64_t> stride(num_variables);
    for (int i = 0; i < num_variables; ++i) {
        int loc = GetVariableIndex(new_order[i]);
        if (loc >= num_variables) {
            fprintf(stderr, "Error in function %s! variable %d is not in the table\n", __FUNCTION__, new_order[i]);
            exit(1);
        }
        nt.var_dims[i] = var_dims[loc];
        stride[i] = cum_levels[loc];
    }
    nt.ConstructCumLevels();
    nt.table_size = table_size;
    nt.potentials.resize(table_size);
    for_each_offset(nt.var_dims, stride, [&](int64_t e, int64_t off) { nt.potentials[e] = potentials[off]; });
    *this = std::move(nt);
}

void PotentialTable::TableMarginalization(const vector<int> &keep) {
    PotentialTable nt;
    nt.vec_related_variables = keep;
    nt.num_variables = (int)keep.size();
    nt.var_dims.resize(nt.num_variables);
    for (int i = 0; i < nt.num_variables; ++i) nt.var_dims[i] = var_dims[GetVariableIndex(keep[i])];
    nt.ConstructCumLevels();
    nt.table_size = nt.num_variables ? nt.cum_levels[0] * nt.var_dims[0] : 1;
    nt.potentials.assign(nt.table_size, 0.0);
    vector<int64_t> stride(num_variables, 0);
    for (int i = 0; i < num_variables; ++i) {
        int loc = nt.GetVariableIndex(vec_related_variables[i]);
        if (loc < nt.num_variables) stride[i] = nt.cum_levels[loc];
    }
    for_each_offset(var_dims, stride, [&](int64_t e, int64_t off) { nt.potentials[off] += potentials[e]; });
    *this = std::move(nt);
}

void PotentialTable::TableDivision(const PotentialTable &second) {
    for (int i = 0; i < table_size; ++i) potentials[i] = second.potentials[i] == 0 ? 0 : potentials[i] / second.potentials[i];
}

void PotentialTable::Normalize() {  // no zero guard, like src/PotentialTableBase.cpp:433-445
    double denom = 0;
    for (double p : potentials) denom += p;
    for (double &p : potentials) p /= denom;
}

Clique::Clique(set<int> set_node_index, Network *net) : clique_variables(std::move(set_node_index)) {
    p_table.ConstructEmptyPotentialTable(clique_variables, net);
}

// Candidate separators are created without a table (with_table = false): the reference allocates a full table of ones
// for each of the O(#cliques^2) candidates (src/JunctionTreeStructure.cpp:231-254); only the chosen ones need one.
Separator::Separator(set<int> set_node_index, Network *net, bool with_table) {
    clique_variables = std::move(set_node_index);
    if (with_table) {
        p_table.ConstructEmptyPotentialTable(clique_variables, net);
    } else {
        p_table.vec_related_variables.assign(clique_variables.begin(), clique_variables.end());
        p_table.num_variables = (int)clique_variables.size();
    }
}
