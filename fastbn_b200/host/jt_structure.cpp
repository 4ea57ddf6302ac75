// JunctionTreeStructure: Bayesian network -> cliques + separators with initial potentials.
// Same decisions as the reference builder (src/JunctionTreeStructure.cpp:12-365):
//   moralise (marry parents, drop directions)                                          :70-114
//   triangulate by repeatedly eliminating the remaining node with the FEWEST remaining neighbours, lowest index on
//     ties; {node + neighbours} becomes a clique unless an earlier clique already contains it      :127-219
//   maximum-weight spanning tree over clique intersections (Prim from clique 0, weight = #shared variables)  :225-308
//   every CPT is multiplied into the FIRST clique (creation order) that covers its family, in node-index order :314-347
// What is different is cost, not outcome: degrees are kept incrementally, Prim keeps one best edge per outside clique
// instead of rescanning O(#cliques^2) heap-allocated candidate separators (each with a full table) on every step, and
// ties between equal-weight candidates resolve by candidate creation order (i < j lexicographic) where the reference
// resolves them by pointer value (set<Separator*>), which is allocation order in practice.
#include <algorithm>
#include <climits>
#include <cstdio>
#include <iostream>

#include "fastbn_host.h"

JunctionTreeStructure::JunctionTreeStructure(Network *net) : network(net) {
    const int n = net->num_nodes;
    vector_clique_ptr_container.reserve(n);
    vector_separator_ptr_container.reserve(n > 0 ? n - 1 : 0);
    vector<vector<char>> adj(n, vector<char>(n, 0));
    Moralize(adj, net);
    Triangulate(adj);
    FormJunctionTree();
    std::cout << "Finish FormJunctionTree, number of cliques = " << vector_clique_ptr_container.size()
              << ", number of separators = " << vector_separator_ptr_container.size() << std::endl;
    AssignPotentials();
    int max_size = 0;
    float mean_size = GetAveAndMaxCliqueSize(max_size);
    std::cout << "Finish GetAveAndMaxCliqueSize, ave clique size = " << mean_size << ", max size = " << max_size << std::endl;
}

JunctionTreeStructure::~JunctionTreeStructure() {
    for (Clique *c : vector_clique_ptr_container) delete c;
    for (Separator *s : vector_separator_ptr_container) delete s;
}

void JunctionTreeStructure::Moralize(vector<vector<char>> &adj, const Network *net) {
    for (auto &kv : net->map_idx_node_ptr) {
        const set<int> &pa = kv.second->set_parent_indexes;
        for (int p : pa) {
            if (p == kv.first) continue;
            adj[p][kv.first] = adj[kv.first][p] = 1;
        }
        for (auto a = pa.begin(); a != pa.end(); ++a)
            for (auto b = std::next(a); b != pa.end(); ++b) adj[*a][*b] = adj[*b][*a] = 1;
    }
}

void JunctionTreeStructure::Triangulate(vector<vector<char>> &adj) {
    const int n = network->num_nodes;
    vector<int> degree(n, 0);
    vector<char> done(n, 0);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) degree[i] += adj[i][j];
    // clique membership lists per variable, to test "contained in an earlier clique" without scanning every clique
    vector<vector<int>> cliques_of_var(n);
    for (int step = 0; step < n; ++step) {
        int pick = -1, best = INT_MAX;
        for (int i = 0; i < n; ++i)
            if (!done[i] && degree[i] < best) { best = degree[i]; pick = i; }
        vector<int> nb;
        for (int j = 0; j < n; ++j)
            if (adj[pick][j]) nb.push_back(j);
        for (size_t a = 0; a < nb.size(); ++a)  // fill-in edges
            for (size_t b = a + 1; b < nb.size(); ++b)
                if (!adj[nb[a]][nb[b]]) {
                    adj[nb[a]][nb[b]] = adj[nb[b]][nb[a]] = 1;
                    ++degree[nb[a]];
                    ++degree[nb[b]];
                }
        set<int> members(nb.begin(), nb.end());
        members.insert(pick);
        // contained in an existing clique <=> some clique holding `pick` also holds every neighbour
        bool redundant = false;
        for (int c : cliques_of_var[pick]) {
            const set<int> &cv = vector_clique_ptr_container[c]->clique_variables;
            if (cv.size() >= members.size() && std::includes(cv.begin(), cv.end(), members.begin(), members.end())) {
                redundant = true;
                break;
            }
        }
        if (!redundant) {
            int id = (int)vector_clique_ptr_container.size();
            vector_clique_ptr_container.push_back(new Clique(members, network));
            for (int v : members) cliques_of_var[v].push_back(id);
        }
        done[pick] = 1;
        for (int j : nb) {
            adj[pick][j] = adj[j][pick] = 0;
            --degree[j];
        }
        degree[pick] = 0;
    }
}

void JunctionTreeStructure::FormJunctionTree() {
    const int nc = (int)vector_clique_ptr_container.size();
    if (nc <= 1) return;
    // cliques as sorted vectors for fast intersections
    vector<vector<int>> cv(nc);
    for (int i = 0; i < nc; ++i) cv[i].assign(vector_clique_ptr_container[i]->clique_variables.begin(), vector_clique_ptr_container[i]->clique_variables.end());
    auto shared = [&](int a, int b) {
        int k = 0;
        size_t i = 0, j = 0;
        while (i < cv[a].size() && j < cv[b].size()) {
            if (cv[a][i] < cv[b][j]) ++i;
            else if (cv[a][i] > cv[b][j]) ++j;
            else { ++k; ++i; ++j; }
        }
        return k;
    };
    // Prim, maximum weight; candidate (i<j) order key = i*nc + j decides ties (earliest-created candidate wins)
    vector<char> in_tree(nc, 0);
    vector<int> best_w(nc, 0), best_to(nc, -1);
    vector<int64_t> best_key(nc, INT64_MAX);
    auto relax = [&](int t) {
        for (int c = 0; c < nc; ++c) {
            if (in_tree[c]) continue;
            int w = shared(c, t);
            if (w == 0) continue;
            int64_t key = (int64_t)std::min(c, t) * nc + std::max(c, t);
            if (w > best_w[c] || (w == best_w[c] && key < best_key[c])) { best_w[c] = w; best_key[c] = key; best_to[c] = t; }
        }
    };
    in_tree[0] = 1;
    relax(0);
    for (int added = 1; added < nc; ++added) {
        int pick = -1;
        for (int c = 0; c < nc; ++c) {
            if (in_tree[c] || best_to[c] < 0) continue;
            if (pick < 0 || best_w[c] > best_w[pick] || (best_w[c] == best_w[pick] && best_key[c] < best_key[pick])) pick = c;
        }
        if (pick < 0) {  // disconnected moral graph: the reference dereferences a null separator here
            fprintf(stderr, "Error in function %s! the network is not connected: no separator joins the remaining cliques\n", __FUNCTION__);
            exit(1);
        }
        int a = std::min(pick, best_to[pick]), b = std::max(pick, best_to[pick]);
        set<int> common;
        std::set_intersection(cv[a].begin(), cv[a].end(), cv[b].begin(), cv[b].end(), std::inserter(common, common.begin()));
        Separator *sep = new Separator(common, network, true);
        sep->is_in_jt = true;
        sep->set_neighbours_ptr.push_back(vector_clique_ptr_container[a]);
        sep->set_neighbours_ptr.push_back(vector_clique_ptr_container[b]);
        vector_separator_ptr_container.push_back(sep);
        in_tree[pick] = 1;
        relax(pick);
    }
    for (Separator *s : vector_separator_ptr_container)
        for (Clique *c : s->set_neighbours_ptr) c->set_neighbours_ptr.push_back(s);
}

void JunctionTreeStructure::AssignPotentials() {
    for (auto &kv : network->map_idx_node_ptr) {  // ascending node index
        DiscreteNode *node = kv.second;
        if (!node->is_discrete) continue;
        PotentialTable cpt(node, network);
        for (Clique *c : vector_clique_ptr_container) {
            if (std::includes(c->clique_variables.begin(), c->clique_variables.end(), cpt.vec_related_variables.begin(),
                              cpt.vec_related_variables.end())) {
                c->p_table.TableMultiplication(cpt);
                break;  // each factor is used exactly once
            }
        }
    }
}

float JunctionTreeStructure::GetAveAndMaxCliqueSize(int &max_size) const {
    int total_size = 0;  // int, like the reference (the printed average inherits its overflow behaviour)
    for (Clique *c : vector_clique_ptr_container) {
        total_size += c->p_table.table_size;
        max_size = std::max(max_size, c->p_table.table_size);
    }
    return (float)total_size / (float)vector_clique_ptr_container.size();
}
