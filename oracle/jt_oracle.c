/*
 * TEST INFRASTRUCTURE ONLY — CPU restatement (plain C) of the reference's junction-tree inference for one test
 * case, used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg as the CHECKER. The product
 * (fastbn_b200/) never links, loads or calls this file.
 *
 * Parity status: PINNED. tests/test_oracle.py checks this file against (a) the reference's shipped golden posteriors
 * dataset/alarm/alarm_1k_pt (7 decimals), (b) full-precision fp64 posteriors + labels produced by the unmodified
 * reference compiled from /root/reference (oracle/_ref, fixtures under tests/golden/ made by tools/make_golden.py).
 *
 * It follows the reference step by step on the same flat tree that include/fastbn_b200.h describes:
 *   reset          JunctionTree::ResetJunctionTree              src/JunctionTree.cpp:302-311
 *   evidence       LoadDiscreteEvidence + LoadEvidenceToNodes    src/JunctionTree.cpp:316-383, :386-457 (zero-masking
 *                  form; verified equivalent to the compacting form the reference runs)
 *   normalise      PotentialTableBase::Normalize                 src/PotentialTableBase.cpp:433-445
 *   collect        Collect / SeparatorLevelCollectionOptimized / CliqueLevelCollection  src/JunctionTree.cpp:1240-1306, :1056-1148, :829-941
 *   distribute     Distribute / SeparatorLevelDistribution / CliqueLevelDistributionOptimized  :1308-1333, :700-816, :1150-1238
 *   posteriors     GetProbabilitiesOneNode                       src/JunctionTree.cpp:1392-1454
 *   label          CalculateMarginalProbability + Inference::ArgMax  src/JunctionTree.cpp:1339-1380, src/Inference.cpp:92-102
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    int nvars;
    int *vars, *dims, *stride; /* storage order, stride == cum_levels (src/PotentialTableBase.cpp:599-607) */
    int size;
    double *p;                 /* working potentials */
} jto_table;

typedef struct jto_tree {
    int n_vars, n_cliques, n_seps, root, query_var, sum_dims, max_depth;
    int *var_dims, *var_off;
    jto_table *clq, *sep;      /* sep[s].p = current separator table */
    double **pristine;         /* backup of clique tables (BackUpJunctionTree, :289-300) */
    int *sep_parent, *sep_child, *depth;
    int **kids, *nkids;        /* per clique: child separators in child_order */
    double *tmp;               /* scratch, max separator size */
    double *old;
    int max_sep;
} jto_tree;

static void table_init(jto_table *t, int n, const int32_t *vars, const int32_t *var_dims) {
    t->nvars = n;
    t->vars = (int *)malloc(sizeof(int) * n);
    t->dims = (int *)malloc(sizeof(int) * n);
    t->stride = (int *)malloc(sizeof(int) * n);
    int sz = 1;
    for (int i = 0; i < n; ++i) {
        t->vars[i] = vars[i];
        t->dims[i] = var_dims[vars[i]];
    }
    for (int i = n - 1; i >= 0; --i) {
        t->stride[i] = sz;
        sz *= t->dims[i];
    }
    t->size = sz;
    t->p = (double *)malloc(sizeof(double) * sz);
}

static int table_find(const jto_table *t, int v) {
    for (int i = 0; i < t->nvars; ++i)
        if (t->vars[i] == v) return i;
    return -1;
}

jto_tree *jto_create(int n_vars, const int32_t *var_dims, int n_cliques, int n_seps, const int32_t *table_nvars,
                     const int32_t *table_vars, const double *clique_pots, const int32_t *sep_parent,
                     const int32_t *sep_child, int root, int query_var, const int32_t *child_order /* may be NULL */) {
    jto_tree *T = (jto_tree *)calloc(1, sizeof(jto_tree));
    T->n_vars = n_vars; T->n_cliques = n_cliques; T->n_seps = n_seps; T->root = root; T->query_var = query_var;
    T->var_dims = (int *)malloc(sizeof(int) * n_vars);
    T->var_off = (int *)malloc(sizeof(int) * (n_vars + 1));
    T->var_off[0] = 0;
    for (int v = 0; v < n_vars; ++v) {
        T->var_dims[v] = var_dims[v];
        T->var_off[v + 1] = T->var_off[v] + var_dims[v];
    }
    T->sum_dims = T->var_off[n_vars];
    T->clq = (jto_table *)calloc(n_cliques, sizeof(jto_table));
    T->sep = (jto_table *)calloc(n_seps > 0 ? n_seps : 1, sizeof(jto_table));
    T->pristine = (double **)calloc(n_cliques, sizeof(double *));
    int64_t pos = 0, ppos = 0;
    for (int c = 0; c < n_cliques; ++c) {
        table_init(&T->clq[c], table_nvars[c], table_vars + pos, var_dims);
        pos += table_nvars[c];
        T->pristine[c] = (double *)malloc(sizeof(double) * T->clq[c].size);
        memcpy(T->pristine[c], clique_pots + ppos, sizeof(double) * T->clq[c].size);
        ppos += T->clq[c].size;
    }
    T->max_sep = 1;
    for (int s = 0; s < n_seps; ++s) {
        table_init(&T->sep[s], table_nvars[n_cliques + s], table_vars + pos, var_dims);
        pos += table_nvars[n_cliques + s];
        if (T->sep[s].size > T->max_sep) T->max_sep = T->sep[s].size;
    }
    T->tmp = (double *)malloc(sizeof(double) * T->max_sep);
    T->old = (double *)malloc(sizeof(double) * T->max_sep);
    T->sep_parent = (int *)malloc(sizeof(int) * (n_seps > 0 ? n_seps : 1));
    T->sep_child = (int *)malloc(sizeof(int) * (n_seps > 0 ? n_seps : 1));
    T->nkids = (int *)calloc(n_cliques, sizeof(int));
    T->kids = (int **)calloc(n_cliques, sizeof(int *));
    for (int s = 0; s < n_seps; ++s) {
        T->sep_parent[s] = sep_parent[s];
        T->sep_child[s] = sep_child[s];
        T->nkids[sep_parent[s]]++;
    }
    for (int c = 0; c < n_cliques; ++c) T->kids[c] = (int *)malloc(sizeof(int) * (T->nkids[c] > 0 ? T->nkids[c] : 1));
    {   /* children in ptr_downstream_cliques order when given, else separator index order */
        int *fill = (int *)calloc(n_cliques, sizeof(int));
        for (int s = 0; s < n_seps; ++s) {
            int p = sep_parent[s];
            if (child_order) T->kids[p][child_order[s]] = s;
            else T->kids[p][fill[p]++] = s;
        }
        free(fill);
    }
    /* depth of every clique below the root (MarkLevel, :137-181): clique level = 2*depth */
    T->depth = (int *)malloc(sizeof(int) * n_cliques);
    for (int c = 0; c < n_cliques; ++c) T->depth[c] = -1;
    int *queue = (int *)malloc(sizeof(int) * n_cliques), qh = 0, qt = 0;
    queue[qt++] = root;
    T->depth[root] = 0;
    while (qh < qt) {
        int c = queue[qh++];
        for (int k = 0; k < T->nkids[c]; ++k) {
            int ch = T->sep_child[T->kids[c][k]];
            T->depth[ch] = T->depth[c] + 1;
            if (T->depth[ch] > T->max_depth) T->max_depth = T->depth[ch];
            queue[qt++] = ch;
        }
    }
    free(queue);
    return T;
}

static void table_free(jto_table *t) { free(t->vars); free(t->dims); free(t->stride); free(t->p); }

void jto_destroy(jto_tree *T) {
    if (!T) return;
    for (int c = 0; c < T->n_cliques; ++c) { table_free(&T->clq[c]); free(T->pristine[c]); free(T->kids[c]); }
    for (int s = 0; s < T->n_seps; ++s) table_free(&T->sep[s]);
    free(T->clq); free(T->sep); free(T->pristine); free(T->kids); free(T->nkids); free(T->sep_parent);
    free(T->sep_child); free(T->depth); free(T->var_dims); free(T->var_off); free(T->tmp); free(T->old);
    free(T);
}

int jto_sum_dims(const jto_tree *T) { return T->sum_dims; }

/* Normalize: p /= sum, no zero guard (src/PotentialTableBase.cpp:433-445) */
static void normalize(double *p, int n) {
    double s = 0;
    for (int i = 0; i < n; ++i) s += p[i];
    for (int i = 0; i < n; ++i) p[i] /= s;
}

/* LoadEvidenceToNodes (:386-457): entries whose digit of `var` differs from `val` become 0, shape kept */
static void mask_table(jto_table *t, int var, int val) {
    int loc = table_find(t, var);
    if (loc < 0) return;
    int st = t->stride[loc], dm = t->dims[loc];
    for (int i = 0; i < t->size; ++i)
        if ((i / st) % dm != val) t->p[i] = 0;
}

/* entry of `sub` addressed by entry e of `big` (TableExtensionMain / TableMarginalizationMain,
 * src/PotentialTable.cpp:603-613, :457-467): digits by successive div/mod, then dot with sub's cum_levels */
static int map_index(const jto_table *big, const jto_table *sub, const int *loc, int e) {
    int idx = 0;
    for (int l = 0; l < sub->nvars; ++l) idx += ((e / big->stride[loc[l]]) % big->dims[loc[l]]) * sub->stride[l];
    return idx;
}

/* One test case. Evidence pairs are applied in ascending (var, val) order like the reference's
 * DiscreteConfig = set<pair<int,int>> (include/common.h:22). Returns 0. */
int jto_run_case(jto_tree *T, int n_ev, const int32_t *ev_vars, const int32_t *ev_vals, double *post, int32_t *label) {
    const int nc = T->n_cliques, ns = T->n_seps;
    /* reset (ResetJunctionTree :302-311): cliques from backup, separators all-ones */
    for (int c = 0; c < nc; ++c) memcpy(T->clq[c].p, T->pristine[c], sizeof(double) * T->clq[c].size);
    for (int s = 0; s < ns; ++s)
        for (int i = 0; i < T->sep[s].size; ++i) T->sep[s].p[i] = 1.0;

    /* evidence (LoadDiscreteEvidence :316-383): sorted, out-of-range variables ignored (:326-331) */
    int *order = (int *)malloc(sizeof(int) * (n_ev > 0 ? n_ev : 1));
    for (int i = 0; i < n_ev; ++i) order[i] = i;
    for (int i = 1; i < n_ev; ++i) { /* insertion sort by (var,val) */
        int k = order[i], j = i - 1;
        while (j >= 0 && (ev_vars[order[j]] > ev_vars[k] || (ev_vars[order[j]] == ev_vars[k] && ev_vals[order[j]] > ev_vals[k]))) {
            order[j + 1] = order[j];
            --j;
        }
        order[j + 1] = k;
    }
    char *is_ev = (char *)calloc(T->n_vars, 1);
    int *n_ev_in = (int *)calloc(nc, sizeof(int)); /* evidence variables per clique: the compacting form drops them */
    for (int i = 0; i < n_ev; ++i) {
        int var = ev_vars[order[i]], val = ev_vals[order[i]];
        if (var >= T->n_vars || var < 0) continue;
        for (int c = 0; c < nc; ++c) {
            if (table_find(&T->clq[c], var) >= 0 && !is_ev[var]) n_ev_in[c]++;
            mask_table(&T->clq[c], var, val);
        }
        for (int s = 0; s < ns; ++s) mask_table(&T->sep[s], var, val);
        is_ev[var] = 1;
    }
    free(order);

    /* "norm" (:1479-1483) */
    for (int c = 0; c < nc; ++c) normalize(T->clq[c].p, T->clq[c].size);

    int loc[64];
    /* Collect (:1240-1306): for depth d = max..1 first the separators above depth-d cliques, then the cliques at
     * depth d-1 absorb their children in rounds k = 0,1,.. */
    for (int d = T->max_depth; d >= 1; --d) {
        for (int s = 0; s < ns; ++s) { /* SeparatorLevelCollectionOptimized (:1056-1148) */
            int ch = T->sep_child[s];
            if (T->depth[ch] != d) continue;
            jto_table *S = &T->sep[s], *C = &T->clq[ch];
            memcpy(T->old, S->p, sizeof(double) * S->size);
            for (int j = 0; j < S->size; ++j) T->tmp[j] = 0;
            for (int k = 0; k < C->size; ++k) T->tmp[k % S->size] += C->p[k]; /* index rule :1108, accumulate PotentialTable.cpp:469-474 */
            for (int j = 0; j < S->size; ++j) S->p[j] = T->old[j] == 0 ? 0 : T->tmp[j] / T->old[j]; /* :1136-1145 */
        }
        int maxk = 0;
        for (int c = 0; c < nc; ++c)
            if (T->depth[c] == d - 1 && T->nkids[c] > maxk) maxk = T->nkids[c];
        for (int k = 0; k < maxk; ++k)
            for (int c = 0; c < nc; ++c) { /* CliqueLevelCollection (:829-941) */
                if (T->depth[c] != d - 1 || T->nkids[c] <= k) continue;
                jto_table *P = &T->clq[c], *S = &T->sep[T->kids[c][k]];
                for (int l = 0; l < S->nvars; ++l) loc[l] = table_find(P, S->vars[l]);
                for (int e = 0; e < P->size; ++e) P->p[e] *= S->p[map_index(P, S, loc, e)];
                normalize(P->p, P->size);
            }
    }
    /* Distribute (:1308-1333) */
    for (int d = 1; d <= T->max_depth; ++d) {
        for (int s = 0; s < ns; ++s) { /* SeparatorLevelDistribution (:700-816) */
            int ch = T->sep_child[s];
            if (T->depth[ch] != d) continue;
            jto_table *S = &T->sep[s], *P = &T->clq[T->sep_parent[s]];
            memcpy(T->old, S->p, sizeof(double) * S->size);
            for (int j = 0; j < S->size; ++j) T->tmp[j] = 0;
            for (int l = 0; l < S->nvars; ++l) loc[l] = table_find(P, S->vars[l]);
            for (int e = 0; e < P->size; ++e) T->tmp[map_index(P, S, loc, e)] += P->p[e];
            for (int j = 0; j < S->size; ++j) S->p[j] = T->old[j] == 0 ? 0 : T->tmp[j] / T->old[j];
        }
        for (int c = 0; c < nc; ++c) { /* CliqueLevelDistributionOptimized (:1150-1238) */
            if (T->depth[c] != d) continue;
            jto_table *C = &T->clq[c];
            int s = -1;
            for (int q = 0; q < ns; ++q)
                if (T->sep_child[q] == c) s = q;
            jto_table *S = &T->sep[s];
            for (int k = 0; k < C->size; ++k) C->p[k] *= S->p[k % S->size];
            normalize(C->p, C->size);
        }
    }

    /* posteriors (GetProbabilitiesOneNode :1392-1454): smallest clique by number of variables still present after the
     * compacting reduction (= nvars - evidence vars in it), first in container order on ties; label (:1339-1380, ArgMax) */
    double m[256];
    for (int v = 0; v < T->n_vars; ++v) {
        int dim = T->var_dims[v];
        double *out = post + T->var_off[v];
        for (int d = 0; d < dim; ++d) out[d] = 0.0;
        if (is_ev[v]) continue;
        int best = -1, best_nv = INT32_MAX;
        for (int c = 0; c < nc; ++c) {
            int nvv = T->clq[c].nvars - n_ev_in[c];
            if (nvv >= best_nv) continue;
            if (table_find(&T->clq[c], v) < 0) continue;
            best_nv = nvv;
            best = c;
        }
        if (best < 0) { free(is_ev); free(n_ev_in); return 1; }
        jto_table *C = &T->clq[best];
        int lc = table_find(C, v);
        for (int d = 0; d < dim; ++d) m[d] = 0;
        for (int e = 0; e < C->size; ++e) m[(e / C->stride[lc]) % dim] += C->p[e];
        normalize(m, dim);
        for (int d = 0; d < dim; ++d) out[d] = m[d];
    }
    if (label) {
        int v = T->query_var, arg = -1;
        double mx = 0;
        if (!is_ev[v])
            for (int d = 0; d < T->var_dims[v]; ++d)
                if (post[T->var_off[v] + d] > mx) { mx = post[T->var_off[v] + d]; arg = d; }
        *label = arg;
    }
    free(is_ev);
    free(n_ev_in);
    return 0;
}

/* sequential loop over cases (PredictUseJTInfer, src/JunctionTree.cpp:1517-1528) */
int jto_run_batch(jto_tree *T, int64_t n_cases, const int64_t *ev_off, const int32_t *ev_vars, const int32_t *ev_vals,
                  double *post, int32_t *labels) {
    double *tmp = post ? NULL : (double *)malloc(sizeof(double) * T->sum_dims);
    for (int64_t i = 0; i < n_cases; ++i) {
        int32_t lab;
        int rc = jto_run_case(T, (int)(ev_off[i + 1] - ev_off[i]), ev_vars + ev_off[i], ev_vals + ev_off[i],
                              post ? post + i * T->sum_dims : tmp, &lab);
        if (rc) { free(tmp); return rc; }
        if (labels) labels[i] = lab;
    }
    free(tmp);
    return 0;
}
