/*
 * fastbn_b200 — C access to the host-side mirror of the reference's front end (fastbn_b200/host/): XMLBIF network ->
 * junction tree (CustomNetwork::GetNetFromXMLBIFFile + JunctionTree ctor, reference src/main.cpp:71-77) flattened to the
 * fbn_tree_desc of fastbn_b200.h; LIBSVM test cases (Dataset::LoadLIBSVMDataKnownNetwork, src/Dataset.cpp:162-262) as CSR;
 * JunctionTree::EvaluateAccuracy (src/JunctionTree.cpp:57-129) on one or more GPUs. Used by the Python harness; C++ callers
 * use the classes in fastbn_b200/host/fastbn_host.h directly. Errors in input files follow the reference's convention
 * (message on stderr, exit(1)).
 */
#ifndef FASTBN_B200_HOST_H
#define FASTBN_B200_HOST_H

#include "fastbn_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fbn_host_model fbn_host_model; /* CustomNetwork + Dataset + JunctionTree */

/* libsvm_path may be NULL/"" (no test cases). quiet != 0 suppresses the reference-style progress lines on stdout. */
int fbn_host_model_create(const char *xmlbif_path, const char *libsvm_path, int quiet, fbn_host_model **out);
void fbn_host_model_destroy(fbn_host_model *m);
/* the flat tree; arrays are owned by the model and live until fbn_host_model_destroy */
const fbn_tree_desc *fbn_host_model_tree(const fbn_host_model *m);
/* out8 = {n_vars, n_cliques, n_separators, sum(table_nvars), sum |clique|, max_level, n_cases, max |clique|} */
int fbn_host_model_info(const fbn_host_model *m, int64_t *out8);
/* CSR evidence + class labels of the loaded test cases (owned by the model) */
int fbn_host_model_cases(const fbn_host_model *m, const int64_t **ev_offsets, const int32_t **ev_vars,
                         const int32_t **ev_vals, const int32_t **truths);
/* Inference::LoadGroundTruthProbabilityTable (src/Inference.cpp:108-146): golden_out = n_cases x sum_dims, first entry of
 * an evidence node's block is -1 */
int fbn_host_model_load_golden(fbn_host_model *m, const char *pt_path, double *golden_out);
/* EvaluateAccuracy(pt_path) over `devices` (NULL: device 0). out5 = {accuracy, average MSE, average HD, jt seconds,
 * kernel ms}. labels_out [n_cases] and posteriors_out [n_cases x sum_dims] may be NULL. */
int fbn_host_model_evaluate(fbn_host_model *m, const char *pt_path, const int32_t *devices, int n_devices,
                            int32_t *labels_out, double *posteriors_out, double *out5);
/* posteriors of the last evaluate (posteriors_out must have been non-NULL) in the reference's pt-file format */
int fbn_host_model_write_pt(fbn_host_model *m, const char *path);

/* Test hooks of the exact text fast paths used for the potential-table files (fastbn_b200/host/textio.cpp):
 * "%.7f" formatting (out352: at least 352 bytes, returns the length, no terminator) and decimal parsing (returns the
 * number of characters consumed, like strtod's end pointer). */
int fbn_host_format_fixed(double x, char *out352);
int fbn_host_parse_double(const char *s, int len, double *out);

#ifdef __cplusplus
}
#endif
#endif
