// BayesianNetwork — CLI drop-in for the reference's `-a 2` path (src/main.cpp:61-83, src/Parameter.cpp:30-66).
// Same flags (-a -t -f0 -f3 -f4; the others are parsed and ignored), same default files, paths relative to
// "../dataset/", same stdout lines. Algorithms other than 2 are outside this build and say so.
// Extensions: -G <n> shards the test cases over n GPUs (or FASTBN_DEVICES=0,1,...); -o <file> writes the posteriors
// in the reference's potential-table format (relative to ../dataset/ too); -D <dir> replaces the "../dataset/" prefix;
// -C keeps a binary cache of the parsed test cases beside the -f3 file (<file>.fbncache, validated by size + mtime).
// -t N is accepted like in the reference; the GPU path has no use for it (the reference's OpenMP parallelism is INSIDE one
// case). The CPU baseline for a given -t is a separate tool that drives the unmodified reference: tools/ref_cpu_baseline.py.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <string>
#include <thread>

#include "../fastbn_host.h"

using namespace std;

int main(int argc, char **argv) {
    int algorithm = 2, num_threads = 1, n_gpus = 0;
    string net_file = "alarm/alarm.xml", test_set_file = "alarm/testing_alarm_1k_p20", pt_file = "alarm/alarm_1k_pt";
    string out_pt, prefix = "../dataset/";
    int i;
    for (i = 1; (i < argc) && ((argv[i])[0] == '-'); i++) {
        auto need = [&](const char *) -> const char * {
            if (i + 1 >= argc) { printf("\n Option %s needs a value!\n", argv[i]); exit(0); }
            return argv[++i];
        };
        switch (argv[i][1]) {
            case 'h':
                cout << "Usage: ./BayesianNetwork -a 2 [-t threads] [-f0 bn.xml] [-f3 test cases] [-f4 potential tables] [-G gpus] [-o out_pt] [-D dataset dir] [-C]" << endl;
                exit(0);
            case 'a': algorithm = atoi(need("a")); break;
            case 't': num_threads = atoi(need("t")); break;
            case 'q': case 'm': case 'l': case 'd': case 'g': need("x"); break;  // sampling / PC-stable options: ignored
            case 'G': n_gpus = atoi(need("G")); break;
            case 'C': setenv("FASTBN_EVIDENCE_CACHE", "1", 1); break;
            case 'o': out_pt = need("o"); break;
            case 'D': prefix = need("D"); if (!prefix.empty() && prefix.back() != '/') prefix += '/'; break;
            case 'f':
                switch (argv[i][2]) {
                    case '0': net_file = need("f0"); break;
                    case '3': test_set_file = need("f3"); break;
                    case '4': pt_file = need("f4"); break;
                    case '1': case '2': need("f"); break;
                }
                break;
            default:
                printf("\n Unrecognized option %s!\n", argv[i]);
                exit(0);
        }
    }
    if (algorithm < 0 || algorithm > 11) {
        cout << "\tError! Please give the right value of -a to specify the functionality/algorithm" << endl;
        exit(0);
    }
    net_file = prefix + net_file;
    test_set_file = prefix + test_set_file;
    pt_file = prefix + pt_file;
    if (algorithm != 2) {
        cout << "This build (fastbn_b200) only provides -a 2: junction tree (JT) for exact inference" << endl;
        return 0;
    }
    cout << "===============================" << endl;
    cout << "Algorithm: junction tree (JT) for exact inference, #threads = " << num_threads << endl;
    cout << "\tBN: " << net_file << endl;
    cout << "\ttesting set: " << test_set_file << endl;
    cout << "\treference potential table: " << pt_file << endl;
    cout << "===============================" << endl;

    // FASTBN_TIMING=1: stage times on stderr (stdout stays the reference's)
    const bool timing = getenv("FASTBN_TIMING") != nullptr;
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double t0 = now();
    // only the GPUs this run uses are made visible (CUDA initialisation costs ~0.6 s per visible GPU on an 8-GPU box);
    // an existing CUDA_VISIBLE_DEVICES or FASTBN_DEVICES choice is respected
    if (!getenv("CUDA_VISIBLE_DEVICES") && !getenv("FASTBN_DEVICES")) {
        string vis;
        for (int d = 0; d < (n_gpus > 0 ? n_gpus : 1); ++d) vis += (d ? "," : "") + std::to_string(d);
        setenv("CUDA_VISIBLE_DEVICES", vis.c_str(), 0);
    }
    // CUDA contexts come up in the background while the input files are parsed. The thread is joined before the process ends,
    // also on the loaders' error exits (exit(1) on a missing file or a malformed line runs the atexit hook).
    static std::thread warm;
    warm = std::thread([n_gpus] { for (int d = 0; d < (n_gpus > 0 ? n_gpus : 1); ++d) fbn_device_warmup(d); });
    atexit([] { if (warm.joinable()) warm.join(); });
    Dataset *tester = new Dataset();
    CustomNetwork *network = new CustomNetwork(true);
    network->GetNetFromXMLBIFFile(net_file);
    double t1 = now();
    tester->LoadLIBSVMDataKnownNetwork(test_set_file, network->num_nodes);
    double t2 = now();

    JunctionTree *jt = new JunctionTree(network, tester, false);
    if (n_gpus > 0) {
        jt->devices.clear();
        for (int d = 0; d < n_gpus; ++d) jt->devices.push_back(d);
    }
    jt->keep_posteriors = !out_pt.empty();
    Inference *inference = jt;
    double t3 = now();
    double accuracy = inference->EvaluateAccuracy(pt_file, num_threads);
    double t4 = now();
    cout << "accuracy = " << accuracy << endl;
    if (!out_pt.empty()) jt->WritePotentialTableFile(prefix + out_pt);
    double t5 = now();
    if (timing)
        fprintf(stderr, "[timing] xmlbif %.3f s, test cases %.3f s, junction tree %.3f s, potential-table file %.3f s, jt (GPU, host buffers) %.3f s, "
                        "write -o %.3f s, total %.3f s\n", t1 - t0, t2 - t1, t3 - t2, (t4 - t3) - jt->jt_seconds, jt->jt_seconds, t5 - t4, t5 - t0);

    delete inference;
    delete network;
    delete tester;
    if (warm.joinable()) warm.join();
    return 0;
}
