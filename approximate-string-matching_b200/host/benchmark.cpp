// benchmark.cpp -- main() of the B200 benchmark driver, the counterpart of GASMA/benchmark/benchmark.cpp:12-32
// (CMake target hurdle-matrix-benchmark): for each error rate generate (or reuse) a simulated data file, then
// benchmark bench(1,1,1,3,N,true); read_string_file; run; print.  Host code stays C++; all compute happens in
// libgasma_b200.so through the C ABI.
//
//   gasma-benchmark                          simulated data, error rates 0.05 0.10 0.15 0.20, 1 M pairs each
//   gasma-benchmark --pairs N --length L     same with other sizes
//   gasma-benchmark --gpus G                 batches dealt to G GPUs of the box (devices 0..G-1), counters summed
//   gasma-benchmark --devices 0,2,3          the same with an explicit device list (one host thread + context per entry)
//   gasma-benchmark --dataset-only ...       only write the simulated_*.seq files (no GPU needed)
//   gasma-benchmark FILE.seq [ANSWERS]       a real data file in the reference's '>read / <ref' format
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <string>
#include <vector>

#include "benchmark_dataset.h"
#include "benchmark_gpu.h"

int main(int argc, char** argv) {
    int pairs = 1000000, length = 100, k = 3, device = 0, gpus = 1;
    long batch = 1000000;
    bool dataset_only = false;
    const char* file = nullptr;
    const char* answers = nullptr;
    std::vector<int> devices;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--pairs") && i + 1 < argc) pairs = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--length") && i + 1 < argc) length = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--k") && i + 1 < argc) k = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--device") && i + 1 < argc) device = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--gpus") && i + 1 < argc) gpus = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--batch") && i + 1 < argc) batch = atol(argv[++i]);
        else if (!strcmp(argv[i], "--devices") && i + 1 < argc) {
            for (const char* p = argv[++i]; *p;) {
                devices.push_back((int)strtol(p, const_cast<char**>(&p), 10));
                if (*p == ',') ++p; else if (*p) break;
            }
        }
        else if (!strcmp(argv[i], "--dataset-only")) dataset_only = true;   // write the simulated files and stop
        else if (!file) file = argv[i];
        else answers = argv[i];
    }
    if (devices.empty())
        for (int g = 0; g < (gpus < 1 ? 1 : gpus); ++g) devices.push_back(device + g);
    if (batch < 1) batch = 1000000;
    try {
        if (!file) {
            const std::vector<float> error_rates = {0.05f, 0.10f, 0.15f, 0.20f};
            for (float error_rate : error_rates) {
                Dataset dataset(pairs, length, error_rate, 0.96f, true);
                const std::string path = dataset.output();
                if (dataset_only) { printf("%s\n", path.c_str()); continue; }
                benchmark_gpu bench(1, 1, 1, k, pairs, true, devices, batch);
                bench.read_string_file(path.c_str());
                bench.run();
                bench.print();
            }
        } else {
            benchmark_gpu bench(1, 1, 1, k, pairs, true, devices, batch);
            bench.read_string_file(file);
            if (answers) bench.read_answer_file(answers);
            bench.run();
            bench.print();
        }
    } catch (const std::exception& ex) {
        fprintf(stderr, "gasma-benchmark: %s\n", ex.what());
        return 1;
    }
    return 0;
}
