// include/compat/benchmark_utils.h -- the reference's harness header name and class name over libgasma_b200.so:
// GASMA/benchmark/benchmark.cpp:12-32 compiles unmodified with -I include/compat.  `benchmark` is benchmark_gpu
// (host/benchmark_gpu.h: same constructor, read_string_file, read_answer_file, run, print; batched on the device).
#ifndef GASMA_COMPAT_BENCHMARK_UTILS_H
#define GASMA_COMPAT_BENCHMARK_UTILS_H
#include "../../approximate-string-matching_b200/host/benchmark_gpu.h"
using benchmark = benchmark_gpu;   // GASMA/benchmark/benchmark_utils.h:47, ctor :263
#endif
