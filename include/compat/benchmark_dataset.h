// include/compat/benchmark_dataset.h -- GASMA/benchmark/benchmark_dataset.h's header name; class Dataset with the
// reference's constructor, output(path) and output() (host/benchmark_dataset.h).
#ifndef GASMA_COMPAT_BENCHMARK_DATASET_H
#define GASMA_COMPAT_BENCHMARK_DATASET_H
#include "../../approximate-string-matching_b200/host/benchmark_dataset.h"
#endif
