// include/compat/benchmark/benchmark_coverage.h -- host versions of the two coverage helpers the reference's drivers
// call directly (GASMA/main.cpp:14-17; GASMA/benchmark/benchmark_coverage.h:26-67 and :73-91).  Same names, arguments
// and results; the batched benchmark computes the same metric on the device (csrc/metrics.cuh).
#ifndef GASMA_COMPAT_BENCHMARK_COVERAGE_H
#define GASMA_COMPAT_BENCHMARK_COVERAGE_H

#include <cstdlib>
#include <string>

// Read characters under the M / '=' runs of `CIGAR` that are at least `threshold` long.  X advances both strings,
// I the read, D the reference; s2 is walked but never read (as in the reference).
inline std::string long_consecutive_matching_substring(const char* s1, const char* s2, const std::string& CIGAR, int threshold = 3) {
    (void)s2;
    std::string out;
    size_t pos = 0, i1 = 0;
    while (pos < CIGAR.size()) {
        char* end = nullptr;
        const long len = strtol(CIGAR.c_str() + pos, &end, 10);
        if (end == CIGAR.c_str() + pos || *end == '\0') break;   // no number / no letter: the stream extraction fails
        const char type = *end;
        pos = (size_t)(end - CIGAR.c_str()) + 1;
        if (type == 'X' || type == 'I') i1 += (size_t)len;
        else if (type == '=' || type == 'M') {
            if (len >= threshold) out.append(s1 + i1, (size_t)len);
            i1 += (size_t)len;
        }
    }
    return out;
}

// true iff s2 is a subsequence of s1 (s1 can be built by inserting characters into s2)
inline bool covers(const std::string& s1, const std::string& s2) {
    if (s1.size() < s2.size()) return false;
    size_t i = 0;
    for (char c : s2) {
        while (i < s1.size() && s1[i] != c) ++i;
        if (i >= s1.size()) return false;
        ++i;
    }
    return true;
}

#endif
