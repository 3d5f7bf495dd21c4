// benchmark_dataset.h -- simulated (read, ref) pair files for the benchmark driver.
//
// Same interface and file format as the reference's Dataset (GASMA/benchmark/benchmark_dataset.h:61-254, itself
// adapted from WFA's generate_dataset.c): Dataset(num_reads, length, error_rate, mismatch_rate, exact, overwrite),
// output() -> "simulated_<n>_<len>_<rate>_{eq,lt_eq}.seq" holding ">read\n<ref\n" per pair.  The edit model is the
// reference's (libc rand(): a uniform draw by rejection, mismatch with probability mismatch_rate else delete or
// insert 50/50, edits applied one after another to the text); the seed is a constructor argument instead of
// time(), so that a file can be regenerated.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>

class Dataset {
public:
    Dataset(int num_reads, int length, float error_rate, float mismatch_rate, bool exact_error_rate = false,
            bool overwrite = false, unsigned seed = 0x47415341u)
        : n_(num_reads), len_(length), err_(error_rate), mm_(mismatch_rate), exact_(exact_error_rate), overwrite_(overwrite),
          seed_(seed) {
        if (err_ < 0 || err_ > 0.7f) { fprintf(stderr, "error rate %f outside [0, 0.7]\n", err_); exit(1); }
        if (mm_ < 0 || mm_ > 1) { fprintf(stderr, "mismatch rate %f outside [0, 1]\n", mm_); exit(1); }
    }

    void output(const char* path) {
        if (!overwrite_) {
            if (FILE* f = fopen(path, "r")) { fclose(f); return; }
        }
        FILE* out = fopen(path, "w");
        if (!out) { fprintf(stderr, "cannot write %s\n", path); exit(1); }
        srand(seed_);
        std::string pattern((size_t)len_, 'A'), text;
        for (int i = 0; i < n_; ++i) {
            for (int j = 0; j < len_; ++j) pattern[(size_t)j] = kAlphabet[uniform(0, 4)];
            fprintf(out, ">%s\n", pattern.c_str());
            mutate(pattern, text);
            fprintf(out, "<%s\n", text.c_str());
        }
        fclose(out);
    }

    std::string output() {
        std::string name = "simulated_" + std::to_string(n_) + "_" + std::to_string(len_) + "_" + std::to_string(err_) + "_" +
                           (exact_ ? "eq.seq" : "lt_eq.seq");
        output(name.c_str());
        return name;
    }

private:
    static constexpr const char* kAlphabet = "ACGT";

    // uniform integer in [lo, hi) from rand(), rejecting the remainder interval (benchmark_dataset.h:85-96)
    static uint64_t uniform(uint64_t lo, uint64_t hi) {
        const uint64_t range = hi - lo, rem = RAND_MAX % range, bucket = RAND_MAX / range;
        for (;;) {
            const int r = rand();
            if ((uint64_t)r < RAND_MAX - rem) return lo + (uint64_t)r / bucket;
        }
    }

    // benchmark_dataset.h:147-186
    void mutate(const std::string& pattern, std::string& text) const {
        const uint64_t nominal = (uint64_t)std::ceil(len_ * err_);
        const uint64_t edits = exact_ ? nominal : uniform(0, nominal);
        text = pattern;
        for (uint64_t i = 0; i < edits; ++i) {
            const float u = (float)rand() / RAND_MAX;
            const int kind = u <= mm_ ? 0 : (int)uniform(1, 3);
            const size_t pos = (size_t)uniform(0, text.size());
            if (kind == 0) text[pos] = kAlphabet[uniform(0, 4)];
            else if (kind == 1) text.erase(pos, 1);
            else text.insert(pos, 1, kAlphabet[uniform(0, 4)]);
        }
    }

    int n_, len_;
    float err_, mm_;
    bool exact_, overwrite_;
    unsigned seed_;
};
