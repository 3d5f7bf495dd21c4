// include/compat/hurdle_matrix.h -- the reference's header name and class names over libgasma_b200.so.
//
// With `-I include/compat` the reference's own drivers compile UNMODIFIED against the B200 library:
//   GASMA/main.cpp:5-19            new hurdle_matrix<int_128bit>(read, ref, 3, GLOBAL, 1, 1, 1); print(); run();
//                                  get_CIGAR(); get_cost(); long_consecutive_matching_substring(...)
//   GASMA/mapper/main.cpp:71-95    hurdle_matrix<int_128bit> per hit: reset(); run(); get_cost(); get_CIGAR()
// This header stands in for GASMA/hurdle_matrix.h (:552-679) and the parts of GASMA/utils.h its callers name
// (alignment_type_t :554-558, MAX_LENGTH :23-25, the int_128bit type :49-271 as a debug value type).
// All alignment work goes through the C ABI to the CUDA kernels (host/hurdle_matrix_gpu.h); the only host arithmetic
// here is the DEBUG view behind print() / operator[] (hurdle_matrix.h:602-607, :458-460): the hurdle lanes rebuilt
// from the 2-bit planes gasma_pack2bit_batch returns.
#ifndef GASMA_COMPAT_HURDLE_MATRIX_H
#define GASMA_COMPAT_HURDLE_MATRIX_H

#include <cstdint>
#include <cstdio>
#include <map>
#include <string>

#ifndef MAX_LENGTH
#define MAX_LENGTH 128   // GASMA/utils.h:23-25
#endif
#ifndef MAX_K
#define MAX_K 50         // GASMA/hurdle_matrix.h:8
#endif

#include "../../approximate-string-matching_b200/host/hurdle_matrix_gpu.h"

// Value type of a lane as callers of operator[] / print() see it (GASMA/utils.h:49-271): 128 bits, bit i = position i.
// A plain value here -- the aligner itself never touches it.
class int_128bit {
public:
    uint64_t w[2];
    int_128bit() : w{1ull, 0ull} {}   // utils.h:57-59: the default value is 1
    int_128bit(uint64_t lo, uint64_t hi) : w{lo, hi} {}
    void print() const {              // utils.h:78-82: bytes in order, bits LSB first == bit 0 .. bit 127
        for (int i = 0; i < 128; ++i) putchar(((w[i >> 6] >> (i & 63)) & 1ull) ? '1' : '0');
        putchar('\n');
    }
    static constexpr int bits = 128;
};
// int_256bit (utils.h:274-549, README "Using AVX2") is NOT offered: in the reference it reads uninitialised stack for
// its upper half (hurdle_matrix.h:262-275) and drops carries at bit 128, so it has no defined result to reproduce.

template <class T>
class hurdle_matrix : public hurdle_matrix_gpu {
    static_assert(T::bits == MAX_LENGTH, "this build of the compat header pairs int_128bit with MAX_LENGTH 128");

public:
    // hurdle_matrix.h:552-562
    explicit hurdle_matrix(alignment_type_t type = GLOBAL, int x = 1, int o = 1, int e = 1, double match_prob = 0.80,
                           double mismatch_prob = 0.20 / 3, double indel_prob = 0.40 / 3)
        : hurdle_matrix_gpu(type, x, o, e, match_prob, mismatch_prob, indel_prob, MAX_LENGTH) {}
    // hurdle_matrix.h:473-484
    hurdle_matrix(const char* read, const char* ref, int error, alignment_type_t type = GLOBAL, int x = 1, int o = 1, int e = 1,
                  double match_prob = 0.95, double mismatch_prob = 0.02, double indel_prob = 0.03)
        : hurdle_matrix_gpu(type, x, o, e, match_prob, mismatch_prob, indel_prob, MAX_LENGTH) {
        reset(read, ref, error);
    }

    void reset(const char* read, const int read_len, const char* ref, const int ref_len, int error) {
        hurdle_matrix_gpu::reset(read, read_len, ref, ref_len, error);
        lanes_.clear();
    }
    void reset(const char* read, const char* ref, int error) {
        hurdle_matrix_gpu::reset(read, ref, error);
        lanes_.clear();
    }

    // hurdle_matrix.h:458-460: the (short-hurdle-flipped) lane; lanes outside the band hold the default value 1
    T& operator[](int lane) {
        build_lanes();
        return lanes_[lane];
    }
    // hurdle_matrix.h:602-607
    void print() {
        build_lanes();
        for (int i = lower_bound(); i <= upper_bound(); ++i) {
            printf("lane %d:", i);
            lanes_[i].print();
        }
    }

private:
    // debug only: _construct_hurdles (hurdle_matrix.h:441-455) on the planes the device packed (bit_convert.cpp:248)
    void build_lanes() {
        if (!lanes_.empty()) return;
        uint64_t a0[2], a1[2], b0[2], b1[2];
        planes(a0, a1, b0, b1);
        for (int l = -MAX_K; l <= MAX_K; ++l) lanes_[l] = T();
        for (int l = lower_bound(); l <= upper_bound(); ++l) {
            uint64_t s0[2], s1[2], v[2];
            const int a = l < 0 ? -l : l;
            shr(l < 0 ? a0 : b0, a, s0);
            shr(l < 0 ? a1 : b1, a, s1);
            for (int j = 0; j < 2; ++j) v[j] = (s0[j] ^ (l < 0 ? b0 : a0)[j]) | (s1[j] ^ (l < 0 ? b1 : a1)[j]);
            // flip_short_hurdles(1), utils.h:200-216
            const uint64_t up[2] = {(v[0] >> 1) | (v[1] << 63), v[1] >> 1}, dn[2] = {v[0] << 1, (v[1] << 1) | (v[0] >> 63)};
            lanes_[l] = T(v[0] & (up[0] | dn[0]), v[1] & (up[1] | dn[1]));
        }
    }
    static void shr(const uint64_t x[2], int s, uint64_t out[2]) {   // the reference's "shift_left": towards bit 0
        if (s == 0) { out[0] = x[0]; out[1] = x[1]; }
        else if (s < 64) { out[0] = (x[0] >> s) | (x[1] << (64 - s)); out[1] = x[1] >> s; }
        else { out[0] = s < 128 ? x[1] >> (s - 64) : 0; out[1] = 0; }
    }
    std::map<int, T> lanes_;
};

#endif
