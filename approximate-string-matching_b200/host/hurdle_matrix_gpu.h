// hurdle_matrix_gpu.h -- the reference's aligner interface on top of libgasma_b200.so.
//
// Drop-in for `hurdle_matrix<int_128bit>` (GASMA/hurdle_matrix.h:552-679) for callers that align one pair at a
// time (benchmark::_run_greedy, benchmark_utils.h:185-201; mapper/main.cpp:71-95): same constructor arguments,
// same method names, same argument meaning.  Every call goes through the C ABI (include/gasma_b200.h) to the
// CUDA kernels; there is no CPU path in here.  One pair per launch wastes a B200 -- callers that can should use
// align_batch() (or benchmark_gpu), which is what the reset/run pair amounts to for a whole batch.
//
// Differences from the reference, all at the boundary:
//   * results are those of the reference in canonical mode (staging tail = 'A'), i.e. they do not
//     depend on the pairs aligned before;
//   * |ref_len - read_len| > MAX_K (out-of-bounds in the reference) and CUDA failures throw std::runtime_error;
//   * MAX_LENGTH is a constructor argument (128 = int_128bit; 192 and 256 are the multi-word extension);
//   * the reference's compile-time band variant `#ifdef CORRECTION` (hurdle_matrix.h:501-512, :634-641) is honoured the
//     same way: compile the caller with -DCORRECTION and the context runs with the widened band, keeping lane slots for
//     |ref_len - read_len| <= GASMA_CORRECTION_MAX_DELTA (default 16; shared memory per alignment grows with it).
#pragma once
#include <algorithm>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/gasma_b200.h"

#ifndef GASMA_CORRECTION_MAX_DELTA
#define GASMA_CORRECTION_MAX_DELTA 16
#endif

#ifndef GASMA_ALIGNMENT_TYPE_T
#define GASMA_ALIGNMENT_TYPE_T
enum alignment_type_t { GLOBAL, SEMI_GLOBAL, LOCAL };   // GASMA/utils.h:554-558
#endif

class hurdle_matrix_gpu {
public:
    // hurdle_matrix(alignment_type, x, o, e, match_prob, mismatch_prob, indel_prob), hurdle_matrix.h:552-562
    explicit hurdle_matrix_gpu(alignment_type_t type = GLOBAL, int x = 1, int o = 1, int e = 1, double match_prob = 0.80,
                               double mismatch_prob = 0.20 / 3, double indel_prob = 0.40 / 3, int max_length = 128,
                               int device = 0)
        : max_length_(max_length) {
        gasma_params p;
        gasma_default_params(&p);
        if (type == LOCAL) throw std::runtime_error("hurdle_matrix_gpu: LOCAL is not supported (nor by the reference)");
        p.alignment_type = type == GLOBAL ? GASMA_GLOBAL : GASMA_SEMI_GLOBAL;
        p.x = x; p.o = o; p.e = e;
        p.max_length = max_length;
        p.match_prob = match_prob; p.mismatch_prob = mismatch_prob; p.indel_prob = indel_prob;
        k_ = p.k;
        if (gasma_create(&p, device, &ctx_) != GASMA_OK)
            throw std::runtime_error(std::string("hurdle_matrix_gpu: ") + gasma_last_error(nullptr));
#ifdef CORRECTION
        check(gasma_set_correction(ctx_, 1, GASMA_CORRECTION_MAX_DELTA));
#endif
    }
    // hurdle_matrix(read, ref, error, alignment_type, x, o, e, 0.95, 0.02, 0.03), hurdle_matrix.h:473-484
    hurdle_matrix_gpu(const char* read, const char* ref, int error, alignment_type_t type = GLOBAL, int x = 1, int o = 1,
                      int e = 1, double match_prob = 0.95, double mismatch_prob = 0.02, double indel_prob = 0.03)
        : hurdle_matrix_gpu(type, x, o, e, match_prob, mismatch_prob, indel_prob) {
        reset(read, ref, error);
    }
    hurdle_matrix_gpu(const hurdle_matrix_gpu&) = delete;
    hurdle_matrix_gpu& operator=(const hurdle_matrix_gpu&) = delete;
    ~hurdle_matrix_gpu() { gasma_destroy(ctx_); }

    // reset(read, read_len, ref, ref_len, error), hurdle_matrix.h:625-665: stages the pair (silently truncated to
    // MAX_LENGTH like :626-627) and the band; the work happens in run().
    void reset(const char* read, const int read_len, const char* ref, const int ref_len, int error) {
        const int m = std::min(max_length_, read_len), n = std::min(max_length_, ref_len);
        read_.assign(read, read + m);
        ref_.assign(ref, ref + n);
        if (error != k_) {
            check(gasma_set_band(ctx_, error));
            k_ = error;
        }
        cost_ = 0;
        cigar_.clear();
    }
    void reset(const char* read, const char* ref, int error) {   // hurdle_matrix.h:667-671
        reset(read, (int)strlen(read), ref, (int)strlen(ref), error);
    }

    // run(), hurdle_matrix.h:568-597
    void run() {
        const uint32_t roff[2] = {0u, (uint32_t)read_.size()}, foff[2] = {0u, (uint32_t)ref_.size()};
        std::vector<gasma_cigar_op> ops((size_t)(2 * max_length_ + 8));   // two ops per step at most
        uint32_t op_off[2] = {0, 0};
        int32_t cost = 0;
        check(gasma_greedy_batch(ctx_, read_.data(), roff, ref_.data(), foff, 1, &cost, ops.data(), (int64_t)ops.size(), op_off));
        cost_ = cost;
        cigar_ = render(ops.data(), (int)op_off[1]);
    }

    int get_cost() const { return cost_; }          // hurdle_matrix.h:677
    std::string get_CIGAR() { return cigar_; }      // hurdle_matrix.h:613-615

    // The batched form of {reset; run; get_cost; get_CIGAR} for many pairs at once (one launch sequence).
    void align_batch(const std::vector<std::string>& reads, const std::vector<std::string>& refs, int error,
                     std::vector<int>& costs, std::vector<std::string>* cigars = nullptr) {
        if (reads.size() != refs.size()) throw std::runtime_error("align_batch: reads/refs size mismatch");
        const int64_t n = (int64_t)reads.size();
        if (error != k_) { check(gasma_set_band(ctx_, error)); k_ = error; }
        std::vector<uint32_t> roff(n + 1, 0), foff(n + 1, 0);
        std::string rbuf, fbuf;
        for (int64_t i = 0; i < n; ++i) {
            rbuf.append(reads[i], 0, (size_t)max_length_);
            fbuf.append(refs[i], 0, (size_t)max_length_);
            roff[i + 1] = (uint32_t)rbuf.size();
            foff[i + 1] = (uint32_t)fbuf.size();
        }
        std::vector<int32_t> c((size_t)n);
        std::vector<gasma_cigar_op> ops(cigars ? (size_t)(32 * n + 64) : 0);
        std::vector<uint32_t> op_off(cigars ? (size_t)n + 1 : 0);
        int rc = gasma_greedy_batch(ctx_, rbuf.data(), roff.data(), fbuf.data(), foff.data(), n, c.data(),
                                    cigars ? ops.data() : nullptr, (int64_t)ops.size(), cigars ? op_off.data() : nullptr);
        if (rc == GASMA_E_CAPACITY) {   // op_off[n] holds the size needed
            ops.resize(op_off[n]);
            rc = gasma_greedy_batch(ctx_, rbuf.data(), roff.data(), fbuf.data(), foff.data(), n, c.data(), ops.data(),
                                    (int64_t)ops.size(), op_off.data());
        }
        check(rc);
        costs.assign(c.begin(), c.end());
        if (cigars) {
            cigars->resize((size_t)n);
            for (int64_t i = 0; i < n; ++i) (*cigars)[i] = render(ops.data() + op_off[i], (int)(op_off[i + 1] - op_off[i]));
        }
    }

    gasma_ctx* context() { return ctx_; }

    // lane range of the staged pair (hurdle_matrix.h:634-644): -k..k, or the widened band of a -DCORRECTION build
    int lower_bound() const {
#ifdef CORRECTION
        if (read_.size() > ref_.size()) return (int)ref_.size() - (int)read_.size() - k_;
#endif
        return -k_;
    }
    int upper_bound() const {
#ifdef CORRECTION
        if (read_.size() <= ref_.size()) return (int)ref_.size() - (int)read_.size() + k_;
#endif
        return k_;
    }

protected:
    // 2-bit planes of the staged pair as the device packs them (sse3_convert2bit1, bit_convert.cpp:248): W words each.
    // For debug views only (include/compat/hurdle_matrix.h: print(), operator[]).
    void planes(uint64_t* a0, uint64_t* a1, uint64_t* b0, uint64_t* b1) {
        const int W = max_length_ / 64;
        std::string both = read_ + ref_;
        const uint32_t off[3] = {0u, (uint32_t)read_.size(), (uint32_t)both.size()};
        std::vector<uint64_t> p0((size_t)2 * W), p1((size_t)2 * W);
        check(gasma_pack2bit_batch(ctx_, both.data(), off, 2, p0.data(), p1.data()));
        for (int j = 0; j < W; ++j) { a0[j] = p0[j]; a1[j] = p1[j]; b0[j] = p0[W + j]; b1[j] = p1[W + j]; }
    }

private:
    static std::string render(const gasma_cigar_op* ops, int n) {
        std::string s((size_t)gasma_render_cigar(ops, n, nullptr, 0), '\0');
        if (!s.empty()) {
            std::vector<char> buf(s.size() + 1);
            gasma_render_cigar(ops, n, buf.data(), (int)buf.size());
            s.assign(buf.data(), s.size());
        }
        return s;
    }
    void check(int rc) {
        if (rc != GASMA_OK) throw std::runtime_error(std::string("hurdle_matrix_gpu: ") + gasma_last_error(ctx_));
    }

    gasma_ctx* ctx_ = nullptr;
    int max_length_, k_ = 3;
    std::string read_, ref_, cigar_;
    int cost_ = 0;
};
