// mapper_gpu.h -- the alignment stage of the reference's read mapper on top of libgasma_b200.so.
//
// Counterpart of the hit loop of map_reads() in GASMA/mapper/main.cpp:71-97: the reference takes every hit the
// FM-index search reports for a read, cuts the window
//     start = hit_begin ? hit_begin - 1 : 0;   text_view = reference[start, start + |query| + 1)      (:75-76)
// aligns it with ONE hurdle_matrix<int_128bit>(GLOBAL,1,1,1) object, band 3 (:71, :78-82), and writes a SAM record with
// ref_offset = start + 2 and mapq = 60 + get_cost() (:85-92).  One pair at a time wastes a B200, so here the hits of a
// whole chunk of reads are COLLECTED (add_hit), aligned as one batch through the C ABI (flush -> gasma_greedy_batch)
// and then written out in hit order.  All compute is on the GPU; there is no CPU path in here.
//
// Differences from the reference, all at the boundary:
//   * the window is clipped at the end of the reference sequence (the reference builds a span past it, :76);
//   * the CIGAR column carries the greedy aligner's CIGAR (I/D/M ops merged into SAM form); the reference writes the
//     ungapped pair {query, text_view} and carries a FIXME for exactly this (:84, :89);
//   * the search stage (SeqAn3 bi_fm_index, absent from this image) is not part of this header: any front end that
//     yields (read, reference id, hit begin) triples can feed add_hit -- mapper_index.h is a small self-contained one,
//     and a SeqAn3 front end is the reference's own loop with `matrix->reset/run` replaced by add_hit.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/gasma_b200.h"

struct mapper_record {
    size_t read_index;      // index into the caller's read table
    size_t ref_id;          // reference sequence
    size_t window_start;    // `start` of mapper/main.cpp:75
    size_t window_len;
    int cost = 0;           // get_cost()
    std::string cigar;      // SAM CIGAR built from the greedy ops
    bool rejected = false;  // |window - read| > MAX_K after clipping at the contig end: not aligned, written as unmapped
};

class mapper_gpu {
public:
    // the aligner the reference constructs at mapper/main.cpp:71 and the band it passes at :82
    explicit mapper_gpu(int band = 3, int device = 0, int max_length = 128) : max_length_(max_length) {
        gasma_params p;
        gasma_default_params(&p);
        p.k = band;
        p.max_length = max_length;
        if (gasma_create(&p, device, &ctx_) != GASMA_OK)
            throw std::runtime_error(std::string("mapper_gpu: ") + gasma_last_error(nullptr));
    }
    mapper_gpu(const mapper_gpu&) = delete;
    mapper_gpu& operator=(const mapper_gpu&) = delete;
    ~mapper_gpu() { gasma_destroy(ctx_); }

    // One search hit (mapper/main.cpp:73-82).  `reference` is the whole reference sequence the hit lies in.
    void add_hit(size_t read_index, const std::string& query, size_t ref_id, const std::string& reference, size_t hit_begin) {
        const size_t start = hit_begin ? hit_begin - 1 : 0;                                  // :75
        size_t len = query.size() + 1;                                                       // :76
        if (start >= reference.size()) len = 0;
        else if (start + len > reference.size()) len = reference.size() - start;
        mapper_record r;
        r.read_index = read_index; r.ref_id = ref_id; r.window_start = start; r.window_len = len;
        // hurdle_matrix::reset cuts both strings at MAX_LENGTH (hurdle_matrix.h:626-627)
        const size_t m = std::min(query.size(), (size_t)max_length_), w = std::min(len, (size_t)max_length_);
        // A window clipped at the end of its contig can differ from the read by more than MAX_K bases: out of bounds in
        // the reference, GASMA_E_BAND for the WHOLE batch here.  Such a hit is not aligned (it travels through the batch
        // as an empty pair) and comes back as an unmapped record instead of aborting the chunk.
        r.rejected = (m > w ? m - w : w - m) > (size_t)GASMA_MAX_K;
        pending_.push_back(r);
        if (!r.rejected) {
            reads_.append(query, 0, m);
            refs_.append(reference, start, w);
        }
        read_off_.push_back((uint32_t)reads_.size());
        ref_off_.push_back((uint32_t)refs_.size());
    }
    size_t pending() const { return pending_.size(); }

    // Aligns everything collected since the last flush as one batch and appends the records to `out`.
    void flush(std::vector<mapper_record>& out) {
        const int64_t n = (int64_t)pending_.size();
        if (n == 0) return;
        cost_.resize((size_t)n);
        op_off_.resize((size_t)n + 1);
        if (ops_.size() < (size_t)(16 * n + 64)) ops_.resize((size_t)(16 * n + 64));
        int rc = gasma_greedy_batch(ctx_, reads_.data(), read_off_.data(), refs_.data(), ref_off_.data(), n, cost_.data(),
                                    ops_.data(), (int64_t)ops_.size(), op_off_.data());
        if (rc == GASMA_E_CAPACITY) {   // op_off[n] holds the size needed
            ops_.resize(op_off_[(size_t)n]);
            rc = gasma_greedy_batch(ctx_, reads_.data(), read_off_.data(), refs_.data(), ref_off_.data(), n, cost_.data(),
                                    ops_.data(), (int64_t)ops_.size(), op_off_.data());
        }
        if (rc != GASMA_OK) {   // nothing of this chunk survives a failed call: do not leave it pending for the next flush
            const std::string why = gasma_last_error(ctx_);
            reset_pending();
            throw std::runtime_error("mapper_gpu::flush: " + why);
        }
        for (int64_t i = 0; i < n; ++i) {
            mapper_record& r = pending_[(size_t)i];
            if (r.rejected) {
                r.cost = 0;
                r.cigar.clear();
            } else {
                r.cost = cost_[(size_t)i];
                r.cigar = sam_cigar(ops_.data() + op_off_[(size_t)i], (int)(op_off_[(size_t)i + 1] - op_off_[(size_t)i]));
            }
            out.push_back(std::move(r));
        }
        reset_pending();
        batches_++;
    }
    int64_t batches() const { return batches_; }
    gasma_ctx* context() { return ctx_; }

    // One SAM line per record, the fields of mapper/main.cpp:56-62: QNAME FLAG RNAME POS MAPQ CIGAR * 0 0 SEQ QUAL.
    // POS is the 1-based form of the reference's ref_offset = start + 2 (:88, its TODO included); MAPQ = 60 + cost (:91).
    static void write_sam_record(FILE* f, const mapper_record& r, const std::string& read_id, const std::string& ref_name,
                                 const std::string& seq, const std::string& qual) {
        if (r.rejected) {   // hit that could not be aligned (window clipped at a contig end): SAM "unmapped" record
            fprintf(f, "%s\t4\t*\t0\t0\t*\t*\t0\t0\t%s\t%s\n", read_id.c_str(), seq.c_str(), qual.empty() ? "*" : qual.c_str());
            return;
        }
        fprintf(f, "%s\t0\t%s\t%zu\t%d\t%s\t*\t0\t0\t%s\t%s\n", read_id.c_str(), ref_name.c_str(), r.window_start + 2 + 1,
                60 + r.cost, r.cigar.empty() ? "*" : r.cigar.c_str(), seq.c_str(), qual.empty() ? "*" : qual.c_str());
    }
    static void write_sam_header(FILE* f, const std::vector<std::string>& names, const std::vector<size_t>& lengths) {
        fprintf(f, "@HD\tVN:1.6\tSO:unknown\n");
        for (size_t i = 0; i < names.size(); ++i) fprintf(f, "@SQ\tSN:%s\tLN:%zu\n", names[i].c_str(), lengths[i]);
        fprintf(f, "@PG\tID:gasma-mapper\tPN:gasma-mapper\n");
    }

    // The aligner's ops are appended step by step and never merged (hurdle_matrix.h:238-251); SAM wants adjacent equal
    // ops merged.  The last M of a traversal that stops short of the destination counts mismatches, not columns
    // (hurdle_matrix.h:581-589, SURVEY hazard 4), so the result is the aligner's CIGAR, not a validated one.
    static std::string sam_cigar(const gasma_cigar_op* ops, int n) {
        std::string s;
        int run_len = 0, run_code = -1;
        auto put = [&]() {
            if (run_len > 0) { s += std::to_string(run_len); s += "MID=X"[run_code]; }
        };
        for (int i = 0; i < n; ++i) {
            const int code = GASMA_OP_CODE(ops[i]), len = GASMA_OP_LEN(ops[i]);
            if (code == run_code) run_len += len;
            else { put(); run_code = code; run_len = len; }
        }
        put();
        return s;
    }

private:
    void reset_pending() {
        pending_.clear();
        reads_.clear(); refs_.clear();
        read_off_.assign(1, 0u); ref_off_.assign(1, 0u);
    }
    gasma_ctx* ctx_ = nullptr;
    int max_length_;
    int64_t batches_ = 0;
    std::vector<mapper_record> pending_;
    std::string reads_, refs_;
    std::vector<uint32_t> read_off_{0u}, ref_off_{0u};
    std::vector<int32_t> cost_;
    std::vector<gasma_cigar_op> ops_;
    std::vector<uint32_t> op_off_;
};
