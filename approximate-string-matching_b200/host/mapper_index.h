// mapper_index.h -- a small self-contained search stage for gasma-mapper: FASTA / FASTQ readers and an exact-seed
// index that proposes ONE best hit per read.
//
// The reference's mapper searches with SeqAn3's bi_fm_index (`search_cfg::max_error_total{errors} | hit_single_best`,
// GASMA/mapper/main.cpp:64-66, index built by mapper/indexer.cpp); SeqAn3 and cereal are not in this image, so this
// header stands in for that stage with the pigeonhole principle: a read with <= e errors cut into e+1 pieces has one
// piece that matches exactly, so the 16-mers at evenly spread read offsets are looked up in a sorted (k-mer, position)
// table and every occurrence votes for the diagonal (reference begin) it implies; the diagonal with the most votes
// is the hit.  This is host-side plumbing in front of the hot path (the alignment of the hits, mapper_gpu.h) and has
// no counterpart in the CUDA library; hits differ from the FM index's only where several placements tie.
#pragma once
#include <algorithm>
#include <cstdint>
#include <fstream>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

struct reference_storage_t {            // mapper/main.cpp:27-31 (ids + sequences; plain strings instead of seqan3::dna5)
    std::vector<std::string> ids;
    std::vector<std::string> seqs;
};

struct read_record_t { std::string id, seq, qual; };

inline std::string first_word(const std::string& s, size_t from) {
    size_t e = s.find_first_of(" \t\r", from);
    return s.substr(from, e == std::string::npos ? std::string::npos : e - from);
}

// read_reference, mapper/main.cpp:33-42 (FASTA; sequence lines concatenated, upper-cased)
inline void read_reference(const std::string& path, reference_storage_t& storage) {
    std::ifstream in(path);
    if (!in.is_open()) throw std::runtime_error("cannot open reference " + path);
    std::string line;
    while (std::getline(in, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (line.empty()) continue;
        if (line[0] == '>') { storage.ids.push_back(first_word(line, 1)); storage.seqs.emplace_back(); }
        else if (!storage.seqs.empty()) {
            for (char& c : line) if (c >= 'a' && c <= 'z') c = (char)(c - 32);
            storage.seqs.back() += line;
        }
    }
}

// FASTQ, four lines per record (what seqan3::sequence_file_input yields at mapper/main.cpp:54,73)
inline std::vector<read_record_t> read_queries(const std::string& path) {
    std::ifstream in(path);
    if (!in.is_open()) throw std::runtime_error("cannot open queries " + path);
    std::vector<read_record_t> out;
    std::string h, s, p, q;
    while (std::getline(in, h)) {
        if (h.empty()) continue;
        if (h[0] != '@' || !std::getline(in, s) || !std::getline(in, p) || !std::getline(in, q))
            throw std::runtime_error("malformed FASTQ record in " + path);
        for (std::string* x : {&h, &s, &q}) if (!x->empty() && x->back() == '\r') x->pop_back();
        for (char& c : s) if (c >= 'a' && c <= 'z') c = (char)(c - 32);
        out.push_back({first_word(h, 1), s, q});
    }
    return out;
}

class seed_index {
public:
    static constexpr int K = 16;
    explicit seed_index(const reference_storage_t& storage) {
        size_t total = 0;
        for (const std::string& s : storage.seqs) total += s.size();
        table_.reserve(total);
        for (size_t r = 0; r < storage.seqs.size(); ++r) {
            const std::string& s = storage.seqs[r];
            uint32_t code = 0; int valid = 0;
            for (size_t i = 0; i < s.size(); ++i) {
                const int b = base(s[i]);
                if (b < 0) { valid = 0; code = 0; continue; }
                code = (code << 2) | (uint32_t)b;
                if (++valid >= K) table_.push_back({code, ((uint64_t)r << 40) | (uint64_t)(i + 1 - K)});
            }
        }
        std::sort(table_.begin(), table_.end());
    }
    // Best placement of `query` with at most `errors` errors: (found, reference id, begin position).
    bool search(const std::string& query, int errors, size_t& ref_id, size_t& begin) const {
        if ((int)query.size() < K) return false;
        const int n_seeds = 2 * (errors + 1);   // pigeonhole needs e+1; twice that tolerates seeds that overlap an edit
        votes_.clear();
        const int span = (int)query.size() - K;
        for (int p = 0; p < n_seeds; ++p) {
            const int off = n_seeds > 1 ? (int)((int64_t)span * p / (n_seeds - 1)) : 0;
            uint32_t code = 0; bool ok = true;
            for (int i = 0; i < K; ++i) {
                const int b = base(query[(size_t)(off + i)]);
                if (b < 0) { ok = false; break; }
                code = (code << 2) | (uint32_t)b;
            }
            if (!ok) continue;
            auto lo = std::lower_bound(table_.begin(), table_.end(), std::make_pair(code, (uint64_t)0));
            for (; lo != table_.end() && lo->first == code; ++lo) {
                const uint64_t pos = lo->second & ((1ull << 40) - 1), rid = lo->second >> 40;
                if (pos < (uint64_t)off) continue;
                // diagonals within +-errors of each other are the same placement seen across an indel
                const uint64_t diag = pos - (uint64_t)off;
                bool merged = false;
                for (auto& v : votes_)
                    if (v.rid == rid && (v.diag > diag ? v.diag - diag : diag - v.diag) <= (uint64_t)errors) { v.count++; v.diag = std::min(v.diag, diag); merged = true; break; }
                if (!merged) votes_.push_back({rid, diag, 1});
            }
        }
        const vote* best = nullptr;
        for (const vote& v : votes_)
            if (!best || v.count > best->count || (v.count == best->count && std::make_pair(v.rid, v.diag) < std::make_pair(best->rid, best->diag))) best = &v;
        if (!best) return false;
        ref_id = (size_t)best->rid; begin = (size_t)best->diag;
        return true;
    }
    size_t size() const { return table_.size(); }

private:
    static int base(char c) {
        switch (c) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; default: return -1; }
    }
    struct vote { uint64_t rid, diag; int count; };
    std::vector<std::pair<uint32_t, uint64_t>> table_;
    mutable std::vector<vote> votes_;
};
