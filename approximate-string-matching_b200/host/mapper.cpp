// mapper.cpp -- main() of gasma-mapper, the B200 counterpart of GASMA/mapper/main.cpp (CMake target my-mapper):
// read the reference (FASTA) and the queries (FASTQ), find one best hit per read, align every hit with the greedy
// aligner (band 3, GLOBAL, x=o=e=1; mapper/main.cpp:71,82) and write SAM with mapq = 60 + cost (:91).
// The hits of a chunk of reads are aligned as ONE batch on the GPU (mapper_gpu.h); the search stage is mapper_index.h
// (SeqAn3's FM index is not in this image).
//
//   gasma-mapper -r ref.fa -q reads.fq [-o out.sam] [-e errors(0..4)] [--chunk N] [--device D]
//   gasma-mapper ... --hits-only      print "read_id ref_id begin" per read and stop (search stage only, no GPU needed)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <string>
#include <vector>

#include "mapper_gpu.h"
#include "mapper_index.h"

int main(int argc, char** argv) {
    std::string ref_path, query_path, sam_path = "out.sam";   // defaults of cmd_arguments, mapper/main.cpp:112-119
    int errors = 0, device = 0;
    long chunk = 1000000;
    bool hits_only = false;
    for (int i = 1; i < argc; ++i) {
        auto val = [&](const char* a, const char* b) { return (!strcmp(argv[i], a) || !strcmp(argv[i], b)) && i + 1 < argc; };
        if (val("-r", "--reference")) ref_path = argv[++i];
        else if (val("-q", "--query")) query_path = argv[++i];
        else if (val("-o", "--output")) sam_path = argv[++i];
        else if (val("-e", "--error")) errors = atoi(argv[++i]);
        else if (val("--chunk", "--chunk")) chunk = atol(argv[++i]);
        else if (val("--device", "--device")) device = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--hits-only")) hits_only = true;
        else { fprintf(stderr, "[PARSER ERROR] unknown argument %s\n", argv[i]); return -1; }
    }
    if (ref_path.empty() || query_path.empty() || errors < 0 || errors > 4 || chunk < 1) {   // arithmetic_range_validator{0, 4}, :139
        fprintf(stderr, "[PARSER ERROR] usage: gasma-mapper -r ref.fa -q reads.fq [-o out.sam] [-e 0..4]\n");
        return -1;
    }
    try {
        reference_storage_t storage;
        read_reference(ref_path, storage);
        const std::vector<read_record_t> reads = read_queries(query_path);
        const seed_index index(storage);
        if (hits_only) {
            for (const read_record_t& r : reads) {
                size_t rid = 0, begin = 0;
                if (index.search(r.seq, errors, rid, begin)) printf("%s %zu %zu\n", r.id.c_str(), rid, begin);
                else printf("%s - -\n", r.id.c_str());
            }
            return 0;
        }
        mapper_gpu mapper(3, device);                         // band 3: mapper/main.cpp:82
        FILE* sam = fopen(sam_path.c_str(), "w");
        if (!sam) throw std::runtime_error("cannot write " + sam_path);
        std::vector<size_t> lengths;
        for (const std::string& s : storage.seqs) lengths.push_back(s.size());
        mapper_gpu::write_sam_header(sam, storage.ids, lengths);
        std::vector<mapper_record> records;
        size_t mapped = 0;
        auto drain = [&]() {
            records.clear();
            mapper.flush(records);
            for (const mapper_record& rec : records) {
                const read_record_t& r = reads[rec.read_index];
                mapper_gpu::write_sam_record(sam, rec, r.id, storage.ids[rec.ref_id], r.seq, r.qual);
            }
            mapped += records.size();
        };
        for (size_t i = 0; i < reads.size(); ++i) {
            size_t rid = 0, begin = 0;
            if (index.search(reads[i].seq, errors, rid, begin)) mapper.add_hit(i, reads[i].seq, rid, storage.seqs[rid], begin);
            if ((long)mapper.pending() >= chunk) drain();
        }
        drain();
        fclose(sam);
        fprintf(stderr, "gasma-mapper: %zu reads, %zu hits aligned in %lld batch(es), SAM written to %s\n", reads.size(), mapped,
                (long long)mapper.batches(), sam_path.c_str());
    } catch (const std::exception& ex) {
        fprintf(stderr, "gasma-mapper: %s\n", ex.what());
        return 1;
    }
    return 0;
}
