// demo.cpp -- the reference's GASMA/main.cpp:5-19 demo pair through hurdle_matrix_gpu (golden vector G1).
#include <cstdio>
#include <exception>

#include "hurdle_matrix_gpu.h"

int main(int argc, char** argv) {
    const char* read = argc > 2 ? argv[1] :
        "TGCGTCGCAGTGTCGAAAAAAATCTAAAGGGCCTACCGGGCCCCTTCATGAAGAGCAGGGAAAGAATCTCTATACTATTACTAATGGAAAGCAGTCTTGG";
    const char* ref = argc > 2 ? argv[2] :
        "TGCGTCGCAGTGTCGAAATAAAATCTAAAGGGCCTACCGGGCCCCTTCATGAAGAGTAGGTAAAGAATCTCTAGTACTACTACTAATGGAAAGCAGTCTTGG";
    try {
        hurdle_matrix_gpu matrix(read, ref, 3, GLOBAL, 1, 1, 1);
        matrix.run();
        printf("CIGAR: %s\ncost: %d\n", matrix.get_CIGAR().c_str(), matrix.get_cost());
    } catch (const std::exception& ex) {
        fprintf(stderr, "demo: %s\n", ex.what());
        return 1;
    }
    return 0;
}
