// Host-side container of a linearized BlobTree (opaque `psgpu_program` of include/psgpu.h).
#ifndef PSGPU_HOST_PROGRAM_H
#define PSGPU_HOST_PROGRAM_H
#include <cstdint>
#include <vector>

struct psgpu_program {
    std::vector<uint32_t> words;  // 4 per quad, see program.h
    std::vector<float> colors;    // 3 per primitive id
    int nPrims = 0;
    int nInstr = 0;
    int maxValueStack = 0;
    int maxPointStack = 0;
    int semantics = 0;
};
#endif
