// TEST INFRASTRUCTURE ONLY: tbb::parallel_reduce / tbb::split for the reference's scalar polygonizer
// (Parsip100/ParsipHaptics/include/CPolyParsipOptimized.h:178 splitting constructor, CPolyParsipOptimized.cpp:402).
// Deterministic: the range is cut into one contiguous piece per worker, every worker owns a body made by the
// splitting constructor, and the pieces are joined left to right.
#ifndef TBBSHIM_PARALLEL_REDUCE_H
#define TBBSHIM_PARALLEL_REDUCE_H
#include "tbbshim_core.h"
namespace tbb {
class split {};
class affinity_partitioner {};
template <typename Range, typename Body>
void parallel_reduce(const Range& range, Body& body) {
    const size_t n = range.size();
    if (n == 0) return;
    int nt = shim::num_threads();
    if ((size_t)nt > n) nt = (int)n;
    if (nt <= 1) { body(range); return; }
    std::vector<std::unique_ptr<Body> > parts;
    for (int t = 1; t < nt; t++) parts.emplace_back(new Body(body, split()));
    std::vector<std::thread> th;
    for (int t = 0; t < nt; t++) {
        const size_t b = n * (size_t)t / (size_t)nt, e = n * (size_t)(t + 1) / (size_t)nt;
        Body* bd = t == 0 ? &body : parts[(size_t)t - 1].get();
        th.emplace_back([=, &range]() { Range sub(range.begin() + b, range.begin() + e); (*bd)(sub); });
    }
    for (size_t t = 0; t < th.size(); t++) th[t].join();
    for (size_t t = 0; t < parts.size(); t++) body.join(*parts[t]);
}
template <typename Range, typename Body, typename Part>
void parallel_reduce(const Range& range, Body& body, const Part&) { parallel_reduce(range, body); }
}
#endif
