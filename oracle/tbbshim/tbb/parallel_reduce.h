// TEST INFRASTRUCTURE ONLY: tbb::parallel_reduce / tbb::split for the reference's scalar polygonizer
// (Parsip100/ParsipHaptics/include/CPolyParsipOptimized.h:178 splitting constructor, CPolyParsipOptimized.cpp:402).
// Deterministic: the range is cut into contiguous pieces (8 per worker, so that an uneven range balances the way work stealing would
// balance it), every piece owns a body made by the splitting constructor, the pool's workers take pieces off an atomic counter, and the
// pieces are joined left to right - the result does not depend on which thread ran which piece.
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
    size_t np = (size_t)nt * 8;
    if (np > n) np = n;
    std::vector<std::unique_ptr<Body> > parts;
    for (size_t t = 1; t < np; t++) parts.emplace_back(new Body(body, split()));
    std::atomic<size_t> next(0);
    struct Ctx { const Range* range; Body* first; std::vector<std::unique_ptr<Body> >* parts; std::atomic<size_t>* next; size_t np, n; };
    Ctx ctx = { &range, &body, &parts, &next, np, n };
    struct Run {
        static void go(void* p) {
            Ctx* c = (Ctx*)p;
            for (;;) {
                const size_t t = c->next->fetch_add(1);
                if (t >= c->np) break;
                const size_t b = c->n * t / c->np, e = c->n * (t + 1) / c->np;
                Range sub(c->range->begin() + b, c->range->begin() + e);
                Body* bd = t == 0 ? c->first : (*c->parts)[t - 1].get();
                (*bd)(sub);
            }
        }
    };
    shim::Pool::get().run(nt, &Run::go, &ctx);
    for (size_t t = 0; t < parts.size(); t++) body.join(*parts[t]);
}
template <typename Range, typename Body, typename Part>
void parallel_reduce(const Range& range, Body& body, const Part&) { parallel_reduce(range, body); }
}
#endif
