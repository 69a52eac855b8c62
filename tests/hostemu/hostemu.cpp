// TEST INFRASTRUCTURE ONLY: the per-thread device functions of the product (eval_program, prune_walk) compiled
// for the host through cuda_shim.h, behind a tiny C API for the pytest suite (tests/test_linearizer.py,
// tests/test_cull_walk.py). Used to check the linearizer + interpreter + pruning logic against the oracle on a
// machine without a GPU. Never loaded by the product.
#include "cuda_shim.h"
#include "../../parsip_b200/csrc/cull_walk.cuh"

using namespace psgpu;

extern "C" {

// field of the quad program at n points
void hostemu_field(const float* quads, const float* xyz, int64_t n, float* f, int sem) {
    const float4* prog = reinterpret_cast<const float4*>(quads);
    for (int64_t i = 0; i < n; i++) {
        float X[1] = {xyz[i * 3]}, Y[1] = {xyz[i * 3 + 1]}, Z[1] = {xyz[i * 3 + 2]}, F[1];
        if (sem == PSGPU_SEM_SIMD) eval_program<1, false, true>(prog, X, Y, Z, F, nullptr, nullptr, sem);
        else eval_program<1, false, false>(prog, X, Y, Z, F, nullptr, nullptr, sem);
        f[i] = F[0];
    }
}

void hostemu_field_color(const float* quads, const float* colorTab, const float* xyz, int64_t n, float* f, float* col, int sem) {
    const float4* prog = reinterpret_cast<const float4*>(quads);
    for (int64_t i = 0; i < n; i++) {
        float X[1] = {xyz[i * 3]}, Y[1] = {xyz[i * 3 + 1]}, Z[1] = {xyz[i * 3 + 2]}, F[1];
        if (sem == PSGPU_SEM_SIMD) eval_program<1, true, true>(prog, X, Y, Z, F, col + i * 3, colorTab, sem);
        else eval_program<1, true, false>(prog, X, Y, Z, F, col + i * 3, colorTab, sem);
        f[i] = F[0];
    }
}

// prunes the program against the box; returns the pruned length in quads (0 = region needs no evaluation)
uint32_t hostemu_prune(const float* quads, const float* lo3, const float* hi3, int cullMode, float iso, float* out, uint32_t outCapQuads,
                       int colorExact) {
    const float4* prog = reinterpret_cast<const float4*>(quads);
    Box3 b;
    for (int a = 0; a < 3; a++) { b.lo[a] = lo3[a]; b.hi[a] = hi3[a]; }
    const uint32_t len = prune_walk<false>(prog, b, cullMode, iso, nullptr, colorExact != 0);
    if (len == 0 || len > outCapQuads || !out) return len;
    const uint32_t len2 = prune_walk<true>(prog, b, cullMode, iso, reinterpret_cast<float4*>(out), colorExact != 0, len);
    if (len2 != len) return 0xFFFFFFFFu;  // the measuring and the emitting pass must agree
    return len;
}

}
