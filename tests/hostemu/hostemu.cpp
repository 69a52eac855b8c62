// TEST INFRASTRUCTURE ONLY: the per-thread device functions of the product (eval_program, prune_walk) compiled
// for the host through cuda_shim.h, behind a tiny C API for the pytest suite (tests/test_linearizer.py,
// tests/test_cull_walk.py). Used to check the linearizer + interpreter + pruning logic against the oracle on a
// machine without a GPU. Never loaded by the product.
#include "cuda_shim.h"
#include "../../parsip_b200/csrc/cull_walk.cuh"
#include "../../parsip_b200/csrc/interp_packed.cuh"

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

// the packed-pair STRICT interpreter (interp_packed.cuh): columns = groups of 8 points sharing x and y (the field kernel's
// z-columns), xy[2*c], z[8*c + k] -> f[8*c + k]
void hostemu_field_packed_columns(const float* quads, const float* xy, const float* z, int64_t nCols, float* f) {
    const float4* prog = reinterpret_cast<const float4*>(quads);
    for (int64_t c = 0; c < nCols; c++) {
        PairPts<4, true> P;
        P.x[0] = xy[c * 2]; P.y[0] = xy[c * 2 + 1];
        for (int i = 0; i < 4; i++) P.z[i] = pk(z[c * 8 + 2 * i], z[c * 8 + 2 * i + 1]);
        F2 F[4];
        eval_program_packed<4, true, false, false, true>(prog, P, F, nullptr, nullptr, true);
        for (int i = 0; i < 4; i++) up(F[i], f[c * 8 + 2 * i], f[c * 8 + 2 * i + 1]);
    }
}

// ... and its general form: 4 arbitrary points per call (the vertex kernel's root / gradient samples); n is rounded down to a multiple of 4
void hostemu_field_packed(const float* quads, const float* xyz, int64_t n, float* f) {
    const float4* prog = reinterpret_cast<const float4*>(quads);
    for (int64_t b = 0; b + 4 <= n; b += 4) {
        PairPts<2, false> P;
        for (int i = 0; i < 2; i++) {
            const float* a = xyz + (b + 2 * i) * 3;
            P.x[i] = pk(a[0], a[3]); P.y[i] = pk(a[1], a[4]); P.z[i] = pk(a[2], a[5]);
        }
        F2 F[2];
        eval_program_packed<2, false, false, false, true>(prog, P, F, nullptr, nullptr, true);  // the 4 points = one SIMD packet of the reference
        for (int i = 0; i < 2; i++) up(F[i], f[b + 2 * i], f[b + 2 * i + 1]);
    }
}

// the packed walk WITH colour (color_kernel<.., NRM>): per point p the 4 samples (p, p + dx, p + dy, p + dz) of the normal pass, colour at p;
// f4[4*i .. 4*i+3] = the four fields, col[3*i ..] = the colour
void hostemu_field_color_packed(const float* quads, const float* colorTab, const float* xyz, int64_t n, float delta, float* f4, float* col) {
    const float4* prog = reinterpret_cast<const float4*>(quads);
    for (int64_t i = 0; i < n; i++) {
        const float px = xyz[i * 3], py = xyz[i * 3 + 1], pz = xyz[i * 3 + 2];
        PairPts<2, false> P;
        P.x[0] = pk(px, px + delta); P.x[1] = pk(px, px);
        P.y[0] = pk(py, py); P.y[1] = pk(py + delta, py);
        P.z[0] = pk(pz, pz); P.z[1] = pk(pz, pz + delta);
        F2 F[2];
        eval_program_packed<2, false, false, true, true>(prog, P, F, col + i * 3, colorTab);
        up(F[0], f4[i * 4], f4[i * 4 + 1]); up(F[1], f4[i * 4 + 2], f4[i * 4 + 3]);
    }
}

// prunes the program against the box; returns the pruned length in quads (0 = region needs no evaluation)
uint32_t hostemu_prune(const float* quads, const float* lo3, const float* hi3, int cullMode, float iso, float* out, uint32_t outCapQuads,
                       int colorExact) {
    const float4* prog = reinterpret_cast<const float4*>(quads);
    Box3 b;
    for (int a = 0; a < 3; a++) { b.lo[a] = lo3[a]; b.hi[a] = hi3[a]; }
    const bool fine = (colorExact & 2) != 0;  // bit 1: the slab prune's fine tiling of elongated boxes (cull_walk.cuh: prim_bound)
    colorExact &= 1;
    const uint32_t len = prune_walk<false>(prog, b, cullMode, iso, nullptr, colorExact != 0, 0xFFFFFFFFu, nullptr, fine);
    if (len == 0 || len > outCapQuads || !out) return len;
    const uint32_t len2 = prune_walk<true>(prog, b, cullMode, iso, reinterpret_cast<float4*>(out), colorExact != 0, len, nullptr, fine);
    if (len2 != len) return 0xFFFFFFFFu;  // the measuring and the emitting pass must agree
    return len;
}

}
