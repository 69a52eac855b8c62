// Device-side interpreter of the postfix quad stream (program.h): evaluates the BlobTree field (and
// optionally the colour) at N points per thread. Every thread of a CTA walks the SAME instruction list, so
// the opcode switch is warp-uniform and parameter reads are shared-memory / L1 broadcasts; register
// blocking over N points amortises the decode.
//
// Restates FieldComputer::fieldValue / computePrimitiveField / fieldValueAndColor
// (PS_SimdPoly/include/PS_Polygonizer.cpp:934-1551) and COMPACTBLOBTREE::fieldvalueOp / fieldvaluePrim /
// baseColorOp / warp* (ParsipHaptics/include/CompactBlobTree.cpp:677-1092, 1124-1294, 1315-1536) as a
// stack machine instead of an explicit-stack tree walk / recursion.
#ifndef PSGPU_INTERP_CUH
#define PSGPU_INTERP_CUH
#ifndef PSGPU_HOST_EMULATION
#include <cuda_runtime.h>
#endif
#include "../../include/psgpu.h"
#include "program.h"

namespace psgpu {

__device__ __forceinline__ float wyvill(float d2) {
    // max(0, (1-d^2)^3): PS_Polygonizer.h:397-407 (== ComputeWyvillFieldValueSquare for d^2 < 1)
    const float t = 1.0f - d2;
    return fmaxf(t * t * t, 0.0f);
}

// ---- STRICT arithmetic: the reference's SSE code has no FMA and a fixed association order, so the PS_SimdPoly
// compat semantics (PSGPU_SEM_SIMD) evaluate every expression with separately rounded mul/add in exactly the order of
// PS_Polygonizer.cpp:934-1179 (the __f*_rn intrinsics are never contracted by nvcc). Field values, cube cases and
// 4-sample roots then come out BIT-IDENTICAL to the reference build instead of within an FMA-rounding tolerance that
// (1-d^2)^3 amplifies near f -> 0. The native semantics keep the fused forms.
#define PS_ADD __fadd_rn
#define PS_SUB __fsub_rn
#define PS_MUL __fmul_rn
#define PS_DIV __fdiv_rn
__device__ __forceinline__ float ps_sqrt_rn(float x) {
#ifdef PSGPU_HOST_EMULATION
    return sqrtf(x);
#else
    return __fsqrt_rn(x);
#endif
}
// _mm_max_ps(a, b) = a > b ? a : b (returns b when either is NaN); _mm_min_ps likewise
__device__ __forceinline__ float sse_max(float a, float b) { return a > b ? a : b; }
__device__ __forceinline__ float sse_min(float a, float b) { return a < b ? a : b; }
__device__ __forceinline__ float dot3_strict(float ax, float ay, float az, float bx, float by, float bz) {
    return PS_ADD(PS_ADD(PS_MUL(ax, bx), PS_MUL(ay, by)), PS_MUL(az, bz));
}
__device__ __forceinline__ float wyvill_strict(float d2) {
    const float r1 = PS_SUB(1.0f, d2);
    return sse_max(0.0f, PS_MUL(PS_MUL(r1, r1), r1));
}

// fast_pow of PS_SIMDVecN.h:122-128 with an exact reciprocal
__device__ __forceinline__ float ricci_simd(float base, float e) { return base / (base + (e - e * base)); }

template <int N, bool STRICT = false>
__device__ __forceinline__ void xform_points(const float4* q, const float (&x)[N], const float (&y)[N], const float (&z)[N],
                                             float (&ox)[N], float (&oy)[N], float (&oz)[N]) {
    const float4 r0 = q[0], r1 = q[1], r2 = q[2];
    if (STRICT) {  // PS_Polygonizer.cpp:951-969: ((m0*x + m1*y) + m2*z) + m3
#pragma unroll
        for (int n = 0; n < N; n++) {
            ox[n] = PS_ADD(dot3_strict(r0.x, r0.y, r0.z, x[n], y[n], z[n]), r0.w);
            oy[n] = PS_ADD(dot3_strict(r1.x, r1.y, r1.z, x[n], y[n], z[n]), r1.w);
            oz[n] = PS_ADD(dot3_strict(r2.x, r2.y, r2.z, x[n], y[n], z[n]), r2.w);
        }
        return;
    }
#pragma unroll
    for (int n = 0; n < N; n++) {
        ox[n] = fmaf(r0.x, x[n], fmaf(r0.y, y[n], fmaf(r0.z, z[n], r0.w)));
        oy[n] = fmaf(r1.x, x[n], fmaf(r1.y, y[n], fmaf(r1.z, z[n], r1.w)));
        oz[n] = fmaf(r2.x, x[n], fmaf(r2.y, y[n], fmaf(r2.z, z[n], r2.w)));
    }
}

// Squared skeleton distance of one primitive at N points. q0 = header quad, q = following quads.
// (Also used by the culling pass, whose bounds rest on these being distances to a set, i.e. 1-Lipschitz.)
// LINES: 0 = both line flavours (culling pass), 1 = clamped segments only (native interpreter; the unclamped line of
// the PS_SimdPoly semantics only exists in the STRICT interpreter, prim_d2_strict) - keeps the hot loops' code small.
template <int N, int LINES = 0>
__device__ __forceinline__ void prim_d2(uint32_t opc, const float4 q0, const float4* __restrict__ q,
                                           const float (&x)[N], const float (&y)[N], const float (&z)[N], float (&d2o)[N]) {
    if (LINES == 1 && opc == OPC_LINE_INF) opc = OPC_NULLPRIM;
    switch (opc) {
    case OPC_POINT: case OPC_QUADRIC: {  // the quadric point's d2 is the plain distance too; its field law differs (prim_field)
#pragma unroll
        for (int n = 0; n < N; n++) {
            const float dx = q0.y - x[n], dy = q0.z - y[n], dz = q0.w - z[n];
            d2o[n] = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
        }
    } break;
    case OPC_SEGMENT: {  // clamped to the segment: PS_GeometryFuncs.cpp:156-179
        const float4 d = q[0];
#pragma unroll
        for (int n = 0; n < N; n++) {
            const float vx = x[n] - q0.y, vy = y[n] - q0.z, vz = z[n] - q0.w;
            const float t = __saturatef(fmaf(vx, d.x, fmaf(vy, d.y, vz * d.z)) * d.w);
            const float rx = fmaf(-t, d.x, vx), ry = fmaf(-t, d.y, vy), rz = fmaf(-t, d.z, vz);
            d2o[n] = fmaf(rx, rx, fmaf(ry, ry, rz * rz));
        }
    } break;
    case OPC_LINE_INF: if (LINES != 1) {  // PS_Polygonizer.cpp:984-1011: projection not clamped
        const float4 d = q[0];
#pragma unroll
        for (int n = 0; n < N; n++) {
            const float vx = x[n] - q0.y, vy = y[n] - q0.z, vz = z[n] - q0.w;
            const float t = fmaf(vx, d.x, fmaf(vy, d.y, vz * d.z)) / d.w;
            const float rx = fmaf(-t, d.x, vx), ry = fmaf(-t, d.y, vy), rz = fmaf(-t, d.z, vz);
            d2o[n] = fmaf(rx, rx, fmaf(ry, ry, rz * rz));
        }
    } break;
    case OPC_CYLINDER: {  // PS_Polygonizer.cpp:1012-1039 == CompactBlobTree.cpp:915-930
        const float4 d = q[0];
        const float height = q[1].x;
#pragma unroll
        for (int n = 0; n < N; n++) {
            const float vx = x[n] - q0.y, vy = y[n] - q0.z, vz = z[n] - q0.w;
            float yy = fmaf(vx, d.x, fmaf(vy, d.y, vz * d.z));
            const float r2 = fmaf(vx, vx, fmaf(vy, vy, vz * vz)) - yy * yy;
            const float xx = fmaxf(sqrtf(fmaxf(r2, 0.0f)) - d.w, 0.0f);
            yy = (yy > 0.0f) ? fmaxf(yy - height, 0.0f) : yy;
            d2o[n] = fmaf(xx, xx, yy * yy);
        }
    } break;
    case OPC_CUBE: {  // PS_Polygonizer.cpp:1059-1097: per-axis overshoot beyond +-side
        const float side = q[0].x;
#pragma unroll
        for (int n = 0; n < N; n++) {
            const float ex = fmaxf(fabsf(x[n] - q0.y) - side, 0.0f);
            const float ey = fmaxf(fabsf(y[n] - q0.z) - side, 0.0f);
            const float ez = fmaxf(fabsf(z[n] - q0.w) - side, 0.0f);
            d2o[n] = fmaf(ex, ex, fmaf(ey, ey, ez * ez));
        }
    } break;
    case OPC_DISC: {  // PS_Polygonizer.cpp:1099-1132 / CompactBlobTree.cpp:996-1017, exact rsqrt (Q7)
        const float4 nr = q[0];
#pragma unroll
        for (int n = 0; n < N; n++) {
            const float vx = x[n] - q0.y, vy = y[n] - q0.z, vz = z[n] - q0.w;
            const float nd = fmaf(nr.x, vx, fmaf(nr.y, vy, nr.z * vz));
            const float ux = fmaf(-nr.x, nd, vx), uy = fmaf(-nr.y, nd, vy), uz = fmaf(-nr.z, nd, vz);
            const float u2 = fmaf(ux, ux, fmaf(uy, uy, uz * uz));
            float d2;
            if (nr.w * nr.w >= u2) d2 = fabsf(fmaf(vx, vx, fmaf(vy, vy, vz * vz)) - u2);
            else {
                const float s = nr.w * rsqrtf(u2);
                const float wx = fmaf(s, ux, -vx), wy = fmaf(s, uy, -vy), wz = fmaf(s, uz, -vz);
                d2 = fmaf(wx, wx, fmaf(wy, wy, wz * wz));
            }
            d2o[n] = d2;
        }
    } break;
    case OPC_RING: {  // PS_Polygonizer.cpp:1134-1173 / CompactBlobTree.cpp:1019-1040
        const float4 nr = q[0];
#pragma unroll
        for (int n = 0; n < N; n++) {
            const float vx = x[n] - q0.y, vy = y[n] - q0.z, vz = z[n] - q0.w;
            const float nd = fmaf(nr.x, vx, fmaf(nr.y, vy, nr.z * vz));
            const float ux = fmaf(-nr.x, nd, vx), uy = fmaf(-nr.y, nd, vy), uz = fmaf(-nr.z, nd, vz);
            const float u2 = fmaf(ux, ux, fmaf(uy, uy, uz * uz));
            float d2;
            if (u2 == 0.0f) d2 = fmaf(nr.w, nr.w, fmaf(vx, vx, fmaf(vy, vy, vz * vz)));
            else {
                const float s = nr.w * rsqrtf(u2);
                const float wx = fmaf(s, ux, -vx), wy = fmaf(s, uy, -vy), wz = fmaf(s, uz, -vz);
                d2 = fmaf(wx, wx, fmaf(wy, wy, wz * wz));
            }
            d2o[n] = d2;
        }
    } break;
    default: {  // OPC_NULLPRIM
#pragma unroll
        for (int n = 0; n < N; n++) d2o[n] = 4.0f;  // field 0
    } break;
    }
}

// The same squared distances in the reference's exact operation order (PS_Polygonizer.cpp:975-1173, SSE, no FMA; see the
// STRICT note above). Only the opcodes the PS_SimdPoly semantics can produce have a strict form; the rest fall through.
// The mask-blend expressions (mask*A + (1-mask)*B) are kept because 0*NaN / 0*inf propagate in the reference too.
template <int N>
__device__ __forceinline__ void prim_d2_strict(uint32_t opc, const float4 q0, const float4* __restrict__ q,
                                               const float (&x)[N], const float (&y)[N], const float (&z)[N], float (&d2o)[N]) {
    switch (opc) {
    case OPC_POINT: {  // :975-983
#pragma unroll
        for (int n = 0; n < N; n++) {
            const float dx = PS_SUB(q0.y, x[n]), dy = PS_SUB(q0.z, y[n]), dz = PS_SUB(q0.w, z[n]);
            d2o[n] = dot3_strict(dx, dy, dz, dx, dy, dz);
        }
    } break;
    case OPC_LINE_INF: {  // :984-1011
        const float4 d = q[0];
#pragma unroll
        for (int n = 0; n < N; n++) {
            float dx = PS_SUB(x[n], q0.y), dy = PS_SUB(y[n], q0.z), dz = PS_SUB(z[n], q0.w);
            const float delta = PS_DIV(dot3_strict(dx, dy, dz, d.x, d.y, d.z), d.w);
            dx = PS_SUB(x[n], PS_ADD(q0.y, PS_MUL(delta, d.x)));
            dy = PS_SUB(y[n], PS_ADD(q0.z, PS_MUL(delta, d.y)));
            dz = PS_SUB(z[n], PS_ADD(q0.w, PS_MUL(delta, d.z)));
            d2o[n] = dot3_strict(dx, dy, dz, dx, dy, dz);
        }
    } break;
    case OPC_CYLINDER: {  // :1012-1039
        const float4 d = q[0];
        const float height = q[1].x;
#pragma unroll
        for (int n = 0; n < N; n++) {
            const float px = PS_SUB(x[n], q0.y), py = PS_SUB(y[n], q0.z), pz = PS_SUB(z[n], q0.w);
            float yy = dot3_strict(px, py, pz, d.x, d.y, d.z);
            const float xx = sse_max(0.0f, PS_SUB(ps_sqrt_rn(PS_SUB(dot3_strict(px, py, pz, px, py, pz), PS_MUL(yy, yy))), d.w));
            const float mask = (yy > 0.0f) ? 1.0f : 0.0f;
            yy = PS_ADD(PS_MUL(mask, sse_max(0.0f, PS_SUB(yy, height))), PS_MUL(PS_SUB(1.0f, mask), yy));
            d2o[n] = PS_ADD(PS_MUL(xx, xx), PS_MUL(yy, yy));
        }
    } break;
    case OPC_CUBE: {  // :1059-1097
        const float side = q[0].x, minusSide = PS_MUL(-1.0f, side);
#pragma unroll
        for (int n = 0; n < N; n++) {
            const float dif[3] = {PS_SUB(x[n], q0.y), PS_SUB(y[n], q0.z), PS_SUB(z[n], q0.w)};
            float acc = 0.0f;
#pragma unroll
            for (int a = 0; a < 3; a++) {
                const float mm = (minusSide > dif[a]) ? 1.0f : 0.0f, mp = (dif[a] > side) ? 1.0f : 0.0f;
                const float delta = PS_ADD(PS_MUL(PS_ADD(dif[a], side), mm), PS_MUL(PS_SUB(dif[a], side), mp));
                acc = (a == 0) ? PS_MUL(delta, delta) : PS_ADD(acc, PS_MUL(delta, delta));
            }
            d2o[n] = acc;
        }
    } break;
    case OPC_DISC:  // :1099-1132 (SimdRSqrt evaluated exactly)
    case OPC_RING: {  // :1134-1173
        const float4 nr = q[0];
        const float radius = nr.w;
#pragma unroll
        for (int n = 0; n < N; n++) {
            const float dX = PS_SUB(x[n], q0.y), dY = PS_SUB(y[n], q0.z), dZ = PS_SUB(z[n], q0.w);
            float dot = dot3_strict(nr.x, nr.y, nr.z, dX, dY, dZ);
            float dirX = PS_SUB(dX, PS_MUL(nr.x, dot)), dirY = PS_SUB(dY, PS_MUL(nr.y, dot)), dirZ = PS_SUB(dZ, PS_MUL(nr.z, dot));
            dot = dot3_strict(dirX, dirY, dirZ, dirX, dirY, dirZ);
            const float rs = PS_DIV(1.0f, ps_sqrt_rn(dot));
            dirX = PS_MUL(dirX, rs); dirY = PS_MUL(dirY, rs); dirZ = PS_MUL(dirZ, rs);
            const float nX = PS_SUB(PS_MUL(radius, dirX), dX), nY = PS_SUB(PS_MUL(radius, dirY), dY), nZ = PS_SUB(PS_MUL(radius, dirZ), dZ);
            const float far2 = dot3_strict(nX, nY, nZ, nX, nY, nZ);
            const float dd = dot3_strict(dX, dY, dZ, dX, dY, dZ);
            if (opc == OPC_DISC) {
                const float mask = (PS_MUL(radius, radius) >= dot) ? 1.0f : 0.0f;
                d2o[n] = PS_ADD(PS_MUL(mask, PS_SUB(dd, dot)), PS_MUL(PS_SUB(1.0f, mask), far2));
            } else {
                const float mask = (dot == 0.0f) ? 1.0f : 0.0f;
                // radius*radius + dX*dX + dY*dY + dZ*dZ, left to right
                const float on = PS_ADD(PS_ADD(PS_ADD(PS_MUL(radius, radius), PS_MUL(dX, dX)), PS_MUL(dY, dY)), PS_MUL(dZ, dZ));
                d2o[n] = PS_ADD(PS_MUL(mask, on), PS_MUL(PS_SUB(1.0f, mask), far2));
            }
        }
    } break;
    default: {  // OPC_NULLPRIM (segments, the native line flavour, do not occur in PS_SimdPoly programs)
#pragma unroll
        for (int n = 0; n < N; n++) d2o[n] = 4.0f;
    } break;
    }
}

template <int N, bool STRICT = false>
__device__ __forceinline__ void prim_field(uint32_t opc, const float4 q0, const float4* __restrict__ q,
                                           const float (&x)[N], const float (&y)[N], const float (&z)[N], float (&f)[N]) {
    float d2[N];
    if (STRICT) {
        prim_d2_strict<N>(opc, q0, q, x, y, z, d2);
#pragma unroll
        for (int n = 0; n < N; n++) f[n] = wyvill_strict(d2[n]);
    } else {
        prim_d2<N, 1>(opc, q0, q, x, y, z, d2);
        if (opc == OPC_QUADRIC) {  // CompactBlobTree.cpp:1051-1064: scale * v * v with v = 1 - d2 / R^2, 0 where v <= 0
            const float4 a = q[0];
#pragma unroll
            for (int n = 0; n < N; n++) { const float v = fmaxf(fmaf(-d2[n], a.x, 1.0f), 0.0f); f[n] = (a.y * v) * v; }
            return;
        }
#pragma unroll
        for (int n = 0; n < N; n++) f[n] = wyvill(d2[n]);
    }
}

// primitive under its own backward matrix (header bit XF): out of line, see eval_program
// sincosf out of line: its argument reduction (with the slow path for large angles) is ~150 SASS instructions, and the warps
// would inline it once per point and call site - csgwarp's vertex kernel stalled 6 cycles per issue on instruction fetch
// Evaluated in double and rounded once: CUDA's sincosf is a 2-ulp routine, the reference's CPU build calls glibc's sinf / cosf
// (< 1 ulp), and a warp multiplies coordinates of size ~5 by these factors - two ulps there showed up as 6e-5 relative in the
// field under three nested warps (tests/test_gpu_compactref.py). One double sincos per warp instruction and point is noise next
// to the primitives below it (B200 keeps full-rate FP64 units; measured on csgwarp@512^3 in profiles/r02_configs.json).
static __device__ __noinline__ void ps_sincosf(float a, float& s, float& c) {
#ifndef PSGPU_SINCOS_F32
    double sd, cd;
    sincos((double)a, &sd, &cd);
    s = (float)sd; c = (float)cd;
#else
    sincosf(a, &s, &c);
#endif
}

template <int N, bool STRICT>
__device__ __noinline__ void prim_field_xf(uint32_t opc, const float4 q0, const float4* __restrict__ q, const float4* __restrict__ m,
                                           const float (&x)[N], const float (&y)[N], const float (&z)[N], float (&f)[N]) {
    float lx[N], ly[N], lz[N];
    // both semantics: row . (x, y, z, 1) summed left to right with separate roundings (PS_Polygonizer.cpp:951-969,
    // CompactBlobTree.cpp:897-905 Vector4d::dot), so the local point has the reference's bits
    xform_points<N, true>(m, x, y, z, lx, ly, lz);
    prim_field<N, STRICT>(opc, q0, q, lx, ly, lz, f);
}

__device__ __forceinline__ float combine_val(uint32_t cmb, float top, float v) {
    switch (cmb) {
    case CMB_ADD: case CMB_RADD: return top + v;
    case CMB_MAX: return fmaxf(top, v);
    case CMB_MIN: return fminf(top, v);
    case CMB_DIF: return fminf(top, 1.0f - v);
    case CMB_RDIF: return fminf(v, 1.0f - top);
    case CMB_RSDIF: return v * (1.0f - top);
    default: return top * (1.0f - v);  // CMB_SDIF
    }
}

// out[n] = a[n] (cmb) b[n] for all N points with ONE warp-uniform dispatch (the per-point form cost one jump-table branch
// per point: ncu r01b, BRX x4 per OPC_COMBINE). Same expressions as combine_val.
template <int N>
__device__ __forceinline__ void combine_n(uint32_t cmb, const float (&a)[N], const float (&b)[N], float (&out)[N]) {
    switch (cmb) {
    case CMB_ADD: case CMB_RADD:
#pragma unroll
        for (int n = 0; n < N; n++) out[n] = a[n] + b[n];
        break;
    case CMB_MAX:
#pragma unroll
        for (int n = 0; n < N; n++) out[n] = fmaxf(a[n], b[n]);
        break;
    case CMB_MIN:
#pragma unroll
        for (int n = 0; n < N; n++) out[n] = fminf(a[n], b[n]);
        break;
    case CMB_DIF:
#pragma unroll
        for (int n = 0; n < N; n++) out[n] = fminf(a[n], 1.0f - b[n]);
        break;
    case CMB_RDIF:
#pragma unroll
        for (int n = 0; n < N; n++) out[n] = fminf(b[n], 1.0f - a[n]);
        break;
    case CMB_RSDIF:
#pragma unroll
        for (int n = 0; n < N; n++) out[n] = b[n] * (1.0f - a[n]);
        break;
    default:  // CMB_SDIF
#pragma unroll
        for (int n = 0; n < N; n++) out[n] = a[n] * (1.0f - b[n]);
        break;
    }
}

// colour of `top (cmb) v` for point 0; (tf, tc) = stack top before the combine, (vf, vc) = incoming value.
// taux = running min used by the scalar twin's (smooth) difference colour pick.
__device__ __forceinline__ void combine_color(int sem, uint32_t cmb, float tf, float (&tc)[3], float& taux, float vf,
                                              const float (&vc)[3], float result) {
    if (sem == PSGPU_SEM_SIMD) {
        float wl, wr;
        if (cmb == CMB_ADD) {  // PS_Polygonizer.cpp:1475-1487: two*(half+f)-one, unnormalised
            wl = __fsub_rn(__fmul_rn(2.0f, __fadd_rn(0.5f, tf)), 1.0f);
            wr = __fsub_rn(__fmul_rn(2.0f, __fadd_rn(0.5f, vf)), 1.0f);
        } else if (cmb == CMB_MAX || cmb == CMB_MIN) {  // :1489-1500
            wl = (result - tf == 0.0f) ? 1.0f : 0.0f;
            wr = (result - vf == 0.0f) ? 1.0f : 0.0f;
        } else if (cmb == CMB_RDIF || cmb == CMB_RSDIF) {  // :1502-1512 with the incoming value as the LEFT operand
            wl = (1.0f - tf == result) ? 1.0f : 0.0f;
            wr = (vf == result) ? 1.0f : 0.0f;
        } else {  // :1502-1512
            wl = (tf == result) ? 1.0f : 0.0f;
            wr = (1.0f - vf == result) ? 1.0f : 0.0f;
        }
#pragma unroll
        for (int a = 0; a < 3; a++) tc[a] = __fadd_rn(__fmul_rn(wl, tc[a]), __fmul_rn(wr, vc[a]));
    } else {  // CompactBlobTree.cpp:1186-1260, kept normalised per entry (association-invariant)
        // taux = what the reference's `best` would be if this entry became the first child of a (smooth) difference: the
        // entry's value so far; under DIF / SDIF the running min(f0, 1 - f_i) that picks the colour; under RADD the raw sum.
        if (cmb == CMB_ADD) {
            const float s = tf + vf;
            if (s != 0.0f) {
                const float inv = 1.0f / s;
#pragma unroll
                for (int a = 0; a < 3; a++) tc[a] = (tf * tc[a] + vf * vc[a]) * inv;
            }
            taux = result;
        } else if (cmb == CMB_MAX) {
            if (vf > tf) { tc[0] = vc[0]; tc[1] = vc[1]; tc[2] = vc[2]; }
            taux = result;
        } else if (cmb == CMB_MIN) {
            if (vf < tf) { tc[0] = vc[0]; tc[1] = vc[1]; tc[2] = vc[2]; }
            taux = result;
        } else {
            const float cf = 1.0f - vf;
            if (cf < taux) { taux = cf; tc[0] = vc[0]; tc[1] = vc[1]; tc[2] = vc[2]; }
        }
    }
}

// PS_SimdPoly colour weight of a blend operand whose sibling was culled as all-zero (program.h: OPC_CSCALE / header bits 28..31)
__device__ __forceinline__ void color_scale(float top0, float (&tc)[3], uint32_t cnt) {
    const float w = __fsub_rn(__fmul_rn(2.0f, __fadd_rn(0.5f, top0)), 1.0f);
    // one separately rounded multiply per dropped sibling and channel, in the reference's order (no power shortcut): unrolled so
    // that a long run (a blend that lost 17 of 20 operands) costs the 3 multiplies per step and little else
#pragma unroll 4
    for (uint32_t k = 0; k < cnt; k++) { tc[0] = __fmul_rn(w, tc[0]); tc[1] = __fmul_rn(w, tc[1]); tc[2] = __fmul_rn(w, tc[2]); }
}

// Evaluates the program at N points. COLOR_T compiles the colour walk in (colour at point 0); `doColor` switches it on at
// run time, so a kernel that needs both plain and coloured evaluations shares ONE copy of the interpreter - the hot loop of
// two inlined copies does not fit the instruction caches (ncu: stall_no_instruction dominated the first vertex kernel).
// STRICT (PS_SimdPoly semantics) also compiles OUT everything those programs cannot contain (warps, powers, segments), and
// the native build drops the unclamped line / fast_pow Ricci / colour-scale opcodes.
// `prog` may point to shared or global memory (generic address).
// WARPS = the program is known to contain warp instructions: their sincosf goes through ONE out-of-line copy (ps_sincosf) so
// the hot loop stays inside the instruction cache. Kernels for warp-free programs (WARPS = false) keep the inline form: it is
// never executed there, and a call in the loop costs the 8-point field kernel 6 % (measured A/B on blob1k, native semantics)
template <int N, bool COLOR_T, bool STRICT = false, bool SMEM = false, bool WARPS = true>
__device__ void eval_program(const float4* __restrict__ prog, const float (&X)[N], const float (&Y)[N], const float (&Z)[N],
                             float (&F)[N], float* rgb, const float* __restrict__ colorTab, int sem, bool doColor = true) {
    // SMEM: the program was staged into shared memory; tells the compiler so the decode path uses LDS instead of generic loads
#ifndef PSGPU_HOST_EMULATION
    if (SMEM) __builtin_assume(__isShared(prog));
#endif
    const bool COLOR = COLOR_T && doColor;
    float x[N], y[N], z[N];
    float top[N];
    // entry 0 of the stacks is a dummy: the first PUSH stores the (unused) initial top there, so PUSH needs no "first" test
    float vs[PS_MAX_VALUE_STACK + 1][N];
    float pstk[STRICT ? 1 : PS_MAX_POINT_STACK][3][N];
    float cstk[COLOR_T ? PS_MAX_VALUE_STACK + 1 : 1][5];
    float tc[3] = {0.f, 0.f, 0.f};
    float taux = 0.f;
    int sp = 0, psp = 0;
#pragma unroll
    for (int n = 0; n < N; n++) { x[n] = X[n]; y[n] = Y[n]; z[n] = Z[n]; top[n] = 0.0f; }
    int pc = 0;
    for (;;) {
        const float4 q0 = prog[pc];
        const uint32_t hdr = __float_as_uint(q0.x);
        const uint32_t opc = PS_HDR_OPC(hdr);
        if (opc == OPC_END) break;
        if (opc < OPC_FIRST_NONPRIM) {
            const uint32_t cmb = PS_HDR_CMB(hdr);
            const int nq = ps_base_quads(opc);
            const bool xf = PS_HDR_XF(hdr) != 0u;
            float f[N];
            // a primitive with its own matrix is evaluated by an out-of-line copy (cold: matrices are rare), so that the hot path
            // never writes x/y/z: where a caller passes the same x and y for all N points (the field kernel's z-columns) the
            // compiler then shares the x/y part of every distance across the points
            // (the native interpreter rewrites x/y/z under warps anyway and measured faster with the matrix applied in place)
            if (STRICT) {
                if (xf) prim_field_xf<N, STRICT>(opc, q0, prog + pc + 1, prog + pc + nq, x, y, z, f);
                else prim_field<N, STRICT>(opc, q0, prog + pc + 1, x, y, z, f);  // ONE inlined copy of the primitive switch (i-cache)
            } else {
                float xsave[3][N];
                if (xf) {
                    float lx[N], ly[N], lz[N];
                    xform_points<N, true>(prog + pc + nq, x, y, z, lx, ly, lz);
#pragma unroll
                    for (int n = 0; n < N; n++) { xsave[0][n] = x[n]; xsave[1][n] = y[n]; xsave[2][n] = z[n]; x[n] = lx[n]; y[n] = ly[n]; z[n] = lz[n]; }
                }
                prim_field<N, STRICT>(opc, q0, prog + pc + 1, x, y, z, f);
                if (xf) {
#pragma unroll
                    for (int n = 0; n < N; n++) { x[n] = xsave[0][n]; y[n] = xsave[1][n]; z[n] = xsave[2][n]; }
                }
            }
            pc += nq + (xf ? 3 : 0);
            float pcol[3];
            if (COLOR) {
                const uint32_t id = PS_HDR_PRIM(hdr);
                pcol[0] = colorTab[id * 3]; pcol[1] = colorTab[id * 3 + 1]; pcol[2] = colorTab[id * 3 + 2];
            }
            if (STRICT && COLOR && PS_HDR_PRE(hdr)) color_scale(f[0], pcol, PS_HDR_CSC(hdr));  // folded one-primitive group (program.h)
            if (cmb == CMB_PUSH) {
#pragma unroll
                for (int n = 0; n < N; n++) { vs[sp][n] = top[n]; top[n] = f[n]; }
                if (COLOR) {
                    cstk[sp][0] = tc[0]; cstk[sp][1] = tc[1]; cstk[sp][2] = tc[2]; cstk[sp][3] = taux;
                    tc[0] = pcol[0]; tc[1] = pcol[1]; tc[2] = pcol[2]; taux = f[0];
                }
                sp++;
            } else {
                if (COLOR) {
                    const float r0 = combine_val(cmb, top[0], f[0]);
                    combine_color(sem, cmb, top[0], tc, taux, f[0], pcol, r0);
                }
                combine_n<N>(cmb, top, f, top);
            }
            if (STRICT && COLOR && PS_HDR_CSC(hdr) && !PS_HDR_PRE(hdr)) color_scale(top[0], tc, PS_HDR_CSC(hdr));
            continue;
        }
        const int pcNext = pc + ps_base_quads(opc);
        switch (opc) {
        case OPC_COMBINE: {
            const uint32_t cmb = PS_HDR_CMB(hdr);
            sp--;
            if (COLOR) {
                float vc[3] = {tc[0], tc[1], tc[2]};
                const float vf = top[0];
                const float vaux = taux;
                tc[0] = cstk[sp][0]; tc[1] = cstk[sp][1]; tc[2] = cstk[sp][2]; taux = cstk[sp][3];
                const float tf = vs[sp][0];
                const float r0 = combine_val(cmb, tf, vf);
                if (cmb == CMB_RADD) {  // weights = raw child fields: taux = raw sum so far, vaux = the incoming child's raw field
                    const float ssum = taux + vaux;
                    combine_color(sem, CMB_ADD, taux, tc, taux, vaux, vc, ssum);
                } else combine_color(sem, cmb, tf, tc, taux, vf, vc, r0);
            }
            {
                float lhs[N];
#pragma unroll
                for (int n = 0; n < N; n++) lhs[n] = vs[sp][n];
                combine_n<N>(cmb, lhs, top, top);
            }
        } break;
        case OPC_POW: if (!STRICT) {
            if (COLOR) taux = (PS_HDR_CMB(hdr) == 1u) ? top[0] : powf(top[0], q0.y);  // raw child field / the blend's value
#pragma unroll
            for (int n = 0; n < N; n++) top[n] = powf(top[n], q0.y);
        } break;
        case OPC_RICCI_SIMD: if (STRICT) {
            sp--;
            if (COLOR) {
                float vc[3] = {tc[0], tc[1], tc[2]};
                const float vf = top[0];
                tc[0] = cstk[sp][0]; tc[1] = cstk[sp][1]; tc[2] = cstk[sp][2]; taux = cstk[sp][3];
                combine_color(sem, CMB_ADD, vs[sp][0], tc, taux, vf, vc, 0.0f);
            }
#pragma unroll
            for (int n = 0; n < N; n++) top[n] = ricci_simd(vs[sp][n] + top[n], q0.y);
        } break;
        case OPC_CSCALE: if (STRICT) {
            if (COLOR) color_scale(top[0], tc, __float_as_uint(q0.y));
        } break;
        case OPC_BOXMASK: if (STRICT) {  // every point is a packet of its own here (broadcast evaluations; program.h)
            const float4 lo = prog[pc + 1], hi = prog[pc + 2];
#pragma unroll
            for (int n = 0; n < N; n++) {
                const bool in = (x[n] >= lo.x && hi.x >= x[n]) || (y[n] >= lo.y && hi.y >= y[n]) || (z[n] >= lo.z && hi.z >= z[n]);
                if (!in) top[n] = 0.0f;
            }
        } break;
        case OPC_CLAMP1: {
#pragma unroll
            for (int n = 0; n < N; n++) top[n] = fminf(top[n], 1.0f);
        } break;
        case OPC_DIF0: {
#pragma unroll
            for (int n = 0; n < N; n++) top[n] = fminf(0.0f, 1.0f - top[n]);
        } break;
        case OPC_POPPOINT: if (!STRICT) {
            psp--;
#pragma unroll
            for (int n = 0; n < N; n++) { x[n] = pstk[psp][0][n]; y[n] = pstk[psp][1][n]; z[n] = pstk[psp][2][n]; }
        } break;
        default: if (!STRICT) {  // warps and operator affine: push the point, transform it
#pragma unroll
            for (int n = 0; n < N; n++) { pstk[psp][0][n] = x[n]; pstk[psp][1][n] = y[n]; pstk[psp][2][n] = z[n]; }
            psp++;
            // Point transforms keep the reference's operation order with separately rounded multiplies and adds (its CPU build
            // has no FMA): the Barr bend subtracts and re-adds 1/k (10-50 for gentle bends), so one fused rounding at that
            // magnitude moved the warped point by ulp(50) ~ 4e-6 and the field by 6e-5 relative (tests/test_gpu_compactref.py,
            // reproduced on the host with -ffp-contract=fast). With this form and the double-rounded sin / cos the warped point
            // has the reference's bits; a handful of instructions per warp and point.
            if (opc == OPC_AFFINE) {
                float ox[N], oy[N], oz[N];
                xform_points<N, true>(prog + pc + 1, x, y, z, ox, oy, oz);
#pragma unroll
                for (int n = 0; n < N; n++) { x[n] = ox[n]; y[n] = oy[n]; z[n] = oz[n]; }
            } else if (opc == OPC_TWIST) {  // CompactBlobTree.cpp:1355-1385
                const float k = q0.y;
                const int axis = (int)q0.z;
#pragma unroll
                for (int n = 0; n < N; n++) {
                    float s, c;
                    const float ang = PS_MUL((axis == 0) ? x[n] : ((axis == 1) ? y[n] : z[n]), k);
                    if (WARPS) ps_sincosf(ang, s, c); else sincosf(ang, &s, &c);
                    if (axis == 0) {
                        const float ny = PS_SUB(PS_MUL(y[n], c), PS_MUL(z[n], s)), nz = PS_ADD(PS_MUL(y[n], s), PS_MUL(z[n], c));
                        y[n] = ny; z[n] = nz;
                    } else if (axis == 1) {
                        const float nx = PS_SUB(PS_MUL(x[n], c), PS_MUL(z[n], s)), nz = PS_ADD(PS_MUL(x[n], s), PS_MUL(z[n], c));
                        x[n] = nx; z[n] = nz;
                    } else {
                        const float nx = PS_SUB(PS_MUL(x[n], c), PS_MUL(y[n], s)), ny = PS_ADD(PS_MUL(x[n], s), PS_MUL(y[n], c));
                        x[n] = nx; y[n] = ny;
                    }
                }
            } else if (opc == OPC_TAPER) {  // CompactBlobTree.cpp:1387-1462 (incl. the taperAlongZ axis mix-up)
                const float k = q0.y;
                const int along = (int)q0.z, taper = (int)q0.w;
#pragma unroll
                for (int n = 0; n < N; n++) {
                    const float a = (along == 0) ? x[n] : ((along == 1) ? y[n] : z[n]);
                    const float s = PS_ADD(1.0f, PS_MUL(a, k));
                    if (along == 0) { if (taper == 2) z[n] = PS_MUL(z[n], s); else y[n] = PS_MUL(y[n], s); }
                    else if (along == 1) { if (taper == 2) z[n] = PS_MUL(z[n], s); else x[n] = PS_MUL(x[n], s); }
                    else { if (taper == 2) y[n] = PS_MUL(y[n], s); else x[n] = PS_MUL(x[n], s); }
                }
            } else if (opc == OPC_BEND) {  // CompactBlobTree.cpp:1315-1353 (Barr bend in y/z)
                const float k = q0.y, y0 = q0.z, left = q0.w;
                const float4 q1 = prog[pc + 1];
                const float right = q1.x, kd = q1.y;
#pragma unroll
                for (int n = 0; n < N; n++) {
                    const float yh = fminf(fmaxf(y[n], left), right);
                    float s, c;
                    const float th = PS_MUL(k, PS_SUB(yh, y0));
                    if (WARPS) ps_sincosf(th, s, c); else sincosf(th, &s, &c);
                    const float dy = PS_SUB(y[n], yh), zz = PS_SUB(z[n], kd);
                    float ny = PS_ADD(PS_MUL(-s, zz), y0);
                    float nz = PS_ADD(PS_MUL(c, zz), kd);
                    if (y[n] < left || y[n] > right) {  // the straight extension outside the bent region (:1339-1350)
                        ny = PS_ADD(ny, PS_MUL(c, dy));
                        nz = PS_ADD(nz, PS_MUL(s, dy));
                    }
                    y[n] = ny; z[n] = nz;
                }
            } else {  // OPC_SHEAR, CompactBlobTree.cpp:1464-1536
                const float k = q0.y;
                const int along = (int)q0.z, dep = (int)q0.w;
#pragma unroll
                for (int n = 0; n < N; n++) {
                    if (along == 1) y[n] = PS_ADD(y[n], PS_MUL(k, (dep == 2) ? z[n] : x[n]));
                    else if (along == 2) z[n] = PS_ADD(z[n], PS_MUL(k, (dep == 1) ? y[n] : x[n]));
                    else x[n] = PS_ADD(x[n], PS_MUL(k, (dep == 2) ? z[n] : y[n]));
                }
            }
        } break;
        }
        if (STRICT && COLOR && opc == OPC_COMBINE && PS_HDR_CSC(hdr)) color_scale(top[0], tc, PS_HDR_CSC(hdr));
        pc = pcNext;
    }
#pragma unroll
    for (int n = 0; n < N; n++) F[n] = top[n];
    if (COLOR) { rgb[0] = tc[0]; rgb[1] = tc[1]; rgb[2] = tc[2]; }
}

}  // namespace psgpu
#endif
